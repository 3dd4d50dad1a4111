set -x
python tools/ncu_br.py 40 4 16384 > gpurun_out/plain_br.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:k_blind_rotate_s -s 1 -c 1 -f -o gpurun_out/r2_prof_brs_v2 python tools/ncu_br.py 40 4 16384 > gpurun_out/ncu_brs.log 2>&1
ncu --metrics sm__pipe_fmaheavy_cycles_active.avg.pct_of_peak_sustained_active,sm__pipe_fmalite_cycles_active.avg.pct_of_peak_sustained_active,sm__inst_executed_pipe_fmaheavy.sum,sm__inst_executed_pipe_fmalite.sum,sm__inst_executed_pipe_alu.sum,sm__inst_executed_pipe_lsu.sum,smsp__inst_executed.sum,smsp__issue_active.avg.per_cycle_active,sm__pipe_alu_cycles_active.avg.pct_of_peak_sustained_active,dram__bytes_read.sum,dram__bytes_write.sum,gpu__time_duration.sum,l1tex__data_pipe_lsu_wavefronts.avg.pct_of_peak_sustained_elapsed,l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum,l1tex__data_pipe_lsu_wavefronts_mem_shared.sum --clock-control none -k regex:k_blind_rotate_s -s 1 -c 1 --csv --log-file gpurun_out/r2_ncu_pipes_k_blind_rotate_s_16384_v2.csv python tools/ncu_br.py 40 4 16384 > gpurun_out/ncu_pipes.log 2>&1
python tools/ncu_ntt.py > gpurun_out/plain_ntt.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:k_ntt_ -s 2 -c 2 -f -o gpurun_out/r2_prof_ntt python tools/ncu_ntt.py > gpurun_out/ncu_ntt.log 2>&1
python bench.py --steps 2 --warmup 1 --no-extras > gpurun_out/plain_bench.log 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r2_ncu_launches_bench_v2.csv python bench.py --steps 2 --warmup 1 --no-extras > gpurun_out/ncu_bench.log 2>&1
ls -la gpurun_out | tail -20
