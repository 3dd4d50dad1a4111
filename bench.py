#!/usr/bin/env python
"""bench.py -- batched NAND gate bootstraps/s (BASELINE.json metric) on N B200s.

A step = one pass of the hot path over one batch: 16384 independent lvl0 NAND gate
bootstraps per GPU (BASELINE config 3, "test_perf shape"), synthetic
random ciphertexts and keys (timing does not depend on the values; correctness is
what tests/ and smoke() check).

  value  gates/s with inputs resident in HBM (b2fhe_gate_batch on device buffers)
  e2e    gates/s through the host-buffer entry point b2fhe_gate_batch_host
         (pinned host buffers, H2D + gate + D2H inside the timed region)
  roofline / int_pipe / cpu_baseline: see DESIGN.md "Measurement".

--impl reference times the CPU implementation of the same path on the host cores
(oracle/_ref TFHEpp build when present, else the oracle port).
"""
import argparse
import ctypes as C
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

GATES_PER_GPU = 16384
METRIC = "gate_bootstraps_per_sec_nand_batched"
UNIT = "gates/s"
# algorithmic figures per NAND bootstrap
KEY_BYTES_PER_GATE = 41_680_896            # NTT-domain bootstrapping key touched per gate
# dram__bytes_read.sum + dram__bytes_write.sum of ONE blind-rotation launch over 16384 gates (round 1's
# kernel, profiles/r1_ncu_traffic_k_blind_rotate16_16384_v7.csv; only used when no profiles/*traffic*.csv
# can be read): keys once + ciphertext I/O, far below the
# algorithmic key bytes because each key row is served to all gates from L2
DRAM_TRAFFIC_FALLBACK = 67_176_704 + 47_980_032


def dram_traffic_per_launch():
    """dram__bytes_read.sum + dram__bytes_write.sum of the dominant blind-rotation launch, from the latest round's
    profiles/*traffic*.csv (an `ncu --csv` of the 16384-gate launch); (bytes, file) or the r1 figure."""
    import csv
    import glob
    files = sorted(glob.glob(os.path.join(ROOT, "profiles", "*traffic*.csv")), key=os.path.basename)     # rNN_ prefix: latest round last
    for path in reversed(files):
        try:
            tot, unit_scale = 0.0, {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}
            best = {}
            for row in csv.reader(open(path)):
                for i, c in enumerate(row):
                    if c in ("dram__bytes_read.sum", "dram__bytes_write.sum") and i + 2 < len(row):
                        v = float(row[i + 2].replace(",", "")) * unit_scale.get(row[i + 1], 1)
                        best[c] = max(best.get(c, 0.0), v)          # the largest launch of the file
            tot = sum(best.values())
            if len(best) == 2 and tot > 0:
                return int(tot), os.path.basename(path)
        except Exception:
            continue
    return DRAM_TRAFFIC_FALLBACK, "profiles/r1_ncu_traffic_k_blind_rotate16_16384_v7.csv (constant)"
# integer multiply-pipe slots per gate of OUR formulation (DESIGN.md "integer roofline"): per CMUX step
# and prime 6 transforms x 5120 Shoup butterflies x 4 slots, 2 x 1024 scale factors x 4, 8192 key
# products at 3.5 slots (a row pair shares one Montgomery reduction: 7 per pair; 5 each before the
# paired kernel, when this figure was 225.3 M) and 1024 CRT lifts x 5
IMAD_SLOTS_PER_GATE = 636 * 2 * (6 * 5120 * 4 + 2 * 1024 * 4 + 8192 * 7 // 2 + 1024 * 5)


def clocks_sampler(stop, out):
    q = "clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown," \
        "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap"
    dev = os.environ.get("LOCAL_RANK", "0")
    while not stop.is_set():
        try:
            r = subprocess.run(["nvidia-smi", "-i", dev, f"--query-gpu={q}", "--format=csv,noheader,nounits"],
                               capture_output=True, text=True, timeout=5).stdout.strip().split(",")
            out.append([x.strip() for x in r])
        except Exception:
            pass
        stop.wait(0.2)


def summarize_clocks(samples):
    sm = sorted(int(s[0]) for s in samples if len(s) >= 6 and s[0].isdigit())
    if not sm:
        return {"sm_mhz": None, "sm_max_mhz": None, "reasons": []}
    names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
    reasons = [n for i, n in enumerate(names) if any(len(s) >= 6 and s[2 + i].lower().startswith("active") for s in samples)]
    mx = max(int(s[1]) for s in samples if len(s) >= 6 and s[1].isdigit())
    return {"sm_mhz": sm[len(sm) // 2], "sm_max_mhz": mx, "reasons": reasons}


def cpu_baseline(sample_gates, threads=0):
    """CPU implementation of the same path, timed on the host cores (bounded sample)."""
    import numpy as np
    ref = os.path.join(ROOT, "oracle", "_ref", "libtfhepp_ref.so")
    cores = threads or os.cpu_count()
    if os.path.exists(ref):
        L = C.CDLL(ref)
        L.ref_bench_nand.restype = C.c_double
        L.ref_bench_nand.argtypes = [C.c_int, C.c_int]
        secs = L.ref_bench_nand(sample_gates, cores)
        return {"value": sample_gates / secs, "unit": UNIT, "cores": cores, "kind": "reference",
                "sample": f"{sample_gates} lvl0 HomNAND (TFHEpp, spqlios FFT, built -march=x86-64-v3: AVX2, no AVX-512) "
                          f"over {cores} threads"}
    from oracle import orc
    keys = orc.Keys(20261104)
    bits = np.arange(sample_gates) & 1
    x0 = orc.encrypt(keys, bits, 0, base_index=1)
    x1 = orc.encrypt(keys, 1 - bits, 0, base_index=100001)
    orc.gate_batch(keys, orc.OP["NAND"], 0, x0[:cores], x1[:cores], threads=cores)   # warm-up (key NTT)
    t = time.perf_counter()
    orc.gate_batch(keys, orc.OP["NAND"], 0, x0, x1, threads=cores)
    secs = time.perf_counter() - t
    return {"value": sample_gates / secs, "unit": UNIT, "cores": cores, "kind": "port",
            "sample": f"{sample_gates} lvl0 NAND gate bootstraps, exact-integer oracle (OpenMP, {cores} threads)"}


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    # a bounded sample of the same workload: ~3 s of CPU work per step on a 16-core host
    sample = max(64, 256 * (os.cpu_count() or 1))
    vals = []
    for i in range(args.warmup + args.steps):
        b = cpu_baseline(sample)
        if i >= args.warmup:
            vals.append(b)
    v = sum(x["value"] for x in vals) / len(vals)
    b = vals[-1]
    line = {"impl": "reference", "metric": METRIC, "value": v, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": sample / v * 1e3, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "u32", "data": "synthetic",
            "config": {"workload": f"batched independent NAND gate bootstraps (lvl0 I/O, TFHEpp 128-bit params); "
                                   f"CPU arm: bounded sample of {sample} gates per step"},
            "cpu_baseline": {"value": v, "unit": UNIT, "cores": b["cores"], "kind": b["kind"], "sample": b["sample"]},
            "e2e": {"value": v, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    print(json.dumps(line), flush=True)


def parity_check(cf, L, orc, keys, local, sp, rank, world, threads):
    """AFTER the timed regions: this rank evaluates 64 NANDs + one mixed batch of all 14 ops taken from ITS OWN
    slice of a global seeded input set, and compares bytes with the oracle (the checker runs on the host,
    outside any timed region).  Returns (ok, gates_checked)."""
    import numpy as np
    import torch
    n_nand, n_mix = 64, 14
    per = n_nand + n_mix
    base = rank * per                                   # contiguous slices of one global batch
    bits0 = (np.arange(base, base + per) * 7 + 3) % 2
    bits1 = (np.arange(base, base + per) * 5 + 1) % 3 % 2
    bits2 = (np.arange(base, base + per) // 2) % 2
    x0 = orc.encrypt(keys, bits0, 0, base_index=100000 + base)
    x1 = orc.encrypt(keys, bits1, 0, base_index=300000 + base)
    x2 = orc.encrypt(keys, bits2, 0, base_index=500000 + base)
    dev = torch.device("cuda", local)
    d = [torch.from_numpy(x.view(np.int16)).to(dev) for x in (x0, x1, x2)]
    out = torch.zeros_like(d[0])
    NAND = cf.OP["NAND"]
    cf._lib.check(L.b2fhe_gate_batch(local, sp, NAND, 0, out.data_ptr(), d[0].data_ptr(), d[1].data_ptr(), None, n_nand))
    ops = np.arange(n_mix, dtype=np.uint8)               # NAND .. COPY, one each (Mux/NMux = two bootstraps)
    off = n_nand * cf.api.LVL0_WORDS * 2
    cf._lib.check(L.b2fhe_gate_batch_mixed(local, sp, ops.ctypes.data, 0, out.data_ptr() + off, d[0].data_ptr() + off,
                                           d[1].data_ptr() + off, d[2].data_ptr() + off, n_mix))
    torch.cuda.synchronize(dev)
    got = out.cpu().numpy().view(np.uint16)
    want_nand = orc.gate_batch(keys, orc.OP["NAND"], 0, x0[:n_nand], x1[:n_nand], threads=threads)
    want_mix = orc.gate_batch(keys, 0, 0, x0[n_nand:], x1[n_nand:], x2[n_nand:], ops=ops, threads=threads)
    ok = bool(np.array_equal(got[:n_nand], want_nand) and np.array_equal(got[n_nand:], want_mix))
    return ok, per


def reference_gpu_baseline(keys, sms):
    """The reference cuFHE GPU library itself (oracle/_ref/libcufhe_ref.so = its unmodified src/*.cu built for
    sm_100), timed with its own method (test/test_util.h:29-72: one gate per call round-robin over streams,
    copies included) on the same GPU.  None when the checker library is absent."""
    path = os.path.join(ROOT, "oracle", "_ref", "libcufhe_ref.so")
    if not os.path.exists(path):
        return None
    R = C.CDLL(path)
    R.cufheref_bench_nand.restype = C.c_double
    R.cufheref_bench_nand.argtypes = [C.c_int, C.c_int]
    if R.cufheref_init(keys.bk.ctypes.data_as(C.c_void_p), keys.ksk.ctypes.data_as(C.c_void_p)) != 0:
        return None
    try:
        count = 8 * sms
        secs = min(R.cufheref_bench_nand(count, 4 * sms) for _ in range(2))
        lat = min(R.cufheref_bench_nand(1, 1) for _ in range(5))
    finally:
        R.cufheref_cleanup()
    return {"gates_per_s": count / secs, "latency_ms": lat * 1e3, "gates": count, "streams": 4 * sms,
            "what": "reference cuFHE GPU build (src/*.cu unmodified, -arch=sm_100), lvl0 Nand per-gate API with "
                    "H2D + kernel + D2H per gate, its own timing method"}


def extra_configs(cf, L, torch, np, local, sp, st, dev, rank, world, dist, hbm_peak, int_peak_gslots):
    """BASELINE configs 2, 4 and 5 outside the headline timed region (same process, same clocks record).
    config 2 on rank 0's GPU only; configs 4 / 5 partitioned across ranks as cf.api.partition."""
    from cufhe_b200 import circuits
    rng = np.random.default_rng(4242 + rank)

    def ev_time(fn, reps):
        fn()
        torch.cuda.synchronize(dev)
        best = 1e30
        for _ in range(reps):
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a.record(st)
            fn()
            b.record(st)
            b.synchronize()
            best = min(best, a.elapsed_time(b))
        return best

    def allmax(ms):
        if dist is None:
            return ms
        t = torch.tensor([ms], device=dev, dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    out = {}
    # ---- config 2: NTT / INTT / negacyclic product, N = 1024, batch 2^16 (3 x 268 MB > L2) ----
    npoly = 1 << 16
    a = torch.from_numpy(rng.integers(0, 2**32, (npoly, 1024), dtype=np.uint64).astype(np.uint32).view(np.int32)).to(dev)
    b = torch.from_numpy(rng.integers(0, 1 << 18 if hasattr(L, "b2fhe_polymul1024_exact_batch") else 256, (npoly, 1024),
                                      dtype=np.int64).astype(np.int32)).to(dev)
    nt = torch.empty((npoly, 2, 1024), dtype=torch.int32, device=dev)
    o = torch.empty((npoly, 1024), dtype=torch.int32, device=dev)
    c2 = {"batch": npoly}
    poly_fn = getattr(L, "b2fhe_polymul1024_exact_batch", None) or L.b2fhe_polymul1024_batch
    # per polynomial: bytes moved / fma-pipe slots (5120 butterflies x 4 per transform and prime)
    for name, fn, nbytes, slots in (
            ("polymul", lambda: cf._lib.check(poly_fn(local, sp, o.data_ptr(), a.data_ptr(), b.data_ptr(), npoly)), 12288,
             (3 if poly_fn is not L.b2fhe_polymul1024_batch else 2) * (3 * 5120 * 4 + 1024 * 12)),
            ("ntt", lambda: cf._lib.check(L.b2fhe_ntt1024_batch(local, sp, nt.data_ptr(), a.data_ptr(), npoly)), 12288, 2 * 5120 * 4),
            ("intt", lambda: cf._lib.check(L.b2fhe_intt1024_batch(local, sp, o.data_ptr(), nt.data_ptr(), npoly)), 12288, 2 * (5120 * 4 + 1024 * 9))):
        ms = ev_time(fn, 5)
        gbs = nbytes * npoly / (ms * 1e-3) * 1e-9
        c2[name] = {"ms": ms, "per_s": npoly / (ms * 1e-3), "gb_per_s": gbs, "frac_of_hbm_measured": gbs / hbm_peak,
                    "frac_of_int_floor": (slots * npoly / (int_peak_gslots * 1e9)) / (ms * 1e-3)}
    c2["b_domain"] = "a: any u32, b < 2^18 (test/test_polynomial_mult_1024.cu:38-39), exact three-prime product" \
        if poly_fn is not L.b2fhe_polymul1024_batch else "a: any u32, |b| < 2^8 (two-prime product, |sum| < 2^52)"
    out["ntt_2e16"] = c2
    del a, b, nt, o

    def rand_ct(*lead):
        x = rng.integers(0, 2**16, (*lead, cf.api.LVL0_WORDS), dtype=np.uint64).astype(np.uint16).view(np.int16)
        return torch.from_numpy(x).to(dev)

    # ---- config 4: all 13 gates + NMux round-robin over ONE batch of 2^18 gates, partitioned ----
    total = 1 << 18
    lo, hi = cf.api.partition(total, world, rank)
    n = hi - lo
    ops = (np.arange(lo, hi) % 14).astype(np.uint8)
    d = [rand_ct(n) for _ in range(3)]
    dout = torch.empty_like(d[0])
    boots = int(np.sum(np.where(ops >= 12, 0, np.where(ops >= 10, 2, 1))))
    ms = allmax(ev_time(lambda: cf._lib.check(L.b2fhe_gate_batch_mixed(local, sp, ops.ctypes.data, 0, dout.data_ptr(),
                                                                       d[0].data_ptr(), d[1].data_ptr(), d[2].data_ptr(), n)), 1))
    tb = boots
    if dist is not None:
        t = torch.tensor([boots], device=dev, dtype=torch.float64)
        dist.all_reduce(t)
        tb = int(t.item())
    out["mixed_2e18"] = {"gates": total, "ms": ms, "gates_per_s": total / (ms * 1e-3), "bootstraps_per_s": tb / (ms * 1e-3),
                         "partition": "cf.api.partition: contiguous slices, %d gates on rank 0" % n, "steps": 1}
    del d, dout
    # ---- config 5: 32-bit ripple-carry adder over 2^14 instances, partitioned ----
    inst_total, nbits = 1 << 14, 32
    if world >= 4:
        lo, hi = cf.api.partition(inst_total, world, rank)
        inst = hi - lo
        wa, wb = circuits.Wires(nbits, inst, 0, local), circuits.Wires(nbits, inst, 0, local)
        wa.t.copy_(rand_ct(nbits, inst))
        wb.t.copy_(rand_ct(nbits, inst))
        stream = cf.Stream(local)
        stream._st = st
        a0, b0 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a0.record(st)
        circuits.ripple_carry_add(wa, wb, stream)
        b0.record(st)
        b0.synchronize()
        ms = allmax(a0.elapsed_time(b0))
        g = inst_total * circuits.adder_gate_count(nbits)
        out["adder_2e14"] = {"instances": inst_total, "bits": nbits, "gates": g, "s_per_addition_batch": ms * 1e-3,
                             "gates_per_s": g / (ms * 1e-3), "steps": 1}
    else:
        out["adder_2e14"] = {"skipped": "2^14 x 32-bit additions = 2.6 M bootstraps (~35 s on one GPU); measured at N >= 4 "
                                        "only so that the default run stays within minutes"}
    return out


def run_product(args):
    # libraries (NCCL's version banner, ...) write to fd 1: keep the real stdout for the ONE JSON line
    sys.stdout.flush()
    real_stdout = os.fdopen(os.dup(1), "w")
    os.dup2(2, 1)
    import numpy as np
    import torch
    import cufhe_b200 as cf
    L = cf._lib.lib()      # raises if the CUDA library is missing: no fallback
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    dist = None
    if world > 1:
        import torch.distributed as dist
        torch.cuda.set_device(local)
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    assert torch.cuda.is_available(), "bench.py needs a CUDA device (no CPU fallback)"
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)

    # REAL keys (the oracle's seeded generator, as smoke()): the timed regions never touch the oracle, but
    # after them every rank byte-checks gates of its own slice against it (parity_check below)
    from oracle import orc
    keys = orc.Keys(20261104)
    if world > 1:
        # rank 0 converts the keys once; every other rank receives them by NCCL broadcast
        keep = cf.InitializeBroadcast(keys.bk if rank == 0 else None, keys.ksk if rank == 0 else None,  # noqa: F841
                                      device=local, src=0)                 # (owns adopted key memory)
    else:
        cf.SetGPUNum(1)
        cf.Initialize(keys.bk, keys.ksk, device=local)

    cnt = args.gates
    rng = np.random.default_rng(99 + rank)
    shape = (cnt, cf.api.LVL0_WORDS)
    h0 = torch.from_numpy(rng.integers(0, 2**16, shape, dtype=np.uint64).astype(np.uint16).view(np.int16)).pin_memory()
    h1 = torch.from_numpy(rng.integers(0, 2**16, shape, dtype=np.uint64).astype(np.uint16).view(np.int16)).pin_memory()
    hout = torch.empty(shape, dtype=torch.int16).pin_memory()
    d0, d1 = h0.to(dev), h1.to(dev)
    dout = torch.empty_like(d0)
    e1 = torch.empty((cnt, cf.api.LVL1_WORDS), dtype=torch.int32, device=dev)
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)     # > 126 MB L2
    st = torch.cuda.current_stream(dev)
    sp = st.cuda_stream
    NAND = cf.OP["NAND"]

    def step_dev():
        cf._lib.check(L.b2fhe_gate_batch(local, sp, NAND, 0, dout.data_ptr(), d0.data_ptr(), d1.data_ptr(), None, cnt))

    def step_host():
        cf._lib.check(L.b2fhe_gate_batch_host(local, sp, NAND, None, 0, hout.data_ptr(), h0.data_ptr(), h1.data_ptr(), None, cnt))

    def step_br():
        cf._lib.check(L.b2fhe_blind_rotate_batch(local, sp, e1.data_ptr(), d0.data_ptr(), 1 << 29, cnt))

    def barrier():
        torch.cuda.synchronize(dev)
        if dist is not None:
            dist.barrier()
        torch.cuda.synchronize(dev)

    def timed(fn, steps, warmup, flush_l2=True):
        """device time of exactly `steps` steps (CUDA events on the launching stream; L2
        flushed between iterations, flush outside the timed intervals); max over ranks."""
        for _ in range(warmup):
            fn()
        barrier()
        total = 0.0
        for _ in range(steps):
            if flush_l2:
                flush.fill_(1)
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a.record(st)
            fn()
            b.record(st)
            b.synchronize()
            total += a.elapsed_time(b)
        barrier()
        if dist is not None:
            t = torch.tensor([total], device=dev, dtype=torch.float64)
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            total = float(t.item())
        return total

    stop, samples = threading.Event(), []
    th = threading.Thread(target=clocks_sampler, args=(stop, samples), daemon=True)
    th.start()
    launches0 = L.b2fhe_kernel_launches()
    ms = timed(step_dev, args.steps, args.warmup)
    launches = (L.b2fhe_kernel_launches() - launches0) * args.steps // (args.steps + args.warmup)
    ms_e2e = timed(step_host, args.steps, max(1, args.warmup // 2))
    ms_br = timed(step_br, args.steps, 1)

    # single-gate latency (the metric's second half): one NAND through the same entry points
    def one_dev():
        cf._lib.check(L.b2fhe_gate_batch(local, sp, NAND, 0, dout.data_ptr(), d0.data_ptr(), d1.data_ptr(), None, 1))

    def one_host():
        cf._lib.check(L.b2fhe_gate_batch_host(local, sp, NAND, None, 0, hout.data_ptr(), h0.data_ptr(), h1.data_ptr(), None, 1))

    lat_dev = timed(one_dev, 10, 3) / 10
    lat_dev_warm = timed(one_dev, 10, 3, flush_l2=False) / 10
    lat_host = timed(one_host, 10, 3) / 10
    stop.set()
    th.join()

    total_gates = cnt * world
    value = total_gates * args.steps / (ms * 1e-3)
    e2e = total_gates * args.steps / (ms_e2e * 1e-3)
    peaks = {}
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            peaks = json.load(f)
    except Exception:
        pass
    hbm_peak = float(peaks.get("hbm_gbs", 6650.0))
    g0, g1 = C.c_double(), C.c_double()
    cf._lib.check(L.b2fhe_int_peak(local, 0, C.byref(g0)))
    cf._lib.check(L.b2fhe_int_peak(local, 1, C.byref(g1)))

    # ---- after the timed regions: byte parity of THIS rank's path against the oracle ----
    threads = max(1, (os.cpu_count() or 1) // world)
    ok, checked = parity_check(cf, L, orc, keys, local, sp, rank, world, threads)
    ranks_ok, gates_checked = int(ok), checked
    if dist is not None:
        t = torch.tensor([int(ok), checked], device=dev, dtype=torch.int64)
        dist.all_reduce(t)
        ranks_ok, gates_checked = int(t[0].item()), int(t[1].item())
    # ---- BASELINE configs 2, 4, 5 (not the headline) ----
    configs = None
    if not args.no_extras:
        configs = extra_configs(cf, L, torch, np, local, sp, st, dev, rank, world, dist, hbm_peak, g0.value)
    line = None
    if rank == 0:
        br_s = ms_br * 1e-3 / args.steps
        achieved = KEY_BYTES_PER_GATE * cnt / br_s * 1e-9
        slots = IMAD_SLOTS_PER_GATE * cnt / br_s * 1e-9
        traffic, traffic_src = dram_traffic_per_launch()
        # ~10-15 s of CPU work: 1024 gates per host thread
        cb = cpu_baseline(max(64, 1024 * (os.cpu_count() or 1))) if world == 1 else None
        refgpu = None
        if not args.no_extras:
            sms = torch.cuda.get_device_properties(local).multi_processor_count
            try:
                refgpu = reference_gpu_baseline(keys, sms)
            except Exception as e:                                   # a checker problem must not lose the bench line
                refgpu = {"unavailable": repr(e)[:200]}
        line = {
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": ms / args.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "u32", "data": "synthetic",
            "config": {"workload": f"{cnt} independent NAND gate bootstraps per GPU (lvl0 in/out, TFHEpp 128-bit "
                                   "params n=636 N=1024 l=2 Bg=2^8 t=7), BASELINE config 3",
                       "gates_per_gpu": cnt, "partition": "contiguous slices of independent gates, keys replicated "
                                                          "by one NCCL broadcast at Initialize, no steady-state collective",
                       "l2": "256 MiB flush write between timed iterations",
                       "keys": "real keys from the oracle's seeded generator (seed 20261104); random-word ciphertexts"},
            "e2e": {"value": e2e, "unit": UNIT, "h2d_bytes_per_step": 2 * cnt * 1274 * world,
                    "d2h_bytes_per_step": cnt * 1274 * world},
            "single_gate_latency_ms": {"device_resident": lat_dev, "host_buffers": lat_host,
                                       "device_resident_keys_in_l2": lat_dev_warm,
                                       "note": "one NAND: k_blind_rotate_cl4s (one gate per 4-CTA cluster, DSMEM) + k_keyswitch_lat; L2 is flushed "
                                               "before every call for the first two figures (keys come from HBM); the third "
                                               "repeats the call without the flush"},
            "gpu_launches": int(launches),
            "clocks": summarize_clocks(samples),
            "parity": {"ranks_ok": ranks_ok, "ranks": world, "gates_checked": gates_checked,
                       "what": "after the timed regions every rank evaluates 64 NANDs + one 14-op mixed batch of its own "
                               "slice through the C ABI and compares bytes with the oracle"},
            # the binding resource is the integer multiply pipe; the HBM figure of the contract is kept beside it
            "roofline": {"bound": "int_pipe", "achieved": slots, "peak": g0.value, "unit": "Gslot/s",
                         "frac": slots / g0.value, "traffic": traffic if cnt == GATES_PER_GPU else None,
                         "traffic_source": traffic_src,
                         "kernel": "k_blind_rotate_s<4> (+ its <3> / <1> launch for the last partial wave)", "kernel_ms": br_s * 1e3,
                         "note": "achieved = algorithmic fma-pipe issue slots (IMAD = 1, IMAD.HI / IMAD.WIDE = 2; 209,707,008 per "
                                 "gate, frozen since round 1) per second of blind-rotation kernel time; peak = the IMAD rate "
                                 "measured on this GPU in this run by b2fhe_int_peak.  DRAM traffic is far below the "
                                 "algorithmic key bytes because the batch-major kernel serves each key row to all gates "
                                 "from L2 -- hence `hbm` below is well under 1 and is NOT the binding resource",
                         "hbm": {"bound": "hbm", "achieved": achieved, "peak": hbm_peak, "unit": "GB/s",
                                 "frac": achieved / hbm_peak,
                                 "note": "algorithmic key bytes (41.68 MB x gates per launch) / kernel time vs the "
                                         "MEASURED HBM copy peak (MEASURED_PEAKS.json)"}},
            "int_pipe": {"achieved_gslots_per_s": slots, "peak_gslots_per_s": g0.value, "frac": slots / g0.value,
                         "butterfly_peak_g_per_s": g1.value,
                         "note": "fma-pipe issue slots (IMAD=1, IMAD.HI/IMAD.WIDE=2) of the algorithm per second vs the "
                                 "IMAD rate measured on this GPU by b2fhe_int_peak"},
        }
        if configs is not None:
            line["configs"] = configs
        if refgpu is not None:
            line["reference_gpu"] = refgpu
        if cb is not None:
            line["cpu_baseline"] = cb
    cf.CleanUp()
    if dist is not None:
        dist.barrier()
        dist.destroy_process_group()
    if line is not None:
        real_stdout.write(json.dumps(line) + "\n")
        real_stdout.flush()
    if ranks_ok != world:
        sys.stderr.write(f"bench.py: PARITY FAILURE: {ranks_ok} of {world} ranks match the oracle\n")
        sys.exit(3)


def run_extra(args):
    """BASELINE configs 4 and 5 (not the default bench line): --workload mixed = all 14 ops
    over one batch of 2^18 gates, --workload adder = 32-bit ripple-carry adder over 2^14
    instances; the batch / the instances are partitioned across ranks (strong scaling)."""
    sys.stdout.flush()
    real_stdout = os.fdopen(os.dup(1), "w")
    os.dup2(2, 1)
    import numpy as np
    import torch
    import cufhe_b200 as cf
    from cufhe_b200 import circuits
    L = cf._lib.lib()
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    dist = None
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        import torch.distributed as dist
        dist.init_process_group("nccl", device_id=dev)
    rng = np.random.default_rng(1234)
    bk = ksk = None
    if rank == 0 or world == 1:
        bk = rng.integers(0, 2**32, cf.api.BK_RAW_WORDS, dtype=np.uint64).astype(np.uint32)
        ksk = rng.integers(0, 2**16, cf.api.KSK_RAW_WORDS, dtype=np.uint64).astype(np.uint16)
    if world > 1:
        keep = cf.InitializeBroadcast(bk, ksk, device=local, src=0)  # noqa: F841
    else:
        cf.SetGPUNum(1)
        cf.Initialize(bk, ksk, device=local)
    st = torch.cuda.current_stream(dev)
    sp = st.cuda_stream
    rng = np.random.default_rng(99 + rank)

    def rand_ct(*lead):
        a = rng.integers(0, 2**16, (*lead, cf.api.LVL0_WORDS), dtype=np.uint64).astype(np.uint16).view(np.int16)
        return torch.from_numpy(a).pin_memory()

    if args.workload == "mixed":
        total = args.gates if args.gates != GATES_PER_GPU else 1 << 18
        b, e = cf.api.partition(total, world, rank)
        cnt = e - b
        ops = (np.arange(b, e) % 14).astype(np.uint8)              # NAND .. COPY round-robin
        h = [rand_ct(cnt) for _ in range(3)]
        hout = torch.empty_like(h[0]).pin_memory()
        d = [x.to(dev) for x in h]
        dout = torch.empty_like(d[0])
        boots = int(np.sum(np.where(ops >= 12, 0, np.where(ops >= 10, 2, 1))))
        units, unit_name = total, "gates"
        desc = f"all 13 gates + NMux round-robin over one batch of {total} gates (BASELINE config 4), lvl0 I/O"

        def step_dev():
            cf._lib.check(L.b2fhe_gate_batch_mixed(local, sp, ops.ctypes.data, 0, dout.data_ptr(), d[0].data_ptr(),
                                                   d[1].data_ptr(), d[2].data_ptr(), cnt))

        def step_host():
            cf._lib.check(L.b2fhe_gate_batch_host(local, sp, 0, ops.ctypes.data, 0, hout.data_ptr(), h[0].data_ptr(),
                                                  h[1].data_ptr(), h[2].data_ptr(), cnt))
        h2d, d2h = 3 * total * 1274, total * 1274
        extra = {"bootstraps_per_step_this_rank": boots}
    else:
        total = args.gates if args.gates != GATES_PER_GPU else 1 << 14
        b, e = cf.api.partition(total, world, rank)
        inst, nbits = e - b, 32
        ha, hb = rand_ct(nbits, inst), rand_ct(nbits, inst)
        hs = torch.empty_like(ha).pin_memory()
        wa, wb = circuits.Wires(nbits, inst, 0, local), circuits.Wires(nbits, inst, 0, local)
        wa.t.copy_(ha)
        wb.t.copy_(hb)
        stream = cf.Stream(local)
        stream._st = st                                              # run on torch's current stream
        units, unit_name = total * circuits.adder_gate_count(nbits), "gates"
        desc = (f"32-bit ripple-carry adder over {total} instances (BASELINE config 5): "
                f"{circuits.adder_gate_count(nbits)} bootstraps per instance in 65 batched levels, device-resident")

        def step_dev():
            circuits.ripple_carry_add(wa, wb, stream)

        def step_host():
            wa.t.copy_(ha, non_blocking=True)
            wb.t.copy_(hb, non_blocking=True)
            s, _ = circuits.ripple_carry_add(wa, wb, stream)
            hs.copy_(s.t, non_blocking=True)
        h2d, d2h = 2 * nbits * total * 1274, nbits * total * 1274
        extra = {"levels": 65}

    def barrier():
        torch.cuda.synchronize(dev)
        if dist is not None:
            dist.barrier()
        torch.cuda.synchronize(dev)

    def timed(fn, steps, warmup):
        for _ in range(warmup):
            fn()
        barrier()
        a, bq = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(st)
        for _ in range(steps):
            fn()
        bq.record(st)
        bq.synchronize()
        total_ms = a.elapsed_time(bq)
        barrier()
        if dist is not None:
            t = torch.tensor([total_ms], device=dev, dtype=torch.float64)
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            total_ms = float(t.item())
        return total_ms

    stop, samples = threading.Event(), []
    th = threading.Thread(target=clocks_sampler, args=(stop, samples), daemon=True)
    th.start()
    l0 = L.b2fhe_kernel_launches()
    ms = timed(step_dev, args.steps, args.warmup)
    launches = (L.b2fhe_kernel_launches() - l0) * args.steps // (args.steps + args.warmup)
    ms_e2e = timed(step_host, args.steps, 1)
    stop.set()
    th.join()
    if rank == 0:
        line = {"metric": METRIC.replace("nand_batched", args.workload), "value": units * args.steps / (ms * 1e-3),
                "unit": f"{unit_name}/s", "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
                "ms_per_step": ms / args.steps, "higher_is_better": True, "scaling": "strong", "vs_baseline": None,
                "dtype": "u32", "data": "synthetic",
                "config": {"workload": desc, "l2": "working set per step exceeds L2 (inputs > 126 MB)", **extra},
                "e2e": {"value": units * args.steps / (ms_e2e * 1e-3), "unit": f"{unit_name}/s",
                        "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h},
                "gpu_launches": int(launches), "clocks": summarize_clocks(samples)}
        real_stdout.write(json.dumps(line) + "\n")
        real_stdout.flush()
    cf.CleanUp()
    if dist is not None:
        dist.barrier()
        dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b2fhe", choices=["b2fhe", "reference"])
    ap.add_argument("--gates", type=int, default=GATES_PER_GPU, help="gates per GPU per step (nand); total gates "
                    "(mixed) / total instances (adder)")
    ap.add_argument("--no-extras", action="store_true", help="skip BASELINE configs 2/4/5 and the reference-GPU baseline")
    ap.add_argument("--workload", default="nand", choices=["nand", "mixed", "adder"],
                    help="nand = BASELINE config 3 (the bench line); mixed = config 4; adder = config 5")
    args = ap.parse_args()
    if args.impl == "reference":
        run_reference(args)
    elif args.workload != "nand":
        run_extra(args)
    else:
        run_product(args)


if __name__ == "__main__":
    main()
