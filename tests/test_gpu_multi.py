"""Single-process multi-GPU path (the reference's model, test/test_gate_gpu_multi.cc:36-72):
SetGPUNum(n) + Initialize replicate the keys with peer copies, gate batches are partitioned
across the devices, results are byte-identical to the oracle.  Needs >= 2 GPUs (skipped on a
one-GPU box; run with `gpurun --gpus 2`)."""
import numpy as np
import pytest
import torch

from oracle import orc

pytestmark = pytest.mark.gpu


def test_partitioned_batch_on_all_gpus(engine, keys):
    n = torch.cuda.device_count() if torch.cuda.is_available() else 0
    if n < 2:
        pytest.skip("needs at least 2 GPUs")
    cf = engine            # session engine (1 GPU): torn down here, restored in `finally`
    cf.CleanUp()
    cf.SetGPUNum(n)
    cf.Initialize(keys.bk, keys.ksk)
    try:
        cnt = 4 * n + 1
        b0 = np.arange(cnt) & 1
        b1 = (np.arange(cnt) >> 1) & 1
        x0 = orc.encrypt(keys, b0, 0, base_index=7000)
        x1 = orc.encrypt(keys, b1, 0, base_index=8000)
        want = orc.gate_batch(keys, orc.OP["XOR"], 0, x0, x1)
        outs = []
        for d in range(n):
            lo, hi = cf.partition(cnt, n, d)
            st = cf.Stream(d).Create()
            a = cf.Ctxt(hi - lo, 0, device=d).set(x0[lo:hi])
            b = cf.Ctxt(hi - lo, 0, device=d).set(x1[lo:hi])
            o = cf.Ctxt(hi - lo, 0, device=d)
            cf.Xor(o, a, b, st)
            outs.append(o)
        cf.Synchronize()
        got = np.concatenate([o.tlwehost for o in outs])
        assert np.array_equal(got, want)
        assert list(orc.decrypt(keys, got, 0)) == list(b0 ^ b1)
    finally:
        cf.CleanUp()
        cf.SetGPUNum(1)
        cf.Initialize(keys.bk, keys.ksk)


def test_initialize_from_archive_and_key_cache(engine, keys):
    """Initialize from a serialised TFHEpp::EvalKey and from the cache of
    converted keys; both must give the byte-identical engine (same converted key bytes, same
    gate outputs as the oracle)."""
    cf = engine
    x0 = orc.encrypt(keys, [0, 1, 1, 0, 1], 0, base_index=9100)
    x1 = orc.encrypt(keys, [1, 1, 0, 0, 1], 0, base_index=9200)
    want = orc.gate_batch(keys, orc.OP["NAND"], 0, x0, x1)

    def nand():
        st = cf.Stream(0).Create()
        a, b, o = cf.Ctxt(5, 0).set(x0), cf.Ctxt(5, 0).set(x1), cf.Ctxt(5, 0)
        cf.Nand(o, a, b, st)
        cf.Synchronize()
        return o.tlwehost.copy()

    cache = cf.ExportKeyCache(0)
    archive = cf.SerializeEvalKey(keys.bk, keys.ksk)
    cf.CleanUp()
    try:
        cf.SetGPUNum(1)
        cf.InitializeFromArchive(archive)
        assert np.array_equal(cf.ExportKeyCache(0), cache)
        assert np.array_equal(nand(), want)
        cf.CleanUp()
        cf.InitializeFromKeyCache(cache)
        assert np.array_equal(nand(), want)
        cf.CleanUp()
        bad = cache.copy(); bad[1000] ^= 1
        with pytest.raises(cf._lib.B2FheError, match="checksum"):
            cf.InitializeFromKeyCache(bad)
        rt = cf.DeserializeCtxts(cf.SerializeCtxts(want, 0), 0)
        assert np.array_equal(rt, want)
    finally:
        cf.CleanUp()
        cf.SetGPUNum(1)
        cf.Initialize(keys.bk, keys.ksk)


def test_deferred_queue_on_second_gpu(engine, keys):
    """The per-gate queue is per device (test/test_gate_gpu_multi.cc issues gates on streams of
    every GPU from one thread): queue a dependent pair on GPU 1 while GPU 0 also has work."""
    import ctypes as C
    n = torch.cuda.device_count() if torch.cuda.is_available() else 0
    if n < 2:
        pytest.skip("needs at least 2 GPUs")
    cf = engine
    L = cf._lib.lib()
    cf.CleanUp()
    cf.SetGPUNum(2)
    cf.Initialize(keys.bk, keys.ksk)
    try:
        x0 = orc.encrypt(keys, [1, 0, 1], 0, base_index=9500)
        x1 = orc.encrypt(keys, [1, 1, 0], 0, base_index=9600)
        nand = orc.gate_batch(keys, orc.OP["NAND"], 0, x0, x1)
        want = orc.gate_batch(keys, orc.OP["XOR"], 0, nand, x1)
        hosts, outs = {}, {}
        for d in (0, 1):
            with torch.cuda.device(d):
                a = torch.from_numpy(x0.view(np.int16)).cuda(d)
                b = torch.from_numpy(x1.view(np.int16)).cuda(d)
                m = torch.zeros_like(a)
                o = torch.zeros_like(a)
                h = torch.zeros(3, 637, dtype=torch.int16).pin_memory()
                torch.cuda.synchronize(d)
                for i in range(3):
                    cf._lib.check(L.b2fhe_defer_gate(d, None, orc.OP["NAND"], 0, m[i].data_ptr(), a[i].data_ptr(), b[i].data_ptr(), None))
                    cf._lib.check(L.b2fhe_defer_gate(d, None, orc.OP["XOR"], 0, o[i].data_ptr(), m[i].data_ptr(), b[i].data_ptr(), None))
                    cf._lib.check(L.b2fhe_defer_d2h(d, h[i].data_ptr(), o[i].data_ptr(), 1274))
                hosts[d], outs[d] = h, (a, b, m, o)
        assert L.b2fhe_defer_pending(0) == 9 and L.b2fhe_defer_pending(1) == 9
        cf.Synchronize()
        for d in (0, 1):
            assert np.array_equal(hosts[d].numpy().view(np.uint16), want), d
    finally:
        cf.CleanUp()
        cf.SetGPUNum(1)
        cf.Initialize(keys.bk, keys.ksk)


def test_torchrun_broadcast_parity():
    """The one-process-per-GPU path every SCALE number comes from: torch.distributed.run --nproc-per-node 2,
    rank 0 converts the keys, rank 1 receives them by NCCL broadcast (InitializeBroadcast -> b2fhe_adopt_keys);
    bench.py's post-timing parity check compares gates of EVERY rank's own slice with the oracle."""
    import json
    import os
    import subprocess
    import sys
    n = torch.cuda.device_count() if torch.cuda.is_available() else 0
    if n < 2:
        pytest.skip("needs at least 2 GPUs")
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2", "--master-addr", "127.0.0.1",
           "--master-port", "29533", os.path.join(root, "bench.py"), "--gpus", "2", "--steps", "1", "--warmup", "3",
           "--gates", "1200", "--no-extras"]
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=900, cwd=root)
    assert r.returncode == 0, r.stderr[-2000:]
    line = json.loads([ln for ln in r.stdout.splitlines() if ln.startswith("{")][-1])
    assert line["parity"]["ranks_ok"] == 2 and line["parity"]["gates_checked"] == 2 * 78
