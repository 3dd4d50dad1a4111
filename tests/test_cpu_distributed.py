"""world_size-2 gloo test of the multi-GPU host logic (no GPU): the key broadcast that
replaces the reference's per-GPU upload loop (src/bootstrap_gpu.cu:116-137) and the static
partition of a gate batch across ranks; each rank evaluates ITS slice with the oracle and
the concatenation equals the single-process result (independent gates, no steady-state
collective)."""
import os
import socket

import numpy as np
import torch
import torch.distributed as dist
import torch.multiprocessing as mp


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, out_dir):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    from cufhe_b200.api import broadcast_key_buffers, partition
    from oracle import orc
    n_bk, n_ksk = 4096, 2048
    if rank == 0:
        bk_t = torch.arange(n_bk, dtype=torch.int64).to(torch.uint8)
        ksk_t = (torch.arange(n_ksk, dtype=torch.int64) * 7).to(torch.uint8)
    else:
        bk_t, ksk_t = torch.zeros(n_bk, dtype=torch.uint8), torch.zeros(n_ksk, dtype=torch.uint8)
    broadcast_key_buffers(bk_t, ksk_t, src=0)
    assert torch.equal(bk_t, torch.arange(n_bk, dtype=torch.int64).to(torch.uint8))
    assert torch.equal(ksk_t, (torch.arange(n_ksk, dtype=torch.int64) * 7).to(torch.uint8))
    # this rank's slice of a 7-gate batch (ragged on purpose)
    keys = orc.Keys(20261104)
    count = 7
    bits0, bits1 = np.arange(count) & 1, (np.arange(count) >> 1) & 1
    x0 = orc.encrypt(keys, bits0, 0, base_index=1)
    x1 = orc.encrypt(keys, bits1, 0, base_index=101)
    b, e = partition(count, world, rank)
    out = orc.gate_batch(keys, orc.OP["NAND"], 0, x0[b:e], x1[b:e], threads=2)
    np.save(os.path.join(out_dir, f"part{rank}.npy"), out)
    dist.barrier()
    dist.destroy_process_group()


def test_two_rank_broadcast_and_partition(tmp_path, keys):
    from oracle import orc
    world = 2
    mp.spawn(_worker, args=(world, _free_port(), str(tmp_path)), nprocs=world, join=True)
    parts = np.concatenate([np.load(tmp_path / f"part{r}.npy") for r in range(world)])
    count = 7
    x0 = orc.encrypt(keys, np.arange(count) & 1, 0, base_index=1)
    x1 = orc.encrypt(keys, (np.arange(count) >> 1) & 1, 0, base_index=101)
    assert np.array_equal(parts, orc.gate_batch(keys, orc.OP["NAND"], 0, x0, x1))
