"""N-sharded MUL_MAT with two ranks on CPU (gloo): a rank's shard is a contiguous byte range of the tile tensor that
is itself a valid [n1-n0][K] tile tensor, and ShardedWeight.gather() rebuilds ggml's dst layout C[m*N + n] from the
per-rank column blocks.  The per-rank "kernel" here is the oracle (test infrastructure) -- no GPU involved; the GPU
kernels are checked against the same oracle in test_gpu_parity.py."""
import os
import socket
import sys

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _worker(rank, world, port, cases, q):
    try:
        sys.path.insert(0, ROOT)
        sys.path.insert(0, os.path.join(ROOT, "tests"))
        import __graft_entry__ as graft
        from conftest import make_acts, make_weights
        os.environ["MASTER_ADDR"] = "127.0.0.1"
        os.environ["MASTER_PORT"] = str(port)
        dist.init_process_group("gloo", rank=rank, world_size=world)
        t4 = graft.load_package()
        orc = graft.load_oracle().Oracle()
        for (M, N, K, align) in cases:
            rng = np.random.default_rng(1000 + N + K)            # same data on both ranks
            tiles = orc.quantize_tiles(make_weights(rng, N, K))
            x = make_acts(rng, M, K)
            full = orc.mul_mat_f32(tiles, x, N)
            sw = t4.ShardedWeight(N, K, world, rank, tiles_full=torch.from_numpy(tiles), align=align)
            n0, n1, off, ln = t4.shard_range(N, K, world, rank, align)
            assert (sw.n0, sw.n1) == (n0, n1) and sw.tiles.numel() == ln == (n1 - n0) // 8 * (K // 256) * 1184
            if n1 > n0:
                yl = orc.mul_mat_f32(sw.tiles.numpy(), x, n1 - n0)           # the shard alone is a valid tile tensor
                assert np.array_equal(yl, full[:, n0:n1]), (rank, M, N, K)
            else:
                yl = np.zeros((M, 0), dtype=np.float32)
            y = sw.gather(torch.from_numpy(np.ascontiguousarray(yl)))
            assert tuple(y.shape) == (M, N)
            assert np.array_equal(y.numpy(), full), (rank, M, N, K, align)
        dist.barrier()
        dist.destroy_process_group()
        q.put((rank, "ok"))
    except Exception as e:  # noqa: BLE001
        import traceback
        q.put((rank, "".join(traceback.format_exception(e))))


def test_two_rank_shard_and_gather():
    cases = [(1, 512, 256, 64),       # decode: contiguous shards
             (5, 512, 512, 128),      # even shards, strided column blocks
             (3, 320, 256, 128),      # uneven: 256 + 64 channels
             (2, 64, 256, 128)]       # one rank owns everything, the other nothing
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, cases, q)) for r in range(2)]
    for p in procs:
        p.start()
    res = [q.get(timeout=300) for _ in procs]
    for p in procs:
        p.join(timeout=60)
    for rank, msg in res:
        assert msg == "ok", f"rank {rank}: {msg}"
