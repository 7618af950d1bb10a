"""Multi-rank parity of the N-sharded MUL_MAT exchange (not collected by pytest: launched by tests/test_multi_gpu.py and by
hand under torchrun).  Every rank holds its contiguous row range of the same TCQ4 tensor; for each shape and each exchange
variant (fused peer stores, NVSwitch multicast stores, copy-engine pushes) the FULL C that lands in this rank's copy of the
symmetric buffer is checked against the CPU oracle (every row, every column) and, bitwise, against the NCCL all-gather.
    python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29531 tests/mp_exchange_check.py"""
import os
import sys

import numpy as np
import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
import __graft_entry__ as graft  # noqa: E402
from conftest import make_acts, make_weights, nmse  # noqa: E402


def main():
    rank, world, lr = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(lr)
    dist.init_process_group("nccl", device_id=torch.device("cuda", lr))
    import torch.distributed._symmetric_memory as symm_mem
    t4 = graft.load_package()
    orc = graft.load_oracle().Oracle()
    pbar = t4.PeerBarrier()
    shapes = [(1, 1024, 512), (8, 2048, 1024), (40, 1536, 512), (300, 2112, 1024), (600, 4096, 512)]
    worst = 0.0
    for (M, N, K) in shapes:
        rng = np.random.default_rng(M + N + K)                       # same data on every rank
        tiles = orc.quantize_tiles(make_weights(rng, N, K))
        x = make_acts(rng, M, K)
        ref = orc.mul_mat(tiles, x, N)["C"]
        sw = t4.ShardedWeight(N, K, world, rank, tiles_full=torch.from_numpy(tiles).cuda(), align=64)
        xd = torch.from_numpy(x).cuda()
        want = sw.forward(xd)                                         # NCCL all-gather of the shard outputs
        buf = symm_mem.empty((M, N), dtype=torch.float32, device="cuda")
        hdl = symm_mem.rendezvous(buf, dist.group.WORLD)
        ptrs = [int(p) for p in hdl.buffer_ptrs]
        mc = int(getattr(hdl, "multicast_ptr", 0) or 0)
        variants = ["peers", "dma"] + (["multicast"] if mc else [])
        for v in variants:
            buf.fill_(float("nan"))
            torch.cuda.synchronize()
            hdl.barrier()
            if sw.n_local:
                if v == "peers":
                    t4.mul_mat_peers(sw.tiles, xd, ptrs, N, sw.n0, sw.n_local)
                elif v == "multicast":
                    t4.mul_mat_multicast(sw.tiles, xd, mc, N, sw.n0, sw.n_local)
                else:
                    t4.mul_mat_peers(sw.tiles, xd, [ptrs[rank]], N, sw.n0, sw.n_local)
                    for r in range(world):
                        if r != rank:
                            t4.push_columns(ptrs[r], ptrs[rank], M, N, sw.n0, sw.n_local)
            pbar()                                    # tcq4_b200_peer_barrier on the compute stream: every rank's block has landed
            got = buf.clone()
            torch.cuda.synchronize()
            e = nmse(got.cpu().numpy(), ref)
            worst = max(worst, e)
            assert torch.isfinite(got).all(), (v, M, N, K, "unwritten elements")
            assert e <= 1e-9, (v, M, N, K, e)
            assert torch.equal(got, want), (v, M, N, K, "differs from the NCCL all-gather")
            hdl.barrier()
        if rank == 0:
            print(f"M={M} N={N} K={K}: {variants} all ranks' full C = oracle (NMSE <= 1e-9) = NCCL all-gather (bitwise)", flush=True)
    torch.cuda.synchronize()
    dist.barrier()
    print(f"MP_EXCHANGE_OK rank {rank} of {world} worst nmse {worst:.2e}", flush=True)
    sys.stdout.flush()
    os._exit(0)


if __name__ == "__main__":
    main()
