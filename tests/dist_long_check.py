"""Multi-GPU check, run under torchrun by tests/test_gpu_parity.py::test_long_horizon_split_over_real_gpus on a box with
>= 2 GPUs (skipped on the one-GPU round-end box), or by hand:
one long-horizon problem split over the ranks, both exchange transports, solution gathered on rank 0 and checked
against the oracle's KKT residual and against the single-GPU solve.

    python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 tests/dist_long_check.py
"""
import os
import sys

import numpy as np
import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import oracle_lib as O          # noqa: E402  (checker only)
import rslqr_b200 as R          # noqa: E402
from rslqr_b200.dist import LongHorizon  # noqa: E402

rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local)
dist.init_process_group("nccl", device_id=torch.device("cuda", local))
n, m, N = 12, 4, 4096
kn, x0 = R.synth_problems(n, m, N, 2024, 0, 1)
for ex in ("auto", "nccl"):      # auto = peer stores + flags when every rank could map symmetric memory, else NCCL
    lh = LongHorizon(n, m, N, rank, world, local, exchange=ex)
    lh.upload(kn[0], x0[0])
    for _ in range(5):            # several epochs: exercises the double-buffered gathered halves
        lh.solve_async()
    z = torch.from_numpy(lh.solution()).cuda()
    parts = [torch.empty_like(z) for _ in range(world)]
    dist.all_gather(parts, z)
    if rank == 0:
        zz = torch.cat(parts).cpu().numpy()
        res = O.kkt_residual(n, m, N, kn[0], x0[0], zz)
        one = R.CudaSolver(n, m, N, batch=1, flags=0)
        one.upload(kn, x0); one.solve()
        diff = np.linalg.norm(one.solution()[0] - zz) / np.linalg.norm(zz)
        one.close()
        print(f"exchange={lh.exchange} world={world}: KKT residual {res:.3e} (|z| = {np.linalg.norm(zz):.3e}), "
              f"rel diff vs single-GPU solve {diff:.3e}", flush=True)
        assert res < 1e-10 * np.linalg.norm(zz) and diff < 1e-11
    lh.close()
    dist.barrier()
dist.destroy_process_group()
