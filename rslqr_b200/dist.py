"""Multi-GPU plumbing (torch.distributed only; no arithmetic here).

* batches of independent problems: contiguous split of problem ids over ranks, no data-path collective
  (first row);
* one long horizon: shard r owns knots [r*N/G, (r+1)*N/G); the only exchange is one all-gather of the per-shard
  payload (boundary entry + [A|B] of the shard's last knot, 7.2 KB at n=12) -- the top log2(G) separator levels are
  then computed redundantly on every rank (second row).
"""
import numpy as np


def shard_range(total, rank, world):
    """contiguous [lo, hi) of `total` items owned by `rank` (sizes differ by at most one)"""
    base, rem = divmod(total, world)
    lo = rank * base + min(rank, rem)
    return lo, lo + base + (1 if rank < rem else 0)


def allgather_payload(send, gathered, group=None):
    """all-gather of equal-sized 1-D tensors (device tensors over NCCL, CPU tensors over gloo)"""
    import torch.distributed as dist
    dist.all_gather_into_tensor(gathered, send, group=group)
    return gathered


class LongHorizon:
    """One rank's part of a single long-horizon solve (BASELINE configs[3])."""

    def __init__(self, n, m, N, rank, world, device):
        import torch
        from .lib import ShardSolver
        self.torch = torch
        self.shard = ShardSolver(n, m, N, rank, world, device=device)
        self.world = world
        pay = self.shard.payload_doubles
        self.d_pay = torch.zeros(pay, dtype=torch.float64, device=f"cuda:{device}")
        self.d_gat = torch.zeros(pay * world, dtype=torch.float64, device=f"cuda:{device}")
        # kernels and the collective are ordered on torch's current stream: no host synchronisation in between
        self.shard.set_stream(torch.cuda.current_stream(device).cuda_stream)

    def upload(self, knots_full, x0):
        self.shard.upload_shard(knots_full, x0)

    def solve_async(self):
        self.shard.reduce(self.d_pay.data_ptr())
        if self.world > 1:
            allgather_payload(self.d_pay, self.d_gat)
        else:
            self.d_gat.copy_(self.d_pay)
        self.shard.finish(self.d_gat.data_ptr())

    def solution(self):
        self.torch.cuda.synchronize()
        return self.shard.solution()[0]

    def close(self):
        self.shard.close()
