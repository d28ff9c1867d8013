"""Multi-GPU plumbing (torch.distributed only; no arithmetic here).

* batches of independent problems: contiguous split of problem ids over ranks, no data-path collective
  (first row);
* one long horizon: shard r owns knots [r*N/G, (r+1)*N/G); the only exchange is one all-gather of the per-shard
  payload (boundary entry + [A|B] of the shard's last knot, 7.2 KB at n=12) -- the top log2(G) separator levels are
  then computed redundantly on every rank (second row). Two transports: peer stores over NVLink into
  symmetric memory with arrival flags, issued by our own kernels (`exchange="p2p"`, no collective launch on the
  critical path), or an NCCL all-gather (`exchange="nccl"`, also the fallback when symmetric memory is unavailable).
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


def agree_all(ok, device=None):
    """1 only if `ok` is true on every rank (MIN all-reduce; CPU tensor over gloo, device tensor over NCCL)"""
    import torch
    import torch.distributed as dist
    if not dist.is_initialized() or dist.get_world_size() == 1:
        return int(bool(ok))
    dev = "cpu" if dist.get_backend() == "gloo" else torch.device("cuda", device)
    t = torch.tensor([int(bool(ok))], dtype=torch.int32, device=dev)
    dist.all_reduce(t, op=dist.ReduceOp.MIN)
    return int(t.item())


class LongHorizon:
    """One rank's part of a single long-horizon solve (BASELINE configs[3])."""

    def __init__(self, n, m, N, rank, world, device, exchange="auto"):
        import torch
        from .lib import ShardSolver
        self.torch = torch
        self.shard = ShardSolver(n, m, N, rank, world, device=device)
        self.world, self.rank = world, rank
        pay = self.pay = self.shard.payload_doubles
        self.d_pay = torch.zeros(pay, dtype=torch.float64, device=f"cuda:{device}")
        self.d_gat = torch.zeros(pay * world, dtype=torch.float64, device=f"cuda:{device}")
        # kernels and the collective are ordered on torch's current stream: no host synchronisation in between
        self.shard.set_stream(torch.cuda.current_stream(device).cuda_stream)
        self.exchange, self.epoch, self.symm, self.p2p_error = "nccl", 0, None, None
        if world > 1 and exchange in ("auto", "p2p"):
            try:
                self._setup_p2p(device)
                ok = 1
            except Exception as e:   # no symmetric memory on this box / build: keep the NCCL all-gather
                ok, self.p2p_error = 0, repr(e)
            # the transport is a property of the JOB, not of a rank: a rank that fell back to NCCL alone would block in
            # all_gather while the others spin on arrival flags. All ranks take p2p only if every rank could set it up.
            ok = agree_all(ok, device)
            if ok:
                self.exchange = "p2p"
            elif exchange == "p2p":
                raise RuntimeError("peer-to-peer exchange requested but not available on every rank: %s" % self.p2p_error)

    def _setup_p2p(self, device):
        """One symmetric allocation per rank: [2][world][pay] gathered payload slots (double-buffered by epoch parity)
        followed by `world` 64-bit arrival flags; every rank maps every other rank's allocation (NVLink peer access)."""
        torch = self.torch
        import torch.distributed as dist
        import torch.distributed._symmetric_memory as symm
        from .lib import MODE_SCAN
        if self.shard.mode != MODE_SCAN:
            raise RuntimeError("the peer-to-peer exchange needs the scan kernels")
        words = 2 * self.world * self.pay + self.world
        buf = symm.empty(words, dtype=torch.float64, device=torch.device("cuda", device))
        buf.zero_()
        hdl = symm.rendezvous(buf, dist.group.WORLD.group_name)
        ptrs = [int(p) for p in hdl.buffer_ptrs]
        self.symm, self.symm_hdl = buf, hdl
        self.peer_base = ptrs
        self.flag_off = 2 * self.world * self.pay * 8
        torch.cuda.synchronize()
        dist.barrier()

    def upload(self, knots_full, x0):
        self.shard.upload_shard(knots_full, x0)

    def solve_async(self):
        if self.exchange == "p2p":
            self.epoch += 1
            half = (self.epoch & 1) * self.world * self.pay * 8
            mine = self.peer_base[self.rank] + half
            self.shard.reduce(mine + self.rank * self.pay * 8)       # payload lands in the own slot of the own buffer
            self.shard.push(mine + self.rank * self.pay * 8, [b + half for b in self.peer_base],
                            [b + self.flag_off for b in self.peer_base], self.epoch)
            self.shard.finish_wait(mine, self.peer_base[self.rank] + self.flag_off, self.epoch)
            return
        self.shard.reduce(self.d_pay.data_ptr())
        if self.world > 1:
            allgather_payload(self.d_pay, self.d_gat)
        else:
            self.d_gat.copy_(self.d_pay)
        self.shard.finish(self.d_gat.data_ptr())

    def check(self):
        """Fail-stop: raise if the last solve on this rank saw a failed factorization (flag 1) or an exchange whose
        arrival flags timed out (flag 2: the cross-GPU tier then merged incomplete data and the solution is garbage)."""
        self.torch.cuda.synchronize()
        flag = int(self.shard.flags()[0])
        if flag:
            from .lib import RslqrError
            why = {1: "a separator or leaf factorization failed (pivot <= 0)",
                   2: "the peer-to-peer exchange timed out waiting for a rank's payload"}.get(flag, "flag %d" % flag)
            raise RslqrError("long-horizon solve failed on rank %d: %s" % (self.rank, why))

    def solution(self):
        self.check()
        return self.shard.solution()[0]

    def close(self):
        self.shard.close()
