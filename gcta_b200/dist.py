"""Rank plumbing for one-process-per-GPU runs (harness: bench.py and tests).

The path shards by row bands of the lower triangle (gcta_b200_partition); ranks never exchange
data-path tensors.  torch.distributed is used only to (a) agree that every rank has finished
(barrier) and (b) reduce per-rank scalars: MAX for device times, SUM for counters.  Works with
the "nccl" backend on GPUs and with "gloo" on CPU (tests/test_dist_cpu.py, world_size 2).
"""
from __future__ import annotations

import os

from . import abi


def env_rank():
    """(rank, world, local_rank) from the torchrun environment; (0, 1, 0) when not launched by it."""
    return int(os.environ.get("RANK", "0")), int(os.environ.get("WORLD_SIZE", "1")), int(os.environ.get("LOCAL_RANK", "0"))


class Ranks:
    def __init__(self, backend: str | None = None, device=None):
        self.rank, self.world, self.local_rank = env_rank()
        self.device = device
        self._dist = None
        if self.world > 1:
            import torch.distributed as dist
            os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
            kw = {}
            if backend == "nccl" and device is not None:
                kw["device_id"] = device
            dist.init_process_group(backend or "gloo", **kw)
            self._dist = dist

    def band(self, n: int):
        """Row band [row_begin, row_end) of the packed lower triangle owned by this rank."""
        return abi.partition(n, self.rank, self.world)

    def band_pairs(self, n: int) -> int:
        b, e = self.band(n)
        return e * (e + 1) // 2 - b * (b + 1) // 2

    def barrier(self):
        if self._dist is not None:
            self._dist.barrier()

    def _reduce(self, x: float, op: str) -> float:
        if self._dist is None:
            return float(x)
        import torch
        t = torch.tensor([x], dtype=torch.float64, device=self.device if self.device is not None else "cpu")
        self._dist.all_reduce(t, op=getattr(self._dist.ReduceOp, op))
        return float(t.item())

    def max(self, x: float) -> float:
        return self._reduce(x, "MAX")

    def min(self, x: float) -> float:
        return self._reduce(x, "MIN")

    def sum(self, x: float) -> float:
        return self._reduce(x, "SUM")

    def gather(self, values):
        """list (one entry per rank) of each rank's list of floats"""
        if self._dist is None:
            return [list(map(float, values))]
        import torch
        dev = self.device if self.device is not None else "cpu"
        t = torch.tensor(list(values), dtype=torch.float64, device=dev)
        out = [torch.zeros_like(t) for _ in range(self.world)]
        self._dist.all_gather(out, t)
        return [o.cpu().tolist() for o in out]

    def broadcast(self, arr, src: int = 0):
        """numpy float64 array of the same shape on every rank -> rank `src`'s values"""
        if self._dist is None:
            return arr
        import numpy as np
        import torch
        dev = self.device if self.device is not None else "cpu"
        t = torch.from_numpy(np.ascontiguousarray(arr, dtype=np.float64)).to(dev)
        self._dist.broadcast(t, src=src)
        return t.cpu().numpy()

    def close(self):
        if self._dist is not None:
            self._dist.destroy_process_group()
            self._dist = None


def job_throughput(ranks: Ranks, n: int, m: int, steps: int, device_ms_this_rank: float) -> float:
    """Whole-job pair*SNP/s: all pairs of the triangle x SNPs x steps over the slowest rank's device time."""
    t = ranks.max(device_ms_this_rank) * 1e-3
    return n * (n + 1) / 2 * m * steps / t
