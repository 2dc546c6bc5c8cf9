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


def parity_subset(n: int, world: int, size: int, seed: int):
    """Individuals of the bench line's parity check (SURVEY.md 8c sub-block identity): the same seeded set on every rank,
    containing the first and last individual and both sides of every band boundary, filled up at random to `size`.
    -> (sorted int64 array, list of the band boundaries it straddles)"""
    import numpy as np
    bounds = sorted({abi.partition(n, r, world)[0] for r in range(world)} - {0})
    pick = {0, n - 1}
    for b in bounds:
        pick |= {b - 2, b - 1, b, b + 1}
    pick = {int(x) for x in pick if 0 <= x < n}
    for x in np.random.default_rng(seed).permutation(n):
        if len(pick) >= min(size, n):
            break
        pick.add(int(x))
    return np.array(sorted(pick), dtype=np.int64), bounds


def band_deltas(A_band, N_band, base: int, row_begin: int, row_end: int, sub, Aref, Nref):
    """This rank's share of the parity check: rows of `sub` inside [row_begin, row_end) against the oracle's block.
    A_band / N_band hold the packed triangle from element `base` on.  -> (max |dA|, N mismatches, sum of diagonal
    deltas, diagonal count)"""
    import numpy as np
    dA_max, n_bad, dsum, dn = 0.0, 0, 0.0, 0
    for t in range(len(sub)):
        i = int(sub[t])
        if not (row_begin <= i < row_end):
            continue
        idx = i * (i + 1) // 2 + sub[:t + 1] - base
        dA_max = max(dA_max, float(np.abs(A_band[idx].astype(np.float64) - Aref[t, :t + 1]).max()))
        n_bad += int((N_band[idx] != np.rint(Nref[t, :t + 1]).astype(np.int32)).sum())
        dsum += float(A_band[idx[-1]]) - Aref[t, t]
        dn += 1
    return dA_max, n_bad, dsum, dn


def job_throughput(ranks: Ranks, n: int, m: int, steps: int, device_ms_this_rank: float) -> float:
    """Whole-job pair*SNP/s: all pairs of the triangle x SNPs x steps over the slowest rank's device time."""
    t = ranks.max(device_ms_this_rank) * 1e-3
    return n * (n + 1) / 2 * m * steps / t
