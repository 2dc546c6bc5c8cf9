"""N>1 host logic on CPU: two gloo ranks agree on disjoint row bands that tile the triangle, and the
whole-job throughput is computed from the slowest rank (what bench.py does with nccl on GPUs)."""
import os
import socket
import subprocess
import sys
import textwrap

from tests.helpers import ROOT

WORKER = textwrap.dedent("""
    import json, sys
    sys.path.insert(0, %r)
    from gcta_b200 import abi, dist
    r = dist.Ranks(backend="gloo")
    n, m = 5000, 1000
    b, e = r.band(n)
    pairs = r.band_pairs(n)
    tiles = len(abi.tile_list(n, b, e))
    r.barrier()
    total_pairs = r.sum(pairs)
    total_tiles = r.sum(tiles)
    fake_ms = 10.0 * (r.rank + 1)                      # rank 1 is the slow one
    thr = dist.job_throughput(r, n, m, 2, fake_ms)
    lo = r.min(b); hi = r.max(e)
    # the exchange of the packed rows, with gloo standing in for the peer copies: every rank "posts" the slices
    # gcta_b200_exchange_slice gives it, "pulls" the others' and must end up with the whole payload
    import numpy as np
    import torch.distributed as tdist
    m_file, stride = 1003, 7
    payload = ((np.arange(m_file * stride, dtype=np.int64) * 2654435761) %% 251).astype(np.uint8).reshape(m_file, stride)
    mine = abi.exchange_slices(m_file, r.world, r.rank)
    buf = np.zeros_like(payload)
    for b0, e0 in mine:
        buf[b0:e0] = payload[b0:e0]
    pieces = [None] * r.world
    tdist.all_gather_object(pieces, [(b0, e0, buf[b0:e0].tobytes()) for b0, e0 in mine])
    for q, lst in enumerate(pieces):
        if q != r.rank:
            for b0, e0, raw in lst:
                buf[b0:e0] = np.frombuffer(raw, dtype=np.uint8).reshape(e0 - b0, stride)
    # the bench line's parity check: the same subset on every rank, the oracle block broadcast from rank 0, every rank
    # compares the rows of its own band - here against a synthetic "GRM" A_ij = i + j / 1000, N_ij = i - j
    sub, bounds = dist.parity_subset(n, r.world, 64, 7)
    ii = sub[:, None].astype(np.float64); jj = sub[None, :].astype(np.float64)
    ref = np.concatenate([(ii + jj / 1000).ravel(), (ii - jj).ravel()]) if r.rank == 0 else np.zeros(2 * len(sub) ** 2)
    ref = r.broadcast(ref, 0)
    Aref, Nref = ref[:len(sub) ** 2].reshape(len(sub), -1), ref[len(sub) ** 2:].reshape(len(sub), -1)
    base = b * (b + 1) // 2
    rows = np.repeat(np.arange(b, e), np.arange(b, e) + 1); cols = np.concatenate([np.arange(i + 1) for i in range(b, e)])
    A_band = (rows + cols / 1000).astype(np.float32); N_band = (rows - cols).astype(np.int32)
    if r.rank == 1:
        bad = int(sub[-1]); N_band[bad * (bad + 1) // 2 + int(sub[0]) - base] += 1      # one wrong count in rank 1's band
    dA, n_bad, dsum, dn = dist.band_deltas(A_band, N_band, base, b, e, sub, Aref, Nref)
    par = dict(sub=[int(x) for x in sub], bounds=bounds, dA=r.max(dA), n_bad=int(r.sum(n_bad)), diag_rows=int(r.sum(dn)))
    print(json.dumps(dict(rank=r.rank, world=r.world, band=[b, e], pairs=pairs, total_pairs=total_pairs, parity=par,
                          total_tiles=total_tiles, thr=thr, lo=lo, hi=hi, exchange_ok=bool((buf == payload).all()),
                          rows_posted=int(sum(e0 - b0 for b0, e0 in mine)), segments=len(mine))))
    r.close()
""") % ROOT


def test_two_gloo_ranks_partition_and_reduce(tmp_path):
    import json
    script = tmp_path / "worker.py"
    script.write_text(WORKER)
    s = socket.socket(); s.bind(("127.0.0.1", 0)); port = s.getsockname()[1]; s.close()
    procs = []
    for rank in range(2):
        env = dict(os.environ, RANK=str(rank), WORLD_SIZE="2", LOCAL_RANK=str(rank), MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
        procs.append(subprocess.Popen([sys.executable, str(script)], env=env, stdout=subprocess.PIPE, stderr=subprocess.PIPE, text=True))
    outs = []
    for p in procs:
        o, e = p.communicate(timeout=120)
        assert p.returncode == 0, e[-2000:]
        outs.append(json.loads(o.strip().splitlines()[-1]))
    outs.sort(key=lambda d: d["rank"])
    n, m = 5000, 1000
    assert outs[0]["band"][0] == 0 and outs[0]["band"][1] == outs[1]["band"][0] and outs[1]["band"][1] == n
    assert outs[0]["pairs"] + outs[1]["pairs"] == n * (n + 1) // 2
    from gcta_b200 import abi
    for d in outs:
        assert d["world"] == 2 and d["total_pairs"] == n * (n + 1) // 2 and d["lo"] == 0 and d["hi"] == n
        assert d["total_tiles"] == len(abi.tile_list(n))
        assert abs(d["thr"] - n * (n + 1) / 2 * m * 2 / 20e-3) < 1e-3 * d["thr"]      # max over ranks = 20 ms
    # the row exchange: both ranks end with the complete payload, having uploaded about half of it each
    assert all(d["exchange_ok"] and d["segments"] == 32 for d in outs)
    assert outs[0]["rows_posted"] + outs[1]["rows_posted"] == 1003 and abs(outs[0]["rows_posted"] - outs[1]["rows_posted"]) <= 32
    # parity plumbing: same subset on both ranks, straddling the band boundary; every row of the subset checked exactly
    # once; the planted wrong count is found; float32 rounding of the synthetic A is all that differs
    p0, p1 = outs[0]["parity"], outs[1]["parity"]
    assert p0 == p1 and len(p0["sub"]) == 64 and p0["bounds"] == [outs[1]["band"][0]]
    bnd = p0["bounds"][0]
    assert {0, n - 1, bnd - 2, bnd - 1, bnd, bnd + 1} <= set(p0["sub"])
    assert p0["diag_rows"] == 64 and p0["n_bad"] == 1 and p0["dA"] < 1e-3
    # bands are balanced by tile count, not by rows
    t0, t1 = (len(abi.tile_list(n, *d["band"])) for d in outs)
    assert abs(t0 - t1) <= 0.15 * (t0 + t1)
