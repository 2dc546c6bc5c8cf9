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
    print(json.dumps(dict(rank=r.rank, world=r.world, band=[b, e], pairs=pairs, total_pairs=total_pairs,
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
    # bands are balanced by tile count, not by rows
    t0, t1 = (len(abi.tile_list(n, *d["band"])) for d in outs)
    assert abs(t0 - t1) <= 0.15 * (t0 + t1)
