"""gcta_b200_exchange_*: the packed rows reach every rank's HBM in pieces - 1/world over PCIe, the rest pulled from
the peers over NVLink by the copy engines - and the GRM band must not change by a bit."""
import os
import subprocess
import sys

import numpy as np
import pytest

from gcta_b200 import abi, synth

pytestmark = pytest.mark.gpu
HERE = os.path.dirname(os.path.abspath(__file__))


def test_exchange_world_1_is_a_segmented_upload():
    """world = 1: no peers, post uploads every segment itself, pull only turns the segment events into the plan's
    per-chunk events; closing the exchange puts the plan back on its own genotype buffer."""
    import torch
    n, m = 1500, 13000
    bed = synth.make_bed_payload(6, n, m, miss_rate=0.02)
    tri = n * (n + 1) // 2
    A0 = np.empty(tri, dtype=np.float32); N0 = np.empty(tri, dtype=np.int32)
    A1 = np.full(tri, np.nan, dtype=np.float32); N1 = np.full(tri, -1, dtype=np.int32)
    A2 = np.full(tri, np.nan, dtype=np.float32); N2 = np.full(tri, -1, dtype=np.int32)
    with abi.Plan(n, m, chunk_snps=4096) as p:
        p.upload(bed); p.run(); p.download(A0, N0)
    h = torch.from_numpy(bed).pin_memory()
    with abi.Plan(n, m, chunk_snps=4096) as p:
        with abi.Exchange(p, 0, 1) as ex:
            assert len(ex.blob) == abi.IPC_HANDLE_BYTES * (1 + 32)
            with pytest.raises(abi.GctaB200Error):
                ex.connect(b"")   # one blob per rank
            ex.connect(ex.blob)
            for _ in range(2):
                ex.post(h.data_ptr())
                ex.pull()
                p.run_to_host(None, A1, N1)
                ex.finish()
                assert np.array_equal(N0, N1) and np.array_equal(A0, A1)
        p.upload(bed); p.run(); p.download(A2, N2)   # own buffer again, no stale chunk events
    assert np.array_equal(N0, N2) and np.array_equal(A0, A2)


def test_exchange_between_two_gpus():
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs two GPUs")
    env = dict(os.environ, MASTER_ADDR="127.0.0.1")
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2", "--master-addr", "127.0.0.1",
           "--master-port", "29533", os.path.join(HERE, "exchange_worker.py")]
    r = subprocess.run(cmd, env=env, capture_output=True, text=True, timeout=240)
    assert r.returncode == 0, r.stdout[-3000:] + r.stderr[-3000:]
    assert r.stdout.count("exchange OK") == 2
