"""One rank of the multi-GPU peer-exchange test (launched by tests/test_gpu_exchange.py under torch.distributed.run).

Every rank holds the same host .bed payload, uploads only its 1/world slices through gcta_b200_exchange_post, pulls the
rest from its peers' HBM (exchange_pull) and computes its row band; the band must be byte-identical to the one a plain
plan_upload + plan_run + plan_download of the whole payload gives.  Host-side synchronisation uses gloo only.
"""
import os
import sys

import numpy as np
import torch
import torch.distributed as tdist

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
from gcta_b200 import abi, synth  # noqa: E402


def main():
    rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
    dev = int(os.environ.get("LOCAL_RANK", rank))
    torch.cuda.set_device(dev)
    tdist.init_process_group("gloo")
    n, m = 2900, 13000
    bed = synth.make_bed_payload(21, n, m, miss_rate=0.01)
    tri = n * (n + 1) // 2
    A0 = np.full(tri, np.nan, dtype=np.float32); N0 = np.full(tri, -1, dtype=np.int32)
    A1 = np.full(tri, np.nan, dtype=np.float32); N1 = np.full(tri, -1, dtype=np.int32)
    # different chunk lengths on the two ranks on purpose: the exchange's segments must not depend on them
    chunk = 4096 if rank == 0 else 3072
    with abi.Plan(n, m, chunk_snps=chunk, device=dev, part_rank=rank, part_world=world) as p:
        p.upload(bed); p.run(); p.download(A0, N0)
    h = torch.from_numpy(bed).pin_memory()
    with abi.Plan(n, m, chunk_snps=chunk, device=dev, part_rank=rank, part_world=world) as p:
        with abi.Exchange(p, rank, world) as ex:
            blobs = [None] * world
            tdist.all_gather_object(blobs, ex.blob)
            ex.connect(b"".join(blobs))
            for step in range(2):
                A1.fill(np.nan); N1.fill(-1)
                tdist.barrier()
                ex.post(h.data_ptr())
                tdist.barrier()
                ex.pull()
                o = p.run_to_host(None, A1, N1)
                ex.finish()
                lo, hi = o.row_begin * (o.row_begin + 1) // 2, o.row_end * (o.row_end + 1) // 2
                assert hi > lo
                assert np.array_equal(N0[lo:hi], N1[lo:hi]), f"rank {rank} step {step}: N differs"
                assert np.array_equal(A0[lo:hi], A1[lo:hi]), f"rank {rank} step {step}: A differs"
                assert (N1[:lo] == -1).all() and (N1[hi:] == -1).all()
            tdist.barrier()   # nobody reads my buffer any more
            ex.disconnect()
            tdist.barrier()   # nobody maps my buffer any more: the exchange may free it
    print(f"rank {rank}: exchange OK rows [{o.row_begin},{o.row_end})", flush=True)
    tdist.destroy_process_group()


if __name__ == "__main__":
    main()
