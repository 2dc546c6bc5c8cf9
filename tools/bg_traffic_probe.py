"""Does DMA traffic into HBM slow the SYRK kernel?  

One GPU, the headline workload resident in HBM.  plan.run() is timed three ways: alone; with host->device copies of a
pinned buffer looping on a side stream at full PCIe rate (copy engine, like the exchange's uploads and pulls); with
device->device copies looping on a side stream (HBM + L2 traffic without PCIe).  Prints ms_syrk per step for each.
Usage: python tools/bg_traffic_probe.py [n m]
"""
import os
import sys
import threading
import time

import torch

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
from gcta_b200 import abi, synth  # noqa: E402


def main():
    n, m = (int(sys.argv[1]), int(sys.argv[2])) if len(sys.argv) >= 3 else (50000, 1000000)
    torch.cuda.set_device(0)
    plan = abi.Plan(n, m)
    plan.synth(1236, synth.snp_thresholds(1236, m, miss_rate=0.0))
    for _ in range(2):
        plan.run()
    host = torch.empty(1 << 30, dtype=torch.uint8, pin_memory=True)
    dev = torch.empty(1 << 30, dtype=torch.uint8, device="cuda")
    dev2 = torch.empty(1 << 30, dtype=torch.uint8, device="cuda")
    side = torch.cuda.Stream()

    def timed(label, pump):
        stop = threading.Event()
        moved = [0]

        def loop():
            torch.cuda.set_device(0)
            with torch.cuda.stream(side):
                while not stop.is_set():
                    for _ in range(4):
                        pump()
                        moved[0] += 1
                    side.synchronize()

        th = threading.Thread(target=loop) if pump else None
        if th:
            th.start()
            time.sleep(0.2)
        t0 = time.perf_counter()
        ms = [plan.run().ms_syrk for _ in range(2)]
        wall = time.perf_counter() - t0
        if th:
            stop.set()
            th.join()
        print(f"{label:28s} ms_syrk/step {sum(ms) / len(ms):8.1f}   background {moved[0] * (1 << 30) / 1e9 / max(wall, 1e-9):7.1f} GB/s", flush=True)

    timed("alone", None)
    timed("with H2D copies (PCIe)", lambda: dev.copy_(host, non_blocking=True))
    timed("with D2D copies (HBM/L2)", lambda: dev2.copy_(dev, non_blocking=True))
    timed("alone again", None)


if __name__ == "__main__":
    main()
