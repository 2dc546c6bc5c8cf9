"""Copy this round's evidence from gpurun_out/ (scratch) into profiles/ (tracked): bench lines, launch list, ncu
summaries, K2 traffic, CLI timings, SASS histogram.    python tools/assemble_profiles.py"""
import collections
import csv
import json
import os
import re
import shutil
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
G, P = os.path.join(ROOT, "gpurun_out"), os.path.join(ROOT, "profiles")


def last_json(path):
    for ln in reversed(open(path).read().strip().split("\n")):
        if ln.startswith("{"):
            return json.loads(ln)
    raise ValueError(path)


def main():
    # bench lines
    lines = [("f_bench_cfg1.json", "cfg1, 1 GPU"), ("f_bench_cfg2.json", "cfg2, 1 GPU"), ("f_bench_cfg3.json", "cfg3, 1 GPU (with --cpu-full)"),
             ("f_bench_cfg5.json", "cfg5, 1 GPU"), ("f_bench_reference.json", "--impl reference"), ("n8b_bench_cfg3_n2.json", "cfg3, 2 GPUs"),
             ("n8b_bench_cfg3_n4.json", "cfg3, 4 GPUs"), ("n8c_bench_cfg3.json", "cfg3, 8 GPUs"), ("n8b_bench_cfg4.json", "cfg4, 8 GPUs"),
             ("f_bench_cfg4_n1.json", "cfg4, 1 GPU (--force: BASELINE names it on 8)"), ("n2b_bench_cfg1.json", "cfg1, 2 GPUs (--force)"),
             ("n2b_bench_cfg2.json", "cfg2, 2 GPUs (--force)"), ("n2b_bench_cfg5.json", "cfg5, 2 GPUs (--force)"), ("n2b_bench_cfg4.json", "cfg4, 2 GPUs (--force)"),
             ("n8_bench_cfg3_prefetch.json", "cfg3, 8 GPUs, GCTA_B200_E2E_PREFETCH=1 (diagnosis run, before the two e2e fixes)")]
    with open(os.path.join(P, "r02_bench_lines.jsonl"), "w") as f:
        for fn, what in lines:
            p = os.path.join(G, fn)
            if os.path.exists(p):
                d = last_json(p)
                d["_run"] = what
                f.write(json.dumps(d) + "\n")
    for g in (1, 2, 8):
        src = os.path.join(G, f"r02_cli_cfg3_gpus{g}.jsonl")
        if os.path.exists(src):
            shutil.copy(src, P)
    if os.path.exists(os.path.join(G, "f_decode_alone.json")):
        shutil.copy(os.path.join(G, "f_decode_alone.json"), os.path.join(P, "r02_decode_alone.json"))
    # launch list + shares
    src = os.path.join(G, "r02_launches_cfg3_n50000_m1000000.csv")
    if os.path.exists(src):
        shutil.copy(src, P)
        rows = list(csv.reader(open(src)))
        h = [i for i, r in enumerate(rows) if r and r[0] == "ID"][0]
        ki, vi = rows[h].index("Kernel Name"), rows[h].index("Metric Value")
        tot, cnt = collections.Counter(), collections.Counter()
        for r in rows[h + 1:]:
            if len(r) > vi:
                name = r[ki].split("(")[0].replace("void ", "")
                tot[name] += float(r[vi].replace(",", ""))
                cnt[name] += 1
        T = sum(v for k, v in tot.items() if "synth" not in k)
        with open(os.path.join(P, "r02_launches_cfg3_shares.txt"), "w") as f:
            f.write("launch list r02_launches_cfg3_n50000_m1000000.csv (4 steps); shares of the steps' kernel time (synthetic generator excluded)\n")
            for k, v in tot.most_common():
                f.write(f"{k:40s} {cnt[k]:5d} launches {v / 1e6:10.2f} ms" + ("" if "synth" in k else f" {100 * v / T:6.2f} %") + "\n")
            if os.path.exists(os.path.join(G, "f_bench_cfg3_plain.json")):
                d = last_json(os.path.join(G, "f_bench_cfg3_plain.json"))
                ph = d["phases_ms_per_step_rank0"]
                tt = sum(ph.values())
                f.write("bench.py CUDA-event phases of the same command without ncu (ms/step): " +
                        ", ".join(f"{k} {v:.2f} ({100 * v / tt:.2f} %)" for k, v in ph.items()) + f"; step {d['ms_per_step']:.1f} ms\n")
    # ncu summaries
    for rep, out in (("r02_decode_final_n50000_m171136.ncu-rep", "r02_decode_ncu_full_n50000_m171136.txt"),
                     ("r02_syrk_final_n50000_m171136.ncu-rep", "r02_syrk_ncu_full_n50000_m171136.txt")):
        src = os.path.join(G, rep)
        if os.path.exists(src):
            txt = subprocess.run([sys.executable, os.path.join(ROOT, "tools", "ncu_summary.py"), src], capture_output=True, text=True).stdout
            open(os.path.join(P, out), "w").write(txt)
            if "syrk" in rep:
                m = {k: float(v) for k, v in re.findall(r"(dram__bytes_(?:read|write)\.sum) \| Gbyte \| ([0-9.]+)", txt)[:2]}
                json.dump({"kernel": "syrk_tiles_kernel", "source": "profiles/" + out + " (ncu --set full, first launch)", "n": 50000,
                           "snps_per_launch": 171136, "chain_snps": 32768, "dram_bytes_read": int(m["dram__bytes_read.sum"] * 1e9),
                           "dram_bytes_write": int(m["dram__bytes_write.sum"] * 1e9), "unique_operand_bytes": 171136 * 50176 * 2,
                           "note": "per launch over one full-length Z chunk (171,136 SNPs = 6 TMEM chains of 32,768) of the whole triangle of n = 50,000 - the largest launch of the default bench"},
                          open(os.path.join(P, "k2_traffic.json"), "w"), indent=1)
    # SASS histogram
    with open(os.path.join(P, "r02_sass_histogram.txt"), "w") as f:
        f.write("cuobjdump -sass of the objects linked into libgcta_b200.so: opcode counts (tensor / TMA / TMEM / memory / reduction opcodes only)\n")
        for o in ("syrk", "syrk2", "decode2"):
            sass = subprocess.run(["cuobjdump", "-sass", os.path.join(ROOT, "gcta_b200", "build", o + ".o")], capture_output=True, text=True).stdout
            c = collections.Counter()
            for ln in sass.split("\n"):
                m = re.match(r"\s+/\*[0-9a-f]{4}\*/\s+(?:@!?U?P\d+\s+)?([A-Za-z0-9_.]+)", ln)
                if m and re.match(r"(UTC|UTMA|LDTM|STTM|SYNCS|REDG|RED|ATOM|FADD2|STG|LDG|LDS|STS|PRMT|ELECT|POPC)", m.group(1)):
                    c[m.group(1)] += 1
            f.write(f"== gcta_b200/csrc/{o}.cu\n")
            for k, v in c.most_common():
                f.write(f"{v:6d} {k}\n")
    print("profiles/ updated")


if __name__ == "__main__":
    main()
