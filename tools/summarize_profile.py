#!/usr/bin/env python
"""Turn one tools/gpu_check.sh run (gpurun_out/<tag>/) into the tracked summaries under profiles/:
  <tag>_launches.csv        the raw ncu launch list (gpu__time_duration.sum per launch, --clock-control none)
  <tag>_launch_summary.txt  per-kernel count / mean / share of the serialised step
  <tag>_<kernel>_ncu.txt    the --set full detail page of the captured kernels (text)
  <tag>_bench.json          the bench line of the same run
  traffic.json              dram bytes per launch of the channel bank (bench.py puts it into roofline.traffic)
usage: python tools/summarize_profile.py r01i"""
import collections
import csv
import json
import os
import shutil
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def main(tag):
    src = os.path.join(ROOT, "gpurun_out", tag)
    dst = os.path.join(ROOT, "profiles")
    os.makedirs(dst, exist_ok=True)
    for name in ("bench.json", "bench_ref.json", "launches.csv"):
        if os.path.exists(os.path.join(src, name)):
            shutil.copy(os.path.join(src, name), os.path.join(dst, f"{tag}_{name}"))
    lp = os.path.join(src, "launches.csv")
    if os.path.exists(lp):
        rows = [r for r in csv.reader(open(lp)) if len(r) > 5]
        hdr = [i for i, r in enumerate(rows) if r[0] == "ID"][0]
        h = rows[hdr]
        ki, vi = h.index("Kernel Name"), h.index("Metric Value")
        agg = collections.OrderedDict()
        for r in rows[hdr + 1:]:
            if r[0].isdigit():
                agg.setdefault(r[ki].split("(")[0], []).append(float(r[vi].replace(",", "")) / 1e3)
        tot = sum(sum(v) / len(v) for v in agg.values())
        with open(os.path.join(dst, f"{tag}_launch_summary.txt"), "w") as f:
            f.write("ncu --metrics gpu__time_duration.sum --clock-control none, bench.py --steps 2 --warmup 3 (cold-cache, serialised)\n")
            f.write(f"{'kernel':44s} {'launches':>8s} {'mean us':>9s} {'min us':>8s} {'share':>7s}\n")
            for k, v in agg.items():
                m = sum(v) / len(v)
                f.write(f"{k[:44]:44s} {len(v):8d} {m:9.2f} {min(v):8.2f} {100 * m / tot:6.1f}%\n")
            f.write(f"{'sum of means (one serialised block)':44s} {'':8s} {tot:9.2f}\n")
    for rep, short in (("bank_full", "bank"), ("head_full", "head")):
        rp = os.path.join(src, rep + ".ncu-rep")
        if not os.path.exists(rp):
            continue
        txt = subprocess.run(["ncu", "-i", rp, "--page", "details"], capture_output=True, text=True).stdout
        with open(os.path.join(dst, f"{tag}_{short}_ncu_details.txt"), "w") as f:
            f.write(txt)
        raw = subprocess.run(["ncu", "-i", rp, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
        rows = list(csv.reader(raw.splitlines()))
        if len(rows) >= 3 and short == "bank":
            hh, uu, vv = rows[0], rows[1], rows[2]

            def val(name):
                i = hh.index(name)
                x = float(vv[i].replace(",", ""))
                return x * {"Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "byte": 1}.get(uu[i], 1)
            rd, wr = val("dram__bytes_read.sum"), val("dram__bytes_write.sum")
            keep = {n: vv[hh.index(n)] + " " + uu[hh.index(n)] for n in (
                "gpu__time_duration.sum", "sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_active",
                "sm__inst_executed.sum", "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum.pct_of_peak_sustained_elapsed",
                "sm__cycles_active.avg", "sm__cycles_elapsed.max") if n in hh}
            with open(os.path.join(dst, "traffic.json"), "w") as f:
                json.dump({"bank_dram_bytes_per_launch": rd + wr, "dram_bytes_read": rd, "dram_bytes_write": wr,
                           "source": f"profiles/{tag}_bank_ncu_details.txt (ncu --set full, one launch of k_bank_fast)",
                           "counters": keep}, f, indent=1)
    print("profiles/ updated from", src)


if __name__ == "__main__":
    main(sys.argv[1])
