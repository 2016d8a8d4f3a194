#!/usr/bin/env python
"""Turn an ncu launch list (--metrics gpu__time_duration.sum --csv) and/or a full capture (.ncu-rep)
into the markdown tables kept under profiles/.

  python scripts/ncu_summary.py --launches gpurun_out/X_launches.csv --rep gpurun_out/prof_X.ncu-rep > profiles/X.md
"""
import argparse
import collections
import csv
import subprocess

COLS = [("gpu__time_duration.sum", "time"), ("dram__bytes_read.sum", "DRAM rd"), ("dram__bytes_write.sum", "DRAM wr"),
        ("launch__registers_per_thread", "regs"), ("launch__grid_size", "grid"),
        ("sm__warps_active.avg.pct_of_peak_sustained_active", "occupancy %"),
        ("sm__inst_issued.avg.pct_of_peak_sustained_active", "issue %"),
        ("gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "DRAM %"),
        ("lts__throughput.avg.pct_of_peak_sustained_elapsed", "L2 %"),
        ("smsp__inst_executed.sum", "warp inst"), ("smsp__thread_inst_executed_per_inst_executed.ratio", "lanes/inst"),
        ("sm__pipe_fp64_cycles_active.avg.pct_of_peak_sustained_active", "fp64 pipe %"),
        ("lts__t_sectors_op_red.sum", "L2 RED sectors"), ("lts__t_sectors_op_atom.sum", "L2 ATOM sectors"),
        ("l1tex__data_pipe_lsu_wavefronts_mem_shared.sum", "smem wavefronts")]
HBM_PEAK_GBS = 6543.7   # MEASURED_PEAKS.json hbm_gbs


def short(name):
    name = name.replace("b200gs::", "").replace("void ", "")
    return name.split("(")[0][:60]


def launches(path):
    rows = []
    with open(path) as f:
        lines = [l for l in f if not l.startswith("==")]
    rd = csv.reader(lines)
    hdr = next(rd)
    ki, vi, ui = hdr.index("Kernel Name"), hdr.index("Metric Value"), hdr.index("Metric Unit")
    agg = collections.OrderedDict()
    for r in rd:
        if len(r) <= vi:
            continue
        v = float(r[vi].replace(",", ""))
        v = v / 1e6 if r[ui] in ("ns", "nsecond") else (v / 1e3 if r[ui] in ("us", "usecond") else v)
        a = agg.setdefault(short(r[ki]), [0, 0.0])
        a[0] += 1
        a[1] += v
    tot = sum(a[1] for a in agg.values())
    print("| kernel | launches | total ms | avg ms | share |\n|---|---|---|---|---|")
    for k, (n, ms) in sorted(agg.items(), key=lambda kv: -kv[1][1]):
        print(f"| `{k}` | {n} | {ms:.3f} | {ms / n:.4f} | {100 * ms / tot:.1f}% |")
    print(f"\nTotal {tot:.2f} ms over {sum(a[0] for a in agg.values())} launches.")


def rep(path):
    if path.endswith(".csv"):       # `ncu -i X.ncu-rep --page raw --csv` exported on the GPU box
        out = open(path).read()
    else:
        out = subprocess.run(["ncu", "-i", path, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(out.splitlines()))
    h, u = rows[0], rows[1]
    print("| # | kernel | " + " | ".join(f"{lab} [{u[h.index(m)]}]" if u[h.index(m)] else lab for m, lab in COLS if m in h) + " | DRAM GB/s | of HBM peak |")
    print("|---|---|" + "---|" * (2 + sum(1 for m, _ in COLS if m in h)))

    def num(r, m):
        v = float(r[h.index(m)].replace(",", ""))
        unit = u[h.index(m)]
        scale = {"Gbyte": 1e9, "Mbyte": 1e6, "Kbyte": 1e3, "byte": 1.0, "ms": 1e-3, "us": 1e-6, "ns": 1e-9, "s": 1.0,
                 "msecond": 1e-3, "usecond": 1e-6, "nsecond": 1e-9, "second": 1.0}.get(unit, 1.0)
        return v * scale
    for n, r in enumerate(rows[2:]):
        vals = []
        for m, _ in COLS:
            if m in h:
                try:
                    vals.append(f"{float(r[h.index(m)].replace(',', '')):.4g}")
                except ValueError:
                    vals.append(r[h.index(m)])
        try:
            gbs = (num(r, "dram__bytes_read.sum") + num(r, "dram__bytes_write.sum")) / num(r, "gpu__time_duration.sum") / 1e9
            extra = f" {gbs:.0f} | {gbs / HBM_PEAK_GBS:.2f} |"
        except Exception:  # noqa: BLE001
            extra = " - | - |"
        print(f"| {n} | `{short(r[h.index('Kernel Name')])}` | " + " | ".join(vals) + " |" + extra)


if __name__ == "__main__":
    ap = argparse.ArgumentParser()
    ap.add_argument("--launches")
    ap.add_argument("--rep", action="append")
    a = ap.parse_args()
    if a.launches:
        launches(a.launches)
        print()
    for path in a.rep or []:
        print(f"\n### {path}\n")
        rep(path)
