"""Turn ncu outputs brought back in gpurun_out/ into the small text summaries committed under profiles/.

  python tools/summarize_ncu.py launches gpurun_out/launches.csv  > profiles/rNN_launches.md
  python tools/summarize_ncu.py full     gpurun_out/prof.ncu-rep  > profiles/rNN_ncu_full.md
"""
import collections
import csv
import re
import subprocess
import sys


def launches(path):
    rows = list(csv.reader(open(path)))
    h = next(i for i, r in enumerate(rows) if "Kernel Name" in r)
    hdr = rows[h]
    ki, vi, ui = hdr.index("Kernel Name"), hdr.index("Metric Value"), hdr.index("Metric Unit")
    agg = collections.OrderedDict()
    for r in rows[h + 1:]:
        if len(r) <= vi:
            continue
        name = re.sub(r"^void ", "", re.sub(r"\(.*", "", r[ki]).strip())
        v = float(r[vi].replace(",", ""))
        v = v / 1000.0 if r[ui] == "ns" else (v * 1000.0 if r[ui] == "ms" else v)
        a = agg.setdefault(name, [0, 0.0])
        a[0] += 1
        a[1] += v
    tot = sum(a[1] for a in agg.values())
    n = sum(a[0] for a in agg.values())
    print(f"ncu --metrics gpu__time_duration.sum --clock-control none: {n} launches, {tot:.1f} us in total")
    print("(cold-cache, serialised replays: compare SHARES, not absolutes)\n")
    print("| kernel | launches | total us | share |")
    print("|---|---:|---:|---:|")
    for k, a in sorted(agg.items(), key=lambda kv: -kv[1][1]):
        ours = "**" if k.startswith("gp_") else ""
        print(f"| {ours}{k[:100]}{ours} | {a[0]} | {a[1]:.1f} | {100 * a[1] / tot:.1f}% |")
    mine = sum(a[1] for k, a in agg.items() if k.startswith("gp_"))
    print(f"\nlibgp_b200 kernels: {100 * mine / tot:.1f}% of the profiled device time")


WANT = [
    ("gpu__time_duration.sum", "duration"),
    ("launch__registers_per_thread", "regs/thread"),
    ("launch__shared_mem_per_block_static", "static smem/block"),
    ("launch__shared_mem_per_block_dynamic", "dynamic smem/block"),
    ("sm__warps_active.avg.pct_of_peak_sustained_active", "achieved occupancy %"),
    ("smsp__issue_active.avg.pct_of_peak_sustained_active", "issue slots busy %"),
    ("sm__inst_executed.sum", "warp instructions"),
    ("sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_active", "FMA pipe %"),
    ("sm__pipe_alu_cycles_active.avg.pct_of_peak_sustained_active", "ALU pipe %"),
    ("sm__pipe_xu_cycles_active.avg.pct_of_peak_sustained_active", "XU (MUFU) pipe %"),
    ("l1tex__data_pipe_lsu_wavefronts_mem_shared.sum", "smem wavefronts"),
    ("l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum", "smem bank conflicts"),
    ("dram__bytes_read.sum", "DRAM read"),
    ("dram__bytes_write.sum", "DRAM written"),
    ("gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "DRAM throughput %"),
    ("lts__t_sector_hit_rate.pct", "L2 hit %"),
    ("sm__throughput.avg.pct_of_peak_sustained_elapsed", "SM throughput %"),
]


def full(path):
    out = subprocess.run(["ncu", "-i", path, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(out.splitlines()))
    hdr, units = rows[0], rows[1]
    idx = {h: i for i, h in enumerate(hdr)}
    stall = [h for h in hdr if h.startswith("smsp__average_warps_issue_stalled") and h.endswith("_per_issue_active.ratio")]
    if not stall:
        stall = [h for h in hdr if "issue_stalled" in h and h.endswith(".ratio")]
    print(f"ncu --set full --clock-control none --import-source on ({path})\n")
    for r in rows[2:]:
        name = re.sub(r"^void ", "", re.sub(r"\(.*", "", r[idx["Kernel Name"]]).strip())
        print(f"### {name}   grid {r[idx['Grid Size']]} block {r[idx['Block Size']]}")
        for key, label in WANT:
            if key in idx and r[idx[key]] not in ("", "n/a"):
                print(f"- {label}: {r[idx[key]]} {units[idx[key]]}")
        vals = []
        for h in stall:
            try:
                vals.append((float(r[idx[h]].replace(",", "")), h))
            except ValueError:
                pass
        vals.sort(reverse=True)
        if vals:
            top = ", ".join(f"{re.sub(r'smsp__average_warps_issue_stalled_|_per_issue_active.ratio', '', h)} {v:.2f}" for v, h in vals[:6])
            print(f"- top stall reasons (warps stalled per issue-active cycle): {top}")
        print()


if __name__ == "__main__":
    {"launches": launches, "full": full}[sys.argv[1]](sys.argv[2])
