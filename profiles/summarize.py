"""Turns ncu outputs into the small text summaries committed under profiles/.
  python profiles/summarize.py launches <launches.csv>          per-kernel launch-time table
  python profiles/summarize.py full <report.ncu-rep>            key metrics of every captured launch
(run here on the CPU box; `ncu -i` needs no GPU)."""
import collections
import csv
import io
import subprocess
import sys

KEYS = [
    ("gpu__time_duration.sum", "duration"),
    ("launch__grid_size", "grid"),
    ("launch__block_size", "block"),
    ("launch__registers_per_thread", "regs/thread"),
    ("sm__warps_active.avg.pct_of_peak_sustained_active", "warps active % of peak"),
    ("smsp__inst_executed.sum", "warp instructions"),
    ("smsp__issue_active.avg.pct_of_peak_sustained_active", "issue slots busy %"),
    ("sm__inst_executed_pipe_fp64.avg.pct_of_peak_sustained_active", "FP64 pipe % of peak (active)"),
    ("sm__pipe_fp64_cycles_active.avg.pct_of_peak_sustained_elapsed", "FP64 pipe % of peak (elapsed)"),
    ("sm__throughput.avg.pct_of_peak_sustained_elapsed", "SM throughput %"),
    ("gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "DRAM throughput %"),
    ("dram__bytes_read.sum", "dram read"),
    ("dram__bytes_write.sum", "dram write"),
    ("lts__t_sector_hit_rate.pct", "L2 hit %"),
    ("sm__cycles_elapsed.max", "SM cycles"),
]
STALLS = ["long_scoreboard", "wait", "short_scoreboard", "barrier", "math_pipe_throttle", "not_selected",
          "no_instruction", "membar", "lg_throttle", "branch_resolving", "dispatch_stall", "mio_throttle"]


def launches(path):
    rows = list(csv.reader(open(path)))
    h = next(i for i, r in enumerate(rows) if r and r[0] == "ID")
    hdr = rows[h]
    ki, vi, ui = hdr.index("Kernel Name"), hdr.index("Metric Value"), hdr.index("Metric Unit")
    agg = collections.OrderedDict()
    for r in rows[h + 1:]:
        if len(r) < len(hdr):
            continue
        name = r[ki].split("(")[0].replace("void ", "").replace("calib::", "")
        v = float(r[vi].replace(",", ""))
        if r[ui] == "ns":
            v /= 1e3
        elif r[ui] in ("ms", "msecond"):
            v *= 1e3
        agg.setdefault(name, []).append(v)
    tot = sum(sum(v) for v in agg.values())
    print("| kernel | launches | total us | share | avg us | min us | max us |")
    print("|---|---|---|---|---|---|---|")
    for n, v in sorted(agg.items(), key=lambda kv: -sum(kv[1])):
        print(f"| {n} | {len(v)} | {sum(v):.1f} | {100 * sum(v) / tot:.1f}% | {sum(v) / len(v):.2f} | {min(v):.2f} | {max(v):.2f} |")
    print(f"\ntotal device time in kernels: {tot:.1f} us over {sum(len(v) for v in agg.values())} launches "
          "(ncu serialises launches and flushes caches: compare shares, not absolutes)")


def full(path):
    out = subprocess.run(["ncu", "-i", path, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(out)))
    hdr, units, data = rows[0], rows[1], rows[2:]
    ni = hdr.index("Kernel Name")
    for r in data:
        print(f"### {r[ni]}")
        for k, label in KEYS:
            if k in hdr:
                i = hdr.index(k)
                print(f"- {label}: {r[i]} {units[i]}   (`{k}`)")
        st = []
        for s in STALLS:
            k = f"smsp__average_warps_issue_stalled_{s}_per_issue_active.ratio"
            if k in hdr:
                st.append((float(r[hdr.index(k)].replace(",", "")), s))
        st.sort(reverse=True)
        print("- stall cycles per issued instruction: " + ", ".join(f"{s} {v:.2f}" for v, s in st[:6]))
        print()


if __name__ == "__main__":
    {"launches": launches, "full": full}[sys.argv[1]](sys.argv[2])
