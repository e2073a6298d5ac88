"""Per-source-line stall samples of one kernel from an ncu report (needs -lineinfo and --import-source on).
  python profiles/srclines.py <report.ncu-rep> <kernel regex> [launch index] [top N]"""
import csv
import io
import subprocess
import sys

rep, pat = sys.argv[1], sys.argv[2]
idx = int(sys.argv[3]) if len(sys.argv) > 3 else 0
top = int(sys.argv[4]) if len(sys.argv) > 4 else 30
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "cuda,sass",
                      "--kernel-name", "regex:" + pat], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(out)))
# the report repeats per launch; split at the first file header that repeats
blocks, cur, seen = [], [], set()
fname = None
for r in rows:
    if r and r[0] == "File Path":
        if r[1] in seen:
            blocks.append(cur); cur = []; seen = set()
        seen.add(r[1]); fname = r[1].split("/")[-1]
        continue
    if r and r[0] in ("Function Name",):
        continue
    if r and r[0] == "Line No":
        hdr = r
        continue
    cur.append((fname, r))
blocks.append(cur)
blk = blocks[min(idx, len(blocks) - 1)]
S = hdr.index("# Samples")
stall = {h: i for i, h in enumerate(hdr) if h.startswith("stall_") and "Not Issued" not in h}
lines = [(f, r) for f, r in blk if r and r[0].strip().isdigit() and r[S].isdigit()]
tot = sum(int(r[S]) for f, r in lines)
print(f"{len(blocks)} launch(es) in report; launch {idx}: {tot} samples on source lines")
for f, r in sorted(lines, key=lambda fr: -int(fr[1][S]))[:top]:
    n = int(r[S])
    if n == 0:
        break
    why = sorted(((int(r[i]), h[6:]) for h, i in stall.items() if r[i].isdigit() and int(r[i]) > 0), reverse=True)[:3]
    print(f"{100 * n / tot:5.1f}%  {f}:{r[0]:>4}  {r[1].strip()[:90]}   [{', '.join(f'{h} {v}' for v, h in why)}]")
