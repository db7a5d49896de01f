"""Instruction distribution per CUDA source line of a kernel from an .ncu-rep (needs -lineinfo + --import-source on).
usage: python tools/ncu_inst.py <file.ncu-rep> [top_n]"""
import csv
import subprocess
import sys

rep = sys.argv[1]
top = int(sys.argv[2]) if len(sys.argv) > 2 else 25
txt = subprocess.run(["ncu", "-i", rep, "--page", "source", "--print-source", "cuda,sass", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(txt.splitlines()))
out, fname, ok = [], "", False
for row in rows:
    if len(row) >= 2 and row[0] == "File Path":
        fname = row[1].split("/")[-1]
    elif len(row) > 6 and row[0] == "Line No":
        i_s, i_i, i_t = row.index("Warp Stall Sampling (All Samples)"), row.index("Instructions Executed"), row.index("Thread Instructions Executed")
        ok = True
    elif ok and len(row) > 6 and row[0] != "":
        try:
            out.append((float(row[i_i] or 0), float(row[i_t] or 0), float(row[i_s] or 0), fname, row[0], row[1]))
        except ValueError:
            pass
ti, ts = sum(o[0] for o in out) or 1, sum(o[2] for o in out) or 1
print(f"warp instructions {ti:.0f}")
for i, t, s, f, ln, src in sorted(out, key=lambda o: -o[0])[:top]:
    print(f"inst {i / ti:6.3f}  stall {s / ts:6.3f}  thr/inst {t / max(i, 1):5.1f}  {f}:{ln:>4s} | {src.strip()[:100]}")
