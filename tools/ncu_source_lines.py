"""Warp-state samples per CUDA source line of one `ncu --set full --import-source on` capture (needs -lineinfo):
    python tools/ncu_source_lines.py <file.ncu-rep> [top]"""
import csv
import io
import subprocess
import sys

rep = sys.argv[1]
top = int(sys.argv[2]) if len(sys.argv) > 2 else 40
txt = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "cuda,sass"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(txt)))
cur_file, out, hdr = "", [], None
for r in rows:
    if len(r) == 2 and r[0] == "File Path":
        cur_file = r[1].split("/")[-1]
    elif r and r[0] == "Line No":
        hdr = {h: i for i, h in enumerate(r)}
    elif hdr and r and r[0].isdigit():  # a CUDA source line (SASS rows have an empty line number)
        num = lambda v: int(v) if v.isdigit() else 0
        out.append((num(r[hdr["# Samples"]]), num(r[hdr["Instructions Executed"]]), cur_file, int(r[0]), r[1].strip()))
tot = sum(o[0] for o in out)
print(f"samples on source lines: {tot}")
for smp, ins, f, ln, src in sorted(out, reverse=True)[:top]:
    print(f"{smp:7d} {100.0 * smp / max(tot, 1):5.1f}%  {ins:10d}  {f}:{ln:<4d} {src[:110]}")
