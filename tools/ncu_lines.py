#!/usr/bin/env python
"""Per CUDA source line totals of an .ncu-rep source page (needs -lineinfo + --import-source on).
usage: python tools/ncu_lines.py x.ncu-rep [min_fraction]"""
import csv, io, subprocess, sys
rep = sys.argv[1]; minf = float(sys.argv[2]) if len(sys.argv) > 2 else 0.004
txt = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "cuda,sass"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(txt)))
out = []; fname = "?"; ix = None
for r in rows:
    if not r: continue
    if r[0] == "File Path": fname = r[1].split("/")[-1]; continue
    if r[0] == "Line No": ix = {h: j for j, h in enumerate(r)}; continue
    if r[0].isdigit() and ix:
        try: out.append((fname, int(r[0]), int(r[ix['# Samples']]), int(r[ix['Instructions Executed']]), int(r[ix['Thread Instructions Executed']]), r[1][:90]))
        except ValueError: pass
tot = sum(o[3] for o in out); samp = sum(o[2] for o in out)
print(f"total warp-instr {tot/1e6:.1f}M, samples {samp}")
for o in sorted(out, key=lambda o: (o[0], o[1])):
    if o[3] > tot * minf or o[2] > samp * minf:
        print(f"{o[0][:22]:22s}:{o[1]:4d} samp={100*o[2]/samp:5.1f}% inst={o[3]/1e6:8.1f}M thr={o[4]/max(o[3],1):5.1f} | {o[5]}")
