"""Samples and executed warp instructions per CUDA source line of the FIRST kernel in an `ncu --set full --import-source on`
report (read here, no GPU):  python tools/ncu_lines.py rep.ncu-rep [top_n]"""
import csv, io, subprocess, sys
rep, top = sys.argv[1], int(sys.argv[2]) if len(sys.argv) > 2 else 40
txt = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "cuda,sass"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(txt)))
cur, samp, inst, files, seen_kernel = None, {}, {}, None, 0
for r in rows:
    if r and r[0] == "Kernel Name":
        seen_kernel += 1
        if seen_kernel > 1:
            break
        continue
    if r and r[0] == "File Name":
        files = r[1].split("/")[-1]
        continue
    if len(r) < 8:
        continue
    if r[0] != "":
        if r[0].isdigit():
            cur = (files, int(r[0]), r[1].strip()[:105])
        continue
    try:
        n, e = int(r[6]), int(r[7])
    except ValueError:
        continue
    samp[cur] = samp.get(cur, 0) + n
    inst[cur] = inst.get(cur, 0) + e
ts, ti = sum(samp.values()), sum(inst.values())
print(f"total samples {ts}, warp instructions {ti}")
print("samples%  inst%   file:line  source")
for k in sorted(samp, key=lambda k: -samp[k])[:top]:
    print(f"{100 * samp[k] / ts:6.2f}  {100 * inst[k] / ti:6.2f}   {k[0]}:{k[1]}  {k[2]}")
