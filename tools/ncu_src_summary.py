"""Summarise `ncu --page source --csv --print-source cuda` output: top source lines by executed instructions."""
import csv, subprocess, sys
rep = sys.argv[1]; top = int(sys.argv[2]) if len(sys.argv) > 2 else 25
txt = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "cuda,sass"], capture_output=True, text=True).stdout
rows = list(csv.reader(txt.splitlines()))
kernel = fname = None; hdr = None; out = {}; agg = {}
for r in rows:
    if not r: continue
    if r[0] == "Kernel Name": kernel = r[1]; continue
    if r[0] == "File Name": fname = r[1].split("/")[-1]; continue
    if r[0] == "Line No": hdr = r; continue
    if hdr and len(r) == len(hdr) and r[0].isdigit() and "Instructions Executed" in hdr:
        i = hdr.index("Instructions Executed")
        try: n = int(r[i] or 0)
        except ValueError: n = 0
        key = (kernel, fname, int(r[0]), r[1].strip()[:110])
        agg[key] = agg.get(key, 0) + n
for (k, f, ln, src), n in agg.items():
    if n: out.setdefault(k, []).append((n, f, ln, src, ""))
for k, v in out.items():
    tot = sum(x[0] for x in v)
    print(f"==== {k}: {tot/1e6:.1f} M warp-instructions")
    for n, f, ln, src, smp in sorted(v, reverse=True)[:top]:
        print(f"{100*n/tot:5.1f}%  {n/1e6:8.1f}M  {f}:{ln:<4} {src}")
