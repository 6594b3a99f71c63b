"""Executed warp-instructions per CUDA source line of one kernel: python tools/ncu_lines.py report.ncu-rep kernel_regex [top]"""
import csv, subprocess, sys, io, collections
rep, kern = sys.argv[1], sys.argv[2]
top = int(sys.argv[3]) if len(sys.argv) > 3 else 30
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "cuda,sass", "--kernel-name", "regex:" + kern], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(out)))
agg = collections.OrderedDict(); cur = None; fpath = ""; ei = None; total = 0
for r in rows:
    if not r: continue
    if r[0] == "File Path": fpath = r[1].split("/")[-1]; continue
    if r[0] == "Line No": ei = r.index("Instructions Executed"); continue
    if r[0] == "Function Name" or ei is None: continue
    if r[0] != "":
        cur = (fpath, r[0], r[1].strip()[:110]); agg.setdefault(cur, 0); continue
    try: n = int(float(r[ei]))
    except (ValueError, IndexError): continue
    agg[cur] += n; total += n
print("total", total)
for (f, ln, src), n in sorted(agg.items(), key=lambda kv: -kv[1])[:top]:
    print(f"{100*n/total:5.1f}% {n/1e6:8.2f}M {f}:{ln}  {src}")
