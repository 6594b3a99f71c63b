"""Opcode mix of one kernel from an ncu report's source page: python tools/sass_mix.py report.ncu-rep kernel_regex [top]"""
import csv, subprocess, sys, collections, io
rep, kern = sys.argv[1], sys.argv[2]
top = int(sys.argv[3]) if len(sys.argv) > 3 else 25
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--kernel-name", "regex:" + kern], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(out)))
# find header row
hi = next(i for i, r in enumerate(rows) if r and r[0] == "Address")
hdr = rows[hi]
si, ei = hdr.index("Source"), hdr.index("Instructions Executed")
mix = collections.Counter(); total = 0
for r in rows[hi + 1:]:
    if r and r[0] in ("Address", "Kernel Name"): break   # the page may list the kernel a second time: one block only
    if len(r) <= ei or not r[ei]: continue
    try: n = int(float(r[ei]))
    except ValueError: continue
    toks = r[si].split()
    op = toks[0] if not toks[0].startswith("@") else (toks[1] if len(toks) > 1 else toks[0])
    op = op.rstrip(";")
    mix[op] += n; total += n
print("total warp-instr", total, "static", len(rows) - hi - 1)
for op, n in mix.most_common(top):
    print(f"{op:28s} {n:12d} {100*n/total:6.2f}%")
