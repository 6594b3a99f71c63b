"""Static SASS of one kernel: lists its loops (backward branches) with their instruction counts, optionally dumps a range
with its opcode histogram.
   python tools/sass_loops.py vk-renderer_b200/build/ssr_trace.o <kernel substring> [dump_start_hex dump_end_hex]"""
import collections
import re
import subprocess
import sys

obj, kern = sys.argv[1], sys.argv[2]
out = subprocess.run(["cuobjdump", "-sass", obj], capture_output=True, text=True).stdout
cur, fn = None, {}
for line in out.split("\n"):
    m = re.search(r"Function : (\S+)", line)
    if m:
        cur = m.group(1)
        fn[cur] = []
        continue
    m = re.match(r"\s+/\*([0-9a-f]{4,5})\*/\s+(.*?);", line)
    if m and cur:
        fn[cur].append((int(m.group(1), 16), m.group(2).strip()))
for name, ins in fn.items():
    if kern not in name:
        continue
    print(name, len(ins), "instructions")
    for a, t in ins:
        m = re.search(r"BRA(?:\.U)?(?:\.DIV)?\s+(?:!?U?P\d,\s*)?(0x[0-9a-f]+)", t)
        if m and int(m.group(1), 16) < a:
            print("  loop", hex(int(m.group(1), 16)), "..", hex(a), (a - int(m.group(1), 16)) // 16 + 1, "instr")
    if len(sys.argv) > 4:
        lo, hi = int(sys.argv[3], 16), int(sys.argv[4], 16)
        mix = collections.Counter()
        for a, t in ins:
            if lo <= a <= hi:
                print(hex(a), t)
                toks = t.split()
                op = toks[1] if toks[0].startswith("@") else toks[0]
                mix[op.split(".")[0]] += 1
        print(dict(mix.most_common()))
    break
