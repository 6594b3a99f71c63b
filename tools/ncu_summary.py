"""Summarise an `ncu --set full` report of the chain into profiles/<name>_summary.json and refresh profiles/traffic.json.

  python tools/ncu_summary.py gpurun_out/r01e_full.ncu-rep profiles/r01e_ncu_full_4k_summary.json "<cmd>" "<note>" [3840x2160] [notraffic]

One entry per kernel (first launch of each name in the report): duration, DRAM bytes, cache hit rates, issue-active,
pipe utilisation, achieved occupancy, registers, executed warp instructions.
"""
import csv
import io
import json
import os
import subprocess
import sys

rep, out, cmd, note = sys.argv[1:5]
size = sys.argv[5] if len(sys.argv) > 5 else "3840x2160"
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
hdr, units = rows[0], rows[1]


def num(d, key, scale=1.0):
    try:
        return float(d[key].replace(",", "")) * scale
    except (KeyError, ValueError):
        return None


def unit_scale(key, want):
    u = units[hdr.index(key)] if key in hdr else ""
    table = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "ns": 1e-3, "us": 1.0, "ms": 1e3, "usecond": 1.0, "nsecond": 1e-3, "msecond": 1e3}
    return table.get(u, 1.0) / (1e6 if want == "MB" else 1.0)


TASK = {"hiz_tile": "DownsampleGbuffer+DownsampleDepth", "sssr_trace": "SSSR_trace", "sssr_filter": "SSSR_filter", "sssr_blur": "SSSR_blur",
        "gtao_main": "GTAO_main", "gtao_filter": "GTAO_filter", "gtao_accumulate": "GTAO_accumulate", "deferred_shading": "DeferedShading",
        "taa_resolve": "TAA"}
seen, kernels, traffic, winst = set(), [], {}, {}
for r in rows[2:]:
    d = dict(zip(hdr, r))
    name = d["Kernel Name"].split("(")[0].replace("vkr::", "").strip()
    if name in seen:
        continue
    seen.add(name)
    rd = num(d, "dram__bytes_read.sum", unit_scale("dram__bytes_read.sum", "MB"))
    wr = num(d, "dram__bytes_write.sum", unit_scale("dram__bytes_write.sum", "MB"))
    kernels.append({
        "kernel": name,
        "time_us": num(d, "gpu__time_duration.sum", unit_scale("gpu__time_duration.sum", "us")),
        "dram_read_MB": rd, "dram_write_MB": wr,
        "l2_hit_pct": num(d, "lts__t_sector_hit_rate.pct"), "l1_hit_pct": num(d, "l1tex__t_sector_hit_rate.pct"),
        "issue_active_pct": num(d, "smsp__issue_active.avg.pct_of_peak_sustained_active"),
        "pipe_fma_cycles_pct": num(d, "sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_active"),
        "pipe_alu_pct": num(d, "sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active"),
        "pipe_lsu_pct": num(d, "sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active"),
        "warps_active_pct": num(d, "sm__warps_active.avg.pct_of_peak_sustained_active"),
        "regs": num(d, "launch__registers_per_thread"),
        "warp_inst": num(d, "smsp__inst_executed.sum"),
    })
    for key, task in TASK.items():
        if key in name and rd is not None and wr is not None:
            traffic[task] = traffic.get(task, 0.0) + (rd + wr) * 1e6       # hiz: tile kernel only (the tail kernel adds ~12 KB)
            winst[task] = winst.get(task, 0.0) + (kernels[-1]["warp_inst"] or 0.0)
json.dump({"cmd": cmd, "size": size, "note": note, "kernels": kernels}, open(out, "w"), indent=1)
if not (len(sys.argv) > 6 and sys.argv[6] == "notraffic"):     # a report of kernels outside the chain leaves traffic.json alone
    tpath = os.path.join(os.path.dirname(out), "traffic.json")
    tj = json.load(open(tpath)) if os.path.exists(tpath) else {}
    tj[size] = traffic
    tj[size + ":warp_inst"] = winst                 # executed warp instructions per launch (smsp__inst_executed.sum)
    tj["source"] = f"{out} (dram__bytes_read.sum + dram__bytes_write.sum per launch)"
    # provenance: bench.py prints traffic_stale when the kernels it times are not the ones profiled here.  The summary is
    # written in the build container right after the capture, from the same tree the capture ran on.
    import glob
    import hashlib
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    hsh = hashlib.sha256()
    for f in sorted(glob.glob(os.path.join(root, "vk-renderer_b200", "csrc", "*.cu*"))):
        hsh.update(os.path.basename(f).encode())
        hsh.update(open(f, "rb").read())
    head = subprocess.run(["git", "-C", root, "rev-parse", "--short", "HEAD"], capture_output=True, text=True).stdout.strip()
    dirty = bool(subprocess.run(["git", "-C", root, "status", "--porcelain", "--", "vk-renderer_b200/csrc"], capture_output=True, text=True).stdout.strip())
    tj["_captured_at"] = {"git_head": head + ("+dirty" if dirty else ""), "csrc_sha16": hsh.hexdigest()[:16]}
    json.dump(tj, open(tpath, "w"), indent=1)
for kk in kernels:
    print(f"{kk['kernel'][:44]:44s} {kk['time_us']:8.1f} us  issue {kk['issue_active_pct']:5.1f}%  fma {kk['pipe_fma_cycles_pct']:5.1f}%  warps {kk['warps_active_pct']:5.1f}%  "
          f"regs {kk['regs']:.0f}  dram {kk['dram_read_MB'] + kk['dram_write_MB']:7.1f} MB  inst {kk['warp_inst'] / 1e6:7.1f}M")
