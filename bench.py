#!/usr/bin/env python
"""bench.py — full image-space chain (Hi-Z, SSR trace/filter/blur, GTAO main/filter/accumulate, deferred lighting, TAA)
on the 3840x2160 synthetic TAA sequence (BASELINE.json configs[2]); metric Gpix/s (ms/frame in ms_per_step).

    python bench.py --gpus N --steps K --warmup W            # this repo: C++ host -> C ABI -> sm_100a kernels
    python bench.py --impl reference --gpus N --steps K ...  # the CPU oracle (restated reference shaders) on the host cores

A step = one frame of the chain.  N > 1 (torchrun): every rank renders its own independent sequence (sequences
shard, frames of one sequence do not: SURVEY §8e) -> weak scaling, no data-path collective.
"""
import argparse
import json
import os
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path[:0] = [os.path.join(ROOT, "vk-renderer_b200")]

# algorithmic bytes per full-res pixel of every task (the fused Hi-Z task = 9 + 1.67)
ALG_BYTES = {"DownsampleGbuffer+DownsampleDepth": 10.67, "DownsampleGbuffer": 9.0, "DownsampleDepth": 1.67, "SSSR_trace": 10.33,
             "SSSR_filter": 16.0, "SSSR_blur": 14.0, "GTAO_main": 13.0, "GTAO_filter": 3.5, "GTAO_accumulate": 5.5,
             "DeferedShading": 23.0, "TAA": 32.0}
KERNELS_PER_FRAME = 10  # hiz tile + hiz tail, trace, filter, blur, gtao main/filter/accumulate, shading, taa


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=64)
    ap.add_argument("--warmup", type=int, default=4)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--width", type=int, default=3840)
    ap.add_argument("--height", type=int, default=2160)
    ap.add_argument("--workload", default="chain", choices=["chain", "probes", "tiles"],
                    help="chain: full chain on the 3840x2160 TAA sequence (BASELINE configs[2], the headline); "
                         "probes: 512 independent 256x256 views, Hi-Z+GTAO+SSR, sharded over the ranks (configs[3]); "
                         "tiles: ONE 7680x4320 frame sequence split into row tiles, halo rows over NVLink peer memory (configs[4])")
    ap.add_argument("--views", type=int, default=512)
    ap.add_argument("--tiles-gather", default="none", choices=["none", "rank0"],
                    help="tiles workload: leave each rank's finished TAA rows resident where they were computed (none), or send them to rank 0 every "
                         "frame (rank0: 265 MB per 8K frame into one GPU)")
    ap.add_argument("--no-tiles-balance", dest="tiles_balance", action="store_false", help="tiles workload: keep the even row split")
    ap.add_argument("--probes-lanes", type=int, default=2, choices=[1, 2],
                    help="probes workload: 2 = SSR filter + blur on a second stream beside the GTAO passes inside the frame's CUDA graph")
    ap.add_argument("--probes-groups", type=int, default=3,
                    help="probes workload: split a rank's views into this many layer ranges, each running its chain on its own streams inside the "
                         "frame's CUDA graph (independent views fill each other's launch gaps and tails)")
    ap.add_argument("--probes-gather", default="fused", choices=["allgather", "p2p", "fused", "none"],
                    help="probes workload: how every rank gets all per-view results each frame — allgather: NCCL all_gather_into_tensor on a "
                         "side stream; p2p: copy-engine pushes into every rank's cudaIpc slab with stream-side flags (vkr_gather.PeerGather: no "
                         "SMs taken from the passes); fused (default): the SSR blur and GTAO accumulate kernels store every result texel into every "
                         "rank's arrays themselves over NVLink (vkr_gather.FusedGather: no collective, no copy; measured 6.96x at 8 GPUs against "
                         "6.70x for allgather and 6.0x for p2p); none: skip it (diagnostic: what the gather costs the frame rate)")
    ap.add_argument("--probes-gather-buffers", type=int, default=4, help="probes workload, p2p gather: landing buffers per tensor on every rank")
    ap.add_argument("--nccl-max-ctas", type=int, default=0,
                    help="probes workload: cap the CTAs NCCL may use per collective (0 = NCCL's default).  The gather moves 64 MB per frame into "
                         "every GPU while the passes want every SM; a small cap leaves the SMs to the passes")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-sub", action="store_true", help="chain workload: skip the probes / tiles sub-records")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    return ap.parse_args()


def peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        return float(json.load(open(p))["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    return 6650.0, "fallback (B200_PROFILING.md 6.65 TB/s)"


class ClockSampler(threading.Thread):
    """Samples SM clock and throttle reasons through NVML while the timed region runs.  Only rank 0 samples (its own GPU; the
    line it prints is the only consumer), every BENCH_CLOCK_INTERVAL_MS (default 20 ms); BENCH_CLOCKS=off / all-ranks-5ms exist
    for the A/B that showed sampling does not affect the result (profiles/r02n_scale8_sampler_ab.txt)."""

    def __init__(self, index):
        super().__init__(daemon=True)
        self.index, self.samples, self.reasons, self.max_mhz, self.stop_flag = index, [], set(), None, False
        mode = os.environ.get("BENCH_CLOCKS", "rank0")
        self.interval = float(os.environ.get("BENCH_CLOCK_INTERVAL_MS", "5" if mode == "all-ranks-5ms" else "20")) * 1e-3
        self.enabled = mode != "off" and (mode == "all-ranks-5ms" or int(os.environ.get("RANK", 0)) == 0)

    def run(self):
        if not self.enabled:
            self.reasons.add("not_sampled_on_this_rank")
            return
        try:
            import pynvml as nv
            nv.nvmlInit()
            h = nv.nvmlDeviceGetHandleByIndex(self.index)
            self.max_mhz = nv.nvmlDeviceGetMaxClockInfo(h, nv.NVML_CLOCK_SM)
            names = {nv.nvmlClocksThrottleReasonSwPowerCap: "sw_power_cap", nv.nvmlClocksThrottleReasonHwSlowdown: "hw_slowdown",
                     nv.nvmlClocksThrottleReasonHwThermalSlowdown: "hw_thermal_slowdown",
                     nv.nvmlClocksThrottleReasonSwThermalSlowdown: "sw_thermal_slowdown",
                     nv.nvmlClocksThrottleReasonHwPowerBrakeSlowdown: "hw_power_brake"}
            while not self.stop_flag:
                self.samples.append(nv.nvmlDeviceGetClockInfo(h, nv.NVML_CLOCK_SM))
                r = nv.nvmlDeviceGetCurrentClocksThrottleReasons(h)
                for bit, name in names.items():
                    if r & bit:
                        self.reasons.add(name)
                time.sleep(self.interval)
        except Exception as e:  # NVML missing: report that instead of inventing numbers
            self.reasons.add(f"nvml_unavailable:{type(e).__name__}")

    def result(self):
        s = sorted(self.samples)
        return {"sm_mhz": s[len(s) // 2] if s else None, "sm_max_mhz": self.max_mhz, "samples": len(s), "reasons": sorted(self.reasons)}


BAND = (0.375, 0.625)   # the CPU legs compute this band of output rows of every pass (a quarter of the frame)


def cpu_chain_rate(width, height, frames, warm, threads=0):
    """Oracle chain on the host cores (the checker doubling as CPU baseline).  Every pass computes the BAND of rows of
    the real frame (the Hi-Z pyramid is built in full); returns (Gpix/s over the band pixels, ms per band-frame, threads, band rows)."""
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import chain
    import vkr
    b = chain.OracleBackend(rows=BAND)
    if threads:
        b.o.lib.orc_set_threads(threads)
    nthreads = b.o.lib.orc_get_max_threads()
    full = chain.OracleBackend()
    ch = chain.Chain(width, height, full, lut_size=1024)     # LUT pre-integration is init-time work, not timed
    ch.b = b
    scene = vkr.synth_scene(0, 64)
    t = 0.0
    for f in range(warm + frames):
        P = ch.begin_frame(scene)                            # G-buffer production is the input, not timed
        t0 = time.perf_counter()
        ch.passes(P)
        dt = time.perf_counter() - t0
        ch.end_frame()
        if f >= warm:
            t += dt
    band_rows = int(BAND[1] * height) - int(BAND[0] * height)
    return width * band_rows * frames / t / 1e9, 1e3 * t / frames, nthreads, band_rows


def run_reference(args, rank, world):
    if rank != 0:
        return
    # This arm measures the CPU side only: lay out images and produce the input G-buffers with oracle/libvkr_hostonly.so (the
    # same host helpers compiled without device code) so that the process never loads libvkr_b200.so.
    os.environ["VKR_HOST_ONLY"] = "1"
    if not os.path.exists(os.path.join(ROOT, "oracle", "libvkr_hostonly.so")):
        import subprocess
        subprocess.run(["make", "-C", os.path.join(ROOT, "oracle"), "libvkr_hostonly.so"], check=True, capture_output=True)
    gpix, ms, threads, band_rows = cpu_chain_rate(args.width, args.height, args.steps, args.warmup, threads=host_threads())
    sample = (f"each step = rows [{int(BAND[0]*args.height)}, {int(BAND[1]*args.height)}) of one {args.width}x{args.height} frame through every pass "
              f"({band_rows} of {args.height} rows; Hi-Z built in full), OpenMP over rows")
    line = {"impl": "reference", "metric": "chain_throughput", "value": gpix, "unit": "Gpix/s", "n_gpus": args.gpus, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": ms, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32",
            "data": "synthetic",
            "config": {"workload": f"full chain (Hi-Z+SSR+GTAO+lighting+TAA), {args.width}x{args.height} TAA sequence", "sample": sample},
            "cpu_baseline": {"value": gpix, "unit": "Gpix/s", "cores": threads, "kind": "port", "sample": sample},
            "e2e": {"value": gpix, "unit": "Gpix/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "native_libraries": sorted({os.path.relpath(ln.split()[-1], ROOT) for ln in open("/proc/self/maps") if "/libvkr" in ln or "/liboracle" in ln}),
            "note": "the reference's GLSL cannot run here (no Vulkan / glslc): this is the CPU restatement of its shaders (oracle/)"}
    print(json.dumps(line))


_NCCL_GROUP = None


def init_dist(args, world):
    """One process per GPU.  The DEFAULT group is gloo: barriers and max-over-ranks carry no GPU data, and the replica and
    row-tile workloads exchange nothing through a collective.  NCCL is created lazily (nccl_group) by the one workload that
    really gathers device tensors, the probe batch ("NCCL only for the final gather").  This is a matter of scope, not of
    speed: the chain runs equally fast with either control plane (profiles/r02p_scale4_matrix.txt).  BENCH_CONTROL=nccl makes
    NCCL the default group."""
    import torch
    import torch.distributed as dist
    local = int(os.environ.get("LOCAL_RANK", os.environ.get("RANK", 0)))
    torch.cuda.set_device(local)
    if world > 1 and not dist.is_initialized():
        if os.environ.get("BENCH_CONTROL", "gloo") == "nccl":
            dist.init_process_group("nccl", device_id=torch.device("cuda", local), pg_options=_nccl_options(args))
        else:
            dist.init_process_group("gloo")
    return local


def _nccl_options(args):
    import torch.distributed as dist
    if args.nccl_max_ctas <= 0:
        return None
    opts = dist.ProcessGroupNCCL.Options()
    opts.config.max_ctas = args.nccl_max_ctas
    opts.config.min_ctas = min(args.nccl_max_ctas, 1)
    return opts


def nccl_group(args):
    """The NCCL group for device-tensor collectives (None = the default group when that already is NCCL)."""
    global _NCCL_GROUP
    import torch.distributed as dist
    if dist.get_backend() == "nccl":
        return None
    if _NCCL_GROUP is None:
        _NCCL_GROUP = dist.new_group(backend="nccl", pg_options=_nccl_options(args))
    return _NCCL_GROUP


def reduce_max(values):
    """MAX over ranks of a list of floats through the default (control) group."""
    import torch
    import torch.distributed as dist
    t = torch.tensor(values, dtype=torch.float64, device="cuda" if dist.get_backend() == "nccl" else "cpu")
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return [float(x) for x in t.tolist()]


def csrc_sha16():
    """Identity of the kernel sources being timed (same function in tools/ncu_summary.py stamps profiles/traffic.json)."""
    import glob
    import hashlib
    hsh = hashlib.sha256()
    for f in sorted(glob.glob(os.path.join(ROOT, "vk-renderer_b200", "csrc", "*.cu*"))):
        hsh.update(os.path.basename(f).encode())
        hsh.update(open(f, "rb").read())
    return hsh.hexdigest()[:16]


def host_threads():
    """Host cores this process may use.  Set explicitly on the oracle: torchrun exports OMP_NUM_THREADS=1, which silently
    turned the CPU arm into a one-core run at N >= 2 in round 1."""
    try:
        return len(os.sched_getaffinity(0))
    except AttributeError:
        return os.cpu_count() or 1


def run_ours(args, rank, world):
    import torch
    import torch.distributed as dist
    local = init_dist(args, world)
    import vkr
    import vkr_host

    w, h, K, W = args.width, args.height, args.steps, args.warmup
    npix = w * h
    stream = torch.cuda.Stream()
    app = vkr_host.App(w, h, device=local, stream=stream)
    for opt in filter(None, os.environ.get("BENCH_APP_OPTIONS", "").split(",")):     # e.g. graph.concurrent_branches=0,graph.frame_graph=0 (experiments)
        name, val = opt.split("=")
        app.set_option(name, float(val))
    # Weak scaling = the SAME work per GPU at every N: every rank renders the same synthetic sequence (seed 0), independently.
    # Rounds 1 and early 2 seeded scene and camera path with the rank; seeds >= 2 hold costlier content (longer marches, wider
    # blurs: 1.31 vs 1.19 ms/frame), the line reports the max over ranks, and that alone was the "0.915 efficiency at N >= 4"
    # (profiles/r02l .. r02r: not the box, not NVML, not NCCL, not host enqueue, not launch gaps, not concurrency).
    seq_seed = int(os.environ.get("BENCH_SEQ_SEED", "0"))
    scene = vkr.synth_scene(seq_seed, 64)
    nres = min(64, K + W)                                 # distinct frames of the sequence kept resident (wraps after 64)
    seq_params = [vkr.synth_camera(seq_seed, f, w, h, True) for f in range(nres)]
    jit = [float(((f * 2654435761) % 1000) / 1000.0 - 0.5) for f in range(nres)]   # stands in for rand()/RAND_MAX - 0.5
    resident = []
    with torch.cuda.stream(stream):
        for f in range(nres):
            gb = vkr.Gbuffer(w, h, device=f"cuda:{local}")
            vkr.synth_gbuffer(scene, seq_params[f], gb, stream=stream)
            resident.append(gb)
    stream.synchronize()

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()

    def max_over_ranks(x):
        if world == 1:
            return x
        return reduce_max([x])[0]

    # ---- device-resident throughput: value ------------------------------------------------------------------------
    def step_resident(i):
        f = i % nres
        app.bind_gbuffer(resident[f])
        app.render_frame(seq_params[f], jit[f])

    for i in range(W):
        step_resident(i)
    barrier()
    if os.environ.get("BENCH_STAGGER_S"):                   # experiment: ranks run their timed regions one after the other
        time.sleep(float(os.environ["BENCH_STAGGER_S"]) * rank)
    sampler = ClockSampler(local)
    sampler.start()
    # Timed region: EXACTLY K frames the way the host API runs them — the rendergraph issues the frame's independent
    # branches (SSR filter + blur | GTAO main, filter, accumulate) on two streams, so tasks overlap.
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(stream)
    t_host0 = time.perf_counter()
    for i in range(W, W + K):
        step_resident(i)
    host_ms = (time.perf_counter() - t_host0) * 1e3 / K       # CPU time to ENQUEUE one frame (not a GPU time): >= the GPU frame time means host-bound
    e1.record(stream)
    stream.synchronize()
    barrier()
    ms_total = max_over_ranks(e0.elapsed_time(e1))
    host_ms = max_over_ranks(host_ms)
    lanes = dict(zip(app.last_tasks(), app.last_lanes()))
    # Per-task durations need tasks that do not overlap: K more frames of the same sequence with task timing on, which puts
    # every task on the one stream with an event pair around it (these frames are not part of `value`).
    app.task_timing(True)
    s0, s1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    s0.record(stream)
    for i in range(W + K, W + 2 * K):
        step_resident(i)
    s1.record(stream)
    stream.synchronize()
    sampler.stop_flag = True
    barrier()
    ms_serial = max_over_ranks(s0.elapsed_time(s1)) / K
    tasks = app.task_times()
    app.task_timing(False)
    sampler.join(timeout=1.0)
    ms_step = ms_total / K
    value = world * npix * K / (ms_total * 1e-3) / 1e9

    # ---- end to end through the host API with host buffers: e2e -----------------------------------------------------
    e2e = None
    if not args.no_e2e:
        nh = min(nres, 8 if K + W > 8 else K + W)          # pinned host frames (the H2D cost does not depend on content)
        host_frames = [vkr.Gbuffer(w, h, device="cpu", pinned=True) for f in range(nh)]
        for f in range(nh):
            for d, s in zip(host_frames[f].input_images(), resident[f].input_images()):
                d.buf.copy_(s.buf)
        out_desc = app.image_desc("taa")
        out_host = vkr.Image(out_desc.width, out_desc.height, out_desc.format, 1, device="cpu", pinned=True)
        h2d = sum(int(i.desc.row_pitch[0]) * i.height for i in host_frames[0].input_images())
        d2h = out_host.nbytes

        # Pipelined through the public host API: three device G-buffers rotate (being uploaded / current / last frame's depth
        # still serving as prev_depth), uploads run on a copy stream, read-backs on another, so the PCIe directions and the
        # chain overlap; events carry the dependencies, the host never blocks inside the timed region.
        import vkr_tiles
        dev_gb = [vkr.Gbuffer(w, h, device=f"cuda:{local}") for _ in range(3)]
        cs, ds = torch.cuda.Stream(), torch.cuda.Stream()
        ev = lambda: torch.cuda.Event()  # noqa: E731
        render_done, rb_done = {}, {}
        taa_alias = {}

        def taa_tensor():
            d = app.image_desc("taa")
            if d.base not in taa_alias:
                taa_alias[d.base] = torch.as_tensor(vkr_tiles._CudaMem(d.base, out_host.nbytes), device=f"cuda:{local}")
            return taa_alias[d.base]

        def step_e2e(i):
            f, g = i % nh, dev_gb[i % 3]
            if i - 2 in render_done:
                cs.wait_event(render_done[i - 2])          # g was current at i-3 and prev_depth at i-2
            with torch.cuda.stream(cs):
                for d, s_ in zip(g.input_images(), host_frames[f].input_images()):
                    d.buf.copy_(s_.buf, non_blocking=True)
                up = ev(); up.record(cs)
            stream.wait_event(up)
            if i - 2 in rb_done:
                stream.wait_event(rb_done[i - 2])           # the TAA target of frame i is the image frame i-2 wrote
            app.bind_gbuffer(g)
            app.render_frame(seq_params[f], jit[f])
            render_done[i] = ev(); render_done[i].record(stream)
            src = taa_tensor()
            ds.wait_event(render_done[i])
            with torch.cuda.stream(ds):
                out_host.buf.copy_(src, non_blocking=True)
                rb_done[i] = ev(); rb_done[i].record(ds)
            render_done.pop(i - 3, None); rb_done.pop(i - 3, None)

        for i in range(W):
            step_e2e(i)
        barrier()
        a0, a1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a0.record(stream)
        for i in range(W, W + K):
            step_e2e(i)
        stream.wait_stream(ds)
        stream.wait_stream(cs)
        a1.record(stream)
        stream.synchronize()
        barrier()
        ms_e2e = max_over_ranks(a0.elapsed_time(a1))
        # The bound of this path on THIS box: the same bytes through the same two copy streams, all ranks at once, nothing rendered.
        src_fixed = taa_tensor()
        barrier()
        c0, c1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        c0.record(stream)
        cs.wait_event(c0); ds.wait_event(c0)
        for i in range(K):
            with torch.cuda.stream(cs):
                for d, s_ in zip(dev_gb[i % 3].input_images(), host_frames[i % nh].input_images()):
                    d.buf.copy_(s_.buf, non_blocking=True)
            with torch.cuda.stream(ds):
                out_host.buf.copy_(src_fixed, non_blocking=True)
        stream.wait_stream(cs); stream.wait_stream(ds)
        c1.record(stream)
        stream.synchronize()
        barrier()
        ms_copy = max_over_ranks(c0.elapsed_time(c1)) / K
        e2e = {"value": world * npix * K / (ms_e2e * 1e-3) / 1e9, "unit": "Gpix/s", "ms_per_step": ms_e2e / K, "h2d_bytes_per_step": h2d,
               "d2h_bytes_per_step": d2h,
               "roofline": {"bound": "host<->device copies", "copy_only_ms_per_step": ms_copy, "copy_only_GBps": world * (h2d + d2h) / (ms_copy * 1e-3) / 1e9,
                            "frac": ms_copy / (ms_e2e / K),
                            "how": "the step's H2D + D2H bytes alone, on the same pinned buffers and copy streams, every rank at once (measured in this run); "
                                   "frac >= 1 means the end-to-end step is at this bound (the two loops differ by the noise of a shared host fabric)"},
               "note": "pinned host G-buffer -> device (copy stream) -> vkr_app_bind_gbuffer -> render_frame -> TAA target -> pinned host "
                                                  "(read-back stream); three device G-buffers rotate so both PCIe directions overlap the chain"}

    fg_stats = app.frame_graph_stats()
    app.close()
    del resident, app
    torch.cuda.empty_cache()
    if rank != 0:
        return None
    peak, peak_src = peaks()
    dom = max(tasks.items(), key=lambda kv: kv[1][0])
    dname, (dms, dcount) = dom
    avg_ms = dms / max(dcount, 1)
    achieved = ALG_BYTES[dname] * npix / (avg_ms * 1e-3) / 1e9
    traffic = None
    traffic_stale, traffic_captured = None, None
    tpath = os.path.join(ROOT, "profiles", "traffic.json")     # dram bytes per launch from the committed ncu --set full capture
    if os.path.exists(tpath):
        tj = json.load(open(tpath))
        traffic = tj.get(f"{w}x{h}", {}).get(dname)
        traffic_captured = tj.get("_captured_at")               # {"git_head", "csrc_sha16"} written by tools/ncu_summary.py
        traffic_stale = (traffic_captured or {}).get("csrc_sha16") != csrc_sha16()   # the profiled kernels are not the timed ones
    chain_alg = 128.0 * npix
    # why the HBM fraction is low: the chain's executed warp instructions (same ncu capture) against the issue slots the
    # frame time offers (SMs x 4 schedulers x the SM clock sampled during the timed region)
    issue = None
    clocks = sampler.result()
    if os.path.exists(tpath) and clocks.get("sm_mhz"):
        winst = json.load(open(tpath)).get(f"{w}x{h}:warp_inst")
        if winst:
            import torch
            sms = torch.cuda.get_device_properties(0).multi_processor_count
            total = float(sum(winst.values()))
            issue = {"warp_inst_per_frame": total, "slots_per_frame": ms_step * 1e-3 * sms * 4 * clocks["sm_mhz"] * 1e6,
                     "frac": total / (ms_step * 1e-3 * sms * 4 * clocks["sm_mhz"] * 1e6), "source": "profiles/traffic.json (smsp__inst_executed.sum)"}
    line = {"metric": "chain_throughput", "value": value, "unit": "Gpix/s", "n_gpus": world, "steps": K, "warmup": W, "ms_per_step": ms_step,
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": {"workload": f"full chain (Hi-Z+SSR+GTAO+lighting+TAA), {w}x{h} TAA sequence with synthetic velocity (BASELINE configs[2])",
                       "frames_resident": nres, "l2": "inputs larger than L2 (one frame touches > 1 GB at 4K)",
                       "parallelism": f"{world} independent replica(s) of the same sequence (seed {seq_seed}), one per GPU: fixed work per GPU; nothing is "
                                      f"exchanged; barrier and max over ranks on {os.environ.get('BENCH_CONTROL', 'gloo')}"},
            "roofline": {"bound": "hbm", "kernel": dname, "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                         "traffic": traffic, "traffic_stale": traffic_stale, "traffic_captured_at": traffic_captured,
                         "peak_source": peak_src, "avg_launch_ms": avg_ms,
                         "chain": {"alg_bytes_per_frame": chain_alg, "achieved": chain_alg / (ms_step * 1e-3) / 1e9,
                                   "frac": chain_alg / (ms_step * 1e-3) / 1e9 / peak, "issue_slots": issue},
                         "per_task_ms": {k: v[0] / max(v[1], 1) for k, v in tasks.items()},
                         "per_task_source": f"{K} further frames with task timing on: one stream, an event pair around every task, "
                                            f"{ms_serial:.4f} ms/frame; in the timed region the rendergraph overlaps tasks on two streams",
                         "task_lanes": lanes},
            "gpu_launches": KERNELS_PER_FRAME * K, "clocks": clocks,
            "host_enqueue_ms_per_frame": host_ms, "frame_graph": fg_stats}
    if e2e:
        line["e2e"] = e2e
    if world == 1 and not args.no_cpu_baseline:
        gpix, ms, threads, band_rows = cpu_chain_rate(w, h, 3, 1, threads=host_threads())
        line["cpu_baseline"] = {"value": gpix, "unit": "Gpix/s", "cores": threads, "kind": "port", "ms_per_sample_frame": ms,
                                "sample": f"rows [{int(BAND[0]*h)}, {int(BAND[1]*h)}) of 3 frames of the same {w}x{h} sequence through every pass after 1 warm-up "
                                          f"({band_rows} of {h} rows; Hi-Z in full)"}
    return line


def run_probes(args, rank, world, emit=True):
    """BASELINE configs[3]: args.views independent 256x256 views (Hi-Z + SSR trace/filter/blur + GTAO main/filter/accumulate),
    contiguous blocks of views per rank, every pass ONE launch per rank (layers on grid.z), outputs all-gathered per frame on a
    side stream (the only collective; it overlaps the next frame).  Strong scaling: the batch is fixed, ranks split it."""
    import torch
    import torch.distributed as dist
    local = init_dist(args, world)
    import vkr
    import vkr_views
    w = h = 256
    K, W = args.steps, args.warmup
    b, e = vkr_views.shard_range(args.views, world, rank)
    stream, side = torch.cuda.Stream(), torch.cuda.Stream()
    lane2 = torch.cuda.Stream() if args.probes_lanes == 2 else None
    gstreams = [(torch.cuda.Stream(), torch.cuda.Stream() if args.probes_lanes == 2 else None) for _ in range(max(0, args.probes_groups - 1))]
    assert e - b > 1, "the batched path needs at least 2 views per rank"
    nres = 4                                                # resident input frames (cycled); > L2 in total

    def allgather_bytes(x):
        out = [None] * world
        dist.all_gather_object(out, x)
        return out

    fg = None
    if world > 1 and args.probes_gather == "fused":         # the gather fused into the blur / accumulate kernels (peer stores)
        import vkr_gather
        fg = vkr_gather.FusedGather(args.views, {"accumulated_ao": (w // 2, h // 2, vkr.R16G16_SFLOAT), "blurred": (w // 2, h // 2, vkr.R8G8B8A8_UNORM)},
                                    rank, world, allgather_bytes, f"cuda:{local}")
    with torch.cuda.stream(stream):
        vb = vkr_views.ViewBatch(w, h, list(range(b, e)), device=f"cuda:{local}", stream=stream, param_sets=nres, fused=fg)
        frames = []
        for f in range(nres):
            vb.frame = f
            vb.synth_inputs()
            frames.append((vb.albedo, vb.normal, vb.material, vb.velocity, vb.depth))
            A = lambda ww, hh, fmt, m=1: vkr.ImageArray(ww, hh, fmt, m, vb.V, device=f"cuda:{local}")  # noqa: E731
            vb.albedo, vb.normal, vb.material = A(w, h, vkr.R8G8B8A8_SRGB), A(w, h, vkr.R16G16_UNORM), A(w, h, vkr.R8G8B8A8_SRGB)
            vb.velocity, vb.depth = A(w, h, vkr.R16G16_SFLOAT), A(w, h, vkr.D24_UNORM_S8_UINT, vkr.full_mip_count(w, h))
        params = []
        for f in range(nres):
            vb.frame, vb.counter = f, f % 16
            params.append(vb.params())
            vb.albedo, vb.normal, vb.material, vb.velocity, vb.depth = frames[f]
            vb.passes(params[f], flags=vkr.EXT_PREPARE_ONLY, param_set=f)       # per-view parameter blocks uploaded once
        vb.frame = vb.counter = 0
    stream.synchronize()
    done = torch.cuda.Event()
    graphs, outs = {}, {}
    ao0, refl0 = vb.outputs()
    eq = len(set(vkr_views.shard_sizes(args.views, world))) == 1
    gout = [torch.empty((args.views, t.shape[1]), dtype=torch.uint8, device=t.device) if (world > 1 and eq) else None for t in (ao0, refl0)]

    def bind(i):
        f = i % nres
        vb.albedo, vb.normal, vb.material, vb.velocity, vb.depth = frames[f]
        if i > 0:
            vb.prev_depth = frames[(f - 1) % nres][4]        # last frame's depth with its pyramid (main.cpp:416)
        return f

    def swap():
        vb.blurred, vb.blurred_history = vb.blurred_history, vb.blurred
        vb.accumulated_ao, vb.accumulated_history = vb.accumulated_history, vb.accumulated_ao

    gather_group = nccl_group(args) if (world > 1 and args.probes_gather == "allgather") else None
    pg = None
    if world > 1 and args.probes_gather == "p2p":
        import vkr_gather
        pg = vkr_gather.PeerGather(args.views, [ao0.shape[1], refl0.shape[1]], rank, world, allgather_bytes, f"cuda:{local}", buffers=args.probes_gather_buffers)

    def gather(ao, refl):
        if pg is not None:
            pg.push([ao, refl], stream)
            return
        if world > 1 and args.probes_gather == "allgather":  # final gather of the per-view results, off the compute stream
            done.record(stream)
            side.wait_event(done)
            with torch.cuda.stream(side):
                vkr_views.gather_blocks(ao, args.views, group=gather_group, out=gout[0])
                vkr_views.gather_blocks(refl, args.views, group=gather_group, out=gout[1])

    def step(i):
        """Steps 0..nres-1 run eagerly, steps nres..2*nres-1 are captured into one CUDA graph each (the image bindings repeat
        with period nres, nres even), later steps replay: 9 kernel launches per frame with no host work between them."""
        f = i % nres
        if pg is not None:
            pg.sources_free(stream)                          # this frame overwrites the result images of two frames ago
        if i >= 2 * nres:
            graphs[f].replay()
            if fg is not None:
                fg.note_replayed()
            gather(*outs[f])
            return
        bind(i)
        if i >= nres:
            g = torch.cuda.CUDAGraph()
            with torch.cuda.graph(g, stream=stream):
                vb.passes(params[f], flags=vkr.EXT_PARAMS_PREPARED, param_set=f, aux_stream=lane2, group_streams=gstreams)
            graphs[f], outs[f] = g, vb.outputs()
            g.replay()
        else:
            vb.passes(params[f], flags=vkr.EXT_PARAMS_PREPARED, param_set=f, aux_stream=lane2, group_streams=gstreams)
        gather(*vb.outputs())
        swap()

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()

    W = max(W, 2 * nres)                                     # eager + capture steps are warm-up
    with torch.cuda.stream(stream):
        for i in range(W):
            step(i)
        barrier()
        sampler = ClockSampler(local)
        sampler.start()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(stream)
        t_host0 = time.perf_counter()
        for i in range(W, W + K):
            step(i)
        host_ms = (time.perf_counter() - t_host0) * 1e3 / K      # CPU time to ENQUEUE one frame (not a GPU time)
        if pg is not None:
            pg.finish()                                      # the last epoch has landed here too
            stream.wait_stream(pg.xfer)
        if fg is not None:
            fg.gathered(fg.epoch, stream)                    # every rank's last frame has landed here too
        stream.wait_stream(side)
        e1.record(stream)
        stream.synchronize()
        sampler.stop_flag = True
        barrier()
    ms = e0.elapsed_time(e1)
    per_rank_ms = [ms / K]
    if world > 1:
        per_rank_ms = [None] * world
        dist.all_gather_object(per_rank_ms, ms / K)          # each rank's own frame time: shows shard imbalance, hidden in the max
        ms = reduce_max([ms])[0]
    sampler.join(timeout=1.0)
    gather_timeouts = 0
    if pg is not None:
        gather_timeouts = pg.wait_failures()
        pg.close(dist.barrier)
    if fg is not None:
        gather_timeouts = fg.wait_failures()
        vb.fused = None
        fg.close(dist.barrier)
    del graphs, outs, frames, vb, gout
    torch.cuda.empty_cache()
    if rank != 0:
        return None
    peak, peak_src = peaks()
    npix = args.views * w * h
    alg = 73.0 * npix                                        # Hi-Z + SSR + GTAO subset of the chain, SURVEY §8d
    line = {"metric": "probe_batch_throughput", "value": npix * K / (ms * 1e-3) / 1e9, "unit": "Gpix/s", "views_per_s": args.views * K / (ms * 1e-3),
            "n_gpus": world, "steps": K, "warmup": W, "ms_per_step": ms / K, "higher_is_better": True, "scaling": "strong", "vs_baseline": None,
            "dtype": "f32", "data": "synthetic",
            "config": {"workload": f"{args.views} probe views 256x256, Hi-Z+GTAO+SSR, contiguous view blocks per rank (BASELINE configs[3])",
                       "frames_resident": nres, "launch": f"one CUDA graph per frame: {args.probes_groups} view range(s) x 9 kernels" + (" (SSR filter+blur beside the GTAO passes)" if args.probes_lanes == 2 else "") + ", parameter blocks prepared once", "l2": f"inputs larger than L2 ({args.views // world} views x 1.4 MB per frame per rank)",
                       "parallelism": f"views sharded over {world} rank(s); " + {"allgather": "all_gather of AO + reflections per frame on a side stream", "p2p": "AO + reflections pushed into every rank's cudaIpc slab per frame by the copy engines, stream-side flags (no collective, no SMs)", "fused": "the blur and the accumulate kernels store every result texel into every rank's arrays themselves (peer-mapped memory over NVLink; epoch flags, no collective, no copy, no SMs beyond the passes')", "none": "NO gather (diagnostic run)"}[args.probes_gather] + (f", NCCL capped at {args.nccl_max_ctas} CTAs" if args.nccl_max_ctas > 0 else "")},
            "roofline": {"bound": "hbm", "kernel": "whole step (8 batched launches)", "achieved": alg / world / (ms / K * 1e-3) / 1e9, "peak": peak, "unit": "GB/s",
                         "frac": alg / world / (ms / K * 1e-3) / 1e9 / peak, "traffic": None, "peak_source": peak_src, "alg_bytes_per_step": alg},
            "gpu_launches": vkr_views.ViewBatch.KERNELS_PER_FRAME * max(1, args.probes_groups) * K, "gather_wait_timeouts": gather_timeouts, "per_rank_ms_per_step": [round(x, 4) for x in per_rank_ms],
            "clocks": sampler.result()}
    if emit:
        print(json.dumps(line))
    return line


def run_tiles(args, rank, world, emit=True):
    """BASELINE configs[4]: one 7680x4320 sequence, full chain, rows split over the ranks (vkr_tiles.RowTiledFrame): inputs and
    Hi-Z replicated, six halo exchanges per frame as direct peer stores + stream-side flags.  The finished TAA rows stay on the rank
    that computed them (--tiles-gather none, the default) or are sent to rank 0 every frame (rank0: NCCL send/recv on a side
    stream; 265 MB per frame into one GPU, which then bounds the frame rate).  Strong scaling."""
    import torch
    import torch.distributed as dist
    local = init_dist(args, world)
    import vkr
    import vkr_tiles
    w, h = (7680, 4320) if args.width == 3840 else (args.width, args.height)
    K, W = args.steps, args.warmup
    stream, side = torch.cuda.Stream(), torch.cuda.Stream()

    def allgather_bytes(b):
        if world == 1:
            return [b]
        out = [None] * world
        dist.all_gather_object(out, b)
        return out

    nres = 3
    jit = [float(((f * 2654435761) % 1000) / 1000.0 - 0.5) for f in range(nres)]
    with torch.cuda.stream(stream):
        scene = vkr.synth_scene(0, 64)
        dps = [vkr.synth_camera(0, f, w, h, True) for f in range(nres)]
        G = []
        for f in range(nres):
            gb = vkr.Gbuffer(w, h, device=f"cuda:{local}")
            vkr.synth_gbuffer(scene, dps[f], gb, stream=stream)
            G.append(gb)
        D = [g.depth for g in G]
        # Cost-balanced tiles: march length and blur radius vary down the frame, so equal rows are not equal work.  Before the
        # timed region every rank times its own tile WITHOUT exchanges, the times are all-gathered and the strip boundaries moved
        # (vkr_tiles.balance_bounds); three rounds.  This is set-up, like choosing a launch configuration: not timed.
        bounds = vkr_tiles.even_bounds(h // 2, world)
        history, cal_ms = [list(bounds)], []
        for _ in range(3 if world > 1 and args.tiles_balance else 0):
            cal = vkr_tiles.RowTiledFrame(w, h, rank, world, allgather_bytes, device=f"cuda:{local}", motion_rows=16, stream=stream, bounds=bounds)
            cal.exchange_enabled = False
            ev = [torch.cuda.Event(enable_timing=True) for _ in range(2)]
            for i in range(2 + 4):
                if i == 2:
                    ev[0].record(stream)
                cal.gb = G[i % nres]
                cal.gb.prev_depth = D[(i - 1) % nres]
                cal.render_frame(dps[i % nres], jit[i % nres], swap_depth=False)
            ev[1].record(stream)
            stream.synchronize()
            t = [None] * world
            dist.all_gather_object(t, ev[0].elapsed_time(ev[1]) / 4)
            cal_ms.append([round(x, 3) for x in t])
            cal.close(dist.barrier)
            bounds = vkr_tiles.balance_bounds(bounds, t)
            history.append(list(bounds))
        tf = vkr_tiles.RowTiledFrame(w, h, rank, world, allgather_bytes, device=f"cuda:{local}", motion_rows=16, stream=stream, bounds=bounds)
        cleared = tf.gb.prev_depth
    stream.synchronize()
    y0, y1 = tf.own_rows(True)
    spans = [None] * world
    if world > 1:
        dist.all_gather_object(spans, (y0, y1))
    done = torch.cuda.Event()

    rows_group = nccl_group(args) if (world > 1 and args.tiles_gather == "rank0") else None

    def rows_of(img, r0, r1):
        pitch = img.desc.row_pitch[0]
        return img.buf[r0 * pitch:r1 * pitch]

    def step(i):
        f = i % nres
        gb = G[f]
        gb.prev_depth = D[(f - 1) % nres] if i > 0 else cleared
        tf.gb = gb
        out = tf.render_frame(dps[f], jit[f], swap_depth=False)
        if world > 1 and args.tiles_gather == "rank0":       # final gather: finished rows to rank 0, off the compute stream
            done.record(stream)
            side.wait_event(done)
            with torch.cuda.stream(side):
                if rank == 0:
                    for r in range(1, world):
                        dist.recv(rows_of(out, *spans[r]), src=r, group=rows_group)
                else:
                    dist.send(rows_of(out, y0, y1), dst=0, group=rows_group)

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()

    with torch.cuda.stream(stream):
        for i in range(W):
            step(i)
        barrier()
        sampler = ClockSampler(local)
        sampler.start()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(stream)
        t_host0 = time.perf_counter()
        for i in range(W, W + K):
            step(i)
        host_ms = (time.perf_counter() - t_host0) * 1e3 / K      # CPU time to ENQUEUE one frame (not a GPU time)
        tf.finish()                                          # the last frame's TAA halo (waited for inside the next frame otherwise)
        stream.wait_stream(side)
        e1.record(stream)
        stream.synchronize()
        sampler.stop_flag = True
        barrier()
    ms = e0.elapsed_time(e1)
    fails = tf.wait_failures()
    per_rank_ms = [ms / K]
    if world > 1:
        per_rank_ms = [None] * world
        dist.all_gather_object(per_rank_ms, ms / K)
        ms, fails = reduce_max([ms, float(fails)])
        fails = int(fails)
    sampler.join(timeout=1.0)
    tf.close(dist.barrier if world > 1 else None)
    halo = tf.halo
    del tf, G, D
    torch.cuda.empty_cache()
    if rank != 0:
        return None
    peak, peak_src = peaks()
    npix = w * h
    alg = 128.0 * npix
    line = {"metric": "tiled_frame_throughput", "value": npix * K / (ms * 1e-3) / 1e9, "unit": "Gpix/s", "n_gpus": world, "steps": K, "warmup": W,
            "ms_per_step": ms / K, "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": {"workload": f"full chain on ONE {w}x{h} sequence, row tiles over {world} rank(s) (BASELINE configs[4])", "frames_resident": nres,
                       "l2": "inputs larger than L2 (one 8K frame touches > 4 GB)",
                       "parallelism": "rows split in 16-row strips; inputs + Hi-Z replicated; 6 halo exchanges per frame as peer stores over NVLink with "
                                      "stream-side flags, pushed on a transfer stream and waited for one independent pass later; " + ("finished TAA rows sent to rank 0 every frame (NCCL send/recv, side stream)"
                                                               if args.tiles_gather == "rank0" else
                                                               "finished TAA rows stay resident on the rank that computed them (SURVEY 8e: shard-resident output)"),
                       "halo_rows": halo, "wait_timeouts": fails, "host_enqueue_ms_per_frame_rank0": round(host_ms, 3),
                       "strip_bounds": {"final": bounds, "even": history[0], "tile_ms_without_exchange_per_round": cal_ms, "balanced": bool(world > 1 and args.tiles_balance),
                                        "note": "16-row half-res strips; boundaries moved by 3 untimed calibration rounds so every rank's tile costs the same"}},
            "roofline": {"bound": "hbm", "kernel": "whole step", "achieved": alg / world / (ms / K * 1e-3) / 1e9, "peak": peak, "unit": "GB/s",
                         "frac": alg / world / (ms / K * 1e-3) / 1e9 / peak, "traffic": None, "peak_source": peak_src, "alg_bytes_per_step": alg},
            "gpu_launches": vkr_tiles.RowTiledFrame.KERNELS_PER_FRAME * K, "per_rank_ms_per_step": [round(x, 4) for x in per_rank_ms],
            "clocks": sampler.result()}
    if emit:
        print(json.dumps(line))
    return line


SUB_STEPS, SUB_WARMUP = 24, 8


def sub_records(args, rank, world):
    """The two workloads north_star asks to SHARD (BASELINE configs[3] probe batch, configs[4] row-tiled 8K frame), measured in
    the same job as the headline line so that the driver's BENCH / SCALE records carry them from one build: at N ranks the
    sharded run, and -- in the same process on rank 0 alone, the other ranks waiting at a barrier -- the same workload on ONE
    GPU, so `speedup_vs_n1` needs no second job.  Bounded: SUB_STEPS timed steps after SUB_WARMUP each."""
    import argparse
    import torch
    import torch.distributed as dist
    out = {}
    sub = argparse.Namespace(**vars(args))
    sub.steps, sub.warmup = SUB_STEPS, SUB_WARMUP
    for name, fn in (("tiles", run_tiles), ("probes", run_probes)):     # tiles first: it needs no NCCL group at all
        solo = None
        if world > 1:
            if rank == 0:
                solo = fn(sub, 0, 1, emit=False)
            torch.cuda.synchronize()
            dist.barrier()
        full = fn(sub, rank, world, emit=False)
        if rank == 0:
            solo = solo or full
            keep = ("metric", "value", "unit", "ms_per_step", "n_gpus", "steps", "warmup", "scaling", "gpu_launches")
            rec = {k: full[k] for k in keep}
            rec["workload"] = full["config"]["workload"]
            rec["parallelism"] = full["config"]["parallelism"]
            rec["roofline_frac"] = full["roofline"]["frac"]
            rec["n1_same_run"] = {"value": solo["value"], "ms_per_step": solo["ms_per_step"]}
            rec["speedup_vs_n1"] = full["value"] / solo["value"]
            for k in ("gather_wait_timeouts", "per_rank_ms_per_step"):
                if k in full:
                    rec[k] = full[k]
            if "wait_timeouts" in full["config"]:
                rec["wait_timeouts"] = full["config"]["wait_timeouts"]
            rec["clocks"] = full["clocks"]
            out[name] = rec
    return out


def main():
    args = parse()
    rank, world = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1))
    if args.impl == "reference":
        run_reference(args, rank, world)
    elif args.workload in ("probes", "tiles"):
        (run_probes if args.workload == "probes" else run_tiles)(args, rank, world)
        if world > 1:
            import torch.distributed as dist
            dist.destroy_process_group()
    else:
        line = run_ours(args, rank, world)
        subs = sub_records(args, rank, world) if not args.no_sub else {}
        if rank == 0:
            line.update(subs)
            print(json.dumps(line))
        if world > 1:
            import torch.distributed as dist
            dist.destroy_process_group()


if __name__ == "__main__":
    main()
