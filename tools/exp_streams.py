"""Experiment: the SSR branch (filter, blur) and the GTAO branch (main, filter, accumulate) of a frame on two streams.
python tools/exp_streams.py W H frames"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [os.path.join(ROOT, "tests"), os.path.join(ROOT, "vk-renderer_b200")]
import torch  # noqa: E402

import chain  # noqa: E402
import vkr  # noqa: E402

w, h, frames = int(sys.argv[1]), int(sys.argv[2]), int(sys.argv[3])
scene = vkr.synth_scene(0, 64)
cc = chain.Chain(w, h, chain.CudaBackend())
for _ in range(3):
    P = cc.begin_frame(scene)
    cc.passes(P)
    cc.end_frame()
P = cc.begin_frame(scene)
st = [t for _, t, _ in cc.steps(P)]
main = torch.cuda.current_stream()
sb = torch.cuda.Stream()


def serial():
    for t in st:
        t()


def forked(order):
    st[0](); st[1]()
    e1 = torch.cuda.Event(); e1.record(main)
    sb.wait_event(e1)
    if order == "gtao_side":
        with torch.cuda.stream(sb):
            st[4](); st[5](); st[6]()
            e2 = torch.cuda.Event(); e2.record(sb)
        st[2](); st[3]()
    else:
        with torch.cuda.stream(sb):
            st[2](); st[3]()
            e2 = torch.cuda.Event(); e2.record(sb)
        st[4](); st[5](); st[6]()
    main.wait_event(e2)
    st[7](); st[8]()


def timed(fn, label):
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(frames):
        fn()
    b.record(); torch.cuda.synchronize()
    print(f"{label:28s} {a.elapsed_time(b) / frames:.4f} ms/frame")


timed(serial, "one stream")
timed(lambda: forked("gtao_side"), "GTAO branch on 2nd stream")
timed(lambda: forked("ssr_side"), "SSR branch on 2nd stream")
timed(serial, "one stream (again)")
