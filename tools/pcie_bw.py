"""Raw pinned-host <-> device copy bandwidth of the box (the ceiling of bench.py's e2e number): python tools/pcie_bw.py [MB]"""
import sys
import torch

mb = int(sys.argv[1]) if len(sys.argv) > 1 else 166
n = mb * 1000 * 1000
h = torch.empty(n, dtype=torch.uint8).pin_memory()
d = torch.empty(n, dtype=torch.uint8, device="cuda")
h2 = torch.empty(n // 2, dtype=torch.uint8).pin_memory()
d2 = torch.empty(n // 2, dtype=torch.uint8, device="cuda")
s1, s2 = torch.cuda.Stream(), torch.cuda.Stream()


def timed(fn, reps=10):
    fn(); torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(reps):
        fn()
    for s in (s1, s2):
        torch.cuda.current_stream().wait_stream(s)
    b.record(); torch.cuda.synchronize()
    return a.elapsed_time(b) / reps


def both():
    s1.wait_stream(torch.cuda.current_stream()); s2.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(s1):
        d.copy_(h, non_blocking=True)
    with torch.cuda.stream(s2):
        h2.copy_(d2, non_blocking=True)


t = timed(lambda: d.copy_(h, non_blocking=True)); print(f"H2D {mb} MB: {t:.3f} ms = {n / t / 1e6:.1f} GB/s")
t = timed(lambda: h.copy_(d, non_blocking=True)); print(f"D2H {mb} MB: {t:.3f} ms = {n / t / 1e6:.1f} GB/s")
t = timed(both); print(f"H2D {mb} MB with D2H {mb // 2} MB in flight: {t:.3f} ms = {n / t / 1e6:.1f} GB/s in, {n / 2 / t / 1e6:.1f} GB/s out")

# the e2e pattern: 8 pinned frames of 5 planes each, uploaded in turn into 3 rotating device buffers, read-back in flight
planes = [[torch.empty(n // 5, dtype=torch.uint8).pin_memory() for _ in range(5)] for _ in range(8)]
dev = [[torch.empty(n // 5, dtype=torch.uint8, device="cuda") for _ in range(5)] for _ in range(3)]
it = [0]


def frames(with_rb):
    def f():
        i = it[0]; it[0] += 1
        s1.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(s1):
            for a, b in zip(dev[i % 3], planes[i % 8]):
                a.copy_(b, non_blocking=True)
        if with_rb:
            s2.wait_stream(torch.cuda.current_stream())
            with torch.cuda.stream(s2):
                h2[: 66 * 1000 * 1000].copy_(d2[: 66 * 1000 * 1000], non_blocking=True)
    return f


t = timed(frames(False), 24); print(f"8 frames x 5 planes, H2D only: {t:.3f} ms/frame = {n / t / 1e6:.1f} GB/s")
t = timed(frames(True), 24); print(f"8 frames x 5 planes + 66 MB D2H in flight: {t:.3f} ms/frame = {n / t / 1e6:.1f} GB/s in")

s3 = torch.cuda.Stream()


def frames2():
    i = it[0]; it[0] += 1
    for s in (s1, s3):
        s.wait_stream(torch.cuda.current_stream())
    for j, (a, b) in enumerate(zip(dev[i % 3], planes[i % 8])):
        with torch.cuda.stream(s1 if j % 2 == 0 else s3):
            a.copy_(b, non_blocking=True)
    s2.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(s2):
        h2[: 66 * 1000 * 1000].copy_(d2[: 66 * 1000 * 1000], non_blocking=True)
    torch.cuda.current_stream().wait_stream(s3)


t = timed(frames2, 24); print(f"same, planes alternating over two upload streams: {t:.3f} ms/frame = {n / t / 1e6:.1f} GB/s in")
