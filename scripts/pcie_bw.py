"""Host<->device copy bandwidth on the box: H2D alone, D2H alone, both directions at once (pinned memory, 48 MiB chunks,
the size of one C2 step's inputs / results).  Bounds the e2e leg of bench.py."""
import torch
dev = torch.device("cuda")
n = 48 << 20
h_in = torch.empty(n, dtype=torch.uint8).pin_memory()
h_out = torch.empty(n, dtype=torch.uint8).pin_memory()
d_in = torch.empty(n, dtype=torch.uint8, device=dev)
d_out = torch.empty(n, dtype=torch.uint8, device=dev)
s1, s2 = torch.cuda.Stream(), torch.cuda.Stream()


def timed(fn, reps=20):
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    s1.wait_event(a); s2.wait_event(a)
    for _ in range(reps):
        fn()
    e1, e2 = torch.cuda.Event(), torch.cuda.Event()
    e1.record(s1); e2.record(s2)
    torch.cuda.current_stream().wait_event(e1); torch.cuda.current_stream().wait_event(e2)
    b.record()
    torch.cuda.synchronize()
    return a.elapsed_time(b) / reps


def h2d():
    with torch.cuda.stream(s1):
        d_in.copy_(h_in, non_blocking=True)


def d2h():
    with torch.cuda.stream(s2):
        h_out.copy_(d_out, non_blocking=True)


def both():
    h2d(); d2h()


for name, fn in (("h2d", h2d), ("d2h", d2h), ("both", both)):
    ms = timed(fn)
    print(f"{name}: {ms:.3f} ms per 48 MiB -> {n / ms / 1e6:.1f} GB/s per direction")

# the same with device work running next to the copies (a bandwidth-bound fill and a compute-bound matmul)
s3 = torch.cuda.Stream()
big = torch.empty(1 << 30, dtype=torch.uint8, device=dev)
a = torch.randn(4096, 4096, device=dev)
for label, work in (("fill 1 GiB (HBM-bound)", lambda: big.zero_()), ("fp32 matmul 4096^3", lambda: a @ a)):
    def both_busy():
        with torch.cuda.stream(s3):
            work()
        both()
    ms = timed(both_busy)
    torch.cuda.synchronize()
    with torch.cuda.stream(s3):
        t0, t1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        t0.record(); work(); t1.record()
    torch.cuda.synchronize()
    print(f"both + {label}: {ms:.3f} ms per iteration (work alone {t0.elapsed_time(t1):.3f} ms)")

# one step's real copy list: 10 H2D tensors / 6 D2H tensors of the C2 sizes
G = 131072
sizes_in = [G * 3 * 4, G * 9 * 4, G * 75 * 4, G * 4, 64 * 4, 36 * 4, 16, 16, 4 * 3 * 65536 * 4, 4 * 65536 * 4]
sizes_out = [4 * 3 * 65536 * 4, 4 * 65536 * 4, G * 3 * 4, G * 9 * 4, G * 75 * 4, G * 4]
hin = [torch.empty(n, dtype=torch.uint8).pin_memory() for n in sizes_in]
din = [torch.empty(n, dtype=torch.uint8, device=dev) for n in sizes_in]
hout = [torch.empty(n, dtype=torch.uint8).pin_memory() for n in sizes_out]
dout = [torch.empty(n, dtype=torch.uint8, device=dev) for n in sizes_out]


def lists():
    with torch.cuda.stream(s1):
        for d, h in zip(din, hin):
            d.copy_(h, non_blocking=True)
    with torch.cuda.stream(s2):
        for d, h in zip(dout, hout):
            h.copy_(d, non_blocking=True)


print(f"both, as the step's 10 + 6 tensors: {timed(lists):.3f} ms per step")
