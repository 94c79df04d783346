"""bench.py — fwd+bwd views/sec of the render path at BASELINE config C2
(RE10K training render step: 131072 Gaussians, 4 target views, 256x256, colour+depth).

  python bench.py --gpus N --steps K --warmup W            # our arm (sm_100a library)
  python bench.py --impl reference --gpus N --steps K ...  # CPU arm: the oracle port on host cores

A "step" = one pass of the hot path over one batch: DecoderSplattingCUDA.forward(depth_mode="depth")
+ backward for 1 scene x 4 views per rank.  Ranks render independent scenes (scene sharding, no
collective) => weak scaling.  At N > 1 the line also carries a `view_sharded` block: ONE scene,
4 N target views split over the ranks, per-Gaussian gradients summed with the path's only collective.
One JSON line is printed by rank 0.
"""
from __future__ import annotations

import argparse
import json
import os
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

G_CTX, HW, VIEWS = 2, 256, 4          # C2: 2 context views of 256x256 -> 131072 Gaussians; 4 target views
N_SCENES = 4                           # distinct resident scenes cycled through (4 x 46 MB of inputs)
METRIC = "fwd+bwd views/sec (256x256, 131072 Gaussians/scene)"
UNIT = "views/s"
WORKLOAD = ("C2: RE10K training render step, 1 scene x 4 target views per GPU, 256x256, 131072 Gaussians (SH degree 4), "
            "colour+depth fwd+bwd")


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=50)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-view-sharded", action="store_true", help="skip the view-sharded block at N > 1")
    ap.add_argument("--c5", action="store_true", help="at N > 1 also time the C5 stress shape view-sharded (8 views per rank)")
    ap.add_argument("--eager", action="store_true", help="launch every step from Python instead of replaying CUDA graphs")
    return ap.parse_args()


def dist_env():
    return int(os.environ.get("RANK", 0)), int(os.environ.get("LOCAL_RANK", 0)), int(os.environ.get("WORLD_SIZE", 1))


def shared_config(G: int, pairs_per_view) -> dict:
    """The `config` object, IDENTICAL in both arms (same workload, same keys, same strings)."""
    return {"workload": WORKLOAD, "gaussians": int(G), "views_per_step_per_gpu": VIEWS, "image": [HW, HW],
            "sh_degree": 4, "depth_mode": "depth", "pairs_per_view": pairs_per_view,
            "l2": "GPU arm: 256 MiB buffer zeroed between timed steps (outside the event pairs), "
                  f"{N_SCENES} scenes cycled; CPU arm: one scene, host caches as they fall"}


def percentiles(ms: list) -> dict:
    s = sorted(ms)
    pick = lambda q: s[min(len(s) - 1, max(0, int(round(q * (len(s) - 1)))))]
    return {"median": round(pick(0.5), 4), "p10": round(pick(0.1), 4), "p90": round(pick(0.9), 4),
            "min": round(s[0], 4), "max": round(s[-1], 4)}


# ------------------------------------------------------------------------------------------------
# clocks: sampled DURING the timed region (NVML; nvidia-smi fallback)
# ------------------------------------------------------------------------------------------------
class ClockSampler:
    def __init__(self, index: int):
        self.index, self.samples, self.reasons, self.max_mhz = index, [], set(), None
        self._stop = threading.Event()
        self._t = None
        try:
            import pynvml
            pynvml.nvmlInit()
            self.nv = pynvml
            self.h = pynvml.nvmlDeviceGetHandleByIndex(index)
            self.max_mhz = pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM)
        except Exception:
            self.nv = None

    def _loop(self):
        nv = self.nv
        names = {0x8: "hw_slowdown", 0x40: "hw_thermal_slowdown", 0x20: "sw_thermal_slowdown", 0x4: "sw_power_cap",
                 0x80: "hw_power_brake_slowdown"}
        while not self._stop.is_set():
            try:
                self.samples.append(nv.nvmlDeviceGetClockInfo(self.h, nv.NVML_CLOCK_SM))
                try:
                    r = nv.nvmlDeviceGetCurrentClocksEventReasons(self.h)
                except Exception:
                    r = nv.nvmlDeviceGetCurrentClocksThrottleReasons(self.h)
                for bit, name in names.items():
                    if r & bit:
                        self.reasons.add(name)
            except Exception:
                pass
            self._stop.wait(0.05)  # NVML queries take the driver lock: polling every 5 ms slowed the launches 8x

    def start(self):
        if self.nv is not None:
            self._t = threading.Thread(target=self._loop, daemon=True)
            self._t.start()

    def stop(self) -> dict:
        self._stop.set()
        if self._t is not None:
            self._t.join()
        s = sorted(self.samples)
        return {"sm_mhz": (s[len(s) // 2] if s else None), "sm_max_mhz": self.max_mhz,
                "reasons": sorted(self.reasons), "samples": len(s)}


# ------------------------------------------------------------------------------------------------
# workload
# ------------------------------------------------------------------------------------------------
def make_inputs(rank: int, n_scenes: int):
    from transplat_b200.synthetic import make_scene, upstream_grads
    scenes = [make_scene(n_ctx=G_CTX, h=HW, w=HW, n_views=VIEWS, seed=1236 + 100 * rank + i) for i in range(n_scenes)]
    target, dLd = upstream_grads(VIEWS, HW, HW, seed=99 + rank)
    return scenes, target, dLd


def bytes_per_view(G: int, D_per_view: float, P: int, T: int) -> float:
    """B_view = 1232 G + 260 D + 48 P + 8 T (fwd+bwd, colour+depth fused, fp32)."""
    return 1232.0 * G + 260.0 * D_per_view + 48.0 * P + 8.0 * T


def timed_pass(run, steps, flush, barrier, world, dev):
    """K steps, one CUDA-event pair per step on the launching stream, L2 flushed between steps (outside the pairs).
    Returns (per-step ms of this rank, ms_per_step = max over ranks of the mean, host wall ms per step)."""
    import torch
    import torch.distributed as dist
    ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(steps)]
    barrier()
    t_wall = time.perf_counter()
    for i in range(steps):
        flush.zero_()
        ev[i][0].record()
        run[i % len(run)]()
        ev[i][1].record()
    barrier()
    t_wall = time.perf_counter() - t_wall
    per = [a.elapsed_time(b) for a, b in ev]
    ms_total = sum(per)
    if world > 1:
        t = torch.tensor([ms_total], device=dev, dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms_total = float(t.item())
    return per, ms_total / steps, t_wall * 1e3 / steps


def graph_or_eager(fns, args, dev, barrier, what):
    """One CUDA graph per callable (transplat_b200.graphs); falls back to eager launches if capture fails."""
    import torch
    if args.eager:
        return list(fns), None
    from transplat_b200.graphs import GraphedStep
    try:
        graphed = [GraphedStep(f, warmup=1, device=dev) for f in fns]
        run = [g.replay for g in graphed]
        for i in range(max(args.warmup, 3)):
            run[i % len(run)]()
        barrier()
        return run, graphed
    except Exception as exc:  # never lose the measurement to a capture problem
        print(f"[bench] CUDA-graph capture of {what} failed ({exc!r}); launching eagerly", file=sys.stderr, flush=True)
        torch.cuda.synchronize()
        return list(fns), None


def run_ours(args):
    import torch
    import torch.distributed as dist
    from transplat_b200 import _lib
    from transplat_b200.decoder import DecoderSplattingCUDA, Gaussians

    rank, local_rank, world = dist_env()
    if not torch.cuda.is_available():
        raise RuntimeError("bench.py (our arm) needs a CUDA device: there is no CPU fallback")
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    _lib.lib()  # fail loudly if libtsplat.so is missing

    W = max(args.warmup, 3)
    scenes, target_h, dLd_h = make_inputs(rank, N_SCENES)
    G = scenes[0].means.shape[0]
    P, T = HW * HW, (HW // 16) ** 2
    decoder = DecoderSplattingCUDA().to(dev)

    def to_dev(sc):
        leaf = lambda t: t.to(dev)
        return dict(means=leaf(sc.means[None]), cov=leaf(sc.covariances[None]), sh=leaf(sc.harmonics[None]),
                    op=leaf(sc.opacities[None]), ext=sc.extrinsics[None].to(dev), intr=sc.intrinsics[None].to(dev),
                    near=sc.near[None].to(dev), far=sc.far[None].to(dev))

    dev_scenes = [to_dev(s) for s in scenes]
    target = target_h[None].to(dev)
    dLd = dLd_h[None].to(dev)
    one = torch.ones((), device=dev)

    def step(s):
        # colour loss = LossMse (loss_mse.py:23-31) fused into the render kernels (forward_mse); depth gets a synthetic
        # upstream gradient (loss_depth.py's role)
        leaves = [s[k].detach().requires_grad_(True) for k in ("means", "cov", "sh", "op")]  # fresh leaves: capture-safe
        out, loss = decoder.forward_mse(Gaussians(*leaves), s["ext"], s["intr"], s["near"], s["far"], (HW, HW), target,
                                        depth_mode="depth")
        grads = torch.autograd.grad([loss, out.depth], leaves, [one, dLd])
        return out, loss, grads, loss.grad_fn.state if hasattr(loss.grad_fn, "state") else None

    def step_dropin(s):
        # the plain drop-in surface: DecoderSplattingCUDA.forward(depth_mode="depth") as the reference calls it
        # (decoder_splatting_cuda.py:41-73) + LossMse as torch ops (loss_mse.py:23-31) + backward
        leaves = [s[k].detach().requires_grad_(True) for k in ("means", "cov", "sh", "op")]
        out = decoder.forward(Gaussians(*leaves), s["ext"], s["intr"], s["near"], s["far"], (HW, HW), depth_mode="depth")
        loss = ((out.color - target) ** 2).mean()
        grads = torch.autograd.grad([loss, out.depth], leaves, [one, dLd])
        return out, loss, grads

    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)  # > 126 MB L2

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    for i in range(W):
        step(dev_scenes[i % N_SCENES])
    barrier()
    # one CUDA graph per resident scene (transplat_b200.graphs): a step is then ONE launch.  Eagerly the host needs
    # 0.3-0.45 ms per step, which starves the GPUs once 8 ranks share the box's 16 vCPUs.
    run, graphed = graph_or_eager([(lambda sc=sc: step(sc)) for sc in dev_scenes], args, dev, barrier, "the step")
    if graphed is None:
        args.eager = True

    # ---- timed region: device-resident inputs, one CUDA-event pair per step, L2 flushed between steps.
    # No per-stage events in here: recording an event between every two kernels of the chain costs 4 % of the step
    # (measured, profiles/r01_experiments.md); the per-kernel durations for the roofline come from a second pass.
    sampler = ClockSampler(local_rank)
    sampler.start()
    per_step, ms_per_step, host_wall = timed_pass(run, args.steps, flush, barrier, world, dev)
    if graphed is not None:  # the pair list of a replayed step must not have overflowed its captured capacity
        from transplat_b200.rasterize import read_status_device
        for gs in graphed:
            assert read_status_device(gs.outputs[3]).overflow == 0, "pair-list overflow inside a captured step"
    value = world * VIEWS / (ms_per_step * 1e-3)

    # ---- the same K steps through the plain drop-in call surface (no fused loss)
    for i in range(W):
        step_dropin(dev_scenes[i % N_SCENES])
    barrier()
    run_d, _ = graph_or_eager([(lambda sc=sc: step_dropin(sc)) for sc in dev_scenes], args, dev, barrier, "the drop-in step")
    per_step_d, ms_dropin, _ = timed_pass(run_d, args.steps, flush, barrier, world, dev)

    # ---- same K steps again with the library's per-stage events on the launching stream (roofline.stage_ms)
    _lib.profile_enable(True)
    for i in range(args.steps):
        flush.zero_()
        step(dev_scenes[i % N_SCENES])
    barrier()
    stages = _lib.profile_read()
    _lib.profile_enable(False)

    # measured pair count D (per call = all 4 views) of scene 0 (the scene the CPU arm renders)
    from transplat_b200.rasterize import _capacity
    _capacity.sweep(wait=True)
    step(dev_scenes[0])
    _capacity.sweep(wait=True)
    D_call = _capacity.last_num_rendered

    # ---- end to end through the public API with HOST buffers (H2D + D2H inside the timed region)
    e2e = None
    if not args.no_e2e:
        e2e = run_e2e(args, dev, decoder, scenes, target_h, dLd_h, flush, barrier, world)
    # ---- the path's only collective: one scene, its views split over the ranks
    view_sharded = None
    if world > 1 and not args.no_view_sharded:
        view_sharded = run_view_sharded(args, dev, decoder, flush, barrier, world, rank, ms_per_step)
    clocks = sampler.stop()  # sampled over the timed passes, the per-stage pass, the e2e leg and the view-sharded leg

    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return

    # ---- roofline of the dominant kernel (SURVEY 8d algorithmic bytes; one launch covers VIEWS views)
    Dv = (D_call or 3 * G * VIEWS) / VIEWS
    alg = {"preprocess": 392.0 * G, "bin_emit": 28.0 * G + 12.0 * Dv, "sort": 152.0 * Dv,
           "ranges_gather": 8.0 * Dv + 8.0 * T, "render_forward": 44.0 * Dv + 24.0 * P,
           "render_backward": 44.0 * Dv + 24.0 * P + 40.0 * G, "preprocess_backward": 772.0 * G}
    kern = {k: stages[k] for k in alg}
    top = max(kern, key=kern.get)
    peaks = {}
    try:
        peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except Exception:
        pass
    peak, peak_src = (peaks["hbm_gbs"], "measured") if "hbm_gbs" in peaks else (6650.0, "fallback")
    top_bytes = alg[top] * VIEWS
    achieved = top_bytes / (kern[top] * 1e-3) / 1e9 if kern[top] > 0 else 0.0
    step_bytes = bytes_per_view(G, Dv, P, T) * VIEWS
    traffic, traffic_src = load_traffic(top)

    def stage_entry(k):
        t, _ = load_traffic(k)
        d = {"ms": round(kern[k], 4), "algorithmic_bytes": round(alg[k] * VIEWS),
             "achieved": round(alg[k] * VIEWS / (kern[k] * 1e-3) / 1e9, 1) if kern[k] > 0 else None,
             # SURVEY 8(d)'s algorithmic bytes count what the REFERENCE's algorithm moves (e.g. a re-read of the Gaussians
             # per view): a fraction above 1 is credit for not moving them, not bandwidth - dram_frac is the bandwidth
             "algorithmic_frac": round(alg[k] * VIEWS / (kern[k] * 1e-3) / 1e9 / peak, 4) if kern[k] > 0 else None,
             "traffic": t, **load_issue(k)}
        # what the kernel actually moves through DRAM (ncu dram__bytes per launch) over its measured duration
        d["achieved_dram_gbs"] = round(t / (kern[k] * 1e-3) / 1e9, 1) if (t and kern[k] > 0) else None
        d["dram_frac"] = round(d["achieved_dram_gbs"] / peak, 4) if d["achieved_dram_gbs"] else None
        return d

    line = {
        "metric": METRIC, "value": round(value, 2), "unit": UNIT, "n_gpus": world, "steps": args.steps,
        "warmup": W, "ms_per_step": round(ms_per_step, 4), "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": shared_config(G, round(Dv)),
        "details": {"parallelism": f"scene-sharded x{world} (no collective)",
                    "api": "DecoderSplattingCUDA.forward_mse(depth_mode='depth') [render + fused MSE colour loss] + torch.autograd.grad"
                           + ("" if args.eager else ", captured once per resident scene in a CUDA graph (transplat_b200.graphs) and replayed")},
        "step_ms": percentiles(per_step),
        "gaussians_per_s": round(world * VIEWS * G / (ms_per_step * 1e-3)),
        "dropin": {"value": round(world * VIEWS / (ms_dropin * 1e-3), 2), "unit": UNIT, "ms_per_step": round(ms_dropin, 4),
                   "step_ms": percentiles(per_step_d),
                   "api": "DecoderSplattingCUDA.forward(depth_mode='depth') + ((color - target) ** 2).mean() as torch ops + "
                          "torch.autograd.grad: the reference's call surface with nothing fused after it"},
        "clocks": clocks,
        "e2e": e2e,
        "gpu_launches": args.steps * launches_per_step(VIEWS, T, graphed=not args.eager),
        "roofline": {"bound": "hbm", "kernel": top, "achieved": round(achieved, 1), "peak": peak, "unit": "GB/s",
                     "frac": round(achieved / peak, 4), "traffic": traffic, "traffic_source": traffic_src,
                     "peak_source": peak_src,
                     "algorithmic_bytes_per_launch": round(top_bytes), "kernel_ms": round(kern[top], 4),
                     "stage_ms": {k: round(v, 4) for k, v in stages.items()},
                     "per_stage": {k: stage_entry(k) for k in alg},
                     "issue": {"what": "the two render kernels have no dense contraction and move little memory: "
                                       "they are bound by instruction issue (ncu --set full capture of this command, "
                                       "profiles/traffic.json: smsp__issue_active, smsp__inst_executed per launch)",
                               **load_issue(top)},
                     "stage_ms_how": "CUDA events around each kernel stage on the launching stream, second pass of the "
                                     "same K steps (the timed pass carries no per-stage events)",
                     "step": {"algorithmic_bytes": round(step_bytes),
                              "achieved": round(step_bytes / (ms_per_step * 1e-3) / 1e9, 1),
                              "frac": round(step_bytes / (ms_per_step * 1e-3) / 1e9 / peak, 4)}},
        "host_wall_ms_per_step": round(host_wall, 4),
    }
    if view_sharded is not None:
        line["view_sharded"] = view_sharded
    if world == 1 and not args.no_cpu_baseline:
        line["cpu_baseline"] = cpu_baseline(scenes[0], target_h, dLd_h, budget_s=12.0)
    print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()


def launches_per_step(n_views: int, T: int, graphed: bool = False) -> int:
    """Kernels of libtsplat.so launched per step (the MSE colour loss runs inside the render kernels).
    bucket path (default): pack_cameras, preprocess_forward, bucket_scan, bucket_emit, tile_sort_gather,
    big_tile_sort_gather, block_lists, render_forward, status_store, render_backward, preprocess_backward = 11
    (the memset node that clears the per-tile pair counters is not a kernel of the library and is not counted);
    onesweep path: pack_cameras, preprocess, bin_emit, sort passes, ranges_gather, block_lists, render_forward,
    status_store, render_backward, preprocess_backward."""
    # a step captured in a CUDA graph carries no status_store: nothing may touch the host under capture, the overflow
    # flag is read from the device after the run (transplat_b200.graphs)
    skip = 1 if graphed else 0
    if os.environ.get("TSPLAT_SORT", "bucket") == "onesweep":
        return 9 + sort_passes(n_views, T) - skip
    return 11 - skip


def sort_passes(n_views: int, T: int) -> int:
    bits = 32 + max(1, (n_views * T - 1).bit_length())
    return (bits + 7) // 8


def load_traffic(kernel: str):
    """(dram bytes per launch, where the number comes from) of a kernel.  The bytes are NOT measured in this run: they
    are read from the committed summary of an `ncu --set full` capture of this same command (profiles/traffic.json)."""
    try:
        d = json.load(open(os.path.join(ROOT, "profiles", "traffic.json")))
        return d.get(kernel), "profiles/traffic.json (" + d.get("_source", "ncu --set full capture of this command") + ")"
    except Exception:
        return None, None


def load_issue(kernel: str) -> dict:
    """issue-slot utilisation / warp instructions per launch of a render kernel from the committed ncu summary."""
    try:
        d = json.load(open(os.path.join(ROOT, "profiles", "traffic.json")))
        return dict(d.get("_issue", {}).get(kernel, {}))
    except Exception:
        return {}


def run_e2e(args, dev, decoder, scenes, target_h, dLd_h, flush, barrier, world):
    """Same step through the public API, but every step starts from PINNED HOST buffers and ends with
    the results (colour, depth, the four gradients) back in pinned host memory.  The copies run on
    their own streams (transplat_b200.streaming.HostPipeline, 3 device slots), so the H2D of step i+1 and
    the D2H of step i-1 overlap the kernels of step i; every step still moves all of its bytes.  Timed
    with two CUDA events around the whole window (first H2D issued -> last D2H finished)."""
    import torch
    import torch.distributed as dist
    from transplat_b200.decoder import Gaussians
    from transplat_b200.streaming import HostPipeline

    from transplat_b200.rasterize import flat_gradients, flat_images
    from transplat_b200.streaming import SlabLayout

    G = scenes[0].means.shape[0]
    # one pinned slab per scene for the inputs, two per result set (images, gradients): a step crosses PCIe as
    # one H2D and two D2H copies (scripts/pcie_bw.py: 16 separate tensor copies cost 20 % more for the same bytes)
    lay_in = SlabLayout(dict(means=(1, G, 3), cov=(1, G, 3, 3), sh=(1, G, 3, 25), op=(1, G), ext=(1, VIEWS, 4, 4),
                             intr=(1, VIEWS, 3, 3), near=(1, VIEWS), far=(1, VIEWS), target=(1, VIEWS, 3, HW, HW),
                             dLd=(1, VIEWS, HW, HW)))
    lay_img = SlabLayout(dict(color=(1, VIEWS, 3, HW, HW), depth=(1, VIEWS, HW, HW)))
    lay_grad = SlabLayout(dict(g_means=(1, G, 3), g_cov=(1, G, 3, 3), g_sh=(1, G, 3, 25), g_op=(1, G)))
    n_out = 4  # result buffers cycled (a consumer would read them between steps)
    # the staging buffers are allocated with this rank bound to its GPU's NUMA node (TSPLAT_NUMA=0: wherever it runs)
    from transplat_b200.streaming import numa_local
    import contextlib
    numa = numa_local(dev) if os.environ.get("TSPLAT_NUMA", "1") != "0" else contextlib.nullcontext()
    with numa:
        host = [dict(slab=lay_in.pack_pinned(dict(
            means=s.means[None], cov=s.covariances[None], sh=s.harmonics[None], op=s.opacities[None], ext=s.extrinsics[None],
            intr=s.intrinsics[None], near=s.near[None], far=s.far[None], target=target_h[None], dLd=dLd_h[None])))
            for s in scenes]
        outs = [dict(images=lay_img.empty_pinned(), grads=lay_grad.empty_pinned()) for _ in range(n_out)]
        for o in outs:  # first touch while bound
            for t in o.values():
                t.zero_()
    numa_info = getattr(numa, "info", {"bound": False, "disabled": True})
    one = torch.ones((), device=dev)
    h2d = host[0]["slab"].numel() * 4
    d2h = sum(t.numel() * 4 for t in outs[0].values())

    def step_fn(dv):
        d = lay_in.views(dv["slab"])
        leaves = [d[k].detach().requires_grad_(True) for k in ("means", "cov", "sh", "op")]
        out, loss = decoder.forward_mse(Gaussians(*leaves), d["ext"], d["intr"], d["near"], d["far"], (HW, HW), d["target"],
                                        depth_mode="depth")
        g = torch.autograd.grad([loss, out.depth], leaves, [one, d["dLd"]])
        images, grads = flat_images(out.color, out.depth), flat_gradients(*g)
        assert images is not None and images.numel() == lay_img.total and grads.numel() == lay_grad.total
        return dict(images=images, grads=grads)

    pipe = HostPipeline(dev, slots=3)
    steps = max(10, min(args.steps, 100))
    if not args.eager:
        # one CUDA graph per device slot (its input slab sits at a fixed address): the kernels of a step are one launch
        from transplat_b200.graphs import GraphedStep
        eager_fn, graphs = step_fn, {}

        def step_fn(dv):
            key = dv["slab"].data_ptr()
            if key not in graphs:
                try:
                    graphs[key] = GraphedStep(lambda: eager_fn(dv), warmup=1, device=dev)
                except Exception as exc:
                    print(f"[bench] e2e graph capture failed ({exc!r}); eager", file=sys.stderr, flush=True)
                    graphs[key] = None
            return graphs[key].replay() if graphs[key] is not None else eager_fn(dv)
    for i in range(6):
        pipe.submit(step_fn, host[i % N_SCENES], outs[i % n_out])
    pipe.drain()
    barrier()
    t0, t1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    t0.record(pipe.h2d)
    for i in range(steps):
        pipe.submit(step_fn, host[i % N_SCENES], outs[i % n_out])
    t1.record(pipe.d2h)
    pipe.drain()
    barrier()
    ms = t0.elapsed_time(t1)
    if world > 1:
        t = torch.tensor([ms], device=dev, dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms = float(t.item())
    ms /= steps
    return {"value": round(world * VIEWS / (ms * 1e-3), 2), "unit": UNIT, "h2d_bytes_per_step": int(h2d),
            "d2h_bytes_per_step": int(d2h), "ms_per_step": round(ms, 4), "steps": steps,
            "how": "pinned host slab (Gaussians, cameras, target) -> H2D stream -> kernels -> D2H stream -> pinned host "
                   "slabs (colour+depth, all four gradients), 3 slots in flight"
                   + ("" if args.eager else ", the kernels of a step replayed from one CUDA graph per slot") + "; "
                   f"{N_SCENES} scenes x {h2d >> 20} MiB cycled (> L2), window timed with CUDA events",
            "numa": numa_info}


# ------------------------------------------------------------------------------------------------
# view sharding (BASELINE configs 4b and 5): ONE scene replicated on every rank, its target views split
# over the ranks, per-Gaussian gradients (85 floats per Gaussian) summed with the path's only collective.
# ------------------------------------------------------------------------------------------------
def run_view_sharded(args, dev, decoder, flush, barrier, world, rank, ms_scene_sharded):
    import torch
    import torch.distributed as dist
    from transplat_b200.decoder import Gaussians
    from transplat_b200.parallel import NVLS_MIN_WORLD, forward_view_sharded
    from transplat_b200.synthetic import make_scene, upstream_grads

    def one_shape(name, hw, per_rank, steps, chunks):
        V = per_rank * world
        sc = make_scene(n_ctx=G_CTX, h=hw, w=hw, n_views=V, seed=4242)   # the same scene on every rank
        tgt, dd = upstream_grads(V, hw, hw, seed=7)
        mine = list(range(rank, V, world))
        tgt, dd = tgt[mine][None].to(dev), dd[mine][None].to(dev)
        G = sc.means.shape[0]
        stat = dict(means=sc.means[None].to(dev), cov=sc.covariances[None].to(dev), sh=sc.harmonics[None].to(dev),
                    op=sc.opacities[None].to(dev))
        cam = [t[None].to(dev) for t in (sc.extrinsics, sc.intrinsics, sc.near, sc.far)]
        one = torch.ones((), device=dev)

        def step(reduce=True, overlap=0, fused=False, nvls=False, ctas=0, default=False):
            leaves = [stat[k].detach().requires_grad_(True) for k in ("means", "cov", "sh", "op")]
            g = Gaussians(*leaves)
            if default:  # what a caller gets without choosing: parallel.forward_view_sharded picks the exchange
                out, _ = forward_view_sharded(decoder, g, *cam, (hw, hw), depth_mode="depth")
            elif reduce:
                out, _ = forward_view_sharded(decoder, g, *cam, (hw, hw), depth_mode="depth", overlap_chunks=overlap,
                                              fused_reduce=fused, nvls_reduce=nvls, nvls_ctas=ctas)
            else:  # the same local views, no collective: what the step costs without the exchange
                idx = torch.tensor(mine, device=dev)
                out = decoder.forward(g, *(c[:, idx] for c in cam), (hw, hw), depth_mode="depth")
            loss = ((out.color - tgt) ** 2).mean()
            return torch.autograd.grad([loss, out.depth], leaves, [one, dd])

        def time_it(fn, k):
            for _ in range(3):
                fn()
            ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(k)]
            barrier()
            for i in range(k):
                flush.zero_()
                ev[i][0].record()
                fn()
                ev[i][1].record()
            barrier()
            t = torch.tensor([sum(a.elapsed_time(b) for a, b in ev) / k], device=dev, dtype=torch.float64)
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            return float(t.item())

        t_plain = time_it(lambda: step(True, 0), steps)
        t_overlap = time_it(lambda: step(True, chunks), steps) if chunks > 1 else None
        try:  # the reduction inside the producing kernel (multimem.red through NVLS multicast memory)
            t_fused = time_it(lambda: step(True, 0, True), steps)
        except Exception as exc:
            print(f"[bench] fused multimem reduction unavailable ({exc!r})", file=sys.stderr, flush=True)
            t_fused = None
        from transplat_b200.rasterize import _pad64
        n_slab = _pad64(G * 3) + _pad64(G * 9) + _pad64(G * 75) + _pad64(G)
        # our own exchange kernel through the switch (csrc/exchange.cu): whole slab behind the kernels, and range by
        # range on a side stream under the per-Gaussian kernel of the next range; then alone, over its grid size
        t_nvls = t_nvls_chunked = t_nvls_chunked2 = t_nvls_chunked4 = None
        nvls_alone = {}
        try:
            from transplat_b200.parallel import SymmetricGradSlab
            t_nvls = time_it(lambda: step(True, 0, nvls=True), steps)
            t_nvls_chunked = time_it(lambda: step(True, chunks, nvls=True), steps) if chunks > 1 else None
            t_nvls_chunked2 = time_it(lambda: step(True, 2 * chunks, nvls=True), steps) if chunks > 1 else None
            t_nvls_chunked4 = time_it(lambda: step(True, 4 * chunks, nvls=True), steps) if chunks > 1 else None
            sym = SymmetricGradSlab.get(None, dev, n_slab)

            def alone(c):
                sym.barrier(0)
                sym.all_reduce([(0, n_slab)], c)
                sym.barrier(0)
            for c in (16, 32, 64, 148, 296):
                nvls_alone[str(c)] = round(time_it(lambda: alone(c), max(steps, 10)), 4)
        except Exception as exc:
            print(f"[bench] NVLS exchange unavailable ({exc!r})", file=sys.stderr, flush=True)
        t_local = time_it(lambda: step(False), steps)
        # the collective alone, on a slab of the same size
        slab = torch.zeros(n_slab, dtype=torch.float32, device=dev)
        t_ar = time_it(lambda: dist.all_reduce(slab), max(steps, 10))
        named = {"one_all_reduce": t_plain, "chunked_overlap": t_overlap, "fused_multimem": t_fused, "nvls_two_shot": t_nvls,
                 "nvls_two_shot_chunked": t_nvls_chunked}
        fastest = min((k for k in named if named[k] is not None), key=lambda k: named[k])
        best = named[fastest]
        t_default = time_it(lambda: step(default=True), steps)
        # the same default step captured once per rank in a CUDA graph (exchange kernels and barriers included) and
        # replayed: at the C2 shape the eager step is bound by the host (Python + ~20 launches), not by the GPU
        t_graphed = None
        try:
            from transplat_b200.graphs import GraphedStep
            gs = GraphedStep(lambda: step(default=True))
            t_graphed = time_it(gs.replay, steps)
            del gs
        except Exception as exc:
            print(f"[bench] graph capture of the view-sharded step unavailable ({exc!r})", file=sys.stderr, flush=True)
        t_eager_default = t_default
        if t_graphed is not None and t_graphed < t_default:
            t_default = t_graphed
        nbytes = n_slab * 4
        return {"shape": name, "views": V, "views_per_rank": per_rank, "gaussians": G, "image": [hw, hw], "steps": steps,
                "ms_per_step": round(t_default, 4), "views_per_s": round(V / (t_default * 1e-3), 1),
                "ms_per_step_one_all_reduce": round(t_plain, 4),
                "ms_per_step_chunked_overlap": round(t_overlap, 4) if t_overlap is not None else None, "chunks": chunks,
                "ms_per_step_fused_multimem": round(t_fused, 4) if t_fused is not None else None,
                "ms_per_step_nvls_two_shot": round(t_nvls, 4) if t_nvls is not None else None,
                "ms_per_step_nvls_two_shot_chunked": round(t_nvls_chunked, 4) if t_nvls_chunked is not None else None,
                "ms_per_step_nvls_two_shot_chunked_twice_the_chunks": round(t_nvls_chunked2, 4) if t_nvls_chunked2 is not None else None,
                "ms_per_step_nvls_two_shot_chunked_four_times_the_chunks": round(t_nvls_chunked4, 4) if t_nvls_chunked4 is not None else None,
                "nvls_exchange_alone_ms_by_ctas": nvls_alone,
                "fastest": fastest,
                "default": "forward_view_sharded() with no options: ncclAllReduce below %d ranks, from there the NVLS "
                           "exchange, range by range under the per-Gaussian kernel when the slab is large" % NVLS_MIN_WORLD,
                "ms_per_step_without_collective": round(t_local, 4),
                "exchange_exposed_ms": round(t_eager_default - t_local, 4),  # eager against eager
                "all_reduce_exposed_ms": round(t_plain - t_local, 4),
                "all_reduce_alone_ms": round(t_ar, 4), "all_reduce_bytes": int(nbytes),
                "all_reduce_bus_gbs": round(2.0 * (world - 1) / world * nbytes / (t_ar * 1e-3) / 1e9, 1),
                "ms_per_step_default_eager": round(t_eager_default, 4),
                "ms_per_step_default_graphed": round(t_graphed, 4) if t_graphed is not None else None,
                "launch": ("ms_per_step: one CUDA-graph replay per step and rank (forward, loss, backward, exchange); the "
                           "variants: eager" if t_default is t_graphed else "eager (one process per GPU)")}

    out = {"what": "one scene, target views split round-robin over the ranks, Gaussians replicated, per-Gaussian gradients "
                   "summed over the ranks: the only exchange on the path (SURVEY 8e).  one_all_reduce = ncclAllReduce of "
                   "the flat gradient slab behind the backward kernels; chunked_overlap = per-Gaussian-range all-reduces "
                   "under the next range's kernels; fused_multimem = preprocess_backward adds its values with "
                   "multimem.red.add.f32 through the multicast alias of a symmetric slab (NVLS: the switch reduces into "
                   "every rank's replica while the kernel computes; no collective follows); nvls_two_shot = the kernels write "
                   "this rank's replica of a symmetric slab, our own kernel (csrc/exchange.cu: multimem.ld_reduce + "
                   "multimem.st) sums the replicas in place through the switch - whole slab, or (chunked) range by range on "
                   "a side stream under the per-Gaussian kernel of the next range.  fwd+bwd through "
                   "transplat_b200.parallel.forward_view_sharded, device time, max over ranks",
           "c2_shape": one_shape("C2 shape: 131072 Gaussians, 256x256, 4 views per rank", HW, VIEWS, min(args.steps, 30), 2)}
    if args.c5:
        out["c5_shape"] = one_shape("C5 stress: 2097152 Gaussians, 1024x1024, 8 views per rank", 1024, 8, 5, 3)
    return out


# ------------------------------------------------------------------------------------------------
# CPU arm: the oracle port (oracle/tsplat_oracle.c, OpenMP) on the host cores.  The reference's own
# arithmetic lives in an external extension whose source is absent, so "reference" here
# is the port, kind "port".
# ------------------------------------------------------------------------------------------------
def oracle_threads() -> int:
    """All the host cores this process may use, set explicitly: torch.distributed.run exports OMP_NUM_THREADS=1."""
    import oracle
    try:
        n = len(os.sched_getaffinity(0))
    except Exception:
        n = os.cpu_count() or 1
    return oracle.set_num_threads(n)


def oracle_step(scene, target, dLd):
    """One C2 step on the CPU: 4 views colour+depth forward, MSE gradient, backward summed over views.
    Returns (gradient accumulators, pairs per view)."""
    import numpy as np
    import torch
    import oracle
    from transplat_b200.cameras import pack_cameras

    cams = pack_cameras(scene.extrinsics, scene.intrinsics, scene.near, scene.far, torch.zeros(VIEWS, 3)).numpy()
    sc = oracle.Scene(scene.means.numpy(), scene.covariances.numpy(), scene.harmonics.numpy(), scene.opacities.numpy(),
                      4, oracle.SH_G3M, oracle.COV_FULL9)
    acc = oracle.GradAccum(sc)
    numel = float(VIEWS * 3 * HW * HW)
    tgt, dd = target.numpy(), dLd.numpy()
    pairs = 0
    for v in range(VIEWS):
        st = oracle.forward(sc, cams[v], HW, HW, depth_mode="depth")
        pairs += st.num_rendered
        dLc = (st.color - tgt[v]) * np.float32(2.0 / numel)
        oracle.backward(sc, cams[v], st, dLc, dd[v], acc=acc, depth_mode="depth")
    return acc, pairs / VIEWS


def cpu_baseline(scene, target, dLd, budget_s: float):
    cores = oracle_threads()
    oracle_step(scene, target, dLd)  # warm-up (builds the library, pages it in)
    t0, n = time.perf_counter(), 0
    while n < 1 or (time.perf_counter() - t0 < budget_s and n < 16):
        oracle_step(scene, target, dLd)
        n += 1
    dt = time.perf_counter() - t0
    return {"value": round(n * VIEWS / dt, 3), "unit": UNIT, "cores": cores, "kind": "port",
            "sample": f"{n} full C2 steps (1 scene x {VIEWS} views fwd+bwd, colour+depth) of the same workload, "
                      f"oracle/tsplat_oracle.c with OpenMP, {dt:.1f} s"}


def run_reference(args):
    rank, _, world = dist_env()
    if rank != 0:
        return
    cores = oracle_threads()
    W = max(args.warmup, 3)
    scenes, target, dLd = make_inputs(0, 1)
    pairs = 0
    for _ in range(W):
        _, pairs = oracle_step(scenes[0], target, dLd)
    per = []
    for _ in range(args.steps):
        t0 = time.perf_counter()
        oracle_step(scenes[0], target, dLd)
        per.append((time.perf_counter() - t0) * 1e3)
    dt = sum(per) * 1e-3
    value = args.steps * VIEWS / dt
    G = int(scenes[0].means.shape[0])
    line = {"impl": "reference", "metric": METRIC, "value": round(value, 3), "unit": UNIT, "n_gpus": args.gpus,
            "steps": args.steps, "warmup": W, "ms_per_step": round(dt * 1e3 / args.steps, 3),
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": shared_config(G, round(pairs)),
            "details": {"note": "CPU port of the rasterizer (the reference's own CUDA extension is an unpinned external "
                                "dependency whose source is absent: the ref-CUDA arm is not available on this box); rank 0 "
                                "only, all host cores (OpenMP team set explicitly)"},
            "step_ms": percentiles(per),
            "gaussians_per_s": round(VIEWS * G / (dt / args.steps)),
            "cpu_baseline": {"value": round(value, 3), "unit": UNIT, "cores": cores, "kind": "port",
                             "sample": f"{args.steps} full C2 steps, {dt:.1f} s"},
            "e2e": {"value": round(value, 3), "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    print(json.dumps(line), flush=True)


if __name__ == "__main__":
    a = parse()
    if a.impl == "reference":
        run_reference(a)
    else:
        run_ours(a)
