"""Host-buffer pipeline for the render path: steps whose inputs live in PINNED HOST memory and whose
results return to pinned host memory, software-pipelined over three CUDA streams (H2D, compute,
D2H) with a small ring of device slots, so the PCIe copies of step i+1 / i-1 overlap the kernels of
step i.  PyTorch is plumbing (streams, events, the caching allocator); the work inside ``step_fn``
is the C-ABI library.

The reference has no such layer: its decoder is fed by the encoder on the device
(src/model/model_wrapper.py:137-145).  This is what a caller uses when the Gaussians come from
somewhere else (a file, another process, another node) - and what ``bench.py`` times as ``e2e``.
"""
from __future__ import annotations

from typing import Callable, Dict

import torch


class SlabLayout:
    """Several float32 tensors packed in ONE flat buffer (segments 256-B aligned), so a step's inputs or results
    cross PCIe as one copy: 10 + 6 separate tensor copies per C2 step cost 1.27 ms against 1.07 ms for the same
    bytes in one copy per direction (scripts/pcie_bw.py on the B200 box)."""

    def __init__(self, shapes: Dict[str, tuple]):
        self.shapes = {n: tuple(int(x) for x in shp) for n, shp in shapes.items()}
        self.offsets, off = {}, 0
        for n, shp in self.shapes.items():
            self.offsets[n] = off
            numel = 1
            for x in shp:
                numel *= x
            off += (numel + 63) // 64 * 64
        self.total = off

    def views(self, slab: torch.Tensor) -> Dict[str, torch.Tensor]:
        out = {}
        for n, shp in self.shapes.items():
            numel = 1
            for x in shp:
                numel *= x
            out[n] = slab[self.offsets[n]:self.offsets[n] + numel].view(shp)
        return out

    def pack_pinned(self, tensors: Dict[str, torch.Tensor]) -> torch.Tensor:
        slab = torch.zeros(self.total, dtype=torch.float32).pin_memory()
        for n, v in self.views(slab).items():
            v.copy_(tensors[n])
        return slab

    def empty_pinned(self) -> torch.Tensor:
        return torch.empty(self.total, dtype=torch.float32).pin_memory()


def _parse_cpulist(text: str) -> set:
    cpus = set()
    for part in text.strip().split(","):
        if not part:
            continue
        lo, _, hi = part.partition("-")
        cpus.update(range(int(lo), int(hi or lo) + 1))
    return cpus


class numa_local:
    """``with numa_local(device): allocate pinned buffers`` - runs the block with this thread bound to the CPUs of the
    NUMA node the GPU hangs off, so that the pinned host memory it allocates is local to that GPU's PCIe root (first
    touch), then restores the affinity.  With one rank per GPU on a two-socket host, unbound ranks put half of their
    staging buffers behind the socket interconnect.  Best effort: without /sys, on a single node or inside a cpuset
    that excludes the node it does nothing; ``info`` says what happened."""

    def __init__(self, device):
        self.device = torch.device(device)
        self.info = {"bound": False}
        self._saved = None

    def __enter__(self):
        import os
        try:
            props = torch.cuda.get_device_properties(self.device)
            bus = "%04x:%02x:%02x.0" % (getattr(props, "pci_domain_id", 0), props.pci_bus_id, props.pci_device_id)
            node = int(open(f"/sys/bus/pci/devices/{bus}/numa_node").read())
            self.info.update(pci=bus, numa_node=node)
            if node >= 0:
                cpus = _parse_cpulist(open(f"/sys/devices/system/node/node{node}/cpulist").read())
                have = os.sched_getaffinity(0)
                want = cpus & have
                self.info.update(node_cpus=len(cpus), usable_cpus=len(want))
                if want and want != have:
                    self._saved = have
                    os.sched_setaffinity(0, want)
                    self.info["bound"] = True
        except Exception as exc:  # no sysfs, no such attribute, no permission: stay as we are
            self.info["error"] = repr(exc)
        return self

    def __exit__(self, *exc):
        import os
        if self._saved is not None:
            try:
                os.sched_setaffinity(0, self._saved)
            except Exception:
                pass
        return False


class HostPipeline:
    def __init__(self, device: torch.device, slots: int = 3, trace: bool = False):
        self.device = device
        self.trace = [] if trace else None  # per step: timing events (h2d0, h2d1, c0, c1, d2h0, d2h1)
        self.slots = slots
        self.h2d = torch.cuda.Stream(device)
        self.d2h = torch.cuda.Stream(device)
        self.compute = torch.cuda.current_stream(device)
        self._dev_in = [None] * slots      # per-slot device input buffers (allocated once per shape)
        self._keep = [None] * slots        # per-slot outputs kept alive until their D2H has finished
        self._done = [None] * slots        # per-slot event: D2H of the step that last used the slot
        self._computed = [None] * slots    # per-slot event: kernels of the step that last used the slot
        self._i = 0

    def _buffers(self, k: int, host_in: Dict[str, torch.Tensor]):
        cur = self._dev_in[k]
        if cur is None or any(cur[n].shape != t.shape or cur[n].dtype != t.dtype for n, t in host_in.items()):
            cur = {n: torch.empty(t.shape, dtype=t.dtype, device=self.device) for n, t in host_in.items()}
            self._dev_in[k] = cur
            # the caching allocator may hand out memory that kernels already queued on the compute stream still
            # use (stream-ordered reuse is only safe on the allocating stream): order the copy stream behind them
            self.h2d.wait_stream(self.compute)
        return cur

    def submit(self, step_fn: Callable[[Dict[str, torch.Tensor]], Dict[str, torch.Tensor]],
               host_in: Dict[str, torch.Tensor], host_out: Dict[str, torch.Tensor]) -> torch.cuda.Event:
        """Enqueue one step.  ``host_in`` / ``host_out`` are pinned tensors; ``step_fn`` maps the device
        copies of ``host_in`` to a dict of device results with the keys of ``host_out``.  Returns the event
        that marks ``host_out`` complete.  The caller must not reuse a ``host_out`` buffer before then."""
        k = self._i % self.slots
        self._i += 1
        if self._done[k] is not None:
            self._done[k].synchronize()  # back-pressure: the slot's previous step has fully drained
        dev = self._buffers(k, host_in)
        # only THIS slot's previous compute must have finished before its inputs are overwritten; the other
        # slots' kernels keep running under this copy (a stream-wide wait would serialise copy and compute)
        if self._computed[k] is not None:
            self.h2d.wait_event(self._computed[k])
        tr = [torch.cuda.Event(enable_timing=True) for _ in range(6)] if self.trace is not None else None
        with torch.cuda.stream(self.h2d):
            if tr:
                tr[0].record(self.h2d)
            for n, t in host_in.items():
                dev[n].copy_(t, non_blocking=True)
            ev_in = torch.cuda.Event()
            ev_in.record(self.h2d)
            if tr:
                tr[1].record(self.h2d)
        self.compute.wait_event(ev_in)
        if tr:
            tr[2].record(self.compute)
        out = step_fn(dev)
        ev_c = torch.cuda.Event()
        ev_c.record(self.compute)
        if tr:
            tr[3].record(self.compute)
        self.d2h.wait_event(ev_c)
        with torch.cuda.stream(self.d2h):
            if tr:
                tr[4].record(self.d2h)
            for n, h in host_out.items():
                h.copy_(out[n].detach(), non_blocking=True)
            ev_out = torch.cuda.Event()
            ev_out.record(self.d2h)
            if tr:
                tr[5].record(self.d2h)
                self.trace.append(tr)
        self._computed[k] = ev_c
        self._keep[k] = out
        self._done[k] = ev_out
        return ev_out

    def drain(self) -> None:
        for ev in self._done:
            if ev is not None:
                ev.synchronize()
        self._keep = [None] * self.slots
