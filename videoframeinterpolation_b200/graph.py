"""CUDA-graph replay of the DAT forward for a fixed input shape.

One 1080p forward is ~360 kernel launches (cuDNN convs, vfi_b200 kernels, elementwise glue); issued eagerly from
Python the host needs longer to enqueue them than the GPU needs to run them.  Capturing the forward once and
replaying it removes the host from the loop (every vfi_b200 entry point is capture-safe: no allocation, no
synchronisation, scratch is caller-supplied)."""
from __future__ import annotations

import torch


class GraphedInference:
    """``g = GraphedInference(net, x0, x1, t); y = g(x0, x1, t)`` -- same result as ``net.inference(x0, x1, t)``.

    The example inputs fix shape, dtype, device and memory format.  ``__call__`` copies its arguments into the static
    input buffers (they may be pinned host tensors: the copy is the H2D transfer), replays the graph and returns the
    static output tensor (valid until the next call)."""

    def __init__(self, net, x0: torch.Tensor, x1: torch.Tensor, t: torch.Tensor, warmup: int = 3, fn=None):
        """``fn(x0, x1, t)``: what to capture instead of ``net.inference`` (e.g. this rank's tiles of
        ``net.inference_tiled``); it must be capture-safe like the forward itself."""
        assert x0.is_cuda and x1.is_cuda and t.is_cuda, "GraphedInference needs CUDA example inputs"
        self.net = net
        fn = fn if fn is not None else net.inference
        self.x0, self.x1, self.t = x0.clone(memory_format=torch.preserve_format), x1.clone(memory_format=torch.preserve_format), t.clone()
        side = torch.cuda.Stream(device=x0.device)
        side.wait_stream(torch.cuda.current_stream(x0.device))
        with torch.cuda.stream(side), torch.no_grad():
            for _ in range(max(warmup, 1)):          # cuDNN autotuning, lazy caches, allocator warm-up: outside the capture
                fn(self.x0, self.x1, self.t)
        torch.cuda.current_stream(x0.device).wait_stream(side)
        torch.cuda.synchronize(x0.device)
        self.graph = torch.cuda.CUDAGraph()
        with torch.cuda.graph(self.graph), torch.no_grad():
            self.out = fn(self.x0, self.x1, self.t)

    def replay(self) -> torch.Tensor:
        self.graph.replay()
        return self.out

    def __call__(self, x0: torch.Tensor, x1: torch.Tensor, t: torch.Tensor = None) -> torch.Tensor:
        self.x0.copy_(x0, non_blocking=True)
        self.x1.copy_(x1, non_blocking=True)
        if t is not None:
            self.t.copy_(t, non_blocking=True)
        return self.replay()


class PipelinedInference:
    """Streaming form of :class:`GraphedInference` for host-resident frames (a video being interpolated): the upload of
    pair i + 1 and the download of result i - 1 run on two copy streams while the graph of pair i runs on the compute stream.

        pipe = PipelinedInference(net, example_x0, example_x1, t, fn=net.interpolate)
        for x0, x1 in pairs:                  # pinned host tensors
            pipe.submit(x0, x1)
            if pipe.pending() > 1:
                frame = pipe.collect()        # pinned host tensor, valid until two more submits
        while pipe.pending(): frame = pipe.collect()

    Two staging slots per direction; every hand-over is a CUDA event, the host never synchronises inside submit().
    Same result as ``fn(x0.cuda(), x1.cuda(), t)``."""

    def __init__(self, net, x0: torch.Tensor, x1: torch.Tensor, t: torch.Tensor, fn=None, warmup: int = 3):
        dev = t.device
        self.dev = dev
        self.g = GraphedInference(net, x0.to(dev), x1.to(dev), t, warmup=warmup, fn=fn)
        self.h2d_stream = torch.cuda.Stream(device=dev)             # separate streams: an upload must not queue behind a
        self.d2h_stream = torch.cuda.Stream(device=dev)             # download that waits for the running graph
        self.in0 = [torch.empty_like(self.g.x0) for _ in range(2)]
        self.in1 = [torch.empty_like(self.g.x1) for _ in range(2)]
        self.out_dev = [torch.empty_like(self.g.out) for _ in range(2)]
        self.out_host = [torch.empty(self.g.out.shape, dtype=self.g.out.dtype).pin_memory() for _ in range(2)]
        mk = lambda: [torch.cuda.Event() for _ in range(2)]
        self.uploaded, self.consumed, self.computed, self.downloaded = mk(), mk(), mk(), mk()
        self.n_sub = self.n_col = 0

    def pending(self) -> int:
        return self.n_sub - self.n_col

    def submit(self, x0: torch.Tensor, x1: torch.Tensor) -> None:
        assert self.pending() < 2, "collect() a result before submitting a third pair"
        s = self.n_sub & 1
        comp = torch.cuda.current_stream(self.dev)
        with torch.cuda.stream(self.h2d_stream):
            if self.n_sub >= 2:
                self.h2d_stream.wait_event(self.consumed[s])         # the graph of pair n - 2 has read this slot
            self.in0[s].copy_(x0, non_blocking=True)
            self.in1[s].copy_(x1, non_blocking=True)
            self.uploaded[s].record(self.h2d_stream)
        comp.wait_event(self.uploaded[s])
        self.g.x0.copy_(self.in0[s], non_blocking=True)             # device to device: the graph reads fixed buffers
        self.g.x1.copy_(self.in1[s], non_blocking=True)
        self.consumed[s].record(comp)
        if self.n_sub >= 2:
            comp.wait_event(self.downloaded[s])                     # result n - 2 has left its device slot
        self.g.replay()
        self.out_dev[s].copy_(self.g.out, non_blocking=True)
        self.computed[s].record(comp)
        with torch.cuda.stream(self.d2h_stream):
            self.d2h_stream.wait_event(self.computed[s])
            self.out_host[s].copy_(self.out_dev[s], non_blocking=True)
            self.downloaded[s].record(self.d2h_stream)
        self.n_sub += 1

    def collect(self) -> torch.Tensor:
        assert self.pending() > 0, "nothing submitted"
        s = self.n_col & 1
        self.downloaded[s].synchronize()
        self.n_col += 1
        return self.out_host[s]
