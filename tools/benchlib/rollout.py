"""BASELINE.json config 5 (makani.inference autoregressive rollout, 40 steps, batch 1, bf16; the loop of
makani/utils/inferencer.py:167-206) for the spectral path: 40 chained forward passes of the eight SpectralConv blocks
under `torch.inference_mode()` (no optimizer between steps, so the packed weights are cached), eager and replayed from
a CUDA graph.  Each pass feeds its full-resolution output back as the next input (autoregression) and the
result of the last pass is copied into a pinned host buffer (a forward-only run has an output to deliver; a pageable
destination made that one 797 MB copy take ~370 ms in round 2's first measurement, 9 of the 14 ms per pass)."""
import torch

from . import workload as W


def measure(blocks, x0, steps=40, warmup=2):
    from makani_b200 import _ext
    out = {"workload": "40 chained forward passes of the 8 SpectralConv blocks (block 7's 721x1440 output is the next "
                       "input), batch 1, bf16, cached packed weights, result of the last pass copied to the host",
           "steps": steps}

    host = torch.empty(x0.shape, dtype=x0.dtype, pin_memory=True)

    def timed(fn, n):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        x = x0
        for _ in range(n):
            x = fn(x)
        host.copy_(x, non_blocking=True)
        e1.record()
        torch.cuda.synchronize()
        return e0.elapsed_time(e1) / n, host

    with torch.inference_mode():
        fwd = lambda x: W.forward_chain(blocks, x)[0]                  # noqa: E731
        for _ in range(warmup):
            y = fwd(x0)
        torch.cuda.synchronize()
        l0 = _ext.launch_count()
        ms_eager, host = timed(fwd, steps)
        out["ms_per_step_eager"] = ms_eager
        out["launches_per_step"] = (_ext.launch_count() - l0) // steps
        out["d2h_bytes"] = host.numel() * host.element_size()

        # CUDA graph of one pass with a static input buffer; the autoregression copies the output back into it
        xin = x0.clone()
        s = torch.cuda.Stream()
        s.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(s):
            fwd(xin)
        torch.cuda.current_stream().wait_stream(s)
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g):
            yg = fwd(xin)
        g.replay()
        torch.cuda.synchronize()
        out["graph_equals_eager"] = bool(torch.equal(yg, y))

        def graph_step(x):
            xin.copy_(x)
            g.replay()
            return yg

        ms_graph, _ = timed(graph_step, steps)
        out["ms_per_step_graph"] = ms_graph
    out["steps_per_s_eager"] = 1e3 / ms_eager
    out["steps_per_s_graph"] = 1e3 / ms_graph
    return out
