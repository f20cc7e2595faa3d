"""CPU reference arm: the SAME workload as the GPU arm (eight chained SpectralConv blocks, forward + backward, batch 1,
bf16 spatial tensors, fp32 spectral arithmetic - what makani/models/common/spectral_convolution.py:124-148 does under
bf16 AMP) through the oracle port of the reference's torch path (oracle/sht_oracle.py; the reference itself cannot be
imported: torch-harmonics / modulus are absent), on all host cores.  Nothing is extrapolated: a step is a step."""
import os
import time

import torch

from . import workload as W


class CpuStep:
    def __init__(self, threads=None, channels=None):
        from oracle import sht_oracle as O
        self.O = O
        self.threads = threads or os.cpu_count() or 1
        torch.set_num_threads(self.threads)
        C = channels or W.C_EMBED
        self.C = C
        torch.manual_seed(0)
        mk = lambda cls, g: cls(g[0], g[1], lmax=W.L_MODES, mmax=W.M_MODES, grid=g[2], dtype=torch.float32)  # noqa: E731
        t = {"sht_in": mk(O.RealSHT, W.GRID_IN), "sht_int": mk(O.RealSHT, W.GRID_INT),
             "isht_int": mk(O.InverseRealSHT, W.GRID_INT), "isht_out": mk(O.InverseRealSHT, W.GRID_IN)}
        self.blocks = []
        for gin, gout in W.block_specs():
            f = t["sht_in"] if gin is W.GRID_IN else t["sht_int"]
            i = t["isht_out"] if gout is W.GRID_IN else t["isht_int"]
            w = O.init_spectral_weight(C, C, W.L_MODES, operator_type="dhconv").requires_grad_(True)
            self.blocks.append((f, i, w))
        dt = torch.bfloat16
        self.x0 = W.make_inputs(1, dt, pinned=False, chained=True, channels=C)[0].requires_grad_(True)
        self.grads = W.make_grads(1, dt, torch.device("cpu"), chained=True, channels=C)

    def step(self):
        """One full step; returns seconds."""
        O = self.O
        self.x0.grad = None
        t0 = time.perf_counter()
        x, outs, gouts = self.x0, [], []
        for (f, i, w), (gy, gr) in zip(self.blocks, self.grads):
            w.grad = None
            y, res = O.spectral_conv_forward(x, f, i, w, "dhconv", compute_dtype=torch.float32)
            if gr is not None:
                outs.append(res)
                gouts.append(gr)
            x = y
        outs.append(x)
        gouts.append(self.grads[-1][0])
        torch.autograd.backward(outs, gouts)
        return time.perf_counter() - t0


    def sample(self):
        """Bounded sample for the `cpu_baseline` leg of the GPU arm: blocks 0, 1 and 7 (the three block kinds), each
        forward + backward once on its own synthetic input.  Returns {kind: seconds}; a whole step is
        t[0] + 6 * t[1] + t[7] (six identical internal blocks)."""
        O, dt = self.O, torch.bfloat16
        xs = W.make_inputs(1, dt, pinned=False, chained=False, channels=self.C)
        gs = W.make_grads(1, dt, torch.device("cpu"), chained=False, channels=self.C)
        out = {}
        for idx in (0, 1, 7):
            f, i, w = self.blocks[idx]
            x = xs[idx].requires_grad_(True)
            gy, gr = gs[idx]
            w.grad = None
            t0 = time.perf_counter()
            y, res = O.spectral_conv_forward(x, f, i, w, "dhconv", compute_dtype=torch.float32)
            if gr is not None:
                torch.autograd.backward([y, res], [gy, gr])
            else:
                torch.autograd.backward([y], [gy])
            out[idx] = time.perf_counter() - t0
        return out


def run(steps, warmup, budget_s=150.0, channels=None):
    """Run up to `steps` timed steps (after `warmup` untimed ones) within about `budget_s` seconds of wall clock.
    Returns (seconds per step [list], steps run, warm-up steps run, threads)."""
    t_start = time.perf_counter()
    cs = CpuStep(channels=channels)
    first = cs.step()                                   # always one untimed step (allocator, table pages, thread pool)
    done_warm = 1
    per = max(first, 1e-3)
    while done_warm < warmup and (time.perf_counter() - t_start) + 3 * per < budget_s:
        per = cs.step()
        done_warm += 1
    times = []
    while len(times) < steps and (not times or (time.perf_counter() - t_start) + per < budget_s):
        per = cs.step()
        times.append(per)
    return times, len(times), done_warm, cs.threads
