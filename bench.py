#!/usr/bin/env python
"""bench.py - SFNO spectral hot path on B200 (contract: the task statement / DESIGN.md section 7 "Measurement").

Workload (BASELINE.json configs[1], `sfno_linear_73chq_sc3_layers8_edim384_asgl2`, batch 1 per GPU, bf16 AMP): the eight
SpectralConv blocks of the model, forward + backward (dX and dW), chained as they are inside the model - see
tools/benchlib/workload.py.  Spatial tensors are bf16 (what autocast hands the module), spectral arithmetic is TF32
tensor-core contractions with fp32 accumulation and fp32 FFTs (the reference runs this path in fp32 with TF32 allowed).

One JSON line on stdout (rank 0):
  value / ms_per_step   device-resident step, CUDA events, max over ranks
  e2e                   the same step fed from pinned host memory every step (+ checksum read back)
  roofline              dominant kernel family (live CUDA events) + `block` (one internal block forward against SURVEY
                        8(d)'s algorithmic bytes) + `step` + `traffic_ratio` (ncu DRAM bytes / algorithmic bytes)
  gpu_baseline          the reference-equivalent torch path (cuFFT + cuBLAS einsums, TF32 allowed) on the same GPU
  cpu_baseline          the oracle port on the host cores, bounded sample (N = 1 only)
  rollout               BASELINE config 5: 40 chained forward passes, eager and CUDA-graph replay (N = 1 only)
  packed_master         the same step with the opt-in packed master weights (no re-pack / unpack passes; N = 1 only)
  full_model            the whole SFNO network of the named config, fwd + bwd under bf16 autocast (N = 1 only)
  config3               BASELINE config 3: global batch 8 split over the N GPUs
  parity / spatial      N >= 2: h-sharded transforms + SpectralConv against the fp64 oracle (exit code != 0 on mismatch)
                        and BASELINE config 4 timings (tools/benchlib/spatial.py)
`--impl reference` runs the SAME step through the oracle port on the host cores (the reference itself cannot be
imported: torch-harmonics / modulus are absent) - whole steps, nothing extrapolated.
"""
import argparse
import json
import os
import subprocess
import sys
import tempfile
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tools"))

import torch  # noqa: E402

from benchlib import workload as W  # noqa: E402
from benchlib.workload import (C_EMBED, GRID_IN, GRID_INT, L_MODES, M_MODES, METRIC, WORKLOAD,  # noqa: E402,F401
                               block_specs, build_gpu, gpu_step, gpu_step_chained, make_grads, make_inputs)


class ClockSampler:
    """nvidia-smi clocks / throttle reasons during the timed region."""
    FIELDS = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
              "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
              "clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.f = tempfile.NamedTemporaryFile("w+", suffix=".csv", delete=False)
        try:
            self.p = subprocess.Popen(["nvidia-smi", f"--id={index}", f"--query-gpu={self.FIELDS}",
                                       "--format=csv,noheader,nounits", "-lms", "20"], stdout=self.f,
                                      stderr=subprocess.DEVNULL)
        except Exception:
            self.p = None

    def stop(self):
        out = {"sm_mhz": None, "sm_max_mhz": None, "reasons": []}
        if self.p is None:
            return out
        self.p.terminate()
        try:
            self.p.wait(timeout=5)
        except Exception:
            self.p.kill()
        self.f.flush()
        self.f.seek(0)
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for line in self.f.read().splitlines():
            parts = [p.strip() for p in line.split(",")]
            if len(parts) < 7:
                continue
            try:
                sm.append(float(parts[0]))
                mx.append(float(parts[1]))
            except ValueError:
                continue
            for n, v in zip(names, parts[3:7]):
                if v.lower().startswith("active"):
                    reasons.add(n)
        os.unlink(self.f.name)
        if sm:
            sm.sort()
            out.update(sm_mhz=sm[len(sm) // 2], sm_max_mhz=max(mx), reasons=sorted(reasons), samples=len(sm))
        return out


def ncu_traffic():
    """profiles/ncu_traffic.json: DRAM bytes per kernel family and for one internal block forward, from the committed
    ncu capture (tools/ncu_traffic.py); {} when absent."""
    try:
        with open(os.path.join(ROOT, "profiles", "ncu_traffic.json")) as f:
            return json.load(f)
    except Exception:
        return {}


def kernel_traffic(table, kernel, shape):
    key = kernel
    if kernel.startswith("dft_"):
        key = f"{kernel}_{shape.get('W')}"
    elif kernel.startswith("legendre_"):
        key = f"{kernel}_{shape.get('K')}"
    e = table.get(key)
    return float(e["bytes"]) if e else None


def measured_peaks():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            p = json.load(f)
        return float(p["hbm_gbs"]), "measured (MEASURED_PEAKS.json)"
    except Exception:
        return 6650.0, "fallback (B200_PROFILING.md)"


def base_config(world, batch):
    return {"workload": WORKLOAD, "per_gpu_batch": batch, "embed_dim": C_EMBED, "modes": [L_MODES, M_MODES],
            "io_dtype": "bf16", "accumulate": "fp32", "parallelism": f"dp{world}",
            "inputs": "chained: block i+1 reads block i's output, one backward pass from block 7 to the input",
            "includes": "fwd + dX + dW of all 8 blocks, weight repack every step; no optimizer update"}


# ------------------------------------------------------------------------------------------------ CPU arm
def run_reference(args):
    """The reference arm: the same chained step (same config keys) through the oracle port on all host cores.  Every
    step is a whole step; as many of the requested `--steps` as fit into about 150 s of wall clock are run and `steps`
    reports how many that was (a step takes tens of seconds on the host)."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    from benchlib import cpu_reference
    times, ran, warm, threads = cpu_reference.run(args.steps, args.warmup, budget_s=args.cpu_budget)
    t = sum(times) / len(times)
    value = 1.0 / t
    sample = (f"{ran} whole step(s) of the 8-block chained workload (block 0 at 721x1440, block 7 up-scaling), fwd+bwd, "
              f"torch CPU fp32 spectral arithmetic on bf16 fields, {threads} threads, after {warm} untimed step(s); "
              f"requested --steps {args.steps} --warmup {args.warmup}, bounded to ~{args.cpu_budget:.0f} s")
    cfg = base_config(args.gpus, 1)
    cfg.update({"host_threads": threads, "steps_requested": args.steps, "warmup_requested": args.warmup,
                "step_seconds": [round(x, 2) for x in times]})
    line = {
        "impl": "reference", "metric": METRIC, "value": value, "unit": "samples/s", "n_gpus": args.gpus,
        "steps": ran, "warmup": warm, "ms_per_step": t * 1e3, "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic", "config": cfg,
        "cpu_baseline": {"value": value, "unit": "samples/s", "cores": threads, "kind": "port", "sample": sample},
        "e2e": {"value": value, "unit": "samples/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line), flush=True)


def cpu_baseline_sample():
    """`cpu_baseline` of the GPU arm (rank 0, N = 1): the three block kinds of the step measured once each on the host."""
    from benchlib import cpu_reference
    cs = cpu_reference.CpuStep()
    t = cs.sample()
    step_s = t[0] + 6 * t[1] + t[7]
    return {"value": 1.0 / step_s, "unit": "samples/s", "cores": cs.threads, "kind": "port",
            "sample": (f"blocks 0, 1, 7 of the step (721x1440 -> 240x480, 240x480 internal, 240x480 -> 721x1440), each fwd+bwd "
                       f"once on the host: {t[0]:.1f} s, {t[1]:.1f} s, {t[7]:.1f} s; step = t0 + 6 t1 + t7 = {step_s:.1f} s "
                       "(six identical internal blocks); torch CPU fp32 (oracle port); the --impl reference arm runs "
                       "whole chained steps")}


# ------------------------------------------------------------------------------------------------ block roofline
def block_roofline(blocks, dev, peak, traffic):
    """SURVEY 8(d)'s unit: ONE internal SpectralConv block forward (240x480, C = 384, B = 1).  Timed through the module
    call with CUDA events over rotating inputs (4 x 88.5 MB bf16 + 283 MB of weights per call: nothing survives in
    the 126 MB L2 from one call to the next)."""
    blk = blocks[1]
    xs = [torch.rand(1, C_EMBED, GRID_INT[0], GRID_INT[1], device=dev).to(torch.bfloat16) for _ in range(4)]
    n = 24

    def timed(fn):
        for i in range(4):
            fn(i % 4)
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for i in range(n):
            fn(i % 4)
        e1.record()
        torch.cuda.synchronize()
        return e0.elapsed_time(e1) / n * 1e3

    with torch.no_grad():
        blk.invalidate()
        us = timed(lambda i: blk(xs[i]))                   # inference: packed weight cached
    xg = [x.clone().requires_grad_(True) for x in xs]
    us_pack = timed(lambda i: blk(xg[i]))                  # training forward: + weight re-pack (layout change)
    b16 = W.block_forward_bytes(GRID_INT, GRID_INT, 1, esize=2)
    b32 = W.block_forward_bytes(GRID_INT, GRID_INT, 1, esize=4)
    out = {"unit": "one internal SpectralConv block forward (240x480, C=384, B=1, bf16 I/O)", "us": us,
           "us_with_weight_repack": us_pack,
           "alg_bytes_bf16_io": b16, "alg_bytes_fp32_survey_8d": b32,
           "formula": "fields in + out (176.9 MB x b at 2 B, 353.9 MB x b at 4 B) + 394.2 MB (weight 283.1 + two tables 55.5)",
           "achieved_gbs": b16 / (us * 1e-6) / 1e9, "frac": b16 / (us * 1e-6) / 1e9 / peak,
           "frac_fp32_formula": b32 / (us * 1e-6) / 1e9 / peak,
           "frac_with_weight_repack": b16 / (us_pack * 1e-6) / 1e9 / peak, "target_frac": 0.70}
    t = traffic.get("block_internal_fwd")
    if t:
        out["dram_bytes_ncu"] = t["bytes"]
        out["traffic_ratio"] = t["bytes"] / b16
        out["traffic_source"] = t.get("source")
    return out


# ------------------------------------------------------------------------------------------------ main
def bind_to_gpu_numa_node(index):
    """N >= 2: run this rank (and first-touch its pinned input buffers) on the CPUs next to its GPU.  Every rank copies
    797 MB per step from pinned host memory; with all ranks' buffers wherever the processes happened to start, the
    end-to-end step at N = 8 was bound by one host's memory system (36 - 39 ms against 15.5 ms at N = 1).  Pure
    Linux: /sys/bus/pci/devices/<gpu>/local_cpulist + sched_setaffinity.  Returns a short description for `config`."""
    try:
        pr = torch.cuda.get_device_properties(index)
        bdf = f"{pr.pci_domain_id:04x}:{pr.pci_bus_id:02x}:{pr.pci_device_id:02x}.0"
        with open(f"/sys/bus/pci/devices/{bdf}/local_cpulist") as f:
            text = f.read().strip()
        cpus = set()
        for part in text.split(","):
            if "-" in part:
                a, b = part.split("-")
                cpus.update(range(int(a), int(b) + 1))
            elif part:
                cpus.add(int(part))
        allowed = os.sched_getaffinity(0)
        use = cpus & allowed
        if use and use != allowed:
            os.sched_setaffinity(0, use)
            return f"rank bound to {len(use)} of {len(allowed)} CPUs local to GPU {bdf}"
        return f"no narrower CPU set local to GPU {bdf} ({len(allowed)} CPUs allowed)"
    except Exception as e:                                  # noqa: BLE001
        return f"not bound ({type(e).__name__})"


def run_ours(args):
    import torch.distributed as dist
    from makani_b200 import _ext

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    dev = torch.device("cuda", local_rank)
    torch.cuda.set_device(dev)
    distributed = world > 1
    numa = bind_to_gpu_numa_node(local_rank) if distributed and os.environ.get("MAKANI_B200_BENCH_NUMA", "1") == "1" else None
    if distributed:
        dist.init_process_group("nccl", device_id=dev)

    dtype = torch.bfloat16
    batch = 1
    t_wall0 = time.perf_counter()

    def barrier():
        if distributed:
            dist.barrier()
        torch.cuda.synchronize()

    # ---- N >= 2: parity of the h-sharded path against the fp64 oracle BEFORE anything is timed (config 4 geometry)
    parity, parity_ok, h, polar = None, True, 1, None
    if distributed and not args.no_spatial:
        from benchlib import spatial
        import makani_b200.distributed as thd
        h, polar, replica = spatial.make_groups(world, rank)
        thd.init(polar, None)
        parity, parity_ok = spatial.parity(dev, rank, h, polar, replica)
        barrier()

    blocks = build_gpu(dev)

    # data-parallel replicas: the only exchange is the average of the weight gradients (DDP's job in the reference,
    # makani/mpu/mappings.py:104-172), started per block as soon as its gradient exists so that it overlaps the
    # remaining backward work: an NCCL all-reduce from post-accumulate hooks (default), or - MAKANI_B200_DP_EXCHANGE=peer -
    # the exchange over NVLink peer memory without any NCCL kernel (distributed/grad_exchange.py), whose result is
    # checked against the NCCL all-reduce first.  Measured equal within 2-3 % at 2 and 4 GPUs
    # (profiles/r2f_dp_gradient_exchange.txt).
    from makani_b200.distributed.grad_sync import GradientAverager
    grads = make_grads(batch, dtype, dev, chained=True)
    xs_host = make_inputs(batch, dtype, pinned=True, chained=True)
    x_dev = xs_host[0].to(dev).requires_grad_(True)
    sync, dp_exchange = None, "none (single GPU)"
    if distributed:
        from makani_b200.distributed import p2p
        sync = GradientAverager([blk.weight for blk in blocks])
        dp_exchange = "NCCL all-reduce (AVG) of view_as_real(weight.grad), from post-accumulate hooks"
        if os.environ.get("MAKANI_B200_DP_EXCHANGE", "nccl") == "peer" and p2p.available(dist.group.WORLD, dev):
            from makani_b200.distributed.grad_exchange import PeerGradientAverager
            # parity of the peer-memory exchange against the NCCL all-reduce on the same step (blocks 0 and 7)
            gpu_step_chained(blocks, x_dev, grads)
            sync.wait()
            want = [blocks[i].weight.grad.clone() for i in (0, 7)]
            sync.remove()
            sync = PeerGradientAverager(blocks)
            gpu_step_chained(blocks, x_dev, grads)
            sync.wait()
            torch.cuda.synchronize()
            err = max(float((blocks[i].weight.grad - w).abs().max() / w.abs().max()) for i, w in zip((0, 7), want))
            errt = torch.tensor([err], device=dev)
            dist.all_reduce(errt, op=dist.ReduceOp.MAX)
            del want
            dp_exchange = ("peer memory (CUDA IPC over NVLink): averaging kernel reads the peers' slices of the packed "
                           "gradient, all-gather folded into the unpack, on a communication stream; no NCCL kernel")
            if parity is None and rank == 0:
                parity = {}
            if parity is not None:
                parity["dp_grad_exchange_vs_nccl_max_rel"] = float(errt)
            if not float(errt) < 1e-5:
                parity_ok = False
                if parity is not None:
                    parity.setdefault("failed", []).append("dp_grad_exchange")

    def step(x=None, g=None):
        out = gpu_step_chained(blocks, x_dev if x is None else x, grads if g is None else g)
        if sync is not None:
            sync.wait()
        return out

    # ---- device-resident timing (`value`)
    for _ in range(max(args.warmup, 3)):
        step()
    barrier()
    launches0 = _ext.launch_count()
    sampler = ClockSampler(local_rank) if rank == 0 else None
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(args.steps):
        step()
    e1.record()
    barrier()
    ms = e0.elapsed_time(e1)
    clocks = sampler.stop() if sampler else None
    launches = _ext.launch_count() - launches0

    # ---- end-to-end timing: pinned host inputs -> device, step, checksum back to host, every step.
    # Two device input sets: while step i computes on one, the copy engine fills the other with the input of step
    # i+1 (a data loader's prefetch); every step's host->device copy and device->host read are inside the timed region.
    sets = [x_dev, x_dev.detach().clone().requires_grad_(True)]
    copy_stream = torch.cuda.Stream()
    ready = [torch.cuda.Event(), torch.cuda.Event()]
    consumed = [torch.cuda.Event(), torch.cuda.Event()]

    def enqueue_copy(k):
        with torch.cuda.stream(copy_stream):
            copy_stream.wait_event(consumed[k])
            sets[k].data.copy_(xs_host[0], non_blocking=True)
            ready[k].record(copy_stream)

    def e2e_step(i, last):
        k = i & 1
        if not last:                                    # hand the next copy to the copy engine before enqueueing this step
            enqueue_copy(k ^ 1)
        torch.cuda.current_stream().wait_event(ready[k])
        out = step(sets[k])
        consumed[k].record()
        return out.cpu()

    for k in range(2):
        consumed[k].record()
    enqueue_copy(0)
    e2e_step(0, True)
    barrier()
    t0, t1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    t0.record()
    enqueue_copy(0)
    for i in range(args.steps):
        chk = e2e_step(i, i == args.steps - 1)
    t1.record()
    barrier()
    ms_e2e = t0.elapsed_time(t1)
    del sets
    h2d = xs_host[0].numel() * xs_host[0].element_size()
    d2h = chk.numel() * chk.element_size()

    times = torch.tensor([ms, ms_e2e], dtype=torch.float64, device=dev)
    if distributed:
        dist.all_reduce(times, op=dist.ReduceOp.MAX)
    ms, ms_e2e = times.tolist()

    # ---- per-kernel numbers, measured live with CUDA events in a separate profiled pass.
    # Every rank runs the pass (the steps contain the gradient all-reduce); only rank 0 instruments and reports.
    rows = None
    if rank == 0:
        from benchlib.profiler import KernelProfiler
        plans = []
        for blk in blocks:
            plans += [blk.forward_transform._plan(dev), blk.inverse_transform._plan(dev)]
        with KernelProfiler(plans) as prof:
            for _ in range(3):
                step()
            rows = prof.summary()
    else:
        for _ in range(3):
            step()
    barrier()

    # ---- BASELINE config 3: global batch 8 over the N GPUs (local batch 8 / N), same step
    config3 = None
    if not args.no_config3 and 8 % world == 0:
        lb = 8 // world
        g3 = make_grads(lb, dtype, dev, chained=True)
        x3 = make_inputs(lb, dtype, pinned=False, chained=True)[0].to(dev).requires_grad_(True)
        for _ in range(2):
            step(x3, g3)
        barrier()
        a0, a1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        n3 = max(2, min(args.steps, 4))
        a0.record()
        for _ in range(n3):
            step(x3, g3)
        a1.record()
        barrier()
        t3 = torch.tensor([a0.elapsed_time(a1) / n3], dtype=torch.float64, device=dev)
        if distributed:
            dist.all_reduce(t3, op=dist.ReduceOp.MAX)
        config3 = {"global_batch": 8, "per_gpu_batch": lb, "ms_per_step": float(t3), "steps": n3,
                   "value": 8.0 / (float(t3) * 1e-3), "unit": "samples/s", "scaling": "strong (global batch fixed)"}
        del g3, x3
        x_dev.grad = None
        torch.cuda.empty_cache()

    # ---- BASELINE config 4: h-sharded SpectralConv blocks (N >= 2)
    spatial_out = None
    if distributed and not args.no_spatial:
        sync.remove()                                   # the single-GPU reference times below must not all-reduce
        spatial_out = spatial.timings(dev, rank, h, polar, blocks, steps=6, warmup=2)
        if world > h:
            spatial_out["layout"] = f"{world // h} data-parallel replicas x h = {h}"
        barrier()

    line = None
    if rank == 0:
        peak, peak_src = measured_peaks()
        traffic = ncu_traffic()
        step_ms_prof = sum(r["total_ms"] for r in rows) / 3
        kernels = []
        for r in rows[:16]:
            per_step = r["total_ms"] / 3
            kernels.append(dict(kernel=r["kernel"], shape=r["shape"], launches_per_step=r["launches"] // 3,
                                ms_per_step=round(per_step, 4), share=round(per_step / step_ms_prof, 3),
                                alg_MB_per_step=round(r["alg_bytes_total"] / 3 / 1e6, 1), gbs=round(r["gbs"], 1),
                                frac=round(r["gbs"] / peak, 3)))
        dom = rows[0]
        # one "launch" of the dominant family = one pass over a whole tensor (its channel windows together)
        passes = max(1, round(dom["channels_total"] / C_EMBED)) if dom["channels_total"] else dom["launches"]
        roofline = {"bound": "hbm", "kernel": dom["kernel"], "shape": dom["shape"], "achieved": dom["gbs"], "peak": peak,
                    "peak_source": peak_src, "unit": "GB/s", "frac": dom["gbs"] / peak,
                    "traffic": kernel_traffic(traffic, dom["kernel"], dom["shape"]),
                    "alg_bytes": dom["alg_bytes_total"] / passes, "avg_us": dom["total_ms"] / passes * 1e3,
                    "kernel_launches_per_pass": dom["launches"] / passes,
                    "share_of_step": dom["total_ms"] / 3 / step_ms_prof}
        sb = W.step_bytes(batch)
        roofline["step"] = {"alg_bytes": sb, "achieved_gbs": sb / (ms / args.steps * 1e-3) / 1e9,
                            "frac": sb / (ms / args.steps * 1e-3) / 1e9 / peak,
                            "definition": "sum over the 8 blocks of fields in/out/residual + weight (+ dW) + tables, fwd + bwd "
                                          "(tools/benchlib/workload.py step_bytes); weight re-pack / gradient unpack count as 0"}
        if world == 1:
            roofline["block"] = block_roofline(blocks, dev, peak, traffic)
            if "traffic_ratio" in roofline["block"]:
                roofline["traffic_ratio"] = roofline["block"]["traffic_ratio"]
    barrier()

    rollout = gpu_base = cpu = full_model = packed_master = None
    if rank == 0 and world == 1:
        if not args.no_rollout:
            from benchlib import rollout as R
            x_dev.grad = None
            rollout = R.measure(blocks, x_dev.detach(), steps=40)
        if not args.no_packed_master:
            # SpectralConv.enable_packed_master(): the trained parameter is the GEMM-layout tensor, so the step has no
            # weight re-pack and no gradient unpack (16 launches fewer); same chained step otherwise
            for blk in blocks:
                blk.enable_packed_master()

            def pstep():
                for blk in blocks:
                    blk.weight_packed.grad = None
                return gpu_step_chained(blocks, x_dev, grads)

            for _ in range(3):
                pstep()
            torch.cuda.synchronize()
            l0 = _ext.launch_count()
            p0, p1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            p0.record()
            for _ in range(args.steps):
                pstep()
            p1.record()
            torch.cuda.synchronize()
            pms = p0.elapsed_time(p1) / args.steps
            packed_master = {"ms_per_step": pms, "value": batch * 1e3 / pms, "unit": "samples/s", "steps": args.steps,
                             "launches_per_step": (_ext.launch_count() - l0) // args.steps,
                             "what": "the same chained step with SpectralConv.enable_packed_master() (opt-in): the trained "
                                     "parameter is the packed planar tensor the GEMMs read, no re-pack / unpack passes; the "
                                     "tensor cores truncate the fp32 master to TF32 instead of reading a rounded copy"}
        if not args.no_gpu_baseline:
            from benchlib import torch_gpu
            del grads
            torch.cuda.empty_cache()
            try:
                gpu_base = torch_gpu.measure(dev, steps=5, warmup=2)
                gpu_base["speedup_of_this_repo"] = gpu_base["ms_per_step"] / (ms / args.steps)
            except Exception as e:                          # noqa: BLE001  (an OOM here must not lose the bench line)
                gpu_base = {"error": f"{type(e).__name__}: {e}"[:300]}
        if not args.no_full_model:
            from benchlib import sfno_model
            torch.cuda.empty_cache()
            try:
                full_model = sfno_model.measure(dev, steps=5, warmup=2)
                full_model["workload"] = (
                    "whole SFNO network of the named config (73 -> 384 encoder, 8 blocks with instance norms, MLPs and skips, "
                    "decoder, big skip; 721x1440, batch 1), forward + backward under bf16 autocast, random weights; restated "
                    "from sfnonet.py with the reference's parameter names (tools/benchlib/sfno_model.py - the reference file "
                    "itself cannot travel to this box).  reference_equivalent_torch: all torch (cuDNN convolutions, torch norms / GELU, "
                    "cuFFT + cuBLAS spectral path with TF32 allowed) = what the reference runs on this GPU; torch_glue: torch convolutions / norms / GELU around this library's "
                    "spectral path; fused_norm: + fused instance norms; fused: + 1x1-convolution stacks as GEMMs on NCHW "
                    "(library GEMMs) with this library's bias / GELU kernels; fused_packed_master: + SpectralConv.enable_packed_master()")
                full_model["spectral_path_share"] = (ms / args.steps) / full_model["fused"]["ms_per_step"]
                ref_ms = full_model.get("reference_equivalent_torch", {}).get("ms_per_step")
                if ref_ms:
                    full_model["speedup_over_reference_equivalent"] = {
                        "fused": ref_ms / full_model["fused"]["ms_per_step"],
                        "fused_packed_master": ref_ms / full_model["fused_packed_master"]["ms_per_step"]}
            except Exception as e:                          # noqa: BLE001
                full_model = {"error": f"{type(e).__name__}: {e}"[:300]}
        if not args.no_cpu_baseline:
            cpu = cpu_baseline_sample()

    if rank == 0:
        samples = batch * world * args.steps
        cfg = base_config(world, batch)
        cfg.update({"l2": f"inputs ({h2d / 1e9:.1f} GB/step), weights (2.27 GB) and intermediates (XM / SP 178-536 MB per kernel) "
                          "are far larger than the 126 MB L2",
                    "window_mb": float(os.environ.get("MAKANI_B200_WINDOW_MB", "0")), "dp_gradient_exchange": dp_exchange,
                    "host_affinity": numa,
                    "e2e": "pinned-host input (block 0's field) copied every step on a copy stream, double buffered "
                           "against compute; the step's results (output field, input gradient, weight gradients) stay on "
                           "the device as in a training step - a 16-byte checksum of output and input gradient is read back",
                    "wall_s": round(time.perf_counter() - t_wall0, 1)})
        line = {
            "metric": METRIC, "value": samples / (ms * 1e-3), "unit": "samples/s", "n_gpus": world, "steps": args.steps,
            "warmup": max(args.warmup, 3), "ms_per_step": ms / args.steps, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "tf32", "data": "synthetic", "config": cfg,
            "e2e": {"value": samples / (ms_e2e * 1e-3), "unit": "samples/s", "h2d_bytes_per_step": h2d,
                    "d2h_bytes_per_step": d2h, "ms_per_step": ms_e2e / args.steps},
            "gpu_launches": launches, "clocks": clocks, "roofline": roofline, "cpu_baseline": cpu,
            "gpu_baseline": gpu_base, "rollout": rollout, "packed_master": packed_master, "full_model": full_model, "config3": config3, "parity": parity, "spatial": spatial_out,
            "kernels": kernels,
        }
    if distributed:
        dist.barrier()
        dist.destroy_process_group()
    if line is not None:
        print(json.dumps(line), flush=True)
    if not parity_ok:
        print("bench: PARITY FAILURE of the h-sharded path against the fp64 oracle (see `parity`)", file=sys.stderr)
        sys.exit(3)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--cpu-budget", type=float, default=150.0, help="--impl reference: wall-clock bound in seconds")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-gpu-baseline", action="store_true")
    ap.add_argument("--no-rollout", action="store_true")
    ap.add_argument("--no-config3", action="store_true")
    ap.add_argument("--no-full-model", action="store_true")
    ap.add_argument("--no-packed-master", action="store_true")
    ap.add_argument("--no-spatial", action="store_true")
    args = ap.parse_args()
    if args.impl == "reference":
        run_reference(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()
