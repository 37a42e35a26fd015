#!/usr/bin/env python
"""Benchmark of the adversarial-SDF step (BASELINE.json metric: stock-months/s per fwd+bwd GAN
step, and % of HBM roofline, next to the reference CPU path).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference]

One step = model.zero_grad(); out = model(panel, phase='conditional'); out['loss'].backward()
(src/train.py:77-84) on the paper-shape synthetic panel T=240, N=10,000, 46 characteristics,
178 macro series (BASELINE.json configs[1]); dropout disabled (the parity configuration).
Under torchrun (N > 1) the panel grows to 10,000*N stocks and is sharded BY MONTH: every rank owns
240/N months of all stocks, i.e. the same 2.4 M stock-months per GPU (weak scaling; the LSTM, replicated
on every rank, keeps its 240 steps); the value is all ranks' stock-months / max-over-ranks device time.
At N > 1 the line also carries BASELINE.json configs[4] (T=600, N=100,000, months dealt over the ranks) and
a parity check of the sharded, graph-replayed step against the unsharded one.

Prints ONE JSON line on rank 0.
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

import numpy as np  # noqa: E402
import torch  # noqa: E402

T_MONTHS, N_STOCKS, F_CHARS, M_MACRO = 240, 10_000, 46, 178
ALGO_BYTES_FWD = 46 * 4 + 4 + 1          # x + R + mask per stock-month
ALGO_BYTES_STEP = 2 * ALGO_BYTES_FWD     # forward + recomputing backward
BWD_FLOP_PER_CELL = 36.9e3               # recompute + dgrad + wgrad per visited stock-month
E2E_RESERVE_SMS = int(os.environ.get("DLAP_E2E_RESERVE", "8"))   # e2e leg: SMs left to the pull kernel of the next panel
REF_DIR = os.path.join(ROOT, "baseline", "_ref")                 # the unmodified reference, staged by build()
METRIC = "stock-months/sec per fwd+bwd GAN step"


def synthetic_panel(T, N, F, M, seed, valid=0.4, device="cpu"):
    """SURVEY.md §8d generator: Bernoulli(0.40) mask, N(0,1) characteristics clipped to [-2,2],
    returns N(0, 0.1^2), both zero-filled where masked (src/data_loader.py:61-65), z-scored macro."""
    g = torch.Generator(device=device).manual_seed(seed)
    mask = torch.rand(T, N, generator=g, device=device) < valid
    x = torch.randn(T, N, F, generator=g, device=device).clamp_(-2, 2)
    x *= mask[:, :, None]
    R = 0.1 * torch.randn(T, N, generator=g, device=device) * mask
    macro = torch.randn(T, M, generator=g, device=device)
    return macro, x, R, mask


def workload_config(world):
    """`config` of the JSON line; the reference arm prints the same dict."""
    T, N = T_MONTHS // world, N_STOCKS * world
    return {"workload": (f"paper-shape panel T={T_MONTHS} N={N} F={F_CHARS} M={M_MACRO} "
                         f"(BASELINE.json configs[1]" + ("" if world == 1 else f", stocks x{world}") +
                         f"); phase 'conditional' fwd+bwd, dropout 0, hidden [64,64], rnn [4], 8 moments; "
                         f"{T * N} stock-months per GPU on {world} GPU(s) (how the panel is dealt over the ranks: "
                         f"`how.parallelism`)"),
            "l2": "inputs (442 MB panel per GPU) larger than the 126 MB L2; no flush needed"}


class ClockSampler(threading.Thread):
    """SM clock + throttle reasons during the timed region (B200_PROFILING.md), read through NVML in-process
    (nvidia_ml_py) every 2 ms; falls back to polling `nvidia-smi` when NVML cannot be loaded.  Rank 0 only: eight
    ranks spawning nvidia-smi compete with the launches they are meant to observe."""
    Q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        super().__init__(daemon=True)
        self.index, self.rows, self._halt = index, [], threading.Event()
        self.nvml = self.handle = None
        try:
            import pynvml
            pynvml.nvmlInit()
            try:
                uuid = "GPU-" + str(torch.cuda.get_device_properties(index).uuid)
                self.handle = pynvml.nvmlDeviceGetHandleByUUID(uuid.encode() if hasattr(uuid, "encode") else uuid)
            except Exception:
                self.handle = pynvml.nvmlDeviceGetHandleByIndex(index)
            self.nvml = pynvml
        except Exception:
            self.nvml = None

    def _sample_nvml(self):
        n, h = self.nvml, self.handle
        sm = n.nvmlDeviceGetClockInfo(h, n.NVML_CLOCK_SM)
        mx = n.nvmlDeviceGetMaxClockInfo(h, n.NVML_CLOCK_SM)
        try:
            bits = n.nvmlDeviceGetCurrentClocksEventReasons(h)
        except Exception:
            bits = n.nvmlDeviceGetCurrentClocksThrottleReasons(h)
        flag = lambda name: "Active" if bits & getattr(n, name, 0) else "Not Active"
        return [str(sm), str(mx), flag("nvmlClocksThrottleReasonHwSlowdown"),
                flag("nvmlClocksThrottleReasonHwThermalSlowdown"), flag("nvmlClocksThrottleReasonSwThermalSlowdown"),
                flag("nvmlClocksThrottleReasonSwPowerCap")]

    def run(self):
        while not self._halt.is_set():
            try:
                if self.nvml is not None:
                    self.rows.append(self._sample_nvml())
                else:
                    out = subprocess.run(["nvidia-smi", f"--id={self.index}", f"--query-gpu={self.Q}",
                                          "--format=csv,noheader,nounits"], capture_output=True, text=True,
                                         timeout=5).stdout.strip()
                    if out:
                        self.rows.append([c.strip() for c in out.split(",")])
            except Exception:
                pass
            self._halt.wait(0.002 if self.nvml is not None else 0.02)

    def stop(self):
        self._halt.set()
        self.join(timeout=6)
        sm = sorted(int(r[0]) for r in self.rows if r and r[0].isdigit())
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = sorted({n for r in self.rows for n, v in zip(names, r[2:6]) if v.lower().startswith("active")})
        return {"sm_mhz": sm[len(sm) // 2] if sm else None,
                "sm_max_mhz": int(self.rows[0][1]) if self.rows and self.rows[0][1].isdigit() else None,
                "reasons": reasons, "samples": len(self.rows)}


def peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        d = json.load(open(p))
        return float(d["hbm_gbs"]), float(d.get("bf16_tflops_sustained", 1394.6)), "measured"
    return 6650.0, 1400.0, "fallback"


def ncu_traffic():
    p = os.path.join(ROOT, "profiles", "dominant_kernel.json")
    if os.path.exists(p):
        return json.load(open(p)).get("dram_bytes_per_launch")
    return None


# ---- the reference on the host cores -----------------------------------------------------------------------
def reference_step_fn(device="cpu"):
    """(step(), kind): one fwd+bwd of phase 'conditional' (src/train.py:77-84) on the paper-shape panel, dropout 0.
    `kind` "reference": the UNMODIFIED AssetPricingGAN of baseline/_ref (staged by __graft_entry__.build() from the
    reference checkout); "port": oracle/torch_port.py, its bit-identical restatement, where that is absent."""
    macro, x, R, mask = [t.to(device) for t in synthetic_panel(T_MONTHS, N_STOCKS, F_CHARS, M_MACRO, seed=1234)]
    if os.path.isdir(os.path.join(REF_DIR, "src")):
        sys.path.insert(0, REF_DIR)
        from src.model import AssetPricingGAN as RefGAN
        torch.manual_seed(42)
        model = RefGAN({"macro_feature_dim": M_MACRO, "individual_feature_dim": F_CHARS, "dropout": 0.0}).to(device)
        model.train()

        def step():
            model.zero_grad(set_to_none=True)
            out = model(macro, x, R, mask, phase="conditional")
            out["loss"].backward()
            return out["loss"]
        return step, "reference"
    from oracle import torch_port as tp
    params = {k: v.to(device) for k, v in tp.init_params(M_MACRO, F_CHARS, seed=42).items()}

    def step():
        out, _ = tp.step(params, macro, x, R, mask, phase="conditional")
        return out["loss"]
    return step, "port"


def cpu_step_rate(budget_s=20.0, min_steps=2, max_steps=5, warmup=1, steps=None):
    """The reference's step timed on ALL host cores (torchrun exports OMP_NUM_THREADS=1: undone here), on the full
    paper-shape panel (needs ~8 GB of host memory)."""
    torch.set_num_threads(os.cpu_count() or 1)
    step, kind = reference_step_fn("cpu")

    def one():
        t0 = time.perf_counter()
        step()
        return time.perf_counter() - t0

    first = [one() for _ in range(max(warmup, 1))][-1]
    if steps is None:
        steps = int(min(max_steps, max(min_steps, budget_s / max(first, 1e-3))))
    times = [one() for _ in range(steps)]
    mean = sum(times) / len(times)
    cells = T_MONTHS * N_STOCKS
    return {"rate_best": cells / min(times), "rate_mean": cells / mean, "ms_per_step": 1e3 * mean, "steps": steps,
            "cores": torch.get_num_threads(), "host_cpus": os.cpu_count(), "kind": kind,
            "sample": f"the full T={T_MONTHS} N={N_STOCKS} panel, {steps} timed steps after {max(warmup, 1)} warm-up"}


def run_reference(args, rank, world):
    if rank != 0:
        return
    r = cpu_step_rate(warmup=args.warmup, steps=args.steps)
    cfg = workload_config(world)
    line = {
        "impl": "reference", "metric": METRIC, "value": r["rate_mean"],
        "unit": "stock-months/s", "n_gpus": args.gpus, "steps": r["steps"], "warmup": max(args.warmup, 1),
        "ms_per_step": r["ms_per_step"], "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "f32", "data": "synthetic", "config": cfg,
        "reference_sample": ("one rank's share of the workload (2.4 M stock-months: the T=240 x N=10,000 panel) on the "
                             "host cores; the CPU path does not shard"),
        "cpu_baseline": {"value": r["rate_mean"], "unit": "stock-months/s", "cores": r["cores"], "kind": r["kind"],
                         "sample": r["sample"], "host_cpus": r["host_cpus"]},
        "e2e": {"value": r["rate_mean"], "unit": "stock-months/s", "h2d_bytes_per_step": 0,
                "d2h_bytes_per_step": 0},
    }
    print(json.dumps(line))


# ---- this repository's arm -----------------------------------------------------------------------------------
def run_ours(args, rank, world):
    from deeplearninginassetpricing_paperreplication_b200 import AssetPricingGAN
    from deeplearninginassetpricing_paperreplication_b200 import functional as Fn
    import torch.distributed as dist

    local = int(os.environ.get("LOCAL_RANK", 0))
    dev = torch.device("cuda", local)
    torch.cuda.set_device(dev)
    if T_MONTHS % world:
        raise SystemExit(f"--gpus must divide {T_MONTHS} months")
    T, N = T_MONTHS // world, N_STOCKS * world           # this rank's months of all N stocks
    T_total = T_MONTHS
    macro_all = torch.randn(T_total, M_MACRO, generator=torch.Generator().manual_seed(99))
    _, x, R, mask = synthetic_panel(T, N, F_CHARS, M_MACRO, seed=1234 + rank)

    def new_model(precision=args.precision, dropout=0.0):
        torch.manual_seed(42)
        return AssetPricingGAN({"macro_feature_dim": M_MACRO, "individual_feature_dim": F_CHARS,
                                "dropout": dropout, "b200_precision": precision}).to(dev)

    model = new_model()
    model.train()
    host = [t.pin_memory() for t in (macro_all, x, R, mask)]
    d_macro, d_x, d_R, d_mask = [t.to(dev) for t in host]
    if world > 1:
        # months dealt round-robin: rank r owns months r, r + world, ... and follows the LSTM recurrence from month 0
        model.month_shard = Fn.make_month_shard(d_mask, rank, T_total, dist.group.WORLD, stride=world)

    def step_of(m, inputs, phase="conditional"):
        def step():
            m.zero_grad(set_to_none=True)
            out = m(*inputs, phase=phase)
            out["loss"].backward()
            return out
        return step

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def timed(fn, k):
        """ms for k calls: CUDA events on the current stream between two barriers, max over ranks."""
        barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(k):
            fn()
        e1.record()
        barrier()
        ms = torch.tensor([e0.elapsed_time(e1)], device=dev)
        if world > 1:
            dist.all_reduce(ms, op=dist.ReduceOp.MAX)
        return float(ms)

    def graphed(m, inputs, phase="conditional", train=False):
        m.train(train)
        gs = m.graphed_step(*inputs, phase=phase)
        for _ in range(3):
            gs.step()
        return gs

    def kernel_ms(m, inputs, name="dlap_cell_bwd", k=20):
        """mean launch duration of one entry point: CUDA events on its stream around every launch of k eager steps"""
        Fn.kernel_timer = Fn.KernelTimer(name)
        timed(step_of(m, inputs), k)
        timer, Fn.kernel_timer = Fn.kernel_timer, None
        return timer.mean_ms()

    resident = (d_macro, d_x, d_R, d_mask)
    eager = step_of(model, resident)
    for _ in range(max(args.warmup, 3)):
        eager()
    k_ms = kernel_ms(model, resident)            # roofline leg: the dominant kernel
    # the timed steps: zero_grad + forward + backward of the resident panel.  Enqueuing the launches of a step from
    # Python costs about as much host time as they run, so the step is captured once in a CUDA graph (GraphedStep)
    # and replayed; --eager times the plain Python loop instead.
    if args.eager:
        run_step, launches_per_step = eager, None
    else:
        gs = graphed(model, resident)
        run_step, launches_per_step = gs.step, gs.launches
    for _ in range(max(args.warmup, 3)):
        run_step()
    sampler = ClockSampler(local) if rank == 0 else None
    if sampler is not None:
        sampler.start()
    l0 = sum(w.launches for w in Fn._workspaces.values())
    ms = timed(run_step, args.steps)
    launches = (launches_per_step * args.steps if launches_per_step is not None
                else sum(w.launches for w in Fn._workspaces.values()) - l0)
    clocks = sampler.stop() if sampler is not None else None
    model.train()
    cells = T * N * world
    valid_cells = int(mask.sum()) if args.precision != "fp32" else T * N
    value = cells * args.steps / (ms * 1e-3)
    ksteps = max(5, min(args.steps, 50))

    extra = {}
    if args.no_legs:
        pass
    elif world == 1:
        extra.update(single_gpu_legs(args, dev, model, new_model, resident, graphed, timed, kernel_ms, step_of, ksteps))
    else:
        extra.update(multi_gpu_legs(args, dev, rank, world, new_model, graphed, timed, ksteps))

    # end to end: every step's panel comes from pinned host memory and its loss goes back to the host.
    # The way a data loader feeds a training loop: two device buffers, the copy of step k+1 runs on a
    # copy stream while step k computes; every step still pays for its own H2D inside the timed region.
    # The per-cell kernels leave E2E_RESERVE_SMS SMs to the pull kernel of the next panel (they take one CTA per
    # SM and all of its registers, so nothing else runs beside them otherwise).
    Fn.reserve_sms = E2E_RESERVE_SMS
    copy_stream = torch.cuda.Stream(dev)
    bufs = [[torch.empty_like(t, device=dev) for t in host] for _ in range(2)]
    ready = [torch.cuda.Event() for _ in range(2)]
    consumed = [torch.cuda.Event() for _ in range(2)]
    state = {"k": 0}

    def upload(slot):
        # the reference's per-epoch `batch[k].to(device)` (src/train.py:69-75): mask, returns and macro series by
        # DMA, the characteristic rows of unmasked stock-months pulled out of pinned host memory by a kernel
        with torch.cuda.stream(copy_stream):
            copy_stream.wait_event(consumed[slot])
            Fn.upload_panel(*host, device=dev, out=bufs[slot], clear=False)    # cleared by e2e_step
            ready[slot].record(copy_stream)

    def e2e_step():
        k = state["k"]; state["k"] = k + 1
        slot = k & 1
        torch.cuda.current_stream(dev).wait_event(ready[slot])
        upload(slot ^ 1)                                   # next step's panel, overlapped with this step
        out = step_of(model, bufs[slot])()
        # the loader's zero-fill of masked rows (src/data_loader.py:61-65) for the panel that will land in this
        # buffer next: done here, straight after the step, when the GPU has SMs to spare, not in front of the pull
        bufs[slot][1].zero_()
        consumed[slot].record(torch.cuda.current_stream(dev))
        return out["loss"].item()

    for b in bufs:
        b[1].zero_()
    for ev in consumed:
        ev.record(torch.cuda.current_stream(dev))
    upload(0)
    for _ in range(3):
        e2e_step()
    k2 = max(5, min(args.steps, 20))
    ms2 = timed(e2e_step, k2)
    torch.cuda.synchronize()
    # bytes that cross the bus per step: macro, returns, mask whole; of x only the rows of unmasked stock-months
    h2d = sum(t.numel() * t.element_size() for t in (host[0], host[2], host[3])) + int(mask.sum()) * F_CHARS * 4
    e2e = cells * k2 / (ms2 * 1e-3)
    Fn.reserve_sms = 0

    if rank != 0:
        return
    hbm_peak, tensor_peak, peak_kind = peaks()
    achieved = ALGO_BYTES_FWD * T * N / (k_ms * 1e-3) / 1e9
    tensor_tflops = 3 * BWD_FLOP_PER_CELL * valid_cells / (k_ms * 1e-3) / 1e12 if args.precision != "fp32" else None
    cfg = workload_config(world)            # identical in both arms
    how = {"precision": args.precision,
           "parallelism": "single GPU" if world == 1 else f"months dealt round-robin over {world} ranks (NCCL)",
           "step": ("python loop (eager launches)" if args.eager else
                    "zero_grad + forward + backward captured once in a CUDA graph (GraphedStep), replayed"),
           "panel": ("resident in HBM: x [T,N,F] fp32 as the reference holds it" + (
               "" if args.precision == "fp32" or not Fn.use_x_image else
               ", plus its X images (dlap_panel_image: the listed rows split into bf16 hi/lo in the tensor-core operand "
               "layout), made once per panel like the valid-cell index -- `panel_image.build_ms`; `gather_path` is the "
               "step without them"))}
    line = {
        "metric": METRIC, "value": value, "unit": "stock-months/s",
        "n_gpus": world, "steps": args.steps, "warmup": max(args.warmup, 3), "ms_per_step": ms / args.steps,
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "f32" if args.precision == "fp32" else "bf16x3 (fp32 split in hi+lo bf16, 3 MMAs per product, f32 accumulate)",
        "data": "synthetic", "config": cfg, "how": how, "clocks": clocks,
        "e2e": {"value": e2e, "unit": "stock-months/s", "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": 4,
                "ms_per_step": ms2 / k2,
                "note": "every step: upload_panel(pinned host panel) + fwd + bwd + loss.item() + zero-fill of the consumed buffer; macro, returns and "
                        "mask go by DMA, the characteristic rows of unmasked stock-months (40 %) are read out of "
                        "pinned host memory by dlap_panel_pull, masked rows never cross the bus; the pull of "
                        "panel k+1 runs on 8 SMs beside the step on panel k (the per-cell kernels use the other 140)"},
        "gpu_launches": launches,
        "roofline": {"bound": "tensor" if args.precision != "fp32" else "fp32 FFMA",
                     "kernel": "cell_bwd (recomputing per-cell backward, %s)" % args.precision,
                     "achieved": achieved, "peak": hbm_peak, "unit": "GB/s", "frac": achieved / hbm_peak,
                     "peak_kind": peak_kind, "kernel_ms": k_ms, "kernel_share_of_step": k_ms * args.steps / ms,
                     "traffic": ncu_traffic(),
                     "frac_visited_cells": achieved * valid_cells / float(T * N) / hbm_peak,
                     "step_frac_per_gpu": value / world * ALGO_BYTES_STEP / 1e9 / hbm_peak,
                     "tensor_tflops": tensor_tflops, "tensor_peak_tflops": tensor_peak,
                     "tensor_frac": tensor_tflops / tensor_peak if tensor_tflops else None,
                     "valid_fraction": valid_cells / float(T * N),
                     "note": "achieved/peak/frac: 189 algorithmic bytes per stock-month of the DENSE panel per launch (x, R, "
                             "mask read once, SURVEY.md 8d) over the measured HBM copy peak, as the metric asks; the bf16x3 "
                             "kernels visit only the unmasked stock-months (valid_fraction; masked cells contribute exactly "
                             "0), frac_visited_cells counts those alone and `traffic` is the real DRAM traffic per launch (with X images the kernel reads 193.5 B per visited stock-month instead of gathering 184-byte rows). "
                             "What binds is the tensor pipe (138 FLOP/B): tensor_frac = issued bf16 MMA work (3 x 36.9 "
                             "kFLOP per visited stock-month) over the sustained dense bf16 peak of MEASURED_PEAKS.json; "
                             "step_frac_per_gpu = whole step at 378 B per stock-month, per GPU"},
    }
    line.update(extra)
    if world > 1 and "by_stock" in extra and extra["by_stock"].get("value", 0) > value:
        bs = extra["by_stock"]
        line["alt_sharding"] = {"parallelism": how["parallelism"], "value": value, "ms_per_step": ms / args.steps,
                                "gpu_launches": launches}
        line["value"], line["ms_per_step"] = bs["value"], bs["ms_per_step"]
        line["gpu_launches"] = bs["launches_per_step"] * args.steps
        how["parallelism"] = bs["parallelism"]
        line["roofline"]["step_frac_per_gpu"] = bs["value"] / world * ALGO_BYTES_STEP / 1e9 / hbm_peak
        line["roofline"]["kernel_share_of_step"] = k_ms / bs["ms_per_step"]
    if not args.no_cpu and world == 1:
        r = cpu_step_rate()
        line["cpu_baseline"] = {"value": r["rate_mean"], "unit": "stock-months/s", "cores": r["cores"],
                                "kind": r["kind"], "sample": r["sample"], "host_cpus": r["host_cpus"]}
    print(json.dumps(line))


def single_gpu_legs(args, dev, model, new_model, resident, graphed, timed, kernel_ms, step_of, k):
    """What SURVEY.md §8d asks for beside the headline: the other phases, the forward-only evaluation, dropout on,
    a fully listed panel (rate per VISITED stock-month), the exact-fp32 path, and the eager-torch step on the same
    GPU (the reference moved to `cuda`, src/train.py:193-199)."""
    from deeplearninginassetpricing_paperreplication_b200 import functional as Fn
    cells = T_MONTHS * N_STOCKS
    rate = lambda ms_total, n: {"value": cells * n / (ms_total * 1e-3), "unit": "stock-months/s", "ms_per_step": ms_total / n}
    out = {}
    legs = {}
    gs = graphed(model, resident, phase="unconditional")                      # phase 1: SDF net on L_u (src/train.py:238-246)
    legs["phase1_unconditional"] = rate(timed(gs.step, k), k)
    for p in model.sdf_net.parameters():                                      # phase 2: SDF net frozen (src/train.py:309-317)
        p.requires_grad = False
    gs = graphed(model, resident, phase="moment")
    legs["phase2_moment"] = rate(timed(gs.step, k), k)
    for p in model.sdf_net.parameters():
        p.requires_grad = True
    m_drop = new_model(dropout=0.05)                                          # the reference's train mode (src/train.py:66)
    gs = graphed(m_drop, resident, train=True)
    legs["phase3_dropout_0.05"] = rate(timed(gs.step, k), k)
    del gs, m_drop

    # evaluate() of src/train.py:106-153: get_weights(normalized=True) + forward(phase='conditional') under no_grad
    model.eval()

    def evaluate():
        with torch.no_grad():
            w, _ = model.get_weights(resident[0], resident[1], resident[3], normalized=True)
            o = model(*resident, phase="conditional")
        return w, o

    for _ in range(3):
        evaluate()
    legs["evaluate_eager"] = rate(timed(evaluate, k), k)
    side = torch.cuda.Stream(dev)
    side.wait_stream(torch.cuda.current_stream(dev))
    with torch.cuda.stream(side):
        evaluate()
    torch.cuda.current_stream(dev).wait_stream(side)
    torch.cuda.synchronize()
    try:
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g):
            evaluate()
        for _ in range(3):
            g.replay()
        legs["evaluate_graph"] = rate(timed(g.replay, k), k)
        out["evaluate_ms"] = legs["evaluate_graph"]["ms_per_step"]
        del g
    except Exception as exc:                                                  # pragma: no cover
        legs["evaluate_graph"] = {"error": repr(exc)[:200]}
        out["evaluate_ms"] = legs["evaluate_eager"]["ms_per_step"]
    model.train()

    # every stock-month listed (mask all ones): the per-cell kernels' rate per visited stock-month
    dmacro, dx, dR, dmask = [t.to(dev) for t in synthetic_panel(T_MONTHS, N_STOCKS, F_CHARS, M_MACRO, seed=77, valid=1.1)]
    dense = (resident[0], dx, dR, dmask)
    for _ in range(3):
        step_of(model, dense)()
    dk_ms = kernel_ms(model, dense, k=10)
    gs = graphed(model, dense)
    leg = rate(timed(gs.step, k), k)
    hbm_peak, tensor_peak, _ = peaks()
    leg.update({"cell_bwd_ms": dk_ms, "cell_bwd_hbm_frac": ALGO_BYTES_FWD * cells / (dk_ms * 1e-3) / 1e9 / hbm_peak,
                "cell_bwd_tensor_frac": (3 * BWD_FLOP_PER_CELL * cells / (dk_ms * 1e-3) / 1e12 / tensor_peak
                                         if args.precision != "fp32" else None),
                "step_hbm_frac": leg["value"] * ALGO_BYTES_STEP / 1e9 / hbm_peak})
    out["dense_panel"] = leg
    model.train()
    del gs, dx, dR, dmask, dense

    if args.precision != "fp32":                                              # the exact-fp32 CUDA-core path
        m32 = new_model(precision="fp32")
        s32 = step_of(m32, resident)
        for _ in range(3):
            s32()
        k3 = max(3, min(k, 10))
        out["fp32_path"] = rate(timed(s32, k3), k3)
        del m32
    if args.precision != "fp32":
        # X images (dlap_panel_image): the resident panel is converted ONCE to the tensor-core operand format (like the
        # valid-cell index it is not part of a step).  What that pass costs, and the step WITHOUT it -- every step
        # gathering and converting the rows of x itself, as a loop that sees each panel once would run.
        try:
            ws0 = Fn.PlanOnly(T_MONTHS, N_STOCKS, F_CHARS, dev)
            idx0 = Fn.panel_index(ws0, resident[3])
            hit = Fn.cached_panel_image(resident[1])
            if hit is not None:
                img_ms = timed(lambda: Fn.rebuild_panel_image(resident[1], hit[0], hit[1]), 5) / 5
                out["panel_image"] = {"build_ms": img_ms, "bytes": int(hit[1].numel()),
                                      "bytes_per_listed_stock_month": hit[1].numel() / float(resident[3].sum().item()),
                                      "what": "dlap_panel_image: once per resident panel, outside the timed steps"}
            Fn.use_x_image = False
            mg = new_model()
            gg = graphed(mg, resident)
            out["gather_path"] = dict(rate(timed(gg.step, k), k), launches=gg.launches,
                                      what="the same step with DLAP_X_IMAGE=0: no X images, every step gathers the 184-byte "
                                           "rows of x and splits them to bf16 hi/lo itself (bit-identical results)")
            del gg, mg
        finally:
            Fn.use_x_image = True
    # a whole epoch of src/train.py:45-103 -- zero_grad, forward, backward, clip_grad_norm_, Adam -- as one graph replay
    # (device time), and through training.train_epoch with its one device-to-host read per epoch (host wall clock)
    from deeplearninginassetpricing_paperreplication_b200 import FusedClipAdam, train_epoch
    m_ep = new_model()
    opt = FusedClipAdam(m_ep.sdf_net.parameters(), lr=1e-3)
    gs = m_ep.train().graphed_step(*resident, phase="conditional", optimizer=opt, max_norm=1.0)
    for _ in range(3):
        gs.step()
    legs["epoch_graph_clip_adam"] = dict(rate(timed(gs.step, k), k), launches=gs.launches)
    data = {"macro_features": resident[0], "individual_features": resident[1], "returns": resident[2], "mask": resident[3]}
    for _ in range(3):
        train_epoch(m_ep, opt, data, dev, phase="conditional", scope="sdf")
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(k):
        train_epoch(m_ep, opt, data, dev, phase="conditional", scope="sdf")
    wall = (time.perf_counter() - t0) / k
    legs["train_epoch_wall"] = {"value": cells / wall, "unit": "stock-months/s", "ms_per_epoch": wall * 1e3,
                                "what": "training.train_epoch(): graph replay + one D2H copy of the epoch's numbers + sync"}
    out["epoch_ms"] = legs["epoch_graph_clip_adam"]["ms_per_step"]
    del gs, m_ep, opt
    out["legs"] = legs

    # the loader stage on the device (src/data_loader.py:42-66): raw [T,N,F+1] -> mask, R, x.
    # HBM-bound byte work: 4(F+1) B read + 4(F+1)+1 B written per stock-month; the raw panel (451 MB) exceeds the L2.
    try:
        from deeplearninginassetpricing_paperreplication_b200 import data as D
        raw = torch.randn(T_MONTHS, N_STOCKS, F_CHARS + 1, device=dev)
        raw[torch.rand(T_MONTHS, N_STOCKS, device=dev) < 0.6] = -99.99
        for _ in range(3):
            D.ingest_panel(raw, dev)
        ms = timed(lambda: D.ingest_panel(raw, dev), 10) / 10
        moved = cells * (8 * (F_CHARS + 1) + 1)
        raw_host = raw.cpu().pin_memory()
        D.ingest_panel(raw_host, dev)
        t0 = time.perf_counter()
        for _ in range(3):
            D.ingest_panel(raw_host, dev)                                     # synchronises: the kernel reads host memory
        host_ms = (time.perf_counter() - t0) / 3 * 1e3
        t0 = time.perf_counter()
        rn = raw_host.numpy()
        m_ref = (rn[:, :, 0] > -98.99) & ~np.isnan(rn[:, :, 0]) & np.all(rn[:, :, 1:] > -98.99, axis=2)
        np.where(m_ref, rn[:, :, 0], 0.0); np.where(m_ref[:, :, None], rn[:, :, 1:], 0.0)
        numpy_ms = (time.perf_counter() - t0) * 1e3
        out["loader"] = {"ingest_ms": ms, "GB/s": moved / (ms * 1e-3) / 1e9, "hbm_frac": moved / (ms * 1e-3) / 1e9 / peaks()[0],
                         "bytes_per_stock_month": 8 * (F_CHARS + 1) + 1,
                         "from_pinned_host_ms": host_ms, "from_pinned_host_GB/s": cells * 4 * (F_CHARS + 1) / (host_ms * 1e-3) / 1e9,
                         "numpy_same_arrays_ms": numpy_ms,
                         "what": "dlap_panel_ingest on the paper-shape raw array: mask from the -99.99 sentinel, zero-filled "
                                 "returns and characteristics, one launch; from_pinned_host: the same kernel reading the raw "
                                 "array over PCIe (upload + mask + zero-fill in one pass, wall clock); numpy_same_arrays: the "
                                 "loader's numpy expressions (src/data_loader.py:53-65) on this host"}
        del raw, raw_host, rn
        torch.cuda.empty_cache()
    except Exception as exc:                                                  # pragma: no cover
        out["loader"] = {"error": repr(exc)[:200]}

    # "the Blackwell kernel to beat" (SURVEY.md §8d): the reference's own eager torch ops on this GPU
    try:
        step, kind = reference_step_fn(dev)
        for _ in range(3):
            step()
        k4 = max(20, min(k, 30))
        ms4 = timed(step, k4)
        out["gpu_eager_baseline"] = dict(rate(ms4, k4), kind=kind,
                                         what="the reference's eager PyTorch step (cuDNN LSTM, cuBLAS GEMMs, ~490 ATen "
                                              "kernels, materialised concats) on the same B200, same panel, CUDA events")
        del step
        torch.cuda.empty_cache()
    except Exception as exc:                                                  # pragma: no cover
        out["gpu_eager_baseline"] = {"error": repr(exc)[:200]}
    return out


def multi_gpu_legs(args, dev, rank, world, new_model, graphed, timed, k):
    """N > 1: (1) the sharded, graph-replayed step must reproduce the unsharded one (small panel, every rank computes
    the unsharded answer itself); (2) BASELINE.json configs[4]: T=600, N=100,000, months dealt over the ranks."""
    from deeplearninginassetpricing_paperreplication_b200 import functional as Fn
    import torch.distributed as dist
    out = {}
    # (1) parity of the NCCL path
    Ts, Ns = 6 * world, 3000
    macro, x, R, mask = [t.to(dev) for t in synthetic_panel(Ts, Ns, F_CHARS, M_MACRO, seed=5)]
    full = new_model()
    o = full(macro, x, R, mask, phase="conditional")
    o["loss"].backward()
    mine = slice(rank, Ts, world)
    part = new_model()
    lm = mask[mine].contiguous()
    part.month_shard = Fn.make_month_shard(lm, rank, Ts, dist.group.WORLD, stride=world)
    gs = graphed(part, (macro, x[mine].contiguous(), R[mine].contiguous(), lm))
    so = gs.step()
    torch.cuda.synchronize()
    rel = lambda a, b: float((a - b).abs().max() / b.abs().max().clamp_min(1e-30))
    errs = torch.tensor([abs(float(so["loss"].detach()) - float(o["loss"].detach())) / abs(float(o["loss"].detach())),
                         rel(so["weights"], o["weights"][mine]),
                         rel(so["portfolio_returns"].detach(), o["portfolio_returns"].detach()),
                         max(float((p.grad - q.grad).norm() / q.grad.norm().clamp_min(1e-30))
                             for (n, p), q in zip(part.named_parameters(), full.parameters())
                             if n != "sdf_net.output_proj.bias")], device=dev, dtype=torch.float64)
    dist.all_reduce(errs, op=dist.ReduceOp.MAX)
    e = [float(v) for v in errs]
    out["sharded_check"] = {"panel": f"T={Ts} N={Ns}, months dealt round-robin over {world} ranks, CUDA-graph replay "
                                     "with the NCCL all-reduces inside, against the unsharded step on every rank",
                            "loss_rel": e[0], "weights_rel": e[1], "portfolio_returns_rel": e[2], "grad_norm_rel": e[3]}
    if not (e[0] < 1e-5 and e[1] < 1e-4 and e[2] < 1e-4 and e[3] < 2e-2):
        raise SystemExit(f"sharded step disagrees with the unsharded one: {out['sharded_check']}")
    del gs, part, o, so
    full.zero_grad(set_to_none=True)
    del full

    # (2) configs[4]
    T5, N5 = 600, 100_000
    if T5 % world == 0:
        Tl = T5 // world
        macro5 = torch.randn(T5, M_MACRO, generator=torch.Generator().manual_seed(98)).to(dev)
        _, x5, R5, m5 = synthetic_panel(Tl, N5, F_CHARS, M_MACRO, seed=4321 + rank, device=dev)
        big = new_model()
        big.month_shard = Fn.make_month_shard(m5, rank, T5, dist.group.WORLD, stride=world)
        gs = graphed(big, (macro5, x5, R5, m5))
        ms = timed(gs.step, k)
        out["scaled_panel"] = {"workload": f"BASELINE.json configs[4]: T={T5} N={N5} F={F_CHARS} M={M_MACRO}, months dealt "
                                           f"round-robin over {world} GPUs ({Tl} months x {N5} stocks = {Tl * N5} stock-months "
                                           "per GPU), phase 'conditional' fwd+bwd, CUDA-graph replay, NCCL inside; panel "
                                           "generated on the device",
                               "value": T5 * N5 * k / (ms * 1e-3), "unit": "stock-months/s", "ms_per_step": ms / k,
                               "scaling": "strong (fixed 60 M stock-months)", "loss": float(gs.out["loss"])}
        del gs, big, x5, R5, m5
        torch.cuda.empty_cache()

    # (3) the same two workloads sharded BY STOCK (SURVEY.md §8e's alternative): every rank keeps all months of its
    # own stocks; three all-reduces of a few kilobytes (per-month sums forward, dS + loss sums, gradients backward)
    bs = {}
    lo = rank * (Ns // world)
    hi = Ns if rank == world - 1 else lo + Ns // world
    full = new_model()
    o = full(macro, x, R, mask, phase="conditional")
    o["loss"].backward()
    part = new_model()
    part.month_shard = Fn.make_stock_shard(dist.group.WORLD, Ns, Ts, dev)
    gs = graphed(part, (macro, x[:, lo:hi].contiguous(), R[:, lo:hi].contiguous(), mask[:, lo:hi].contiguous()))
    so = gs.step()
    torch.cuda.synchronize()
    errs = torch.tensor([abs(float(so["loss"].detach()) - float(o["loss"].detach())) / abs(float(o["loss"].detach())),
                         rel(so["weights"], o["weights"][:, lo:hi]),
                         rel(so["portfolio_returns"].detach(), o["portfolio_returns"].detach()),
                         max(float((p.grad - q.grad).norm() / q.grad.norm().clamp_min(1e-30))
                             for (n, p), q in zip(part.named_parameters(), full.parameters())
                             if n != "sdf_net.output_proj.bias")], device=dev, dtype=torch.float64)
    dist.all_reduce(errs, op=dist.ReduceOp.MAX)
    e = [float(v) for v in errs]
    bs["sharded_check"] = {"loss_rel": e[0], "weights_rel": e[1], "portfolio_returns_rel": e[2], "grad_norm_rel": e[3]}
    if not (e[0] < 1e-5 and e[1] < 1e-4 and e[2] < 1e-4 and e[3] < 2e-2):
        raise SystemExit(f"stock-sharded step disagrees with the unsharded one: {bs['sharded_check']}")
    del gs, part, full, o, so
    # weak scaling: every rank the paper-shape panel of its own 10,000 stocks
    _, xs, Rs, ms_ = synthetic_panel(T_MONTHS, N_STOCKS, F_CHARS, M_MACRO, seed=1234 + rank, device=dev)
    macro_w = torch.randn(T_MONTHS, M_MACRO, generator=torch.Generator().manual_seed(99)).to(dev)
    wk = new_model()
    wk.month_shard = Fn.make_stock_shard(dist.group.WORLD, N_STOCKS * world, T_MONTHS, dev)
    gs = graphed(wk, (macro_w, xs, Rs, ms_))
    ms = timed(gs.step, args.steps)
    bs.update({"value": T_MONTHS * N_STOCKS * world * args.steps / (ms * 1e-3), "unit": "stock-months/s",
               "ms_per_step": ms / args.steps, "launches_per_step": gs.launches,
               "parallelism": (f"stocks dealt over {world} ranks (all months each); per-month sums and dS / loss sums "
                               "exchanged INSIDE the reduction kernel over NVLink peer memory (flag barrier + P2P loads), "
                               "gradients by one NCCL all-reduce; all inside the CUDA graph"
                               if wk.month_shard.peers is not None else
                               f"stocks dealt over {world} ranks (all months each), NCCL all-reduces of [T,4], [T,2]+2 "
                               "doubles and the gradients inside the CUDA graph")})
    del gs, wk, xs, Rs, ms_
    torch.cuda.empty_cache()
    if N5 % world == 0:
        macro5 = torch.randn(T5, M_MACRO, generator=torch.Generator().manual_seed(98)).to(dev)
        _, x5, R5, m5 = synthetic_panel(T5, N5 // world, F_CHARS, M_MACRO, seed=8765 + rank, device=dev)
        big = new_model()
        big.month_shard = Fn.make_stock_shard(dist.group.WORLD, N5, T5, dev)
        gs = graphed(big, (macro5, x5, R5, m5))
        ms = timed(gs.step, k)
        bs["scaled_panel"] = {"value": T5 * N5 * k / (ms * 1e-3), "unit": "stock-months/s", "ms_per_step": ms / k,
                              "workload": f"T={T5} N={N5}: {N5 // world} stocks x {T5} months per GPU"}
        del gs, big, x5, R5, m5
        torch.cuda.empty_cache()
    out["by_stock"] = bs
    return out


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=500)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--precision", default="bf16x3", choices=["bf16x3", "fp32"],
                    help="arithmetic of the per-cell contractions: bf16x3 = tcgen05 BF16 MMAs on hi/lo-split "
                         "fp32 operands (3 products per fp32 product, fp32 accumulate); fp32 = CUDA-core FFMA")
    ap.add_argument("--no-cpu", action="store_true", help="skip the cpu_baseline leg")
    ap.add_argument("--no-legs", action="store_true", help="headline, e2e and roofline only (development runs)")
    ap.add_argument("--eager", action="store_true", help="time the plain Python loop instead of the CUDA-graph replay")
    args = ap.parse_args()
    rank = int(os.environ.get("RANK", 0))
    world = int(os.environ.get("WORLD_SIZE", 1))
    if args.impl == "reference":
        run_reference(args, rank, world)
        return
    if world > 1:
        import torch.distributed as dist
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=torch.device("cuda", int(os.environ.get("LOCAL_RANK", 0))))
    run_ours(args, rank, world)
    if world > 1:
        # the process group is not torn down: with symmetric-memory buffers alive destroy_process_group() does not
        # return (seen on this image); every rank has finished its work and rank 0 has printed the line
        import torch.distributed as dist
        torch.cuda.synchronize()
        dist.barrier()
        sys.stdout.flush()
        sys.stderr.flush()
        os._exit(0)


if __name__ == "__main__":
    main()
