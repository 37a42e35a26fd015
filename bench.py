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

import torch  # noqa: E402

T_MONTHS, N_STOCKS, F_CHARS, M_MACRO = 240, 10_000, 46, 178
ALGO_BYTES_FWD = 46 * 4 + 4 + 1          # x + R + mask per stock-month
ALGO_BYTES_STEP = 2 * ALGO_BYTES_FWD     # forward + recomputing backward
CPU_SAMPLE_N = 2_500                     # stocks in the CPU sample (cost is linear in T*N)
E2E_RESERVE_SMS = int(os.environ.get("DLAP_E2E_RESERVE", "8"))                      # e2e leg: SMs left to the pull kernel of the next panel


def synthetic_panel(T, N, F, M, seed, valid=0.4):
    """SURVEY.md §8d generator: Bernoulli(0.40) mask, N(0,1) characteristics clipped to [-2,2],
    returns N(0, 0.1^2), both zero-filled where masked (src/data_loader.py:61-65), z-scored macro."""
    g = torch.Generator().manual_seed(seed)
    mask = torch.rand(T, N, generator=g) < valid
    x = torch.randn(T, N, F, generator=g).clamp_(-2, 2) * mask[:, :, None]
    R = 0.1 * torch.randn(T, N, generator=g) * mask
    macro = torch.randn(T, M, generator=g)
    return macro, x, R, mask


class ClockSampler(threading.Thread):
    """SM clock + throttle reasons during the timed region (B200_PROFILING.md), read through NVML in-process
    (nvidia_ml_py); falls back to polling `nvidia-smi` when NVML cannot be loaded.  Rank 0 only: eight ranks
    spawning nvidia-smi every 20 ms compete with the launches they are meant to observe."""
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
            self._halt.wait(0.02)

    def stop(self):
        self._halt.set()
        self.join(timeout=6)
        sm = sorted(int(r[0]) for r in self.rows if r and r[0].isdigit())
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = sorted({n for r in self.rows for n, v in zip(names, r[2:6]) if v.lower().startswith("active")})
        return {"sm_mhz": sm[len(sm) // 2] if sm else None,
                "sm_max_mhz": int(self.rows[0][1]) if self.rows and self.rows[0][1].isdigit() else None,
                "reasons": reasons, "samples": len(self.rows)}


def hbm_peak():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        return float(json.load(open(p))["hbm_gbs"]), "measured"
    return 6650.0, "fallback"


def ncu_traffic():
    p = os.path.join(ROOT, "profiles", "dominant_kernel.json")
    if os.path.exists(p):
        return json.load(open(p)).get("dram_bytes_per_launch")
    return None


def cpu_step_rate(threads=None, budget_s=20.0, min_steps=2, max_steps=5, warmup=1, steps=None):
    """The oracle port (torch CPU restatement of src/model.py, autograd backward) timed on the host
    cores: stock-months/s of fwd+bwd, phase 'conditional', dropout off."""
    from oracle import torch_port as tp
    if threads:
        torch.set_num_threads(threads)
    macro, x, R, mask = synthetic_panel(T_MONTHS, CPU_SAMPLE_N, F_CHARS, M_MACRO, seed=1234)
    params = tp.init_params(M_MACRO, F_CHARS, seed=42)

    def one():
        t0 = time.perf_counter()
        tp.step(params, macro, x, R, mask, phase="conditional")
        return time.perf_counter() - t0

    first = [one() for _ in range(max(warmup, 1))][-1]
    if steps is None:
        steps = int(min(max_steps, max(min_steps, budget_s / max(first, 1e-3))))
    times = [one() for _ in range(steps)]
    best, mean = min(times), sum(times) / len(times)
    cells = T_MONTHS * CPU_SAMPLE_N
    return {"rate_best": cells / best, "rate_mean": cells / mean, "ms_per_step": 1e3 * mean, "steps": steps,
            "cores": torch.get_num_threads(), "host_cpus": os.cpu_count(),
            "sample": f"T={T_MONTHS} N={CPU_SAMPLE_N} slice of the workload, {steps} timed steps after "
                      f"{max(warmup, 1)} warm-up (cost is linear in T*N)"}


def run_reference(args, rank):
    if rank != 0:
        return
    r = cpu_step_rate(warmup=args.warmup, steps=args.steps)
    line = {
        "impl": "reference", "metric": "stock-months/sec per fwd+bwd GAN step", "value": r["rate_mean"],
        "unit": "stock-months/s", "n_gpus": args.gpus, "steps": r["steps"], "warmup": max(args.warmup, 1),
        "ms_per_step": r["ms_per_step"], "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "f32", "data": "synthetic",
        "config": {"workload": f"paper-shape panel T={T_MONTHS} N={N_STOCKS} F={F_CHARS} M={M_MACRO}, "
                               "phase 'conditional' fwd+bwd, dropout 0; timed on a bounded sample"},
        "cpu_baseline": {"value": r["rate_mean"], "unit": "stock-months/s", "cores": r["cores"], "kind": "port",
                         "sample": r["sample"]},
        "e2e": {"value": r["rate_mean"], "unit": "stock-months/s", "h2d_bytes_per_step": 0,
                "d2h_bytes_per_step": 0},
    }
    print(json.dumps(line))


def run_ours(args, rank, world):
    from deeplearninginassetpricing_paperreplication_b200 import AssetPricingGAN
    from deeplearninginassetpricing_paperreplication_b200 import functional as Fn
    import torch.distributed as dist

    local = int(os.environ.get("LOCAL_RANK", 0))
    dev = torch.device("cuda", local)
    torch.cuda.set_device(dev)
    if T_MONTHS % world:
        raise SystemExit(f"--gpus must divide {T_MONTHS} months")
    T, N = T_MONTHS // world, N_STOCKS * world           # this rank's slab: T months of all N stocks
    T_total = T_MONTHS
    macro_all = torch.randn(T_total, M_MACRO, generator=torch.Generator().manual_seed(99))
    _, x, R, mask = synthetic_panel(T, N, F_CHARS, M_MACRO, seed=1234 + rank)
    torch.manual_seed(42)
    model = AssetPricingGAN({"macro_feature_dim": M_MACRO, "individual_feature_dim": F_CHARS,
                             "dropout": 0.0, "b200_precision": args.precision}).to(dev)
    model.train()
    host = [t.pin_memory() for t in (macro_all, x, R, mask)]
    d_macro, d_x, d_R, d_mask = [t.to(dev) for t in host]
    if world > 1:
        # months dealt round-robin: rank r owns months r, r + world, ... and follows the LSTM recurrence from month 0
        model.month_shard = Fn.make_month_shard(d_mask, rank, T_total, dist.group.WORLD, stride=world)

    def step(inputs):
        model.zero_grad(set_to_none=True)
        out = model(*inputs, phase="conditional")
        out["loss"].backward()
        return out

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def timed(fn, k):
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

    resident = (d_macro, d_x, d_R, d_mask)
    for _ in range(max(args.warmup, 3)):
        step(resident)
    # roofline leg: the dominant kernel bracketed by CUDA events on its stream, over a short eager run
    Fn.kernel_timer = Fn.KernelTimer("dlap_cell_bwd")
    timed(lambda: step(resident), 20)
    timer, Fn.kernel_timer = Fn.kernel_timer, None
    # the timed steps: zero_grad + forward + backward of the resident panel.  Enqueuing the ~16 launches of a
    # step from Python costs about as much host time as they run, so the step is captured once in a CUDA
    # graph (GraphedStep) and replayed; --eager times the plain Python loop instead.
    if args.eager:
        run_step, launches_per_step = (lambda: step(resident)), None
    else:
        model.eval()                                   # dropout is 0 anyway; GraphedStep insists on it
        gs = model.graphed_step(*resident, phase="conditional")
        run_step, launches_per_step = gs.step, gs.launches
        for _ in range(3):
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
        out = step(bufs[slot])
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
    peak, peak_kind = hbm_peak()
    k_ms = timer.mean_ms()
    achieved = ALGO_BYTES_FWD * T * N / (k_ms * 1e-3) / 1e9
    line = {
        "metric": "stock-months/sec per fwd+bwd GAN step", "value": value, "unit": "stock-months/s",
        "n_gpus": world, "steps": args.steps, "warmup": max(args.warmup, 3), "ms_per_step": ms / args.steps,
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "f32" if args.precision == "fp32" else "bf16x3 (fp32 split in hi+lo bf16, 3 MMAs per product, f32 accumulate)",
        "data": "synthetic",
        "config": {"workload": (f"paper-shape panel T={T_total} N={N} F={F_CHARS} M={M_MACRO} "
                                f"(BASELINE.json configs[1]" + ("" if world == 1 else f", stocks x{world}") +
                                f"); phase 'conditional' fwd+bwd, dropout 0, hidden [64,64], rnn [4], 8 moments; "
                                f"sharded by month, {T} months x {N} stocks = {T * N} stock-months per GPU "
                                f"on {world} GPU(s)"),
                   "l2": "inputs (442 MB panel per GPU) larger than the 126 MB L2; no flush needed",
                   "precision": args.precision,
                   "step": ("python loop (eager launches)" if args.eager else
                            "zero_grad + forward + backward captured once in a CUDA graph (GraphedStep), replayed")},
        "clocks": clocks,
        "e2e": {"value": e2e, "unit": "stock-months/s", "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": 4,
                "ms_per_step": ms2 / k2,
                "note": "every step: upload_panel(pinned host panel) + fwd + bwd + loss.item() + zero-fill of the consumed buffer; macro, returns and "
                        "mask go by DMA, the characteristic rows of unmasked stock-months (40 %) are read out of "
                        "pinned host memory by dlap_panel_pull, masked rows never cross the bus; the pull of "
                        "panel k+1 runs on 8 SMs beside the step on panel k (the per-cell kernels use the other 140)"},
        "gpu_launches": launches,
        "roofline": {"bound": "hbm", "kernel": "cell_bwd (recomputing per-cell backward, %s)" % args.precision,
                     "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                     "peak_kind": peak_kind, "kernel_ms": k_ms, "kernel_share_of_step": k_ms * args.steps / ms,
                     "traffic": ncu_traffic(),
                     "step_frac": value * ALGO_BYTES_STEP / 1e9 / peak,
                     "tensor_tflops": (3 * 36.9e3 * valid_cells / (k_ms * 1e-3) / 1e12
                                       if args.precision != "fp32" else None),
                     "valid_fraction": valid_cells / float(T * N),
                     "note": "algorithmic bytes = 189 B per stock-month of the dense panel per launch (x, R, mask read "
                             "once, SURVEY.md 8d); the bf16x3 kernels visit only the unmasked stock-months "
                             "(valid_fraction; masked cells contribute exactly 0), so their real HBM traffic is lower "
                             "(`traffic`); step_frac = whole step at 378 B per stock-month; the kernel is bound by "
                             "tensor-core operand bandwidth + CUDA-core epilogue (138 FLOP/B), not by HBM; "
                             "tensor_tflops = issued bf16 MMA work (3 x 36.9 kFLOP per visited stock-month)"},
    }
    if args.precision != "fp32" and world == 1:       # the exact-fp32 CUDA-core path, for reference
        m32 = AssetPricingGAN({"macro_feature_dim": M_MACRO, "individual_feature_dim": F_CHARS,
                               "dropout": 0.0, "b200_precision": "fp32"}).to(dev)
        m32.load_state_dict(model.state_dict())

        def step32():
            m32.zero_grad(set_to_none=True)
            m32(*resident, phase="conditional")["loss"].backward()

        for _ in range(3):
            step32()
        k3 = max(3, min(args.steps, 10))
        ms3 = timed(step32, k3)
        line["fp32_path"] = {"value": T * N * k3 / (ms3 * 1e-3), "unit": "stock-months/s", "ms_per_step": ms3 / k3}
    if not args.no_cpu:
        r = cpu_step_rate()
        line["cpu_baseline"] = {"value": r["rate_mean"], "unit": "stock-months/s", "cores": r["cores"],
                                "kind": "port", "sample": r["sample"], "host_cpus": r["host_cpus"]}
    print(json.dumps(line))


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
    ap.add_argument("--eager", action="store_true", help="time the plain Python loop instead of the CUDA-graph replay")
    args = ap.parse_args()
    rank = int(os.environ.get("RANK", 0))
    world = int(os.environ.get("WORLD_SIZE", 1))
    if args.impl == "reference":
        run_reference(args, rank)
        return
    if world > 1:
        import torch.distributed as dist
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=torch.device("cuda", int(os.environ.get("LOCAL_RANK", 0))))
    try:
        run_ours(args, rank, world)
    finally:
        if world > 1:
            import torch.distributed as dist
            dist.destroy_process_group()


if __name__ == "__main__":
    main()
