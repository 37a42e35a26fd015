#!/usr/bin/env python
"""Warm CUDA-event time of every C-ABI entry point inside the fwd+bwd step (paper shape, bf16x3 by default)."""
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402

import bench  # noqa: E402
from deeplearninginassetpricing_paperreplication_b200 import AssetPricingGAN  # noqa: E402
from deeplearninginassetpricing_paperreplication_b200 import functional as Fn  # noqa: E402

world, rank = int(os.environ.get("WORLD_SIZE", 1)), int(os.environ.get("RANK", 0))
dev = torch.device("cuda", int(os.environ.get("LOCAL_RANK", 0)))
torch.cuda.set_device(dev)
prec = sys.argv[1] if len(sys.argv) > 1 else "bf16x3"
T, N = 240 // world, 10000 * world                      # bench.py's weak-scaling shard: months split, stocks x world
macro = torch.randn(240, 178, generator=torch.Generator().manual_seed(99)).to(dev)
_, x, R, mask = [t.to(dev) for t in bench.synthetic_panel(T, N, 46, 178, seed=1234 + rank)]
torch.manual_seed(42)
model = AssetPricingGAN({"macro_feature_dim": 178, "individual_feature_dim": 46, "dropout": 0.0,
                         "b200_precision": prec}).to(dev)
if world > 1:
    import torch.distributed as dist
    os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
    dist.init_process_group("nccl", device_id=dev)
    model.month_shard = Fn.make_month_shard(mask, rank, 240, dist.group.WORLD, stride=world)
names = ["dlap_cell_fwd", "dlap_step_reduce", "dlap_cell_bwd", "dlap_macro_bwd"]
if world > 1:
    names += ["dlap_grad_finish", "exchange_fwd", "exchange_bwd"]
res = {}
for name in names:
    for it in range(23):
        if it == 3:
            Fn.kernel_timer = Fn.KernelTimer(name)
        model.zero_grad(set_to_none=True)
        model(macro, x, R, mask)["loss"].backward()
    torch.cuda.synchronize()
    res[name] = round(Fn.kernel_timer.mean_ms() * 1e3, 1)
    Fn.kernel_timer = None
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(50):
    model.zero_grad(set_to_none=True)
    model(macro, x, R, mask)["loss"].backward()
e1.record(); torch.cuda.synchronize()
import time  # noqa: E402
torch.cuda.synchronize()
t0 = time.perf_counter()
for _ in range(50):
    model.zero_grad(set_to_none=True)
    model(macro, x, R, mask)["loss"].backward()
host_us = (time.perf_counter() - t0) / 50 * 1e6          # time to ENQUEUE a step (no synchronisation)
torch.cuda.synchronize()
if world == 1:
    model.eval()
    gs = model.graphed_step(macro, x, R, mask)
    for _ in range(3):
        gs.step()
    g0, g1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    g0.record()
    for _ in range(100):
        gs.step()
    g1.record(); torch.cuda.synchronize()
    graph_us = g0.elapsed_time(g1) / 100 * 1e3
res["sum_us"] = round(sum(res.values()), 1)
if world == 1:
    res["graphed_step_us"] = round(graph_us, 1)
res["host_enqueue_us"] = round(host_us, 1)
res["step_us"] = round(e0.elapsed_time(e1) / 50 * 1e3, 1)
if rank == 0:
    print(json.dumps(res))
if world > 1:
    dist.destroy_process_group()
