#!/usr/bin/env python
"""configs[3] on N GPUs: `torchrun --nproc-per-node N scripts/ensemble_gpu.py --members 9 [--train-epochs E]`.
Nine seeds of the model (random-init, or trained for a few epochs with training.train_epoch when --train-epochs > 0),
one model per GPU at a time, weights averaged with one all-reduce per split; prints one JSON line on rank 0.
Panel: the paper shape, synthetic (bench.synthetic_panel), train / valid / test = 240 / 60 / 300 months."""
import argparse
import json
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402
import torch.distributed as dist  # noqa: E402

import bench  # noqa: E402
from deeplearninginassetpricing_paperreplication_b200 import AssetPricingGAN, FusedClipAdam, train_epoch  # noqa: E402
from deeplearninginassetpricing_paperreplication_b200.ensemble import ensemble_weights  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--members", type=int, default=9)
ap.add_argument("--train-epochs", type=int, default=0)
ap.add_argument("--stocks", type=int, default=10_000)
args = ap.parse_args()
world, rank = int(os.environ.get("WORLD_SIZE", 1)), int(os.environ.get("RANK", 0))
dev = torch.device("cuda", int(os.environ.get("LOCAL_RANK", 0)))
torch.cuda.set_device(dev)
group = None
if world > 1:
    os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
    dist.init_process_group("nccl", device_id=dev)
    group = dist.group.WORLD
cfg = {"macro_feature_dim": 178, "individual_feature_dim": 46, "dropout": 0.05}
panels = {}
for name, T, seed in (("train", 240, 1), ("valid", 60, 2), ("test", 300, 3)):
    macro, x, R, mask = bench.synthetic_panel(T, args.stocks, 46, 178, seed=seed, device=dev)
    panels[name] = {"macro_features": macro, "individual_features": x, "returns": R, "mask": mask}
members = []
t0 = time.perf_counter()
for i in range(args.members):
    torch.manual_seed(1000 + i)
    m = AssetPricingGAN(cfg).to(dev)
    if args.train_epochs > 0 and i % world == rank:
        opt = FusedClipAdam(m.sdf_net.parameters(), lr=1e-3)
        for _ in range(args.train_epochs):
            train_epoch(m, opt, panels["train"], dev, phase="unconditional", scope="sdf")
    members.append({k: v.detach().clone() for k, v in m.state_dict().items()})
if world > 1 and args.train_epochs > 0:          # every rank needs every member's parameters only to index them
    for i, sd in enumerate(members):
        for v in sd.values():
            dist.broadcast(v, src=i % world)
torch.cuda.synchronize()
t1 = time.perf_counter()
res = ensemble_weights(members, cfg, panels, dev, group)
torch.cuda.synchronize()
t2 = time.perf_counter()
if rank == 0:
    print(json.dumps({"members": args.members, "gpus": world, "train_s": round(t1 - t0, 3), "ensemble_eval_s": round(t2 - t1, 4),
                      "sharpe": {k: v["sharpe"] for k, v in res.items()},
                      "stock_months_per_s": args.members * sum(p["returns"].numel() for p in panels.values()) / (t2 - t1)}))
if world > 1:
    dist.destroy_process_group()
