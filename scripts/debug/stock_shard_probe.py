import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import torch, torch.distributed as dist
from deeplearninginassetpricing_paperreplication_b200 import AssetPricingGAN
from deeplearninginassetpricing_paperreplication_b200 import functional as Fn
os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT="29533")
dist.init_process_group("gloo", rank=0, world_size=1)
dev = torch.device("cuda:0")
g = torch.Generator().manual_seed(21)
T, N, M = 22, 700, 8
mask = torch.rand(T, N, generator=g) < 0.4
x = torch.randn(T, N, 46, generator=g).clamp_(-2, 2) * mask[:, :, None]
R = 0.1 * torch.randn(T, N, generator=g) * mask
macro = torch.randn(T, M, generator=g)
for prec in ("fp32", "bf16x3"):
    torch.manual_seed(4)
    a = AssetPricingGAN({"macro_feature_dim": M, "individual_feature_dim": 46, "dropout": 0.0, "b200_precision": prec}).to(dev)
    b = AssetPricingGAN({"macro_feature_dim": M, "individual_feature_dim": 46, "dropout": 0.0, "b200_precision": prec}).to(dev)
    b.load_state_dict(a.state_dict())
    b.month_shard = Fn.StockShard(None, N)
    args = (macro.to(dev), x.to(dev), R.to(dev), mask.to(dev))
    oa = a(*args); ob = b(*args)
    for k in ("loss", "loss_unconditional", "sharpe"):
        print(prec, k, float(oa[k]), float(ob[k]))
    print(prec, "pr diff", float((oa["portfolio_returns"] - ob["portfolio_returns"]).abs().max()),
          "w diff", float((oa["weights"] - ob["weights"]).abs().max()))
dist.destroy_process_group()
