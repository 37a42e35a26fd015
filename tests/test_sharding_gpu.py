"""Month-sharded step on the GPU: two processes must reproduce the single-process losses, weights and gradients.
On one GPU the two ranks share cuda:0 and gloo carries the all-reduces; with two GPUs or more the same check
runs over NCCL, one rank per GPU, eagerly AND as a replay of the captured CUDA graph (GraphedStep, the NCCL
all-reduces inside the graph) -- the path bench.py times at N > 1."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

pytestmark = pytest.mark.gpu


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _panel(T, N, M):
    g = torch.Generator().manual_seed(21)
    mask = torch.rand(T, N, generator=g) < 0.4
    x = torch.randn(T, N, 46, generator=g).clamp_(-2, 2) * mask[:, :, None]
    R = 0.1 * torch.randn(T, N, generator=g) * mask
    macro = torch.randn(T, M, generator=g)
    return macro, x, R, mask


def _shape(layout):
    # "wide": so many stocks that the per-stock sums come in ONE month chunk (plan.col_splits == 1), which is what
    # the eight-GPU weak-scaling shards of the bench have (the exchange then reduces the chunk in place)
    return (6, 76100, 8) if layout == "round_robin_wide" else (22, 700, 8)


def _stocks(rank, N):
    """Stock numbers of a rank (layout "stocks"): two uneven slices."""
    cut = (3 * N) // 7
    return slice(0, cut) if rank == 0 else slice(cut, N)


def _months(rank, T, layout):
    """Month numbers of a rank: two uneven slabs, or every second month."""
    if layout == "slabs":
        return list(range(0, 9)) if rank == 0 else list(range(9, T))
    return list(range(rank, T, 2))


def _worker(rank, world, port, precision, layout, out_q, backend="gloo"):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dev = torch.device("cuda", rank if backend == "nccl" else 0)
    torch.cuda.set_device(dev)
    if backend == "nccl":
        dist.init_process_group("nccl", rank=rank, world_size=world, device_id=dev)
    else:
        dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        from deeplearninginassetpricing_paperreplication_b200 import AssetPricingGAN
        from deeplearninginassetpricing_paperreplication_b200 import functional as Fn
        T, N, M = _shape(layout)
        macro, x, R, mask = _panel(T, N, M)
        torch.manual_seed(4)
        model = AssetPricingGAN({"macro_feature_dim": M, "individual_feature_dim": 46, "dropout": 0.0,
                                 "b200_precision": precision}).to(dev)
        if layout == "stocks":
            mine = _stocks(rank, N)
            # NCCL, one rank per GPU: both forward exchanges run inside the reduction kernel over peer memory
            model.month_shard = (Fn.make_stock_shard(None, N, T, dev) if backend == "nccl" else Fn.StockShard(None, N))
            if backend == "nccl":
                assert model.month_shard.peers is not None
            panel = (macro.to(dev), x[:, mine].contiguous().to(dev), R[:, mine].contiguous().to(dev),
                     mask[:, mine].contiguous().to(dev))
        else:
            mine = _months(rank, T, layout)
            model.month_shard = Fn.make_month_shard(mask[mine].to(dev), mine[0], T, None,
                                                    stride=1 if layout == "slabs" else 2)
            panel = (macro.to(dev), x[mine].contiguous().to(dev), R[mine].contiguous().to(dev),
                     mask[mine].contiguous().to(dev))
        out = model(*panel, phase="conditional")
        out["loss"].backward()
        torch.cuda.synchronize()
        if backend == "nccl":          # the same step as a CUDA-graph replay must say the same, to the bit
            eager = {k: p.grad.clone() for k, p in model.named_parameters()}
            e_loss, e_w = out["loss"].detach().clone(), out["weights"].clone()
            del out             # an earlier graph that is still alive pins the parameters' AccumulateGrad stream
            model.eval()
            gs = model.graphed_step(*panel, phase="conditional")
            out = gs.step(); out = gs.step()
            torch.cuda.synchronize()
            assert torch.equal(out["loss"], e_loss) and torch.equal(out["weights"], e_w)
            for k, p in model.named_parameters():
                assert torch.equal(p.grad, eager[k]), k
        res = {"loss": float(out["loss"].detach()), "loss_u": float(out["loss_unconditional"].detach()),
               "sharpe": float(out["sharpe"]),
               "pr": out["portfolio_returns"].detach().cpu().numpy(), "w": out["weights"].detach().cpu().numpy(),
               "grads": {k: p.grad.cpu().numpy() for k, p in model.named_parameters()}}
        out_q.put((rank, res))
        if backend == "nccl" and layout == "stocks":      # tearing the group down hangs with symmetric memory alive
            out_q.close(); out_q.join_thread()
            torch.cuda.synchronize()
            os._exit(0)
    except Exception as exc:  # pragma: no cover
        import traceback
        out_q.put((rank, "ERR " + traceback.format_exc()))
    finally:
        dist.destroy_process_group()


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs two GPUs (NCCL, one rank per GPU)")
@pytest.mark.parametrize("precision,layout", [("bf16x3", "round_robin"), ("bf16x3", "slabs"), ("fp32", "round_robin"),
                                              ("bf16x3", "round_robin_wide"), ("bf16x3", "stocks")])
def test_two_month_shards_nccl_eager_and_graph(precision, layout):
    test_two_month_shards_match_single_process(precision, layout, backend="nccl")


@pytest.mark.parametrize("precision,layout", [("fp32", "slabs"), ("bf16x3", "slabs"), ("fp32", "round_robin"),
                                              ("bf16x3", "round_robin"), ("bf16x3", "round_robin_wide"),
                                              ("fp32", "stocks"), ("bf16x3", "stocks")])
def test_two_month_shards_match_single_process(precision, layout, backend="gloo"):
    from deeplearninginassetpricing_paperreplication_b200 import AssetPricingGAN
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, precision, layout, q, backend)) for r in range(2)]
    for p in procs:
        p.start()
    res = dict(q.get(timeout=120) for _ in procs)
    for p in procs:
        p.join(timeout=20)
        if p.is_alive():            # a rank that failed may sit in the process group's teardown
            p.terminate()
    assert all(not isinstance(v, str) for v in res.values()), res
    dev = torch.device("cuda:0")
    T, N, M = _shape(layout)
    macro, x, R, mask = _panel(T, N, M)
    torch.manual_seed(4)
    model = AssetPricingGAN({"macro_feature_dim": M, "individual_feature_dim": 46, "dropout": 0.0,
                             "b200_precision": precision}).to(dev)
    ref = model(macro.to(dev), x.to(dev), R.to(dev), mask.to(dev), phase="conditional")
    ref["loss"].backward()
    for r in (0, 1):
        assert abs(res[r]["loss"] - float(ref["loss"])) <= 1e-5 * abs(float(ref["loss"]))
        assert abs(res[r]["loss_u"] - float(ref["loss_unconditional"])) <= 1e-5 * abs(float(ref["loss_unconditional"]))
        assert np.allclose(res[r]["pr"], ref["portfolio_returns"].detach().cpu().numpy(), rtol=1e-4, atol=1e-7)
        assert abs(res[r]["sharpe"] - float(ref["sharpe"])) <= 1e-4 * abs(float(ref["sharpe"])) + 1e-6
    wref = ref["weights"].detach().cpu().numpy()
    w = np.empty_like(wref)
    for r in (0, 1):
        if layout == "stocks":
            w[:, _stocks(r, N)] = res[r]["w"]
        else:
            w[_months(r, T, layout)] = res[r]["w"]
    assert np.abs(w - wref).max() <= 1e-4 * np.abs(wref).max()
    tol = 2e-4 if precision == "fp32" else 2e-2
    for k, p in model.named_parameters():
        if k == "sdf_net.output_proj.bias":
            continue
        gr = p.grad.cpu().numpy()
        for r in (0, 1):                      # every rank ends with the full gradient
            err = np.abs(res[r]["grads"][k] - gr).max() / max(np.abs(gr).max(), 1e-30)
            assert err < tol, (k, r, err)
