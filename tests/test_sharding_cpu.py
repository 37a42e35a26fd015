"""Month-sharding algebra and its host-side plumbing on CPU: two gloo ranks, each owning half of the
months.  The per-shard arithmetic is the fp64 oracle; what is under test is that the two exchanges of
DESIGN.md §5 (per-stock column sums forward; da_t|dc_t rows + per-cell weight gradients backward, N_bar
once) reproduce the unsharded result -- with exchange 1 laid out and rounded as the kernels do it: ONE fp32 buffer,
the [10,N] per-stock sums followed by the series' portfolio returns (own months filled in, zeros elsewhere) -- and
that functional.make_month_shard moves the right numbers."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from _dlap_testlib import golden_params, load_golden
from oracle import closed_form as cf


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _worker(rank, world, port, out_q):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        from deeplearninginassetpricing_paperreplication_b200 import functional as Fn
        g = load_golden("shipped_sub")
        params = {k: v.astype(np.float64) for k, v in golden_params(g).items()}
        macro, x, R, mask = g["macro"], g["x"], g["R"], g["mask"]
        T = x.shape[0]
        t0, t1 = (0, T // 2) if rank == 0 else (T // 2, T)
        shard = Fn.make_month_shard(torch.from_numpy(mask[t0:t1]), t0, T, None)
        n_bar_ref = np.maximum(mask.sum(axis=1), 1).mean()
        assert abs(float(shard.n_bar) - n_bar_ref) < 1e-4 and (shard.t0, shard.t1, shard.T_total) == (t0, t1, T)
        # full forward of the oracle gives every per-cell quantity; each rank only USES its slab
        full = cf.forward(params, macro, x, R, mask)
        s = full["_saved"]
        sl = slice(t0, t1)
        # exchange 1: per-stock sums over the rank's months, all-reduced
        S_loc = 1.0 + (full["weights"][sl] * R[sl] * mask[sl]).sum(axis=1) / np.maximum(mask[sl].sum(axis=1), 1) \
            * float(shard.n_bar)
        rm = (R * mask)[sl]
        col = np.concatenate([np.einsum("tnk,tn->kn", s["q"][sl], rm * S_loc[:, None]),
                              (rm * S_loc[:, None]).sum(axis=0, keepdims=True),
                              mask[sl].sum(axis=0, keepdims=True).astype(np.float64)], axis=0)
        N = x.shape[1]

        def exchange(col_local, pr_local, months):
            """dlap_step_reduce's xbuf: fp32 [10 N + T], summed over the ranks"""
            xb = np.zeros(10 * N + T, dtype=np.float32)
            xb[:10 * N] = col_local.reshape(-1).astype(np.float32)
            xb[10 * N + np.asarray(months)] = pr_local.astype(np.float32)
            xt = torch.from_numpy(xb)
            dist.all_reduce(xt)
            return xt.numpy()[:10 * N].reshape(10, N).astype(np.float64), xt.numpy()[10 * N:]

        colt, pr_full = exchange(col, full["portfolio_returns"][sl], np.arange(t0, t1))
        t_i = np.maximum(colt[9], 1.0)
        e = colt[:8] / t_i
        e0 = colt[8] / t_i
        assert np.allclose(e, s["e"], rtol=1e-5, atol=1e-9) and np.allclose(e0, s["e0"], rtol=1e-5, atol=1e-9)
        l_c = (e ** 2).mean(axis=1).mean()
        assert abs(l_c - full["loss_conditional"]) < 1e-5 * abs(full["loss_conditional"])
        # the series' portfolio returns rode along; Sharpe of the whole series (src/model.py:551)
        assert np.allclose(pr_full, full["portfolio_returns"], rtol=1e-5, atol=1e-8)
        sharpe = pr_full.astype(np.float64).mean() / (pr_full.astype(np.float64).std(ddof=1) + 1e-8)
        assert abs(float(sharpe) - float(full["sharpe"])) < 1e-4 * abs(float(full["sharpe"]))
        # exchange 2: da_t rows (zero outside the slab) summed over ranks == the full table
        gfull = cf.backward(params, macro, x, R, mask, full)
        G = 2.0 * e / (8 * x.shape[1] * t_i)
        gq = np.einsum("kn,tnk->tn", G, s["q"][sl])
        dS = (rm * gq).sum(axis=1)
        dP = dS * float(shard.n_bar) / s["n_t"][sl]
        rbar = rm.sum(axis=1) / s["n_t"][sl]
        du = mask[sl] * dP[:, None] * (R[sl] - rbar[:, None])
        dz2 = du[:, :, None] * params[cf.K_W3][0][None, None, :] * (s["z2"][sl] > 0)
        dz1 = (dz2 @ params[cf.K_W2]) * (s["z1"][sl] > 0)
        da = np.zeros((T, 64)); da[sl] = dz1.sum(axis=1)
        w2 = np.einsum("tnj,tnk->jk", dz2, s["h1"][sl])
        buf = torch.from_numpy(np.concatenate([da.reshape(-1), w2.reshape(-1)]))
        dist.all_reduce(buf)
        da_all = buf[: T * 64].numpy().reshape(T, 64)
        assert np.allclose(buf[T * 64:].numpy().reshape(64, 64), gfull[cf.K_W2], rtol=1e-4, atol=1e-12)
        assert np.allclose(da_all.sum(axis=0), gfull[cf.K_B1], rtol=1e-4, atol=1e-12)
        # months dealt round-robin: rank r owns months r, r + 2, ...
        rr = Fn.make_month_shard(torch.from_numpy(mask[rank::2]), rank, T, None, stride=2)
        assert abs(float(rr.n_bar) - n_bar_ref) < 1e-4 and rr.stride == 2 and rr.t1 <= T
        assert torch.equal(rr.rows(torch.arange(T)), torch.arange(T)[rank::2])
        _, pr_rr = exchange(np.zeros((10, N)), full["portfolio_returns"][rank::2], np.arange(rank, T, 2))
        assert np.allclose(pr_rr, full["portfolio_returns"], rtol=1e-5, atol=1e-8)
        out_q.put((rank, "ok"))
    except Exception:  # pragma: no cover
        import traceback
        out_q.put((rank, traceback.format_exc()))
    finally:
        dist.destroy_process_group()


def test_month_sharding_two_gloo_ranks():
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    res = [q.get(timeout=180) for _ in procs]
    for p in procs:
        p.join(timeout=60)
    assert sorted(res) == [(0, "ok"), (1, "ok")], res
