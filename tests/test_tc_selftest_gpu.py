"""Pins the tcgen05 operand-layout conventions of csrc/tc_sm100.cuh (no-swizzle canonical layout,
[group of 8][other dim][8] bf16) against numpy through dlap_umma_selftest.  If a convention were
wrong the test says which LBO/SBO assignment does match."""
import itertools
import os

import numpy as np
import pytest
import torch

from deeplearninginassetpricing_paperreplication_b200 import _lib

pytestmark = pytest.mark.gpu


def pack(mat):
    """[other, grouped] fp32 -> bf16 image [grouped/8, other, 8] and the bf16-rounded values."""
    R, C = mat.shape
    b = mat.to(torch.bfloat16)
    return b.view(R, C // 8, 8).permute(1, 0, 2).contiguous(), b.float()


def run(a_img, b_img, M, N, K, a_mn, b_mn, a_str, b_str):
    dev = a_img.device
    D = torch.full((128, N), float("nan"), device=dev)
    _lib.call("dlap_umma_selftest", a_img.data_ptr(), a_img.numel() * 2, b_img.data_ptr(), b_img.numel() * 2,
              M, N, K, a_mn, b_mn, *a_str, *b_str, D.data_ptr(), torch.cuda.current_stream().cuda_stream)
    torch.cuda.synchronize()
    return D


def strides(other, mn_major):
    """(lbo, sbo, kstep) in bytes for an operand stored [g][other][8]."""
    if mn_major:      # grouped dim = MN, other dim = K
        return (128, other * 16, 256)
    return (other * 16, 128, 2 * other * 16)   # grouped dim = K, other dim = MN


CASES = [  # name, N, K, a_mn, b_mn
    ("fwd_l2   A=h1[cell,unit] K-major, B=W2[j,k] K-major", 64, 64, 0, 0),
    ("fwd_l1   N=80 K=48", 80, 48, 0, 0),
    ("dgrad    A=dz2[cell,j] K-major, B=W2[j,k] as MN-major (N=k, K=j)", 64, 64, 0, 1),
    ("wgrad    A=dz[cell,unit] MN-major, B=[h1|x][cell,unit] MN-major, K=cells", 112, 128, 1, 1),
    ("wgrad_q  N=16", 16, 128, 1, 1),
]


@pytest.mark.parametrize("name,N,K,a_mn,b_mn", CASES)
def test_umma_layout_conventions(name, N, K, a_mn, b_mn):
    dev = torch.device("cuda:0")
    g = torch.Generator().manual_seed(N * 1000 + K + a_mn * 7 + b_mn * 13)
    M = 128
    # stored matrices are [other, grouped]
    if a_mn:   # stored [K, M]; logical A[m,k] = T[k,m]
        Ta = torch.randn(K, M, generator=g); a_img, a_val = pack(Ta); A = a_val.t()
        a_other = K
    else:      # stored [M, K]
        Ta = torch.randn(M, K, generator=g); a_img, a_val = pack(Ta); A = a_val
        a_other = M
    if b_mn:   # stored [K, N]; logical B[n,k] = T[k,n]
        Tb = torch.randn(K, N, generator=g); b_img, b_val = pack(Tb); B = b_val.t()
        b_other = K
    else:
        Tb = torch.randn(N, K, generator=g); b_img, b_val = pack(Tb); B = b_val
        b_other = N
    ref = (A.double() @ B.double().t()).float()
    a_img, b_img = a_img.to(dev), b_img.to(dev)
    D = run(a_img, b_img, M, N, K, a_mn, b_mn, strides(a_other, a_mn), strides(b_other, b_mn)).cpu()
    err = (D - ref).abs().max().item()
    if not err < 1e-3 * ref.abs().max().item():
        lines = [f"{name}: documented convention FAILS, max err {err:.3e}"]
        for sa, sb in itertools.product((0, 1), (0, 1)):
            sA = strides(a_other, a_mn); sB = strides(b_other, b_mn)
            if sa: sA = (sA[1], sA[0], sA[2])
            if sb: sB = (sB[1], sB[0], sB[2])
            Dx = run(a_img, b_img, M, N, K, a_mn, b_mn, sA, sB).cpu()
            lines.append(f"  swapA={sa} swapB={sb}: max err {(Dx - ref).abs().max().item():.3e} "
                         f"nan={int(torch.isnan(Dx).sum())}")
        os.makedirs("gpurun_out", exist_ok=True)
        open("gpurun_out/umma_selftest_fail.txt", "a").write("\n".join(lines) + "\n")
        pytest.fail("\n".join(lines))


def test_umma_m64_row_mapping_is_recorded():
    """M=64 accumulators use a different TMEM lane map; record it (the kernels only use M=128)."""
    dev = torch.device("cuda:0")
    g = torch.Generator().manual_seed(5)
    Ta = torch.randn(64, 32, generator=g); Tb = torch.randn(16, 32, generator=g)
    a_img, a_val = pack(Ta); b_img, b_val = pack(Tb)
    ref = a_val @ b_val.t()
    D = run(a_img.to(dev), b_img.to(dev), 64, 16, 32, 0, 0, strides(64, 0), strides(16, 0)).cpu()
    rows = []
    for r in range(64):
        hit = [l for l in range(128) if torch.allclose(D[l], ref[r], atol=1e-2, rtol=1e-2)]
        rows.append(hit[0] if hit else -1)
    os.makedirs("gpurun_out", exist_ok=True)
    open("gpurun_out/umma_m64_map.txt", "w").write(" ".join(map(str, rows)) + "\n")
    assert sum(r >= 0 for r in rows) == 64


@pytest.mark.parametrize("N,K,b_mn", [(64, 64, 0), (64, 64, 1), (80, 48, 0)])
def test_umma_a_from_tmem(N, K, b_mn):
    """A operand in tensor memory (lane = row, 32-bit column = two consecutive K values)."""
    dev = torch.device("cuda:0")
    g = torch.Generator().manual_seed(N + K + b_mn)
    A = torch.randn(128, K, generator=g).to(torch.bfloat16)
    if b_mn:
        Tb = torch.randn(K, N, generator=g); b_img, b_val = pack(Tb); B = b_val.t(); b_other = K
    else:
        Tb = torch.randn(N, K, generator=g); b_img, b_val = pack(Tb); B = b_val; b_other = N
    ref = (A.double() @ B.double().t()).float()
    D = torch.full((128, N), float("nan"), device=dev)
    b_img = b_img.to(dev); A_d = A.contiguous().to(dev)
    _lib.call("dlap_umma_ts_selftest", A_d.data_ptr(), b_img.data_ptr(), b_img.numel() * 2, N, K, b_mn,
              *strides(b_other, b_mn), 1, D.data_ptr(), None, None)
    torch.cuda.synchronize()
    err = (D.cpu() - ref).abs().max().item()
    assert err < 1e-3 * ref.abs().max().item(), err
    cyc = torch.zeros(1, dtype=torch.int64, device=dev)
    out = []
    for rep in (1, 64):
        _lib.call("dlap_umma_ts_selftest", A_d.data_ptr(), b_img.data_ptr(), b_img.numel() * 2, N, K, b_mn,
                  *strides(b_other, b_mn), rep, D.data_ptr(), cyc.data_ptr(), None)
        torch.cuda.synchronize()
        out.append(int(cyc))
    per = (out[1] - out[0]) / (63 * (K // 16))
    os.makedirs("gpurun_out", exist_ok=True)
    open("gpurun_out/umma_ts_cycles.txt", "a").write(f"TS N={N} K={K} b_mn={b_mn}: {per:.1f} cycles per MMA\n")
