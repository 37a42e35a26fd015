#!/usr/bin/env python
"""Cycles per tcgen05.mma (bf16, K=16) for the operand shapes/majors the fused kernels use."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from deeplearninginassetpricing_paperreplication_b200 import _lib

dev = torch.device("cuda:0")
def strides(other, mn):
    return (128, other * 16, 256) if mn else (other * 16, 128, 2 * other * 16)
cases = [("L2   M128 N64 K64 KK", 128, 64, 64, 0, 0), ("L1   M128 N80 K48 KK", 128, 80, 48, 0, 0),
         ("DG   M128 N64 K64 K/MN", 128, 64, 64, 0, 1), ("WG2  M64 N64 K128 MN/MN", 64, 64, 128, 1, 1),
         ("WG1  M64 N48 K128 MN/MN", 64, 48, 128, 1, 1), ("MQ   M64 N8 K128 MN/MN", 64, 8, 128, 1, 1),
         ("WGx  M128 N112 K128 MN/MN", 128, 112, 128, 1, 1), ("big  M128 N256 K64 KK", 128, 256, 64, 0, 0),
         ("WG2' M64 N64 K128 KK-layout", 64, 64, 128, 0, 0)]
for name, M, N, K, amn, bmn in cases:
    a_other = K if amn else 128
    b_other = K if bmn else N
    a = torch.randn((128 if amn else K) // 8, a_other, 8).to(torch.bfloat16).to(dev)
    b = torch.randn((N if bmn else K) // 8, b_other, 8).to(torch.bfloat16).to(dev)
    D = torch.zeros(128, N, device=dev); cyc = torch.zeros(1, dtype=torch.int64, device=dev)
    out = []
    for rep in (1, 64):
        _lib.call("dlap_umma_cycles", a.data_ptr(), a.numel() * 2, b.data_ptr(), b.numel() * 2, M, N, K, amn, bmn,
                  *strides(a_other, amn), *strides(b_other, bmn), rep, D.data_ptr(), cyc.data_ptr(), None)
        torch.cuda.synchronize()
        out.append(int(cyc))
    n1, n64 = K // 16, 64 * (K // 16)
    per = (out[1] - out[0]) / (n64 - n1)
    _lib.call("dlap_umma_cycles", a.data_ptr(), a.numel() * 2, b.data_ptr(), b.numel() * 2, M, N, K, amn, bmn,
              *strides(a_other, amn), *strides(b_other, bmn), -64, D.data_ptr(), cyc.data_ptr(), None)
    torch.cuda.synchronize()
    per2 = (int(cyc) - out[0]) / (n64 - n1)
    print(f"{name:32s} 1 pass: {out[0]:6d} cyc ({n1} MMAs)   steady: {per:7.1f} cyc/MMA   2 accumulators: {per2:6.1f}   floor {max(M,128)*N/256:5.1f}")
