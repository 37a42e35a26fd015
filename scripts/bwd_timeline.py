#!/usr/bin/env python
"""Per-phase cycle counts of the tensor-core per-cell kernels (CTA 0), from dlap_debug_timeline stamps."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import bench
from deeplearninginassetpricing_paperreplication_b200 import AssetPricingGAN, _lib

dev = torch.device("cuda:0")
macro, x, R, mask = [t.to(dev) for t in bench.synthetic_panel(240, 10000, 46, 178, seed=1234)]
torch.manual_seed(42)
model = AssetPricingGAN({"macro_feature_dim": 178, "individual_feature_dim": 46, "dropout": 0.0,
                         "b200_precision": "bf16x3"}).to(dev)
for _ in range(3):
    model.zero_grad(); model(macro, x, R, mask)["loss"].backward()
st = torch.zeros(4 * 256 * 16, dtype=torch.int64, device=dev)
_lib.call("dlap_debug_timeline", st.data_ptr())
model.zero_grad(); model(macro, x, R, mask)["loss"].backward()
torch.cuda.synchronize()
_lib.call("dlap_debug_timeline", None)
s = st[:256 * 16].view(256, 16).cpu().double()
f = st[256 * 16:2 * 256 * 16].view(256, 16).cpu().double()
gb = st[2 * 256 * 16:3 * 256 * 16].view(256, 16).cpu().double()
bp = st[3 * 256 * 16:].view(256, 16).cpu()
ntile = int((s[:, 0] > 0).sum())
lo, hi = 5, max(ntile - 5, 7)
print('tiles of CTA 0:', ntile)
def d(a, b, shift=0):
    return float((s[lo + shift:hi + shift, b] - s[lo:hi, a]).mean())
print("tile period (epilogue E2 start -> next E2 start):", d(0, 0, 1))
print("epilogue: wait D2_FULL", d(0, 1), "| E2 body", d(1, 2), "| E1(i+1) incl. waits", d(2, 3),
      "| wait D3_FULL", d(3, 4), "| E3 body", d(4, 5), "| tail (flush etc.)", d(5, 0, 1))
print("MMA: wait DZ2_FULL -> issued WG2+DG", d(8, 9), "| L2(i+1) incl. wait h1", d(9, 10), "| wait DZ1_FULL", d(10, 11),
      "| WG1 + L1(i+2) issue", d(11, 12), "| next wait DZ2_FULL", d(12, 8, 1))

def gd(a, b, shift=0):
    nb = int((gb[:, 0] > 0).sum())
    idx = torch.arange(4, max(nb - 6, 6))
    return float((gb[idx + shift, b] - gb[idx, a]).mean())
print("BWD gather warp per 16-row batch (8 batches per tile over 3 warps): period", gd(0, 0, 1), "| wait XOP_EMPTY", gd(0, 1),
      "| convert (waits for the row loads) + arrive", gd(1, 2), "| issue next loads", gd(2, 3), "| loop tail", gd(3, 0, 1))


def fd(a, b, shift=0, step=1):
    idx = torch.arange(lo, hi - 2, step)
    return float((f[idx + shift, b] - f[idx, a]).mean())
print("FWD gather warp 4 per 32-row batch (4 batches per tile over 7 warps): period", fd(0, 0, 1), "| wait XOP_EMPTY", fd(0, 1),
      "| convert (waits for the row loads) + arrive", fd(1, 2), "| issue next loads", fd(2, 3), "| loop tail", fd(3, 0, 1))
print("FWD MMA: wait XOP_FULL(i) after L2(i-2).. L1 issue", fd(8, 9), "| L1 issued -> H1_FULL(i-1) waited", fd(9, 10, -1) if lo > 0 else 0,
      "| L2 issue", fd(10, 11), "| period", fd(8, 8, 1))
print("FWD epilogue (even tiles): D1_FULL -> E1 done", fd(12, 13, 0, 2), "| -> D2_FULL", fd(13, 14, 0, 2), "| E2+outputs", fd(14, 15, 0, 2),
      "| period per 2 tiles", fd(12, 12, 2, 2))
print("FWD chain for tile i:", "| L1 issued(9) -> D1_FULL seen(12)", fd(9, 12, 0, 2),
      "| E1 done(13) -> L2 start(10)", fd(13, 10, 0, 2), "| L2 issued(11) -> D2_FULL seen(14)", fd(11, 14, 0, 2))
nf = int((f[:, 15] > 0).sum())
first = float(min(x for x in f[0, [8, 9, 12]].tolist() if x > 0))
print("FWD CTA 0: tiles", nf, "| first stamp -> last epilogue stamp:", float(f[nf - 1, 15]) - first, "cycles | per tile",
      (float(f[nf - 1, 15]) - first) / max(nf, 1), "| first 6 tile ends:", [float(f[k, 15]) - first for k in range(6)])
nb_ = int((s[:, 5] > 0).sum())
print("BWD CTA 0: tiles", nb_, "| first E2 start -> last E3 end:", float(s[nb_ - 1, 5] - s[0, 0]), "cycles | per tile",
      float(s[nb_ - 1, 5] - s[0, 0]) / max(nb_, 1), "| first 5 tile E3 ends:", [float(s[k, 5] - s[0, 0]) for k in range(5)],
      "| last 3 periods:", [float(s[k, 0] - s[k - 1, 0]) for k in range(nb_ - 3, nb_)])
print("BWD CTA 0: entry -> set-up done", float(s[255, 1] - s[255, 0]), "| set-up done -> first E2 start", float(s[0, 0] - s[255, 1]),
      "| last E3 end -> all roles done", float(s[255, 2] - s[nb_ - 1, 5]), "| entry -> end", float(s[255, 2] - s[255, 0]))
cta = st[2 * 4096 + 2048:2 * 4096 + 2048 + 148].cpu()
cyc = (cta >> 8).double(); nt = (cta & 255)
print("BWD per-CTA active cycles: min", float(cyc.min()), "median", float(cyc.median()), "max", float(cyc.max()),
      "| tiles per CTA min/max", int(nt.min()), int(nt.max()), "| slowest CTAs", torch.topk(cyc, 5).indices.tolist())

t0 = int(bp[0, 0])
if t0:
    print("BPTT recurrence CTA (ns from its start): dadc rows complete", int(bp[0, 1]) - t0, "| records built", int(bp[0, 2]) - t0,
          "| scan done", int(bp[0, 3]) - t0, "| dpre written", int(bp[0, 4]) - t0, "| flag released", int(bp[0, 5]) - t0)
    for c in (0, 5, 11, 12, 16, 30, 48):
        r = bp[1 + c]
        if int(r[0]):
            print(f"BPTT summing CTA {c}: start {int(r[0]) - t0} | sums done {int(r[1]) - t0} | rows final seen {int(r[2]) - t0} | dpre seen {int(r[3]) - t0} | tasks done {int(r[4]) - t0}")

rd = st[3 * 256 * 16 + 2048:3 * 256 * 16 + 2048 + 64].cpu()
if int(rd[0]):
    names = ["start", "p0 done", "sync0", "cta0 stats", "p1 done", "sync1", "p2 done", "sync2", "p3 done", "sync3", "p4 done"]
    r0 = int(rd[0])
    for off, who in ((0, "CTA 0"), (32, "last CTA")):
        print("REDUCE", who, "(ns):", " | ".join(f"{n} {int(rd[off + k]) - r0}" for k, n in enumerate(names)))

if int(rd[0]):
    ends = st[3 * 256 * 16 + 2048 + 64:3 * 256 * 16 + 2048 + 64 + 592].cpu()
    smid = st[3 * 256 * 16 + 2048 + 64 + 640:3 * 256 * 16 + 2048 + 64 + 640 + 592].cpu()
    live = ends > 0
    e = (ends[live] - r0).double() / 1000.0
    print("REDUCE p1 end per CTA (us): n", int(live.sum()), "min %.1f p10 %.1f median %.1f p90 %.1f max %.1f" % (
        float(e.min()), float(e.kthvalue(max(1, int(0.1 * e.numel()))).values), float(e.median()),
        float(e.kthvalue(int(0.9 * e.numel())).values), float(e.max())))
    order = torch.argsort(e, descending=True)[:12]
    idx = torch.nonzero(live).flatten()[order]
    print("REDUCE slowest CTAs (cta, sm, end):", [(int(c), int(smid[c]), round(float((ends[c] - r0)) / 1000, 1)) for c in idx])
    per_sm = {}
    for c in torch.nonzero(live).flatten().tolist():
        per_sm.setdefault(int(smid[c]), []).append(c)
    print("REDUCE CTAs per SM: min", min(len(v) for v in per_sm.values()), "max", max(len(v) for v in per_sm.values()), "SMs used", len(per_sm))
