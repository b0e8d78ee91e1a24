"""Where does the idle tail of burgers_pair_loss_unit_kernel come from?  Needs a library built with -DF2_TIMING
(scripts/diag/build_variants.sh burgers_fast2.cu timing:-DF2_TIMING; PDEC_LIB_PATH=pde_control_b200/variants/libpdec_timing.so).
Every warp records (SM id, start ns, end ns); printed: spread of warp durations, finish time per SM, and whether the slow SMs are
the same ones launch after launch (hardware position) or move with the input set (data)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import numpy as np, torch
from benchlib.burgers import BurgersConfig2

wl = BurgersConfig2(torch.device("cuda", 0), 0, 1)
for _ in range(20): wl.step()
torch.cuda.synchronize()
W = 148 * 2 * 14
per_sm = []
for rep in range(6):
    wl.i = rep % 2                      # input set 0 / 1 alternately
    wl.step(); torch.cuda.synchronize()
    raw = wl.ws.view(torch.int32).cpu().numpy().astype(np.int64) & 0xffffffff
    d = raw[65536 * 2: 65536 * 2 + 4 * W].reshape(W, 4)
    smid, t0, t1 = d[:, 0], d[:, 1], d[:, 2]
    start = t0.min()
    dur = (t1 - t0) / 1e3
    end = (t1 - start) / 1e3
    fin = np.array([end[smid == m].max() if (smid == m).any() else 0 for m in range(148)])
    mean_dur = np.array([dur[smid == m].mean() if (smid == m).any() else 0 for m in range(148)])
    per_sm.append(mean_dur)
    print("set %d: warp duration us min/mean/max %.1f %.1f %.1f | last warp ends %.1f | SM finish min/mean/max %.1f %.1f %.1f | start skew %.1f"
          % (rep % 2, dur.min(), dur.mean(), dur.max(), end.max(), fin.min(), fin.mean(), fin.max(), (t0.max() - start) / 1e3))
per_sm = np.array(per_sm)
print("correlation of per-SM mean warp duration, same input set, different launches:", np.corrcoef(per_sm[0], per_sm[2])[0, 1], np.corrcoef(per_sm[1], per_sm[3])[0, 1])
print("correlation between the two input sets:", np.corrcoef(per_sm[0], per_sm[1])[0, 1], np.corrcoef(per_sm[2], per_sm[3])[0, 1])
order = np.argsort(per_sm.mean(0))
print("fastest SMs", order[:10], per_sm.mean(0)[order[:10]].round(1))
print("slowest SMs", order[-10:], per_sm.mean(0)[order[-10:]].round(1))
# per-warp duration vs position inside the grid (range index)
d_sorted = dur.reshape(-1, 14).mean(1)
print("mean warp duration of CTA 0..9:", d_sorted[:10].round(1), " last 10:", d_sorted[-10:].round(1))
cta_mean = dur.reshape(-1, 14).mean(1)
print("mean warp duration by CTA octile:", [round(float(cta_mean[i * 37:(i + 1) * 37].mean()), 1) for i in range(8)])
cta_sm = smid.reshape(-1, 14)[:, 0]
same = sum(1 for i in range(148) if cta_sm[i] == cta_sm[i + 148])
print("CTA i and CTA i + 148 on the same SM:", same, "of 148;  SM of CTA 0..7:", cta_sm[:8], " of CTA 148..155:", cta_sm[148:156])
begin = ((t0 - start) / 1e3).reshape(-1, 14)[:, 0]
print("CTA start us, octiles:", [round(float(begin[i * 37:(i + 1) * 37].mean()), 2) for i in range(8)])
