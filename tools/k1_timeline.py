"""Reads the per-warp clock stamps a -DEWN_EXP_TIMELINE build of k_vstep left behind (tools/k1_sweep.py with EWN_K1_TIMELINE=file.npy)
and prints, per event, the distribution over the warps of the LAST launch (ns since the first warp entered; clock64 scaled by the SM clock,
SMs aligned on %globaltimer at entry).   python tools/k1_timeline.py file.npy [blocks warps_per_block]"""
import sys
import numpy as np
tl = np.load(sys.argv[1]).astype(np.int64)
blocks, wpb = (int(sys.argv[2]), int(sys.argv[3])) if len(sys.argv) > 3 else (1480, 4)
nw = blocks * wpb
t = tl[:nw]
smid, gt = t[:, 14], t[:, 15]
clk = 1.965
ev = t[:, :13].astype(float)
out = np.zeros_like(ev)
g0 = gt.min()
for s in np.unique(smid):
    m = smid == s
    c0 = ev[m, 0].min()
    out[m] = (ev[m] - c0) / clk + (gt[m][np.argmin(ev[m, 0])] - g0)
names = ["entry", "after_wait", "u0", "u1", "u2", "u3", "u4", "u5", "u6", "u7", "loop_end", "resv_done", "flush_end"]
for k, nm in enumerate(names):
    valid = t[:, k] > 0
    if not valid.any():
        continue
    col = out[valid, k]
    print(f"{nm:10s} min {col.min():8.0f} p10 {np.percentile(col, 10):8.0f} med {np.median(col):8.0f} p90 {np.percentile(col, 90):8.0f} max {col.max():8.0f} ns  warps {valid.sum()}")
units = t[:, 13]
print("units per warp: ", {int(k): int((units == k).sum()) for k in np.unique(units)}, " total", int(units.sum()))
