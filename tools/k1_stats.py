"""Developer tool (CPU): workload statistics of the epipolar search on the bench scene --
candidates per pixel x neighbour, row-run structure, gate pass rates.  Uses the oracle (dev tool)."""
import importlib, os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from __graft_entry__ import load_package, PKG_NAME
pkg = load_package()
pipeline = importlib.import_module(PKG_NAME + ".pipeline"); synth = importlib.import_module(PKG_NAME + ".synth")
from oracle import pm_oracle as po

cam = sys.argv[1] if len(sys.argv) > 1 else "720p"
nkf, n = 24, int(sys.argv[2]) if len(sys.argv) > 2 else 10
scene = synth.Scene(cam, seed=1234)
plan = pipeline.SequencePlan(scene, nkf, n)
k = 12
ora = po.Oracle(threads=8)
ids = [k] + list(plan.nbrs[k])
for j in ids:
    ora.add_keyframe(int(j), scene.render(int(j)), scene.K, plan.poses[int(j)])
W, H = scene.w, scene.h
g1 = ora.plane(k, po.GRAD); th1 = ora.plane(k, po.THETA)
ys, xs = np.nonzero(g1[2:H-2, 2:W-2] >= 8); ys += 2; xs += 2
print("active", len(xs), "of", W*H, "prior", plan.prior[k])
mind, maxd = plan.prior[k]
fx, fy, cx, cy = [np.float32(v) for v in scene.K]
tot_c = tot_s = tot_g1 = tot_runs = tot_px = 0
tot_q = 0.0
for jn, j in enumerate(plan.nbrs[k]):
    F, R, t = po.pair_geometry(scene.K, plan.poses[k], scene.K, plan.poses[int(j)])
    xf = xs.astype(np.float32); yf = ys.astype(np.float32)
    a = xf*F[0,0] + yf*F[1,0] + F[2,0]; b = xf*F[0,1] + yf*F[1,1] + F[2,1]; c = xf*F[0,2] + yf*F[1,2] + F[2,2]
    ab = a/b; cb = c/b
    ok = ~((ab < -4) | (ab > 4))
    xp0 = (xf-cx)/fx; xp1 = (yf-cy)/fy
    q0 = R[0,0]*xp0 + R[0,1]*xp1 + R[0,2]; q2 = R[2,0]*xp0 + R[2,1]*xp1 + R[2,2]
    def proj(r): return fx*(q0*np.float32(r)+t[0])/(q2*np.float32(r)+t[2]) + cx
    u1, u2 = proj(mind), proj(maxd)
    umin = np.clip(np.minimum(u1, u2), 0, W); umax = np.clip(np.maximum(u1, u2), 0, W)
    lo = np.ceil(umin).astype(int); hi = np.floor(umax).astype(int)
    cnt = np.where(ok, np.maximum(hi-lo+1, 0), 0)
    v_lo = -np.trunc(ab*lo + cb).astype(int); v_hi = -np.trunc(ab*hi + cb).astype(int)
    runs = np.where(cnt > 0, np.abs(v_hi - v_lo) + 1, 0)
    g2 = ora.plane(int(j), po.GRAD); t2 = ora.plane(int(j), po.THETA)
    rot = plan.rots[k][jn]
    # sample 4000 pixels for gate statistics
    rng = np.random.default_rng(0); sel = rng.choice(len(xs), 4000, replace=False)
    c_all = s_all = g1_all = q_all = 0
    for i in sel:
        if cnt[i] == 0: continue
        uj = np.arange(lo[i], hi[i]+1); vj = -np.trunc(ab[i]*uj.astype(np.float32) + cb[i]).astype(int)
        m = (vj > 0) & (vj < H)
        uj, vj = uj[m], vj[m]; ujc = np.minimum(uj, W-1)
        G = g2[vj, ujc]; T = t2[vj, ujc]
        p1 = G >= 8
        the = np.degrees(np.arctan2(-ab[i], 1.0)) % 360
        d = (T - the) % 360; d = np.where(d > 180, 360-d, d); d = np.where(d > 90, 180-d, d)
        e = (T - (th1[ys[i], xs[i]] + rot)) % 360; e = np.where(e > 180, 360-e, e)
        sv = p1 & (d <= 80) & (e <= 45)
        apr = (th1[ys[i], xs[i]] + rot) % 360
        cen = (np.floor(apr / 11.25) + 0.5) * 11.25
        cd = np.abs((T - cen + 180) % 360 - 180)
        q_all += (p1 & (cd <= 45 + 5.625 + 0.05)).sum()
        c_all += len(uj); g1_all += p1.sum(); s_all += sv.sum()
    print(f"nbr {int(j):3d} rot {rot:+.2f}: mean cand {cnt.mean():6.1f} max {cnt.max():4d}  p50/p90/p99 {np.percentile(cnt,[50,90,99])}  mean runs {runs.mean():.2f}  |ab| mean {np.abs(ab).mean():.4f} max {np.abs(ab).max():.3f}  gate1 {g1_all/max(c_all,1):.3f} surv {s_all/max(c_all,1):.4f} queued {q_all/max(c_all,1):.4f}")
    tot_q += q_all/max(c_all,1) * cnt.sum()
    tot_c += cnt.sum(); tot_runs += runs.sum(); tot_px += len(xs)
    tot_s += s_all / max(c_all, 1) * cnt.sum()
print(f"total candidates/KF {tot_c/1e6:.1f} M  per pixel-nbr {tot_c/tot_px:.1f}  runs per pixel-nbr {tot_runs/tot_px:.2f}  survivors/KF ~{tot_s/1e6:.2f} M")

print("queued by the arc planes /KF ~%.2f M" % (tot_q/1e6))
