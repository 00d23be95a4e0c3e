"""Why `jump_ahead`'s advance() calls diverged: numpy float32 emulation of advance<false>() (grid.cuh) on 4096 rays of
the bench scene, counting loop trips per (ray, axis) and per warp.  TIES=0: rounding ties stepped literally (the
round's first state); TIES=1 (default): ties taken linearly from an even mantissa.  No GPU needed.

    python profiles/advance_tie_emulation.py        # after
    TIES=0 python profiles/advance_tie_emulation.py # before
"""
import os
TIES = os.environ.get("TIES", "1") != "0"
import numpy as np, sys
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), '..'))
from bench import make_rays, make_grid, AABB, RES
f = np.float32
o, d = make_rays(4096, 1234)
bins = make_grid()[0]
idx = np.argwhere(bins); lo_b, hi_b = idx.min(0), idx.max(0)
bmin, bmax = AABB[:3], AABB[3:]
stats = {"trips": [], "single": [], "linear": [], "k": []}
why = {}
def advance(t, dt, target):
    trips = single = linear = 0; steps = 0
    if not (dt > 0) or target != target: return t, 0, 0, 0, 0
    while True:
        if t >= target: break
        trips += 1
        t1 = f(t + dt)
        if not (t1 > t): break
        q = f(t1 - t)
        tb = np.float32(t).view(np.uint32); E = int(tb >> 23)
        if 24 <= E < 254:
            u = np.uint32((E - 23) << 23).view(np.float32); top = np.uint32((E + 1) << 23).view(np.float32)
            r = f(dt - q)
            ti = int((tb & 0x7fffff) | 0x800000)
            if t1 <= top and (f(abs(r) * f(2)) != u or (TIES and (ti & 1) == 0)):
                qi = int(np.rint(f(q * np.uint32((277 - E) << 23).view(np.float32))))
                kmax = (0x1000000 - ti) // qi
                k = kmax
                t_at = lambda k: f(f(ti + k * qi) * u)
                if kmax > 1 and t_at(kmax) >= target:
                    gap = f(target - t); kf = np.ceil(f(gap / q))
                    k = kmax if kf >= kmax else (1 if kf <= 1 else int(kf))
                    while k > 1 and t_at(k - 1) >= target: k -= 1
                    while not (t_at(k) >= target): k += 1
                linear += 1; steps += k; t = t_at(k)
                if not (t >= target):
                    t2 = f(t + dt)
                    if t2 > t: steps += 1; t = t2
                continue
            else:
                key = "cross" if not (t1 <= top) else "tie"
        else:
            key = "E%d" % E
        why[key] = why.get(key, 0) + 1
        single += 1; steps += 1; t = t1
    return t, trips, single, linear, steps
tot = []
for i in range(o.shape[0]):
    oi, di = o[i], d[i]
    inv = f(1) / di
    t0 = (bmin - oi) * inv; t1 = (bmax - oi) * inv
    tmin = f(max(np.minimum(t0, t1).max(), 0)); tmax = f(np.maximum(t0, t1).min())
    if not tmin < tmax: continue
    vox = (bmax - bmin) / f(RES)
    plo = bmin + (lo_b - 1).astype(f) * vox; phi = bmin + (hi_b + 2).astype(f) * vox
    a = (plo - oi) / di; b = (phi - oi) / di
    t_in = np.minimum(a, b).max(); t_out = np.maximum(a, b).min()
    if t_in >= t_out or not t_in > tmin: continue
    ts = f(tmin + f(1e-6)); rs = di * ts + oi
    cur = np.clip(((rs - bmin) / (bmax - bmin) * f(RES)).astype(int), 0, RES - 1)
    st = cur + (di > 0)
    tm = (bmin + (st.astype(f) * vox - rs)) * inv + tmin
    delta = np.abs(vox * inv)
    row = []
    for k in range(3):
        _, trips, single, linear, steps = advance(f(tm[k]), f(delta[k]), f(t_in))
        row.append((trips, single, linear, steps))
    tot.append(row)
tot = np.array(tot)
print("rays", len(tot), "mean trips/axis", tot[:, :, 0].mean(), "single", tot[:, :, 1].mean(), "linear", tot[:, :, 2].mean(), "steps", tot[:, :, 3].mean())
print("max trips per ray-axis", tot[:, :, 0].max(), "p99", np.percentile(tot[:, :, 0], 99), "hist", np.bincount(tot[:, :, 0].ravel())[:12])
print("why", why)
# per-warp cost: sum over axes of max over 32 lanes
w = tot[: len(tot) // 32 * 32].reshape(-1, 32, 3, 4)
print("per warp: sum_axes max_lane trips", w[:, :, :, 0].max(1).sum(1).mean(), " mean lane trips", w[:, :, :, 0].sum(2).mean())
