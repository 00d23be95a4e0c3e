"""Generate golden_gpu.npz by running the reference's OWN CUDA kernels on a B200.

    gpurun -- 'python tests/golden/make_golden_gpu.py'        (writes gpurun_out/golden_gpu.npz;
                                                               copy it to tests/golden/ and commit)

Uses oracle/_ref/nerfacc_cuda.so (the unmodified reference extension, built for sm_100a by
oracle/build_ref.py) through tests/_refcuda.py.  These are the paths the reference cannot run
on CPU: packed scans fwd/bwd, traverse_grids, ray_aabb_intersect,
importance_sampling, searchsorted.  tests/test_oracle_golden.py checks the CPU oracle against
this file (bit-exact for the traversal), which is what pins the oracle for those paths.
"""
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from tests import _refcuda as R  # noqa: E402
from tests._util import packed_info_from_counts, ragged_counts  # noqa: E402

G = {}
dev = torch.device("cuda:0")


def put(name, t):
    G[name] = t.detach().cpu().numpy() if isinstance(t, torch.Tensor) else np.asarray(t)


def T(a):
    return torch.from_numpy(np.ascontiguousarray(a)).to(dev)


def enlarge(aabb, f):
    c, e = (aabb[:3] + aabb[3:]) / 2, (aabb[3:] - aabb[:3]) / 2
    return np.concatenate([c - e * f, c + e * f]).astype(np.float32)


def main():
    rng = np.random.default_rng(2024)

    # ---- packed scans, forward and backward (csrc/scan.cu)
    counts = ragged_counts(300, mean=20, seed=3, long_rays=((7, 3000), (150, 1), (151, 33)))
    pi = packed_info_from_counts(counts)
    n = int(counts.sum())
    x = (rng.random(n) * 0.9 + 0.05).astype(np.float32)
    g = rng.random(n).astype(np.float32)
    put("scan/x", x); put("scan/g", g); put("scan/packed_info", pi)
    tx, tg, tpi = T(x), T(g), T(pi)
    for op in ("sum", "prod"):
        for excl in (False, True):
            name = f"scan/{'exclusive' if excl else 'inclusive'}_{op}"
            y = R.scan(tx, tpi, op, excl)
            put(name, y)
            if op == "sum":
                put(name + "_grad", R.scan(tg, tpi, "sum", excl, backward=True))
            else:
                put(name + "_grad", R.prod_backward(tg, tx, y, tpi, excl))

    # ---- traversal (csrc/grid.cu)
    def nested(n_rays, levels, res, seed, p_occ=0.5):
        r = np.random.default_rng(seed)
        o = r.standard_normal((n_rays, 3)).astype(np.float32)
        d = r.standard_normal((n_rays, 3)).astype(np.float32)
        d /= np.linalg.norm(d, axis=-1, keepdims=True)
        base = np.array([-1, -1, -1, 1, 1, 1], np.float32)
        aabbs = np.stack([enlarge(base, 2**i) for i in range(levels)])
        bins = r.random((levels, res, res, res)) < p_occ
        return o, d.astype(np.float32), bins, aabbs

    def shell(n_rays, res, seed):
        r = np.random.default_rng(seed)
        lin = (np.arange(res) + 0.5) / res * 3.0 - 1.5
        X, Y, Z = np.meshgrid(lin, lin, lin, indexing="ij")
        rad = np.sqrt(X**2 + Y**2 + Z**2)
        bins = ((rad >= 0.8) & (rad <= 0.9))[None]
        o = r.standard_normal((n_rays, 3)); o = 4.0 * o / np.linalg.norm(o, axis=-1, keepdims=True)
        tgt = r.standard_normal((n_rays, 3)); tgt = tgt / np.linalg.norm(tgt, axis=-1, keepdims=True)
        tgt *= 1.5 * r.random((n_rays, 1)) ** (1 / 3)
        d = tgt - o; d /= np.linalg.norm(d, axis=-1, keepdims=True)
        return o.astype(np.float32), d.astype(np.float32), bins, np.array([[-1.5] * 3 + [1.5] * 3], np.float32)

    cases = {
        "nested": (nested(24, 4, 32, 42), dict(step_size=5e-3, cone_angle=0.0, limit=-1), True),
        "cone": (nested(24, 4, 32, 43), dict(step_size=0.013, cone_angle=0.004, limit=-1), True),
        "voxel": (nested(24, 3, 16, 44), dict(step_size=0.0, cone_angle=0.0, limit=-1), True),
        "limit": (nested(24, 4, 32, 45), dict(step_size=2e-2, cone_angle=0.0, limit=37), True),
        "shell": (shell(512, 128, 1234), dict(step_size=5e-3, cone_angle=0.0, limit=-1), False),
    }
    for name, ((o, d, bins, aabbs), kw, with_nf) in cases.items():
        p = f"trav/{name}/"
        put(p + "o", o); put(p + "d", d); put(p + "bins", np.packbits(bins).astype(np.uint8))
        put(p + "bins_shape", np.array(bins.shape)); put(p + "aabbs", aabbs)
        put(p + "step_size", kw["step_size"]); put(p + "cone_angle", kw["cone_angle"]); put(p + "limit", kw["limit"])
        near = far = None
        if with_nf:
            near = (np.random.default_rng(1).random(o.shape[0]) * 0.2).astype(np.float32)
            far = np.full(o.shape[0], 6.0, np.float32)
            put(p + "near", near); put(p + "far", far)
        iv, sm, term = R.traverse_grids(T(o), T(d), T(bins), T(aabbs), None if near is None else T(near),
                                        None if far is None else T(far), kw["step_size"], kw["cone_angle"],
                                        None if kw["limit"] < 0 else kw["limit"])
        for k, v in iv.items():
            put(f"{p}intervals/{k}", v)
        for k, v in sm.items():
            put(f"{p}samples/{k}", v)
        put(p + "termination_planes", term)
        print(name, "samples", sm["vals"].shape[0], "edges", iv["vals"].shape[0])

    # ---- ray_aabb_intersect
    o = rng.random((500, 3)).astype(np.float32)
    d = rng.standard_normal((500, 3)).astype(np.float32); d /= np.linalg.norm(d, axis=-1, keepdims=True)
    lo = rng.random((20, 3)).astype(np.float32)
    bb = np.concatenate([lo, lo + rng.random((20, 3)).astype(np.float32)], -1)
    tmin, tmax, hits = R.ray_aabb_intersect(T(o), T(d), T(bb))
    put("aabb/o", o); put("aabb/d", d); put("aabb/bb", bb); put("aabb/tmin", tmin); put("aabb/tmax", tmax)
    put("aabb/hits", hits)

    # ---- importance_sampling / searchsorted, batched (csrc/pdf.cu)
    vals = np.sort(rng.random((40, 65)).astype(np.float32), -1)
    cdfs = np.sort(rng.random((40, 65)).astype(np.float32), -1)
    cdfs[:, 0] = 0.0; cdfs[:, -1] = 1.0
    cdfs[3, 10:20] = cdfs[3, 10]  # flat stretch: degenerate bins (pdf.cu:159-160)
    edges, centers = R.importance_sampling(T(vals), T(cdfs), 48)
    query = np.sort(rng.random((40, 49)).astype(np.float32) * 1.2 - 0.1, -1)
    il, ir = R.searchsorted(T(vals), T(query))
    put("pdf/vals", vals); put("pdf/cdfs", cdfs); put("pdf/edges", edges); put("pdf/centers", centers)
    put("pdf/query", query); put("pdf/ids_left", il); put("pdf/ids_right", ir)

    out_dir = os.path.join(ROOT, "gpurun_out")
    os.makedirs(out_dir, exist_ok=True)
    out = os.path.join(out_dir, "golden_gpu.npz")
    np.savez_compressed(out, **G)
    print(f"wrote {out}: {len(G)} arrays, {os.path.getsize(out) / 1024:.0f} KiB")


if __name__ == "__main__":
    main()
