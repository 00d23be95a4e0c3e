"""Shared input generators and comparison helpers for the parity tests."""
import numpy as np


def ragged_counts(n_rays, mean=12.0, sigma=0.9, p_empty=0.2, max_count=2048, seed=0, long_rays=()):
    """log-normal sample counts, a share of empty rays, optional giants."""
    rng = np.random.default_rng(seed)
    c = np.clip(np.round(np.exp(rng.normal(np.log(mean), sigma, n_rays))), 0, max_count).astype(np.int64)
    c[rng.random(n_rays) < p_empty] = 0
    for i, L in long_rays:
        c[i] = L
    return c


def packed_info_from_counts(counts):
    counts = np.asarray(counts, dtype=np.int64)
    starts = np.cumsum(counts) - counts
    return np.stack([starts, counts], -1)


def ray_indices_from_counts(counts):
    return np.repeat(np.arange(len(counts), dtype=np.int64), counts)


def assert_close(actual, expected, rtol, atol, what=""):
    actual = np.asarray(actual, dtype=np.float64)
    expected = np.asarray(expected, dtype=np.float64)
    assert actual.shape == expected.shape, (what, actual.shape, expected.shape)
    err = np.abs(actual - expected)
    tol = atol + rtol * np.abs(expected)
    bad = err > tol
    if bad.any():
        i = int(np.argmax(err - tol))
        raise AssertionError(
            f"{what}: {int(bad.sum())}/{bad.size} mismatches; worst at {i}: got {actual.flat[i]!r} "
            f"expected {expected.flat[i]!r} (err {err.flat[i]:.3e}, tol {tol.flat[i]:.3e})")


def record(name, values):
    """Append measured figures (max errors of ours / of the reference kernels against float64, ...)
    to gpurun_out/measured_errors.json — the source of the tolerance table in DESIGN.md section 2.
    Never fails a test."""
    import json
    import os

    try:
        root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
        path = os.path.join(root, "gpurun_out", "measured_errors.json")
        os.makedirs(os.path.dirname(path), exist_ok=True)
        data = {}
        if os.path.exists(path):
            with open(path) as f:
                data = json.load(f)
        data[name] = values
        with open(path, "w") as f:
            json.dump(data, f, indent=1, sort_keys=True)
    except Exception:
        pass


def err_vs(x, truth):
    """max |x - t| / (|t| + 1e-2 rms(t)) (the metric of bench_configs.err_vs, numpy)."""
    x, t = np.asarray(x, np.float64), np.asarray(truth, np.float64)
    scale = max(float(np.sqrt(np.mean(t * t))), 1e-30)
    return float(np.max(np.abs(x - t) / (np.abs(t) + 1e-2 * scale))) if t.size else 0.0


def rms_err(x, truth):
    x, t = np.asarray(x, np.float64), np.asarray(truth, np.float64)
    return float(np.sqrt(np.mean((x - t) ** 2)) / max(np.sqrt(np.mean(t * t)), 1e-300)) if t.size else 0.0


def assert_as_close_as_reference(ours, ref32, truth64, what="", tol=1e-5):
    """north_star's float bar with numbers behind it: `ours` within 1e-5 of a float64 evaluation,
    or as close to it as the reference's fp32 arithmetic (`ref32`: the oracle's sequential fp32
    restatement, or the reference's own kernels) is: max error within 2x, RMS error within 1.25x."""
    e_o, e_r = err_vs(ours, truth64), err_vs(ref32, truth64)
    r_o, r_r = rms_err(ours, truth64), rms_err(ref32, truth64)
    ok = e_o <= max(tol, 2.0 * e_r) and r_o <= max(0.1 * tol, 1.25 * r_r)
    assert ok, f"{what}: max err ours {e_o:.3e} vs reference fp32 {e_r:.3e}; rms ours {r_o:.3e} vs {r_r:.3e}"
    return e_o, e_r
