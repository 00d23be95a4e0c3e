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
