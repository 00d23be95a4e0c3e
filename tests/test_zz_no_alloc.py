"""nfb_compact_samples_static — the allocation-free variant of nfb_compact_samples (its look-back
scratch is a file-scope `static __device__` array inside the library, include/nerfacc_b200.h).

No counterpart in the reference: the entry point exists for callers that cannot call cudaMalloc
(CUDA-graph capture, serving loops).  It must give exactly what nfb_compact_samples gives — the
same kernel, only the scratch differs — on calls made back to back (the array keeps the previous
call's contents, so stale look-back states must not leak into the next call), whether this file
runs alone in a fresh process or after every other test.  Named test_zz_* so pytest collects it
last.
"""
import pytest
import torch

from nerfacc_b200 import _C

MAX_N = 1 << 26  # NFB_COMPACT_STATIC_MAX_SAMPLES


def _inputs(n, keep, seed, dev):
    g = torch.Generator(device=dev).manual_seed(seed)
    mask = torch.rand(n, device=dev, generator=g) < keep
    ray_indices = torch.sort(torch.randint(0, max(n // 30, 1), (n,), device=dev, generator=g))[0]
    t_starts = torch.rand(n, device=dev, generator=g)
    t_ends = t_starts + torch.rand(n, device=dev, generator=g)
    return mask, ray_indices, t_starts, t_ends


def _outputs(ray_indices, t_starts, t_ends):
    """Outputs pre-filled with a sentinel so untouched slots show; total on the device."""
    return (torch.full_like(ray_indices, -7), torch.full_like(t_starts, -7.0), torch.full_like(t_ends, -7.0),
            torch.full((1,), -1, dtype=torch.long, device=ray_indices.device))


def _call(static, mask, ray_indices, t_starts, t_ends, out, ws=None):
    """One call through the C ABI (nothing but the call: the caller owns every buffer)."""
    dev, n = mask.device, mask.numel()
    o_ray, o_ts, o_te, total = out
    lib = _C.lib()
    with torch.cuda.device(dev):
        if static:
            rc = lib.nfb_compact_samples_static(
                mask.data_ptr(), ray_indices.data_ptr(), t_starts.data_ptr(), t_ends.data_ptr(), n,
                o_ray.data_ptr(), o_ts.data_ptr(), o_te.data_ptr(), total.data_ptr(), _C.stream_ptr(dev))
        else:
            rc = lib.nfb_compact_samples(
                mask.data_ptr(), ray_indices.data_ptr(), t_starts.data_ptr(), t_ends.data_ptr(), n,
                o_ray.data_ptr(), o_ts.data_ptr(), o_te.data_ptr(), total.data_ptr(), ws.data_ptr(),
                ws.numel(), _C.stream_ptr(dev))
    _C.check(rc, "compact_samples_static" if static else "compact_samples")
    torch.cuda.synchronize(dev)


def _run(static, mask, ray_indices, t_starts, t_ends):
    out = _outputs(ray_indices, t_starts, t_ends)
    ws = None
    if not static:
        ws = torch.empty(int(_C.lib().nfb_compact_workspace_bytes(mask.numel())), dtype=torch.uint8,
                         device=mask.device)
    _call(static, mask, ray_indices, t_starts, t_ends, out, ws)
    return out[0], out[1], out[2], int(out[3].item())


@pytest.mark.gpu
def test_static_scratch_compaction_matches_the_workspace_entry_point(device):
    # two calls in a row, then a third: a big one (1 465 tiles), a small one whose tiles were all
    # left in the "prefix published" state by the first, and one in between with another density
    cases = [(3_000_017, 0.37, 11), (70_001, 0.9, 12), (1_234_567, 0.05, 13)]
    for i, (n, keep, seed) in enumerate(cases):
        mask, ri, ts, te = _inputs(n, keep, seed, device)
        ref = _run(False, mask, ri, ts, te)
        out = _outputs(ri, ts, te)
        torch.cuda.synchronize(device)
        free0 = torch.cuda.mem_get_info(device)[0]
        _call(True, mask, ri, ts, te, out)
        free1 = torch.cuda.mem_get_info(device)[0]
        got = (out[0], out[1], out[2], int(out[3].item()))
        k = ref[3]
        assert got[3] == k == int(mask.sum().item()), (n, got[3], k)
        # bit-exact (integer ids, floats copied unchanged), including the untouched tail
        for a, b, what in zip(got[:3], ref[:3], ("ray_indices", "t_starts", "t_ends")):
            assert torch.equal(a, b), (n, what)
        assert torch.equal(got[0][:k], ri[mask]) and torch.equal(got[1][:k], ts[mask])
        if i == 2:  # (by now every code path of the call is warm: nothing may be allocated)
            assert free1 == free0, f"device memory moved by {free0 - free1} bytes across the call"


@pytest.mark.gpu
def test_static_scratch_compaction_empty_and_all_kept(device):
    mask, ri, ts, te = _inputs(4099, 2.0, 5, device)  # keep everything
    got = _run(True, mask, ri, ts, te)
    assert got[3] == 4099 and torch.equal(got[0], ri) and torch.equal(got[2], te)
    got = _run(True, torch.zeros_like(mask), ri, ts, te)  # keep nothing, right after a full call
    assert got[3] == 0 and bool((got[0] == -7).all())
    total = torch.full((1,), -1, dtype=torch.long, device=device)
    rc = _C.lib().nfb_compact_samples_static(None, None, None, None, 0, None, None, None,
                                             total.data_ptr(), _C.stream_ptr(device))
    assert rc == 0 and int(total.item()) == 0


def test_static_scratch_compaction_rejects_sizes_beyond_its_scratch():
    """Argument check only (returns before any CUDA call, so it also runs on the CPU-only box)."""
    rc = _C.lib().nfb_compact_samples_static(None, None, None, None, MAX_N + 1, None, None, None, None, None)
    assert rc == 1
    msg = _C.last_error()
    assert "exceeds" in msg and str(MAX_N) in msg, msg
    assert _C.lib().nfb_compact_samples_static(None, None, None, None, -1, None, None, None, None, None) == 1
