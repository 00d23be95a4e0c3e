// composite.cu — the per-ray compositing of nerfacc.rendering() (volrend.py:140-156) and its
// backward, fused.
//
// The reference calls accumulate_along_rays three times (rgb, opacity, depth): three
// [S,D] temporaries and three index_add_ passes forward, three gathers + two adds backward,
// then the scan backward.  Here:
//   composite_accum_kernel   colors[R,3], opacity[R], depth[R] = segmented sums of
//                            w*rgb, w, w*(ts+te)/2 in ONE pass over (w, rgb, ts, te):
//                            deterministic (no atomics), every ray row written exactly once;
//   composite_bwd_kernel     one reverse pass: forms the total gradient on the weights
//                              g_w = g_w_ext + <g_colors[r], rgb> + g_opacity[r] + g_depth[r]*tmid
//                            from the per-ray rows (gathered through ray_indices), writes
//                            g_rgb = w * g_colors[r], and runs the weight backward of
//                            volrend.cu on g_w in the same tile (reverse segmented scan),
//                            writing g_sigmas (or g_alphas).
#include "seg_tile.cuh"

namespace nfb {

// 4 consecutive items x 3 channels = 12 consecutive floats = three aligned float4.
template <bool REV>
__device__ __forceinline__ void load_rgb(const TileGeom<REV>& g, const float* __restrict__ rgb,
                                         bool vec_ok, int k, float (&x)[4][3]) {
    const long long gi = g.base + g.voff(k);
    if (vec_ok && gi + 3 < g.n) {
        float flat[12];
        const float* p = rgb + gi * 3;
#pragma unroll
        for (int q = 0; q < 3; ++q) {
            const float4 t = ldg_stream4(p + 4 * q);
            flat[4 * q] = t.x; flat[4 * q + 1] = t.y; flat[4 * q + 2] = t.z; flat[4 * q + 3] = t.w;
        }
#pragma unroll
        for (int c = 0; c < 4; ++c)
#pragma unroll
            for (int d = 0; d < 3; ++d) x[REV ? 3 - c : c][d] = flat[c * 3 + d];
    } else {
#pragma unroll
        for (int c = 0; c < 4; ++c) {
            const long long i = gi + (REV ? 3 - c : c);
#pragma unroll
            for (int d = 0; d < 3; ++d) x[c][d] = (i < g.n) ? ldg_stream(rgb + i * 3 + d) : 0.0f;
        }
    }
}

// One channel of the composite: segmented inclusive sum of p, the value at every segment tail
// goes to s_tail[pos]; then the rays that end in this tile read their total.
__device__ __forceinline__ void reduce_channel(float (&p)[kVec][4], const uint32_t (&f)[kVec],
                                               const TileGeom<false>& g, int slot,
                                               unsigned long long* status, uint32_t epoch, TileSmem& sm,
                                               float* s_tail, const int64_t* __restrict__ packed_info,
                                               float* __restrict__ out, int out_stride, int out_off) {
    const uint8_t* fb = reinterpret_cast<const uint8_t*>(sm.flags);
    float E[kVec];
    tile_segscan<SumOp>(p, f, E, status, slot, epoch, sm);
#pragma unroll
    for (int k = 0; k < kVec; ++k) {
        float I[4];
        finish_inclusive<SumOp>(p[k], f[k], E[k], I);
        const int j0 = g.voff(k);
#pragma unroll
        for (int c = 0; c < 4; ++c) {
            const long long gi = g.base + j0 + c;
            const bool next_head = (c < 3) ? flag_at(f[k], c + 1) : (fb[j0 + 4] != 0);
            if (gi < g.n && (next_head || gi == g.n - 1)) s_tail[j0 + c] = I[c];
        }
    }
    __syncthreads();
    const long long r_lo = sm.range[2], r_hi = sm.range[3];
    for (long long r = r_lo + threadIdx.x; r < r_hi; r += kThreads) {
        const RayChunk c = load_chunk(packed_info, r);
        out[r * out_stride + out_off] = (c.cnt > 0) ? s_tail[c.start + c.cnt - 1 - g.base] : 0.0f;
    }
    __syncthreads();
}

__global__ void __launch_bounds__(kThreads)
composite_accum_kernel(const float* __restrict__ weights, const float* __restrict__ rgbs,
                       const float* __restrict__ t_starts, const float* __restrict__ t_ends,
                       const int64_t* __restrict__ packed_info, long long n_rays, long long n,
                       long long n_tiles, float* __restrict__ colors, float* __restrict__ opacity,
                       float* __restrict__ depth, unsigned long long* status, uint32_t epoch,
                       bool vec_ok) {
    __shared__ TileSmem sm;
    __shared__ float s_tail[kTile];
    const TileGeom<false> g(n, n_tiles);
    float w[kVec][4];
#pragma unroll
    for (int k = 0; k < kVec; ++k) g.load(weights, vec_ok, k, 0.0f, w[k]);
    const SegSpec seg{packed_info, n_rays, 0};
    scatter_heads<false, true, true>(seg, g.base, n, sm.flags, sm.range);
    uint32_t f[kVec];
#pragma unroll
    for (int k = 0; k < kVec; ++k) f[k] = vec_flags(sm.flags, k);
    float p[kVec][4];
    {   // colours: one channel at a time keeps the register footprint of one scan
        float x[kVec][4][3];
#pragma unroll
        for (int k = 0; k < kVec; ++k) load_rgb<false>(g, rgbs, vec_ok, k, x[k]);
#pragma unroll
        for (int d = 0; d < 3; ++d) {
#pragma unroll
            for (int k = 0; k < kVec; ++k)
#pragma unroll
                for (int c = 0; c < 4; ++c) p[k][c] = __fmul_rn(w[k][c], x[k][c][d]);
            reduce_channel(p, f, g, d, status, epoch, sm, s_tail, packed_info, colors, 3, d);
        }
    }
    // opacity
#pragma unroll
    for (int k = 0; k < kVec; ++k)
#pragma unroll
        for (int c = 0; c < 4; ++c) p[k][c] = w[k][c];
    reduce_channel(p, f, g, 3, status, epoch, sm, s_tail, packed_info, opacity, 1, 0);
    // depth: w * (t_starts + t_ends) / 2   (volrend.py:148: (t_starts + t_ends)[..., None] / 2.0)
#pragma unroll
    for (int k = 0; k < kVec; ++k) {
        float ts[4], te[4];
        g.load(t_starts, vec_ok, k, 0.0f, ts);
        g.load(t_ends, vec_ok, k, 0.0f, te);
#pragma unroll
        for (int c = 0; c < 4; ++c) p[k][c] = __fmul_rn(w[k][c], __fmul_rn(__fadd_rn(ts[c], te[c]), 0.5f));
    }
    reduce_channel(p, f, g, 4, status, epoch, sm, s_tail, packed_info, depth, 1, 0);
}

struct CompositeBwdArgs {
    const float* g_colors;   // [n_rays, 3]  (may be null = zero)
    const float* g_opacity;  // [n_rays]
    const float* g_depth;    // [n_rays]     gradient wrt sum w*tmid (before normalisation)
    const float* g_weights;  // optional extra upstream grads on the per-sample outputs
    const float* g_trans;
    const float* g_alphas;
    const float* alphas;     // alpha mode input
    const float* t_starts;   // always given (tmid); density mode also uses them for delta
    const float* t_ends;
    const float* sigmas;     // density mode input
    const float* rgbs;       // [n, 3]
    const float* trans;      // saved forward output
    const int64_t* ray_indices;
    float* g_in;             // grad wrt sigmas / alphas
    float* g_rgbs;           // [n, 3] (may be null)
    SegSpec seg;
    long long n, n_tiles;
    unsigned long long* status;
    uint32_t epoch;
    bool vec_ok;
};

template <bool DENSITY>
__global__ void __launch_bounds__(kThreads) composite_bwd_kernel(const CompositeBwdArgs p) {
    __shared__ TileSmem sm;
    const TileGeom<true> g(p.n, p.n_tiles);
    float v[kVec][4];    // (g_T + g_w*a) * T, reverse-exclusive-summed
    float dir[kVec][4];  // direct term
    float sc[kVec][4];   // alpha: max(1-a, 1e-10) ; density: te - ts
#pragma unroll
    for (int k = 0; k < kVec; ++k) {
        float T[4], ts[4], te[4], a[4], e[4], gw[4], gT[4];
        g.load(p.trans, p.vec_ok, k, 0.0f, T);
        g.load(p.t_starts, p.vec_ok, k, 0.0f, ts);
        g.load(p.t_ends, p.vec_ok, k, 0.0f, te);
        if (DENSITY) {
            float s[4];
            g.load(p.sigmas, p.vec_ok, k, 0.0f, s);
#pragma unroll
            for (int c = 0; c < 4; ++c) {
                sc[k][c] = __fsub_rn(te[c], ts[c]);
                e[c] = expf(-__fmul_rn(s[c], sc[k][c]));
                a[c] = __fsub_rn(1.0f, e[c]);
            }
        } else {
            g.load(p.alphas, p.vec_ok, k, 0.0f, a);
#pragma unroll
            for (int c = 0; c < 4; ++c) { sc[k][c] = fmaxf(__fsub_rn(1.0f, a[c]), 1e-10f); e[c] = 1.0f; }
        }
        if (p.g_weights) g.load(p.g_weights, p.vec_ok, k, 0.0f, gw);
        else { gw[0] = gw[1] = gw[2] = gw[3] = 0.0f; }
        if (p.g_trans) g.load(p.g_trans, p.vec_ok, k, 0.0f, gT);
        else { gT[0] = gT[1] = gT[2] = gT[3] = 0.0f; }
        // per-ray rows through ray_indices (logical component order is reversed in REV tiles)
        float x[4][3];
        load_rgb<true>(g, p.rgbs, p.vec_ok, k, x);
        const long long gi = g.base + g.voff(k);
        float grgb[4][3];
#pragma unroll
        for (int c = 0; c < 4; ++c) {
            const long long i = gi + (3 - c);
            float gc0 = 0.0f, gc1 = 0.0f, gc2 = 0.0f, go = 0.0f, gd = 0.0f;
            if (i < p.n) {
                const long long r = __ldg(p.ray_indices + i);
                if (p.g_colors) { gc0 = __ldg(p.g_colors + 3 * r); gc1 = __ldg(p.g_colors + 3 * r + 1); gc2 = __ldg(p.g_colors + 3 * r + 2); }
                if (p.g_opacity) go = __ldg(p.g_opacity + r);
                if (p.g_depth) gd = __ldg(p.g_depth + r);
            }
            const float wgt = __fmul_rn(T[c], a[c]);
            const float tmid = __fmul_rn(__fadd_rn(ts[c], te[c]), 0.5f);
            // same association as autograd's accumulation of the three accumulate backward terms
            float acc = __fadd_rn(__fadd_rn(__fmul_rn(gc0, x[c][0]), __fmul_rn(gc1, x[c][1])), __fmul_rn(gc2, x[c][2]));
            acc = __fadd_rn(acc, go);
            acc = __fadd_rn(acc, __fmul_rn(gd, tmid));
            gw[c] = __fadd_rn(gw[c], acc);
            grgb[c][0] = __fmul_rn(wgt, gc0); grgb[c][1] = __fmul_rn(wgt, gc1); grgb[c][2] = __fmul_rn(wgt, gc2);
        }
        if (p.g_rgbs) {
            if (p.vec_ok && gi + 3 < p.n) {
                float flat[12];
#pragma unroll
                for (int c = 0; c < 4; ++c)
#pragma unroll
                    for (int d = 0; d < 3; ++d) flat[(3 - c) * 3 + d] = grgb[c][d];
                float* q = p.g_rgbs + gi * 3;
#pragma unroll
                for (int t = 0; t < 3; ++t)
                    stg_stream4(q + 4 * t, make_float4(flat[4 * t], flat[4 * t + 1], flat[4 * t + 2], flat[4 * t + 3]));
            } else {
#pragma unroll
                for (int c = 0; c < 4; ++c) {
                    const long long i = gi + (3 - c);
                    if (i < p.n) {
#pragma unroll
                        for (int d = 0; d < 3; ++d) stg_stream(p.g_rgbs + i * 3 + d, grgb[c][d]);
                    }
                }
            }
        }
#pragma unroll
        for (int c = 0; c < 4; ++c) {
            dir[k][c] = __fmul_rn(gw[c], T[c]);
            if (DENSITY) dir[k][c] = __fmul_rn(dir[k][c], e[c]);
            v[k][c] = __fmul_rn(__fadd_rn(gT[c], __fmul_rn(gw[c], a[c])), T[c]);
        }
        if (DENSITY && p.g_alphas) {
            float ga[4];
            g.load(p.g_alphas, p.vec_ok, k, 0.0f, ga);
#pragma unroll
            for (int c = 0; c < 4; ++c)
                dir[k][c] = __fmul_rn(__fadd_rn(ga[c], __fmul_rn(gw[c], T[c])), e[c]);
        }
    }
    scatter_heads<true>(p.seg, g.base, p.n, sm.flags, sm.range);
    uint32_t f[kVec];
#pragma unroll
    for (int k = 0; k < kVec; ++k) f[k] = vec_flags(sm.flags, k);
    float E[kVec];
    tile_segscan<SumOp>(v, f, E, p.status, 0, p.epoch, sm);
#pragma unroll
    for (int k = 0; k < kVec; ++k) {
        float R[4], o[4];
        finish_exclusive<SumOp>(v[k], f[k], E[k], R);
#pragma unroll
        for (int c = 0; c < 4; ++c) {
            if (DENSITY) o[c] = __fmul_rn(__fsub_rn(dir[k][c], R[c]), sc[k][c]);
            else o[c] = __fsub_rn(dir[k][c], __fdiv_rn(R[c], sc[k][c]));
        }
        g.store(p.g_in, p.vec_ok, k, o);
    }
}

static int comp_prep_ws(void* ws, size_t ws_bytes, long long n, uint32_t& epoch, cudaStream_t s) {
    const size_t need = nfb_workspace_bytes(n);
    if (ws == nullptr || ws_bytes < need) {
        set_error("workspace too small: need %zu bytes, got %zu", need, ws_bytes);
        return NFB_ERR_WORKSPACE;
    }
    if (epoch == 0) {
        if (cudaMemsetAsync(ws, 0, need, s) != cudaSuccess) { set_error("workspace memset failed"); return NFB_ERR_CUDA; }
        epoch = kEpochMemset;
    } else if (epoch >= kEpochMemset) {
        set_error("epoch must be in [1, 2^30-2]");
        return NFB_ERR_INVALID_ARGUMENT;
    }
    return NFB_OK;
}

static bool comp_aligned(std::initializer_list<const void*> ps) {
    for (const void* q : ps)
        if (q != nullptr && !aligned16(q)) return false;
    return true;
}

}  // namespace nfb

extern "C" {

NFB_API int nfb_composite_fwd(const float* weights, const float* rgbs, const float* t_starts,
                              const float* t_ends, const int64_t* packed_info, int64_t n_rays,
                              int64_t n, float* colors, float* opacity, float* depth,
                              void* workspace, size_t workspace_bytes, uint32_t epoch,
                              nfb_stream_t stream) {
    using namespace nfb;
    NFB_CHECK_ARG(n >= 0 && n_rays >= 0, "negative size");
    if (n_rays == 0) return NFB_OK;
    NFB_CHECK_ARG(colors && opacity && depth, "null output");
    cudaStream_t s = static_cast<cudaStream_t>(stream);
    if (n == 0) {
        cudaMemsetAsync(colors, 0, sizeof(float) * 3 * n_rays, s);
        cudaMemsetAsync(opacity, 0, sizeof(float) * n_rays, s);
        cudaMemsetAsync(depth, 0, sizeof(float) * n_rays, s);
        return NFB_OK;
    }
    NFB_CHECK_ARG(weights && rgbs && t_starts && t_ends && packed_info, "null pointer");
    NFB_CHECK_ARG(aligned16(packed_info), "packed_info must be 16-byte aligned");
    int rc = comp_prep_ws(workspace, workspace_bytes, n, epoch, s);
    if (rc != NFB_OK) return rc;
    const long long nt = num_tiles(n);
    const bool vec_ok = comp_aligned({weights, rgbs, t_starts, t_ends});
    composite_accum_kernel<<<(unsigned)nt, kThreads, 0, s>>>(
        weights, rgbs, t_starts, t_ends, packed_info, n_rays, n, nt, colors, opacity, depth,
        static_cast<unsigned long long*>(workspace), epoch, vec_ok);
    count_launch();
    NFB_CHECK_LAUNCH("composite_fwd");
    return NFB_OK;
}

NFB_API int nfb_composite_bwd(int from_density, const float* g_colors, const float* g_opacity,
                              const float* g_depth, const float* g_weights, const float* g_trans,
                              const float* g_alphas, const float* alphas, const float* t_starts,
                              const float* t_ends, const float* sigmas, const float* rgbs,
                              const float* trans, const int64_t* ray_indices,
                              const int64_t* packed_info, int64_t n_rays, int64_t n, float* g_in,
                              float* g_rgbs, void* workspace, size_t workspace_bytes,
                              uint32_t epoch, nfb_stream_t stream) {
    using namespace nfb;
    NFB_CHECK_ARG(n >= 0 && n_rays >= 0, "negative size");
    if (n == 0) return NFB_OK;
    NFB_CHECK_ARG(packed_info && ray_indices && t_starts && t_ends && rgbs && trans && g_in, "null pointer");
    NFB_CHECK_ARG(aligned16(packed_info), "packed_info must be 16-byte aligned");
    if (from_density) NFB_CHECK_ARG(sigmas != nullptr, "density mode needs sigmas");
    else NFB_CHECK_ARG(alphas != nullptr, "alpha mode needs alphas");
    cudaStream_t s = static_cast<cudaStream_t>(stream);
    int rc = comp_prep_ws(workspace, workspace_bytes, n, epoch, s);
    if (rc != NFB_OK) return rc;
    CompositeBwdArgs a{};
    a.g_colors = g_colors; a.g_opacity = g_opacity; a.g_depth = g_depth;
    a.g_weights = g_weights; a.g_trans = g_trans; a.g_alphas = g_alphas;
    a.alphas = alphas; a.t_starts = t_starts; a.t_ends = t_ends; a.sigmas = sigmas; a.rgbs = rgbs;
    a.trans = trans; a.ray_indices = ray_indices; a.g_in = g_in; a.g_rgbs = g_rgbs;
    a.seg = SegSpec{packed_info, n_rays, 0};
    a.n = n; a.n_tiles = num_tiles(n);
    a.status = static_cast<unsigned long long*>(workspace); a.epoch = epoch;
    a.vec_ok = comp_aligned({g_weights, g_trans, g_alphas, alphas, t_starts, t_ends, sigmas, rgbs, trans, g_in, g_rgbs});
    if (from_density) composite_bwd_kernel<true><<<(unsigned)a.n_tiles, kThreads, 0, s>>>(a);
    else composite_bwd_kernel<false><<<(unsigned)a.n_tiles, kThreads, 0, s>>>(a);
    count_launch();
    NFB_CHECK_LAUNCH("composite_bwd");
    return NFB_OK;
}

}  // extern "C"
