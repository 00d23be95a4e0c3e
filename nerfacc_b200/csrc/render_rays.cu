// render_rays.cu — nerfacc.rendering() for packed rays as ONE forward and ONE backward kernel.
//
// Reference composition (nerfacc/volrend.py:14-158): render_weight_from_density/alpha (5-7
// elementwise kernels around a packed scan), then accumulate_along_rays three times (rgb,
// opacity, depth: three [S,D] temporaries and three index_add_ passes), and autograd replays
// all of it backwards.  Here a group of L lanes (L = 4..32, chosen from the mean samples per
// ray) owns one ray and walks its samples L at a time:
//   forward   sd = sigma*(te-ts); a = 1-exp(-sd); T = exp(-exclusive_sum(sd)); w = T*a
//             (alpha mode: T = exclusive_prod(1-a)); writes w, T, a and accumulates
//             sum w*rgb, sum w, sum w*(ts+te)/2 in registers -> colors, opacity, depth.
//             36 B/sample of traffic (read 24, write 12) instead of 48 in two kernels.
//   backward  walks the ray from its last sample: the per-ray gradient rows are loaded once per
//             group (no ray_indices gather: 8 B/sample less), g_w is formed, g_rgb written, the
//             reverse exclusive sum of (g_T + g_w*a)*T carried in a register, g_sigma / g_alpha
//             written.  44 B/sample.
// Rays are independent, chunks are addressed through packed_info (start, count): the kernels do
// not need a canonical layout, a workspace or a look-back.  A ray much longer than the others
// is walked by its L lanes alone; the flat tile kernels (volrend.cu, composite.cu) remain the
// load-balanced path and the host picks them when the mean count per ray is large.
#include "common.cuh"

namespace nfb {

constexpr int kRayThreads = 256;

struct RenderRaysArgs {
    // inputs
    const float* __restrict__ alphas;    // alpha mode
    const float* __restrict__ t_starts;
    const float* __restrict__ t_ends;
    const float* __restrict__ sigmas;    // density mode
    const float* __restrict__ rgbs;      // [n, 3]
    const int64_t* __restrict__ packed_info;
    long long n_rays;
    // forward outputs
    float* __restrict__ weights;
    float* __restrict__ trans;
    float* __restrict__ alphas_out;  // density mode only
    float* __restrict__ colors;      // [n_rays, 3]
    float* __restrict__ opacity;     // [n_rays]
    float* __restrict__ depth;       // [n_rays]
    // backward inputs (each may be null = zero) and outputs
    const float* __restrict__ trans_in;
    const float* __restrict__ g_colors;
    const float* __restrict__ g_opacity;
    const float* __restrict__ g_depth;
    const float* __restrict__ g_weights;
    const float* __restrict__ g_trans;
    const float* __restrict__ g_alphas;
    float* __restrict__ g_in;
    float* __restrict__ g_rgbs;  // may be null
};

template <int L>
__device__ __forceinline__ unsigned group_mask(int lane) {
    if constexpr (L == 32) return 0xffffffffu;
    else return ((1u << L) - 1u) << ((lane / L) * L);
}

// W rays per group can be walked together (their loads issued back to back before the dependent
// math).  Measured on B200 at 35 samples/ray: W = 2 costs 80-110 registers and the lost
// occupancy outweighs the extra bytes in flight (fwd 0.53 ms vs 0.45 ms), so W = 1.
constexpr int kRaysPerGroup = 1;

template <int L, bool DENSITY>
__global__ void __launch_bounds__(kRayThreads) render_rays_fwd_kernel(const RenderRaysArgs p) {
    constexpr int G = 32 / L, W = kRaysPerGroup;
    const int lane = lane_id();
    const int sub = lane % L;
    const unsigned gm = group_mask<L>(lane);
    const long long warps = (long long)gridDim.x * (kRayThreads / 32);
    for (long long rb = (((long long)blockIdx.x * (kRayThreads / 32) + warp_id()) * G + lane / L) * W; rb < p.n_rays;
         rb += warps * G * W) {
        RayChunk c[W];
        float carry[W], acc[W][5];
        long long maxcnt = 0;
#pragma unroll
        for (int w = 0; w < W; ++w) {
            c[w].start = 0; c[w].cnt = 0;
            if (rb + w < p.n_rays) c[w] = load_chunk(p.packed_info, rb + w);
            carry[w] = DENSITY ? 0.0f : 1.0f;  // exclusive sum / product up to this step
#pragma unroll
            for (int k = 0; k < 5; ++k) acc[w][k] = 0.0f;
            maxcnt = c[w].cnt > maxcnt ? c[w].cnt : maxcnt;
        }
#pragma unroll 2
        for (long long i = 0; i < maxcnt; i += L) {
            bool in[W];
            float ts[W], te[W], fv[W], r0[W], r1[W], r2[W];
#pragma unroll
            for (int w = 0; w < W; ++w) {
                in[w] = i + sub < c[w].cnt;
                const long long s = c[w].start + i + sub;
                ts[w] = te[w] = fv[w] = r0[w] = r1[w] = r2[w] = 0.0f;
                if (in[w]) {
                    ts[w] = __ldg(p.t_starts + s);
                    te[w] = __ldg(p.t_ends + s);
                    fv[w] = __ldg((DENSITY ? p.sigmas : p.alphas) + s);
                    r0[w] = __ldg(p.rgbs + 3 * s);
                    r1[w] = __ldg(p.rgbs + 3 * s + 1);
                    r2[w] = __ldg(p.rgbs + 3 * s + 2);
                }
            }
#pragma unroll
            for (int w = 0; w < W; ++w) {
                if (i >= c[w].cnt) continue;  // uniform over the group
                const long long s = c[w].start + i + sub;
                float a, x;
                if (DENSITY) {
                    x = in[w] ? __fmul_rn(fv[w], __fsub_rn(te[w], ts[w])) : 0.0f;
                    a = __fsub_rn(1.0f, expf(-x));
                } else {
                    a = fv[w];
                    x = __fsub_rn(1.0f, a);  // out-of-range lanes: a = 0, x = 1
                }
                // inclusive scan of x over the group's lanes
                float inc = x;
#pragma unroll
                for (int d = 1; d < L; d <<= 1) {
                    const float y = __shfl_up_sync(gm, inc, d, L);
                    if (sub >= d) inc = DENSITY ? __fadd_rn(inc, y) : __fmul_rn(inc, y);
                }
                float before = __shfl_up_sync(gm, inc, 1, L);
                if (sub == 0) before = DENSITY ? 0.0f : 1.0f;
                const float excl = DENSITY ? __fadd_rn(carry[w], before) : __fmul_rn(carry[w], before);
                const float total = __shfl_sync(gm, inc, L - 1, L);
                carry[w] = DENSITY ? __fadd_rn(carry[w], total) : __fmul_rn(carry[w], total);
                const float T = DENSITY ? expf(-excl) : excl;
                const float wt = __fmul_rn(T, a);
                if (in[w]) {
                    p.weights[s] = wt;
                    p.trans[s] = T;
                    if (DENSITY) p.alphas_out[s] = a;
                    acc[w][0] = __fmaf_rn(wt, r0[w], acc[w][0]);
                    acc[w][1] = __fmaf_rn(wt, r1[w], acc[w][1]);
                    acc[w][2] = __fmaf_rn(wt, r2[w], acc[w][2]);
                    acc[w][3] = __fadd_rn(acc[w][3], wt);
                    acc[w][4] = __fmaf_rn(wt, __fmul_rn(__fadd_rn(ts[w], te[w]), 0.5f), acc[w][4]);
                }
            }
        }
#pragma unroll
        for (int w = 0; w < W; ++w) {
#pragma unroll
            for (int d = L / 2; d >= 1; d >>= 1)
#pragma unroll
                for (int k = 0; k < 5; ++k) acc[w][k] = __fadd_rn(acc[w][k], __shfl_xor_sync(gm, acc[w][k], d, L));
            const long long r = rb + w;
            if (sub == 0 && r < p.n_rays) {
                p.colors[3 * r] = acc[w][0];
                p.colors[3 * r + 1] = acc[w][1];
                p.colors[3 * r + 2] = acc[w][2];
                p.opacity[r] = acc[w][3];
                p.depth[r] = acc[w][4];
            }
        }
    }
}

template <int L, bool DENSITY>
__global__ void __launch_bounds__(kRayThreads) render_rays_bwd_kernel(const RenderRaysArgs p) {
    constexpr int G = 32 / L, W = kRaysPerGroup;
    const int lane = lane_id();
    const int sub = lane % L;
    const unsigned gm = group_mask<L>(lane);
    const long long warps = (long long)gridDim.x * (kRayThreads / 32);
    for (long long rb = (((long long)blockIdx.x * (kRayThreads / 32) + warp_id()) * G + lane / L) * W; rb < p.n_rays;
         rb += warps * G * W) {
        RayChunk c[W];
        float gc0[W], gc1[W], gc2[W], go[W], gd[W], carry[W];
        long long maxcnt = 0;
#pragma unroll
        for (int w = 0; w < W; ++w) {
            const long long r = rb + w;
            c[w].start = 0; c[w].cnt = 0;
            gc0[w] = gc1[w] = gc2[w] = go[w] = gd[w] = 0.0f;
            carry[w] = 0.0f;  // sum of v over the samples behind (later than) this step
            if (r < p.n_rays) {
                c[w] = load_chunk(p.packed_info, r);
                if (p.g_colors) { gc0[w] = __ldg(p.g_colors + 3 * r); gc1[w] = __ldg(p.g_colors + 3 * r + 1); gc2[w] = __ldg(p.g_colors + 3 * r + 2); }
                if (p.g_opacity) go[w] = __ldg(p.g_opacity + r);
                if (p.g_depth) gd[w] = __ldg(p.g_depth + r);
            }
            maxcnt = c[w].cnt > maxcnt ? c[w].cnt : maxcnt;
        }
        for (long long i = 0; i < maxcnt; i += L) {
            bool in[W];
            float ts[W], te[W], T[W], fv[W], r0[W], r1[W], r2[W], gw[W], gT[W], ga[W];
#pragma unroll
            for (int w = 0; w < W; ++w) {
                in[w] = i + sub < c[w].cnt;
                const long long s = c[w].start + (c[w].cnt - 1 - i - sub);  // lane 0 takes the last sample
                ts[w] = te[w] = T[w] = fv[w] = r0[w] = r1[w] = r2[w] = gw[w] = gT[w] = ga[w] = 0.0f;
                if (in[w]) {
                    ts[w] = __ldg(p.t_starts + s);
                    te[w] = __ldg(p.t_ends + s);
                    T[w] = __ldg(p.trans_in + s);
                    fv[w] = __ldg((DENSITY ? p.sigmas : p.alphas) + s);
                    r0[w] = __ldg(p.rgbs + 3 * s);
                    r1[w] = __ldg(p.rgbs + 3 * s + 1);
                    r2[w] = __ldg(p.rgbs + 3 * s + 2);
                    if (p.g_weights) gw[w] = __ldg(p.g_weights + s);
                    if (p.g_trans) gT[w] = __ldg(p.g_trans + s);
                    if (DENSITY && p.g_alphas) ga[w] = __ldg(p.g_alphas + s);
                }
            }
#pragma unroll
            for (int w = 0; w < W; ++w) {
                if (i >= c[w].cnt) continue;  // uniform over the group
                const long long s = c[w].start + (c[w].cnt - 1 - i - sub);
                float a, e = 1.0f, sc;
                if (DENSITY) {
                    sc = __fsub_rn(te[w], ts[w]);
                    e = expf(-__fmul_rn(fv[w], sc));
                    a = __fsub_rn(1.0f, e);
                } else {
                    a = fv[w];
                    sc = fmaxf(__fsub_rn(1.0f, a), 1e-10f);
                }
                const float wgt = __fmul_rn(T[w], a);
                const float tmid = __fmul_rn(__fadd_rn(ts[w], te[w]), 0.5f);
                // same association as autograd's accumulation of the three accumulate backward terms
                float t3 = __fadd_rn(__fadd_rn(__fmul_rn(gc0[w], r0[w]), __fmul_rn(gc1[w], r1[w])), __fmul_rn(gc2[w], r2[w]));
                t3 = __fadd_rn(t3, go[w]);
                t3 = __fadd_rn(t3, __fmul_rn(gd[w], tmid));
                const float g = __fadd_rn(gw[w], t3);
                if (in[w] && p.g_rgbs) {
                    p.g_rgbs[3 * s] = __fmul_rn(wgt, gc0[w]);
                    p.g_rgbs[3 * s + 1] = __fmul_rn(wgt, gc1[w]);
                    p.g_rgbs[3 * s + 2] = __fmul_rn(wgt, gc2[w]);
                }
                float dir;
                if (DENSITY) dir = __fmul_rn(__fadd_rn(ga[w], __fmul_rn(g, T[w])), e);
                else dir = __fmul_rn(g, T[w]);
                const float v = in[w] ? __fmul_rn(__fadd_rn(gT[w], __fmul_rn(g, a)), T[w]) : 0.0f;
                float inc = v;
#pragma unroll
                for (int d = 1; d < L; d <<= 1) {
                    const float y = __shfl_up_sync(gm, inc, d, L);
                    if (sub >= d) inc = __fadd_rn(inc, y);
                }
                float before = __shfl_up_sync(gm, inc, 1, L);
                if (sub == 0) before = 0.0f;
                const float R = __fadd_rn(carry[w], before);
                carry[w] = __fadd_rn(carry[w], __shfl_sync(gm, inc, L - 1, L));
                if (in[w]) p.g_in[s] = DENSITY ? __fmul_rn(__fsub_rn(dir, R), sc) : __fsub_rn(dir, __fdiv_rn(R, sc));
            }
        }
    }
}

static inline unsigned rays_grid(long long n_rays, int L) {
    const long long per_block = (long long)(kRayThreads / 32) * (32 / L) * kRaysPerGroup;
    long long b = ceil_div<long long>(n_rays, per_block);
    const long long cap = (long long)sm_count() * 8;
    if (b > cap) b = cap;
    if (b < 1) b = 1;
    return (unsigned)b;
}

// lanes per ray: the smallest group that walks a mean-length ray in one step.  Measured at 35
// samples/ray: L = 32 0.45 ms, 16 0.48, 8 0.51, 4 0.58 forward — the per-ray set-up (chunk load,
// the final 5-channel reduce, row stores) is paid once per ray whatever L is, and wider groups
// need fewer dependent steps per ray.
static inline int pick_group(long long n, long long n_rays, int forced) {
    if (forced == 4 || forced == 8 || forced == 16 || forced == 32) return forced;
    const double mean = (double)n / (double)(n_rays > 0 ? n_rays : 1);
    int L = 4;
    while (L < 32 && (double)L < mean) L <<= 1;
    return L;
}

template <bool BWD, bool DENSITY>
static void launch_rays(const RenderRaysArgs& a, int L, cudaStream_t s) {
    const unsigned g = rays_grid(a.n_rays, L);
#define NFB_RAYS_CASE(LL)                                                                      \
    case LL:                                                                                   \
        if (BWD) render_rays_bwd_kernel<LL, DENSITY><<<g, kRayThreads, 0, s>>>(a);             \
        else render_rays_fwd_kernel<LL, DENSITY><<<g, kRayThreads, 0, s>>>(a);                 \
        break;
    switch (L) {
        NFB_RAYS_CASE(4)
        NFB_RAYS_CASE(8)
        NFB_RAYS_CASE(16)
        default:
        NFB_RAYS_CASE(32)
    }
#undef NFB_RAYS_CASE
}

}  // namespace nfb

extern "C" {

NFB_API int nfb_render_rays_fwd(int from_density, const float* alphas, const float* t_starts,
                                const float* t_ends, const float* sigmas, const float* rgbs,
                                const int64_t* packed_info, int64_t n_rays, int64_t n,
                                float* weights, float* trans, float* alphas_out, float* colors,
                                float* opacity, float* depth, int lanes_per_ray,
                                nfb_stream_t stream) {
    using namespace nfb;
    NFB_CHECK_ARG(n >= 0 && n_rays >= 0, "negative size");
    if (n_rays == 0) return NFB_OK;
    NFB_CHECK_ARG(colors && opacity && depth && packed_info, "null pointer");
    NFB_CHECK_ARG(aligned16(packed_info), "packed_info must be 16-byte aligned");
    if (n > 0) {
        NFB_CHECK_ARG(t_starts && t_ends && rgbs && weights && trans, "null pointer");
        if (from_density) NFB_CHECK_ARG(sigmas && alphas_out, "density mode needs sigmas and alphas_out");
        else NFB_CHECK_ARG(alphas != nullptr, "alpha mode needs alphas");
    }
    RenderRaysArgs a{};
    a.alphas = alphas; a.t_starts = t_starts; a.t_ends = t_ends; a.sigmas = sigmas; a.rgbs = rgbs;
    a.packed_info = packed_info; a.n_rays = n_rays;
    a.weights = weights; a.trans = trans; a.alphas_out = alphas_out;
    a.colors = colors; a.opacity = opacity; a.depth = depth;
    const int L = pick_group(n, n_rays, lanes_per_ray);
    cudaStream_t s = static_cast<cudaStream_t>(stream);
    if (from_density) launch_rays<false, true>(a, L, s);
    else launch_rays<false, false>(a, L, s);
    count_launch();
    NFB_CHECK_LAUNCH("render_rays_fwd");
    return NFB_OK;
}

NFB_API int nfb_render_rays_bwd(int from_density, const float* g_colors, const float* g_opacity,
                                const float* g_depth, const float* g_weights, const float* g_trans,
                                const float* g_alphas, const float* alphas, const float* t_starts,
                                const float* t_ends, const float* sigmas, const float* rgbs,
                                const float* trans, const int64_t* packed_info, int64_t n_rays,
                                int64_t n, float* g_in, float* g_rgbs, int lanes_per_ray,
                                nfb_stream_t stream) {
    using namespace nfb;
    NFB_CHECK_ARG(n >= 0 && n_rays >= 0, "negative size");
    if (n == 0 || n_rays == 0) return NFB_OK;
    NFB_CHECK_ARG(packed_info && t_starts && t_ends && rgbs && trans && g_in, "null pointer");
    NFB_CHECK_ARG(aligned16(packed_info), "packed_info must be 16-byte aligned");
    if (from_density) NFB_CHECK_ARG(sigmas != nullptr, "density mode needs sigmas");
    else NFB_CHECK_ARG(alphas != nullptr, "alpha mode needs alphas");
    RenderRaysArgs a{};
    a.alphas = alphas; a.t_starts = t_starts; a.t_ends = t_ends; a.sigmas = sigmas; a.rgbs = rgbs;
    a.packed_info = packed_info; a.n_rays = n_rays; a.trans_in = trans;
    a.g_colors = g_colors; a.g_opacity = g_opacity; a.g_depth = g_depth;
    a.g_weights = g_weights; a.g_trans = g_trans; a.g_alphas = from_density ? g_alphas : nullptr;
    a.g_in = g_in; a.g_rgbs = g_rgbs;
    const int L = pick_group(n, n_rays, lanes_per_ray);
    cudaStream_t s = static_cast<cudaStream_t>(stream);
    if (from_density) launch_rays<true, true>(a, L, s);
    else launch_rays<true, false>(a, L, s);
    count_launch();
    NFB_CHECK_LAUNCH("render_rays_bwd");
    return NFB_OK;
}

}  // extern "C"
