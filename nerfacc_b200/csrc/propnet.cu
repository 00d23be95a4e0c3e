// propnet.cu — the batched-row glue of PropNetEstimator as kernels.
//
// Between two importance_sampling calls the reference's PropNetEstimator.sampling runs, per
// level and over [n_rays, n] rows (nerfacc/estimators/prop_net.py:96-114): the s -> t transform
// (5 elementwise kernels, folded into nfb_importance_sampling here), the user's proposal network,
// render_transmittance_from_density (3 elementwise kernels + cumsum on two strided views) and
// `1 - cat([trans, 0])` (2 more) — each a full pass over >= 1 GiB at 2^20 rays.  Its loss
// (_pdf_loss, prop_net.py:232-256) is searchsorted + two int64 gathers + 5 elementwise kernels,
// and backward a scatter_add with atomics.  Here:
//   cdf_from_density_{fwd,bwd}   t_vals [R,n+1], sigmas [R,n] -> cdfs [R,n+1] = 1 - [T, 0],
//                                T_i = exp(-sum_{j<i} sigma_j (t_{j+1} - t_j)); one warp per row,
//                                the row in registers, one pass each way (12 / 16 B per sample);
//   pdf_loss_{fwd,bwd}           searchsorted of the query edges in the key edges, envelope mass
//                                w_outer, loss = clip(w - w_outer, 0)^2 / (w + eps) per query bin;
//                                backward writes the dense gradient row of cdfs_key without atomics
//                                (the ids are monotone along the row, so the bins that hit one key
//                                edge sit in neighbouring lanes: a segmented shuffle sum).
// Elementwise arithmetic is the reference's torch sequence op for op (round-to-nearest each).
#include "common.cuh"

namespace nfb {

constexpr int kRowWarps = 8;

// ------------------------------------------------------------------ cdf from density
// lane l of the warp holds elements l, l+32, ... of its row ("chunks" of 32)
template <int NC>
__global__ void __launch_bounds__(kRowWarps * 32)
cdf_from_density_fwd_kernel(const float* __restrict__ t_vals, const float* __restrict__ sigmas,
                            float* __restrict__ cdfs, long long n_rays, int n) {
    const int lane = lane_id();
    const long long r = (long long)blockIdx.x * kRowWarps + warp_id();
    if (r >= n_rays) return;
    const float* __restrict__ t = t_vals + r * (n + 1);
    const float* __restrict__ s = sigmas + r * n;
    float* __restrict__ c = cdfs + r * (n + 1);
    float sd[NC];
#pragma unroll
    for (int k = 0; k < NC; ++k) {
        const int i = 32 * k + lane;
        float v = 0.0f;
        if (i < n) v = __fmul_rn(ldg_stream(s + i), __fsub_rn(__ldg(t + i + 1), __ldg(t + i)));
        sd[k] = v;
    }
    float carry = 0.0f;  // sum of everything before this chunk
#pragma unroll
    for (int k = 0; k < NC; ++k) {
        float x = sd[k];
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) {
            const float y = __shfl_up_sync(kFull, x, d);
            if (lane >= d) x = __fadd_rn(y, x);
        }
        const float incl = __fadd_rn(carry, x);
        float excl = __shfl_up_sync(kFull, incl, 1);
        if (lane == 0) excl = carry;
        carry = __shfl_sync(kFull, incl, 31);
        const int i = 32 * k + lane;
        if (i < n) stg_stream(c + i, __fsub_rn(1.0f, expf(-excl)));
    }
    if (lane == 0) stg_stream(c + n, 1.0f);  // 1 - 0: the closing edge (prop_net.py:112-114)
}

// g_sigmas_j = (t_{j+1} - t_j) * sum_{j < i <= n-1} g_cdfs_i * T_i   (g_cdfs_n meets a constant)
template <int NC>
__global__ void __launch_bounds__(kRowWarps * 32)
cdf_from_density_bwd_kernel(const float* __restrict__ t_vals, const float* __restrict__ sigmas,
                            const float* __restrict__ g_cdfs, float* __restrict__ g_sigmas,
                            long long n_rays, int n) {
    const int lane = lane_id();
    const long long r = (long long)blockIdx.x * kRowWarps + warp_id();
    if (r >= n_rays) return;
    const float* __restrict__ t = t_vals + r * (n + 1);
    const float* __restrict__ s = sigmas + r * n;
    const float* __restrict__ g = g_cdfs + r * (n + 1);
    float* __restrict__ o = g_sigmas + r * n;
    float dt[NC], v[NC];
#pragma unroll
    for (int k = 0; k < NC; ++k) {
        const int i = 32 * k + lane;
        dt[k] = 0.0f;
        v[k] = 0.0f;
        if (i < n) {
            dt[k] = __fsub_rn(__ldg(t + i + 1), __ldg(t + i));
            v[k] = __fmul_rn(ldg_stream(s + i), dt[k]);
        }
    }
    // forward: T_i, then a_i = g_cdfs_i * T_i
    float carry = 0.0f;
#pragma unroll
    for (int k = 0; k < NC; ++k) {
        float x = v[k];
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) {
            const float y = __shfl_up_sync(kFull, x, d);
            if (lane >= d) x = __fadd_rn(y, x);
        }
        const float incl = __fadd_rn(carry, x);
        float excl = __shfl_up_sync(kFull, incl, 1);
        if (lane == 0) excl = carry;
        carry = __shfl_sync(kFull, incl, 31);
        const int i = 32 * k + lane;
        v[k] = (i < n) ? __fmul_rn(ldg_stream(g + i), expf(-excl)) : 0.0f;
    }
    // reverse exclusive sum of a over the row
    carry = 0.0f;  // sum of everything after this chunk
#pragma unroll
    for (int k = NC - 1; k >= 0; --k) {
        float x = v[k];
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) {
            const float y = __shfl_down_sync(kFull, x, d);
            if (lane + d < 32) x = __fadd_rn(x, y);
        }
        const float incl = __fadd_rn(x, carry);
        float excl = __shfl_down_sync(kFull, incl, 1);
        if (lane == 31) excl = carry;
        carry = __shfl_sync(kFull, incl, 0);
        const int i = 32 * k + lane;
        if (i < n) stg_stream(o + i, __fmul_rn(excl, dt[k]));
    }
}

// ------------------------------------------------------------------ proposal loss
struct PdfLossArgs {
    const float* q_vals;   // [R, Q]   query edges
    const float* q_cdfs;   // [R, Q]
    const float* k_vals;   // [R, K]   key edges
    const float* k_cdfs;   // [R, K]
    const float* g_loss;   // [R, Q-1] (backward)
    float* loss;           // [R, Q-1] (forward)
    float* g_k_cdfs;       // [R, K]   (backward)
    long long n_rays;
    int Q, K;
    float eps;
};

// A CTA owns kRowWarps consecutive rays: their key rows are one contiguous span, staged with
// block-wide 128-bit loads; then one warp per ray.  Query edges are taken 32 at a time with a
// stride of 31, so that lane l holds edge e0 + l and bin e0 + l (l < 31) finds its right edge's id
// in the next lane.  The search is the bounds-check-free ladder of common.cuh.
// Backward: every bin's +-d goes to the key edge it gathered from; the ids are non-decreasing along
// the row, so lanes that hit the same edge are neighbours: a segmented shuffle sum, one writer per
// edge, rows accumulated in shared memory in a fixed order (deterministic) and written once.
// Shared memory: [kRowWarps][K] key vals, [kRowWarps][K] key cdfs, (backward) [kRowWarps][K] gradient rows.
template <bool BWD>
__global__ void __launch_bounds__(kRowWarps * 32) pdf_loss_kernel(const PdfLossArgs p) {
    extern __shared__ float s_loss[];
    const int lane = lane_id(), w = warp_id(), tid = threadIdx.x;
    const int Q = p.Q, K = p.K, last = K - 1, nb = Q - 1;
    float* s_kv = s_loss;
    float* s_kc = s_kv + kRowWarps * K;
    float* s_g = s_kc + kRowWarps * K;
    int top = 1, ltop = 0;  // largest power of two <= max(last, 1)
    while (2 * top <= last) { top <<= 1; ++ltop; }
    const long long n_tiles = (p.n_rays + kRowWarps - 1) / kRowWarps;
    for (long long tile = blockIdx.x; tile < n_tiles; tile += gridDim.x) {
        const long long r0 = tile * kRowWarps;
        const int rays = (int)((p.n_rays - r0 < kRowWarps) ? (p.n_rays - r0) : kRowWarps);
        const float* __restrict__ gv = p.k_vals + r0 * K;
        const float* __restrict__ gc = p.k_cdfs + r0 * K;
        const int total = rays * K;
        if (((reinterpret_cast<uintptr_t>(gv) | reinterpret_cast<uintptr_t>(gc)) & 15u) == 0) {
            const int n4 = total >> 2;
            for (int i = tid; i < n4; i += kRowWarps * 32) {
                reinterpret_cast<float4*>(s_kv)[i] = ldg_stream4(gv + 4 * i);
                reinterpret_cast<float4*>(s_kc)[i] = ldg_stream4(gc + 4 * i);
            }
            for (int i = 4 * n4 + tid; i < total; i += kRowWarps * 32) { s_kv[i] = ldg_stream(gv + i); s_kc[i] = ldg_stream(gc + i); }
        } else {
            for (int i = tid; i < total; i += kRowWarps * 32) { s_kv[i] = ldg_stream(gv + i); s_kc[i] = ldg_stream(gc + i); }
        }
        if (BWD)
            for (int i = tid; i < total; i += kRowWarps * 32) s_g[i] = 0.0f;
        __syncthreads();
        if (w < rays) {
            const long long r = r0 + w;
            const float* kv = s_kv + w * K;
            const float* kc = s_kc + w * K;
            float* gk = s_g + w * K;
            const float* __restrict__ qv = p.q_vals + r * Q;
            const float* __restrict__ qc = p.q_cdfs + r * Q;
            for (int e0 = 0; e0 < nb; e0 += 31) {
                const int e = e0 + lane;  // my edge; my bin if lane < 31
                const bool on = e < Q;
                const float v = on ? __ldg(qv + e) : 0.0f;
                const float c_lo = on ? __ldg(qc + e) : 0.0f;
                // ids of nfb_searchsorted (csrc/pdf.cu:245-286): q = upper_bound(key[0..last), v)
                const int qq = last > 0 ? ladder_upper_bound(kv, last, top, ltop, v) : 0;
                const int left = max(qq - 1, 0), right = qq;  // (qq <= last)
                const int right_n = __shfl_down_sync(kFull, right, 1);
                const float c_hi = __shfl_down_sync(kFull, c_lo, 1);
                const bool bin = lane < 31 && e < nb;
                float d = 0.0f;
                if (bin) {
                    const float wq = __fsub_rn(c_hi, c_lo);                        // prop_net.py:244
                    const float w_outer = __fsub_rn(kc[right_n], kc[left]);        // :252-254
                    const float c = fmaxf(__fsub_rn(wq, w_outer), 0.0f);           // clip(w - w_outer, min=0)
                    const float den = __fadd_rn(wq, p.eps);
                    if (!BWD) {
                        stg_stream(p.loss + r * nb + e, __fdiv_rn(__fmul_rn(c, c), den));  // ** 2 / (w + eps)
                    } else {
                        // autograd of the above: g / den, times 2 c (pow), through clip and the minus
                        d = __fmul_rn(__fdiv_rn(ldg_stream(p.g_loss + r * nb + e), den), __fmul_rn(2.0f, c));
                    }
                }
                if (BWD) {
                    // -d to cdfs_key[right id of the bin's right edge], +d to cdfs_key[left id of its left edge]
#pragma unroll
                    for (int pass = 0; pass < 2; ++pass) {
                        const int id = bin ? (pass == 0 ? right_n : left) : -1 - lane;
                        float x = pass == 0 ? -d : d;
#pragma unroll
                        for (int dd = 1; dd < 32; dd <<= 1) {
                            const float y = __shfl_up_sync(kFull, x, dd);
                            const int idy = __shfl_up_sync(kFull, id, dd);
                            if (lane >= dd && idy == id) x = __fadd_rn(y, x);
                        }
                        const int idn = __shfl_down_sync(kFull, id, 1);
                        // (one writer per edge and pass when the query edges are sorted; atomic so that
                        // unsorted input cannot lose an update)
                        if (bin && (lane == 31 || idn != id)) atomicAdd(gk + id, x);
                        __syncwarp();
                    }
                }
            }
            if (BWD) {
                __syncwarp();
                float* __restrict__ out = p.g_k_cdfs + r * K;
                for (int k = lane; k < K; k += 32) stg_stream(out + k, gk[k]);
            }
        }
        __syncthreads();
    }
}

static inline unsigned row_blocks(long long n_rays) {
    return (unsigned)ceil_div<long long>(n_rays, kRowWarps);
}

}  // namespace nfb

extern "C" {

#define NFB_ROW_DISPATCH(KERNEL, NC, ...)                                              \
    switch (NC) {                                                                      \
        case 1: KERNEL<1><<<grid, kRowWarps * 32, 0, st>>>(__VA_ARGS__); break;        \
        case 2: KERNEL<2><<<grid, kRowWarps * 32, 0, st>>>(__VA_ARGS__); break;        \
        case 3: KERNEL<3><<<grid, kRowWarps * 32, 0, st>>>(__VA_ARGS__); break;        \
        case 4: KERNEL<4><<<grid, kRowWarps * 32, 0, st>>>(__VA_ARGS__); break;        \
        case 5: case 6: KERNEL<6><<<grid, kRowWarps * 32, 0, st>>>(__VA_ARGS__); break; \
        case 7: case 8: KERNEL<8><<<grid, kRowWarps * 32, 0, st>>>(__VA_ARGS__); break; \
        default: KERNEL<16><<<grid, kRowWarps * 32, 0, st>>>(__VA_ARGS__); break;      \
    }

NFB_API int nfb_cdf_from_density_fwd(const float* t_vals, const float* sigmas, int64_t n_rays, int64_t n,
                                     float* cdfs, nfb_stream_t stream) {
    using namespace nfb;
    NFB_CHECK_ARG(n_rays >= 0 && n >= 0, "negative size");
    if (n_rays == 0) return NFB_OK;
    NFB_CHECK_ARG(n >= 1 && n <= 512, "rows of 1..512 samples");
    NFB_CHECK_ARG(t_vals && sigmas && cdfs, "null pointer");
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    const unsigned grid = row_blocks(n_rays);
    const int nc = (int)((n + 31) / 32);
    NFB_ROW_DISPATCH(cdf_from_density_fwd_kernel, nc, t_vals, sigmas, cdfs, n_rays, (int)n)
    count_launch();
    NFB_CHECK_LAUNCH("cdf_from_density_fwd");
    return NFB_OK;
}

NFB_API int nfb_cdf_from_density_bwd(const float* t_vals, const float* sigmas, const float* g_cdfs,
                                     int64_t n_rays, int64_t n, float* g_sigmas, nfb_stream_t stream) {
    using namespace nfb;
    NFB_CHECK_ARG(n_rays >= 0 && n >= 0, "negative size");
    if (n_rays == 0) return NFB_OK;
    NFB_CHECK_ARG(n >= 1 && n <= 512, "rows of 1..512 samples");
    NFB_CHECK_ARG(t_vals && sigmas && g_cdfs && g_sigmas, "null pointer");
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    const unsigned grid = row_blocks(n_rays);
    const int nc = (int)((n + 31) / 32);
    NFB_ROW_DISPATCH(cdf_from_density_bwd_kernel, nc, t_vals, sigmas, g_cdfs, g_sigmas, n_rays, (int)n)
    count_launch();
    NFB_CHECK_LAUNCH("cdf_from_density_bwd");
    return NFB_OK;
}

static int launch_pdf_loss(bool bwd, const nfb::PdfLossArgs& a, cudaStream_t st) {
    using namespace nfb;
    const size_t smem = (size_t)kRowWarps * a.K * (bwd ? 3 : 2) * sizeof(float);
    if (smem > 200 * 1024) {
        set_error("pdf_loss: rows too long for shared memory (K = %d)", a.K);
        return NFB_ERR_INVALID_ARGUMENT;
    }
    long long blocks = ceil_div<long long>(a.n_rays, kRowWarps);
    const long long cap = (long long)sm_count() * 1024;
    if (blocks > cap) blocks = cap;
    if (bwd) {
        if (smem > 48 * 1024) cudaFuncSetAttribute(pdf_loss_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        pdf_loss_kernel<true><<<(unsigned)blocks, kRowWarps * 32, smem, st>>>(a);
    } else {
        if (smem > 48 * 1024) cudaFuncSetAttribute(pdf_loss_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        pdf_loss_kernel<false><<<(unsigned)blocks, kRowWarps * 32, smem, st>>>(a);
    }
    count_launch();
    NFB_CHECK_LAUNCH("pdf_loss");
    return NFB_OK;
}

NFB_API int nfb_pdf_loss_fwd(const float* q_vals, const float* q_cdfs, int64_t n_q, const float* k_vals,
                             const float* k_cdfs, int64_t n_k, int64_t n_rays, float eps, float* loss,
                             nfb_stream_t stream) {
    using namespace nfb;
    NFB_CHECK_ARG(n_rays >= 0, "negative size");
    if (n_rays == 0) return NFB_OK;
    NFB_CHECK_ARG(n_q >= 2 && n_k >= 1 && n_q < (1 << 20) && n_k < (1 << 20), "need n_q >= 2 query edges and n_k >= 1 key edges");
    NFB_CHECK_ARG(q_vals && q_cdfs && k_vals && k_cdfs && loss, "null pointer");
    PdfLossArgs a{};
    a.q_vals = q_vals; a.q_cdfs = q_cdfs; a.k_vals = k_vals; a.k_cdfs = k_cdfs; a.loss = loss;
    a.n_rays = n_rays; a.Q = (int)n_q; a.K = (int)n_k; a.eps = eps;
    return launch_pdf_loss(false, a, static_cast<cudaStream_t>(stream));
}

NFB_API int nfb_pdf_loss_bwd(const float* q_vals, const float* q_cdfs, int64_t n_q, const float* k_vals,
                             const float* k_cdfs, int64_t n_k, int64_t n_rays, float eps, const float* g_loss,
                             float* g_k_cdfs, nfb_stream_t stream) {
    using namespace nfb;
    NFB_CHECK_ARG(n_rays >= 0, "negative size");
    if (n_rays == 0) return NFB_OK;
    NFB_CHECK_ARG(n_q >= 2 && n_k >= 1 && n_q < (1 << 20) && n_k < (1 << 20), "need n_q >= 2 query edges and n_k >= 1 key edges");
    NFB_CHECK_ARG(q_vals && q_cdfs && k_vals && k_cdfs && g_loss && g_k_cdfs, "null pointer");
    PdfLossArgs a{};
    a.q_vals = q_vals; a.q_cdfs = q_cdfs; a.k_vals = k_vals; a.k_cdfs = k_cdfs; a.g_loss = g_loss;
    a.g_k_cdfs = g_k_cdfs; a.n_rays = n_rays; a.Q = (int)n_q; a.K = (int)n_k; a.eps = eps;
    return launch_pdf_loss(true, a, static_cast<cudaStream_t>(stream));
}

}  // extern "C"
