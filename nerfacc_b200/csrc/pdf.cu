// pdf.cu — inverse-CDF importance sampling and per-ray searchsorted behind the C ABI.
//
// Replaces csrc/pdf.cu of the reference: importance_sampling_kernel + compute_intervels_kernel
// (pdf.cu:98-241, one thread per output sample, a global-memory binary search each, and a second
// kernel that re-reads the samples to make the edges) and searchsorted_kernel (pdf.cu:245-286).
// Here one WARP owns a ray:
//   * the ray's CDF and edge rows are staged once in shared memory (coalesced), every search
//     runs there;
//   * the sample centres go to shared memory as well, so the interval edges (midpoints of
//     neighbouring centres, clamped to the ray's first / last edge) come out of the same kernel;
//   * outputs are written as coalesced rows.
// Arithmetic is spelled with round-to-nearest intrinsics exactly as nvcc compiles the reference
// (u = fma(sid + bias, u_step, u_floor), t = fma(u - u_lower, scaling, t_lower), IEEE divisions,
// the 1e-10 degenerate-bin rule of pdf.cu:159-160), so positions agree to the last bit where the
// reference is well defined.
#include "common.cuh"

namespace nfb {

constexpr int kPdfWarps = 8;

// first index in [lo, hi) whose value is > val (pdf.cu:44-63), on any pointer
__device__ __forceinline__ int upper_bound(const float* data, int lo, int hi, float val) {
    while (lo < hi) {
        const int mid = lo + ((hi - lo) >> 1);
        if (!(data[mid] > val)) lo = mid + 1; else hi = mid;
    }
    return lo;
}

struct PdfArgs {
    const float* vals;            // input edges: [n_rays, n_in] or packed
    const float* cdfs;            // same layout
    const int64_t* in_info;       // packed input chunks or nullptr (batched, n_in per ray)
    long long n_in;               // batched row length
    const int64_t* n_out_per_ray; // optional per-ray sample counts (packed output), else n_out
    long long n_out;
    const int64_t* edge_info;     // packed output chunks (with n_out_per_ray), else rows of n_out+1
    const int64_t* sample_info;   //                                            else rows of n_out
    const float* bias;            // per-ray stratification offset in (0,1], or nullptr = 0.5
    float* out_edges;             // may be nullptr
    float* out_centres;
    long long n_rays;
    int smem_floats_per_warp;     // staging capacity; longer rows are searched in global memory
};

__global__ void __launch_bounds__(kPdfWarps * 32) importance_sampling_kernel(const PdfArgs p) {
    extern __shared__ float s_pdf[];
    const int lane = lane_id(), w = warp_id();
    float* s_row = s_pdf + (size_t)w * p.smem_floats_per_warp;
    for (long long r = (long long)blockIdx.x * kPdfWarps + w; r < p.n_rays;
         r += (long long)gridDim.x * kPdfWarps) {
        long long in0, cnt;
        if (p.in_info) { const RayChunk c = load_chunk(p.in_info, r); in0 = c.start; cnt = c.cnt; }
        else { in0 = r * p.n_in; cnt = p.n_in; }
        long long e0, s0;
        int n_out;
        if (p.n_out_per_ray) {
            n_out = (int)p.n_out_per_ray[r];
            e0 = p.edge_info[2 * r];
            s0 = p.sample_info[2 * r];
        } else {
            n_out = (int)p.n_out;
            e0 = r * (p.n_out + 1);
            s0 = r * p.n_out;
        }
        if (cnt <= 0 || n_out <= 0) continue;
        const int n_in = (int)cnt, last = n_in - 1;
        // stage the rows: [cdfs n_in][vals n_in][centres n_out]
        // (without a centres output the edges need the shared-memory copy: host guarantees it fits)
        const bool staged = 2 * n_in + n_out <= p.smem_floats_per_warp;
        const float* cdf = p.cdfs + in0;
        const float* val = p.vals + in0;
        float* cen = nullptr;
        if (staged) {
            for (int i = lane; i < n_in; i += 32) { s_row[i] = __ldg(cdf + i); s_row[n_in + i] = __ldg(val + i); }
            __syncwarp();
            cdf = s_row; val = s_row + n_in; cen = s_row + 2 * n_in;
        }
        const float u_floor = cdf[0], u_ceil = cdf[last];
        const float u_step = __fdiv_rn(__fsub_rn(u_ceil, u_floor), (float)n_out);
        const float b = p.bias ? __ldg(p.bias + r) : 0.5f;
        const float t_min = val[0], t_max = val[last];
        for (int sid = lane; sid < n_out; sid += 32) {
            const float u = __fmaf_rn(__fadd_rn((float)sid, b), u_step, u_floor);
            const int q = upper_bound(cdf, 0, last, u);
            const int p0 = max(min(q - 1, last), 0), p1 = max(min(q, last), 0);
            const float u_lower = cdf[p0], u_upper = cdf[p1], t_lower = val[p0], t_upper = val[p1];
            float t;
            if (__fsub_rn(u_upper, u_lower) < 1e-10f) {
                t = __fmul_rn(__fadd_rn(t_lower, t_upper), 0.5f);
            } else {
                const float scaling = __fdiv_rn(__fsub_rn(t_upper, t_lower), __fsub_rn(u_upper, u_lower));
                t = __fmaf_rn(__fsub_rn(u, u_lower), scaling, t_lower);
            }
            if (p.out_centres) p.out_centres[s0 + sid] = t;
            if (cen) cen[sid] = t;
        }
        if (p.out_edges == nullptr) { __syncwarp(); continue; }
        __syncwarp();  // centres visible (shared memory, or this warp's own global writes)
        const float* c = cen ? cen : p.out_centres + s0;
        // compute_intervels_kernel, pdf.cu:169-241
        for (int sid = lane; sid < n_out; sid += 32) {
            if (sid == 0) {
                const float t = c[0], t_next = n_out > 1 ? c[1] : c[0];
                const float half = __fmul_rn(__fsub_rn(t_next, t), 0.5f);
                p.out_edges[e0] = fmaxf(__fsub_rn(t, half), t_min);
                if (n_out == 1) p.out_edges[e0 + 1] = fminf(__fadd_rn(t, half), t_max);
            } else {
                const float t = c[sid], t_prev = c[sid - 1];
                p.out_edges[e0 + sid] = __fmul_rn(__fadd_rn(t, t_prev), 0.5f);
                if (sid == n_out - 1) {
                    const float half = __fmul_rn(__fsub_rn(t, t_prev), 0.5f);
                    p.out_edges[e0 + sid + 1] = fminf(__fadd_rn(t, half), t_max);
                }
            }
        }
        __syncwarp();
    }
}

// Batched rows ([n_rays, n_in] -> [n_rays, n_out]): a CTA owns a tile of RT consecutive rays.
// Their input rows are one contiguous span of global memory, staged with block-wide coalesced
// loads (all in flight at once); each warp then resamples its rays out of shared memory.  The
// interval edges need the neighbouring centres: those travel between lanes by shuffle (lane 0
// carries the last centre of the previous 32), so centres and edges go straight from registers
// to global memory as coalesced row stores — no second pass, no staging of the outputs.
// With few input bins (n_in - 1 <= 32 and well below n_out) the per-bin slope
// (t_upper - t_lower) / (u_upper - u_lower) is computed once per bin instead of once per sample
// (same operands, same IEEE quotient).  Shared memory: RT * 2 * n_in floats (+ 64 per warp).
struct PdfTileArgs {
    const float* vals; const float* cdfs; const float* bias;
    float* out_edges; float* out_centres;  // either may be nullptr
    // optional: the edges mapped from s in [0,1] to ray distance t as well (PropNetEstimator's
    // _transform_stot, prop_net.py:215-229): 1 uniform: s*c1 + (1-s)*c0 with (c0, c1) = (t_min, t_max);
    // 2 lindisp: 1 / (s*c1 + (1-s)*c0) with (c0, c1) = (1/t_min, 1/t_max)
    float* out_t_edges;
    int stot;
    float stot_c0, stot_c1;
    long long n_rays;
    int n_in, n_out, rt;
    bool bin_slopes;
};

__device__ __forceinline__ float stot_apply(int mode, float s, float c0, float c1) {
    const float a = __fadd_rn(__fmul_rn(s, c1), __fmul_rn(__fsub_rn(1.0f, s), c0));
    return mode == 2 ? __frcp_rn(a) : a;  // (the correctly rounded reciprocal: what 1.0f / a is)
}

// BINS: 0 general (ladder search + one IEEE division per sample), 1 a single input bin
// (n_in == 2: PropNetEstimator's first level, the uniform CDF — no search, the bin's constants in
// registers), 2 few bins (n_in - 1 <= 32, fewer than half
// as many as samples): search, slopes from a per-ray table.  STOT: 0 off, 1 uniform, 2 lindisp.
// The loop body is straight-line: all lanes store every iteration, and the first and last edge
// (pdf.cu:178-196, 224-238) are patched by lane 0 after the loop.
template <int STOT, bool CENTRES, int BINS, bool HALF>
__global__ void __launch_bounds__(kPdfWarps * 32) importance_sampling_tile_kernel(const PdfTileArgs p) {
    extern __shared__ __align__(16) float s_pdf[];
    // HALF: a ray per half-warp (16 lanes), two rays per warp at a time: rows whose last chunk of 32
    // samples would be at most half full (48 samples: 3 rounds for two rays instead of 4)
    constexpr int L = HALF ? 16 : 32;   // lanes per ray
    constexpr int RPW = HALF ? 2 : 1;   // rays per warp at a time
    const int lane = lane_id(), w = warp_id(), tid = threadIdx.x;
    const int ll = lane & (L - 1), sub = HALF ? lane >> 4 : 0;
    const int n_in = p.n_in, n_out = p.n_out, last = n_in - 1;
    uint64_t* bar = reinterpret_cast<uint64_t*>(s_pdf);  // 16 bytes reserved
    float* s_cdf = s_pdf + 4;                          // [rt][n_in]
    float* s_val = s_cdf + (size_t)p.rt * n_in;        // [rt][n_in]   (rt * n_in is a multiple of 4)
    float* s_bin = s_val + (size_t)p.rt * n_in + w * 64;  // per warp: [32] slope, [32] degenerate midpoint
    int top = 1, ltop = 0;  // largest power of two <= last
    while (2 * top <= last) { top <<= 1; ++ltop; }
    if (tid == 0) mbar_init(bar, 1);
    __syncthreads();
    uint32_t phase = 0;
    const int n_tiles = (int)((p.n_rays + p.rt - 1) / p.rt);
    for (int tile = blockIdx.x; tile < n_tiles; tile += gridDim.x) {
        const long long r0 = (long long)tile * p.rt;
        const int rays = (int)((p.n_rays - r0 < p.rt) ? (p.n_rays - r0) : p.rt);
        const float* __restrict__ gc = p.cdfs + r0 * n_in;
        const float* __restrict__ gv = p.vals + r0 * n_in;
        const int total = rays * n_in;
        if ((((reinterpret_cast<uintptr_t>(gc) | reinterpret_cast<uintptr_t>(gv)) & 15u) | (total & 3)) == 0) {
            // the tile's rows are one contiguous, 16-byte aligned span of each array
            if (tid == 0) {
                // (the rows were last touched through the generic proxy: the previous tile's reads)
                asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
                mbar_expect_tx(bar, 2u * total * 4u);
                bulk_g2s(s_cdf, gc, total * 4u, bar);
                bulk_g2s(s_val, gv, total * 4u, bar);
            }
            mbar_wait(bar, phase);
            phase ^= 1u;
        } else {
            for (int i = tid; i < total; i += kPdfWarps * 32) { s_cdf[i] = __ldg(gc + i); s_val[i] = __ldg(gv + i); }
            __syncthreads();
        }
        for (int jj = w * RPW; jj < rays; jj += kPdfWarps * RPW) {
            // (uniform loop: the second half-warp of an odd tail repeats the first one's ray and stores nothing)
            const bool live = jj + sub < rays;
            const int j = live ? jj + sub : jj;
            const float* cdf = s_cdf + j * n_in;
            const float* val = s_val + j * n_in;
            float* __restrict__ ge = p.out_edges + (r0 + j) * (n_out + 1) + ll;
            float* __restrict__ gt = STOT ? p.out_t_edges + (r0 + j) * (n_out + 1) + ll : nullptr;
            float* __restrict__ gs = CENTRES ? p.out_centres + (r0 + j) * n_out + ll : nullptr;
            const float u_floor = cdf[0], u_ceil = cdf[last];
            const float u_step = __fdiv_rn(__fsub_rn(u_ceil, u_floor), (float)n_out);
            const float b = p.bias ? __ldg(p.bias + r0 + j) : 0.5f;
            const float t_min = val[0], t_max = val[last];
            float t0 = 0.0f, t1 = 0.0f, tl = 0.0f, tl2 = 0.0f;  // centres 0, 1, n_out-1, n_out-2
            if (BINS == 1) {
                const float du = __fsub_rn(u_ceil, u_floor);
                const bool flat1 = du < 1e-10f;  // pdf.cu:159-160
                const float slope1 = __fdiv_rn(__fsub_rn(t_max, t_min), du);
                const float mid1 = __fmul_rn(__fadd_rn(t_min, t_max), 0.5f);
                // centre of sample k (any k, also -1: only its use at sid == 0 is patched below)
                auto centre = [&](int k) {
                    const float u = __fmaf_rn(__fadd_rn((float)k, b), u_step, u_floor);
                    // upper_bound over the single leading edge: p0 = 0, p1 = 1 when cdf[0] <= u, else both 0
                    float t = flat1 ? mid1 : __fmaf_rn(__fsub_rn(u, u_floor), slope1, t_min);
                    if (u_floor > u) t = __fmul_rn(__fadd_rn(t_min, t_min), 0.5f);  // degenerate pair (0, 0): du = 0 < 1e-10
                    return t;
                };
                // the ordinary ray: a bin that is not flat and an increasing u (u_step >= 0 makes every
                // u >= cdf[0]) — the two selects above cannot fire, so its loop does not carry them
                auto centre_plain = [&](int k) {
                    const float u = __fmaf_rn(__fadd_rn((float)k, b), u_step, u_floor);
                    return __fmaf_rn(__fsub_rn(u, u_floor), slope1, t_min);
                };
                auto rounds = [&](auto&& centre_of) {
                    float carry_prev = 0.0f;  // centre of sample i0 - 1
                    for (int i0 = 0; i0 < n_out; i0 += L) {
                        const int sid = i0 + ll;
                        const float t = centre_of(sid);
                        float t_prev = __shfl_up_sync(kFull, t, 1, L);
                        if (ll == 0) t_prev = carry_prev;
                        carry_prev = __shfl_sync(kFull, t, L - 1, L);
                        const float e = __fmul_rn(__fadd_rn(t, t_prev), 0.5f);
                        if (sid < n_out && live) {
                            if (CENTRES) gs[i0] = t;
                            ge[i0] = e;  // (edge 0 is patched below)
                            if (STOT) gt[i0] = stot_apply(STOT, e, p.stot_c0, p.stot_c1);
                        }
                    }
                };
                if (!flat1 && u_step >= 0.0f) rounds(centre_plain);
                else rounds(centre);
                t0 = centre(0); t1 = centre(1); tl = centre(n_out - 1); tl2 = centre(n_out - 2);
            } else {
                if (BINS == 2) {
                    if (lane < last) {
                        const float du = __fsub_rn(cdf[lane + 1], cdf[lane]);
                        const bool flat = du < 1e-10f;
                        // a flat bin keeps its midpoint in the second slot; other bins a NaN there
                        s_bin[lane] = __fdiv_rn(__fsub_rn(val[lane + 1], val[lane]), du);
                        s_bin[32 + lane] = flat ? __fmul_rn(__fadd_rn(val[lane], val[lane + 1]), 0.5f) : __int_as_float(0x7fc00000);
                    }
                    __syncwarp();
                }
                float carry_prev = 0.0f;  // centre of sample i0 - 1
                for (int i0 = 0; i0 < n_out; i0 += L) {
                    const int sid = i0 + ll;
                    const float u = __fmaf_rn(__fadd_rn((float)sid, b), u_step, u_floor);
                    const int q = ladder_upper_bound(cdf, last, top, ltop, u);  // <= last
                    const int p0 = max(q - 1, 0);
                    const float u_lower = cdf[p0], t_lower = val[p0];
                    float t;
                    if (BINS == 2 && q == p0 + 1) {
                        const float mid = s_bin[32 + p0];
                        t = (mid == mid) ? mid : __fmaf_rn(__fsub_rn(u, u_lower), s_bin[p0], t_lower);
                    } else {
                        const float u_upper = cdf[q], t_upper = val[q];
                        const float du = __fsub_rn(u_upper, u_lower);
                        const float scaling = __fdiv_rn(__fsub_rn(t_upper, t_lower), du);
                        t = (du < 1e-10f) ? __fmul_rn(__fadd_rn(t_lower, t_upper), 0.5f)
                                          : __fmaf_rn(__fsub_rn(u, u_lower), scaling, t_lower);
                    }
                    float t_prev = __shfl_up_sync(kFull, t, 1, L);
                    if (ll == 0) t_prev = carry_prev;
                    if (i0 == 0) { t0 = __shfl_sync(kFull, t, 0, L); t1 = __shfl_sync(kFull, t, 1, L); }
                    if (i0 + L >= n_out) {  // the last chunk: centres n_out-1 and n_out-2
                        const int l = n_out - 1 - i0;
                        tl = __shfl_sync(kFull, t, l, L);
                        tl2 = (l > 0) ? __shfl_sync(kFull, t, l > 0 ? l - 1 : 0, L) : carry_prev;
                    }
                    carry_prev = __shfl_sync(kFull, t, L - 1, L);
                    // compute_intervels_kernel, pdf.cu:169-241: interior edges are midpoints of neighbouring centres
                    const float e = __fmul_rn(__fadd_rn(t, t_prev), 0.5f);
                    if (sid < n_out && live) {
                        if (CENTRES) gs[i0] = t;
                        ge[i0] = e;  // (edge 0 is patched below)
                        if (STOT) gt[i0] = stot_apply(STOT, e, p.stot_c0, p.stot_c1);
                    }
                }
            }
            if (ll == 0 && live) {
                const float half0 = __fmul_rn(__fsub_rn(n_out > 1 ? t1 : t0, t0), 0.5f);
                const float e0 = fmaxf(__fsub_rn(t0, half0), t_min);
                const float halfn = (n_out > 1) ? __fmul_rn(__fsub_rn(tl, tl2), 0.5f) : half0;
                const float en = fminf(__fadd_rn(tl, halfn), t_max);
                ge[0] = e0;
                ge[n_out] = en;
                if (STOT) { gt[0] = stot_apply(STOT, e0, p.stot_c0, p.stot_c1); gt[n_out] = stot_apply(STOT, en, p.stot_c0, p.stot_c1); }
            }
            __syncwarp();
        }
        __syncthreads();
    }
}

template <int STOT, bool CENTRES>
static void launch_is_tile(const PdfTileArgs& t, int bins, unsigned blocks, size_t smem, cudaStream_t st) {
    // a ray per half-warp when that saves rounds (general path only: the slope tables are per warp)
    const bool half = bins == 0 && (t.n_out + 15) / 16 < 2 * ((t.n_out + 31) / 32);
    if (bins == 1) importance_sampling_tile_kernel<STOT, CENTRES, 1, false><<<blocks, kPdfWarps * 32, smem, st>>>(t);
    else if (bins == 2) importance_sampling_tile_kernel<STOT, CENTRES, 2, false><<<blocks, kPdfWarps * 32, smem, st>>>(t);
    else if (half) importance_sampling_tile_kernel<STOT, CENTRES, 0, true><<<blocks, kPdfWarps * 32, smem, st>>>(t);
    else importance_sampling_tile_kernel<STOT, CENTRES, 0, false><<<blocks, kPdfWarps * 32, smem, st>>>(t);
}

struct SearchArgs {
    const float* query; const int64_t* q_info; long long n_q;   // packed (q_info) or batched
    const float* key; const int64_t* k_info; long long n_k;
    int64_t* ids_left; int64_t* ids_right;
    long long n_rays;
    int smem_floats_per_warp;
};

__global__ void __launch_bounds__(kPdfWarps * 32) searchsorted_kernel(const SearchArgs p) {
    extern __shared__ float s_pdf[];
    const int lane = lane_id(), w = warp_id();
    float* s_row = s_pdf + (size_t)w * p.smem_floats_per_warp;
    for (long long r = (long long)blockIdx.x * kPdfWarps + w; r < p.n_rays;
         r += (long long)gridDim.x * kPdfWarps) {
        long long k0, kc, q0, qc;
        if (p.k_info) { const RayChunk c = load_chunk(p.k_info, r); k0 = c.start; kc = c.cnt; }
        else { k0 = r * p.n_k; kc = p.n_k; }
        if (p.q_info) { const RayChunk c = load_chunk(p.q_info, r); q0 = c.start; qc = c.cnt; }
        else { q0 = r * p.n_q; qc = p.n_q; }
        if (qc <= 0) continue;
        const int n_k = (int)kc, last = n_k - 1;
        const float* key = p.key + k0;
        const bool staged = n_k > 0 && n_k <= p.smem_floats_per_warp;
        if (staged) {
            for (int i = lane; i < n_k; i += 32) s_row[i] = __ldg(key + i);
            __syncwarp();
            key = s_row;
        }
        const int ltop = last > 0 ? 31 - __clz(last) : 0, top = 1 << ltop;  // largest power of two <= last
        // batched query: ids relative to the key row; packed query: absolute positions (pdf.cu:278-284)
        const long long off = p.q_info ? k0 : 0;
        for (long long i = lane; i < qc; i += 32) {
            const float v = __ldg(p.query + q0 + i);
            const int q = (n_k <= 0) ? 0 : (staged && last > 0 ? ladder_upper_bound(key, last, top, ltop, v)
                                                                : upper_bound(key, 0, last, v));
            const int l = max(min(q - 1, last), 0), h = max(min(q, last), 0);
            p.ids_left[q0 + i] = off + l;
            p.ids_right[q0 + i] = off + h;
        }
        __syncwarp();
    }
}

static inline unsigned pdf_blocks(long long n_rays) {
    long long b = ceil_div<long long>(n_rays, kPdfWarps);
    const long long cap = (long long)sm_count() * 32;
    if (b > cap) b = cap;
    if (b < 1) b = 1;
    return (unsigned)b;
}

// shared-memory floats per warp: what the longest row needs when the caller knows it, a 6 KiB
// default otherwise, capped so that 8 warps stay under 96 KiB (2 CTAs/SM).  Rows that do not fit
// are searched in global memory (the kernels decide per ray).
static inline int pdf_smem_floats(long long want) {
    const long long cap = (96 * 1024) / (kPdfWarps * 4);
    if (want <= 0) return 1536;
    return (int)(want <= cap ? want : cap);
}

}  // namespace nfb

extern "C" {

static int importance_sampling_impl(const float* vals, const float* cdfs, const int64_t* in_info,
                                    int64_t n_rays, int64_t n_in, int64_t max_in,
                                    const int64_t* n_out_per_ray, int64_t n_out, int64_t max_out,
                                    const int64_t* edge_info, const int64_t* sample_info,
                                    const float* bias, float* out_edges, float* out_centres,
                                    float* out_t_edges, int stot, float stot_c0, float stot_c1,
                                    nfb_stream_t stream) {
    using namespace nfb;
    NFB_CHECK_ARG(n_rays >= 0, "negative size");
    if (n_rays == 0) return NFB_OK;
    NFB_CHECK_ARG(vals && cdfs && (out_centres || out_edges), "null pointer");
    NFB_CHECK_ARG(in_info != nullptr || n_in > 0, "need in_info or n_in");
    NFB_CHECK_ARG(in_info == nullptr || aligned16(in_info), "packed_info must be 16-byte aligned");
    if (n_out_per_ray) NFB_CHECK_ARG(edge_info && sample_info, "per-ray counts need edge_info and sample_info");
    else NFB_CHECK_ARG(n_out > 0, "n_out must be positive");
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    if (in_info == nullptr && n_out_per_ray == nullptr) {
        // batched in, batched out: CTA tiles of rays when a useful tile fits in shared memory
        const long long per_ray = 2 * n_in;
        const long long fixed = (long long)kPdfWarps * 64 + 4;  // per-warp slope tables + the copy barrier
        int rt = 64;
        while (rt > 4 && (per_ray * rt + fixed) * 4 > 48 * 1024) rt >>= 1;
        if ((per_ray * rt + fixed) * 4 <= 48 * 1024 && n_in < (1 << 20) && n_out < (1 << 20) && out_edges != nullptr &&
            n_in >= 2) {
            PdfTileArgs t{};
            t.vals = vals; t.cdfs = cdfs; t.bias = bias; t.out_edges = out_edges; t.out_centres = out_centres;
            t.out_t_edges = out_t_edges; t.stot = stot; t.stot_c0 = stot_c0; t.stot_c1 = stot_c1;
            t.n_rays = n_rays; t.n_in = (int)n_in; t.n_out = (int)n_out; t.rt = rt;
            t.bin_slopes = (n_in - 1 <= 32) && (2 * (n_in - 1) <= n_out);
            const size_t smem = (size_t)(per_ray * rt + fixed) * 4;
            long long blocks = ceil_div<long long>(n_rays, rt);
            // one CTA per tile of rays (measured: a persistent grid of 8 CTAs per SM leaves a tail
            // and keeps the load and compute phases of neighbouring tiles from overlapping)
            const long long cap = (long long)sm_count() * 4096;
            if (blocks > cap) blocks = cap;
            const int bins = (n_in == 2) ? 1 : (t.bin_slopes ? 2 : 0);
            const bool cen = out_centres != nullptr;
            const unsigned nb = (unsigned)blocks;
            if (stot == 2) { if (cen) launch_is_tile<2, true>(t, bins, nb, smem, st); else launch_is_tile<2, false>(t, bins, nb, smem, st); }
            else if (stot == 1) { if (cen) launch_is_tile<1, true>(t, bins, nb, smem, st); else launch_is_tile<1, false>(t, bins, nb, smem, st); }
            else { if (cen) launch_is_tile<0, true>(t, bins, nb, smem, st); else launch_is_tile<0, false>(t, bins, nb, smem, st); }
            count_launch();
            NFB_CHECK_LAUNCH("importance_sampling_tile");
            return NFB_OK;
        }
    }
    NFB_CHECK_ARG(out_t_edges == nullptr, "the s -> t output needs batched rows that fit the tile kernel");
    PdfArgs a{};
    a.vals = vals; a.cdfs = cdfs; a.in_info = in_info; a.n_in = n_in;
    a.n_out_per_ray = n_out_per_ray; a.n_out = n_out; a.edge_info = edge_info; a.sample_info = sample_info;
    a.bias = bias; a.out_edges = out_edges; a.out_centres = out_centres; a.n_rays = n_rays;
    const long long mi = in_info ? max_in : n_in, mo = n_out_per_ray ? max_out : n_out;
    a.smem_floats_per_warp = pdf_smem_floats((mi > 0 && mo > 0) ? 2 * mi + mo : 0);
    NFB_CHECK_ARG(out_centres != nullptr || out_edges != nullptr, "nothing to compute");
    NFB_CHECK_ARG(out_centres != nullptr || (mi > 0 && mo > 0 && 2 * mi + mo <= a.smem_floats_per_warp),
                  "rows too long to resample without a centres output");
    const size_t smem = (size_t)a.smem_floats_per_warp * 4 * kPdfWarps;
    if (smem > 48 * 1024)
        cudaFuncSetAttribute(importance_sampling_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    importance_sampling_kernel<<<pdf_blocks(n_rays), kPdfWarps * 32, smem, st>>>(a);
    count_launch();
    NFB_CHECK_LAUNCH("importance_sampling");
    return NFB_OK;
}

NFB_API int nfb_importance_sampling(const float* vals, const float* cdfs, const int64_t* in_info,
                                    int64_t n_rays, int64_t n_in, int64_t max_in,
                                    const int64_t* n_out_per_ray, int64_t n_out, int64_t max_out,
                                    const int64_t* edge_info, const int64_t* sample_info,
                                    const float* bias, float* out_edges, float* out_centres,
                                    nfb_stream_t stream) {
    return importance_sampling_impl(vals, cdfs, in_info, n_rays, n_in, max_in, n_out_per_ray, n_out, max_out,
                                    edge_info, sample_info, bias, out_edges, out_centres, nullptr, 0, 0.0f, 0.0f,
                                    stream);
}

NFB_API int nfb_importance_sampling_stot(const float* vals, const float* cdfs, int64_t n_rays, int64_t n_in,
                                         int64_t n_out, const float* bias, float* out_edges,
                                         float* out_t_edges, int stot_mode, float c0, float c1,
                                         nfb_stream_t stream) {
    using namespace nfb;
    NFB_CHECK_ARG(stot_mode == 1 || stot_mode == 2, "stot_mode: 1 uniform, 2 lindisp");
    NFB_CHECK_ARG(out_edges && out_t_edges, "null pointer");
    return importance_sampling_impl(vals, cdfs, nullptr, n_rays, n_in, n_in, nullptr, n_out, n_out, nullptr,
                                    nullptr, bias, out_edges, nullptr, out_t_edges, stot_mode, c0, c1, stream);
}

NFB_API int nfb_searchsorted(const float* query, const int64_t* q_info, int64_t n_q,
                             const float* key, const int64_t* k_info, int64_t n_k, int64_t max_k,
                             int64_t n_rays, int64_t* ids_left, int64_t* ids_right,
                             nfb_stream_t stream) {
    using namespace nfb;
    NFB_CHECK_ARG(n_rays >= 0, "negative size");
    if (n_rays == 0) return NFB_OK;
    NFB_CHECK_ARG(query && key && ids_left && ids_right, "null pointer");
    NFB_CHECK_ARG((q_info == nullptr || aligned16(q_info)) && (k_info == nullptr || aligned16(k_info)),
                  "packed_info must be 16-byte aligned");
    SearchArgs a{};
    a.query = query; a.q_info = q_info; a.n_q = n_q; a.key = key; a.k_info = k_info; a.n_k = n_k;
    a.ids_left = ids_left; a.ids_right = ids_right; a.n_rays = n_rays;
    const long long mk = k_info ? max_k : n_k;
    a.smem_floats_per_warp = pdf_smem_floats(mk);
    const size_t smem = (size_t)a.smem_floats_per_warp * 4 * kPdfWarps;
    if (smem > 48 * 1024)
        cudaFuncSetAttribute(searchsorted_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    searchsorted_kernel<<<pdf_blocks(n_rays), kPdfWarps * 32, smem, static_cast<cudaStream_t>(stream)>>>(a);
    count_launch();
    NFB_CHECK_LAUNCH("searchsorted");
    return NFB_OK;
}

}  // extern "C"
