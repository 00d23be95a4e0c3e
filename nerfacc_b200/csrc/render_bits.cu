// render_bits.cu — nerfacc.rendering() for a dense packed layout on warp tiles (warp_tile.cuh).
//
// Same composition as render_rays.cu (volrend.py:14-158: weights from density / alpha, then
// accumulate_along_rays for rgb, opacity and depth, and the autograd of all of it), but the
// samples are cut into 512-item warp tiles independent of the rays, so every access is a
// coalesced 512-byte (1536-byte for rgb) warp transaction and nothing on the critical path
// depends on the length of a ray:
//   forward   one segmented scan gives T and w (written with alpha); five more segmented sums
//             (w*r, w*g, w*b, w, w*t_mid) give the per-ray totals, which the lane holding a
//             ray's last sample writes to colors / opacity / depth.  The ray id of that sample
//             is a sparse load from ray_indices (one per ray, not per sample).  36 B/sample.
//   backward  walks the same tiles in reverse: the ray id of every sample is propagated from
//             the (reverse) heads by one shuffle per vector, the five per-ray gradient values
//             are gathered through it (L1-resident rows), g_rgb is written and one reverse
//             segmented sum gives g_sigma / g_alpha.  44 B/sample, no dense ray_indices read.
// Rows of rays without samples are zero: the entry point clears the outputs first.
#include <type_traits>

#include "warp_tile.cuh"

namespace nfb {

constexpr int kRbSlots = 6;  // look-back words per warp tile: T/w scan + five channel sums

struct RenderBitsArgs {
    const float* __restrict__ alphas;
    const float* __restrict__ t_starts;
    const float* __restrict__ t_ends;
    const float* __restrict__ sigmas;
    const float* __restrict__ rgbs;  // [n, 3]
    const int64_t* __restrict__ ray_indices;
    const uint32_t* __restrict__ bits;
    long long n, n_tiles;
    unsigned long long* status;
    uint32_t epoch;
    bool vec_ok;
    // forward outputs
    float* __restrict__ weights;
    float* __restrict__ trans;
    float* __restrict__ alphas_out;
    float* __restrict__ colors;
    float* __restrict__ opacity;
    float* __restrict__ depth;
    // backward
    const float* __restrict__ trans_in;
    const float* __restrict__ g_colors;
    const float* __restrict__ g_opacity;
    const float* __restrict__ g_depth;
    const float* __restrict__ g_weights;
    const float* __restrict__ g_trans;
    const float* __restrict__ g_alphas;
    float* __restrict__ g_in;
    float* __restrict__ g_rgbs;
};

// 128-bit load that may stay in L1: a lane's three rgb vectors share their sectors with the
// neighbouring lanes' (48-byte stride), so the second touch of a sector should hit.
__device__ __forceinline__ void ldg_l1_4_if(bool on, const float* p, float& a, float& b, float& c, float& d) {
    asm volatile("{\n\t.reg .pred q;\n\tsetp.ne.s32 q, %5, 0;\n\t"
                 "@q ld.global.nc.v4.f32 {%0,%1,%2,%3}, [%4];\n\t}"
                 : "+f"(a), "+f"(b), "+f"(c), "+f"(d) : "l"(p), "r"((int)on));
}

// rgb rows of the four items of vector k, logical component order.
template <bool REV>
__device__ __forceinline__ void load_rgb(const WarpTile<REV>& wt, const float* __restrict__ rgbs, int k,
                                         float (&q)[4][3]) {
    const long long gg = wt.g(k);
    const bool whole = wt.vec_ok && gg >= 0 && gg + 3 < wt.n;
    float t[12];
#pragma unroll
    for (int j = 0; j < 12; ++j) t[j] = 0.0f;
    const float* __restrict__ b = rgbs + 3 * gg;
    ldg_l1_4_if(whole, b, t[0], t[1], t[2], t[3]);
    ldg_l1_4_if(whole, b + 4, t[4], t[5], t[6], t[7]);
    ldg_l1_4_if(whole, b + 8, t[8], t[9], t[10], t[11]);
    if (wt.slow) {
#pragma unroll
        for (int j = 0; j < 12; ++j) {
            const long long gi = gg + j / 3;
            ldg_stream_if(!whole && gi >= 0 && gi < wt.n, b + j, t[j]);
        }
    }
#pragma unroll
    for (int c = 0; c < 4; ++c) {
        const int pc = REV ? 3 - c : c;
#pragma unroll
        for (int ch = 0; ch < 3; ++ch) q[c][ch] = t[3 * pc + ch];
    }
}

// writes the rgb-shaped rows of the components this warp owns
template <bool REV>
__device__ __forceinline__ void store_rgb(const WarpTile<REV>& wt, float* __restrict__ out, int k,
                                          const float (&q)[4][3]) {
    const long long gg = wt.g(k);
    const uint32_t m = wt.mask(k);
    const bool whole = wt.vec_ok && gg >= 0 && gg + 3 < wt.n && m == 0xfu;
    float t[12];
#pragma unroll
    for (int c = 0; c < 4; ++c) {
        const int pc = REV ? 3 - c : c;
#pragma unroll
        for (int ch = 0; ch < 3; ++ch) t[3 * pc + ch] = q[c][ch];
    }
    float* __restrict__ b = out + 3 * gg;
    stg_stream4_if(whole, b, t[0], t[1], t[2], t[3]);
    stg_stream4_if(whole, b + 4, t[4], t[5], t[6], t[7]);
    stg_stream4_if(whole, b + 8, t[8], t[9], t[10], t[11]);
    if (wt.slow || (m != 0xfu && m != 0u)) {
#pragma unroll
        for (int j = 0; j < 12; ++j) {
            const int pc = j / 3;
            const int c = REV ? 3 - pc : pc;
            const long long gi = gg + pc;
            stg_stream_if(!whole && ((m >> c) & 1u) && gi >= 0 && gi < wt.n, b + j, t[j]);
        }
    }
}

// Forward layouts only: bit c set <=> item g + c is the last sample of its ray (head bit of the
// item after it; bit n of the head bits is always set).
__device__ __forceinline__ uint32_t end_nibble(const uint32_t* __restrict__ bits, long long g, long long n) {
    if (g >= n || g < 0) return 0u;
    const long long h = g + 1;
    const uint32_t lo = __ldg(bits + (h >> 5)), hi = __ldg(bits + (h >> 5) + 1);
    return __funnelshift_r(lo, hi, (unsigned)(h & 31)) & 0xfu;
}

// ---------------------------------------------------------------------------------------------
// A tile is walked vector by vector (128 items) with the running sums carried in registers, so
// nothing per-item outlives its vector: few registers (4 CTAs/SM forward, 3 backward), straight-line code
// per vector, and the head logic of a vector computed once for all the sums that use it.  A tile
// with no ray boundary in its first 128 items continues a long ray: it first waits for the
// running sums its predecessor publishes when it is done (a chain as long as the ray is in
// tiles, met only by rays of hundreds of samples).
struct VecHeads {
    bool h0, h1, h2, h3;  // head flags of my four items (scan order)
    bool o0, o1, o2, o3;  // no head at or before the component inside my four items
    unsigned hmask;       // lanes holding a head
    int lim;              // a partial from d lanes back is absorbed iff d <= lim
    bool nhb;             // no head in the lanes before mine
    __device__ __forceinline__ void init(uint32_t nib, int lane) {
        h0 = nib & 1u; h1 = nib & 2u; h2 = nib & 4u; h3 = nib & 8u;
        o0 = !h0; o1 = o0 && !h1; o2 = o1 && !h2; o3 = o2 && !h3;
        const unsigned lt_mask = (1u << lane) - 1u;
        hmask = __ballot_sync(kFull, nib != 0u);
        const unsigned le = hmask & (lt_mask | (1u << lane));
        lim = le ? (lane - 31 + __clz(le)) : lane;
        nhb = (hmask & lt_mask) == 0u;
    }
};

// Segmented inclusive scan of one vector.  in: u = values, wc = running value since the warp's
// last head.  out: u = lane-local scans, e = value carried into the lane (full result of
// component c: o_c ? Op(e, u_c) : u_c), wc updated.
template <class Op>
__device__ __forceinline__ void vec_segscan(float& u0, float& u1, float& u2, float& u3, float& e, float& wc,
                                            const VecHeads& H, int lane) {
    if (!H.h1) u1 = Op::apply(u0, u1);
    if (!H.h2) u2 = Op::apply(u1, u2);
    if (!H.h3) u3 = Op::apply(u2, u3);
    float t = u3;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
        const float y = __shfl_up_sync(kFull, t, d);
        if (d <= H.lim) t = Op::apply(y, t);
    }
    e = __shfl_up_sync(kFull, t, 1);
    if (lane == 0) e = wc;
    else if (H.nhb) e = Op::apply(wc, e);
    const float t31 = __shfl_sync(kFull, t, 31);
    wc = (H.hmask == 0u) ? Op::apply(wc, t31) : t31;
}

// Plain (schedulable) 128-bit accesses for tiles that lie wholly inside the arrays.
__device__ __forceinline__ void ldg_nc4(const float* p, float& a, float& b, float& c, float& d) {
    asm volatile("ld.global.nc.L1::no_allocate.v4.f32 {%0,%1,%2,%3}, [%4];" : "=f"(a), "=f"(b), "=f"(c), "=f"(d) : "l"(p));
}
__device__ __forceinline__ void ldg_l1_4(const float* p, float& a, float& b, float& c, float& d) {
    asm volatile("ld.global.nc.v4.f32 {%0,%1,%2,%3}, [%4];" : "=f"(a), "=f"(b), "=f"(c), "=f"(d) : "l"(p));
}

// Per-vector access for the streaming kernels.  FAST: the tile and its overhang lie inside the
// arrays and everything is 16-byte aligned — no range predicates, addresses are the lane's base
// pointer plus a constant; only the two boundary vectors (k = 0 behind a predecessor's
// overhang, k = kWtVec) can be partially owned.
template <bool FAST, bool REV>
struct VecIO {
    const WarpTile<REV>& wt;
    __device__ __forceinline__ void load(const float* __restrict__ p, int k, float (&v)[4]) const {
        if (FAST && k == kWtVec) {
            // the overhang: only the lanes holding items before the next tile's first head
            const float* q = p + wt.g0 + k * (REV ? -128 : 128);
            v[0] = v[1] = v[2] = v[3] = 0.0f;
            const bool need = wt.over && 4 * wt.lane < wt.p_nx;
            if (REV) ldg_stream4_if(need, q, v[3], v[2], v[1], v[0]);
            else ldg_stream4_if(need, q, v[0], v[1], v[2], v[3]);
        } else if (FAST) {
            const float* q = p + wt.g0 + k * (REV ? -128 : 128);
            if (REV) ldg_nc4(q, v[3], v[2], v[1], v[0]);
            else ldg_nc4(q, v[0], v[1], v[2], v[3]);
        } else {
            wt.load(p, k, 0.0f, v);
        }
    }
    __device__ __forceinline__ void load3(const float* __restrict__ p, int k, float (&q)[4][3]) const {
        if (FAST) {
            const float* b = p + 3 * (wt.g0 + k * (REV ? -128 : 128));
            float t[12];
            if (k == kWtVec) {
                const bool need = wt.over && 4 * wt.lane < wt.p_nx;
#pragma unroll
                for (int j = 0; j < 12; ++j) t[j] = 0.0f;
                ldg_l1_4_if(need, b, t[0], t[1], t[2], t[3]);
                ldg_l1_4_if(need, b + 4, t[4], t[5], t[6], t[7]);
                ldg_l1_4_if(need, b + 8, t[8], t[9], t[10], t[11]);
            } else {
                ldg_l1_4(b, t[0], t[1], t[2], t[3]);
                ldg_l1_4(b + 4, t[4], t[5], t[6], t[7]);
                ldg_l1_4(b + 8, t[8], t[9], t[10], t[11]);
            }
#pragma unroll
            for (int c = 0; c < 4; ++c)
#pragma unroll
                for (int ch = 0; ch < 3; ++ch) q[c][ch] = t[3 * (REV ? 3 - c : c) + ch];
        } else {
            load_rgb<REV>(wt, p, k, q);
        }
    }
    __device__ __forceinline__ bool partial(int k) const {
        return (k == 0 && wt.by_pred) || k == kWtVec;
    }
    __device__ __forceinline__ void store(float* __restrict__ p, int k, const float (&v)[4]) const {
        if (FAST && !partial(k)) {
            float* q = p + wt.g0 + k * (REV ? -128 : 128);
            if (REV) stg_stream4(q, make_float4(v[3], v[2], v[1], v[0]));
            else stg_stream4(q, make_float4(v[0], v[1], v[2], v[3]));
        } else {
            wt.store(p, k, v);
        }
    }
    __device__ __forceinline__ void store3(float* __restrict__ p, int k, const float (&q)[4][3]) const {
        if (FAST && !partial(k)) {
            float* b = p + 3 * (wt.g0 + k * (REV ? -128 : 128));
            float t[12];
#pragma unroll
            for (int c = 0; c < 4; ++c)
#pragma unroll
                for (int ch = 0; ch < 3; ++ch) t[3 * (REV ? 3 - c : c) + ch] = q[c][ch];
            stg_stream4(b, make_float4(t[0], t[1], t[2], t[3]));
            stg_stream4(b + 4, make_float4(t[4], t[5], t[6], t[7]));
            stg_stream4(b + 8, make_float4(t[8], t[9], t[10], t[11]));
        } else {
            store_rgb<REV>(wt, p, k, q);
        }
    }
};

// The running sums a tile that continues a long ray starts from: its predecessor's final ones.
__device__ __forceinline__ float wait_final(const unsigned long long* st, uint32_t epoch) {
    while (true) {
        const unsigned long long w = ld_relaxed_u64(st);
        const uint32_t meta = (uint32_t)(w >> 32);
        if ((meta >> 2) == epoch && (meta & 3u) == 2u) return __uint_as_float((uint32_t)w);
        __nanosleep(200);
    }
}

template <bool DENSITY, bool FAST, bool AHEAD>
__device__ __forceinline__ void render_fwd_stream(const RenderBitsArgs& p, const WarpTile<false>& wt) {
    using TOp = typename std::conditional<DENSITY, SumOp, ProdOp>::type;
    const int lane = wt.lane;
    const VecIO<FAST, false> io{wt};
    float wcT = TOp::identity();
    float wc[5] = {0.0f, 0.0f, 0.0f, 0.0f, 0.0f};
    if (wt.need_x) {  // no ray boundary in my first 128 items: carry on from the previous tile
        const unsigned long long* st = p.status + (wt.t - 1) * kRbSlots;
        wcT = wait_final(st, p.epoch);
#pragma unroll
        for (int ch = 0; ch < 5; ++ch) wc[ch] = wait_final(st + 1 + ch, p.epoch);
    }
    // whole tile into L2 now; the vector loads below then run one vector ahead of the math
    wt.prefetch_l2(p.t_starts, 4);
    wt.prefetch_l2(p.t_ends, 4);
    wt.prefetch_l2(DENSITY ? p.sigmas : p.alphas, 4);
    wt.prefetch_l2(p.rgbs, 12);
    float bts[2][4], bte[2][4], bfv[2][4], brgb[2][4][3];
    io.load(p.t_starts, 0, bts[0]);
    io.load(p.t_ends, 0, bte[0]);
    io.load(DENSITY ? p.sigmas : p.alphas, 0, bfv[0]);
    io.load3(p.rgbs, 0, brgb[0]);
#pragma unroll 2  // (fully unrolled the hot path is 94 KB of straight-line code: 13 % of the stalls were instruction fetch; rolled, the next vector's loads no longer overlap this one's math: 0.307 / 0.297 / 0.322 ms for 5 / 2 / 1)
    for (int k = 0; k < kWtNV; ++k) {
        const long long gg = wt.g(k);
        if (AHEAD && k + 1 < kWtNV) {
            io.load(p.t_starts, k + 1, bts[(k + 1) & 1]);
            io.load(p.t_ends, k + 1, bte[(k + 1) & 1]);
            io.load(DENSITY ? p.sigmas : p.alphas, k + 1, bfv[(k + 1) & 1]);
            io.load3(p.rgbs, k + 1, brgb[(k + 1) & 1]);
        }
        if (!AHEAD && k > 0) {
            io.load(p.t_starts, k, bts[0]);
            io.load(p.t_ends, k, bte[0]);
            io.load(DENSITY ? p.sigmas : p.alphas, k, bfv[0]);
            io.load3(p.rgbs, k, brgb[0]);
        }
        float (&ts)[4] = bts[AHEAD ? (k & 1) : 0], (&te)[4] = bte[AHEAD ? (k & 1) : 0], (&fv)[4] = bfv[AHEAD ? (k & 1) : 0];
        float (&rgb)[4][3] = brgb[AHEAD ? (k & 1) : 0];
        // this vector's ray ends (mine to write) and the ray id of the first of them
        uint32_t em, hn = 0u;
        if (k < kWtVec) {
            // one 5-bit window of the head bits: heads of my four items and of the one after them
            uint32_t win = 0u;
            if (gg >= 0 && gg < p.n) {
                const uint32_t lo = __ldg(p.bits + (gg >> 5)), hi = __ldg(p.bits + (gg >> 5) + 1);
                win = __funnelshift_r(lo, hi, (unsigned)(gg & 31));
            }
            hn = win & 0xfu;
            em = (win >> 1) & 0xfu & wt.mask(k);
        } else {
            const int d = wt.p_nx - 1 - 4 * lane;  // the overhang's one end: item p_nx - 1
            em = (wt.over && d >= 0 && d < 4) ? (1u << d) : 0u;
        }
        const int cf = __ffs(em) - 1;
        long long rid = 0;
        if (em) rid = __ldg(p.ray_indices + gg + cf);
        VecHeads H;
        H.init(hn, lane);
        // transmittance and weights
        float al[4], x[4], tm[4];
#pragma unroll
        for (int c = 0; c < 4; ++c) {
            if (DENSITY) {
                x[c] = __fmul_rn(fv[c], __fsub_rn(te[c], ts[c]));
                al[c] = __fsub_rn(1.0f, expf(-x[c]));
            } else {
                al[c] = fv[c];
                x[c] = __fsub_rn(1.0f, al[c]);
            }
            tm[c] = __fmul_rn(__fadd_rn(ts[c], te[c]), 0.5f);
        }
        float eT;
        vec_segscan<TOp>(x[0], x[1], x[2], x[3], eT, wcT, H, lane);
        float T[4], w[4];
        {
            const float i0 = H.o0 ? TOp::apply(eT, x[0]) : x[0], i1 = H.o1 ? TOp::apply(eT, x[1]) : x[1],
                        i2 = H.o2 ? TOp::apply(eT, x[2]) : x[2];
            T[0] = H.h0 ? TOp::identity() : eT;
            T[1] = H.h1 ? TOp::identity() : i0;
            T[2] = H.h2 ? TOp::identity() : i1;
            T[3] = H.h3 ? TOp::identity() : i2;
        }
#pragma unroll
        for (int c = 0; c < 4; ++c) {
            if (DENSITY) T[c] = expf(-T[c]);
            w[c] = __fmul_rn(T[c], al[c]);
        }
        io.store(p.trans, k, T);
        if (DENSITY) io.store(p.alphas_out, k, al);
        io.store(p.weights, k, w);
        // the five per-ray sums, scanned together: five independent shuffle chains in flight
        // instead of one (a lone chain leaves the warp waiting on each shuffle's latency)
        const bool of = cf == 0 ? H.o0 : (cf == 1 ? H.o1 : (cf == 2 ? H.o2 : H.o3));
        float sel[5], run[5], uf[5];
#pragma unroll
        for (int ch = 0; ch < 5; ++ch) {
            float u0, u1, u2, u3;
            if (ch < 3) {
                u0 = __fmul_rn(w[0], rgb[0][ch < 3 ? ch : 0]); u1 = __fmul_rn(w[1], rgb[1][ch < 3 ? ch : 0]);
                u2 = __fmul_rn(w[2], rgb[2][ch < 3 ? ch : 0]); u3 = __fmul_rn(w[3], rgb[3][ch < 3 ? ch : 0]);
            } else if (ch == 3) {
                u0 = w[0]; u1 = w[1]; u2 = w[2]; u3 = w[3];
            } else {
                u0 = __fmul_rn(w[0], tm[0]); u1 = __fmul_rn(w[1], tm[1]);
                u2 = __fmul_rn(w[2], tm[2]); u3 = __fmul_rn(w[3], tm[3]);
            }
            if (!H.h1) u1 = __fadd_rn(u0, u1);
            if (!H.h2) u2 = __fadd_rn(u1, u2);
            if (!H.h3) u3 = __fadd_rn(u2, u3);
            float f = u3;
            if (cf == 2) f = u2;
            if (cf == 1) f = u1;
            if (cf == 0) f = u0;
            uf[ch] = f;
            run[ch] = u3;
        }
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) {
            const bool take = d <= H.lim;
#pragma unroll
            for (int ch = 0; ch < 5; ++ch) {
                const float y = __shfl_up_sync(kFull, run[ch], d);
                if (take) run[ch] = __fadd_rn(y, run[ch]);
            }
        }
#pragma unroll
        for (int ch = 0; ch < 5; ++ch) {
            float e = __shfl_up_sync(kFull, run[ch], 1);
            if (lane == 0) e = wc[ch];
            else if (H.nhb) e = __fadd_rn(wc[ch], e);
            const float t31 = __shfl_sync(kFull, run[ch], 31);
            wc[ch] = (H.hmask == 0u) ? __fadd_rn(wc[ch], t31) : t31;
            sel[ch] = of ? __fadd_rn(e, uf[ch]) : uf[ch];
        }
        if (em) {
            float* __restrict__ pc = p.colors + 3 * rid;
            pc[0] = sel[0]; pc[1] = sel[1]; pc[2] = sel[2];
            p.opacity[rid] = sel[3];
            p.depth[rid] = sel[4];
            uint32_t more = em & (em - 1u);
            // further ends in the same lane: rays that start and end inside my four items
            while (more) {
                const int c = __ffs(more) - 1;
                more &= more - 1u;
                // (rare; its operands are read back rather than kept in registers across the scans:
                // the weights this thread has just written, the inputs from L1/L2)
                float a0 = 0.0f, a1 = 0.0f, a2 = 0.0f, a3 = 0.0f, a4 = 0.0f;
#pragma unroll
                for (int j = 1; j < 4; ++j) {
                    const bool hj = j == 1 ? H.h1 : (j == 2 ? H.h2 : H.h3);
                    if (j <= c) {
                        if (hj) { a0 = a1 = a2 = a3 = a4 = 0.0f; }
                        const long long gi = gg + j;
                        const float wj = __ldcg(p.weights + gi);
                        const float tmj = __fmul_rn(__fadd_rn(__ldg(p.t_starts + gi), __ldg(p.t_ends + gi)), 0.5f);
                        a0 = __fadd_rn(a0, __fmul_rn(wj, __ldg(p.rgbs + 3 * gi)));
                        a1 = __fadd_rn(a1, __fmul_rn(wj, __ldg(p.rgbs + 3 * gi + 1)));
                        a2 = __fadd_rn(a2, __fmul_rn(wj, __ldg(p.rgbs + 3 * gi + 2)));
                        a3 = __fadd_rn(a3, wj);
                        a4 = __fadd_rn(a4, __fmul_rn(wj, tmj));
                    }
                }
                const long long r = __ldg(p.ray_indices + gg + c);
                p.colors[3 * r] = a0; p.colors[3 * r + 1] = a1; p.colors[3 * r + 2] = a2;
                p.opacity[r] = a3;
                p.depth[r] = a4;
            }
        }
        if (k == kWtVec - 1 && wt.has_next && wt.p_nx == 128 && lane == 0) {
            // the next tile continues my last ray beyond its first 128 items: it waits for these
            st_relaxed_u64(p.status + wt.t * kRbSlots, pack_status(wcT, p.epoch, 2u));
#pragma unroll
            for (int ch = 0; ch < 5; ++ch)
                st_relaxed_u64(p.status + wt.t * kRbSlots + 1 + ch, pack_status(wc[ch], p.epoch, 2u));
        }
    }
}

template <bool DENSITY, int MINB, bool AHEAD>
__global__ void __launch_bounds__(kThreads, MINB) render_bits_fwd_stream_kernel(const RenderBitsArgs p) {
    WarpTile<false> wt;
    if (!wt.init(p.bits, p.n, p.n_tiles, p.vec_ok)) return;
    if (!wt.slow) render_fwd_stream<DENSITY, true, AHEAD>(p, wt);
    else render_fwd_stream<DENSITY, false, AHEAD>(p, wt);
}

template <bool DENSITY, bool FAST, bool AHEAD>
__device__ __forceinline__ void render_bwd_stream(const RenderBitsArgs& p, const WarpTile<true>& wt) {
    const int lane = wt.lane;
    const unsigned lt_mask = (1u << lane) - 1u;
    const VecIO<FAST, true> io{wt};
    const float* __restrict__ field = DENSITY ? p.sigmas : p.alphas;
    wt.prefetch_l2(p.t_starts, 4);
    wt.prefetch_l2(p.t_ends, 4);
    wt.prefetch_l2(p.trans_in, 4);
    wt.prefetch_l2(field, 4);
    wt.prefetch_l2(p.rgbs, 12);
    // ray ids are read at the heads only: pull in just those 32-byte sectors (a lane's four items)
#pragma unroll
    for (int k = 0; k < kWtVec; ++k)
        if (wt.f[k] != 0u) asm volatile("prefetch.global.L2 [%0];" :: "l"(p.ray_indices + wt.g(k)));
    if (p.g_weights) wt.prefetch_l2(p.g_weights, 4);
    float wc = 0.0f;  // sum of v over the samples of the open ray behind (later than) the current one
    if (wt.need_x) wc = wait_final(p.status + (wt.t - 1), p.epoch);
    // ray of the tile's first item in scan order (the physically last one)
    long long first = wt.item(0, 0);
    first = __shfl_sync(kFull, first, 0);
    int carry_rid = (int)__ldg(p.ray_indices + (first < wt.n ? first : wt.n - 1));
    float bts[2][4], bte[2][4], bT[2][4], bfv[2][4], brgb[2][4][3];
    uint32_t bnib[2];
    int bat[2][4];  // ray ids at my heads (-1: not a head)
    auto fetch = [&](int k, int b) {
        io.load(p.t_starts, k, bts[b]);
        io.load(p.t_ends, k, bte[b]);
        io.load(p.trans_in, k, bT[b]);
        io.load(field, k, bfv[b]);
        io.load3(p.rgbs, k, brgb[b]);
        bnib[b] = k < kWtVec ? head_nibble<true>(p.bits, wt.g(k), p.n) : 0u;
#pragma unroll
        for (int c = 0; c < 4; ++c) {
            bat[b][c] = -1;
            if ((bnib[b] >> c) & 1u) bat[b][c] = (int)__ldg(p.ray_indices + wt.item(k, c));
        }
    };
    fetch(0, 0);
#pragma unroll
    for (int k = 0; k < kWtNV; ++k) {
        if (AHEAD && k + 1 < kWtNV) fetch(k + 1, (k + 1) & 1);
        if (!AHEAD && k > 0) fetch(k, 0);
        constexpr int B = AHEAD ? 1 : 0;
        float (&ts)[4] = bts[k & B], (&te)[4] = bte[k & B], (&T)[4] = bT[k & B], (&fv)[4] = bfv[k & B];
        float (&rgb)[4][3] = brgb[k & B];
        int (&at)[4] = bat[k & B];
        VecHeads H;
        H.init(bnib[k & B], lane);
        // ray id of every item: the nearest head at or before it, else the ray carried in
        int rid[4];
        {
            const int last = at[3] >= 0 ? at[3] : (at[2] >= 0 ? at[2] : (at[1] >= 0 ? at[1] : at[0]));
            const unsigned below = H.hmask & lt_mask;
            const int got = __shfl_sync(kFull, last, below ? 31 - __clz(below) : 0);
            int run = below ? got : carry_rid;
#pragma unroll
            for (int c = 0; c < 4; ++c) {
                if (at[c] >= 0) run = at[c];
                rid[c] = run;
            }
            carry_rid = __shfl_sync(kFull, run, 31);
        }
        float gw[4] = {0.0f, 0.0f, 0.0f, 0.0f}, gT[4] = {0.0f, 0.0f, 0.0f, 0.0f}, ga[4] = {0.0f, 0.0f, 0.0f, 0.0f};
        if (p.g_weights) io.load(p.g_weights, k, gw);
        if (p.g_trans) io.load(p.g_trans, k, gT);
        if (DENSITY && p.g_alphas) io.load(p.g_alphas, k, ga);
        float v[4], dir[4], sc[4], grgb[4][3];
#pragma unroll
        for (int c = 0; c < 4; ++c) {
            const long long r = rid[c];
            float gc0 = 0.0f, gc1 = 0.0f, gc2 = 0.0f, go = 0.0f, gd = 0.0f;
            if (p.g_colors) { gc0 = __ldg(p.g_colors + 3 * r); gc1 = __ldg(p.g_colors + 3 * r + 1); gc2 = __ldg(p.g_colors + 3 * r + 2); }
            if (p.g_opacity) go = __ldg(p.g_opacity + r);
            if (p.g_depth) gd = __ldg(p.g_depth + r);
            float a, e = 1.0f;
            if (DENSITY) {
                sc[c] = __fsub_rn(te[c], ts[c]);
                e = expf(-__fmul_rn(fv[c], sc[c]));
                a = __fsub_rn(1.0f, e);
            } else {
                a = fv[c];
                sc[c] = fmaxf(__fsub_rn(1.0f, a), 1e-10f);
            }
            const float wgt = __fmul_rn(T[c], a);
            const float tmid = __fmul_rn(__fadd_rn(ts[c], te[c]), 0.5f);
            // same association as autograd's accumulation of the three accumulate backward terms
            float t3 = __fadd_rn(__fadd_rn(__fmul_rn(gc0, rgb[c][0]), __fmul_rn(gc1, rgb[c][1])), __fmul_rn(gc2, rgb[c][2]));
            t3 = __fadd_rn(t3, go);
            t3 = __fadd_rn(t3, __fmul_rn(gd, tmid));
            const float g = __fadd_rn(gw[c], t3);
            grgb[c][0] = __fmul_rn(wgt, gc0);
            grgb[c][1] = __fmul_rn(wgt, gc1);
            grgb[c][2] = __fmul_rn(wgt, gc2);
            if (DENSITY) dir[c] = __fmul_rn(__fadd_rn(ga[c], __fmul_rn(g, T[c])), e);
            else dir[c] = __fmul_rn(g, T[c]);
            v[c] = __fmul_rn(__fadd_rn(gT[c], __fmul_rn(g, a)), T[c]);
        }
        if (p.g_rgbs) io.store3(p.g_rgbs, k, grgb);
        float e;
        vec_segscan<SumOp>(v[0], v[1], v[2], v[3], e, wc, H, lane);
        float R[4], o[4];
        {
            const float i0 = H.o0 ? __fadd_rn(e, v[0]) : v[0], i1 = H.o1 ? __fadd_rn(e, v[1]) : v[1],
                        i2 = H.o2 ? __fadd_rn(e, v[2]) : v[2];
            R[0] = H.h0 ? 0.0f : e;
            R[1] = H.h1 ? 0.0f : i0;
            R[2] = H.h2 ? 0.0f : i1;
            R[3] = H.h3 ? 0.0f : i2;
        }
#pragma unroll
        for (int c = 0; c < 4; ++c) {
            if (DENSITY) o[c] = __fmul_rn(__fsub_rn(dir[c], R[c]), sc[c]);
            else o[c] = __fsub_rn(dir[c], __fdiv_rn(R[c], sc[c]));
        }
        io.store(p.g_in, k, o);
        if (k == kWtVec - 1 && wt.has_next && wt.p_nx == 128 && lane == 0)
            st_relaxed_u64(p.status + wt.t, pack_status(wc, p.epoch, 2u));
    }
}

template <bool DENSITY, int MINB, bool AHEAD>
__global__ void __launch_bounds__(kThreads, MINB) render_bits_bwd_stream_kernel(const RenderBitsArgs p) {
    WarpTile<true> wt;
    if (!wt.init(p.bits, p.n, p.n_tiles, p.vec_ok)) return;
    if (!wt.slow && wt.has_next) render_bwd_stream<DENSITY, true, AHEAD>(p, wt);  // (the overhang lies before the tile)
    else render_bwd_stream<DENSITY, false, AHEAD>(p, wt);
}

// ---------------------------------------------------------------------------------------------
// accumulate_along_rays (volrend.py:483-547) on the same streaming warp tiles: out[r, :] = sum of
// weights[i] * values[i, :] over the samples of ray r (values == NULL: sum of weights), D <= 4.
struct AccBitsArgs {
    const float* __restrict__ weights;
    const float* __restrict__ values;  // [n, D] or NULL
    const int64_t* __restrict__ ray_indices;
    const uint32_t* __restrict__ bits;
    long long n, n_tiles;
    unsigned long long* status;
    uint32_t epoch;
    bool vec_ok, add_into;
    float* __restrict__ out;  // [n_rays, D]
};
constexpr int kAccSlots = 4;

template <int D, bool HAS_V, bool FAST>
__device__ __forceinline__ void accumulate_stream(const AccBitsArgs& p, const WarpTile<false>& wt) {
    const int lane = wt.lane;
    const VecIO<FAST, false> io{wt};
    float wc[D];
#pragma unroll
    for (int ch = 0; ch < D; ++ch) wc[ch] = 0.0f;
    if (wt.need_x) {
#pragma unroll
        for (int ch = 0; ch < D; ++ch) wc[ch] = wait_final(p.status + (wt.t - 1) * kAccSlots + ch, p.epoch);
    }
    wt.prefetch_l2(p.weights, 4);
    if (HAS_V) wt.prefetch_l2(p.values, 4 * D);
#pragma unroll
    for (int k = 0; k < kWtNV; ++k) {
        const long long gg = wt.g(k);
        float w[4], val[4][D];
        io.load(p.weights, k, w);
        if (HAS_V) {
            // D consecutive floats per item, 4 items per lane: D 128-bit loads
            float t[4 * D];
#pragma unroll
            for (int j = 0; j < 4 * D; ++j) t[j] = 0.0f;
            const float* b = p.values + D * gg;
            const bool need = k < kWtVec || (wt.over && 4 * lane < wt.p_nx);
            const bool whole = wt.vec_ok && gg >= 0 && gg + 3 < wt.n && need;
#pragma unroll
            for (int j = 0; j < D; ++j) ldg_l1_4_if(whole, b + 4 * j, t[4 * j], t[4 * j + 1], t[4 * j + 2], t[4 * j + 3]);
            if (!FAST) {
#pragma unroll
                for (int j = 0; j < 4 * D; ++j) {
                    const long long gi = gg + j / D;
                    ldg_stream_if(!whole && need && gi >= 0 && gi < wt.n, b + j, t[j]);
                }
            }
#pragma unroll
            for (int c = 0; c < 4; ++c)
#pragma unroll
                for (int ch = 0; ch < D; ++ch) val[c][ch] = t[D * c + ch];
        }
        uint32_t em;
        if (k < kWtVec) {
            em = end_nibble(p.bits, gg, p.n) & wt.mask(k);
        } else {
            const int d = wt.p_nx - 1 - 4 * lane;  // the overhang's one end: item p_nx - 1
            em = (wt.over && d >= 0 && d < 4) ? (1u << d) : 0u;
        }
        const int cf = __ffs(em) - 1;
        long long rid = 0;
        if (em) rid = __ldg(p.ray_indices + gg + cf);
        VecHeads H;
        H.init(k < kWtVec ? head_nibble<false>(p.bits, gg, p.n) : 0u, lane);
        const bool of = cf == 0 ? H.o0 : (cf == 1 ? H.o1 : (cf == 2 ? H.o2 : H.o3));
        float sel[D];
#pragma unroll
        for (int ch = 0; ch < D; ++ch) {
            float u0 = w[0], u1 = w[1], u2 = w[2], u3 = w[3], e;
            if (HAS_V) {
                u0 = __fmul_rn(u0, val[0][ch]); u1 = __fmul_rn(u1, val[1][ch]);
                u2 = __fmul_rn(u2, val[2][ch]); u3 = __fmul_rn(u3, val[3][ch]);
            }
            vec_segscan<SumOp>(u0, u1, u2, u3, e, wc[ch], H, lane);
            const float uf = cf == 0 ? u0 : (cf == 1 ? u1 : (cf == 2 ? u2 : u3));
            sel[ch] = of ? __fadd_rn(e, uf) : uf;
        }
        if (em) {
            float* __restrict__ o = p.out + D * rid;
#pragma unroll
            for (int ch = 0; ch < D; ++ch) o[ch] = p.add_into ? __fadd_rn(o[ch], sel[ch]) : sel[ch];
            uint32_t more = em & (em - 1u);
            while (more) {  // rays that start and end inside my four items
                const int c = __ffs(more) - 1;
                more &= more - 1u;
                float a[D];
#pragma unroll
                for (int ch = 0; ch < D; ++ch) a[ch] = 0.0f;
#pragma unroll
                for (int j = 1; j < 4; ++j) {
                    const bool hj = j == 1 ? H.h1 : (j == 2 ? H.h2 : H.h3);
                    if (j <= c) {
#pragma unroll
                        for (int ch = 0; ch < D; ++ch) {
                            if (hj) a[ch] = 0.0f;
                            a[ch] = __fadd_rn(a[ch], HAS_V ? __fmul_rn(w[j], val[j][ch]) : w[j]);
                        }
                    }
                }
                float* __restrict__ o2 = p.out + D * __ldg(p.ray_indices + gg + c);
#pragma unroll
                for (int ch = 0; ch < D; ++ch) o2[ch] = p.add_into ? __fadd_rn(o2[ch], a[ch]) : a[ch];
            }
        }
        if (k == kWtVec - 1 && wt.has_next && wt.p_nx == 128 && lane == 0) {
#pragma unroll
            for (int ch = 0; ch < D; ++ch)
                st_relaxed_u64(p.status + wt.t * kAccSlots + ch, pack_status(wc[ch], p.epoch, 2u));
        }
    }
}

template <int D, bool HAS_V>
__global__ void __launch_bounds__(kThreads, 4) accumulate_bits_kernel(const AccBitsArgs p) {
    WarpTile<false> wt;
    if (!wt.init(p.bits, p.n, p.n_tiles, p.vec_ok)) return;
    if (!wt.slow) accumulate_stream<D, HAS_V, true>(p, wt);
    else accumulate_stream<D, HAS_V, false>(p, wt);
}

// depth = sum(w t_mid) / max(sum(w), eps) per ray (volrend.py:147-150) and its gradient, as one
// kernel each way over the [n_rays] rows instead of a dozen torch elementwise launches.
constexpr float kDepthEps = 1.1920928955078125e-07f;  // torch.finfo(torch.float32).eps
__global__ void depth_normalize_fwd_kernel(const float* __restrict__ opacity, float* __restrict__ depth, long long n) {
    for (long long r = (long long)blockIdx.x * blockDim.x + threadIdx.x; r < n; r += (long long)gridDim.x * blockDim.x)
        depth[r] = __fdiv_rn(depth[r], fmaxf(__ldg(opacity + r), kDepthEps));
}
// g_d0 = g_depth / den; g_o0 = g_opacity - g_depth * depth / den where opacity >= eps (clamp_min)
__global__ void depth_normalize_bwd_kernel(const float* __restrict__ g_opacity, const float* __restrict__ g_depth,
                                           const float* __restrict__ opacity, const float* __restrict__ depth,
                                           long long n, float* __restrict__ g_o0, float* __restrict__ g_d0) {
    for (long long r = (long long)blockIdx.x * blockDim.x + threadIdx.x; r < n; r += (long long)gridDim.x * blockDim.x) {
        const float o0 = __ldg(opacity + r);
        const float den = fmaxf(o0, kDepthEps);
        const float gdn = __fdiv_rn(__ldg(g_depth + r), den);
        float go = g_opacity ? __ldg(g_opacity + r) : 0.0f;
        if (o0 >= kDepthEps) go = __fsub_rn(go, __fmul_rn(gdn, __ldg(depth + r)));
        g_o0[r] = go;
        g_d0[r] = gdn;
    }
}

static int prep_ws_rb(void* ws, size_t ws_bytes, size_t need, uint32_t& epoch, cudaStream_t s) {
    if (ws == nullptr || ws_bytes < need) {
        set_error("workspace too small: need %zu bytes, got %zu", need, ws_bytes);
        return NFB_ERR_WORKSPACE;
    }
    if (epoch == 0) {
        if (cudaMemsetAsync(ws, 0, need, s) != cudaSuccess) {
            set_error("workspace memset failed");
            return NFB_ERR_CUDA;
        }
        epoch = kEpochMemset;
    } else if (epoch >= kEpochMemset) {
        set_error("epoch must be in [1, 2^30-2]");
        return NFB_ERR_INVALID_ARGUMENT;
    }
    return NFB_OK;
}

static bool all_aligned16_rb(std::initializer_list<const void*> ps) {
    for (const void* q : ps)
        if (q != nullptr && !aligned16(q)) return false;
    return true;
}

}  // namespace nfb

extern "C" {

NFB_API size_t nfb_render_bits_workspace_bytes(int64_t n) {
    return nfb_workspace_bytes((n > 0 ? n : 1) * 8);  // >= kRbSlots words per 512-item warp tile
}

NFB_API int nfb_render_bits_fwd(int from_density, const float* alphas, const float* t_starts,
                                const float* t_ends, const float* sigmas, const float* rgbs,
                                const int64_t* ray_indices, const uint32_t* head_bits, int64_t n_rays,
                                int64_t n, float* weights, float* trans, float* alphas_out,
                                float* colors, float* opacity, float* depth, void* workspace,
                                size_t workspace_bytes, uint32_t epoch, nfb_stream_t stream) {
    using namespace nfb;
    NFB_CHECK_ARG(n >= 0 && n_rays >= 0, "negative size");
    if (n_rays == 0) return NFB_OK;
    NFB_CHECK_ARG(colors && opacity && depth, "null pointer");
    cudaStream_t s = static_cast<cudaStream_t>(stream);
    // rows of rays without samples stay zero
    bool ok;
    if (opacity == colors + 3 * n_rays && depth == opacity + n_rays) {
        ok = cudaMemsetAsync(colors, 0, sizeof(float) * 5 * (size_t)n_rays, s) == cudaSuccess;
    } else {
        ok = cudaMemsetAsync(colors, 0, sizeof(float) * 3 * (size_t)n_rays, s) == cudaSuccess &&
             cudaMemsetAsync(opacity, 0, sizeof(float) * (size_t)n_rays, s) == cudaSuccess &&
             cudaMemsetAsync(depth, 0, sizeof(float) * (size_t)n_rays, s) == cudaSuccess;
    }
    if (!ok) {
        set_error("render_bits_fwd: clearing the outputs failed");
        return NFB_ERR_CUDA;
    }
    if (n == 0) return NFB_OK;
    NFB_CHECK_ARG(t_starts && t_ends && rgbs && weights && trans && ray_indices && head_bits, "null pointer");
    if (from_density) NFB_CHECK_ARG(sigmas && alphas_out, "density mode needs sigmas and alphas_out");
    else NFB_CHECK_ARG(alphas != nullptr, "alpha mode needs alphas");
    const size_t ws_need = nfb_render_bits_workspace_bytes(n);
    int rc = prep_ws_rb(workspace, workspace_bytes, ws_need, epoch, s);
    if (rc != NFB_OK) return rc;
    RenderBitsArgs a{};
    a.alphas = alphas; a.t_starts = t_starts; a.t_ends = t_ends; a.sigmas = sigmas; a.rgbs = rgbs;
    a.ray_indices = ray_indices; a.bits = head_bits;
    a.n = n; a.n_tiles = num_warp_tiles(n);
    a.status = static_cast<unsigned long long*>(workspace); a.epoch = epoch;
    a.weights = weights; a.trans = trans; a.alphas_out = alphas_out;
    a.colors = colors; a.opacity = opacity; a.depth = depth;
    a.vec_ok = all_aligned16_rb({alphas, t_starts, t_ends, sigmas, rgbs, weights, trans, alphas_out});
    const unsigned grid = (unsigned)ceil_div<long long>(a.n_tiles, kWarps);
    if (from_density) render_bits_fwd_stream_kernel<true, 4, false><<<grid, kThreads, 0, s>>>(a);
    else render_bits_fwd_stream_kernel<false, 4, false><<<grid, kThreads, 0, s>>>(a);
    count_launch();
    NFB_CHECK_LAUNCH("render_bits_fwd");
    return NFB_OK;
}

NFB_API int nfb_render_bits_bwd(int from_density, const float* g_colors, const float* g_opacity,
                                const float* g_depth, const float* g_weights, const float* g_trans,
                                const float* g_alphas, const float* alphas, const float* t_starts,
                                const float* t_ends, const float* sigmas, const float* rgbs,
                                const float* trans, const int64_t* ray_indices,
                                const uint32_t* head_bits, int64_t n, float* g_in, float* g_rgbs,
                                void* workspace, size_t workspace_bytes, uint32_t epoch,
                                nfb_stream_t stream) {
    using namespace nfb;
    NFB_CHECK_ARG(n >= 0, "negative size");
    if (n == 0) return NFB_OK;
    NFB_CHECK_ARG(ray_indices && head_bits && t_starts && t_ends && rgbs && trans && g_in, "null pointer");
    if (from_density) NFB_CHECK_ARG(sigmas != nullptr, "density mode needs sigmas");
    else NFB_CHECK_ARG(alphas != nullptr, "alpha mode needs alphas");
    cudaStream_t s = static_cast<cudaStream_t>(stream);
    int rc = prep_ws_rb(workspace, workspace_bytes, nfb_workspace_bytes(n), epoch, s);
    if (rc != NFB_OK) return rc;
    RenderBitsArgs a{};
    a.alphas = alphas; a.t_starts = t_starts; a.t_ends = t_ends; a.sigmas = sigmas; a.rgbs = rgbs;
    a.ray_indices = ray_indices; a.bits = head_bits; a.trans_in = trans;
    a.n = n; a.n_tiles = num_warp_tiles(n);
    a.status = static_cast<unsigned long long*>(workspace); a.epoch = epoch;
    a.g_colors = g_colors; a.g_opacity = g_opacity; a.g_depth = g_depth;
    a.g_weights = g_weights; a.g_trans = g_trans; a.g_alphas = from_density ? g_alphas : nullptr;
    a.g_in = g_in; a.g_rgbs = g_rgbs;
    a.vec_ok = all_aligned16_rb({alphas, t_starts, t_ends, sigmas, rgbs, trans, g_weights, g_trans, g_alphas,
                                 g_in, g_rgbs});
    const unsigned grid = (unsigned)ceil_div<long long>(a.n_tiles, kWarps);
    if (from_density) render_bits_bwd_stream_kernel<true, 3, false><<<grid, kThreads, 0, s>>>(a);
    else render_bits_bwd_stream_kernel<false, 3, false><<<grid, kThreads, 0, s>>>(a);
    count_launch();
    NFB_CHECK_LAUNCH("render_bits_bwd");
    return NFB_OK;
}

/* accumulate_along_rays for a dense canonical layout given as head bits + ray_indices, dim <= 4. */
NFB_API int nfb_accumulate_bits_fwd(const float* weights, const float* values, int64_t dim,
                                    const int64_t* ray_indices, const uint32_t* head_bits,
                                    int64_t n_rays, int64_t n, float* out, int add_into,
                                    void* workspace, size_t workspace_bytes, uint32_t epoch,
                                    nfb_stream_t stream) {
    using namespace nfb;
    NFB_CHECK_ARG(n >= 0 && n_rays >= 0 && dim >= 1 && dim <= 4, "bad sizes (dim must be 1..4)");
    NFB_CHECK_ARG(values != nullptr || dim == 1, "dim must be 1 when values is NULL");
    cudaStream_t s = static_cast<cudaStream_t>(stream);
    if (n_rays == 0) return NFB_OK;
    NFB_CHECK_ARG(out != nullptr, "null out");
    if (!add_into && cudaMemsetAsync(out, 0, sizeof(float) * n_rays * dim, s) != cudaSuccess) {
        set_error("accumulate_bits_fwd: clearing the output failed");
        return NFB_ERR_CUDA;
    }
    if (n == 0) return NFB_OK;
    NFB_CHECK_ARG(weights && ray_indices && head_bits, "null pointer");
    int rc = prep_ws_rb(workspace, workspace_bytes, nfb_render_bits_workspace_bytes(n), epoch, s);
    if (rc != NFB_OK) return rc;
    AccBitsArgs a{};
    a.weights = weights; a.values = values; a.ray_indices = ray_indices; a.bits = head_bits;
    a.n = n; a.n_tiles = num_warp_tiles(n);
    a.status = static_cast<unsigned long long*>(workspace); a.epoch = epoch;
    a.vec_ok = all_aligned16_rb({weights, values});
    a.add_into = add_into != 0; a.out = out;
    const unsigned grid = (unsigned)ceil_div<long long>(a.n_tiles, kWarps);
    if (values == nullptr) accumulate_bits_kernel<1, false><<<grid, kThreads, 0, s>>>(a);
    else if (dim == 1) accumulate_bits_kernel<1, true><<<grid, kThreads, 0, s>>>(a);
    else if (dim == 2) accumulate_bits_kernel<2, true><<<grid, kThreads, 0, s>>>(a);
    else if (dim == 3) accumulate_bits_kernel<3, true><<<grid, kThreads, 0, s>>>(a);
    else accumulate_bits_kernel<4, true><<<grid, kThreads, 0, s>>>(a);
    count_launch();
    NFB_CHECK_LAUNCH("accumulate_bits_fwd");
    return NFB_OK;
}

NFB_API int nfb_depth_normalize_fwd(const float* opacity, float* depth, int64_t n_rays, nfb_stream_t stream) {
    using namespace nfb;
    NFB_CHECK_ARG(n_rays >= 0, "negative size");
    if (n_rays == 0) return NFB_OK;
    NFB_CHECK_ARG(opacity && depth, "null pointer");
    long long b = ceil_div<long long>(n_rays, 256);
    if (b > (long long)sm_count() * 16) b = (long long)sm_count() * 16;
    depth_normalize_fwd_kernel<<<(unsigned)b, 256, 0, static_cast<cudaStream_t>(stream)>>>(opacity, depth, n_rays);
    count_launch();
    NFB_CHECK_LAUNCH("depth_normalize_fwd");
    return NFB_OK;
}

NFB_API int nfb_depth_normalize_bwd(const float* g_opacity, const float* g_depth, const float* opacity,
                                    const float* depth, int64_t n_rays, float* g_opacity_sum,
                                    float* g_depth_sum, nfb_stream_t stream) {
    using namespace nfb;
    NFB_CHECK_ARG(n_rays >= 0, "negative size");
    if (n_rays == 0) return NFB_OK;
    NFB_CHECK_ARG(g_depth && opacity && depth && g_opacity_sum && g_depth_sum, "null pointer");
    long long b = ceil_div<long long>(n_rays, 256);
    if (b > (long long)sm_count() * 16) b = (long long)sm_count() * 16;
    depth_normalize_bwd_kernel<<<(unsigned)b, 256, 0, static_cast<cudaStream_t>(stream)>>>(
        g_opacity, g_depth, opacity, depth, n_rays, g_opacity_sum, g_depth_sum);
    count_launch();
    NFB_CHECK_LAUNCH("depth_normalize_bwd");
    return NFB_OK;
}

}  // extern "C"
