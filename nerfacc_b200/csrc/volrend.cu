// volrend.cu — fused volume-rendering weights: alpha / density -> (weights, transmittance,
// alphas, visibility) forward, and the matching reverse-mode backward, one pass each.
//
// The reference composes these from 5-7 elementwise torch kernels around one packed scan
// (nerfacc/volrend.py:161-480) and lets autograd keep 3-4 sample-sized temporaries.  Here the
// elementwise math is folded into the load and store of the flat segmented scan tile
// (seg_tile.cuh):
//   forward  alpha  : T = exclusive_prod(1-a) [*prefix];           w = T*a        12 B/sample
//   forward  density: sd = s*(te-ts); a = 1-exp(-sd);
//                     T = exp(-exclusive_sum(sd)) [*prefix];        w = T*a        24 B/sample
//   backward alpha  : g_a = g_w*T - revexcl_sum((g_T+g_w*a)*T) / max(1-a,1e-10)    16 B/sample
//   backward density: g_s = ((g_a+g_w*T)*e - revexcl_sum((g_T+g_w*a)*T)) * (te-ts) 24 B/sample
// The backward formulas are exactly what torch autograd derives from the reference's
// composite, including its clamp_min(1e-10) division in the product-scan backward
// (csrc/scan.cu:299-300), so gradients match the reference, not the "ideal" derivative.
#include "warp_tile.cuh"

namespace nfb {

struct WeightFwdArgs {
    const float* alphas;    // alpha mode input
    const float* t_starts;  // density mode inputs
    const float* t_ends;
    const float* sigmas;
    const float* prefix;    // optional prefix_trans
    float* weights;         // optional outputs
    float* trans;
    float* alphas_out;
    uint8_t* vis;
    float early_stop_eps, alpha_thre;
    SegSpec seg;
    const uint32_t* bits;  // head bits (warp-tile kernels) instead of seg
    long long n, n_tiles;
    unsigned long long* status;
    uint32_t epoch;
    bool vec_ok;
};

template <bool DENSITY>
__global__ void __launch_bounds__(kThreads) weight_fwd_kernel(const WeightFwdArgs p) {
    __shared__ TileSmem sm;
    const TileGeom<false> g(p.n, p.n_tiles);
    float v[kVec][4];   // scanned quantity: (1-a) or sd
    float al[kVec][4];  // alphas
#pragma unroll
    for (int k = 0; k < kVec; ++k) {
        if (DENSITY) {
            float ts[4], te[4], s[4];
            g.load(p.t_starts, p.vec_ok, k, 0.0f, ts);
            g.load(p.t_ends, p.vec_ok, k, 0.0f, te);
            g.load(p.sigmas, p.vec_ok, k, 0.0f, s);
#pragma unroll
            for (int c = 0; c < 4; ++c) {
                v[k][c] = __fmul_rn(s[c], __fsub_rn(te[c], ts[c]));
                al[k][c] = __fsub_rn(1.0f, expf(-v[k][c]));
            }
        } else {
            g.load(p.alphas, p.vec_ok, k, 0.0f, al[k]);
#pragma unroll
            for (int c = 0; c < 4; ++c) v[k][c] = __fsub_rn(1.0f, al[k][c]);
        }
    }
    scatter_heads<false>(p.seg, g.base, p.n, sm.flags, sm.range);
    uint32_t f[kVec];
#pragma unroll
    for (int k = 0; k < kVec; ++k) f[k] = vec_flags(sm.flags, k);
    float E[kVec];
    if (DENSITY) tile_segscan<SumOp>(v, f, E, p.status, 0, p.epoch, sm);
    else tile_segscan<ProdOp>(v, f, E, p.status, 0, p.epoch, sm);
#pragma unroll
    for (int k = 0; k < kVec; ++k) {
        float T[4];
        if (DENSITY) {
            finish_exclusive<SumOp>(v[k], f[k], E[k], T);
#pragma unroll
            for (int c = 0; c < 4; ++c) T[c] = expf(-T[c]);
        } else {
            finish_exclusive<ProdOp>(v[k], f[k], E[k], T);
        }
        if (p.prefix) {
            float pf[4];
            g.load(p.prefix, p.vec_ok, k, 1.0f, pf);
#pragma unroll
            for (int c = 0; c < 4; ++c) T[c] = __fmul_rn(T[c], pf[c]);
        }
        if (p.trans) g.store(p.trans, p.vec_ok, k, T);
        if (p.alphas_out) g.store(p.alphas_out, p.vec_ok, k, al[k]);
        if (p.weights) {
            float w[4];
#pragma unroll
            for (int c = 0; c < 4; ++c) w[c] = __fmul_rn(T[c], al[k][c]);
            g.store(p.weights, p.vec_ok, k, w);
        }
        if (p.vis) {  // volrend.py:414-417, 476-479: T >= eps [& a >= thre if thre > 0]
            const long long gi = g.base + g.voff(k);
            uint8_t m[4];
#pragma unroll
            for (int c = 0; c < 4; ++c) {
                bool ok = T[c] >= p.early_stop_eps;
                if (p.alpha_thre > 0.0f) ok = ok && (al[k][c] >= p.alpha_thre);
                m[c] = ok ? 1 : 0;
            }
            if (p.vec_ok && gi + 3 < p.n) {
                *reinterpret_cast<uchar4*>(p.vis + gi) = make_uchar4(m[0], m[1], m[2], m[3]);
            } else {
#pragma unroll
                for (int c = 0; c < 4; ++c)
                    if (gi + c < p.n) p.vis[gi + c] = m[c];
            }
        }
    }
}

// ---------------------------------------------------------------- visibility -> compaction
// One pass for what OccGridEstimator.sampling does after the user's sigma_fn / alpha_fn
// (nerfacc/estimators/occ_grid.py:186-220: render_visibility_* -> a bool mask -> three boolean
// selects -> later a pack_info over the survivors): the tile computes T with the same segmented
// scan as weight_fwd_kernel, thresholds, ranks the survivors inside the tile with ballots, gets
// its global offset by a second decoupled look-back (integer counts, status slot 1), and writes
// (ray_indices, t_starts, t_ends) compacted.  The survivors' packed_info comes from the same
// ranks: new_start[r] = rank of the item at old_start[r] (also right for empty rays, whose start
// is the next ray's); pi_counts_kernel then turns neighbouring starts into counts.
struct VisCompactArgs {
    WeightFwdArgs w;
    const int64_t* ray_indices;
    const float* t_starts;  // (== w.t_starts in density mode)
    const float* t_ends;
    int64_t* o_ray;
    float* o_ts;
    float* o_te;
    int64_t* o_packed_info;  // optional [n_rays, 2]
    long long* total;        // device-visible survivor count
};

__device__ __forceinline__ unsigned long long pack_count(uint32_t v, uint32_t epoch, uint32_t state) {
    return ((unsigned long long)((epoch << 2) | state) << 32) | (unsigned long long)v;
}

// Survivors before `tile`: look back to the nearest tile that has published its inclusive prefix
// (tile 0 always has), one warp reading 32 predecessors per round; all lanes call it and get the result.
__device__ __forceinline__ long long warp_lookback_count(unsigned long long* st, int slot, long long tile,
                                                         uint32_t epoch) {
    const int lane = threadIdx.x & 31;
    long long X = 0;
    long long k = tile - 1;
    while (true) {
        const long long idx = k - lane;
        uint32_t state = 2u, val = 0u;  // before tile 0: an inclusive prefix of zero
        if (idx >= 0) {
            const unsigned long long wd = ld_relaxed_u64(st + idx * kStatusPerTile + slot);
            const uint32_t meta = (uint32_t)(wd >> 32);
            state = (meta >> 2) == epoch ? (meta & 3u) : 0u;
            val = (uint32_t)wd;
        }
        const unsigned incl = __ballot_sync(kFull, state == 2u);
        const unsigned ready = __ballot_sync(kFull, state != 0u);
        const unsigned need = incl ? (0xffffffffu >> (32 - __ffs(incl))) : 0xffffffffu;
        if ((ready & need) != need) { __nanosleep(20); continue; }
        long long v = ((need >> lane) & 1u) ? (long long)val : 0;
#pragma unroll
        for (int d = 16; d > 0; d >>= 1) v += __shfl_xor_sync(kFull, v, d);
        X += v;
        if (incl) return X;
        k -= 32;
    }
}

template <bool DENSITY>
__global__ void __launch_bounds__(kThreads, 5) vis_compact_kernel(const VisCompactArgs q) {
    const WeightFwdArgs& p = q.w;
    __shared__ TileSmem sm;
    __shared__ uint32_t s_mask[kWarps][kVec][4];  // survivors of (warp, vector, component) by lane
    __shared__ int s_vbase[kWarps][kVec];         // survivors of the tile before vector (w, k)
    __shared__ int s_wtot[kWarps];
    __shared__ long long s_tile_base;
    const TileGeom<false> g(p.n, p.n_tiles);
    const int lane = lane_id(), w = warp_id();
    // what the survivors' copies will read after the two look-backs: pull it into L2 now, so those
    // late, dependent loads do not start a DRAM round trip (one 128-byte line per thread)
    {
        const long long left = p.n - g.base;
        const long long items = left < kTile ? left : kTile;
        const char* ri = reinterpret_cast<const char*>(q.ray_indices + g.base);
        if ((long long)threadIdx.x * 128 < items * 8) asm volatile("prefetch.global.L2 [%0];" :: "l"(ri + threadIdx.x * 128));
        if (!DENSITY && (long long)threadIdx.x * 128 < items * 8) {
            const int half = threadIdx.x >> 1;  // 128 lines each for t_starts and t_ends
            const char* tp = reinterpret_cast<const char*>(((threadIdx.x & 1) ? q.t_ends : q.t_starts) + g.base);
            if ((long long)half * 128 < items * 4) asm volatile("prefetch.global.L2 [%0];" :: "l"(tp + half * 128));
        }
    }
    float v[kVec][4], al[kVec][4];
#pragma unroll
    for (int k = 0; k < kVec; ++k) {
        if (DENSITY) {
            float ts[4], te[4], s[4];
            g.load(p.t_starts, p.vec_ok, k, 0.0f, ts);
            g.load(p.t_ends, p.vec_ok, k, 0.0f, te);
            g.load(p.sigmas, p.vec_ok, k, 0.0f, s);
#pragma unroll
            for (int c = 0; c < 4; ++c) {
                v[k][c] = __fmul_rn(s[c], __fsub_rn(te[c], ts[c]));
                al[k][c] = __fsub_rn(1.0f, expf(-v[k][c]));
            }
        } else {
            g.load(p.alphas, p.vec_ok, k, 0.0f, al[k]);
#pragma unroll
            for (int c = 0; c < 4; ++c) v[k][c] = __fsub_rn(1.0f, al[k][c]);
        }
    }
    scatter_heads<false>(p.seg, g.base, p.n, sm.flags, sm.range);
    uint32_t f[kVec];
#pragma unroll
    for (int k = 0; k < kVec; ++k) f[k] = vec_flags(sm.flags, k);
    float E[kVec];
    if (DENSITY) tile_segscan<SumOp>(v, f, E, p.status, 0, p.epoch, sm);
    else tile_segscan<ProdOp>(v, f, E, p.status, 0, p.epoch, sm);

    const unsigned lt = (1u << lane) - 1u;
    uint32_t keep[kVec];
    int within[kVec], vtot[kVec], wtot = 0;
#pragma unroll
    for (int k = 0; k < kVec; ++k) {
        float T[4];
        if (DENSITY) {
            finish_exclusive<SumOp>(v[k], f[k], E[k], T);
#pragma unroll
            for (int c = 0; c < 4; ++c) T[c] = expf(-T[c]);
        } else {
            finish_exclusive<ProdOp>(v[k], f[k], E[k], T);
        }
        if (p.prefix) {
            float pf[4];
            g.load(p.prefix, p.vec_ok, k, 1.0f, pf);
#pragma unroll
            for (int c = 0; c < 4; ++c) T[c] = __fmul_rn(T[c], pf[c]);
        }
        const long long gi = g.base + g.voff(k);
        keep[k] = 0u;
        within[k] = 0;
        vtot[k] = 0;
#pragma unroll
        for (int c = 0; c < 4; ++c) {  // volrend.py:414-417, 476-479
            bool ok = T[c] >= p.early_stop_eps && gi + c < p.n;
            if (p.alpha_thre > 0.0f) ok = ok && (al[k][c] >= p.alpha_thre);
            const unsigned m = __ballot_sync(kFull, ok);
            if (lane == 0) s_mask[w][k][c] = m;
            keep[k] |= ok ? (1u << c) : 0u;
            within[k] += __popc(m & lt);
            vtot[k] += __popc(m);
        }
        wtot += vtot[k];
        if (p.vis) {
            uint8_t m4[4];
#pragma unroll
            for (int c = 0; c < 4; ++c) m4[c] = (keep[k] >> c) & 1u;
            if (p.vec_ok && gi + 3 < p.n) {
                *reinterpret_cast<uchar4*>(p.vis + gi) = make_uchar4(m4[0], m4[1], m4[2], m4[3]);
            } else {
#pragma unroll
                for (int c = 0; c < 4; ++c)
                    if (gi + c < p.n) p.vis[gi + c] = m4[c];
            }
        }
    }
    if (lane == 0) s_wtot[w] = wtot;
    __syncthreads();
    int before = 0, agg = 0;
#pragma unroll
    for (int j = 0; j < kWarps; ++j) { if (j == w) before = agg; agg += s_wtot[j]; }
    if (lane == 0) {
        int run = before;
#pragma unroll
        for (int k = 0; k < kVec; ++k) { s_vbase[w][k] = run; run += vtot[k]; }
    }
    if (w == 0) {
        const long long tile = blockIdx.x;
        unsigned long long* mine = p.status + tile * kStatusPerTile + 1;
        long long X = 0;
        if (tile == 0) {
            if (lane == 0) st_relaxed_u64(mine, pack_count((uint32_t)agg, p.epoch, 2u));
        } else {
            if (lane == 0) st_relaxed_u64(mine, pack_count((uint32_t)agg, p.epoch, 1u));
            X = warp_lookback_count(p.status, 1, tile, p.epoch);
            if (lane == 0) st_relaxed_u64(mine, pack_count((uint32_t)(X + agg), p.epoch, 2u));
        }
        if (lane == 0) {
            s_tile_base = X;
            if (tile == (long long)gridDim.x - 1 && q.total) *q.total = X + agg;
        }
    }
    __syncthreads();
    const long long tile_base = s_tile_base;
    int vb = before;
#pragma unroll
    for (int k = 0; k < kVec; ++k) {
        long long pos = tile_base + vb + within[k];
        vb += vtot[k];
        const long long gi = g.base + g.voff(k);
#pragma unroll
        for (int c = 0; c < 4; ++c) {
            if ((keep[k] >> c) & 1u) {
                q.o_ray[pos] = __ldg(q.ray_indices + gi + c);
                q.o_ts[pos] = __ldg(q.t_starts + gi + c);
                q.o_te[pos] = __ldg(q.t_ends + gi + c);
                ++pos;
            }
        }
    }
    if (q.o_packed_info) {
        const long long r_lo = sm.range[0], r_hi = sm.range[1];
        for (long long r = r_lo + threadIdx.x; r < r_hi; r += kThreads) {
            const int P = (int)(load_chunk(p.seg.packed_info, r).start - g.base);  // in [0, kTile)
            const int pw = P >> 9, pk = (P >> 7) & 3, pl = (P >> 2) & 31, pc = P & 3;
            const unsigned below = (1u << pl) - 1u;
            int rank = s_vbase[pw][pk];
#pragma unroll
            for (int c = 0; c < 4; ++c) {
                const unsigned m = s_mask[pw][pk][c];
                rank += __popc(m & below) + ((c < pc) ? (int)((m >> pl) & 1u) : 0);
            }
            q.o_packed_info[2 * r] = tile_base + rank;
        }
        if (blockIdx.x == gridDim.x - 1)  // rays that start at or past the end of the array (empty)
            for (long long r = r_hi + threadIdx.x; r < p.seg.n_rays; r += kThreads)
                q.o_packed_info[2 * r] = tile_base + agg;
    }
}

// counts of the survivors' packed_info from neighbouring starts (the last ray: up to the total)
__global__ void pi_counts_kernel(int64_t* __restrict__ pi, long long n_rays, const long long* __restrict__ total) {
    for (long long r = (long long)blockIdx.x * blockDim.x + threadIdx.x; r < n_rays;
         r += (long long)gridDim.x * blockDim.x) {
        const long long a = pi[2 * r];
        const long long b = (r + 1 < n_rays) ? pi[2 * (r + 1)] : *total;
        pi[2 * r + 1] = b - a;
    }
}

struct WeightBwdArgs {
    const float* g_weights;  // any of the three upstream grads may be null (= zero)
    const float* g_trans;
    const float* g_alphas;
    const float* alphas;     // alpha mode
    const float* t_starts;   // density mode
    const float* t_ends;
    const float* sigmas;
    const float* trans;      // saved forward output (prefix already applied)
    float* g_in;             // grad wrt alphas (alpha mode) or sigmas (density mode)
    SegSpec seg;
    const uint32_t* bits;
    long long n, n_tiles;
    unsigned long long* status;
    uint32_t epoch;
    bool vec_ok;
};

template <bool DENSITY>
__global__ void __launch_bounds__(kThreads) weight_bwd_kernel(const WeightBwdArgs p) {
    __shared__ TileSmem sm;
    const TileGeom<true> g(p.n, p.n_tiles);
    float v[kVec][4];    // s = (g_T + g_w*a) * T, reverse-exclusive-summed
    float dir[kVec][4];  // direct term
    float sc[kVec][4];   // alpha: max(1-a, 1e-10) ; density: te - ts
#pragma unroll
    for (int k = 0; k < kVec; ++k) {
        float T[4], gw[4], gT[4], a[4];
        g.load(p.trans, p.vec_ok, k, 0.0f, T);
        if (p.g_weights) g.load(p.g_weights, p.vec_ok, k, 0.0f, gw);
        else { gw[0] = gw[1] = gw[2] = gw[3] = 0.0f; }
        if (p.g_trans) g.load(p.g_trans, p.vec_ok, k, 0.0f, gT);
        else { gT[0] = gT[1] = gT[2] = gT[3] = 0.0f; }
        if (DENSITY) {
            float ts[4], te[4], s[4], ga[4];
            g.load(p.t_starts, p.vec_ok, k, 0.0f, ts);
            g.load(p.t_ends, p.vec_ok, k, 0.0f, te);
            g.load(p.sigmas, p.vec_ok, k, 0.0f, s);
            if (p.g_alphas) g.load(p.g_alphas, p.vec_ok, k, 0.0f, ga);
            else { ga[0] = ga[1] = ga[2] = ga[3] = 0.0f; }
#pragma unroll
            for (int c = 0; c < 4; ++c) {
                const float dt = __fsub_rn(te[c], ts[c]);
                const float e = expf(-__fmul_rn(s[c], dt));
                a[c] = __fsub_rn(1.0f, e);
                sc[k][c] = dt;
                dir[k][c] = __fmul_rn(__fadd_rn(ga[c], __fmul_rn(gw[c], T[c])), e);
            }
        } else {
            g.load(p.alphas, p.vec_ok, k, 0.0f, a);
#pragma unroll
            for (int c = 0; c < 4; ++c) {
                sc[k][c] = fmaxf(__fsub_rn(1.0f, a[c]), 1e-10f);
                dir[k][c] = __fmul_rn(gw[c], T[c]);
            }
        }
#pragma unroll
        for (int c = 0; c < 4; ++c)
            v[k][c] = __fmul_rn(__fadd_rn(gT[c], __fmul_rn(gw[c], a[c])), T[c]);
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

// ---------------------------------------------------------------- the same on warp tiles
// Dense packed layouts with head bits (warp_tile.cuh): identical math per sample, no block
// barrier and no look-back wait on the critical path.  n_tiles counts warp tiles here.
template <bool DENSITY>
__global__ void __launch_bounds__(kThreads, DENSITY ? 3 : 4) weight_fwd_bits_kernel(const WeightFwdArgs p) {
    WarpTile<false> wt;
    if (!wt.init(p.bits, p.n, p.n_tiles, p.vec_ok)) return;
    float v[kWtNV][4];   // scanned quantity: (1-a) or sd
    float al[kWtNV][4];  // alphas
#pragma unroll
    for (int k = 0; k < kWtNV; ++k) {
        if (DENSITY) {
            float ts[4], te[4], s[4];
            wt.load(p.t_starts, k, 0.0f, ts);
            wt.load(p.t_ends, k, 0.0f, te);
            wt.load(p.sigmas, k, 0.0f, s);
#pragma unroll
            for (int c = 0; c < 4; ++c) {
                v[k][c] = __fmul_rn(s[c], __fsub_rn(te[c], ts[c]));
                al[k][c] = __fsub_rn(1.0f, expf(-v[k][c]));
            }
        } else {
            wt.load(p.alphas, k, 0.0f, al[k]);
#pragma unroll
            for (int c = 0; c < 4; ++c) v[k][c] = __fsub_rn(1.0f, al[k][c]);
        }
    }
    float E[kWtNV];
    if (DENSITY) wt.template scan<SumOp>(v, E, p.status, 1, 0, p.epoch);
    else wt.template scan<ProdOp>(v, E, p.status, 1, 0, p.epoch);
#pragma unroll
    for (int k = 0; k < kWtNV; ++k) {
        if (k == kWtVec && !wt.over) break;
        float T[4];
        if (DENSITY) {
            finish_exclusive<SumOp>(v[k], wt.f[k], E[k], T);
#pragma unroll
            for (int c = 0; c < 4; ++c) T[c] = expf(-T[c]);
        } else {
            finish_exclusive<ProdOp>(v[k], wt.f[k], E[k], T);
        }
        if (p.prefix) {
            float pf[4];
            wt.load(p.prefix, k, 1.0f, pf);
#pragma unroll
            for (int c = 0; c < 4; ++c) T[c] = __fmul_rn(T[c], pf[c]);
        }
        if (p.trans) wt.store(p.trans, k, T);
        if (p.alphas_out) wt.store(p.alphas_out, k, al[k]);
        if (p.weights) {
            float w[4];
#pragma unroll
            for (int c = 0; c < 4; ++c) w[c] = __fmul_rn(T[c], al[k][c]);
            wt.store(p.weights, k, w);
        }
        if (p.vis) {  // volrend.py:414-417, 476-479: T >= eps [& a >= thre if thre > 0]
            const long long gi = wt.g(k);
            const uint32_t m = wt.mask(k);
            uint8_t q[4];
#pragma unroll
            for (int c = 0; c < 4; ++c) {
                bool ok = T[c] >= p.early_stop_eps;
                if (p.alpha_thre > 0.0f) ok = ok && (al[k][c] >= p.alpha_thre);
                q[c] = ok ? 1 : 0;
            }
            if (p.vec_ok && m == 0xfu && gi >= 0 && gi + 3 < p.n) {
                *reinterpret_cast<uchar4*>(p.vis + gi) = make_uchar4(q[0], q[1], q[2], q[3]);
            } else {
#pragma unroll
                for (int c = 0; c < 4; ++c)
                    if (((m >> c) & 1u) && gi + c >= 0 && gi + c < p.n) p.vis[gi + c] = q[c];
            }
        }
    }
}

template <bool DENSITY>
__global__ void __launch_bounds__(kThreads, DENSITY ? 3 : 4) weight_bwd_bits_kernel(const WeightBwdArgs p) {
    WarpTile<true> wt;
    if (!wt.init(p.bits, p.n, p.n_tiles, p.vec_ok)) return;
    float v[kWtNV][4];    // s = (g_T + g_w*a) * T, reverse-exclusive-summed
    float dir[kWtNV][4];  // direct term
    float sc[kWtNV][4];   // alpha: max(1-a, 1e-10) ; density: te - ts
#pragma unroll
    for (int k = 0; k < kWtNV; ++k) {
        float T[4], gw[4], gT[4], a[4];
        wt.load(p.trans, k, 0.0f, T);
        if (p.g_weights) wt.load(p.g_weights, k, 0.0f, gw);
        else { gw[0] = gw[1] = gw[2] = gw[3] = 0.0f; }
        if (p.g_trans) wt.load(p.g_trans, k, 0.0f, gT);
        else { gT[0] = gT[1] = gT[2] = gT[3] = 0.0f; }
        if (DENSITY) {
            float ts[4], te[4], s[4], ga[4];
            wt.load(p.t_starts, k, 0.0f, ts);
            wt.load(p.t_ends, k, 0.0f, te);
            wt.load(p.sigmas, k, 0.0f, s);
            if (p.g_alphas) wt.load(p.g_alphas, k, 0.0f, ga);
            else { ga[0] = ga[1] = ga[2] = ga[3] = 0.0f; }
#pragma unroll
            for (int c = 0; c < 4; ++c) {
                const float dt = __fsub_rn(te[c], ts[c]);
                const float e = expf(-__fmul_rn(s[c], dt));
                a[c] = __fsub_rn(1.0f, e);
                sc[k][c] = dt;
                dir[k][c] = __fmul_rn(__fadd_rn(ga[c], __fmul_rn(gw[c], T[c])), e);
            }
        } else {
            wt.load(p.alphas, k, 0.0f, a);
#pragma unroll
            for (int c = 0; c < 4; ++c) {
                sc[k][c] = fmaxf(__fsub_rn(1.0f, a[c]), 1e-10f);
                dir[k][c] = __fmul_rn(gw[c], T[c]);
            }
        }
#pragma unroll
        for (int c = 0; c < 4; ++c)
            v[k][c] = __fmul_rn(__fadd_rn(gT[c], __fmul_rn(gw[c], a[c])), T[c]);
    }
    float E[kWtNV];
    wt.template scan<SumOp>(v, E, p.status, 1, 0, p.epoch);
#pragma unroll
    for (int k = 0; k < kWtNV; ++k) {
        if (k == kWtVec && !wt.over) break;
        float R[4], o[4];
        finish_exclusive<SumOp>(v[k], wt.f[k], E[k], R);
#pragma unroll
        for (int c = 0; c < 4; ++c) {
            if (DENSITY) o[c] = __fmul_rn(__fsub_rn(dir[k][c], R[c]), sc[k][c]);
            else o[c] = __fsub_rn(dir[k][c], __fdiv_rn(R[c], sc[k][c]));
        }
        wt.store(p.g_in, k, o);
    }
}

static int prep_ws(void* ws, size_t ws_bytes, long long n, uint32_t& epoch, cudaStream_t s) {
    const size_t need = nfb_workspace_bytes(n);
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

static bool all_aligned16(std::initializer_list<const void*> ps) {
    for (const void* q : ps)
        if (q != nullptr && !aligned16(q)) return false;
    return true;
}

}  // namespace nfb

extern "C" {

NFB_API int nfb_render_weight_fwd(int from_density, const float* alphas, const float* t_starts,
                                  const float* t_ends, const float* sigmas, const float* prefix_trans,
                                  const int64_t* packed_info, int64_t n_rays, int64_t uniform_len,
                                  int64_t n, float* weights, float* trans, float* alphas_out,
                                  uint8_t* vis, float early_stop_eps, float alpha_thre,
                                  void* workspace, size_t workspace_bytes, uint32_t epoch,
                                  nfb_stream_t stream) {
    using namespace nfb;
    NFB_CHECK_ARG(n >= 0 && n_rays >= 0, "negative size");
    if (n == 0) return NFB_OK;
    NFB_CHECK_ARG(packed_info != nullptr || uniform_len > 0, "need packed_info or uniform_len");
    NFB_CHECK_ARG(packed_info == nullptr || aligned16(packed_info), "packed_info must be 16-byte aligned");
    if (from_density) NFB_CHECK_ARG(t_starts && t_ends && sigmas, "density mode needs t_starts, t_ends, sigmas");
    else NFB_CHECK_ARG(alphas != nullptr, "alpha mode needs alphas");
    cudaStream_t s = static_cast<cudaStream_t>(stream);
    int rc = prep_ws(workspace, workspace_bytes, n, epoch, s);
    if (rc != NFB_OK) return rc;
    WeightFwdArgs a{};
    a.alphas = alphas; a.t_starts = t_starts; a.t_ends = t_ends; a.sigmas = sigmas; a.prefix = prefix_trans;
    a.weights = weights; a.trans = trans; a.alphas_out = alphas_out; a.vis = vis;
    a.early_stop_eps = early_stop_eps; a.alpha_thre = alpha_thre;
    a.seg = SegSpec{packed_info, n_rays, uniform_len};
    a.n = n; a.n_tiles = num_tiles(n);
    a.status = static_cast<unsigned long long*>(workspace); a.epoch = epoch;
    a.vec_ok = all_aligned16({alphas, t_starts, t_ends, sigmas, prefix_trans, weights, trans, alphas_out}) &&
               (vis == nullptr || (reinterpret_cast<uintptr_t>(vis) & 3u) == 0);
    if (from_density) weight_fwd_kernel<true><<<(unsigned)a.n_tiles, kThreads, 0, s>>>(a);
    else weight_fwd_kernel<false><<<(unsigned)a.n_tiles, kThreads, 0, s>>>(a);
    count_launch();
    NFB_CHECK_LAUNCH("render_weight_fwd");
    return NFB_OK;
}

NFB_API int nfb_render_weight_bwd(int from_density, const float* g_weights, const float* g_trans,
                                  const float* g_alphas, const float* alphas, const float* t_starts,
                                  const float* t_ends, const float* sigmas, const float* trans,
                                  const int64_t* packed_info, int64_t n_rays, int64_t uniform_len,
                                  int64_t n, float* g_in, void* workspace, size_t workspace_bytes,
                                  uint32_t epoch, nfb_stream_t stream) {
    using namespace nfb;
    NFB_CHECK_ARG(n >= 0 && n_rays >= 0, "negative size");
    if (n == 0) return NFB_OK;
    NFB_CHECK_ARG(packed_info != nullptr || uniform_len > 0, "need packed_info or uniform_len");
    NFB_CHECK_ARG(packed_info == nullptr || aligned16(packed_info), "packed_info must be 16-byte aligned");
    NFB_CHECK_ARG(trans && g_in, "null pointer");
    if (from_density) NFB_CHECK_ARG(t_starts && t_ends && sigmas, "density mode needs t_starts, t_ends, sigmas");
    else NFB_CHECK_ARG(alphas != nullptr, "alpha mode needs alphas");
    cudaStream_t s = static_cast<cudaStream_t>(stream);
    int rc = prep_ws(workspace, workspace_bytes, n, epoch, s);
    if (rc != NFB_OK) return rc;
    WeightBwdArgs a{};
    a.g_weights = g_weights; a.g_trans = g_trans; a.g_alphas = g_alphas; a.alphas = alphas;
    a.t_starts = t_starts; a.t_ends = t_ends; a.sigmas = sigmas; a.trans = trans; a.g_in = g_in;
    a.seg = SegSpec{packed_info, n_rays, uniform_len};
    a.n = n; a.n_tiles = num_tiles(n);
    a.status = static_cast<unsigned long long*>(workspace); a.epoch = epoch;
    a.vec_ok = all_aligned16({g_weights, g_trans, g_alphas, alphas, t_starts, t_ends, sigmas, trans, g_in});
    if (from_density) weight_bwd_kernel<true><<<(unsigned)a.n_tiles, kThreads, 0, s>>>(a);
    else weight_bwd_kernel<false><<<(unsigned)a.n_tiles, kThreads, 0, s>>>(a);
    count_launch();
    NFB_CHECK_LAUNCH("render_weight_bwd");
    return NFB_OK;
}

/* nfb_render_weight_fwd / _bwd for a dense packed layout given as head bits (nfb_head_bits). */
NFB_API int nfb_render_weight_bits_fwd(int from_density, const float* alphas, const float* t_starts,
                                       const float* t_ends, const float* sigmas,
                                       const float* prefix_trans, const uint32_t* head_bits, int64_t n,
                                       float* weights, float* trans, float* alphas_out, uint8_t* vis,
                                       float early_stop_eps, float alpha_thre, void* workspace,
                                       size_t workspace_bytes, uint32_t epoch, nfb_stream_t stream) {
    using namespace nfb;
    NFB_CHECK_ARG(n >= 0, "negative size");
    if (n == 0) return NFB_OK;
    NFB_CHECK_ARG(head_bits != nullptr, "null head_bits");
    if (from_density) NFB_CHECK_ARG(t_starts && t_ends && sigmas, "density mode needs t_starts, t_ends, sigmas");
    else NFB_CHECK_ARG(alphas != nullptr, "alpha mode needs alphas");
    cudaStream_t s = static_cast<cudaStream_t>(stream);
    int rc = prep_ws(workspace, workspace_bytes, n, epoch, s);
    if (rc != NFB_OK) return rc;
    WeightFwdArgs a{};
    a.alphas = alphas; a.t_starts = t_starts; a.t_ends = t_ends; a.sigmas = sigmas; a.prefix = prefix_trans;
    a.weights = weights; a.trans = trans; a.alphas_out = alphas_out; a.vis = vis;
    a.early_stop_eps = early_stop_eps; a.alpha_thre = alpha_thre;
    a.bits = head_bits;
    a.n = n; a.n_tiles = num_warp_tiles(n);
    a.status = static_cast<unsigned long long*>(workspace); a.epoch = epoch;
    a.vec_ok = all_aligned16({alphas, t_starts, t_ends, sigmas, prefix_trans, weights, trans, alphas_out}) &&
               (vis == nullptr || (reinterpret_cast<uintptr_t>(vis) & 3u) == 0);
    const unsigned grid = (unsigned)ceil_div<long long>(a.n_tiles, kWarps);
    if (from_density) weight_fwd_bits_kernel<true><<<grid, kThreads, 0, s>>>(a);
    else weight_fwd_bits_kernel<false><<<grid, kThreads, 0, s>>>(a);
    count_launch();
    NFB_CHECK_LAUNCH("render_weight_bits_fwd");
    return NFB_OK;
}

NFB_API int nfb_render_weight_bits_bwd(int from_density, const float* g_weights, const float* g_trans,
                                       const float* g_alphas, const float* alphas, const float* t_starts,
                                       const float* t_ends, const float* sigmas, const float* trans,
                                       const uint32_t* head_bits, int64_t n, float* g_in, void* workspace,
                                       size_t workspace_bytes, uint32_t epoch, nfb_stream_t stream) {
    using namespace nfb;
    NFB_CHECK_ARG(n >= 0, "negative size");
    if (n == 0) return NFB_OK;
    NFB_CHECK_ARG(head_bits && trans && g_in, "null pointer");
    if (from_density) NFB_CHECK_ARG(t_starts && t_ends && sigmas, "density mode needs t_starts, t_ends, sigmas");
    else NFB_CHECK_ARG(alphas != nullptr, "alpha mode needs alphas");
    cudaStream_t s = static_cast<cudaStream_t>(stream);
    int rc = prep_ws(workspace, workspace_bytes, n, epoch, s);
    if (rc != NFB_OK) return rc;
    WeightBwdArgs a{};
    a.g_weights = g_weights; a.g_trans = g_trans; a.g_alphas = g_alphas; a.alphas = alphas;
    a.t_starts = t_starts; a.t_ends = t_ends; a.sigmas = sigmas; a.trans = trans; a.g_in = g_in;
    a.bits = head_bits;
    a.n = n; a.n_tiles = num_warp_tiles(n);
    a.status = static_cast<unsigned long long*>(workspace); a.epoch = epoch;
    a.vec_ok = all_aligned16({g_weights, g_trans, g_alphas, alphas, t_starts, t_ends, sigmas, trans, g_in});
    const unsigned grid = (unsigned)ceil_div<long long>(a.n_tiles, kWarps);
    if (from_density) weight_bwd_bits_kernel<true><<<grid, kThreads, 0, s>>>(a);
    else weight_bwd_bits_kernel<false><<<grid, kThreads, 0, s>>>(a);
    count_launch();
    NFB_CHECK_LAUNCH("render_weight_bits_bwd");
    return NFB_OK;
}

/* OccGridEstimator.sampling's "skip invisible space" step (occ_grid.py:186-220) in one pass. */
NFB_API int nfb_visibility_compact(int from_density, const float* alphas, const float* t_starts,
                                   const float* t_ends, const float* sigmas, const float* prefix_trans,
                                   const int64_t* packed_info, int64_t n_rays, int64_t n,
                                   const int64_t* ray_indices, float early_stop_eps, float alpha_thre,
                                   int64_t* out_ray_indices, float* out_t_starts, float* out_t_ends,
                                   int64_t* out_packed_info, uint8_t* vis, int64_t* total,
                                   void* workspace, size_t workspace_bytes, uint32_t epoch,
                                   nfb_stream_t stream) {
    using namespace nfb;
    NFB_CHECK_ARG(n >= 0 && n_rays >= 0, "negative size");
    cudaStream_t s = static_cast<cudaStream_t>(stream);
    NFB_CHECK_ARG(total != nullptr, "null total");
    if (n == 0) {
        cudaMemsetAsync(total, 0, sizeof(int64_t), s);
        if (out_packed_info && n_rays > 0) cudaMemsetAsync(out_packed_info, 0, sizeof(int64_t) * 2 * n_rays, s);
        return NFB_OK;
    }
    NFB_CHECK_ARG(packed_info != nullptr && aligned16(packed_info), "packed_info (canonical, dense) is required, 16-byte aligned");
    NFB_CHECK_ARG(ray_indices && t_starts && t_ends && out_ray_indices && out_t_starts && out_t_ends, "null pointer");
    if (from_density) NFB_CHECK_ARG(sigmas != nullptr, "density mode needs sigmas");
    else NFB_CHECK_ARG(alphas != nullptr, "alpha mode needs alphas");
    int rc = prep_ws(workspace, workspace_bytes, n, epoch, s);
    if (rc != NFB_OK) return rc;
    VisCompactArgs a{};
    a.w.alphas = alphas; a.w.t_starts = t_starts; a.w.t_ends = t_ends; a.w.sigmas = sigmas; a.w.prefix = prefix_trans;
    a.w.vis = vis; a.w.early_stop_eps = early_stop_eps; a.w.alpha_thre = alpha_thre;
    a.w.seg = SegSpec{packed_info, n_rays, 0};
    a.w.n = n; a.w.n_tiles = num_tiles(n);
    a.w.status = static_cast<unsigned long long*>(workspace); a.w.epoch = epoch;
    a.w.vec_ok = all_aligned16({alphas, t_starts, t_ends, sigmas, prefix_trans}) &&
                 (vis == nullptr || (reinterpret_cast<uintptr_t>(vis) & 3u) == 0);
    a.ray_indices = ray_indices; a.t_starts = t_starts; a.t_ends = t_ends;
    a.o_ray = out_ray_indices; a.o_ts = out_t_starts; a.o_te = out_t_ends;
    a.o_packed_info = out_packed_info; a.total = reinterpret_cast<long long*>(total);
    if (from_density) vis_compact_kernel<true><<<(unsigned)a.w.n_tiles, kThreads, 0, s>>>(a);
    else vis_compact_kernel<false><<<(unsigned)a.w.n_tiles, kThreads, 0, s>>>(a);
    count_launch();
    if (out_packed_info && n_rays > 0) {
        long long b = ceil_div<long long>(n_rays, 256);
        if (b > (long long)sm_count() * 8) b = (long long)sm_count() * 8;
        pi_counts_kernel<<<(unsigned)b, 256, 0, s>>>(out_packed_info, n_rays, a.total);
        count_launch();
    }
    NFB_CHECK_LAUNCH("visibility_compact");
    return NFB_OK;
}

}  // extern "C"
