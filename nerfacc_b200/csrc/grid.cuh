// grid.cuh — per-ray multi-level DDA march shared by the count / fill kernels.
//
// Semantics follow the reference kernel line by line (csrc/grid.cu:98-281 with
// include/utils_grid.cuh:10-142, data_spec_packed.cuh:43-72, utils_contraction.cuh:19-24):
// sample counts are decided by float compares, so every operation is spelled with a
// round-to-nearest intrinsic.  The three places where nvcc contracts the reference's
// expressions into FFMA are explicit __fmaf_rn here; nothing else may
// fuse (the library is also built with -fmad=false).
//
// What is NOT like the reference: how the marching recurrence is evaluated.  The reference
// advances t by repeated addition, one dependent FADD per step, both to skip empty space
//     while (t + dt*0.5f < target) t += dt;                          (grid.cu:157-161, 199-203)
// and to emit samples (grid.cu:208-262).  A closed form t0 + k*dt is not allowed (it rounds
// differently), but the recurrence itself has an exact closed form per binade: while t stays
// in [2^e, 2^(e+1)] it is a multiple of u = ulp(2^e), so RN(t + dt) = t + q with the constant
// q = RN_u(dt) (ties excepted), and t_k = t + k*q holds EXACTLY.  advance() below finds the
// number of steps with integer mantissa arithmetic plus an exact predicate check, takes the
// binade crossings (and the rare tie case) as single real steps, and hands the marcher
// "runs" (t0, q, n) of samples whose edges are t0 + j*q — bit-identical to the serial loop,
// O(1) per grid cell instead of O(cell length / dt), and expandable by a whole warp.
//
// Differences from the reference, all outside its defined behaviour:
//  * occupancy is read from a bit-packed copy of `binaries` (1 bit/cell, L1/L2 resident);
//  * the DDA stops if an index leaves the grid (the reference would read out of bounds);
//  * a march that can no longer advance (t + dt == t, or NaN) stops (the reference spins).
#pragma once

#include <cmath>
#include <cstring>

#include "common.cuh"

namespace nfb {

constexpr int kMaxLevels = 8;  // MAX_GRID_LEVELS in csrc/grid.cu:18

struct GridSpec {
    const uint32_t* __restrict__ occ_bits;  // [n_grids * cells_per_level] bits, C order [m,x,y,z]
    const float* __restrict__ aabbs;        // [n_grids, 6]
    int n_grids;
    int res[3];
    long long cells_per_level;
    // "any occupied" mip: one bit per (2^shift)^3 block of cells, [n_grids, mres0, mres1, mres2]
    // in C order.  `mip` points to global memory in the kernel parameters; marching kernels copy
    // it to shared memory and march with `mip` redirected there (nullptr: no mip).  97 % of the
    // cells of a trained scene are empty: their lookups then never leave the SM.
    const uint32_t* mip;
    int mip_shift;
    int mres[3];
    int mip_cells_per_level;
    int mip_words;
};

// Copy the mip into shared memory (all threads of the CTA; ends with a barrier) and return a
// GridSpec whose `mip` points at the copy.
__device__ __forceinline__ GridSpec stage_mip(const GridSpec& G, uint32_t* smem) {
    GridSpec L = G;
    if (G.mip != nullptr) {
        for (int i = threadIdx.x; i < G.mip_words; i += blockDim.x) smem[i] = __ldg(G.mip + i);
        L.mip = smem;
    }
    __syncthreads();
    return L;
}

struct MarchOpts {
    float step_size;
    float cone_angle;
    int limit;  // traverse_steps_limit, <= 0: none
};

// Sorted ray/box events of one ray: 2*m values, the index (< m: entering level idx, else
// leaving level idx-m) and the per-level hit bits.
struct RayEvents {
    float t[2 * kMaxLevels];
    int8_t idx[2 * kMaxLevels];
    uint32_t hit_mask;
};

__device__ __forceinline__ float calc_dt(float t, float cone, float dt_min) {
    // grid.cu:23-28 with clamp(f,a,b) = fmaxf(a, fminf(f, b)) (utils_math.cuh:1167-1170)
    return fmaxf(dt_min, fminf(__fmul_rn(t, cone), 1e10f));
}

// utils_grid.cuh:10-55.  Returns hit; tmin/tmax clipped to [near, far].
__device__ __forceinline__ bool slab_test(const float (&o)[3], const float (&inv)[3],
                                          const float* __restrict__ bb, float near, float far,
                                          float& tmin, float& tmax) {
    float lo[3], hi[3];
#pragma unroll
    for (int k = 0; k < 3; ++k) {
        const float a = __fmul_rn(__fsub_rn(bb[k], o[k]), inv[k]);
        const float b = __fmul_rn(__fsub_rn(bb[3 + k], o[k]), inv[k]);
        if (inv[k] >= 0) { lo[k] = a; hi[k] = b; }
        else { lo[k] = b; hi[k] = a; }
    }
    tmin = lo[0]; tmax = hi[0];
    if (tmin > hi[1] || lo[1] > tmax) return false;
    if (lo[1] > tmin) tmin = lo[1];
    if (hi[1] < tmax) tmax = hi[1];
    if (tmin > hi[2] || lo[2] > tmax) return false;
    if (lo[2] > tmin) tmin = lo[2];
    if (hi[2] < tmax) tmax = hi[2];
    if (tmax <= 0) return false;
    tmin = fmaxf(tmin, near);
    tmax = fminf(tmax, far);
    return true;
}

// grid.py:158-162 fused: intersect all levels (near=-inf, far=inf, miss=inf), then sort the
// 2m values ascending (stable on ties, like a stable argsort).
__device__ __forceinline__ void compute_events(const float (&o)[3], const float (&inv)[3],
                                               const GridSpec& G, RayEvents& ev) {
    const int m = G.n_grids;
    ev.hit_mask = 0;
    const float inf = __int_as_float(0x7f800000);
    for (int l = 0; l < m; ++l) {
        float t0, t1;
        const bool hit = slab_test(o, inv, G.aabbs + 6 * l, -inf, inf, t0, t1);
        ev.t[l] = hit ? t0 : inf;
        ev.t[m + l] = hit ? t1 : inf;
        ev.idx[l] = (int8_t)l;
        ev.idx[m + l] = (int8_t)(m + l);
        if (hit) ev.hit_mask |= (1u << l);
    }
    for (int i = 1; i < 2 * m; ++i) {  // stable insertion sort, <= 16 entries
        const float tv = ev.t[i];
        const int8_t iv = ev.idx[i];
        int j = i - 1;
        while (j >= 0 && ev.t[j] > tv) { ev.t[j + 1] = ev.t[j]; ev.idx[j + 1] = ev.idx[j]; --j; }
        ev.t[j + 1] = tv;
        ev.idx[j + 1] = iv;
    }
}

__device__ __forceinline__ int clampi(int v, int lo, int hi) { return max(lo, min(v, hi)); }

// ---------------------------------------------------------------------------------------------
// advance(): exact evaluation of
//     n = 0; while (n < cap && !(fmaf(dt, .5f, t) >= target)) { t1 = t + dt; if (!(t1 > t)) break;
//                                                              piece(t, t1 - t, 1); t = t1; ++n; }
// in O(#binades crossed) instead of O(n).  `piece(t0, q, k)` is called for every maximal stretch
// of k steps with constant increment q (t0 + j*q are exactly the serial values); `steps` is
// incremented by the total.  dt must be > 0.
// Round-to-nearest primitives that also compile for the host, so tests/host/advance_check.cu
// can run advance() against the serial loop on the CPU (x86-64 SSE arithmetic is IEEE; the
// host build uses -ffp-contract=off).
#ifdef __CUDA_ARCH__
#define NFB_HD __host__ __device__ __forceinline__
NFB_HD float rn_add(float a, float b) { return __fadd_rn(a, b); }
NFB_HD float rn_sub(float a, float b) { return __fsub_rn(a, b); }
NFB_HD float rn_mul(float a, float b) { return __fmul_rn(a, b); }
NFB_HD float rn_div(float a, float b) { return __fdiv_rn(a, b); }
NFB_HD float rn_fma(float a, float b, float c) { return __fmaf_rn(a, b, c); }
NFB_HD int f2i_rn(float a) { return __float2int_rn(a); }
NFB_HD float i2f_rn(int a) { return __int2float_rn(a); }
NFB_HD uint32_t f_bits(float a) { return __float_as_uint(a); }
NFB_HD float bits_f(uint32_t a) { return __uint_as_float(a); }
#else
#define NFB_HD __host__ __device__ inline
NFB_HD float rn_add(float a, float b) { volatile float r = a + b; return r; }
NFB_HD float rn_sub(float a, float b) { volatile float r = a - b; return r; }
NFB_HD float rn_mul(float a, float b) { volatile float r = a * b; return r; }
NFB_HD float rn_div(float a, float b) { volatile float r = a / b; return r; }
NFB_HD float rn_fma(float a, float b, float c) { return ::fmaf(a, b, c); }
NFB_HD int f2i_rn(float a) { return (int)::lrintf(a); }
NFB_HD float i2f_rn(int a) { return (float)a; }
NFB_HD uint32_t f_bits(float a) { uint32_t r; memcpy(&r, &a, 4); return r; }
NFB_HD float bits_f(uint32_t a) { float r; memcpy(&r, &a, 4); return r; }
#endif

template <class Piece>
NFB_HD float advance(float t, const float dt, const float target, const int cap, int& steps,
                     Piece&& piece) {
    if (!(dt > 0.0f) || target != target) return t;
    while (steps < cap) {
        if (rn_fma(dt, 0.5f, t) >= target) break;
        const float t1 = rn_add(t, dt);
        if (!(t1 > t)) break;  // stuck at this magnitude (or NaN / inf)
        const float q = rn_sub(t1, t);  // exact: t1 is a grid point reached from t
        const uint32_t tb = f_bits(t);
        const int E = (int)(tb >> 23);  // sign bit set (t < 0) makes E >= 256: excluded below
        if (E >= 24 && E < 254) {
            const float u = bits_f((uint32_t)(E - 23) << 23);   // ulp of t's binade
            const float top = bits_f((uint32_t)(E + 1) << 23);  // 2^(e+1)
            const float r = rn_sub(dt, q);                      // exact, |r| <= u/2
            if (t1 <= top && rn_mul(fabsf(r), 2.0f) != u) {
                // linear regime: t_k = t + k*q for every k with t + k*q <= top
                const int ti = (int)((tb & 0x7fffffu) | 0x800000u);
                const int qi = f2i_rn(rn_mul(q, bits_f((uint32_t)(277 - E) << 23)));
                int kmax = (0x1000000 - ti) / qi;  // >= 1 because t1 <= top
                if (kmax > cap - steps) kmax = cap - steps;
                auto t_at = [&](int k) { return rn_mul(i2f_rn(ti + k * qi), u); };
                auto done = [&](int k) { return rn_fma(dt, 0.5f, t_at(k)) >= target; };
                int k = kmax;
                if (kmax > 1 && done(kmax)) {
                    // smallest k in [1, kmax] with done(k): estimate, then fix up (monotone)
                    const float gap = rn_sub(target, rn_fma(dt, 0.5f, t));
                    const float kf = ceilf(rn_div(gap, q));
                    k = (kf >= (float)kmax) ? kmax : ((kf <= 1.0f) ? 1 : (int)kf);
                    while (k > 1 && done(k - 1)) --k;
                    while (!done(k)) ++k;
                }
                piece(t, q, k);
                steps += k;
                t = t_at(k);
                continue;
            }
        }
        piece(t, q, 1);  // binade crossing, tie case, tiny / negative t: one real step
        steps += 1;
        t = t1;
    }
    return t;
}

struct NoPiece {
    NFB_HD void operator()(float, float, int) const {}
};

// ---------------------------------------------------------------------------------------------
// The marching engine.
//
// The reference walks every ray with one thread: DDA step, occupancy test, and — inside the
// same loop — the skip / emit recurrences.  On a GPU that interleaving is what hurts: lanes of
// a warp reach occupied cells at different times, so the expensive emit code runs with two or
// three lanes active.  Here the work of a ray is split into a PRODUCER and a CONSUMER that a
// warp executes in lock step:
//   produce   a tight, branch-free DDA loop (<= kRoundSteps cells per round) that touches only
//             the grid and appends EVENTS (kind, t) to a per-lane queue in shared memory.  With
//             a constant dt consecutive cells of the same kind fold into one event, because both
//             recurrences are idempotent in their target: a ray through a trained scene yields
//             a handful of events, not hundreds;
//   consume   when a queue fills up, all lanes replay their queued events together through
//             ONE advance() call site.  The queue also carries the ray boundaries (EV_RAY), so
//             the producer of a lane may be several rays ahead of its consumer and the replay
//             finds every lane with work.
// Lanes whose DDA has finished a ray fetch the next one from a global counter (persistent
// threads).  Results are addressed by ray id; nothing depends on which lane marched which ray.
//
// Event kinds (t = exit of the cell / start of the segment):
//   EV_EMPTY  empty cell(s):     skip to t, continuous = false        (grid.cu:189-206)
//   EV_OCC    occupied cell(s):  emit samples up to t                 (grid.cu:208-262)
//   EV_SEG    segment start:     if (!continuous) skip to t           (grid.cu:150-163)
//   EV_RAY    a new ray starts:  payload = ray id; retires the previous ray of the lane
constexpr int kQ = 16;           // events queued per lane
constexpr int kRoundSteps = 32;  // DDA cells per lane and round
enum { EV_EMPTY = 0, EV_OCC = 1, EV_SEG = 2, EV_RAY = 3 };

// Policy interface (see grid.cu):
//   typename Sink;  bool accept(r);  void skipped(r);  Sink begin(r);
//   void end(r, Sink&, nS, nI, t_last);
// Sink: run(t0, q, k, continuous, nS, nI) — k samples with edges t0 + j*q (constant-dt runs) —
// and one(t_last, t_next, continuous, nS, nI); `continuous`, nS, nI are the state before the call.
template <bool MIP, class Policy>
__device__ __forceinline__ void march_engine(const float* __restrict__ rays_o,
                                             const float* __restrict__ rays_d,
                                             const float* __restrict__ near_arr, float near_s,
                                             const float* __restrict__ far_arr, float far_s,
                                             const float* __restrict__ t_sorted,
                                             const int64_t* __restrict__ t_indices,
                                             const uint8_t* __restrict__ hits, long long n_rays,
                                             const GridSpec& G, const MarchOpts& opt, Policy& P,
                                             float* __restrict__ s_evt /* [kQ][blockDim.x] */,
                                             unsigned long long* __restrict__ ray_counter) {
    const float eps = 1e-6f;
    const int m = G.n_grids;
    const float step = opt.step_size, cone = opt.cone_angle;
    const int cap = opt.limit > 0 ? opt.limit : 0x7fffffff;
    const bool const_dt = (step > 0.0f) && (cone == 0.0f);
    const bool fold_empty = const_dt || step <= 0.0f;  // consecutive empty cells fold
    const bool fold_occ = const_dt;                    // consecutive occupied cells fold
    const bool eager = opt.limit > 0;  // with a sample cap, replay every round (bounds wasted DDA work)
    const int lane = lane_id();
    const int tid = threadIdx.x, nthr = blockDim.x;

    // ---- producer state (the DDA of the ray being walked)
    bool have_ray = false, exhausted = false;
    float far = 0.0f, near = 0.0f;
    long long r = -1;
    RayEvents ev;
    int seg_i = 0, seg_steps = 0;
    const int max_seg_steps = G.res[0] + G.res[1] + G.res[2] + 8;  // a DDA cannot take more
    bool in_seg = false, first_seg = true;
    float tdist[3], delta[3], tmax_seg = 0.0f;
    int cur[3], stepi[3], stop[3], cstride[3];
    int cell = 0;  // linear index of the current cell (host side guarantees it fits 31 bits)
    int mip_base = 0;
    int pend_kind = -1;
    float pend_t = 0.0f;
    // ---- queue
    int nq = 0;
    uint32_t kinds = 0;
    // ---- consumer state (the t_last recurrence of the ray being replayed)
    bool c_have = false, continuous = false, capped = false;
    long long c_r = -1;
    float t_last = 0.0f;
    int nS = 0, nI = 0, dummy = 0;
    typename Policy::Sink sink;

    auto push = [&](int kind, float t) {
        s_evt[nq * nthr + tid] = t;
        kinds |= (uint32_t)kind << (2 * nq);
        ++nq;
    };
    auto push_pending = [&]() {
        if (pend_kind >= 0) { push(pend_kind, pend_t); pend_kind = -1; }
    };

    auto consume = [&]() {
        const int maxq = __reduce_max_sync(kFull, nq);
#pragma unroll 1
        for (int j = 0; j < maxq; ++j) {
            if (j >= nq) continue;
            const int kind = (int)((kinds >> (2 * j)) & 3u);
            const float t = s_evt[j * nthr + tid];
            if (kind == EV_RAY) {
                if (c_have) P.end(c_r, sink, nS, nI, t_last);
                c_r = (long long)__float_as_int(t);
                t_last = near_arr ? __ldg(near_arr + c_r) : near_s;
                continuous = false; capped = false; nS = 0; nI = 0;
                sink = P.begin(c_r);
                c_have = true;
                continue;
            }
            if (capped) continue;
            const bool emit = (kind == EV_OCC);
            const bool doit = !(kind == EV_SEG && continuous);
            if (step <= 0.0f) {  // one interval per occupied voxel crossing
                if (!emit) {
                    if (doit) t_last = t;
                } else {
                    sink.one(t_last, t, continuous, nS, nI);
                    nI += continuous ? 1 : 2;
                    nS += 1;
                    continuous = true;
                    t_last = t;
                }
            } else if (const_dt || !emit) {
                if (doit) {
                    const float dt = const_dt ? step : calc_dt(t_last, cone, step);
                    if (emit) {
                        t_last = advance(t_last, dt, t, cap, nS, [&](float t0, float q, int k) {
                            sink.run(t0, q, k, continuous, nS, nI);
                            nI += continuous ? k : k + 1;
                            continuous = true;
                        });
                    } else {
                        t_last = advance(t_last, dt, t, 0x7fffffff, dummy, NoPiece());
                    }
                }
            } else {  // cone_angle > 0: dt follows t, a true serial recurrence
                while (nS < cap) {
                    const float dt = calc_dt(t_last, cone, step);
                    if (__fmaf_rn(dt, 0.5f, t_last) >= t) break;
                    const float t_next = __fadd_rn(t_last, dt);
                    if (!(t_next > t_last)) break;  // stuck
                    sink.one(t_last, t_next, continuous, nS, nI);
                    nI += continuous ? 1 : 2;
                    nS += 1;
                    continuous = true;
                    t_last = t_next;
                    if (t_next >= t) break;
                }
            }
            if (kind == EV_EMPTY) continuous = false;
            if (nS >= cap) {
                capped = true;
                if (have_ray && r == c_r) { have_ray = false; pend_kind = -1; }  // stop walking it
            }
        }
        nq = 0;
        kinds = 0;
    };

    while (true) {
        // ------------------------------------------------------------------ refill
        // (needs room for the EV_RAY marker and the first events of the ray)
#pragma unroll 1
        for (int attempt = 0; attempt < 4; ++attempt) {
            const bool need = !have_ray && !exhausted && nq <= kQ - 4;
            const unsigned mask = __ballot_sync(kFull, need);
            if (mask == 0u) break;
            unsigned long long base = 0;
            if (lane == __ffs(mask) - 1) base = atomicAdd(ray_counter, (unsigned long long)__popc(mask));
            base = __shfl_sync(kFull, base, __ffs(mask) - 1);
            if (need) {
                const long long cand = (long long)base + __popc(mask & ((1u << lane) - 1u));
                if (cand >= n_rays) {
                    exhausted = true;
                } else if (!P.accept(cand)) {
                    P.skipped(cand);
                } else {
                    r = cand;
                    float o[3], d[3], inv[3];
#pragma unroll
                    for (int k = 0; k < 3; ++k) {
                        o[k] = __ldg(rays_o + 3 * r + k);
                        d[k] = __ldg(rays_d + 3 * r + k);
                        inv[k] = __fdiv_rn(1.0f, d[k]);
                    }
                    near = near_arr ? __ldg(near_arr + r) : near_s;
                    far = far_arr ? __ldg(far_arr + r) : far_s;
                    if (t_sorted != nullptr) {
                        ev.hit_mask = 0;
                        for (int l = 0; l < m; ++l)
                            if (hits[r * m + l]) ev.hit_mask |= (1u << l);
                        for (int i = 0; i < 2 * m; ++i) {
                            ev.t[i] = __ldg(t_sorted + r * 2 * m + i);
                            ev.idx[i] = (int8_t)__ldg(t_indices + r * 2 * m + i);
                        }
                    } else {
                        compute_events(o, inv, G, ev);
                    }
                    have_ray = true; in_seg = false; seg_i = 0; pend_kind = -1; first_seg = true;
                    push(EV_RAY, __int_as_float((int)r));
                }
            }
        }
        if (!__any_sync(kFull, have_ray)) {
            if (__all_sync(kFull, exhausted || nq > kQ - 4)) {
                consume();
                if (__all_sync(kFull, exhausted)) break;
            }
            continue;
        }
        // ------------------------------------------------------------------ produce
        int steps = 0;
        while (have_ray && steps < kRoundSteps && nq <= kQ - 3) {
            if (!in_seg) {
                // next ray/box segment (grid.cu:128-149)
                bool found = false;
                float tmin = 0.0f;
                int level = 0;
                while (seg_i < 2 * m - 1) {
                    const int i = seg_i++;
                    const bool entering = ev.idx[i] < m;
                    level = ev.idx[i] % m;
                    if (!((ev.hit_mask >> level) & 1u)) continue;
                    if (!entering) {
                        if (ev.idx[i + 1] < m) continue;  // next event enters a box: we are in a gap
                        level = ev.idx[i + 1] % m;
                        if (!((ev.hit_mask >> level) & 1u)) continue;
                    }
                    tmin = fmaxf(ev.t[i], near);
                    tmax_seg = fminf(ev.t[i + 1], far);
                    if (tmin >= tmax_seg) continue;
                    found = true;
                    break;
                }
                if (!found) { push_pending(); have_ray = false; break; }
                // "if (!continuous) skip to tmin": right after the ray starts or after empty
                // cells `continuous` is known to be false and the event is a plain skip
                if (first_seg) {
                    pend_kind = EV_EMPTY; pend_t = tmin; first_seg = false;
                } else if (pend_kind == EV_EMPTY && fold_empty) {
                    pend_t = tmin;
                } else {
                    push_pending();
                    push(EV_SEG, tmin);
                }
                // setup_traversal, utils_grid.cuh:58-114 (the ray is re-read: it is not kept in
                // registers across the DDA loop)
                const float* __restrict__ bb = G.aabbs + 6 * level;
                const float ts = __fadd_rn(tmin, eps), te = __fsub_rn(tmax_seg, eps);
#pragma unroll
                for (int k = 0; k < 3; ++k) {
                    const float ok = __ldg(rays_o + 3 * r + k), dk = __ldg(rays_d + 3 * r + k);
                    const float invk = __fdiv_rn(1.0f, dk);
                    const float bmin = bb[k], bmax = bb[3 + k];
                    const float resf = (float)G.res[k];
                    const float ext = __fsub_rn(bmax, bmin);
                    const float vox = __fdiv_rn(ext, resf);
                    const float rs = __fmaf_rn(dk, ts, ok);  // FFMA site 1
                    const float re = __fmaf_rn(dk, te, ok);
                    cur[k] = clampi(__float2int_rz(__fmul_rn(__fdiv_rn(__fsub_rn(rs, bmin), ext), resf)), 0, G.res[k] - 1);
                    const int fin = clampi(__float2int_rz(__fmul_rn(__fdiv_rn(__fsub_rn(re, bmin), ext), resf)), 0, G.res[k] - 1);
                    const int start_index = cur[k] + (dk > 0.0f ? 1 : 0);
                    // FFMA site 2: fma(bmin + fma(float(idx), vox, -rs), inv, tmin)
                    const float tm = __fmaf_rn(__fadd_rn(bmin, __fmaf_rn((float)start_index, vox, -rs)), invk, tmin);
                    const float sf = (dk == 0.0f) ? 0.0f : (dk > 0.0f ? 1.0f : -1.0f);
                    tdist[k] = (dk == 0.0f) ? tmax_seg : tm;
                    stepi[k] = (int)sf;
                    delta[k] = (dk == 0.0f) ? tmax_seg : __fmul_rn(__fmul_rn(vox, invk), sf);
                    // The walk leaves when the index reaches fin + step (utils_grid.cuh:123) or the
                    // grid edge, whichever lies ahead: one stop value per axis.
                    const int ovf = fin + stepi[k];
                    const bool ahead = stepi[k] > 0 ? ovf > cur[k] : (stepi[k] < 0 ? ovf < cur[k] : true);
                    stop[k] = ahead ? ovf : (stepi[k] > 0 ? G.res[k] : -1);
                }
                cstride[0] = stepi[0] * G.res[1] * G.res[2];
                cstride[1] = stepi[1] * G.res[2];
                cstride[2] = stepi[2];
                cell = level * (int)G.cells_per_level + (cur[0] * G.res[1] + cur[1]) * G.res[2] + cur[2];
                mip_base = level * G.mip_cells_per_level;
                in_seg = true;
                seg_steps = 0;
                continue;
            }
            // the DDA proper (grid.cu:183-281), branch-free; leaves the loop when the segment
            // ends, the round is over or the queue is (nearly) full
            do {
                const float t_trav = fminf(fminf(tdist[0], fminf(tdist[1], tdist[2])), tmax_seg);
                uint32_t cbit = 1u;
                if (MIP) {
                    const int ci = mip_base + ((cur[0] >> G.mip_shift) * G.mres[1] + (cur[1] >> G.mip_shift)) * G.mres[2] +
                                   (cur[2] >> G.mip_shift);
                    cbit = (G.mip[ci >> 5] >> (ci & 31)) & 1u;
                }
                // fine bit, fetched only under a set mip bit
                uint32_t word = 0u;
                if (cbit) word = __ldg(G.occ_bits + (cell >> 5));
                const int kind = (int)((word >> (cell & 31)) & 1u);  // EV_OCC == 1, EV_EMPTY == 0
                if (kind != pend_kind || !(kind == EV_OCC ? fold_occ : fold_empty)) { push_pending(); pend_kind = kind; }
                pend_t = t_trav;
                // single_traversal, utils_grid.cuh:116-142: x if strictly smallest, else y if < z, else z
                const bool px = tdist[0] < tdist[1] && tdist[0] < tdist[2];
                const bool py = !px && tdist[1] < tdist[2];
                const bool pz = !px && !py;
                cur[0] += px ? stepi[0] : 0;
                cur[1] += py ? stepi[1] : 0;
                cur[2] += pz ? stepi[2] : 0;
                tdist[0] = px ? __fadd_rn(tdist[0], delta[0]) : tdist[0];
                tdist[1] = py ? __fadd_rn(tdist[1], delta[1]) : tdist[1];
                tdist[2] = pz ? __fadd_rn(tdist[2], delta[2]) : tdist[2];
                cell += px ? cstride[0] : (py ? cstride[1] : cstride[2]);
                const bool leave = (px && cur[0] == stop[0]) || (py && cur[1] == stop[1]) || (pz && cur[2] == stop[2]);
                if (leave || ++seg_steps > max_seg_steps) in_seg = false;  // (the bound only trips on NaN rays)
                ++steps;
            } while (in_seg && steps < kRoundSteps && nq <= kQ - 3);
        }
        // ------------------------------------------------------------------ consume
        if (eager || __any_sync(kFull, nq > kQ - 6)) consume();
    }
    if (c_have) P.end(c_r, sink, nS, nI, t_last);
}

}  // namespace nfb
