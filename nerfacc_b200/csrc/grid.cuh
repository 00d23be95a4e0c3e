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
    // per level the inclusive index box (x0,y0,z0,x1,y1,z1) of the occupied cells, or nullptr
    const int* __restrict__ bounds;
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
    int batch;  // rays a warp fetches at a time (1..32; set by the launcher from the launch size)
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

// HALF = false: the same with the stop test `t >= target` (the DDA's own recurrence tdist += delta,
// consumed up to a time; see jump_ahead).
template <bool HALF = true, class Piece>
NFB_HD float advance(float t, const float dt, const float target, const int cap, int& steps,
                     Piece&& piece) {
    if (!(dt > 0.0f) || target != target) return t;
    while (steps < cap) {
        if ((HALF ? rn_fma(dt, 0.5f, t) : t) >= target) break;
        const float t1 = rn_add(t, dt);
        if (!(t1 > t)) break;  // stuck at this magnitude (or NaN / inf)
        const float q = rn_sub(t1, t);  // exact: t1 is a grid point reached from t
        const uint32_t tb = f_bits(t);
        const int E = (int)(tb >> 23);  // sign bit set (t < 0) makes E >= 256: excluded below
        if (E >= 24 && E < 254) {
            const float u = bits_f((uint32_t)(E - 23) << 23);   // ulp of t's binade
            const float top = bits_f((uint32_t)(E + 1) << 23);  // 2^(e+1)
            const float r = rn_sub(dt, q);                      // exact, |r| <= u/2
            const int ti = (int)((tb & 0x7fffffu) | 0x800000u);
            // A tie (dt exactly halfway between two points of the binade's grid) rounds to the even
            // mantissa: from an even mantissa every step adds the same even q (the stretch below is exact
            // then too); from an odd one, a single literal step lands on an even one.  One (ray, axis) in
            // 64 has such a delta at t in [2, 4), and stepping it literally held its whole warp back.
            if (t1 <= top && (rn_mul(fabsf(r), 2.0f) != u || (ti & 1) == 0)) {
                // linear regime: t_k = t + k*q for every k with t + k*q <= top
                const int qi = f2i_rn(rn_mul(q, bits_f((uint32_t)(277 - E) << 23)));
                int kmax = (0x1000000 - ti) / qi;  // >= 1 because t1 <= top
                if (kmax > cap - steps) kmax = cap - steps;
                auto t_at = [&](int k) { return rn_mul(i2f_rn(ti + k * qi), u); };
                auto done = [&](int k) { return (HALF ? rn_fma(dt, 0.5f, t_at(k)) : t_at(k)) >= target; };
                int k = kmax;
                if (kmax > 1 && done(kmax)) {
                    // smallest k in [1, kmax] with done(k): estimate, then fix up (monotone)
                    const float gap = rn_sub(target, HALF ? rn_fma(dt, 0.5f, t) : t);
                    const float kf = ceilf(rn_div(gap, q));
                    k = (kf >= (float)kmax) ? kmax : ((kf <= 1.0f) ? 1 : (int)kf);
                    while (k > 1 && done(k - 1)) --k;
                    while (!done(k)) ++k;
                }
                piece(t, q, k);
                steps += k;
                t = t_at(k);
                // The step after a stretch — usually the one that crosses into the next binade — is taken
                // right here, literally, instead of costing a trip around the loop (with 32 lanes, one of
                // them nearly always crosses a binade, and the others wait for it).
                if (steps < cap && !((HALF ? rn_fma(dt, 0.5f, t) : t) >= target)) {
                    const float t2 = rn_add(t, dt);
                    if (t2 > t) {
                        piece(t, rn_sub(t2, t), 1);
                        steps += 1;
                        t = t2;
                    }
                }
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

// Checkpoints of the common trajectory.  With a constant dt and one near plane for all rays,
// every ray's t_last walks the SAME sequence near, near+dt, ... (rounded as the hardware rounds):
// rays differ only in where they stop.  A CTA records a point of that sequence per stretch of
// constant increment (about one per binade) once; a long skip — a ray's first one runs from the
// near plane to the scene, a dozen binades — then starts from the last checkpoint that is not
// past its target instead of re-deriving every binade on the way.
constexpr int kMaxCk = 64;
__device__ __forceinline__ int make_checkpoints(float near, float dt, float* __restrict__ ck) {
    int n = 0, cap_hits = 0;
    float t = near;
    while (n < kMaxCk && cap_hits < 4) {
        ck[n++] = t;
        // to the end of t's binade (or 4096 steps, whichever comes first)
        const uint32_t E = __float_as_uint(t) >> 23;
        const bool binade = t > 0.0f && E >= 1u && E < 254u;  // else (zero, negative, denormal): single steps
        int cnt = 0;
        float t2 = t;
        if (binade) t2 = advance(t, dt, __uint_as_float((E + 1u) << 23), 4096, cnt, NoPiece());
        if (cnt >= 4096) ++cap_hits;
        float t3 = t2;
        if (cnt == 0) {  // already within dt/2 of the binade's top: take the crossing step itself
            t3 = __fadd_rn(t2, dt);
            if (!(t3 > t2)) break;
        }
        t = t3;
    }
    return n;
}
// last checkpoint c with !(fma(dt, .5, c) >= target), or `none` if even the first one is past
__device__ __forceinline__ float checkpoint_before(const float* __restrict__ ck, int n, float dt, float target,
                                                   float none) {
    if (n == 0 || __fmaf_rn(dt, 0.5f, ck[0]) >= target) return none;
    int lo = 0, hi = n;
    while (hi - lo > 1) {
        const int mid = (lo + hi) >> 1;
        if (__fmaf_rn(dt, 0.5f, ck[mid]) >= target) hi = mid;
        else lo = mid;
    }
    return ck[lo];
}

// ---------------------------------------------------------------------------------------------
// The marching engine.
//
// The reference walks every ray with one thread: DDA step, occupancy test, and — inside the
// same loop — the skip / emit recurrences.  On a GPU that interleaving is what hurts: lanes of
// a warp reach occupied cells at different times, so the expensive emit code runs with two or
// three lanes active, and lanes whose ray is finished idle until the longest ray of the warp
// ends.  Here a warp runs three convergent phases in lock step:
//   prepare   when the warp's pool of ready rays is empty, all 32 lanes fetch one ray each
//             (one atomicAdd of 32 on a global counter), intersect it with the level boxes, find
//             its first segment and set up the DDA — and park the result in a per-warp pool in
//             shared memory.  The ~500 instructions of set-up run with 32 active lanes instead
//             of the 2-3 that happen to need a ray at that moment;
//   produce   a tight, branch-free DDA loop (<= kRoundSteps cells per round) that touches only
//             the grid and appends EVENTS (kind, t) to a per-lane queue in shared memory.  A
//             lane whose ray has ended pops a prepared ray from the pool at the next round (a
//             dozen shared-memory loads), so the DDA keeps ~all lanes busy.  With a constant dt
//             consecutive cells of the same kind fold into one event, because both recurrences
//             are idempotent in their target: a ray through a trained scene yields a handful of
//             events, not hundreds;
//   consume   when a queue fills up, all lanes replay their queued events together through
//             ONE advance() call site.  The queue also carries the ray boundaries (EV_RAY), so
//             the producer of a lane may be several rays ahead of its consumer and the replay
//             finds every lane with work.
// Results are addressed by ray id; nothing depends on which lane marched which ray.
//
// Event kinds (t = exit of the cell / start of the segment):
//   EV_EMPTY  empty cell(s):     skip to t, continuous = false        (grid.cu:189-206)
//   EV_OCC    occupied cell(s):  emit samples up to t                 (grid.cu:208-262)
//   EV_SEG    segment start:     if (!continuous) skip to t           (grid.cu:150-163)
//   EV_RAY    a new ray starts:  payload = ray id; retires the previous ray of the lane
constexpr int kQ = 16;          // events queued per lane
constexpr int kRoundSteps = 32; // DDA cells per lane between two refill checks
// Cone stepping, count pass with a piece log (grid.cu, PieceSink): the samples of a ray come in
// pieces of <= kPieceLen consecutive samples that follow the recurrence t += calc_dt(t) from the
// piece's first t; the count pass appends one 16-byte record per piece to a log (chunks of
// kPieceChunk records handed out to warps), and pass 2 replays the pieces instead of marching again.
constexpr int kPieceLen = 32;
constexpr int kPieceChunk = 256;
constexpr int kConeSteps = 4;    // cone stepping: recurrence steps per lane between two cell visits
constexpr int kConeLiteral = 16; // ... literal skip steps in one cell before the closed form takes over
constexpr int kPoolWords = 16;  // words per prepared ray (x 32 slots per warp)
enum { EV_EMPTY = 0, EV_OCC = 1, EV_SEG = 2, EV_RAY = 3 };

// DDA state of one ray segment (utils_grid.cuh:58-114).
struct Dda {
    float tdist[3], delta[3];
    int cur[3], stepi[3], stop[3];
    int cm;  // bit k: stop[k] was pulled in to the occupied box (leaving there skips empty cells)
};

// Next ray/box segment at or after event seg_i (grid.cu:128-149).
__device__ __forceinline__ bool find_segment(const RayEvents& ev, int m, int& seg_i, float near, float far,
                                             float& tmin, float& tmax, int& level) {
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
        tmax = fminf(ev.t[i + 1], far);
        if (tmin >= tmax) continue;
        return true;
    }
    return false;
}

// setup_traversal, utils_grid.cuh:58-114.
// `clip`: the walk may also stop where only empty cells can follow — once it is past the box of
// the level's occupied cells on an axis it moves away along.  Allowed when trailing empty cells
// leave no trace in the output (their skip is folded or dropped; see march_engine).
__device__ __forceinline__ void setup_dda(const float (&o)[3], const float (&d)[3], const GridSpec& G, int level,
                                          float tmin, float tmax, Dda& S, bool clip) {
    const float eps = 1e-6f;
    const float* __restrict__ bb = G.aabbs + 6 * level;
    const float ts = __fadd_rn(tmin, eps), te = __fsub_rn(tmax, eps);
    S.cm = 0;
#pragma unroll
    for (int k = 0; k < 3; ++k) {
        const float ok = o[k], dk = d[k];
        const float invk = __fdiv_rn(1.0f, dk);
        const float bmin = bb[k], bmax = bb[3 + k];
        const float resf = (float)G.res[k];
        const float ext = __fsub_rn(bmax, bmin);
        const float vox = __fdiv_rn(ext, resf);
        const float rs = __fmaf_rn(dk, ts, ok);  // FFMA site 1
        const float re = __fmaf_rn(dk, te, ok);
        S.cur[k] = clampi(__float2int_rz(__fmul_rn(__fdiv_rn(__fsub_rn(rs, bmin), ext), resf)), 0, G.res[k] - 1);
        const int fin = clampi(__float2int_rz(__fmul_rn(__fdiv_rn(__fsub_rn(re, bmin), ext), resf)), 0, G.res[k] - 1);
        const int start_index = S.cur[k] + (dk > 0.0f ? 1 : 0);
        // FFMA site 2: fma(bmin + fma(float(idx), vox, -rs), inv, tmin)
        const float tm = __fmaf_rn(__fadd_rn(bmin, __fmaf_rn((float)start_index, vox, -rs)), invk, tmin);
        const float sf = (dk == 0.0f) ? 0.0f : (dk > 0.0f ? 1.0f : -1.0f);
        S.tdist[k] = (dk == 0.0f) ? tmax : tm;
        S.stepi[k] = (int)sf;
        S.delta[k] = (dk == 0.0f) ? tmax : __fmul_rn(__fmul_rn(vox, invk), sf);
        // The walk leaves when the index reaches fin + step (utils_grid.cuh:123) or the grid
        // edge, whichever lies ahead: one stop value per axis.
        const int ovf = fin + S.stepi[k];
        const bool ahead = S.stepi[k] > 0 ? ovf > S.cur[k] : (S.stepi[k] < 0 ? ovf < S.cur[k] : true);
        S.stop[k] = ahead ? ovf : (S.stepi[k] > 0 ? G.res[k] : -1);
        if (clip && G.bounds != nullptr && S.stepi[k] != 0) {
            // first index past the occupied box in my direction, but at least one step ahead (the
            // leave test is an equality on the index just stepped to)
            if (S.stepi[k] > 0) {
                const int past = max(__ldg(G.bounds + 6 * level + 3 + k) + 1, S.cur[k] + 1);
                if (past < S.stop[k]) { S.stop[k] = past; S.cm |= 1 << k; }
            } else {
                const int past = min(__ldg(G.bounds + 6 * level + k) - 1, S.cur[k] - 1);
                if (past > S.stop[k]) { S.stop[k] = past; S.cm |= 1 << k; }
            }
        }
    }
}

// The cells a fresh segment crosses before the ray reaches the box of the occupied cells are
// empty, but the walk needs the state it would have after them — and tdist[k] += delta[k] is the
// same float recurrence as t += dt, so advance<false>() consumes, in closed form per binade, every
// crossing of axis k with a time before t_in (the ray's entry into the occupied box grown by one
// cell: the margin dwarfs the slab test's rounding).  All axes stop at the same time, so the state
// reached is one the serial walk passes through (it takes crossings in time order).  t_first
// becomes the exit time of the last cell skipped: the t of the empty span that stands for them.
// Returns true when the segment cannot meet an occupied cell at all.
__device__ __forceinline__ bool jump_ahead(const float (&o)[3], const float (&d)[3], const GridSpec& G, int level,
                                           float tmin, float tmax, Dda& S, float& t_first) {
    if (G.bounds == nullptr) return false;
    const float inf = __int_as_float(0x7f800000);
    const float* __restrict__ bb = G.aabbs + 6 * level;
    float t_in = -inf, t_out = inf;
#pragma unroll
    for (int k = 0; k < 3; ++k) {
        const float vox = (bb[3 + k] - bb[k]) / (float)G.res[k];
        const float plo = bb[k] + (float)(__ldg(G.bounds + 6 * level + k) - 1) * vox;
        const float phi = bb[k] + (float)(__ldg(G.bounds + 6 * level + 3 + k) + 2) * vox;
        if (d[k] == 0.0f) {
            if (o[k] < plo || o[k] > phi) t_out = -inf;  // never inside
            continue;
        }
        const float a = (plo - o[k]) / d[k], b = (phi - o[k]) / d[k];
        t_in = fmaxf(t_in, fminf(a, b));
        t_out = fminf(t_out, fmaxf(a, b));
    }
    if (t_in >= t_out || t_out <= tmin || t_in >= tmax) {
        // The ray does not meet the (grown) box inside this segment: every cell of the segment is
        // empty.  Let the walk leave after its first step (an early leave at a pulled-in stop,
        // which stands in for the cells not walked with one empty span).
#pragma unroll
        for (int k = 0; k < 3; ++k)
            if (S.stepi[k] != 0 && S.stop[k] != S.cur[k] + S.stepi[k]) { S.stop[k] = S.cur[k] + S.stepi[k]; S.cm |= 1 << k; }
        return true;
    }
    if (!(t_in > tmin)) return false;  // starts inside the box
    int n[3];
    float tn[3], last = -inf;
#pragma unroll
    for (int k = 0; k < 3; ++k) {
        n[k] = 0;
        tn[k] = S.tdist[k];
        if (S.stepi[k] == 0) continue;
        float lastk = -inf;
        tn[k] = advance<false>(S.tdist[k], S.delta[k], t_in, 0x7fffffff, n[k],
                               [&](float t0, float q, int kk) { lastk = __fmaf_rn((float)(kk - 1), q, t0); });
        const int room = S.stop[k] > S.cur[k] ? S.stop[k] - S.cur[k] : S.cur[k] - S.stop[k];
        if (n[k] >= room) return false;  // would step out of the segment: leave it to the walk
        last = fmaxf(last, lastk);
    }
    if (!(last > -inf)) return false;  // nothing to skip
#pragma unroll
    for (int k = 0; k < 3; ++k) {
        S.cur[k] += n[k] * S.stepi[k];
        S.tdist[k] = tn[k];
    }
    t_first = fminf(last, tmax);
    return false;
}

__device__ __forceinline__ void load_events(const float (&o)[3], const float (&d)[3], long long r,
                                            const float* __restrict__ t_sorted,
                                            const int64_t* __restrict__ t_indices,
                                            const uint8_t* __restrict__ hits, const GridSpec& G,
                                            RayEvents& ev) {
    const int m = G.n_grids;
    if (t_sorted != nullptr) {
        ev.hit_mask = 0;
        for (int l = 0; l < m; ++l)
            if (hits[r * m + l]) ev.hit_mask |= (1u << l);
        for (int i = 0; i < 2 * m; ++i) {
            ev.t[i] = __ldg(t_sorted + r * 2 * m + i);
            ev.idx[i] = (int8_t)__ldg(t_indices + r * 2 * m + i);
        }
    } else {
        float inv[3];
#pragma unroll
        for (int k = 0; k < 3; ++k) inv[k] = __fdiv_rn(1.0f, d[k]);
        compute_events(o, inv, G, ev);
    }
}

// Policy interface (see grid.cu):
//   typename Sink;  bool accept(r);  void skipped(r);  Sink begin(r);
//   void end(r, Sink&, nS, nI, t_last);  bool wants_term();
// Sink: run(P, t0, q, k, continuous, nS, nI) — k samples with edges t0 + j*q (constant-dt runs) —
// and one(P, t_last, t_next, continuous, nS, nI); `continuous`, nS, nI are the state before the call.
//
// What the walk of a ray starts from: its first segment, the DDA set up in it and moved to the occupied
// box (jump_ahead).  16 words: [0] 1 = has a segment, 2 = misses the occupied box; [1] level, segment
// index, step signs, pulled-in-stop mask; [2] end of the empty stretch before the first cell walked;
// [3] the segment's t_max; [4..15] tdist, delta, cur, stop.  Computed inside the march (prepare) or, for
// the plain single-level count pass, by march_prepare_kernel ahead of it: a thread per ray at full
// occupancy instead of 500 instructions inlined into a kernel that is short of registers.
__device__ __forceinline__ void prepare_ray(const float* __restrict__ rays_o, const float* __restrict__ rays_d,
                                            const float* __restrict__ near_arr, float near_s,
                                            const float* __restrict__ far_arr, float far_s,
                                            const float* __restrict__ t_sorted, const int64_t* __restrict__ t_indices,
                                            const uint8_t* __restrict__ hits, const GridSpec& G, long long cand,
                                            bool clip, uint32_t (&w)[16]) {
    const int m = G.n_grids;
    float o[3], d[3];
#pragma unroll
    for (int k = 0; k < 3; ++k) { o[k] = __ldg(rays_o + 3 * cand + k); d[k] = __ldg(rays_d + 3 * cand + k); }
    const float near = near_arr ? __ldg(near_arr + cand) : near_s;
    const float far = far_arr ? __ldg(far_arr + cand) : far_s;
    RayEvents pev;
    load_events(o, d, cand, t_sorted, t_indices, hits, G, pev);
    Dda S = {};
    float tmin = 0.0f, tmax = 0.0f, t_first = 0.0f;
    int level = 0, si = 0;
    bool miss = false;
    const bool found = find_segment(pev, m, si, near, far, tmin, tmax, level);
    if (found) {
        setup_dda(o, d, G, level, tmin, tmax, S, clip);
        t_first = tmin;
        if (clip) miss = jump_ahead(o, d, G, level, tmin, tmax, S, t_first);
    }
    w[0] = (found ? 1u : 0u) | (miss ? 2u : 0u);
    w[1] = (found ? 1u : 0u) | ((uint32_t)level << 1) | ((uint32_t)si << 8) | ((uint32_t)(S.stepi[0] + 1) << 16) |
           ((uint32_t)(S.stepi[1] + 1) << 18) | ((uint32_t)(S.stepi[2] + 1) << 20) | ((uint32_t)S.cm << 22);
    w[2] = __float_as_uint(t_first);  // end of the (empty) stretch before the first cell walked
    w[3] = __float_as_uint(tmax);
#pragma unroll
    for (int k = 0; k < 3; ++k) {
        w[4 + k] = __float_as_uint(S.tdist[k]);
        w[7 + k] = __float_as_uint(S.delta[k]);
        w[10 + k] = (uint32_t)S.cur[k];
        w[13 + k] = (uint32_t)S.stop[k];
    }
}

// Shared memory: s_evt = [kQ][blockDim.x] floats (event queue), s_pool = [warps][kPoolWords][32]
// words (prepared rays).
template <bool MIP, bool DIRECT, class Policy, bool PREPPED = false>
__device__ __forceinline__ void march_engine(const float* __restrict__ rays_o,
                                             const float* __restrict__ rays_d,
                                             const float* __restrict__ near_arr, float near_s,
                                             const float* __restrict__ far_arr, float far_s,
                                             const float* __restrict__ t_sorted,
                                             const int64_t* __restrict__ t_indices,
                                             const uint8_t* __restrict__ hits,
                                             const int* __restrict__ ray_list, long long n_work,
                                             const GridSpec& G, const MarchOpts& opt, Policy& P,
                                             float* __restrict__ s_evt, uint32_t* __restrict__ s_pool,
                                             const float* __restrict__ s_ck, int n_ck,
                                             unsigned long long* __restrict__ ray_counter,
                                             const uint4* __restrict__ prepped = nullptr) {
    const int m = G.n_grids;
    const float step = opt.step_size, cone = opt.cone_angle;
    const int cap = opt.limit > 0 ? opt.limit : 0x7fffffff;
    const bool const_dt = (step > 0.0f) && (cone == 0.0f);
    const bool fold_empty = const_dt || step <= 0.0f;  // consecutive empty cells fold
    // consecutive occupied cells fold whenever there is a step size: the emit loop's state is
    // t_last alone (dt is re-derived from it every step), so stopping at a cell's exit and resuming
    // in the next occupied cell visits the same t values as running on to the later exit
    const bool fold_occ = step > 0.0f;
    // trailing empty cells of a segment leave no trace when their skip folds into the next event
    // or is dropped at the end of the ray: the walk may then stop at the occupied box (setup_dda)
    const bool clip = fold_empty && !P.wants_term();
    const bool eager = opt.limit > 0;  // with a sample cap, replay every round (bounds wasted DDA work)
    const int lane = lane_id();
    const int tid = threadIdx.x, nthr = blockDim.x;
    const unsigned lt_mask = (1u << lane) - 1u;
    uint32_t* __restrict__ pool = s_pool + warp_id() * (kPoolWords * 32);

    // ---- warp-uniform pool state
    int pool_head = 0, pool_cnt = 0;
    bool exhausted = false;
    // ---- producer state (the DDA of the ray being walked)
    bool have_ray = false;
    long long r = -1;
    RayEvents ev;
    bool ev_valid = false;
    int seg_i = 0, seg_steps = 0;
    const int max_seg_steps = G.res[0] + G.res[1] + G.res[2] + 8;  // a DDA cannot take more
    bool in_seg = false;
    Dda D = {};
    float tmax_seg = 0.0f;
    int cstride[3];
    int cell = 0;  // linear index of the current cell (host side guarantees it fits 31 bits)
    int mip_base = 0;
    int pend_kind = -1;
    float pend_t = 0.0f;
    // DIRECT: the cell being marched through — its exit time (-inf: no cell; every step's "half a step
    // ahead reaches the exit" test is then true and the step changes nothing), whether it emits, the
    // skip's dt fixed at its entry, and the loop iterations spent skipping in it
    constexpr float kNoCell = -1e38f * 1e38f;  // -inf
    float ct = kNoCell, cdt = 0.0f;
    bool cemit = false;
    int lit = 0;
    uint32_t nword = 0u;  // DIRECT: occupancy word of the cell the DDA stands in (fetched a cell ahead of its use)
    // ---- queue
    int nq = 0;
    uint32_t kinds = 0;
    // ---- consumer state (the t_last recurrence of the ray being replayed)
    bool c_have = false, continuous = false, capped = false;
    long long c_r = -1;
    float t_last = 0.0f;
    int nS = 0, nI = 0, dummy = 0;
    typename Policy::Sink sink;
    // piece log (Policy::Sink::kPieces): the chunk this warp appends to and how much of it is used
    int log_chunk = -1, log_used = kPieceChunk;
    auto append_piece = [&](bool mine, float t0, int n, int s0, int rid) {
        if constexpr (Policy::Sink::kPieces) {
            const unsigned m = __ballot_sync(kFull, mine);
            if (m == 0u) return;
            const int cnt = __popc(m);
            if (log_used + cnt > kPieceChunk) {  // (a chunk is closed with a few slots unused rather than split)
                int c = 0;
                if (lane == 0) {
                    if (log_chunk >= 0) P.pieces.chunk_fill[log_chunk] = log_used;
                    c = (int)atomicAdd(P.pieces.cursor, 1ull);
                    if (c >= P.pieces.n_chunks) { c = -1; *P.pieces.failed = 1ull; }
                }
                log_chunk = __shfl_sync(kFull, c, 0);
                log_used = 0;
            }
            if (mine && log_chunk >= 0)
                P.pieces.log[(long long)log_chunk * kPieceChunk + log_used + __popc(m & lt_mask)] =
                    make_int4(__float_as_int(t0), n, s0, rid);
            log_used += cnt;
        }
    };

    auto push = [&](int kind, float t) {
        s_evt[nq * nthr + tid] = t;
        kinds |= (uint32_t)kind << (2 * nq);
        ++nq;
    };
    auto push_pending = [&]() {
        if (pend_kind >= 0) { push(pend_kind, pend_t); pend_kind = -1; }
    };
    // fine occupancy word of the current cell, fetched only under a set mip bit
    auto occupancy_word = [&]() -> uint32_t {
        uint32_t cbit = 1u;
        if (MIP) {
            const int ci = mip_base + ((D.cur[0] >> G.mip_shift) * G.mres[1] + (D.cur[1] >> G.mip_shift)) * G.mres[2] +
                           (D.cur[2] >> G.mip_shift);
            cbit = (G.mip[ci >> 5] >> (ci & 31)) & 1u;
        }
        // a predicated load, not a branch around one (the compiler will not speculate a load: its branch
        // with the reconvergence bookkeeping cost 14 instructions per cell against 4)
        uint32_t word = 0u;
        asm("{\n\t.reg .pred p;\n\tsetp.ne.u32 p, %1, 0;\n\t@p ld.global.nc.u32 %0, [%2];\n\t}"
            : "+r"(word) : "r"(cbit), "l"(G.occ_bits + (cell >> 5)));
        return word;
    };
    auto enter_segment = [&](int level) {
        cstride[0] = D.stepi[0] * G.res[1] * G.res[2];
        cstride[1] = D.stepi[1] * G.res[2];
        cstride[2] = D.stepi[2];
        cell = level * (int)G.cells_per_level + (D.cur[0] * G.res[1] + D.cur[1]) * G.res[2] + D.cur[2];
        mip_base = level * G.mip_cells_per_level;
        in_seg = true;
        seg_steps = 0;
        if constexpr (DIRECT) nword = occupancy_word();
    };

    // ---- the t_last recurrence: one ray boundary / one event
    auto begin_ray = [&](long long rid) {
        if (c_have) P.end(c_r, sink, nS, nI, t_last);
        c_r = rid;
        t_last = near_arr ? __ldg(near_arr + c_r) : near_s;
        continuous = false; capped = false; nS = 0; nI = 0;
        sink = P.begin(c_r);
        c_have = true;
    };
    auto apply = [&](int kind, float t) {
        if (capped) return;
        const bool emit = (kind == EV_OCC);
        const bool doit = !(kind == EV_SEG && continuous);
        // (launch_march picks DIRECT exactly when step > 0 and cone != 0: the queue kernels do not carry
        // the cone branches — the march is a 90 KB kernel with 10 % of its stalls on instruction fetch;
        // the same pruning of the DIRECT kernels changed their register allocation for the worse)
        if (step <= 0.0f) {  // one interval per occupied voxel crossing
            if (!emit) {
                if (doit) t_last = t;
            } else {
                sink.one(P, t_last, t, continuous, nS, nI);
                nI += continuous ? 1 : 2;
                nS += 1;
                continuous = true;
                t_last = t;
            }
        } else if (DIRECT ? const_dt : true) {  // constant dt (the only case the queue kernels see with a step)
            if (doit) {
                if (emit) {
                    t_last = advance(t_last, step, t, cap, nS, [&](float t0, float q, int k) {
                        sink.run(P, t0, q, k, continuous, nS, nI);
                        nI += continuous ? k : k + 1;
                        continuous = true;
                    });
                } else {
                    // a skip over more than a binade: start from the common trajectory's last
                    // checkpoint before the target (see make_checkpoints)
                    if (n_ck > 0 && t > t_last && t_last >= 0.0f &&
                        (__float_as_uint(t) >> 23) > (__float_as_uint(t_last) >> 23) + 1u) {
                        const float c = checkpoint_before(s_ck, n_ck, step, t, t_last);
                        if (c > t_last) t_last = c;
                    }
                    t_last = advance(t_last, step, t, 0x7fffffff, dummy, NoPiece());
                }
            }
        } else if (!emit) {
            // cone_angle > 0 (DIRECT), skip: dt is fixed at the cell's entry (grid.cu:157-161, 199-203).
            // Cells are about as long as dt: a few literal steps first, the closed form for
            // what remains.
            if (doit) {
                const float dt = calc_dt(t_last, cone, step);
                bool more = true;
#pragma unroll 1
                for (int i = 0; i < 12; ++i) {
                    if (__fmaf_rn(dt, 0.5f, t_last) >= t) { more = false; break; }
                    const float t1 = __fadd_rn(t_last, dt);
                    if (!(t1 > t_last)) { more = false; break; }
                    t_last = t1;
                }
                if (more) t_last = advance(t_last, dt, t, 0x7fffffff, dummy, NoPiece());
            }
        } else {  // cone_angle > 0, emit: dt follows t, a true serial recurrence
            while (nS < cap) {
                const float dt = calc_dt(t_last, cone, step);
                if (__fmaf_rn(dt, 0.5f, t_last) >= t) break;
                const float t_next = __fadd_rn(t_last, dt);
                if (!(t_next > t_last)) break;  // stuck
                sink.one(P, t_last, t_next, continuous, nS, nI);
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
    };

    // ------------------------------------------------------------------ prepare (all 32 lanes)
    auto prepare = [&]() {
        unsigned long long base = 0;
        const int batch = opt.batch;
        if (lane == 0) base = atomicAdd(ray_counter, (unsigned long long)batch);
        base = __shfl_sync(kFull, base, 0);
        if ((long long)base + batch >= n_work) exhausted = true;
        bool ok = lane < batch && (long long)base + lane < n_work;
        // work item -> ray id (identity unless a work list is given)
        const long long cand = !ok ? 0 : (ray_list ? (long long)__ldg(ray_list + base + lane) : (long long)base + lane);
        if (ok && !P.accept(cand)) { P.skipped(cand); ok = false; }
        uint32_t w[16];
        if (ok) {
            if constexpr (PREPPED) {
                const uint4 a0 = __ldg(prepped + 4 * cand), a1 = __ldg(prepped + 4 * cand + 1),
                            a2 = __ldg(prepped + 4 * cand + 2), a3 = __ldg(prepped + 4 * cand + 3);
                w[0] = a0.x; w[1] = a0.y; w[2] = a0.z; w[3] = a0.w; w[4] = a1.x; w[5] = a1.y; w[6] = a1.z; w[7] = a1.w;
                w[8] = a2.x; w[9] = a2.y; w[10] = a2.z; w[11] = a2.w; w[12] = a3.x; w[13] = a3.y; w[14] = a3.z; w[15] = a3.w;
            } else {
                prepare_ray(rays_o, rays_d, near_arr, near_s, far_arr, far_s, t_sorted, t_indices, hits, G, cand, clip, w);
            }
            const bool found = w[0] & 1u, miss = w[0] & 2u;
            // A ray that cannot yield anything (no segment, or — one level — one that misses the
            // occupied box) is retired here: no marker in the event queue, no cell walked.
            if (!found || (miss && m == 1)) {
                typename Policy::Sink s0 = P.begin(cand);
                P.end(cand, s0, 0, 0, near_arr ? __ldg(near_arr + cand) : near_s);
                ok = false;
            }
        }
        const unsigned okm = __ballot_sync(kFull, ok);
        if (ok) {
            uint32_t* __restrict__ e = pool + __popc(okm & lt_mask);
            e[0 * 32] = (uint32_t)(int)cand;
#pragma unroll
            for (int k = 1; k < 16; ++k) e[k * 32] = w[k];
        }
        pool_head = 0;
        pool_cnt = __popc(okm);
        __syncwarp();
    };

    // ------------------------------------------------------------------ pop (warp-collective)
    auto refill = [&]() {
        const bool need = !have_ray && nq <= kQ - 4;  // room for EV_RAY and the first events
        const unsigned mask = __ballot_sync(kFull, need);
        if (mask == 0u) return;
        if (pool_head == pool_cnt) {
            if (exhausted) return;
            prepare();
        }
        const int k = __popc(mask & lt_mask);
        const int avail = pool_cnt - pool_head;
        if constexpr (Policy::Sink::kPieces)  // the open piece of the ray that ends here
            append_piece(need && k < avail && c_have && nS > sink.p_s0, sink.p_t0, nS - sink.p_s0, sink.p_s0, (int)c_r);
        if (need && k < avail) {
            const uint32_t* __restrict__ e = pool + pool_head + k;
            r = (long long)(int)e[0];
            const uint32_t misc = e[32];
            if constexpr (DIRECT) begin_ray(r);
            else push(EV_RAY, __int_as_float((int)r));
            ev_valid = false;
            seg_i = (int)((misc >> 8) & 0xffu);
            pend_kind = -1;
            in_seg = false;
            if (misc & 1u) {
                const float tmin = __uint_as_float(e[2 * 32]);
                tmax_seg = __uint_as_float(e[3 * 32]);
#pragma unroll
                for (int a = 0; a < 3; ++a) {
                    D.tdist[a] = __uint_as_float(e[(4 + a) * 32]);
                    D.delta[a] = __uint_as_float(e[(7 + a) * 32]);
                    D.cur[a] = (int)e[(10 + a) * 32];
                    D.stop[a] = (int)e[(13 + a) * 32];
                    D.stepi[a] = (int)((misc >> (16 + 2 * a)) & 3u) - 1;
                }
                D.cm = (int)((misc >> 22) & 7u);
                // right after the ray starts `continuous` is false: the segment start is a plain skip
                have_ray = true;
                if constexpr (DIRECT) {
                    apply(EV_EMPTY, tmin);
                } else {
                    pend_kind = EV_EMPTY;
                    pend_t = tmin;
                }
                enter_segment((int)((misc >> 1) & 0x7fu));
            }
            // a ray without any segment is just its EV_RAY marker (zero samples)
        }
        const int want = __popc(mask);
        pool_head += want < avail ? want : avail;
        __syncwarp();
    };

    // Replay of the queued events.  Each lane keeps its own cursor: at every step the warp runs
    // the ONE code path (ray boundary / skip / emit) that most lanes are waiting for, and only
    // those lanes move on.  Lanes whose rays are at different phases thereby fall into step
    // (skip, emit, skip, emit, ...) instead of splitting every queue position three ways.
    auto consume = [&]() {
        int jq = 0;
#pragma unroll 1
        while (true) {
            int kind = 0, path = 0;  // path: 0 nothing left, 1 ray boundary, 2 skip, 3 emit, 4 no-op
            float t = 0.0f;
            if (jq < nq) {
                kind = (int)((kinds >> (2 * jq)) & 3u);
                t = s_evt[jq * nthr + tid];
                if (kind == EV_RAY) path = 1;
                else if (capped || (kind == EV_SEG && continuous)) path = 4;
                else path = (kind == EV_OCC) ? 3 : 2;
            }
            if (path == 4) ++jq;  // nothing to do for this event: it costs no turn
            const unsigned m_ray = __ballot_sync(kFull, path == 1);
            const unsigned m_skip = __ballot_sync(kFull, path == 2);
            const unsigned m_emit = __ballot_sync(kFull, path == 3);
            if ((m_ray | m_skip | m_emit) == 0u) {
                if (__ballot_sync(kFull, path == 4) == 0u) break;
                continue;
            }
            const int c_ray = __popc(m_ray), c_skip = __popc(m_skip), c_emit = __popc(m_emit);
            const int chosen = (c_skip >= c_emit && c_skip >= c_ray) ? 2 : (c_emit >= c_ray ? 3 : 1);
            if (path == chosen) {
                if (chosen == 1) begin_ray((long long)__float_as_int(t));
                else apply(kind, t);
                ++jq;
            }
        }
        nq = 0;
        kinds = 0;
    };
    // DIRECT (cone stepping: no event folds, dt follows t): events are applied where they arise,
    // as the reference does, instead of going through the queue
    auto event = [&](int kind, float t) {
        if constexpr (DIRECT) apply(kind, t);
        else push(kind, t);
    };

    while (true) {
        // ------------------------------------------------------------------ switch / retire
        // One round = (a) the lanes whose segment ended switch to their next segment or retire
        // the ray, then (b) all lanes walk up to kRoundSteps cells together.  Keeping (a) out of
        // the walk matters for nested grids: a lane that switched inside the walk would run the
        // rest of its round alone.
        if (have_ray && !in_seg && ct == kNoCell && nq <= kQ - 3) {
            // Left at the occupied box rather than at the end of the segment (the axis stepped last
            // sits on a pulled-in stop): the cells not walked are empty — stand in for them with
            // an empty span, whose t is replaced by the next segment's start or dropped below.
            if (D.cm != 0) {
                const bool early = ((D.cm & 1) && D.cur[0] == D.stop[0]) || ((D.cm & 2) && D.cur[1] == D.stop[1]) ||
                                   ((D.cm & 4) && D.cur[2] == D.stop[2]);
                if (early && pend_kind != EV_EMPTY) { push_pending(); pend_kind = EV_EMPTY; }
                D.cm = 0;
            }
            // a later segment of a multi-level march (the first one came prepared)
            bool found = false;
            float tmin = 0.0f;
            int level = 0;
            float o[3], d[3];
            if (m > 1) {
#pragma unroll
                for (int k = 0; k < 3; ++k) { o[k] = __ldg(rays_o + 3 * r + k); d[k] = __ldg(rays_d + 3 * r + k); }
                if (!ev_valid) { load_events(o, d, r, t_sorted, t_indices, hits, G, ev); ev_valid = true; }
                const float near = near_arr ? __ldg(near_arr + r) : near_s;
                const float far = far_arr ? __ldg(far_arr + r) : far_s;
                found = find_segment(ev, m, seg_i, near, far, tmin, tmax_seg, level);
            }
            if (!found) {
                // the ray is over: a trailing skip only moves t_last, which nobody reads unless the
                // termination planes are wanted
                if (pend_kind == EV_EMPTY && !P.wants_term()) pend_kind = -1;
                push_pending();
                have_ray = false;
            } else {
                // "if (!continuous) skip to tmin": after empty cells `continuous` is known to be
                // false and the event is a plain skip
                if (pend_kind == EV_EMPTY && fold_empty) {
                    pend_t = tmin;
                } else {
                    push_pending();
                    event(EV_SEG, tmin);
                }
                setup_dda(o, d, G, level, tmin, tmax_seg, D, clip);
                enter_segment(level);
            }
        }
        // ------------------------------------------------------------------ refill
        refill();
        if (!__any_sync(kFull, have_ray)) {
            // nothing to walk: queues are full (the replay frees them), the pool came back empty
            // (masked rays; try again), or the work is done
            if constexpr (!DIRECT) consume();
            if (exhausted && pool_head == pool_cnt) break;
            continue;
        }
        // ------------------------------------------------------------------ produce
        int steps = 0;
        if constexpr (DIRECT) {
            // Cone stepping: dt follows t, nothing folds, every cell costs a few steps of a serial
            // recurrence — a different number per lane.  Walking cell by cell in lock step (visit,
            // then each lane's emit / skip loop to the end) left 9 of 32 lanes busy; one step per
            // iteration left the visit running for the 6 lanes that had just finished a cell.
            // Every iteration of ONE loop visits a cell (the lanes that need one) and then takes up
            // to kConeSteps steps of its recurrence: most cells are over within one iteration, so
            // most lanes visit in most iterations.  The occupancy word of the cell the DDA stands
            // in is fetched when the DDA gets there, one visit ahead of its use.  Same arithmetic
            // in the same order per ray as apply() (grid.cu:189-262).
            // (the loop is warp-uniform: lanes without a cell or a ray run it predicated off — the steps
            // change nothing when ct is -inf — so the piece log can use full-mask collectives inside)
            if (__any_sync(kFull, have_ray && (in_seg || ct != kNoCell))) {
                do {
                    if (ct == kNoCell && in_seg && have_ray) {
                        // visit the cell the DDA stands in, then step past it (grid.cu:183-281)
                        ct = fminf(fminf(D.tdist[0], fminf(D.tdist[1], D.tdist[2])), tmax_seg);
                        cemit = (nword >> (cell & 31)) & 1u;
                        // an empty cell ends a run of samples: nothing reads `continuous` before the cell is over
                        continuous = continuous && cemit;
                        lit = 0;
                        cdt = calc_dt(t_last, cone, step);  // the skip's dt, fixed at the cell's entry (grid.cu:199-203)
                        const bool px = D.tdist[0] < D.tdist[1] && D.tdist[0] < D.tdist[2];
                        const bool py = !px && D.tdist[1] < D.tdist[2];
                        const bool pz = !px && !py;
                        D.cur[0] += px ? D.stepi[0] : 0;
                        D.cur[1] += py ? D.stepi[1] : 0;
                        D.cur[2] += pz ? D.stepi[2] : 0;
                        D.tdist[0] = px ? __fadd_rn(D.tdist[0], D.delta[0]) : D.tdist[0];
                        D.tdist[1] = py ? __fadd_rn(D.tdist[1], D.delta[1]) : D.tdist[1];
                        D.tdist[2] = pz ? __fadd_rn(D.tdist[2], D.delta[2]) : D.tdist[2];
                        cell += px ? cstride[0] : (py ? cstride[1] : cstride[2]);
                        const bool leave = (px && D.cur[0] == D.stop[0]) || (py && D.cur[1] == D.stop[1]) ||
                                           (pz && D.cur[2] == D.stop[2]);
                        if (leave || ++seg_steps > max_seg_steps) in_seg = false;
                        else nword = occupancy_word();
                    }
                    [[maybe_unused]] const int nS0 = nS;
                    [[maybe_unused]] const float t_i0 = t_last;
                    [[maybe_unused]] const bool cont0 = continuous;
#pragma unroll
                    for (int j = 0; j < kConeSteps; ++j) {
                        // one step of the cell's recurrence, branch-free: lanes at an emit, at a skip
                        // and between cells all run the same instructions.  (The reference also leaves
                        // an emitting cell right after a sample that reaches its exit; here the next
                        // step's test sees that, with the same result.)
                        const float dt = cemit ? calc_dt(t_last, cone, step) : cdt;
                        const float t_next = __fadd_rn(t_last, dt);
                        const bool stop = __fmaf_rn(dt, 0.5f, t_last) >= ct || !(t_next > t_last) || nS >= cap;
                        const bool em = !stop && cemit;
                        if (em) sink.one(P, t_last, t_next, continuous, nS, nI);
                        nI += em ? (continuous ? 1 : 2) : 0;
                        nS += em ? 1 : 0;
                        continuous = continuous || em;
                        t_last = stop ? t_last : t_next;
                        ct = stop ? kNoCell : ct;
                    }
                    if constexpr (Policy::Sink::kPieces) {
                        // Samples emitted in this iteration start with its first step (a cell that does not emit
                        // at once is over).  They open a new piece after a gap, or when the open piece could
                        // otherwise outgrow kPieceLen; the piece before it is complete and goes to the log.
                        const bool newp = nS > nS0 && (!cont0 || nS0 - sink.p_s0 > kPieceLen - kConeSteps);
                        append_piece(newp && nS0 > sink.p_s0, sink.p_t0, nS0 - sink.p_s0, sink.p_s0, (int)c_r);
                        if (newp) { sink.p_t0 = t_i0; sink.p_s0 = nS0; }
                    }
                    if (nS >= cap) { capped = true; have_ray = false; }
                    if (ct != kNoCell && !cemit && ++lit >= kConeLiteral / kConeSteps) {
                        // a long skip: the closed form for what remains
                        t_last = advance(t_last, cdt, ct, 0x7fffffff, dummy, NoPiece());
                        ct = kNoCell;
                    }
                    ++steps;
                } while (__any_sync(kFull, (in_seg || ct != kNoCell) && have_ray) && steps < kRoundSteps);
                if (!have_ray) ct = kNoCell;  // capped: the ray is over
            }
        } else if (have_ray && in_seg && nq <= kQ - 3) {
            // the DDA proper (grid.cu:183-281), branch-free; leaves the loop when the segment
            // ends, the round is over or the queue is (nearly) full
            do {
                const float t_trav = fminf(fminf(D.tdist[0], fminf(D.tdist[1], D.tdist[2])), tmax_seg);
                // fine bit, fetched only under a set mip bit.  (Fetching it one cell ahead, as the cone loop
                // does, was measured here: 0.788 ms against 0.754 ms — one more live register in this loop.)
                const int kind = (int)((occupancy_word() >> (cell & 31)) & 1u);  // EV_OCC == 1, EV_EMPTY == 0
                if constexpr (DIRECT) {
                    apply(kind, t_trav);
                } else {
                    if (kind != pend_kind || !(kind == EV_OCC ? fold_occ : fold_empty)) { push_pending(); pend_kind = kind; }
                    pend_t = t_trav;
                }
                // single_traversal, utils_grid.cuh:116-142: x if strictly smallest, else y if < z, else z
                const bool px = D.tdist[0] < D.tdist[1] && D.tdist[0] < D.tdist[2];
                const bool py = !px && D.tdist[1] < D.tdist[2];
                const bool pz = !px && !py;
                D.cur[0] += px ? D.stepi[0] : 0;
                D.cur[1] += py ? D.stepi[1] : 0;
                D.cur[2] += pz ? D.stepi[2] : 0;
                D.tdist[0] = px ? __fadd_rn(D.tdist[0], D.delta[0]) : D.tdist[0];
                D.tdist[1] = py ? __fadd_rn(D.tdist[1], D.delta[1]) : D.tdist[1];
                D.tdist[2] = pz ? __fadd_rn(D.tdist[2], D.delta[2]) : D.tdist[2];
                cell += px ? cstride[0] : (py ? cstride[1] : cstride[2]);
                const bool leave = (px && D.cur[0] == D.stop[0]) || (py && D.cur[1] == D.stop[1]) ||
                                   (pz && D.cur[2] == D.stop[2]);
                if (leave || ++seg_steps > max_seg_steps) in_seg = false;  // (the bound only trips on NaN rays)
                ++steps;
            } while (in_seg && steps < kRoundSteps && nq <= kQ - 3 && (!DIRECT || have_ray));
        }
        // ------------------------------------------------------------------ consume
        if constexpr (!DIRECT) {
            if (eager || __any_sync(kFull, nq > kQ - 4)) consume();
        }
    }
    if constexpr (Policy::Sink::kPieces) {
        append_piece(c_have && nS > sink.p_s0, sink.p_t0, nS - sink.p_s0, sink.p_s0, (int)c_r);
        if (lane == 0 && log_chunk >= 0) P.pieces.chunk_fill[log_chunk] = log_used;
    }
    if (c_have) P.end(c_r, sink, nS, nI, t_last);
}

}  // namespace nfb
