// grid.cuh — per-ray multi-level DDA march shared by the count and fill kernels.
//
// Semantics follow the reference kernel line by line (csrc/grid.cu:98-281 with
// include/utils_grid.cuh:10-142, data_spec_packed.cuh:43-72, utils_contraction.cuh:19-24):
// sample counts are decided by float compares, so every operation is spelled with a
// round-to-nearest intrinsic.  The three places where nvcc contracts the reference's
// expressions into FFMA are explicit __fmaf_rn here; nothing else may
// fuse (the library is also built with -fmad=false).
//
// Differences from the reference, all outside its defined behaviour:
//  * occupancy is read from a bit-packed copy of `binaries` (1 bit/cell, L1/L2 resident);
//  * the DDA stops if an index leaves the grid (the reference would read out of bounds);
//  * a march that can no longer advance (t + dt == t) stops (the reference spins forever).
#pragma once

#include "common.cuh"

namespace nfb {

constexpr int kMaxLevels = 8;  // MAX_GRID_LEVELS in csrc/grid.cu:18

struct GridSpec {
    const uint32_t* __restrict__ occ_bits;  // [n_grids * cells_per_level] bits, C order [m,x,y,z]
    const float* __restrict__ aabbs;        // [n_grids, 6]
    int n_grids;
    int res[3];
    long long cells_per_level;
};

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

// The march of one ray.  `emit(t_last, t_next, continuous)` is called for every sample, in
// order; the caller's functor counts or stores.  Returns through n_samples / n_intervals /
// t_last exactly what the reference kernel leaves in chunk_cnts and terminate_planes.
template <class Emit>
__device__ __forceinline__ void march_ray(const float (&o)[3], const float (&d)[3], float near,
                                          float far, const GridSpec& G, const RayEvents& ev,
                                          const MarchOpts& opt, Emit& emit, int& n_samples,
                                          int& n_intervals, float& t_term) {
    const float eps = 1e-6f;
    const int m = G.n_grids;
    const float step = opt.step_size, cone = opt.cone_angle;
    const int limit = opt.limit;
    float inv[3];
#pragma unroll
    for (int k = 0; k < 3; ++k) inv[k] = __fdiv_rn(1.0f, d[k]);
    int nS = 0, nI = 0;
    float t_last = near;
    bool continuous = false;
    for (int i = 0; i < 2 * m - 1; ++i) {
        const bool entering = ev.idx[i] < m;
        int level = ev.idx[i] % m;
        if (!((ev.hit_mask >> level) & 1u)) continue;
        if (!entering) {
            if (ev.idx[i + 1] < m) continue;  // next event enters a box: we are in a gap
            level = ev.idx[i + 1] % m;
            if (!((ev.hit_mask >> level) & 1u)) continue;
        }
        const float tmin = fmaxf(ev.t[i], near);
        const float tmax = fminf(ev.t[i + 1], far);
        if (tmin >= tmax) continue;
        if (!continuous) {
            if (step <= 0.0f) t_last = tmin;
            else {
                const float dt = calc_dt(t_last, cone, step);  // once, even when cone > 0
                while (!(__fmaf_rn(dt, 0.5f, t_last) >= tmin)) {
                    const float nt = __fadd_rn(t_last, dt);
                    if (nt == t_last) break;
                    t_last = nt;
                }
            }
        }
        // setup_traversal, utils_grid.cuh:58-114
        const float* __restrict__ bb = G.aabbs + 6 * level;
        float tdist[3], delta[3];
        int stepi[3], cur[3], ovf[3];
        const float ts = __fadd_rn(tmin, eps), te = __fsub_rn(tmax, eps);
#pragma unroll
        for (int k = 0; k < 3; ++k) {
            const float bmin = bb[k], bmax = bb[3 + k];
            const float resf = (float)G.res[k];
            const float ext = __fsub_rn(bmax, bmin);
            const float vox = __fdiv_rn(ext, resf);
            const float rs = __fmaf_rn(d[k], ts, o[k]);  // FFMA site 1
            const float re = __fmaf_rn(d[k], te, o[k]);
            cur[k] = clampi(__float2int_rz(__fmul_rn(__fdiv_rn(__fsub_rn(rs, bmin), ext), resf)), 0, G.res[k] - 1);
            const int fin = clampi(__float2int_rz(__fmul_rn(__fdiv_rn(__fsub_rn(re, bmin), ext), resf)), 0, G.res[k] - 1);
            const int start_index = cur[k] + (d[k] > 0.0f ? 1 : 0);
            // FFMA site 2: fma(bmin + fma(float(idx), vox, -rs), inv, tmin)
            const float tm = __fmaf_rn(__fadd_rn(bmin, __fmaf_rn((float)start_index, vox, -rs)), inv[k], tmin);
            const float sf = (d[k] == 0.0f) ? 0.0f : (d[k] > 0.0f ? 1.0f : -1.0f);
            tdist[k] = (d[k] == 0.0f) ? tmax : tm;
            stepi[k] = (int)sf;
            delta[k] = (d[k] == 0.0f) ? tmax : __fmul_rn(__fmul_rn(vox, inv[k]), sf);
            ovf[k] = fin + stepi[k];
        }
        const long long level_base = (long long)level * G.cells_per_level;
        while (limit <= 0 || nS < limit) {
            const float t_trav = fminf(fminf(tdist[0], fminf(tdist[1], tdist[2])), tmax);
            const long long cell = level_base + ((long long)cur[0] * G.res[1] + cur[1]) * G.res[2] + cur[2];
            const bool occupied = (__ldg(G.occ_bits + (cell >> 5)) >> (cell & 31)) & 1u;
            if (!occupied) {
                if (step <= 0.0f) t_last = t_trav;
                else {
                    const float dt = calc_dt(t_last, cone, step);
                    while (!(__fmaf_rn(dt, 0.5f, t_last) >= t_trav)) {
                        const float nt = __fadd_rn(t_last, dt);
                        if (nt == t_last) break;
                        t_last = nt;
                    }
                }
                continuous = false;
            } else {
                while (limit <= 0 || nS < limit) {
                    float t_next;
                    if (step <= 0.0f) t_next = t_trav;
                    else {
                        const float dt = calc_dt(t_last, cone, step);
                        if (__fmaf_rn(dt, 0.5f, t_last) >= t_trav) break;
                        t_next = __fadd_rn(t_last, dt);
                        if (!(t_next > t_last)) break;  // stuck
                    }
                    emit(t_last, t_next, continuous, nS, nI);
                    nI += continuous ? 1 : 2;
                    nS += 1;
                    continuous = true;
                    t_last = t_next;
                    if (t_next >= t_trav) break;
                }
            }
            // single_traversal, utils_grid.cuh:116-142: x if strictly smallest, else y if < z, else z
            const int ax = (tdist[0] < tdist[1] && tdist[0] < tdist[2]) ? 0 : (tdist[1] < tdist[2] ? 1 : 2);
            // (written without dynamic register indexing)
            if (ax == 0) { cur[0] += stepi[0]; tdist[0] = __fadd_rn(tdist[0], delta[0]); if (cur[0] == ovf[0] || cur[0] < 0 || cur[0] >= G.res[0]) break; }
            else if (ax == 1) { cur[1] += stepi[1]; tdist[1] = __fadd_rn(tdist[1], delta[1]); if (cur[1] == ovf[1] || cur[1] < 0 || cur[1] >= G.res[1]) break; }
            else { cur[2] += stepi[2]; tdist[2] = __fadd_rn(tdist[2], delta[2]); if (cur[2] == ovf[2] || cur[2] < 0 || cur[2] >= G.res[2]) break; }
        }
    }
    n_samples = nS;
    n_intervals = nI;
    t_term = t_last;
}

}  // namespace nfb
