// accumulate.cu — accumulate_along_rays as a segmented reduction.
//
// The reference materialises weights[:,None]*values ([S,D] temporary) and calls
// index_add_, i.e. one float atomicAdd per (sample, channel), contended because
// consecutive samples hit the same output row (nerfacc/volrend.py:532-547).  Here:
//   * accumulate_flat_kernel<D>  (canonical packed_info, D <= 4): the w*v product is formed
//     in registers, reduced with the flat segmented-scan tile, and every ray's total is
//     written exactly once, without atomics, by the tile in which the ray ends (empty rays
//     get their zeros from the same tile) — deterministic, no memset, no temporary;
//   * accumulate_runs_kernel<D>  (only ray_indices known, any order, D <= 4): runs of equal
//     ray index are reduced inside the tile and each run issues one atomicAdd per channel;
//   * accumulate_atomic_kernel   (any D): plain per-element atomics, for wide feature dims;
//   * accumulate_bwd_kernel      : g_w = <g_out[ray], v>, g_v = w * g_out[ray].
#include "seg_tile.cuh"

namespace nfb {

// Load the D values of this thread's 4 items of vector k: 4*D consecutive floats.
template <int D, bool REV>
__device__ __forceinline__ void load_values(const TileGeom<REV>& g, const float* __restrict__ values,
                                            bool vec_ok, int k, float (&x)[4][D]) {
    const long long gi = g.base + g.voff(k);
    if (vec_ok && gi + 3 < g.n) {
        float flat[4 * D];
        const float* p = values + gi * D;
#pragma unroll
        for (int q = 0; q < D; ++q) {
            const float4 t = ldg_stream4(p + 4 * q);
            flat[4 * q] = t.x; flat[4 * q + 1] = t.y; flat[4 * q + 2] = t.z; flat[4 * q + 3] = t.w;
        }
#pragma unroll
        for (int c = 0; c < 4; ++c)
#pragma unroll
            for (int d = 0; d < D; ++d) x[REV ? 3 - c : c][d] = flat[c * D + d];
    } else {
#pragma unroll
        for (int c = 0; c < 4; ++c) {
            const long long i = gi + (REV ? 3 - c : c);
#pragma unroll
            for (int d = 0; d < D; ++d) x[c][d] = (i < g.n) ? ldg_stream(values + i * D + d) : 0.0f;
        }
    }
}

template <int D>
__global__ void __launch_bounds__(kThreads)
accumulate_flat_kernel(const float* __restrict__ weights, const float* __restrict__ values,
                       const int64_t* __restrict__ packed_info, long long n_rays, long long n,
                       long long n_tiles, float* __restrict__ out, bool add_into,
                       unsigned long long* status, uint32_t epoch, bool vec_ok) {
    __shared__ TileSmem sm;
    extern __shared__ float s_tail[];  // [D][kTile]: inclusive sum at each segment tail
    const TileGeom<false> g(n, n_tiles);
    float p[D][kVec][4];
#pragma unroll
    for (int k = 0; k < kVec; ++k) {
        float w[4];
        g.load(weights, vec_ok, k, 0.0f, w);
        if (values != nullptr) {
            float x[4][D];
            load_values<D, false>(g, values, vec_ok, k, x);
#pragma unroll
            for (int c = 0; c < 4; ++c)
#pragma unroll
                for (int d = 0; d < D; ++d) p[d][k][c] = __fmul_rn(w[c], x[c][d]);
        } else {
#pragma unroll
            for (int c = 0; c < 4; ++c) p[0][k][c] = w[c];
        }
    }
    const SegSpec seg{packed_info, n_rays, 0};
    scatter_heads<false, true, true>(seg, g.base, n, sm.flags, sm.range);
    uint32_t f[kVec];
#pragma unroll
    for (int k = 0; k < kVec; ++k) f[k] = vec_flags(sm.flags, k);
    const uint8_t* fb = reinterpret_cast<const uint8_t*>(sm.flags);
#pragma unroll
    for (int d = 0; d < D; ++d) {
        float E[kVec];
        tile_segscan<SumOp>(p[d], f, E, status, d, epoch, sm);
#pragma unroll
        for (int k = 0; k < kVec; ++k) {
            float I[4];
            finish_inclusive<SumOp>(p[d][k], f[k], E[k], I);
            const int j0 = g.voff(k);
#pragma unroll
            for (int c = 0; c < 4; ++c) {
                const long long gi = g.base + j0 + c;
                const bool next_head = (c < 3) ? flag_at(f[k], c + 1) : (fb[j0 + 4] != 0);
                if (gi < n && (next_head || gi == n - 1)) s_tail[d * kTile + j0 + c] = I[c];
            }
        }
    }
    __syncthreads();
    // rays that end in this tile: end in (a, b]  (tile 0 also takes end == 0)
    const long long r_lo = sm.range[2], r_hi = sm.range[3];
    for (long long r = r_lo + threadIdx.x; r < r_hi; r += kThreads) {
        const RayChunk c = load_chunk(packed_info, r);
        const long long j = c.start + c.cnt - 1 - g.base;
#pragma unroll
        for (int d = 0; d < D; ++d) {
            const float t = (c.cnt > 0) ? s_tail[d * kTile + j] : 0.0f;
            float* o = out + r * D + d;
            if (add_into) { if (c.cnt > 0) *o = __fadd_rn(*o, t); }
            else *o = t;
        }
    }
}

// Only ray_indices known (any order).  Runs of equal ray index inside the tile are reduced
// with a tile-local segmented scan; each run ends with one atomicAdd per channel.
template <int D>
__global__ void __launch_bounds__(kThreads)
accumulate_runs_kernel(const float* __restrict__ weights, const float* __restrict__ values,
                       const int64_t* __restrict__ ray_indices, long long n, long long n_rays,
                       long long n_tiles, float* __restrict__ out, bool vec_ok) {
    __shared__ TileSmem sm;
    const TileGeom<false> g(n, n_tiles);
    float p[D][kVec][4];
    long long rid[kVec][4];
    uint32_t f[kVec];
    uint32_t tail[kVec];
#pragma unroll
    for (int k = 0; k < kVec; ++k) {
        float w[4];
        g.load(weights, vec_ok, k, 0.0f, w);
        if (values != nullptr) {
            float x[4][D];
            load_values<D, false>(g, values, vec_ok, k, x);
#pragma unroll
            for (int c = 0; c < 4; ++c)
#pragma unroll
                for (int d = 0; d < D; ++d) p[d][k][c] = __fmul_rn(w[c], x[c][d]);
        } else {
#pragma unroll
            for (int c = 0; c < 4; ++c) p[0][k][c] = w[c];
        }
        const long long gi = g.base + g.voff(k);
        long long prev = (gi > g.base && gi - 1 < n) ? __ldg(ray_indices + gi - 1) : -1;
        f[k] = 0; tail[k] = 0;
        long long nxt[4];
#pragma unroll
        for (int c = 0; c < 4; ++c) rid[k][c] = (gi + c < n) ? __ldg(ray_indices + gi + c) : -1;
        const bool last_in_tile = (g.voff(k) + 4 == kTile);
        const long long after = (!last_in_tile && gi + 4 < n) ? __ldg(ray_indices + gi + 4) : -2;
#pragma unroll
        for (int c = 0; c < 4; ++c) nxt[c] = (c < 3) ? rid[k][c + 1] : after;
#pragma unroll
        for (int c = 0; c < 4; ++c) {
            if (gi + c < n) {
                if (rid[k][c] != prev) f[k] |= (1u << (8 * c));
                if (nxt[c] != rid[k][c]) tail[k] |= (1u << c);
            }
            prev = rid[k][c];
        }
    }
#pragma unroll
    for (int d = 0; d < D; ++d) {
        float E[kVec];
        tile_segscan<SumOp, false>(p[d], f, E, nullptr, d, 0u, sm);
#pragma unroll
        for (int k = 0; k < kVec; ++k) {
            float I[4];
            finish_inclusive<SumOp>(p[d][k], f[k], E[k], I);
#pragma unroll
            for (int c = 0; c < 4; ++c)
                if ((tail[k] >> c) & 1u) {
                    const long long r = rid[k][c];
                    if (r >= 0 && r < n_rays) atomicAdd(out + r * D + d, I[c]);
                }
        }
    }
}

// Any D: one thread per (sample, channel), plain atomics (the reference's own strategy).
__global__ void accumulate_atomic_kernel(const float* __restrict__ weights,
                                         const float* __restrict__ values,
                                         const int64_t* __restrict__ ray_indices, long long n,
                                         long long D, long long n_rays, float* __restrict__ out) {
    const long long total = n * D;
    for (long long t = (long long)blockIdx.x * blockDim.x + threadIdx.x; t < total;
         t += (long long)gridDim.x * blockDim.x) {
        const long long i = t / D, d = t - i * D;
        const long long r = __ldg(ray_indices + i);
        if (r < 0 || r >= n_rays) continue;
        const float w = weights[i];
        atomicAdd(out + r * D + d, values ? __fmul_rn(w, values[t]) : w);
    }
}

// g_w[i] = sum_d g_out[ray(i), d] * v[i, d];  g_v[i, d] = w[i] * g_out[ray(i), d].
__global__ void accumulate_bwd_kernel(const float* __restrict__ g_out,
                                      const float* __restrict__ weights,
                                      const float* __restrict__ values,
                                      const int64_t* __restrict__ ray_indices, long long n,
                                      long long D, float* __restrict__ g_weights,
                                      float* __restrict__ g_values) {
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n;
         i += (long long)gridDim.x * blockDim.x) {
        const long long r = __ldg(ray_indices + i);
        const float w = (g_values != nullptr) ? weights[i] : 0.0f;
        float acc = 0.0f;
        for (long long d = 0; d < D; ++d) {
            const float go = __ldg(g_out + r * D + d);
            if (values != nullptr) {
                acc = __fadd_rn(acc, __fmul_rn(go, values[i * D + d]));
                if (g_values != nullptr) g_values[i * D + d] = __fmul_rn(w, go);
            } else {
                acc = __fadd_rn(acc, go);
            }
        }
        if (g_weights != nullptr) g_weights[i] = acc;
    }
}

}  // namespace nfb

namespace nfb {
static int acc_prep_ws(void* ws, size_t ws_bytes, long long n, uint32_t& epoch, cudaStream_t s) {
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

template <int D>
static int launch_flat(const float* w, const float* v, const int64_t* pi, long long n_rays, long long n,
                       float* out, bool add_into, unsigned long long* st, uint32_t epoch, bool vec_ok,
                       cudaStream_t s) {
    const size_t dyn = sizeof(float) * D * kTile;
    cudaFuncSetAttribute(accumulate_flat_kernel<D>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)dyn);
    const long long nt = num_tiles(n);
    accumulate_flat_kernel<D><<<(unsigned)nt, kThreads, dyn, s>>>(w, v, pi, n_rays, n, nt, out, add_into, st, epoch, vec_ok);
    return NFB_OK;
}
}  // namespace nfb

extern "C" {

NFB_API int nfb_accumulate_fwd(const float* weights, const float* values, int64_t dim,
                               const int64_t* packed_info, const int64_t* ray_indices,
                               int64_t n_rays, int64_t n, float* out, int add_into,
                               void* workspace, size_t workspace_bytes, uint32_t epoch,
                               nfb_stream_t stream) {
    using namespace nfb;
    NFB_CHECK_ARG(n >= 0 && n_rays >= 0 && dim >= 1, "bad sizes");
    NFB_CHECK_ARG(values != nullptr || dim == 1, "dim must be 1 when values is NULL");
    cudaStream_t s = static_cast<cudaStream_t>(stream);
    if (n_rays == 0) return NFB_OK;
    NFB_CHECK_ARG(out != nullptr, "null out");
    if (n == 0) {
        if (!add_into) cudaMemsetAsync(out, 0, sizeof(float) * n_rays * dim, s);
        return NFB_OK;
    }
    NFB_CHECK_ARG(weights != nullptr, "null weights");
    NFB_CHECK_ARG(packed_info != nullptr || ray_indices != nullptr, "need packed_info or ray_indices");
    const bool vec_ok = aligned16(weights) && (values == nullptr || aligned16(values));
    if (packed_info != nullptr && dim <= 4) {
        NFB_CHECK_ARG(aligned16(packed_info), "packed_info must be 16-byte aligned");
        int rc = acc_prep_ws(workspace, workspace_bytes, n, epoch, s);
        if (rc != NFB_OK) return rc;
        auto* st = static_cast<unsigned long long*>(workspace);
        switch (dim) {
            case 1: launch_flat<1>(weights, values, packed_info, n_rays, n, out, add_into != 0, st, epoch, vec_ok, s); break;
            case 2: launch_flat<2>(weights, values, packed_info, n_rays, n, out, add_into != 0, st, epoch, vec_ok, s); break;
            case 3: launch_flat<3>(weights, values, packed_info, n_rays, n, out, add_into != 0, st, epoch, vec_ok, s); break;
            default: launch_flat<4>(weights, values, packed_info, n_rays, n, out, add_into != 0, st, epoch, vec_ok, s); break;
        }
        count_launch();
        NFB_CHECK_LAUNCH("accumulate_flat");
        return NFB_OK;
    }
    NFB_CHECK_ARG(ray_indices != nullptr, "this configuration needs ray_indices");
    if (!add_into) cudaMemsetAsync(out, 0, sizeof(float) * n_rays * dim, s);
    if (dim <= 4) {
        const long long nt = num_tiles(n);
        switch (dim) {
            case 1: accumulate_runs_kernel<1><<<(unsigned)nt, kThreads, 0, s>>>(weights, values, ray_indices, n, n_rays, nt, out, vec_ok); break;
            case 2: accumulate_runs_kernel<2><<<(unsigned)nt, kThreads, 0, s>>>(weights, values, ray_indices, n, n_rays, nt, out, vec_ok); break;
            case 3: accumulate_runs_kernel<3><<<(unsigned)nt, kThreads, 0, s>>>(weights, values, ray_indices, n, n_rays, nt, out, vec_ok); break;
            default: accumulate_runs_kernel<4><<<(unsigned)nt, kThreads, 0, s>>>(weights, values, ray_indices, n, n_rays, nt, out, vec_ok); break;
        }
    } else {
        long long blocks = ceil_div<long long>(n * dim, 256);
        const long long cap = (long long)sm_count() * 32;
        if (blocks > cap) blocks = cap;
        accumulate_atomic_kernel<<<(unsigned)blocks, 256, 0, s>>>(weights, values, ray_indices, n, dim, n_rays, out);
    }
    count_launch();
    NFB_CHECK_LAUNCH("accumulate");
    return NFB_OK;
}

NFB_API int nfb_accumulate_bwd(const float* g_out, const float* weights, const float* values,
                               int64_t dim, const int64_t* ray_indices, int64_t n,
                               float* g_weights, float* g_values, nfb_stream_t stream) {
    using namespace nfb;
    NFB_CHECK_ARG(n >= 0 && dim >= 1, "bad sizes");
    if (n == 0) return NFB_OK;
    NFB_CHECK_ARG(g_out && ray_indices, "null pointer");
    NFB_CHECK_ARG(g_values == nullptr || (values != nullptr && weights != nullptr), "g_values needs weights and values");
    long long blocks = ceil_div<long long>(n, 256);
    const long long cap = (long long)sm_count() * 32;
    if (blocks > cap) blocks = cap;
    accumulate_bwd_kernel<<<(unsigned)blocks, 256, 0, static_cast<cudaStream_t>(stream)>>>(
        g_out, weights, values, ray_indices, n, dim, g_weights, g_values);
    count_launch();
    NFB_CHECK_LAUNCH("accumulate_bwd");
    return NFB_OK;
}

}  // extern "C"
