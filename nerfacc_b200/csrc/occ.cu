// occ.cu — occupancy-grid EMA update, binarisation and visibility compaction.
//
// The reference does these with ~15 small torch kernels per level (gather of an int64
// coordinate table, rand_like, index_put with duplicate indices = last-writer-wins, masked
// mean with a boolean-index temporary, three boolean-mask compactions; nerfacc/estimators/
// occ_grid.py:216-220, 367-404).  Here:
//  * cell_points_kernel   cell id + U[0,1)^3 offset -> world position, no coordinate table;
//  * ema_*_kernel         occs[c] = max(occs[c]*decay, max_k occ_k): scatter-max semantics
//                         (the author's intent, occ_grid.py:399-402), deterministic under
//                         duplicate cell ids; the decay is applied exactly once per cell;
//  * occ_sum/binarize     masked mean in fp64 with a fixed reduction tree, threshold, bool
//                         grid — two launches, no host sync, no temporaries;
//  * compact_kernel       single-pass stream compaction (decoupled look-back) of
//                         (ray_indices, t_starts, t_ends) by the visibility mask.
#include <atomic>

#include "seg_tile.cuh"

namespace nfb {

// occ_grid.py:385-392: x = aabb_min + ((ijk + u) / res) * (aabb_max - aabb_min)
__global__ void cell_points_kernel(const int64_t* __restrict__ cells, const float* __restrict__ u,
                                   long long n, int rx, int ry, int rz,
                                   const float* __restrict__ aabb, float* __restrict__ x) {
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n;
         i += (long long)gridDim.x * blockDim.x) {
        const long long c = cells[i];
        const int res[3] = {rx, ry, rz};
        const long long ijk[3] = {c / ((long long)ry * rz), (c / rz) % ry, c % rz};
#pragma unroll
        for (int k = 0; k < 3; ++k) {
            const float g = __fdiv_rn(__fadd_rn((float)ijk[k], u[3 * i + k]), (float)res[k]);
            x[3 * i + k] = __fadd_rn(aabb[k], __fmul_rn(g, __fsub_rn(aabb[3 + k], aabb[k])));
        }
    }
}

__global__ void ema_gather_kernel(const float* __restrict__ occs, const int64_t* __restrict__ ids,
                                  long long n, float decay, float* __restrict__ tmp) {
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n;
         i += (long long)gridDim.x * blockDim.x)
        tmp[i] = __fmul_rn(occs[ids[i]], decay);
}
__global__ void ema_store_kernel(float* __restrict__ occs, const int64_t* __restrict__ ids,
                                 long long n, const float* __restrict__ tmp) {
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n;
         i += (long long)gridDim.x * blockDim.x)
        occs[ids[i]] = tmp[i];  // duplicates store the same value
}
__device__ __forceinline__ void atomic_max_float(float* addr, float v) {
    if (v != v) return;
    if (v >= 0.0f) atomicMax(reinterpret_cast<int*>(addr), __float_as_int(v));
    else atomicMin(reinterpret_cast<unsigned int*>(addr), __float_as_uint(v));
}
__global__ void ema_max_kernel(float* __restrict__ occs, const int64_t* __restrict__ ids,
                               const float* __restrict__ occ, long long n) {
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n;
         i += (long long)gridDim.x * blockDim.x)
        atomic_max_float(occs + ids[i], occ[i]);
}

// ---- binarise: thre = min(mean(occs[occs >= 0]), occ_thre); binaries = occs > thre ----------
constexpr int kSumBlocks = 592;  // 4 per SM on a 148-SM part; fixed so the tree is deterministic
__global__ void __launch_bounds__(256)
occ_sum_kernel(const float* __restrict__ occs, long long n, double* __restrict__ partial) {
    double s = 0.0;
    long long c = 0;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n;
         i += (long long)gridDim.x * blockDim.x) {
        const float v = occs[i];
        if (v >= 0.0f) { s += (double)v; ++c; }
    }
    __shared__ double ss[256];
    __shared__ long long sc[256];
    ss[threadIdx.x] = s; sc[threadIdx.x] = c;
    __syncthreads();
    for (int d = 128; d > 0; d >>= 1) {
        if (threadIdx.x < d) { ss[threadIdx.x] += ss[threadIdx.x + d]; sc[threadIdx.x] += sc[threadIdx.x + d]; }
        __syncthreads();
    }
    if (threadIdx.x == 0) { partial[2 * blockIdx.x] = ss[0]; partial[2 * blockIdx.x + 1] = (double)sc[0]; }
}
__global__ void __launch_bounds__(256)
occ_binarize_kernel(const float* __restrict__ occs, long long n, const double* __restrict__ partial,
                    int n_partial, float occ_thre, uint8_t* __restrict__ binaries,
                    float* __restrict__ thre_out) {
    __shared__ double ss[256], sc[256];
    double s = 0.0, c = 0.0;
    for (int i = threadIdx.x; i < n_partial; i += 256) { s += partial[2 * i]; c += partial[2 * i + 1]; }
    ss[threadIdx.x] = s; sc[threadIdx.x] = c;
    __syncthreads();
    for (int d = 128; d > 0; d >>= 1) {
        if (threadIdx.x < d) { ss[threadIdx.x] += ss[threadIdx.x + d]; sc[threadIdx.x] += sc[threadIdx.x + d]; }
        __syncthreads();
    }
    const float mean = (float)(ss[0] / sc[0]);  // 0/0 -> NaN, like torch's mean of an empty tensor
    const float thre = (mean != mean) ? mean : fminf(mean, occ_thre);
    if (blockIdx.x == 0 && threadIdx.x == 0 && thre_out) *thre_out = thre;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n;
         i += (long long)gridDim.x * blockDim.x)
        binaries[i] = occs[i] > thre ? 1 : 0;
}

// ---- compaction ---------------------------------------------------------------------------
// One tile = 256 threads x 8 items.  status word: count << 2 | state.
constexpr int kCompItems = 8;
constexpr int kCompTile = kThreads * kCompItems;
// Order-preserving output position of this thread's first kept item: an intra-tile scan of the
// per-thread counts, then the tile's carry by decoupled look-back over `status` (zeroed by the
// caller).  The last tile writes the number of kept items to `total`.  All threads call it.
__device__ __forceinline__ long long tile_lookback_position(int cnt, unsigned long long* status,
                                                            long long* __restrict__ total,
                                                            const long long* base = nullptr) {
    __shared__ int s_warp[kWarps];
    __shared__ long long s_carry;
    const long long tile = blockIdx.x;
    int t = cnt;
    const int lane = lane_id(), w = warp_id();
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
        const int y = __shfl_up_sync(kFull, t, d);
        if (lane >= d) t += y;
    }
    if (lane == 31) s_warp[w] = t;
    __syncthreads();
    int before = 0, agg = 0;
#pragma unroll
    for (int j = 0; j < kWarps; ++j) { if (j == w) before = agg; agg += s_warp[j]; }
    if (w == 0) {  // the first warp looks back, 32 predecessors per round
        long long X = 0;
        if (tile == 0) {
            if (lane == 0) st_relaxed_u64(status, ((unsigned long long)agg << 2) | 2ull);
        } else {
            if (lane == 0) st_relaxed_u64(status + tile, ((unsigned long long)agg << 2) | 1ull);
            X = warp_lookback_u64(status, tile);
            if (lane == 0) st_relaxed_u64(status + tile, ((unsigned long long)(X + agg) << 2) | 2ull);
        }
        if (lane == 0) {
            s_carry = X + (base ? *base : 0);
            if (tile == (long long)gridDim.x - 1 && total) *total = X + agg;
        }
    }
    __syncthreads();
    return s_carry + before + (t - cnt);
}

__global__ void __launch_bounds__(kThreads)
compact_kernel(const uint8_t* __restrict__ mask, const int64_t* __restrict__ ray_indices,
               const float* __restrict__ t_starts, const float* __restrict__ t_ends, long long n,
               int64_t* __restrict__ o_ray, float* __restrict__ o_ts, float* __restrict__ o_te,
               long long* __restrict__ total, unsigned long long* status) {
    const long long tile = blockIdx.x;
    const long long i0 = tile * kCompTile + (long long)threadIdx.x * kCompItems;
    uint8_t m[kCompItems];
    int cnt = 0;
    if (i0 + kCompItems <= n && ((reinterpret_cast<uintptr_t>(mask + i0) & 7u) == 0)) {
        const unsigned long long w = *reinterpret_cast<const unsigned long long*>(mask + i0);
#pragma unroll
        for (int k = 0; k < kCompItems; ++k) { m[k] = (uint8_t)((w >> (8 * k)) & 0xff); cnt += m[k] ? 1 : 0; }
    } else {
#pragma unroll
        for (int k = 0; k < kCompItems; ++k) { m[k] = (i0 + k < n) ? mask[i0 + k] : 0; cnt += m[k] ? 1 : 0; }
    }
    long long pos = tile_lookback_position(cnt, status, total);
#pragma unroll
    for (int k = 0; k < kCompItems; ++k) {
        if (m[k]) {
            const long long i = i0 + k;
            o_ray[pos] = ray_indices[i];
            o_ts[pos] = t_starts[i];
            o_te[pos] = t_ends[i];
            ++pos;
        }
    }
}

// ---- cell selection of the update (occ_grid.py:344-365) -------------------------------------
// The two data-dependent selections `_sample_uniform_and_occupied_cells` makes per level, each one
// launch, appended to one list: CANDIDATES: keep the drawn cells whose occupancy is not negative
// (`uniform[occs[uniform] >= 0]`); otherwise the indices of the occupied cells in order
// (`nonzero(binaries)`).  `base` (device, may be null): where this call's output starts in `out`
// — the first call's `total`, so the second selection lands behind the first with no host step
// between them.  `total`: kept items of THIS call.
template <bool CANDIDATES>
__global__ void __launch_bounds__(kThreads)
select_cells_kernel(const float* __restrict__ occs, const int64_t* __restrict__ candidates,
                    const uint8_t* __restrict__ binaries, long long n, const long long* __restrict__ base,
                    int64_t* __restrict__ out, long long* __restrict__ total, unsigned long long* status) {
    const long long i0 = (long long)blockIdx.x * kCompTile + (long long)threadIdx.x * kCompItems;
    int64_t v[kCompItems];
    unsigned keep = 0;
    if constexpr (CANDIDATES) {
#pragma unroll
        for (int k = 0; k < kCompItems; ++k) v[k] = (i0 + k < n) ? candidates[i0 + k] : -1;
#pragma unroll
        for (int k = 0; k < kCompItems; ++k)
            if (v[k] >= 0 && __ldg(occs + v[k]) >= 0.0f) keep |= 1u << k;
    } else {
        if (i0 + kCompItems <= n && ((reinterpret_cast<uintptr_t>(binaries + i0) & 7u) == 0)) {
            const unsigned long long w = *reinterpret_cast<const unsigned long long*>(binaries + i0);
#pragma unroll
            for (int k = 0; k < kCompItems; ++k) if ((w >> (8 * k)) & 0xff) keep |= 1u << k;
        } else {
#pragma unroll
            for (int k = 0; k < kCompItems; ++k) if (i0 + k < n && binaries[i0 + k]) keep |= 1u << k;
        }
#pragma unroll
        for (int k = 0; k < kCompItems; ++k) v[k] = i0 + k;
    }
    long long pos = tile_lookback_position(__popc(keep), status, total, base);
#pragma unroll
    for (int k = 0; k < kCompItems; ++k)
        if (keep & (1u << k)) out[pos++] = v[k];
}

// ---- mark_invisible_cells (occ_grid.py:263-332) ------------------------------------------------
// One thread per cell, all cameras in a loop (camera rows staged in shared memory, 21 floats
// each: world->camera rotation, translation, intrinsics).  The reference runs ~12 torch kernels
// and two batched matmuls per chunk of 32^3 cells; the arithmetic here is the same sequence of
// fp32 operations (x = ijk / (res - 1); world = lo + x * ext; cam = R . world + T with the dot
// product as an FMA chain, the way a K = 3 matmul accumulates; pix = K . cam; u, v by IEEE division).
constexpr int kCamChunk = 256;
__global__ void __launch_bounds__(256)
mark_invisible_kernel(const float* __restrict__ cams /*[n_cams,21]*/, int n_cams, int rx, int ry, int rz,
                      int levels, const float* __restrict__ aabbs, float width, float height, float near,
                      float* __restrict__ occs) {
    __shared__ float s_cam[kCamChunk * 21];
    const long long per_lvl = (long long)rx * ry * rz;
    const long long n = per_lvl * levels;
    // grid-stride in lockstep so that every thread of the CTA reaches the barriers
    for (long long base = (long long)blockIdx.x * blockDim.x; base < n; base += (long long)gridDim.x * blockDim.x) {
        const long long i = base + threadIdx.x;
        const bool live = i < n && occs[i] >= 0.0f;  // cells already invisible are left alone
        float w[3] = {0.0f, 0.0f, 0.0f};
        if (live) {
            const int lvl = (int)(i / per_lvl);
            const long long c = i - (long long)lvl * per_lvl;
            const int ijk[3] = {(int)(c / ((long long)ry * rz)), (int)((c / rz) % ry), (int)(c % rz)};
            const int res[3] = {rx, ry, rz};
#pragma unroll
            for (int k = 0; k < 3; ++k) {
                const float x = __fdiv_rn((float)ijk[k], (float)(res[k] - 1));
                const float lo = aabbs[6 * lvl + k], hi = aabbs[6 * lvl + 3 + k];
                w[k] = __fadd_rn(lo, __fmul_rn(x, __fsub_rn(hi, lo)));
            }
        }
        int seen = 0;
        bool too_near = false;
        for (int c0 = 0; c0 < n_cams; c0 += kCamChunk) {
            const int nc = min(kCamChunk, n_cams - c0);
            __syncthreads();
            for (int j = threadIdx.x; j < nc * 21; j += blockDim.x) s_cam[j] = __ldg(cams + (long long)c0 * 21 + j);
            __syncthreads();
            if (!live) continue;
            for (int j = 0; j < nc; ++j) {
                const float* __restrict__ m = s_cam + j * 21;
                float cam[3], pix[3];
#pragma unroll
                for (int k = 0; k < 3; ++k)
                    cam[k] = __fadd_rn(__fmaf_rn(m[3 * k + 2], w[2], __fmaf_rn(m[3 * k + 1], w[1], __fmul_rn(m[3 * k], w[0]))), m[9 + k]);
#pragma unroll
                for (int k = 0; k < 3; ++k)
                    pix[k] = __fmaf_rn(m[12 + 3 * k + 2], cam[2], __fmaf_rn(m[12 + 3 * k + 1], cam[1], __fmul_rn(m[12 + 3 * k], cam[0])));
                const float u = __fdiv_rn(pix[0], pix[2]), v = __fdiv_rn(pix[1], pix[2]);
                const bool in_image = pix[2] >= 0.0f && u >= 0.0f && u < width && v >= 0.0f && v < height;
                seen += (in_image && pix[2] >= near) ? 1 : 0;
                too_near = too_near || (in_image && pix[2] < near);
            }
        }
        if (live) occs[i] = (seen > 0 && !too_near) ? 0.0f : -1.0f;
    }
}

// ---- allocation-free compaction ------------------------------------------------------------
// nfb_compact_samples_static keeps the look-back status words of compact_kernel in this
// file-scope array instead of a caller-provided workspace, for callers that cannot allocate
// (CUDA-graph capture, serving loops).  Every GPU that loads the library holds the 256 KiB until
// exit.  `g_compact_status` is a host-side stand-in in host code, never a device address: the
// launch below passes what cudaGetSymbolAddress returned for the current device.
constexpr long long kStaticCompactMaxN = NFB_COMPACT_STATIC_MAX_SAMPLES;  // samples per call
constexpr long long kStaticCompactTiles = kStaticCompactMaxN / kCompTile;
static __device__ __align__(256) unsigned long long g_compact_status[kStaticCompactTiles];

constexpr int kMaxDevices = 64;
static std::atomic<unsigned long long*> g_compact_status_addr[kMaxDevices];

// Device address of g_compact_status on the CURRENT device; looked up once per device.
static cudaError_t compact_status_address(unsigned long long** out) {
    int dev = -1;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess) return e;
    if (dev < 0 || dev >= kMaxDevices) return cudaErrorInvalidDevice;
    unsigned long long* p = g_compact_status_addr[dev].load(std::memory_order_acquire);
    if (p == nullptr) {
        e = cudaGetSymbolAddress(reinterpret_cast<void**>(&p), g_compact_status);
        if (e != cudaSuccess) return e;
        if (p == nullptr) return cudaErrorInvalidSymbol;
        g_compact_status_addr[dev].store(p, std::memory_order_release);  // racing threads store the same value
    }
    *out = p;
    return cudaSuccess;
}

static inline unsigned blocks_for(long long n, int block = 256, int per_sm = 16) {
    long long b = ceil_div<long long>(n, block);
    const long long cap = (long long)sm_count() * per_sm;
    if (b > cap) b = cap;
    if (b < 1) b = 1;
    return (unsigned)b;
}

}  // namespace nfb

extern "C" {

NFB_API int nfb_occ_cell_points(const int64_t* cells, const float* u, int64_t n, const int32_t* res_host,
                                const float* aabb, float* x, nfb_stream_t stream) {
    using namespace nfb;
    NFB_CHECK_ARG(n >= 0, "negative size");
    if (n == 0) return NFB_OK;
    NFB_CHECK_ARG(cells && u && res_host && aabb && x, "null pointer");
    cell_points_kernel<<<blocks_for(n), 256, 0, static_cast<cudaStream_t>(stream)>>>(
        cells, u, n, res_host[0], res_host[1], res_host[2], aabb, x);
    count_launch();
    NFB_CHECK_LAUNCH("occ_cell_points");
    return NFB_OK;
}

NFB_API int nfb_occ_ema_update(float* occs, const int64_t* cell_ids, int64_t n, const float* occ,
                               int64_t eval_lo, int64_t eval_hi, float decay, float* tmp,
                               nfb_stream_t stream) {
    using namespace nfb;
    NFB_CHECK_ARG(n >= 0 && eval_lo >= 0 && eval_hi >= eval_lo && eval_hi <= n, "bad range");
    if (n == 0) return NFB_OK;
    NFB_CHECK_ARG(occs && cell_ids && tmp, "null pointer");
    cudaStream_t s = static_cast<cudaStream_t>(stream);
    ema_gather_kernel<<<blocks_for(n), 256, 0, s>>>(occs, cell_ids, n, decay, tmp);
    ema_store_kernel<<<blocks_for(n), 256, 0, s>>>(occs, cell_ids, n, tmp);
    count_launch(2);
    if (eval_hi > eval_lo) {
        NFB_CHECK_ARG(occ != nullptr, "null occ");
        ema_max_kernel<<<blocks_for(eval_hi - eval_lo), 256, 0, s>>>(occs, cell_ids + eval_lo, occ, eval_hi - eval_lo);
        count_launch();
    }
    NFB_CHECK_LAUNCH("occ_ema_update");
    return NFB_OK;
}

NFB_API size_t nfb_occ_binarize_workspace_bytes(void) { return sizeof(double) * 2 * nfb::kSumBlocks; }

NFB_API int nfb_occ_binarize(const float* occs, int64_t n_cells, float occ_thre, uint8_t* binaries,
                             float* thre_out, void* workspace, size_t workspace_bytes,
                             nfb_stream_t stream) {
    using namespace nfb;
    NFB_CHECK_ARG(n_cells >= 0, "negative size");
    if (n_cells == 0) return NFB_OK;
    NFB_CHECK_ARG(occs && binaries && workspace, "null pointer");
    NFB_CHECK_ARG(workspace_bytes >= nfb_occ_binarize_workspace_bytes(), "workspace too small");
    cudaStream_t s = static_cast<cudaStream_t>(stream);
    double* partial = static_cast<double*>(workspace);
    occ_sum_kernel<<<kSumBlocks, 256, 0, s>>>(occs, n_cells, partial);
    occ_binarize_kernel<<<blocks_for(n_cells), 256, 0, s>>>(occs, n_cells, partial, kSumBlocks, occ_thre, binaries, thre_out);
    count_launch(2);
    NFB_CHECK_LAUNCH("occ_binarize");
    return NFB_OK;
}

NFB_API size_t nfb_compact_workspace_bytes(int64_t n) {
    return (size_t)nfb::ceil_div<long long>(n > 0 ? n : 1, nfb::kCompTile) * sizeof(unsigned long long);
}

NFB_API int nfb_compact_samples(const uint8_t* mask, const int64_t* ray_indices, const float* t_starts,
                                const float* t_ends, int64_t n, int64_t* out_ray_indices,
                                float* out_t_starts, float* out_t_ends, int64_t* total,
                                void* workspace, size_t workspace_bytes, nfb_stream_t stream) {
    using namespace nfb;
    NFB_CHECK_ARG(n >= 0, "negative size");
    cudaStream_t s = static_cast<cudaStream_t>(stream);
    if (n == 0) { if (total) cudaMemsetAsync(total, 0, sizeof(int64_t), s); return NFB_OK; }
    NFB_CHECK_ARG(mask && ray_indices && t_starts && t_ends && out_ray_indices && out_t_starts && out_t_ends, "null pointer");
    const size_t need = nfb_compact_workspace_bytes(n);
    NFB_CHECK_ARG(workspace && workspace_bytes >= need, "workspace too small");
    cudaMemsetAsync(workspace, 0, need, s);
    const long long nt = ceil_div<long long>(n, kCompTile);
    compact_kernel<<<(unsigned)nt, kThreads, 0, s>>>(mask, ray_indices, t_starts, t_ends, n, out_ray_indices,
                                                    out_t_starts, out_t_ends, reinterpret_cast<long long*>(total),
                                                    static_cast<unsigned long long*>(workspace));
    count_launch();
    NFB_CHECK_LAUNCH("compact_samples");
    return NFB_OK;
}

NFB_API int nfb_occ_select_cells(const float* occs, const int64_t* candidates, const uint8_t* binaries,
                                 int64_t n, const int64_t* base, int64_t* out, int64_t* total,
                                 void* workspace, size_t workspace_bytes, nfb_stream_t stream) {
    using namespace nfb;
    NFB_CHECK_ARG(n >= 0, "negative size");
    NFB_CHECK_ARG(total, "null pointer");
    NFB_CHECK_ARG((candidates != nullptr) != (binaries != nullptr), "give either candidates (+ occs) or binaries");
    NFB_CHECK_ARG(!candidates || occs, "candidates need occs");
    cudaStream_t s = static_cast<cudaStream_t>(stream);
    if (n == 0) { cudaMemsetAsync(total, 0, sizeof(int64_t), s); return NFB_OK; }
    NFB_CHECK_ARG(out, "null pointer");
    const size_t need = nfb_compact_workspace_bytes(n);
    NFB_CHECK_ARG(workspace && workspace_bytes >= need, "workspace too small");
    cudaMemsetAsync(workspace, 0, need, s);
    const unsigned nt = (unsigned)ceil_div<long long>(n, kCompTile);
    auto* st = static_cast<unsigned long long*>(workspace);
    auto* b = reinterpret_cast<const long long*>(base);
    auto* t = reinterpret_cast<long long*>(total);
    if (candidates) select_cells_kernel<true><<<nt, kThreads, 0, s>>>(occs, candidates, nullptr, n, b, out, t, st);
    else select_cells_kernel<false><<<nt, kThreads, 0, s>>>(nullptr, nullptr, binaries, n, b, out, t, st);
    count_launch();
    NFB_CHECK_LAUNCH("occ_select_cells");
    return NFB_OK;
}

NFB_API int nfb_compact_samples_static(const uint8_t* mask, const int64_t* ray_indices,
                                       const float* t_starts, const float* t_ends, int64_t n,
                                       int64_t* out_ray_indices, float* out_t_starts,
                                       float* out_t_ends, int64_t* total, nfb_stream_t stream) {
    using namespace nfb;
    NFB_CHECK_ARG(n >= 0, "negative size");
    NFB_CHECK_ARG(n <= kStaticCompactMaxN,
                  "compact_samples_static: n = %lld exceeds the built-in scratch (at most %lld samples per call); "
                  "use nfb_compact_samples with a caller-provided workspace",
                  (long long)n, (long long)kStaticCompactMaxN);
    cudaStream_t s = static_cast<cudaStream_t>(stream);
    if (n == 0) { if (total) cudaMemsetAsync(total, 0, sizeof(int64_t), s); return NFB_OK; }
    NFB_CHECK_ARG(mask && ray_indices && t_starts && t_ends && out_ray_indices && out_t_starts && out_t_ends, "null pointer");
    unsigned long long* status = nullptr;
    const cudaError_t e = compact_status_address(&status);
    if (e != cudaSuccess) {
        set_error("compact_samples_static: cudaGetSymbolAddress: %s", cudaGetErrorString(e));
        return NFB_ERR_CUDA;
    }
    const long long nt = ceil_div<long long>(n, kCompTile);
    // the array keeps the previous call's states: the tiles this call uses must start at zero
    cudaMemsetAsync(status, 0, (size_t)nt * sizeof(unsigned long long), s);
    compact_kernel<<<(unsigned)nt, kThreads, 0, s>>>(mask, ray_indices, t_starts, t_ends, n, out_ray_indices,
                                                    out_t_starts, out_t_ends, reinterpret_cast<long long*>(total),
                                                    status);
    count_launch();
    NFB_CHECK_LAUNCH("compact_samples_static");
    return NFB_OK;
}

NFB_API int nfb_mark_invisible_cells(const float* cams, int64_t n_cams, const int32_t* res_host, int32_t levels,
                                     const float* aabbs, float width, float height, float near_plane,
                                     float* occs, nfb_stream_t stream) {
    using namespace nfb;
    NFB_CHECK_ARG(n_cams >= 0 && levels >= 0, "negative size");
    NFB_CHECK_ARG(res_host != nullptr, "null resolution");
    const long long n = (long long)res_host[0] * res_host[1] * res_host[2] * levels;
    if (n == 0) return NFB_OK;
    NFB_CHECK_ARG(n_cams < (1ll << 31), "too many cameras");
    NFB_CHECK_ARG((cams || n_cams == 0) && aabbs && occs, "null pointer");
    mark_invisible_kernel<<<blocks_for(n, 256, 8), 256, 0, static_cast<cudaStream_t>(stream)>>>(
        cams, (int)n_cams, res_host[0], res_host[1], res_host[2], levels, aabbs, width, height, near_plane, occs);
    count_launch();
    NFB_CHECK_LAUNCH("mark_invisible_cells");
    return NFB_OK;
}

}  // extern "C"
