// pack.cu — pack_info (ray_indices -> packed_info) and the per-ray exclusive scan that
// turns per-ray counts into packed_info starts (shared with the traversal).
//
// Reference: nerfacc/pack.py:43-46 = index_add_ of ones (one int64 atomic per sample) +
// cumsum + stack (4 kernels, 3 temporaries).  Here: one histogram pass that issues one
// atomic per *run* of equal indices, and one single-pass look-back scan over the rays that
// writes (start, count) interleaved in place.
#include "seg_tile.cuh"

namespace nfb {

// counts[r] += length of every run of r in ray_indices; counts live at packed_info[2r+1].
__global__ void __launch_bounds__(kThreads)
pack_hist_kernel(const int64_t* __restrict__ ray_indices, long long n, long long n_rays,
                 long long n_tiles, int64_t* __restrict__ packed_info) {
    __shared__ TileSmem sm;
    const TileGeom<false> g(n, n_tiles);
    float ones[kVec][4];
    long long rid[kVec][4];
    uint32_t f[kVec], tail[kVec];
#pragma unroll
    for (int k = 0; k < kVec; ++k) {
        const long long gi = g.base + g.voff(k);
        long long prev = (gi > g.base && gi - 1 < n) ? __ldg(ray_indices + gi - 1) : -1;
#pragma unroll
        for (int c = 0; c < 4; ++c) {
            rid[k][c] = (gi + c < n) ? __ldg(ray_indices + gi + c) : -1;
            ones[k][c] = (gi + c < n) ? 1.0f : 0.0f;
        }
        const bool last_in_tile = (g.voff(k) + 4 == kTile);
        const long long after = (!last_in_tile && gi + 4 < n) ? __ldg(ray_indices + gi + 4) : -2;
        f[k] = 0; tail[k] = 0;
#pragma unroll
        for (int c = 0; c < 4; ++c) {
            const long long nx = (c < 3) ? ((gi + c + 1 < n) ? rid[k][(c + 1) & 3] : -2) : after;
            if (gi + c < n) {
                if (rid[k][c] != prev) f[k] |= (1u << (8 * c));
                if (nx != rid[k][c]) tail[k] |= (1u << c);
            }
            prev = rid[k][c];
        }
    }
    float E[kVec];
    tile_segscan<SumOp, false>(ones, f, E, nullptr, 0, 0u, sm);  // run lengths <= 4096: exact in fp32
#pragma unroll
    for (int k = 0; k < kVec; ++k) {
        float I[4];
        finish_inclusive<SumOp>(ones[k], f[k], E[k], I);
#pragma unroll
        for (int c = 0; c < 4; ++c)
            if ((tail[k] >> c) & 1u) {
                const long long r = rid[k][c];
                if (r >= 0 && r < n_rays)
                    atomicAdd(reinterpret_cast<unsigned long long*>(packed_info + 2 * r + 1),
                              (unsigned long long)(I[c] + 0.5f));
            }
    }
}

// Exclusive scan over rays, in place on packed_info: reads counts at [2r+1], writes starts
// at [2r]; the grand total goes to *total_out.  2048 rays per CTA, int64 look-back
// (status word = value << 2 | state; the status array must be zeroed before launch).
constexpr int kRayTile = 2048;
__global__ void __launch_bounds__(kThreads)
ray_scan_kernel(int64_t* __restrict__ packed_info, long long n_rays, long long* __restrict__ total_out,
                unsigned long long* status) {
    __shared__ long long s_warp[kWarps];
    __shared__ long long s_carry;
    const long long tile = blockIdx.x;
    const long long r0 = tile * kRayTile + (long long)threadIdx.x * 8;
    long long c[8];
    long long run = 0;
#pragma unroll
    for (int i = 0; i < 8; ++i) {
        c[i] = (r0 + i < n_rays) ? packed_info[2 * (r0 + i) + 1] : 0;
        run += c[i];
    }
    long long t = run;
    const int lane = lane_id(), w = warp_id();
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
        const long long y = __shfl_up_sync(kFull, t, d);
        if (lane >= d) t += y;
    }
    if (lane == 31) s_warp[w] = t;
    __syncthreads();
    long long before = 0, agg = 0;
#pragma unroll
    for (int j = 0; j < kWarps; ++j) { if (j == w) before = agg; agg += s_warp[j]; }
    if (threadIdx.x == 0) {
        long long X = 0;
        if (tile == 0) {
            st_relaxed_u64(status, ((unsigned long long)agg << 2) | 2ull);
        } else {
            st_relaxed_u64(status + tile, ((unsigned long long)agg << 2) | 1ull);
            long long k = tile - 1;
            while (true) {
                const unsigned long long s = ld_relaxed_u64(status + k);
                if ((s & 3ull) == 0ull) { __nanosleep(40); continue; }
                X += (long long)(s >> 2);
                if ((s & 3ull) == 2ull) break;
                --k;
            }
            st_relaxed_u64(status + tile, ((unsigned long long)(X + agg) << 2) | 2ull);
        }
        s_carry = X;
        if (tile == (long long)gridDim.x - 1 && total_out != nullptr) *total_out = X + agg;
    }
    __syncthreads();
    long long start = s_carry + before + (t - run);
#pragma unroll
    for (int i = 0; i < 8; ++i) {
        if (r0 + i < n_rays) packed_info[2 * (r0 + i)] = start;
        start += c[i];
    }
}

size_t ray_scan_status_bytes(long long n_rays) {
    return (size_t)ceil_div<long long>(n_rays > 0 ? n_rays : 1, kRayTile) * sizeof(unsigned long long);
}

// Launch the in-place per-ray scan; `status` must hold ray_scan_status_bytes(n_rays).
int launch_ray_scan(int64_t* packed_info, long long n_rays, long long* total_out,
                    void* status, cudaStream_t s) {
    const long long nt = ceil_div<long long>(n_rays, kRayTile);
    if (cudaMemsetAsync(status, 0, (size_t)nt * sizeof(unsigned long long), s) != cudaSuccess) {
        set_error("ray scan: memset failed");
        return NFB_ERR_CUDA;
    }
    ray_scan_kernel<<<(unsigned)nt, kThreads, 0, s>>>(packed_info, n_rays, total_out,
                                                      static_cast<unsigned long long*>(status));
    count_launch();
    NFB_CHECK_LAUNCH("ray_scan");
    return NFB_OK;
}

}  // namespace nfb

extern "C" {

NFB_API int nfb_pack_info(const int64_t* ray_indices, int64_t n, int64_t n_rays,
                          int64_t* packed_info, void* workspace, size_t workspace_bytes,
                          nfb_stream_t stream) {
    using namespace nfb;
    NFB_CHECK_ARG(n >= 0 && n_rays >= 0, "negative size");
    if (n_rays == 0) return NFB_OK;
    NFB_CHECK_ARG(packed_info != nullptr && (n == 0 || ray_indices != nullptr), "null pointer");
    NFB_CHECK_ARG(workspace != nullptr && workspace_bytes >= ray_scan_status_bytes(n_rays),
                  "workspace too small for %lld rays", (long long)n_rays);
    cudaStream_t s = static_cast<cudaStream_t>(stream);
    cudaMemsetAsync(packed_info, 0, sizeof(int64_t) * 2 * n_rays, s);
    if (n > 0) {
        const long long nt = num_tiles(n);
        pack_hist_kernel<<<(unsigned)nt, kThreads, 0, s>>>(ray_indices, n, n_rays, nt, packed_info);
        count_launch();
        NFB_CHECK_LAUNCH("pack_hist");
    }
    return launch_ray_scan(packed_info, n_rays, nullptr, workspace, s);
}

}  // extern "C"
