// pack.cu — pack_info (ray_indices -> packed_info) and the per-ray exclusive scan that
// turns per-ray counts into packed_info starts (shared with the traversal).
//
// Reference: nerfacc/pack.py:43-46 = index_add_ of ones (one int64 atomic per sample) +
// cumsum + stack (4 kernels, 3 temporaries).  Here: one pass that compares neighbours and issues
// two atomics per *run* of equal indices (run end minus run start), and one single-pass
// look-back scan over the rays that writes (start, count) interleaved in place.
#include "seg_tile.cuh"

namespace nfb {

// counts[r] += length of every run of r in ray_indices; counts live at packed_info[2r+1].
// A run [a, b) of equal indices contributes b - a: the thread holding its first element adds -a,
// the one holding its last adds b (two 64-bit atomics per run, none per sample; any number of
// runs per ray).  Every thread takes four consecutive indices as two 128-bit loads and gets its
// neighbours' edge elements by shuffle; only the first and last lane of a warp read one more.
constexpr int kHistPerThread = 4;
__global__ void __launch_bounds__(kThreads)
pack_hist_kernel(const int64_t* __restrict__ ray_indices, long long n, long long n_rays,
                 int64_t* __restrict__ packed_info, bool vec_ok, const unsigned long long* __restrict__ only_if) {
    if (only_if != nullptr && *only_if == 0ull) return;  // the sorted pass has done the work
    const int lane = lane_id();
    const long long per_block = (long long)kThreads * kHistPerThread;
    for (long long base = (long long)blockIdx.x * per_block; base < n; base += (long long)gridDim.x * per_block) {
        const long long i0 = base + (long long)threadIdx.x * kHistPerThread;
        long long v[kHistPerThread];
        if (vec_ok && i0 + kHistPerThread <= n) {
            const longlong2 a = __ldg(reinterpret_cast<const longlong2*>(ray_indices + i0));
            const longlong2 b = __ldg(reinterpret_cast<const longlong2*>(ray_indices + i0 + 2));
            v[0] = a.x; v[1] = a.y; v[2] = b.x; v[3] = b.y;
        } else {
#pragma unroll
            for (int c = 0; c < kHistPerThread; ++c) v[c] = (i0 + c < n) ? __ldg(ray_indices + i0 + c) : -1;
        }
        long long prev = __shfl_up_sync(kFull, v[kHistPerThread - 1], 1);
        long long next = __shfl_down_sync(kFull, v[0], 1);
        if (lane == 0) prev = (i0 > 0 && i0 - 1 < n) ? __ldg(ray_indices + i0 - 1) : -1;
        if (lane == 31) next = (i0 + kHistPerThread < n) ? __ldg(ray_indices + i0 + kHistPerThread) : -1;
#pragma unroll
        for (int c = 0; c < kHistPerThread; ++c) {
            const long long i = i0 + c;
            if (i >= n) break;
            const long long r = v[c];
            if (r < 0 || r >= n_rays) continue;
            const long long before = c == 0 ? prev : v[c - 1];
            const long long after = (c == kHistPerThread - 1) ? next : ((i + 1 < n) ? v[c + 1] : -1);
            long long add = 0;
            if (before != r) add -= i;
            if (after != r) add += i + 1;
            if (before != r || after != r)
                atomicAdd(reinterpret_cast<unsigned long long*>(packed_info + 2 * r + 1), (unsigned long long)add);
        }
    }
}

// The same for SORTED ray_indices (what every producer on this path emits), without zero-filling,
// atomics or a scan: at every position i where the index changes from a to b, the rays in (a, b] start
// at i and the rays in [a, b) end at i — plain stores, each ray written once; pack_finish_kernel turns
// (start, end) into (start, count) and fills in the rays after the last index.  An index that
// decreases (or lies outside [0, n_rays)) raises *flag: pack_finish_kernel then zero-fills instead and
// the general pass above (pack_hist_kernel + ray scan, launched behind with only_if = flag) takes over —
// the caller never has to know, and no host round trip is needed to find out.
constexpr int kRunsPerThread = 4;  // (8 per thread: no faster — the pass is bound by its scattered 8-byte stores)
__global__ void __launch_bounds__(kThreads)
pack_runs_kernel(const int64_t* __restrict__ ray_indices, long long n, long long n_rays,
                 int64_t* __restrict__ packed_info, bool vec_ok, unsigned long long* __restrict__ flag) {
    const int lane = lane_id();
    const long long per_block = (long long)kThreads * kRunsPerThread;
    for (long long base = (long long)blockIdx.x * per_block; base < n; base += (long long)gridDim.x * per_block) {
        const long long i0 = base + (long long)threadIdx.x * kRunsPerThread;
        long long v[kRunsPerThread];
        if (vec_ok && i0 + kRunsPerThread <= n) {
#pragma unroll
            for (int c = 0; c < kRunsPerThread; c += 2) {
                const longlong2 a = __ldg(reinterpret_cast<const longlong2*>(ray_indices + i0 + c));
                v[c] = a.x; v[c + 1] = a.y;
            }
        } else {
#pragma unroll
            for (int c = 0; c < kRunsPerThread; ++c) v[c] = (i0 + c < n) ? __ldg(ray_indices + i0 + c) : -1;
        }
        long long prev = __shfl_up_sync(kFull, v[kRunsPerThread - 1], 1);
        if (lane == 0) prev = (i0 > 0 && i0 - 1 < n) ? __ldg(ray_indices + i0 - 1) : -1;
#pragma unroll
        for (int c = 0; c < kRunsPerThread; ++c) {
            const long long i = i0 + c;
            if (i >= n) break;
            const long long r = v[c];
            const long long before = c == 0 ? prev : v[c - 1];
            if (r < 0 || r >= n_rays || before > r) { *flag = 1ull; continue; }
            if (before != r) {
                packed_info[2 * r] = i;                             // r starts here
                if (before >= 0) packed_info[2 * before + 1] = i;   // `before` ends here
                if (r > before + 1) {                               // rays without samples in between (rare)
                    for (long long rr = before + 1; rr < r; ++rr)
                        *reinterpret_cast<longlong2*>(packed_info + 2 * rr) = make_longlong2(i, i);
                }
            }
        }
    }
}

__global__ void __launch_bounds__(256)
pack_finish_kernel(const int64_t* __restrict__ ray_indices, long long n, long long n_rays,
                   int64_t* __restrict__ packed_info, const unsigned long long* __restrict__ flag) {
    const long long r = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (r >= n_rays) return;
    longlong2* __restrict__ e = reinterpret_cast<longlong2*>(packed_info) + r;
    if (*flag != 0ull) { *e = make_longlong2(0, 0); return; }  // zero-filled for the general pass
    const long long r_last = n > 0 ? __ldg(ray_indices + n - 1) : -1;
    if (r > r_last) { *e = make_longlong2(n, 0); return; }
    const longlong2 se = *e;
    *e = make_longlong2(se.x, (r == r_last ? n : se.y) - se.x);
}

// ---- sample positions (the callers' side of the callbacks) -----------------
// What every sigma_fn / rgb_sigma_fn of the reference's examples computes first
// (examples/utils.py:82-112):
//     t_origins = rays_o[ray_indices]; t_dirs = rays_d[ray_indices]
//     positions = t_origins + t_dirs * (t_starts + t_ends)[:, None] / 2.0
// — two gathers that materialise 24 B/sample, three elementwise kernels, 100+ B/sample of traffic.
// One pass here: a thread takes four consecutive samples (two 128-bit loads of ray ids, one each
// of t_starts / t_ends), gathers its rays' rows (neighbouring samples share a ray: L1 hits) and
// writes 48 contiguous bytes of positions (and of directions, if asked) as three 128-bit stores.
// Same fp32 operations in the same order as the torch expression: bit-identical.
__global__ void __launch_bounds__(256)
sample_positions_kernel(const float* __restrict__ rays_o, const float* __restrict__ rays_d,
                        const int64_t* __restrict__ ray_indices, const float* __restrict__ t_starts,
                        const float* __restrict__ t_ends, long long n, float* __restrict__ positions,
                        float* __restrict__ dirs, bool vec_ok) {
    for (long long i0 = ((long long)blockIdx.x * blockDim.x + threadIdx.x) * 4; i0 < n;
         i0 += (long long)gridDim.x * blockDim.x * 4) {
        long long r[4];
        float ts[4], te[4];
        const bool whole = vec_ok && i0 + 4 <= n;
        if (whole) {
            const longlong2 a = __ldg(reinterpret_cast<const longlong2*>(ray_indices + i0));
            const longlong2 b = __ldg(reinterpret_cast<const longlong2*>(ray_indices + i0 + 2));
            r[0] = a.x; r[1] = a.y; r[2] = b.x; r[3] = b.y;
            const float4 s = ldg_stream4(t_starts + i0), e = ldg_stream4(t_ends + i0);
            ts[0] = s.x; ts[1] = s.y; ts[2] = s.z; ts[3] = s.w;
            te[0] = e.x; te[1] = e.y; te[2] = e.z; te[3] = e.w;
        } else {
#pragma unroll
            for (int c = 0; c < 4; ++c) {
                const bool on = i0 + c < n;
                r[c] = on ? __ldg(ray_indices + i0 + c) : 0;
                ts[c] = on ? __ldg(t_starts + i0 + c) : 0.0f;
                te[c] = on ? __ldg(t_ends + i0 + c) : 0.0f;
            }
        }
        float p[12], dd[12];
#pragma unroll
        for (int c = 0; c < 4; ++c) {
            const float tsum = __fadd_rn(ts[c], te[c]);
#pragma unroll
            for (int k = 0; k < 3; ++k) {
                const float o = __ldg(rays_o + 3 * r[c] + k), d = __ldg(rays_d + 3 * r[c] + k);
                dd[3 * c + k] = d;
                p[3 * c + k] = __fadd_rn(o, __fmul_rn(__fmul_rn(d, tsum), 0.5f));  // o + (d * (ts + te)) / 2
            }
        }
        if (whole) {
#pragma unroll
            for (int v = 0; v < 3; ++v) {
                stg_stream4(positions + 3 * i0 + 4 * v, make_float4(p[4 * v], p[4 * v + 1], p[4 * v + 2], p[4 * v + 3]));
                if (dirs) stg_stream4(dirs + 3 * i0 + 4 * v, make_float4(dd[4 * v], dd[4 * v + 1], dd[4 * v + 2], dd[4 * v + 3]));
            }
        } else {
            for (int c = 0; c < 4 && i0 + c < n; ++c)
#pragma unroll
                for (int k = 0; k < 3; ++k) {
                    positions[3 * (i0 + c) + k] = p[3 * c + k];
                    if (dirs) dirs[3 * (i0 + c) + k] = dd[3 * c + k];
                }
        }
    }
}

__global__ void zero_words_kernel(unsigned long long* __restrict__ p, long long n) {
    const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) p[i] = 0ull;
}

// Exclusive scan over rays, in place on packed_info: reads counts at [2r+1], writes starts
// at [2r]; the grand total goes to *total_out.  2048 rays per CTA, int64 look-back
// (status word = value << 2 | state; the status array must be zeroed before launch).
constexpr int kRayTile = 2048;
__global__ void __launch_bounds__(kThreads)
ray_scan_kernel(int64_t* __restrict__ packed_info, long long n_rays, long long* __restrict__ total_out,
                unsigned long long* status, const unsigned long long* __restrict__ only_if,
                const unsigned long long* __restrict__ also_src, long long* __restrict__ also_dst) {
    if (only_if != nullptr && *only_if == 0ull) return;
    // a second word for the host, written beside the total (the caller's read-back then needs no copy
    // on the stream: with a large H2D transfer in flight an 8-byte cudaMemcpyAsync took 25 us)
    if (also_dst != nullptr && blockIdx.x == gridDim.x - 1 && threadIdx.x == 32) *also_dst = (long long)*also_src;
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
            s_carry = X;
            if (tile == (long long)gridDim.x - 1 && total_out != nullptr) *total_out = X + agg;
        }
    }
    __syncthreads();
    long long start = s_carry + before + (t - run);
#pragma unroll
    for (int i = 0; i < 8; ++i) {
        if (r0 + i < n_rays) packed_info[2 * (r0 + i)] = start;
        start += c[i];
    }
}

// one status word per tile of rays + one spare word (pack_info's "not sorted" flag)
size_t ray_scan_status_bytes(long long n_rays) {
    return (size_t)(ceil_div<long long>(n_rays > 0 ? n_rays : 1, kRayTile) + 1) * sizeof(unsigned long long);
}

// Launch the in-place per-ray scan; `status` must hold ray_scan_status_bytes(n_rays).
int launch_ray_scan(int64_t* packed_info, long long n_rays, long long* total_out,
                    void* status, cudaStream_t s, const unsigned long long* only_if,
                    const unsigned long long* also_src, long long* also_dst) {
    const long long nt = ceil_div<long long>(n_rays, kRayTile);
    // (a kernel, not cudaMemsetAsync: behind a large H2D transfer the memset of these few KB took 12 us)
    zero_words_kernel<<<(unsigned)ceil_div<long long>(nt, 256), 256, 0, s>>>(static_cast<unsigned long long*>(status), nt);
    ray_scan_kernel<<<(unsigned)nt, kThreads, 0, s>>>(packed_info, n_rays, total_out,
                                                      static_cast<unsigned long long*>(status), only_if,
                                                      also_src, also_dst);
    count_launch(2);
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
    NFB_CHECK_ARG(aligned16(packed_info), "packed_info must be 16-byte aligned");
    // sorted pass (pack_runs + pack_finish); the general pass behind it runs only if an index decreased
    const long long nt = ceil_div<long long>(n_rays, kRayTile);
    unsigned long long* flag = static_cast<unsigned long long*>(workspace) + nt;  // the status array's spare word
    cudaMemsetAsync(flag, 0, sizeof(unsigned long long), s);
    long long nb = ceil_div<long long>(n > 0 ? n : 1, (long long)kThreads * kHistPerThread);
    const long long cap = (long long)sm_count() * 16;
    if (nb > cap) nb = cap;
    if (n > 0) {
        long long nr = ceil_div<long long>(n, (long long)kThreads * kRunsPerThread);
        if (nr > cap) nr = cap;
        pack_runs_kernel<<<(unsigned)nr, kThreads, 0, s>>>(ray_indices, n, n_rays, packed_info, aligned16(ray_indices), flag);
        count_launch();
        NFB_CHECK_LAUNCH("pack_runs");
    }
    pack_finish_kernel<<<(unsigned)ceil_div<long long>(n_rays, 256), 256, 0, s>>>(ray_indices, n, n_rays, packed_info, flag);
    count_launch();
    NFB_CHECK_LAUNCH("pack_finish");
    if (n == 0) return NFB_OK;
    pack_hist_kernel<<<(unsigned)nb, kThreads, 0, s>>>(ray_indices, n, n_rays, packed_info, aligned16(ray_indices), flag);
    count_launch();
    NFB_CHECK_LAUNCH("pack_hist");
    return launch_ray_scan(packed_info, n_rays, nullptr, workspace, s, flag, nullptr, nullptr);
}

NFB_API int nfb_sample_positions(const float* rays_o, const float* rays_d, const int64_t* ray_indices,
                                 const float* t_starts, const float* t_ends, int64_t n, int64_t n_rays,
                                 float* positions, float* dirs, nfb_stream_t stream) {
    using namespace nfb;
    NFB_CHECK_ARG(n >= 0 && n_rays >= 0, "negative size");
    if (n == 0) return NFB_OK;
    NFB_CHECK_ARG(rays_o && rays_d && ray_indices && t_starts && t_ends && positions, "null pointer");
    const bool vec_ok = aligned16(ray_indices) && aligned16(t_starts) && aligned16(t_ends) && aligned16(positions) &&
                        (dirs == nullptr || aligned16(dirs));
    long long b = ceil_div<long long>(ceil_div<long long>(n, 4), 256);
    const long long cap = (long long)sm_count() * 32;
    if (b > cap) b = cap;
    sample_positions_kernel<<<(unsigned)b, 256, 0, static_cast<cudaStream_t>(stream)>>>(
        rays_o, rays_d, ray_indices, t_starts, t_ends, n, positions, dirs, vec_ok);
    count_launch();
    NFB_CHECK_LAUNCH("sample_positions");
    return NFB_OK;
}

}  // extern "C"
