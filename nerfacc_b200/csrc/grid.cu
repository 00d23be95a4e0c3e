// grid.cu — ray_aabb_intersect, occupancy bit-packing and the two-pass grid traversal
// (count -> per-ray scan -> fill) behind the C ABI.
//
// Replaces csrc/grid.cu of the reference.  What changed structurally:
//  * ray/box intersection + event sort are fused into the traversal when the caller does
//    not pass precomputed t_sorted/t_indices/hits (the reference runs ray_aabb_intersect,
//    torch.cat and torch.sort with int64 indices first, grid.py:158-162);
//  * occupancy is a 1-bit/cell field (256 KiB per 128^3 level) instead of 1-byte bools;
//  * per-ray counts go straight into the int64 packed_info layout and one look-back scan
//    produces starts and the grand totals (the reference: 2x torch.cumsum + 2x .item());
//  * OccGridEstimator.sampling gets exactly its API outputs (ray_indices, t_starts, t_ends)
//    from the fill pass; the wider traverse_grids outputs are a second emission mode.
#include "grid.cuh"

namespace nfb {

int launch_ray_scan(int64_t* packed_info, long long n_rays, long long* total_out, void* status,
                    cudaStream_t s);
size_t ray_scan_status_bytes(long long n_rays);

struct RayArgs {
    const float* __restrict__ rays_o;
    const float* __restrict__ rays_d;
    const uint8_t* __restrict__ rays_mask;   // optional
    const float* __restrict__ near_arr;      // optional, else near_s
    const float* __restrict__ far_arr;       // optional, else far_s
    float near_s, far_s;
    const float* __restrict__ t_sorted;      // optional precomputed events [n_rays, 2m]
    const int64_t* __restrict__ t_indices;   // [n_rays, 2m]
    const uint8_t* __restrict__ hits;        // [n_rays, m]
    long long n_rays;
};

__device__ __forceinline__ void load_ray(const RayArgs& A, const GridSpec& G, long long r,
                                         float (&o)[3], float (&d)[3], float& near, float& far,
                                         RayEvents& ev) {
#pragma unroll
    for (int k = 0; k < 3; ++k) { o[k] = __ldg(A.rays_o + 3 * r + k); d[k] = __ldg(A.rays_d + 3 * r + k); }
    near = A.near_arr ? __ldg(A.near_arr + r) : A.near_s;
    far = A.far_arr ? __ldg(A.far_arr + r) : A.far_s;
    const int m = G.n_grids;
    if (A.t_sorted != nullptr) {
        ev.hit_mask = 0;
        for (int l = 0; l < m; ++l)
            if (A.hits[r * m + l]) ev.hit_mask |= (1u << l);
        for (int i = 0; i < 2 * m; ++i) {
            ev.t[i] = __ldg(A.t_sorted + r * 2 * m + i);
            ev.idx[i] = (int8_t)__ldg(A.t_indices + r * 2 * m + i);
        }
    } else {
        float inv[3];
#pragma unroll
        for (int k = 0; k < 3; ++k) inv[k] = __fdiv_rn(1.0f, d[k]);
        compute_events(o, inv, G, ev);
    }
}

struct CountEmit {
    __device__ __forceinline__ void operator()(float, float, bool, int, int) const {}
};

// Pass 1: counts (into packed_info[2r+1]) and termination planes.
__global__ void __launch_bounds__(256)
traverse_count_kernel(const RayArgs A, const GridSpec G, const MarchOpts opt,
                      int64_t* __restrict__ sample_info, int64_t* __restrict__ interval_info,
                      float* __restrict__ term) {
    for (long long r = (long long)blockIdx.x * blockDim.x + threadIdx.x; r < A.n_rays;
         r += (long long)gridDim.x * blockDim.x) {
        float o[3], d[3], near, far;
        RayEvents ev;
        int nS = 0, nI = 0;
        float tt;
        if (A.rays_mask != nullptr && !A.rays_mask[r]) {
            tt = A.near_arr ? A.near_arr[r] : A.near_s;
        } else {
            load_ray(A, G, r, o, d, near, far, ev);
            CountEmit e;
            march_ray(o, d, near, far, G, ev, opt, e, nS, nI, tt);
        }
        if (sample_info) sample_info[2 * r + 1] = nS;
        if (interval_info) interval_info[2 * r + 1] = nI;
        if (term) term[r] = tt;
    }
}

struct SamplingEmit {  // OccGridEstimator.sampling outputs (occ_grid.py:174-177)
    int64_t* __restrict__ ray_indices;
    float* __restrict__ t_starts;
    float* __restrict__ t_ends;
    long long base, r;
    __device__ __forceinline__ void operator()(float t_last, float t_next, bool, int nS, int) const {
        const long long i = base + nS;
        ray_indices[i] = r;
        t_starts[i] = t_last;
        t_ends[i] = t_next;
    }
};

struct FullEmit {  // traverse_grids outputs (grid.cu:218-255)
    float* __restrict__ iv_vals; int64_t* __restrict__ iv_ray;
    uint8_t* __restrict__ iv_left; uint8_t* __restrict__ iv_right;
    float* __restrict__ sm_vals; int64_t* __restrict__ sm_ray; uint8_t* __restrict__ sm_valid;
    long long iv_base, sm_base, r;
    __device__ __forceinline__ void operator()(float t_last, float t_next, bool continuous, int nS, int nI) const {
        long long e = iv_base + nI;
        if (!continuous) {
            iv_vals[e] = t_last; iv_ray[e] = r; iv_left[e] = 1; iv_right[e] = 0;
            ++e;
        } else {
            iv_left[e - 1] = 1;
        }
        iv_vals[e] = t_next; iv_ray[e] = r; iv_left[e] = 0; iv_right[e] = 1;
        const long long s = sm_base + nS;
        sm_vals[s] = __fmul_rn(__fadd_rn(t_next, t_last), 0.5f);
        sm_ray[s] = r;
        sm_valid[s] = 1;
    }
};

struct FillArgs {
    // where each ray's data starts: info[r * stride]; packed_info (stride 2) or a plain array
    const int64_t* sm_starts; int sm_stride;
    const int64_t* iv_starts; int iv_stride;
    // two-pass: skip rays whose count (info[2r+1]) is zero; single pass: write the counts
    int64_t* sample_info; int64_t* interval_info; bool single_pass;
    float* term;
    SamplingEmit se; FullEmit fe; bool full;
};

__global__ void __launch_bounds__(256)
traverse_fill_kernel(const RayArgs A, const GridSpec G, const MarchOpts opt, const FillArgs F) {
    for (long long r = (long long)blockIdx.x * blockDim.x + threadIdx.x; r < A.n_rays;
         r += (long long)gridDim.x * blockDim.x) {
        if (A.rays_mask != nullptr && !A.rays_mask[r]) {
            if (F.single_pass) {
                if (F.sample_info) F.sample_info[2 * r + 1] = 0;
                if (F.interval_info) F.interval_info[2 * r + 1] = 0;
                if (F.term) F.term[r] = A.near_arr ? A.near_arr[r] : A.near_s;
            }
            continue;
        }
        if (!F.single_pass && F.sample_info[2 * r + 1] == 0) continue;  // grid.cu:103-106
        float o[3], d[3], near, far;
        RayEvents ev;
        load_ray(A, G, r, o, d, near, far, ev);
        int nS = 0, nI = 0;
        float tt;
        if (F.full) {
            FullEmit e = F.fe;
            e.r = r; e.sm_base = F.sm_starts[r * F.sm_stride]; e.iv_base = F.iv_starts[r * F.iv_stride];
            march_ray(o, d, near, far, G, ev, opt, e, nS, nI, tt);
        } else {
            SamplingEmit e = F.se;
            e.r = r; e.base = F.sm_starts[r * F.sm_stride];
            march_ray(o, d, near, far, G, ev, opt, e, nS, nI, tt);
        }
        if (F.single_pass) {
            if (F.sample_info) F.sample_info[2 * r + 1] = nS;
            if (F.interval_info) F.interval_info[2 * r + 1] = nI;
        }
        if (F.term) F.term[r] = tt;
    }
}

// csrc/grid.cu:284-313
__global__ void ray_aabb_kernel(const float* __restrict__ rays_o, const float* __restrict__ rays_d,
                                const float* __restrict__ aabbs, long long n_rays, int n_aabbs,
                                float near, float far, float miss, float* __restrict__ t_mins,
                                float* __restrict__ t_maxs, uint8_t* __restrict__ hits) {
    const long long total = n_rays * n_aabbs;
    for (long long t = (long long)blockIdx.x * blockDim.x + threadIdx.x; t < total;
         t += (long long)gridDim.x * blockDim.x) {
        const long long r = t / n_aabbs;
        const int b = (int)(t - r * n_aabbs);
        float o[3], inv[3];
#pragma unroll
        for (int k = 0; k < 3; ++k) { o[k] = __ldg(rays_o + 3 * r + k); inv[k] = __fdiv_rn(1.0f, __ldg(rays_d + 3 * r + k)); }
        float t0, t1;
        const bool hit = slab_test(o, inv, aabbs + 6 * b, near, far, t0, t1);
        t_mins[t] = hit ? t0 : miss;
        t_maxs[t] = hit ? t1 : miss;
        hits[t] = hit ? 1 : 0;
    }
}

// bool bytes -> bit field, 32 cells per word
__global__ void pack_bits_kernel(const uint8_t* __restrict__ binaries, long long n_cells,
                                 uint32_t* __restrict__ bits) {
    const long long n_words = (n_cells + 31) >> 5;
    const int lane = threadIdx.x & 31;
    for (long long w = ((long long)blockIdx.x * blockDim.x + threadIdx.x) >> 5; w < n_words;
         w += ((long long)gridDim.x * blockDim.x) >> 5) {
        const long long c = w * 32 + lane;
        const bool on = (c < n_cells) && binaries[c] != 0;
        const unsigned m = __ballot_sync(kFull, on);
        if (lane == 0) bits[w] = m;
    }
}

static inline unsigned grid_for(long long n, int block, int per_sm = 16) {
    long long b = ceil_div<long long>(n, block);
    const long long cap = (long long)sm_count() * per_sm;
    if (b > cap) b = cap;
    if (b < 1) b = 1;
    return (unsigned)b;
}

}  // namespace nfb

namespace nfb {
static int check_common(const nfb_rays_t* rays, const nfb_grid_t* grid, const nfb_march_t* march) {
    NFB_CHECK_ARG(rays && grid && march, "null descriptor");
    NFB_CHECK_ARG(rays->n_rays >= 0, "negative n_rays");
    NFB_CHECK_ARG(grid->n_grids >= 1 && grid->n_grids <= kMaxLevels, "n_grids must be in [1, %d]", kMaxLevels);
    NFB_CHECK_ARG(grid->res[0] > 0 && grid->res[1] > 0 && grid->res[2] > 0, "bad resolution");
    if (rays->n_rays > 0) {
        NFB_CHECK_ARG(rays->rays_o && rays->rays_d && grid->occ_bits && grid->aabbs, "null pointer");
        const int pre = (rays->t_sorted != nullptr) + (rays->t_indices != nullptr) + (rays->hits != nullptr);
        NFB_CHECK_ARG(pre == 0 || pre == 3, "t_sorted, t_indices and hits go together");
    }
    return NFB_OK;
}
static RayArgs to_args(const nfb_rays_t* r) {
    RayArgs a;
    a.rays_o = r->rays_o; a.rays_d = r->rays_d; a.rays_mask = r->rays_mask;
    a.near_arr = r->near_planes; a.far_arr = r->far_planes; a.near_s = r->near_plane; a.far_s = r->far_plane;
    a.t_sorted = r->t_sorted; a.t_indices = r->t_indices; a.hits = r->hits; a.n_rays = r->n_rays;
    return a;
}
static GridSpec to_grid(const nfb_grid_t* g) {
    GridSpec G;
    G.occ_bits = g->occ_bits; G.aabbs = g->aabbs; G.n_grids = g->n_grids;
    G.res[0] = g->res[0]; G.res[1] = g->res[1]; G.res[2] = g->res[2];
    G.cells_per_level = (long long)g->res[0] * g->res[1] * g->res[2];
    return G;
}
static MarchOpts to_opts(const nfb_march_t* m) { return MarchOpts{m->step_size, m->cone_angle, m->traverse_steps_limit}; }

static int scan_both(int64_t* sample_info, int64_t* interval_info, long long n_rays, int64_t* totals,
                     void* scratch, size_t scratch_bytes, cudaStream_t s) {
    const size_t one = ray_scan_status_bytes(n_rays);
    NFB_CHECK_ARG(scratch != nullptr && scratch_bytes >= 2 * one, "scratch too small");
    int rc = launch_ray_scan(sample_info, n_rays, reinterpret_cast<long long*>(totals), scratch, s);
    if (rc != NFB_OK) return rc;
    if (interval_info != nullptr)
        rc = launch_ray_scan(interval_info, n_rays, totals ? reinterpret_cast<long long*>(totals) + 1 : nullptr,
                             static_cast<char*>(scratch) + one, s);
    return rc;
}
}  // namespace nfb

extern "C" {

NFB_API int nfb_ray_aabb_intersect(const float* rays_o, const float* rays_d, const float* aabbs,
                                   int64_t n_rays, int64_t n_aabbs, float near_plane,
                                   float far_plane, float miss_value, float* t_mins,
                                   float* t_maxs, uint8_t* hits, nfb_stream_t stream) {
    using namespace nfb;
    NFB_CHECK_ARG(n_rays >= 0 && n_aabbs >= 0, "negative size");
    if (n_rays * n_aabbs == 0) return NFB_OK;
    NFB_CHECK_ARG(rays_o && rays_d && aabbs && t_mins && t_maxs && hits, "null pointer");
    ray_aabb_kernel<<<grid_for(n_rays * n_aabbs, 256), 256, 0, static_cast<cudaStream_t>(stream)>>>(
        rays_o, rays_d, aabbs, n_rays, (int)n_aabbs, near_plane, far_plane, miss_value, t_mins, t_maxs, hits);
    count_launch();
    NFB_CHECK_LAUNCH("ray_aabb_intersect");
    return NFB_OK;
}

NFB_API int nfb_pack_occupancy(const uint8_t* binaries, int64_t n_cells, uint32_t* bits,
                               nfb_stream_t stream) {
    using namespace nfb;
    NFB_CHECK_ARG(n_cells >= 0, "negative size");
    if (n_cells == 0) return NFB_OK;
    NFB_CHECK_ARG(binaries && bits, "null pointer");
    pack_bits_kernel<<<grid_for(n_cells, 256), 256, 0, static_cast<cudaStream_t>(stream)>>>(binaries, n_cells, bits);
    count_launch();
    NFB_CHECK_LAUNCH("pack_occupancy");
    return NFB_OK;
}

NFB_API int nfb_traverse_count(const nfb_rays_t* rays, const nfb_grid_t* grid,
                               const nfb_march_t* march, int64_t* sample_info,
                               int64_t* interval_info, float* termination_planes,
                               int64_t* totals, void* scratch, size_t scratch_bytes,
                               nfb_stream_t stream) {
    using namespace nfb;
    int rc = check_common(rays, grid, march);
    if (rc != NFB_OK) return rc;
    cudaStream_t s = static_cast<cudaStream_t>(stream);
    if (rays->n_rays == 0) {
        if (totals) cudaMemsetAsync(totals, 0, 2 * sizeof(int64_t), s);
        return NFB_OK;
    }
    NFB_CHECK_ARG(sample_info != nullptr, "null sample_info");
    RayArgs A = to_args(rays);
    A.rays_mask = nullptr;  // the two-pass path ignores rays_mask (grid.cu:418,450)
    traverse_count_kernel<<<grid_for(rays->n_rays, 256, 64), 256, 0, s>>>(
        A, to_grid(grid), to_opts(march), sample_info, interval_info, termination_planes);
    count_launch();
    NFB_CHECK_LAUNCH("traverse_count");
    return scan_both(sample_info, interval_info, rays->n_rays, totals, scratch, scratch_bytes, s);
}

NFB_API int nfb_traverse_fill_sampling(const nfb_rays_t* rays, const nfb_grid_t* grid,
                                       const nfb_march_t* march, const int64_t* sample_info,
                                       int64_t* ray_indices, float* t_starts, float* t_ends,
                                       nfb_stream_t stream) {
    using namespace nfb;
    int rc = check_common(rays, grid, march);
    if (rc != NFB_OK) return rc;
    if (rays->n_rays == 0) return NFB_OK;
    NFB_CHECK_ARG(sample_info != nullptr, "null sample_info");
    RayArgs A = to_args(rays);
    A.rays_mask = nullptr;
    FillArgs F{};
    F.sm_starts = sample_info; F.sm_stride = 2;
    F.sample_info = const_cast<int64_t*>(sample_info); F.single_pass = false; F.full = false;
    F.se.ray_indices = ray_indices; F.se.t_starts = t_starts; F.se.t_ends = t_ends;
    traverse_fill_kernel<<<grid_for(rays->n_rays, 256, 64), 256, 0, static_cast<cudaStream_t>(stream)>>>(
        A, to_grid(grid), to_opts(march), F);
    count_launch();
    NFB_CHECK_LAUNCH("traverse_fill_sampling");
    return NFB_OK;
}

NFB_API int nfb_traverse_fill(const nfb_rays_t* rays, const nfb_grid_t* grid,
                              const nfb_march_t* march, int over_allocate,
                              const int64_t* sample_starts, const int64_t* interval_starts,
                              int64_t* sample_info, int64_t* interval_info,
                              float* termination_planes, int64_t* totals, float* iv_vals,
                              int64_t* iv_ray_indices, uint8_t* iv_is_left, uint8_t* iv_is_right,
                              float* sm_vals, int64_t* sm_ray_indices, uint8_t* sm_is_valid,
                              void* scratch, size_t scratch_bytes, nfb_stream_t stream) {
    using namespace nfb;
    int rc = check_common(rays, grid, march);
    if (rc != NFB_OK) return rc;
    cudaStream_t s = static_cast<cudaStream_t>(stream);
    if (rays->n_rays == 0) return NFB_OK;
    NFB_CHECK_ARG(sample_info && interval_info, "null packed_info");
    RayArgs A = to_args(rays);
    FillArgs F{};
    F.full = true;
    F.sample_info = sample_info; F.interval_info = interval_info; F.term = termination_planes;
    F.fe.iv_vals = iv_vals; F.fe.iv_ray = iv_ray_indices; F.fe.iv_left = iv_is_left; F.fe.iv_right = iv_is_right;
    F.fe.sm_vals = sm_vals; F.fe.sm_ray = sm_ray_indices; F.fe.sm_valid = sm_is_valid;
    if (over_allocate) {
        NFB_CHECK_ARG(march->traverse_steps_limit > 0, "traverse_steps_limit must be > 0 when over_allocate is true");
        NFB_CHECK_ARG(sample_starts && interval_starts, "over_allocate needs the allocated starts");
        F.single_pass = true;
        F.sm_starts = sample_starts; F.sm_stride = 1;
        F.iv_starts = interval_starts; F.iv_stride = 1;
    } else {
        A.rays_mask = nullptr;
        F.single_pass = false;
        F.term = nullptr;  // already written by nfb_traverse_count for every ray
        F.sm_starts = sample_info; F.sm_stride = 2;
        F.iv_starts = interval_info; F.iv_stride = 2;
    }
    traverse_fill_kernel<<<grid_for(rays->n_rays, 256, 64), 256, 0, s>>>(A, to_grid(grid), to_opts(march), F);
    count_launch();
    NFB_CHECK_LAUNCH("traverse_fill");
    if (over_allocate) return scan_both(sample_info, interval_info, rays->n_rays, totals, scratch, scratch_bytes, s);
    return NFB_OK;
}

}  // extern "C"
