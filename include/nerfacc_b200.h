/*
 * nerfacc_b200.h — C ABI of libnerfacc_b200.so: the B200-native (sm_100a) replacement for
 * nerfacc's private CUDA extension on the rendering hot path.
 *
 * This is the drop-in boundary.  The reference reaches its kernels through the pybind
 * module registered at nerfacc/cuda/csrc/nerfacc.cpp:100-129 (torch::Tensor in/out);
 * each function below names the reference entry point it replaces.  Here the signatures
 * carry no torch types: plain device pointers, sizes, scalar options and the CUDA stream.
 *
 * Conventions
 *  - every pointer is a DEVICE pointer unless its name ends in `_host`;
 *  - the caller owns all memory (inputs and pre-allocated outputs); the library never
 *    allocates, frees or synchronises, and launches on `stream` only;
 *  - `packed_info` is int64 [n_rays, 2] = (start, count) per ray, nerfacc's own layout
 *    (nerfacc/pack.py:43-46); `ray_indices` is int64; masks are 1 byte per element;
 *  - return value 0 = success, otherwise an NFB_ERR_* code; nfb_last_error() gives text;
 *  - "canonical" packed_info means starts == exclusive cumsum of counts (what pack_info
 *    and traverse_grids produce).  Kernels that need it take a `canonical` flag and fall
 *    back to a general (any non-overlapping chunks) kernel when it is 0.
 */
#ifndef NERFACC_B200_H_
#define NERFACC_B200_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#if defined(__GNUC__)
#define NFB_API __attribute__((visibility("default")))
#else
#define NFB_API
#endif

typedef void* nfb_stream_t; /* cudaStream_t */

enum {
    NFB_OK = 0,
    NFB_ERR_INVALID_ARGUMENT = 1,
    NFB_ERR_CUDA = 2,
    NFB_ERR_WORKSPACE = 3
};

enum { NFB_OP_SUM = 0, NFB_OP_PROD = 1 };

/* Library version (major*10000 + minor*100 + patch) and the compiled SM architecture (100). */
NFB_API int nfb_version(void);
NFB_API int nfb_compiled_arch(void);
/* Thread-local text of the last error returned on this host thread. */
NFB_API const char* nfb_last_error(void);
/* Number of kernels this library has launched from this process (bench.py's gpu_launches). */
NFB_API long long nfb_launch_count(void);

/*
 * Look-back workspace.  The flat (canonical) kernels chain 4096-element tiles with
 * decoupled look-back and need `nfb_workspace_bytes(n)` bytes of device scratch for an
 * n-element array.  The scratch must be zero-filled once when allocated; after that it is
 * reused without clearing as long as the caller passes a fresh non-zero `epoch`
 * (1, 2, 3, ... per workspace, wrapping below 2^30) to each call.  epoch == 0 makes the
 * call clear the scratch itself with a memset node (use under CUDA-graph capture).
 */
NFB_API size_t nfb_workspace_bytes(int64_t n_elements);

/* ------------------------------------------------------------------------------------
 * Packed segmented scans.
 * Replaces inclusive_sum / exclusive_sum (csrc/scan.cu:9-125; `reverse`=1 is their
 * backward=True launch on reverse iterators) and inclusive/exclusive_prod_forward
 * (csrc/scan.cu:127-167, 217-257).  out[i] for i outside every chunk is left untouched.
 * `normalize` divides each chunk by max(total, 1e-10) (utils_scan.cuh:102-110); sum only.
 */
NFB_API int nfb_scan_f32(const float* in, float* out, const int64_t* packed_info,
                         int64_t n_rays, int64_t n, int op, int exclusive, int reverse,
                         int normalize, int canonical, void* workspace, size_t workspace_bytes,
                         uint32_t epoch, nfb_stream_t stream);

/* Replaces inclusive/exclusive_prod_backward (csrc/scan.cu:169-214, 259-304):
 * grad_in = reverse_sum_scan(grad_out * out) / max(in, 1e-10), fused in one pass. */
NFB_API int nfb_prod_scan_bwd_f32(const float* grad_out, const float* in, const float* out,
                                  float* grad_in, const int64_t* packed_info, int64_t n_rays,
                                  int64_t n, int exclusive, int canonical, void* workspace,
                                  size_t workspace_bytes, uint32_t epoch, nfb_stream_t stream);

/*
 * Head bits: the ray boundaries of a canonical packed_info as one bit per sample (bit i set iff
 * sample i is the first of a ray; bit n always set), `nfb_head_bits_bytes(n)` bytes.  Built once
 * per packed_info (the call clears `bits` itself) and shared by every scan over that layout:
 * the `_bits` scans below read 1 bit per sample instead of searching packed_info per tile, and
 * every warp works on its own (warp-granular look-back, no block barrier).  Same results as
 * nfb_scan_f32 / nfb_prod_scan_bwd_f32 with canonical=1 (reference: csrc/scan.cu:9-304).
 */
NFB_API size_t nfb_head_bits_bytes(int64_t n_elements);
NFB_API int nfb_head_bits(const int64_t* packed_info, int64_t n_rays, int64_t n, uint32_t* bits,
                          nfb_stream_t stream);
NFB_API int nfb_scan_bits_f32(const float* in, float* out, const uint32_t* head_bits, int64_t n,
                              int op, int exclusive, int reverse, void* workspace,
                              size_t workspace_bytes, uint32_t epoch, nfb_stream_t stream);
NFB_API int nfb_prod_scan_bwd_bits_f32(const float* grad_out, const float* in, const float* out,
                                       float* grad_in, const uint32_t* head_bits, int64_t n,
                                       int exclusive, void* workspace, size_t workspace_bytes,
                                       uint32_t epoch, nfb_stream_t stream);

/* ------------------------------------------------------------------------------------
 * Fused volume-rendering weights (forward).
 * Replaces the torch composites render_transmittance_from_alpha/density,
 * render_weight_from_alpha/density and render_visibility_from_alpha/density
 * (nerfacc/volrend.py:161-480) and the packed exclusive_prod / exclusive_sum inside them.
 *   from_density == 0: inputs `alphas`;      T = exclusive_prod(1-alphas) [* prefix_trans]
 *   from_density == 1: inputs t_starts,t_ends,sigmas; alphas = 1-exp(-sigma*(te-ts)),
 *                      T = exp(-exclusive_sum(sigma*(te-ts))) [* prefix_trans]
 * Outputs (each optional, may be NULL): weights = T*alphas, trans = T, alphas_out,
 * vis[i] = T >= early_stop_eps && (alpha_thre <= 0 || alphas >= alpha_thre).
 * Segments: canonical `packed_info` [n_rays,2], or packed_info == NULL and `uniform_len` > 0
 * for a batched [n_rays, uniform_len] tensor.
 */
NFB_API int nfb_render_weight_fwd(int from_density, const float* alphas, const float* t_starts,
                                  const float* t_ends, const float* sigmas,
                                  const float* prefix_trans, const int64_t* packed_info,
                                  int64_t n_rays, int64_t uniform_len, int64_t n, float* weights,
                                  float* trans, float* alphas_out, uint8_t* vis,
                                  float early_stop_eps, float alpha_thre, void* workspace,
                                  size_t workspace_bytes, uint32_t epoch, nfb_stream_t stream);

/* Backward of the above, equal to what torch autograd derives from the reference composite
 * (volrend.py:203-205, 259-263, 308, 361 with scan.py:233-242, 282-288 and
 * csrc/scan.cu:259-304).  g_weights / g_trans / g_alphas may each be NULL (= zero).
 * `trans` is the forward output (prefix already applied).  g_in receives d/d alphas
 * (from_density == 0) or d/d sigmas (from_density == 1). */
NFB_API int nfb_render_weight_bwd(int from_density, const float* g_weights, const float* g_trans,
                                  const float* g_alphas, const float* alphas,
                                  const float* t_starts, const float* t_ends, const float* sigmas,
                                  const float* trans, const int64_t* packed_info, int64_t n_rays,
                                  int64_t uniform_len, int64_t n, float* g_in, void* workspace,
                                  size_t workspace_bytes, uint32_t epoch, nfb_stream_t stream);

/* The two calls above for a dense packed layout given as head bits (nfb_head_bits) instead of
 * packed_info: same per-sample math, warp tiles instead of 4096-sample block tiles. */
NFB_API int nfb_render_weight_bits_fwd(int from_density, const float* alphas, const float* t_starts,
                                       const float* t_ends, const float* sigmas,
                                       const float* prefix_trans, const uint32_t* head_bits,
                                       int64_t n, float* weights, float* trans, float* alphas_out,
                                       uint8_t* vis, float early_stop_eps, float alpha_thre,
                                       void* workspace, size_t workspace_bytes, uint32_t epoch,
                                       nfb_stream_t stream);
NFB_API int nfb_render_weight_bits_bwd(int from_density, const float* g_weights, const float* g_trans,
                                       const float* g_alphas, const float* alphas,
                                       const float* t_starts, const float* t_ends,
                                       const float* sigmas, const float* trans,
                                       const uint32_t* head_bits, int64_t n, float* g_in,
                                       void* workspace, size_t workspace_bytes, uint32_t epoch,
                                       nfb_stream_t stream);

/* ------------------------------------------------------------------------------------
 * accumulate_along_rays: out[r, :] (+)= sum_{i in ray r} weights[i] * values[i, :].
 * Replaces the temporary + index_add_ atomics of nerfacc/volrend.py:532-547 (and the
 * in-place variant :550-573 with add_into = 1).  values == NULL accumulates the weights
 * (dim must be 1).  With a canonical `packed_info` and dim <= 4 the reduction is
 * deterministic and atomic-free and writes every row of `out` (zeros for empty rays);
 * otherwise `ray_indices` (any order) is required and run totals are added atomically.
 */
NFB_API int nfb_accumulate_fwd(const float* weights, const float* values, int64_t dim,
                               const int64_t* packed_info, const int64_t* ray_indices,
                               int64_t n_rays, int64_t n, float* out, int add_into,
                               void* workspace, size_t workspace_bytes, uint32_t epoch,
                               nfb_stream_t stream);

/* Backward of accumulate_along_rays (what autograd derives from volrend.py:537-544):
 * g_weights[i] = <g_out[ray_i, :], values[i, :]>, g_values[i, :] = weights[i] * g_out[ray_i, :].
 * Either output may be NULL. */
NFB_API int nfb_accumulate_bwd(const float* g_out, const float* weights, const float* values,
                               int64_t dim, const int64_t* ray_indices, int64_t n,
                               float* g_weights, float* g_values, nfb_stream_t stream);

/* ------------------------------------------------------------------------------------
 * Compositing of nerfacc.rendering() (nerfacc/volrend.py:140-156), fused.
 * Forward: colors[n_rays,3] = sum w*rgb, opacity[n_rays] = sum w, depth[n_rays] = sum
 * w*(t_starts+t_ends)/2 per ray in one pass (replaces three accumulate_along_rays calls);
 * canonical packed_info required; deterministic, every row written (zeros for empty rays). */
NFB_API int nfb_composite_fwd(const float* weights, const float* rgbs, const float* t_starts,
                              const float* t_ends, const int64_t* packed_info, int64_t n_rays,
                              int64_t n, float* colors, float* opacity, float* depth,
                              void* workspace, size_t workspace_bytes, uint32_t epoch,
                              nfb_stream_t stream);
/* Backward of rendering() down to the radiance field outputs, one reverse pass: what autograd
 * derives from volrend.py:140-151 (three accumulate_along_rays) chained with the weight backward
 * (nfb_render_weight_bwd).  g_colors [n_rays,3], g_opacity [n_rays], g_depth [n_rays] (wrt the
 * un-normalised depth sum) and the optional per-sample g_weights / g_trans / g_alphas may each
 * be NULL.  Writes g_in (d/d sigmas or d/d alphas) and g_rgbs [n,3] (optional). */
NFB_API int nfb_composite_bwd(int from_density, const float* g_colors, const float* g_opacity,
                              const float* g_depth, const float* g_weights, const float* g_trans,
                              const float* g_alphas, const float* alphas, const float* t_starts,
                              const float* t_ends, const float* sigmas, const float* rgbs,
                              const float* trans, const int64_t* ray_indices,
                              const int64_t* packed_info, int64_t n_rays, int64_t n, float* g_in,
                              float* g_rgbs, void* workspace, size_t workspace_bytes,
                              uint32_t epoch, nfb_stream_t stream);

/* ------------------------------------------------------------------------------------
 * pack_info: ray_indices [n] -> packed_info [n_rays, 2] (nerfacc/pack.py:43-46: histogram of
 * the indices, starts = exclusive cumsum).  `workspace` needs 8 bytes per 2048 rays (rounded up)
 * plus 8; packed_info must be 16-byte aligned.  Sorted indices — what every producer on this path
 * emits — take two plain passes (run boundaries, then counts: no zero-fill, no atomics, no scan); an
 * index that decreases is noticed on the device and the general pass (atomics per run + a scan over
 * the rays) redoes the work behind it, without a host round trip.
 */
NFB_API int nfb_pack_info(const int64_t* ray_indices, int64_t n, int64_t n_rays,
                          int64_t* packed_info, void* workspace, size_t workspace_bytes,
                          nfb_stream_t stream);

/* ------------------------------------------------------------------------------------
 * Occupancy-grid traversal.
 * Replaces ray_aabb_intersect and traverse_grids of the reference's extension
 * (csrc/grid.cu:284-313, 320-519; bound at csrc/nerfacc.cpp:42-70) and the
 * intersect + cat + sort prologue of nerfacc/grid.py:158-162.
 */
typedef struct nfb_rays {
    const float* rays_o;        /* [n_rays, 3] */
    const float* rays_d;        /* [n_rays, 3] */
    const uint8_t* rays_mask;   /* [n_rays] or NULL (honoured like the reference: single-pass mode) */
    const float* near_planes;   /* [n_rays] or NULL -> near_plane */
    const float* far_planes;    /* [n_rays] or NULL -> far_plane */
    float near_plane, far_plane;
    const float* t_sorted;      /* [n_rays, 2m] precomputed sorted ray/box events, or NULL: computed */
    const int64_t* t_indices;   /* [n_rays, 2m]   in the kernel (near=-inf, far=inf, miss=inf,   */
    const uint8_t* hits;        /* [n_rays, m]     stable sort), grid.py:158-162                  */
    int64_t n_rays;
} nfb_rays_t;

typedef struct nfb_grid {
    const uint32_t* occ_bits;   /* bit field from nfb_pack_occupancy: [n_grids*rx*ry*rz] bits */
    const uint32_t* occ_mip;    /* optional "any occupied" mip from nfb_pack_occupancy, or NULL */
    int32_t mip_shift;          /* one mip bit per (2^mip_shift)^3 block of cells */
    const float* aabbs;         /* [n_grids, 6] */
    int32_t n_grids;            /* <= 8 (MAX_GRID_LEVELS, csrc/grid.cu:18) */
    int32_t res[3];
    const int32_t* occ_bounds;  /* optional, from nfb_occupancy_bounds: per level the index box
                                   (x0,y0,z0,x1,y1,z1) holding every occupied cell, or NULL */
} nfb_grid_t;

typedef struct nfb_march {
    float step_size;            /* <= 0: one interval per occupied voxel crossing */
    float cone_angle;
    int32_t traverse_steps_limit; /* <= 0: none */
} nfb_march_t;

/* t_mins/t_maxs [n_rays, n_aabbs] float, hits [n_rays, n_aabbs] bytes. */
NFB_API int nfb_ray_aabb_intersect(const float* rays_o, const float* rays_d, const float* aabbs,
                                   int64_t n_rays, int64_t n_aabbs, float near_plane,
                                   float far_plane, float miss_value, float* t_mins,
                                   float* t_maxs, uint8_t* hits, nfb_stream_t stream);

/* binaries [n_grids, rx, ry, rz]: 1 byte per cell (torch.bool) -> bits: ceil(n_cells/32) uint32
 * words, and optionally (mip_bits != NULL) the "any occupied" mip with one bit per
 * (2^mip_shift)^3 block: ceil(n_grids * prod(ceil(res/2^mip_shift)) / 32) words.  Marching
 * kernels keep the mip in shared memory (used when it is <= 160 KiB).  `res_host`: HOST int[3]. */
NFB_API int nfb_pack_occupancy(const uint8_t* binaries, int32_t n_grids, const int32_t* res_host,
                               uint32_t* bits, uint32_t* mip_bits, int32_t mip_shift,
                               nfb_stream_t stream);

/* bounds [n_grids, 6] int32 (device): per level the inclusive index box of the occupied cells,
 * (res, res, res, -1, -1, -1) for a level without any.  A march that is past the box on an axis
 * it moves away along can stop walking: only empty cells follow (the reference walks them to
 * the end of the grid; nothing is emitted there).  `res_host`: HOST int[3]. */
NFB_API int nfb_occupancy_bounds(const uint32_t* bits, int32_t n_grids, const int32_t* res_host,
                                 int32_t* bounds, nfb_stream_t stream);

/* Scratch needed by the traversal entry points: per-ray scan status words, ray counters and 64 bytes
 * per ray for the prepared-ray records of the plain single-level count pass (16-byte aligned). */
NFB_API size_t nfb_traverse_scratch_bytes(int64_t n_rays);

/* Pass 1 (csrc/grid.cu:413-438 + data_spec.hpp:90-93): per-ray counts -> complete, canonical
 * packed_info for samples and for interval edges (int64 [n_rays,2] each, interval_info may be
 * NULL), termination planes [n_rays] (may be NULL; defined for every ray), and the grand
 * totals {n_samples, n_edges, n_rays whose run list overflowed} written to totals[0..2] (any
 * device-visible memory, e.g. pinned host memory: read it after synchronising the stream).
 * `scratch`: nfb_traverse_scratch_bytes(n_rays) bytes. */
NFB_API int nfb_traverse_count(const nfb_rays_t* rays, const nfb_grid_t* grid,
                               const nfb_march_t* march, int64_t* sample_info,
                               int64_t* interval_info, float* termination_planes,
                               int64_t* totals, void* scratch, size_t scratch_bytes,
                               void* run_buf, size_t run_buf_bytes, nfb_stream_t stream);

/* Optional run buffer (nfb_traverse_run_bytes(n_rays) bytes, 16-byte aligned, contents need
 * not be initialised).  When step_size > 0 and cone_angle == 0 the marched samples of a ray are
 * a few arithmetic progressions ("runs": the reference's t += dt recurrence is exactly linear
 * inside one float binade); pass 1 records them here and nfb_traverse_fill_sampling, given the
 * same buffer, expands them warp-cooperatively instead of marching a second time. */
NFB_API size_t nfb_traverse_run_bytes(int64_t n_rays);

/* The run-buffer part of pass 2 alone, for outputs that hold `capacity` samples (writes past
 * them are dropped): lets a caller that remembers the previous step's total expand the runs
 * BEFORE it reads this step's totals back, so the device does not idle during that host round
 * trip; if the total turns out larger than `capacity` the caller runs nfb_traverse_fill_sampling
 * on exact buffers instead. */
NFB_API int nfb_traverse_expand(int64_t n_rays, const int64_t* sample_info, const void* run_buf,
                                int64_t capacity, int64_t* ray_indices, float* t_starts, float* t_ends,
                                void* scratch, size_t scratch_bytes, nfb_stream_t stream);

/* The same for cone stepping (step_size > 0 and cone_angle != 0; csrc/grid.cu:23-28, 208-262): dt
 * follows t, so a run of samples has no closed form, but it is still the pure recurrence
 * t += calc_dt(t) from its first t.  Given a piece log as `run_buf` (nfb_traverse_piece_bytes(n_rays)
 * bytes, 16-byte aligned, contents need not be initialised) pass 1 cuts every run into pieces of at
 * most 32 samples and logs one 16-byte record per piece; nfb_traverse_expand_cone — or
 * nfb_traverse_fill_sampling given the same buffer — replays the pieces instead of marching the grid
 * again.  totals[2] != 0 after pass 1: the log ran full (more than 32 pieces per ray on average);
 * call nfb_traverse_fill_sampling with run_buf = NULL, which marches.  After
 * nfb_traverse_expand_cone has run, nfb_traverse_fill_sampling takes n_overflowed = -1 and does
 * nothing. */
NFB_API size_t nfb_traverse_piece_bytes(int64_t n_rays);
NFB_API int nfb_traverse_expand_cone(int64_t n_rays, const int64_t* sample_info, const void* piece_buf,
                                     float step_size, float cone_angle, int64_t capacity,
                                     int64_t* ray_indices, float* t_starts, float* t_ends,
                                     void* scratch, size_t scratch_bytes, nfb_stream_t stream);

/* Pass 2, OccGridEstimator.sampling outputs (estimators/occ_grid.py:174-177): for every sample,
 * in packed order, ray_indices, t_starts (= left edge) and t_ends (= right edge).
 * n_overflowed = totals[2] of pass 1; or, after nfb_traverse_expand has expanded the runs,
 * -(totals[2]) - 1 (< 0) to march only the rays whose runs overflowed the buffer. */
NFB_API int nfb_traverse_fill_sampling(const nfb_rays_t* rays, const nfb_grid_t* grid,
                                       const nfb_march_t* march, const int64_t* sample_info,
                                       int64_t* ray_indices, float* t_starts, float* t_ends,
                                       const void* run_buf /* or NULL */,
                                       int64_t n_overflowed,
                                       void* scratch, size_t scratch_bytes, nfb_stream_t stream);

/* Pass 2, traverse_grids outputs (csrc/grid.cu:445-470): interval edges {vals, ray_indices,
 * is_left, is_right} and samples {vals, ray_indices, is_valid}.
 * over_allocate == 0: two-pass layout, data at the starts recorded in sample_info/interval_info
 *   by nfb_traverse_count; every slot is written.
 * over_allocate == 1 (csrc/grid.cu:364-404): single pass; ray r's data goes to
 *   sample_starts[r] / interval_starts[r] (the caller's over-allocated layout, arrays must be
 *   zero-filled), rays_mask is honoured, and sample_info/interval_info/termination_planes/
 *   totals are produced here with the *compact* starts, as the reference does. */
NFB_API int nfb_traverse_fill(const nfb_rays_t* rays, const nfb_grid_t* grid,
                              const nfb_march_t* march, int over_allocate,
                              const int64_t* sample_starts, const int64_t* interval_starts,
                              int64_t* sample_info, int64_t* interval_info,
                              float* termination_planes, int64_t* totals, float* iv_vals,
                              int64_t* iv_ray_indices, uint8_t* iv_is_left, uint8_t* iv_is_right,
                              float* sm_vals, int64_t* sm_ray_indices, uint8_t* sm_is_valid,
                              void* scratch, size_t scratch_bytes, nfb_stream_t stream);

/* ------------------------------------------------------------------------------------
 * The test-time rendering loop of the reference's callers (examples/utils.py:252-425,
 * render_image_with_occgrid_test): per iteration the alive rays march up to k
 * more samples from their termination planes (traverse_grids(over_allocate=True,
 * traverse_steps_limit=k, rays_mask=alive, near_planes=termination_planes)), the radiance field is
 * queried, the chunk is composited behind what the rays already hold (render_weight_from_density
 * with prefix_trans = 1 - opacity, the alpha_thre filter, accumulate_along_rays_ x3) and rays that
 * are opaque or did not fill their chunk retire.  Here the loop state — alive list, per-ray near
 * planes, k — stays on the device: an iteration is nfb_chunk_march, the caller's field, and
 * nfb_chunk_composite, with ONE host read (the chunk's sample total) in between.
 *   state: nfb_chunk_state_bytes() device bytes;  plan_host: 2 int64 in pinned (device-visible)
 *   memory, receives (rays alive, k) of the next chunk whenever a plan is made;
 *   alive / alive_next: int32 [n_rays] double buffer (alive starts as 0..n_rays-1);
 *   slot_of: int64 [n_rays], a ray's position in the alive list (starts as 0..n_rays-1);
 *   sample_info: int64 [n_rays,2], the chunk's packed_info (zero counts for retired rays);
 *   slot_t_starts/ends: float [n_rays * max(min_samples, 1)] marching slots;
 *   ray_indices / t_starts / t_ends: the dense chunk the field sees, same capacity. */
NFB_API size_t nfb_chunk_state_bytes(void);
NFB_API int nfb_chunk_init(void* state, int64_t n_rays, int32_t min_samples, int64_t* plan_host,
                           nfb_stream_t stream);
/* rays->near_planes [n_rays] is read and overwritten in place with the termination planes of the
 * rays marched; n_alive_max: an upper bound of the alive list's length (sizes the launch);
 * totals[0] = samples of the chunk.  scratch: nfb_traverse_scratch_bytes(n_rays). */
NFB_API int nfb_chunk_march(const nfb_rays_t* rays, const nfb_grid_t* grid, const nfb_march_t* march,
                            const int32_t* alive, int64_t n_alive_max, const void* state,
                            const int64_t* slot_of, int64_t* sample_info, float* slot_t_starts,
                            float* slot_t_ends, int64_t* ray_indices, float* t_starts, float* t_ends,
                            int64_t* totals, void* scratch, size_t scratch_bytes, nfb_stream_t stream);
/* rgb [n_rays,3], opacity [n_rays], depth [n_rays] accumulate in place; a ray stays alive iff
 * opacity <= opacity_thre (= 1 - early_stop_eps) and it marched its full k samples; then plans the
 * next chunk: k = max(min(n_rays / alive, 64), min_samples). */
NFB_API int nfb_chunk_composite(const int32_t* alive, int32_t* alive_next, int64_t n_alive_max, void* state,
                                int64_t* sample_info, int64_t* slot_of, const float* t_starts,
                                const float* t_ends, const float* sigmas, const float* rgbs, int64_t n_rays,
                                float* rgb, float* opacity, float* depth, float alpha_thre,
                                float opacity_thre, int32_t min_samples, int64_t* plan_host,
                                nfb_stream_t stream);

/* ------------------------------------------------------------------------------------
 * Occupancy-grid update (nerfacc/estimators/occ_grid.py:367-404) and visibility compaction
 * (occ_grid.py:216-220).
 */
/* x[i,:] = aabb[:3] + ((ijk(cells[i]) + u[i,:]) / res) * (aabb[3:] - aabb[:3])  (occ_grid.py:385-392).
 * `res_host` is a HOST pointer to 3 ints; cell id = (i*res_y + j)*res_z + k. */
NFB_API int nfb_occ_cell_points(const int64_t* cells, const float* u, int64_t n,
                                const int32_t* res_host, const float* aabb, float* x,
                                nfb_stream_t stream);
/* occs[c] = max(occs[c]*decay, max over k in [eval_lo,eval_hi) with cell_ids[k]==c of occ[k-eval_lo])
 * for every c in cell_ids[0..n): scatter-max semantics of occ_grid.py:395-402, the decay applied
 * once per touched cell.  `tmp`: n floats of scratch. */
NFB_API int nfb_occ_ema_update(float* occs, const int64_t* cell_ids, int64_t n, const float* occ,
                               int64_t eval_lo, int64_t eval_hi, float decay, float* tmp,
                               nfb_stream_t stream);
/* thre = min(mean(occs[occs >= 0]), occ_thre); binaries[i] = occs[i] > thre (occ_grid.py:403-404).
 * thre_out (device, optional) receives the threshold. */
NFB_API size_t nfb_occ_binarize_workspace_bytes(void);
NFB_API int nfb_occ_binarize(const float* occs, int64_t n_cells, float occ_thre, uint8_t* binaries,
                             float* thre_out, void* workspace, size_t workspace_bytes,
                             nfb_stream_t stream);
/* OccGridEstimator.mark_invisible_cells (nerfacc/estimators/occ_grid.py:263-332): occs[c] = 0 for
 * the cells (occs[c] >= 0 on entry) whose corner-lattice point projects into at least one camera's
 * image at depth >= near_plane and into none at 0 <= depth < near_plane, else -1.
 * cams: [n_cams, 21] floats per camera = world->camera rotation (9, row-major), translation (3),
 * intrinsics K (9, row-major); occs: [levels * rx*ry*rz]; res_host: HOST pointer to 3 ints. */
NFB_API int nfb_mark_invisible_cells(const float* cams, int64_t n_cams, const int32_t* res_host,
                                     int32_t levels, const float* aabbs, float width, float height,
                                     float near_plane, float* occs, nfb_stream_t stream);
/* The two data-dependent cell selections of OccGridEstimator._sample_uniform_and_occupied_cells
 * (nerfacc/estimators/occ_grid.py:344-365), one launch each, order preserved:
 *   candidates != NULL: the drawn cells with occs[candidates[i]] >= 0 (`uniform[occs[uniform] >= 0]`;
 *                       occs = that level's slice, binaries = NULL);
 *   binaries   != NULL: the indices i of the n cells with binaries[i] != 0 (`nonzero(binaries)`;
 *                       occs = candidates = NULL).
 * out[(base ? *base : 0) + k] = the k-th kept cell; *total = kept cells of THIS call.  `base` and
 * `total` are device-visible; passing the first call's `total` as the second call's `base` appends
 * the occupied cells behind the uniform ones with no host step in between (the reference: two
 * boolean selects, each with its own host read-back, and a cat).
 * workspace: nfb_compact_workspace_bytes(n). */
NFB_API int nfb_occ_select_cells(const float* occs, const int64_t* candidates, const uint8_t* binaries,
                                 int64_t n, const int64_t* base, int64_t* out, int64_t* total,
                                 void* workspace, size_t workspace_bytes, nfb_stream_t stream);
/* Keep the samples with mask[i] != 0, order preserved; *total (device-visible) = how many. */
NFB_API size_t nfb_compact_workspace_bytes(int64_t n);
NFB_API int nfb_compact_samples(const uint8_t* mask, const int64_t* ray_indices,
                                const float* t_starts, const float* t_ends, int64_t n,
                                int64_t* out_ray_indices, float* out_t_starts, float* out_t_ends,
                                int64_t* total, void* workspace, size_t workspace_bytes,
                                nfb_stream_t stream);
/* The whole "skip invisible space" step of OccGridEstimator.sampling
 * (nerfacc/estimators/occ_grid.py:186-220: render_visibility_from_{alpha,density} -> bool mask ->
 * ray_indices[mask], t_starts[mask], t_ends[mask]; nerfacc/volrend.py:365-480) in ONE pass over the
 * samples: T by the segmented scan of nfb_render_weight_fwd, vis = T >= early_stop_eps
 * [& alpha >= alpha_thre if alpha_thre > 0], survivors ranked in the tile, global offset by
 * decoupled look-back, (ray_indices, t_starts, t_ends) written compacted in order; *total
 * (device-visible) = how many.  `packed_info` must be canonical and dense over the n samples
 * (what sampling / traverse_grids / pack_info produce).  Optional outputs: out_packed_info
 * [n_rays,2] = pack_info of the survivors (what nerfacc/pack.py:38-46 would compute from
 * out_ray_indices); vis [n] = the mask itself.  workspace: nfb_workspace_bytes(n). */
NFB_API int nfb_visibility_compact(int from_density, const float* alphas, const float* t_starts,
                                   const float* t_ends, const float* sigmas, const float* prefix_trans,
                                   const int64_t* packed_info, int64_t n_rays, int64_t n,
                                   const int64_t* ray_indices, float early_stop_eps, float alpha_thre,
                                   int64_t* out_ray_indices, float* out_t_starts, float* out_t_ends,
                                   int64_t* out_packed_info, uint8_t* vis, int64_t* total,
                                   void* workspace, size_t workspace_bytes, uint32_t epoch,
                                   nfb_stream_t stream);
/* nfb_compact_samples without a workspace argument, for callers that cannot allocate (code under
 * CUDA-graph capture, tight serving loops, harnesses that forbid cudaMalloc).  NO COUNTERPART IN THE
 * REFERENCE.  The look-back status words live in a
 * file-scope `static __device__` array of the library (256 KiB, held by every GPU that loads the
 * library until exit), so one call handles at most NFB_COMPACT_STATIC_MAX_SAMPLES samples; a larger
 * n returns NFB_ERR_INVALID_ARGUMENT.  Calls no cudaMalloc / cudaFree / allocating routine; the
 * array's device address is looked up with cudaGetSymbolAddress once per device, at the first call
 * made with that device current (make that call outside capture).  Every call shares the array and
 * clears the part it uses on `stream` first: calls that overlap on the GPU (different streams, or
 * graphs replayed concurrently) race — serialise them.  Outputs are identical to nfb_compact_samples. */
#define NFB_COMPACT_STATIC_MAX_SAMPLES (1ll << 26)
NFB_API int nfb_compact_samples_static(const uint8_t* mask, const int64_t* ray_indices,
                                       const float* t_starts, const float* t_ends, int64_t n,
                                       int64_t* out_ray_indices, float* out_t_starts,
                                       float* out_t_ends, int64_t* total, nfb_stream_t stream);

/* ------------------------------------------------------------------------------------
 * nerfacc.rendering() for packed rays in one forward and one backward kernel (a group of
 * `lanes_per_ray` lanes walks one ray; 0 = choose from the mean count; else 4, 8, 16 or 32).
 * Forward replaces render_weight_from_density/alpha + 3x accumulate_along_rays
 * (nerfacc/volrend.py:116-150): writes weights, trans (+ alphas_out in density mode) per sample
 * and colors [n_rays,3] = sum w*rgb, opacity [n_rays] = sum w, depth [n_rays] = sum w*(ts+te)/2
 * per ray (rows of empty rays are zero).  `packed_info` may be any non-overlapping chunks.
 * Backward is what autograd derives from that composition (see nfb_composite_bwd); it needs no
 * ray_indices.  Same results as nfb_render_weight_fwd + nfb_composite_fwd / nfb_composite_bwd
 * up to fp32 summation order. */
NFB_API int nfb_render_rays_fwd(int from_density, const float* alphas, const float* t_starts,
                                const float* t_ends, const float* sigmas, const float* rgbs,
                                const int64_t* packed_info, int64_t n_rays, int64_t n,
                                float* weights, float* trans, float* alphas_out, float* colors,
                                float* opacity, float* depth, int lanes_per_ray,
                                nfb_stream_t stream);
NFB_API int nfb_render_rays_bwd(int from_density, const float* g_colors, const float* g_opacity,
                                const float* g_depth, const float* g_weights, const float* g_trans,
                                const float* g_alphas, const float* alphas, const float* t_starts,
                                const float* t_ends, const float* sigmas, const float* rgbs,
                                const float* trans, const int64_t* packed_info, int64_t n_rays,
                                int64_t n, float* g_in, float* g_rgbs, int lanes_per_ray,
                                nfb_stream_t stream);

/* The same rendering() on warp tiles, for a DENSE canonical layout given as head bits
 * (nfb_head_bits) plus the API's ray_indices: samples are cut into 512-item tiles independent of
 * the rays (coalesced 128-bit accesses throughout, cost independent of ray lengths).  Forward:
 * one segmented scan for T / w and five segmented sums for the per-ray totals; the ray id of a
 * ray's last sample is a sparse ray_indices load.  colors / opacity / depth are cleared by the
 * call (rows of rays without samples are zero).  Backward: ray ids are propagated from the
 * heads inside the tile, the per-ray gradient rows are gathered through them; no dense
 * ray_indices read.  Replaces nerfacc/volrend.py:116-150 and its autograd, like
 * nfb_render_rays_fwd / _bwd.  Workspace: nfb_render_bits_workspace_bytes(n). */
NFB_API size_t nfb_render_bits_workspace_bytes(int64_t n);
NFB_API int nfb_render_bits_fwd(int from_density, const float* alphas, const float* t_starts,
                                const float* t_ends, const float* sigmas, const float* rgbs,
                                const int64_t* ray_indices, const uint32_t* head_bits, int64_t n_rays,
                                int64_t n, float* weights, float* trans, float* alphas_out,
                                float* colors, float* opacity, float* depth, void* workspace,
                                size_t workspace_bytes, uint32_t epoch, nfb_stream_t stream);
NFB_API int nfb_render_bits_bwd(int from_density, const float* g_colors, const float* g_opacity,
                                const float* g_depth, const float* g_weights, const float* g_trans,
                                const float* g_alphas, const float* alphas, const float* t_starts,
                                const float* t_ends, const float* sigmas, const float* rgbs,
                                const float* trans, const int64_t* ray_indices,
                                const uint32_t* head_bits, int64_t n, float* g_in, float* g_rgbs,
                                void* workspace, size_t workspace_bytes, uint32_t epoch,
                                nfb_stream_t stream);

/* accumulate_along_rays (nerfacc/volrend.py:483-547) on the same warp tiles, for a dense canonical
 * layout given as head bits + the API's ray_indices: out[r, :] (+)= sum over the samples of ray r
 * of weights[i] * values[i, :] (values == NULL: the sum of weights), dim in 1..4.  Without
 * add_into the call clears out first (rows of rays without samples are zero).
 * Workspace: nfb_render_bits_workspace_bytes(n). */
NFB_API int nfb_accumulate_bits_fwd(const float* weights, const float* values, int64_t dim,
                                    const int64_t* ray_indices, const uint32_t* head_bits,
                                    int64_t n_rays, int64_t n, float* out, int add_into,
                                    void* workspace, size_t workspace_bytes, uint32_t epoch,
                                    nfb_stream_t stream);

/* depth[r] = depth_sum[r] / max(opacity[r], FLT_EPSILON) in place (volrend.py:147-150), and the
 * gradient of that quotient: g_depth_sum = g_depth / den, g_opacity_sum = g_opacity (NULL = 0)
 * - g_depth * depth / den where opacity >= eps — one launch each instead of the reference's
 * elementwise torch ops and autograd nodes on [n_rays] tensors. */
NFB_API int nfb_depth_normalize_fwd(const float* opacity, float* depth, int64_t n_rays, nfb_stream_t stream);
NFB_API int nfb_depth_normalize_bwd(const float* g_opacity, const float* g_depth, const float* opacity,
                                    const float* depth, int64_t n_rays, float* g_opacity_sum,
                                    float* g_depth_sum, nfb_stream_t stream);

/* ------------------------------------------------------------------------------------
 * Inverse-CDF resampling and per-ray search (nerfacc/cuda/csrc/pdf.cu).
 */
/* importance_sampling: replaces csrc/pdf.cu:294-456 (the three host overloads) and their kernels
 * importance_sampling_kernel (pdf.cu:98-167) + compute_intervels_kernel (pdf.cu:169-241).
 * For each ray draws n_out sample centres u_k = cdf_first + (k + bias) * (cdf_last-cdf_first)/n_out
 * through the piecewise-linear inverse of (vals, cdfs) and, when out_edges != NULL, the n_out+1
 * interval edges (midpoints of neighbouring centres, clamped to the ray's first / last edge).
 *   input rows : in_info == NULL -> [n_rays, n_in] batched;  else packed chunks (start,count) int64
 *   output rows: n_out_per_ray == NULL -> [n_rays, n_out] centres, [n_rays, n_out+1] edges;
 *                else per-ray counts with packed chunks edge_info / sample_info ([n_rays,2] int64)
 *   bias       : per-ray offset in (0,1] (the reference's `1 - curand_uniform`, drawn by the
 *                caller), or NULL for the unstratified 0.5
 *   max_in / max_out: upper bounds of the per-ray counts (size shared-memory staging; 0 = unknown,
 *                search in global memory).
 *   out_centres: may be NULL when only the edges are wanted (batched rows, or rows that fit the
 *                shared-memory staging). */
NFB_API int nfb_importance_sampling(const float* vals, const float* cdfs, const int64_t* in_info,
                                    int64_t n_rays, int64_t n_in, int64_t max_in,
                                    const int64_t* n_out_per_ray, int64_t n_out, int64_t max_out,
                                    const int64_t* edge_info, const int64_t* sample_info,
                                    const float* bias, float* out_edges, float* out_centres,
                                    nfb_stream_t stream);
/* The first lines of every sigma_fn / rgb_sigma_fn in the reference's examples
 * (examples/utils.py:82-112), the callers' side of the callbacks:
 *   positions[i,:] = rays_o[ray_indices[i],:] + rays_d[ray_indices[i],:] * (t_starts[i] + t_ends[i]) / 2
 *   dirs[i,:]      = rays_d[ray_indices[i],:]                       (optional: dirs may be NULL)
 * in one pass, bit-identical to the torch expression.  ray_indices must lie in [0, n_rays). */
NFB_API int nfb_sample_positions(const float* rays_o, const float* rays_d, const int64_t* ray_indices,
                                 const float* t_starts, const float* t_ends, int64_t n, int64_t n_rays,
                                 float* positions, float* dirs, nfb_stream_t stream);

/* Batched nfb_importance_sampling that writes the interval edges twice: in the normalised
 * coordinate s (out_edges [n_rays, n_out+1], as above) and mapped to ray distance t
 * (out_t_edges, same shape) the way PropNetEstimator._transform_stot does
 * (nerfacc/estimators/prop_net.py:215-229), op for op:
 *   stot_mode 1 "uniform": t = s*c1 + (1-s)*c0      with c0 = t_min,   c1 = t_max
 *   stot_mode 2 "lindisp": t = 1 / (s*c1 + (1-s)*c0) with c0 = 1/t_min, c1 = 1/t_max (rounded to fp32). */
NFB_API int nfb_importance_sampling_stot(const float* vals, const float* cdfs, int64_t n_rays,
                                         int64_t n_in, int64_t n_out, const float* bias,
                                         float* out_edges, float* out_t_edges, int stot_mode, float c0,
                                         float c1, nfb_stream_t stream);

/* ------------------------------------------------------------------------------------
 * PropNetEstimator's batched-row glue (nerfacc/estimators/prop_net.py:96-114, 232-256).
 */
/* cdfs [n_rays, n+1] = 1 - cat([T, 0]) with T_i = exp(-sum_{j<i} sigmas_j * (t_vals_{j+1} - t_vals_j))
 * (render_transmittance_from_density on t_starts = t_vals[:, :-1], t_ends = t_vals[:, 1:],
 * nerfacc/volrend.py:209-264, then prop_net.py:112-114); 1 <= n <= 512. */
NFB_API int nfb_cdf_from_density_fwd(const float* t_vals, const float* sigmas, int64_t n_rays, int64_t n,
                                     float* cdfs, nfb_stream_t stream);
/* g_sigmas [n_rays, n] from g_cdfs [n_rays, n+1] (what autograd derives from the composite above). */
NFB_API int nfb_cdf_from_density_bwd(const float* t_vals, const float* sigmas, const float* g_cdfs,
                                     int64_t n_rays, int64_t n, float* g_sigmas, nfb_stream_t stream);
/* _pdf_loss for batched rows (prop_net.py:232-256): query edges/cdfs [n_rays, n_q], key edges/cdfs
 * [n_rays, n_k]; loss [n_rays, n_q-1] = clip(w - w_outer, 0)^2 / (w + eps), w = the query bin's mass,
 * w_outer = cdfs_key[ids_right of the bin's right edge] - cdfs_key[ids_left of its left edge]
 * (ids as nfb_searchsorted). */
NFB_API int nfb_pdf_loss_fwd(const float* q_vals, const float* q_cdfs, int64_t n_q, const float* k_vals,
                             const float* k_cdfs, int64_t n_k, int64_t n_rays, float eps, float* loss,
                             nfb_stream_t stream);
/* g_k_cdfs [n_rays, n_k]: gradient of the loss rows w.r.t. cdfs_key given g_loss [n_rays, n_q-1]
 * (the reference: gather backward = scatter_add with atomics; here every key edge sums its
 * contiguous range of query bins, deterministic). */
NFB_API int nfb_pdf_loss_bwd(const float* q_vals, const float* q_cdfs, int64_t n_q, const float* k_vals,
                             const float* k_cdfs, int64_t n_k, int64_t n_rays, float eps,
                             const float* g_loss, float* g_k_cdfs, nfb_stream_t stream);

/* searchsorted: replaces csrc/pdf.cu:245-286 + 458-503.  For every query value the pair of
 * neighbouring key positions (ids_left, ids_right) with key[left] <= v < key[right], clamped to
 * the row.  Batched query (q_info == NULL): ids are relative to the key row; packed query:
 * absolute positions in key.  q_info / k_info: packed chunks, or NULL for rows of n_q / n_k. */
NFB_API int nfb_searchsorted(const float* query, const int64_t* q_info, int64_t n_q,
                             const float* key, const int64_t* k_info, int64_t n_k, int64_t max_k,
                             int64_t n_rays, int64_t* ids_left, int64_t* ids_right,
                             nfb_stream_t stream);

#ifdef __cplusplus
}
#endif
#endif /* NERFACC_B200_H_ */
