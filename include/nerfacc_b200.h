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
 * pack_info: ray_indices [n] -> packed_info [n_rays, 2] (nerfacc/pack.py:43-46: histogram of
 * the indices, starts = exclusive cumsum).  `workspace` needs 8 bytes per 2048 rays.
 */
NFB_API int nfb_pack_info(const int64_t* ray_indices, int64_t n, int64_t n_rays,
                          int64_t* packed_info, void* workspace, size_t workspace_bytes,
                          nfb_stream_t stream);

#ifdef __cplusplus
}
#endif
#endif /* NERFACC_B200_H_ */
