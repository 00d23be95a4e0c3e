// scan.cu — packed segmented scans (sum / prod, inclusive / exclusive, forward / reverse)
// and the fused product-scan backward.
//
// Replaces csrc/scan.cu + include/utils_scan.cuh of the reference (one 16-thread row per
// ray, Blelloch sweep in shared memory, ~11 block barriers per 32 elements).  Two kernels:
//   * scan_flat_kernel  — canonical packed_info: flat 4096-element tiles chained with
//     decoupled look-back (seg_tile.cuh); bandwidth-bound for any ray-length mix.
//   * scan_rows_kernel  — any non-overlapping packed_info (gaps, arbitrary order) and the
//     `normalize` option: one warp per ray, shuffle scan over 32-element strips.
#include "warp_tile.cuh"

namespace nfb {

enum ScanMode { kPlain = 0, kProdBwd = 1 };

// MODE kPlain  : out = scan(in)
// MODE kProdBwd: out = scan_sum(in * aux_y) / max(aux_x, 1e-10)   (scan.cu:199-210, 289-300)
template <class Op, bool REV, bool EXCL, int MODE>
__global__ void __launch_bounds__(kThreads)
scan_flat_kernel(const float* __restrict__ in, const float* __restrict__ aux_y,
                 const float* __restrict__ aux_x, float* __restrict__ out,
                 const int64_t* __restrict__ packed_info, long long n_rays, long long n,
                 long long n_tiles, unsigned long long* status, uint32_t epoch, bool vec_ok) {
    __shared__ TileSmem sm;
    const TileGeom<REV> g(n, n_tiles);
    float v[kVec][4];
#pragma unroll
    for (int k = 0; k < kVec; ++k) g.load(in, vec_ok, k, Op::identity(), v[k]);
    if (MODE == kProdBwd) {
#pragma unroll
        for (int k = 0; k < kVec; ++k) {
            float y[4];
            g.load(aux_y, vec_ok, k, 0.0f, y);
#pragma unroll
            for (int c = 0; c < 4; ++c) v[k][c] = __fmul_rn(v[k][c], y[c]);
        }
    }
    const SegSpec seg{packed_info, n_rays, 0};
    scatter_heads<REV>(seg, g.base, n, sm.flags, sm.range);
    uint32_t f[kVec];
#pragma unroll
    for (int k = 0; k < kVec; ++k) f[k] = vec_flags(sm.flags, k);
    float E[kVec];
    tile_segscan<Op>(v, f, E, status, 0, epoch, sm);
#pragma unroll
    for (int k = 0; k < kVec; ++k) {
        float o[4];
        if (EXCL) finish_exclusive<Op>(v[k], f[k], E[k], o);
        else finish_inclusive<Op>(v[k], f[k], E[k], o);
        if (MODE == kProdBwd) {
            float x[4];
            g.load(aux_x, vec_ok, k, 1.0f, x);
#pragma unroll
            for (int c = 0; c < 4; ++c) o[c] = __fdiv_rn(o[c], fmaxf(x[c], 1e-10f));
        }
        g.store(out, vec_ok, k, o);
    }
}

// ---------------------------------------------------------------- scans from head bits
// Same results as scan_flat_kernel, on warp tiles (warp_tile.cuh).
template <class Op, bool REV, bool EXCL, int MODE>
__global__ void __launch_bounds__(kThreads, 6)
scan_bits_kernel(const float* __restrict__ in, const float* __restrict__ aux_y,
                 const float* __restrict__ aux_x, float* __restrict__ out,
                 const uint32_t* __restrict__ bits, long long n, long long n_wt,
                 unsigned long long* status, uint32_t epoch, bool vec_ok) {
    WarpTile<REV> wt;
    float v[kWtNV][4];
    if (!wt.init(bits, n, n_wt, vec_ok)) return;
#pragma unroll
    for (int k = 0; k < kWtNV; ++k) wt.load(in, k, Op::identity(), v[k]);
    if (MODE == kProdBwd) {
#pragma unroll
        for (int k = 0; k < kWtNV; ++k) {
            float y[4];
            wt.load(aux_y, k, 0.0f, y);
#pragma unroll
            for (int c = 0; c < 4; ++c) v[k][c] = __fmul_rn(v[k][c], y[c]);
        }
    }
    float E[kWtNV];
    wt.template scan<Op>(v, E, status, 1, 0, epoch);
#pragma unroll
    for (int k = 0; k < kWtNV; ++k) {
        if (k == kWtVec && !wt.over) break;
        float o[4];
        if (EXCL) finish_exclusive<Op>(v[k], wt.f[k], E[k], o);
        else finish_inclusive<Op>(v[k], wt.f[k], E[k], o);
        if (MODE == kProdBwd) {
            float x[4];
            wt.load(aux_x, k, 1.0f, x);
#pragma unroll
            for (int c = 0; c < 4; ++c) o[c] = __fdiv_rn(o[c], fmaxf(x[c], 1e-10f));
        }
        wt.store(out, k, o);
    }
}

// bit start[r] for every non-empty ray, and bit n
__global__ void __launch_bounds__(256)
head_bits_kernel(const int64_t* __restrict__ packed_info, long long n_rays, long long n, uint32_t* __restrict__ bits) {
    const long long stride = (long long)gridDim.x * blockDim.x;
    for (long long r = (long long)blockIdx.x * blockDim.x + threadIdx.x; r < n_rays; r += stride) {
        const RayChunk c = load_chunk(packed_info, r);
        if (c.cnt > 0 && c.start >= 0 && c.start < n) atomicOr(bits + (c.start >> 5), 1u << (unsigned)(c.start & 31));
    }
    if (blockIdx.x == 0 && threadIdx.x == 0) atomicOr(bits + (n >> 5), 1u << (unsigned)(n & 31));
}

template <class Op, bool EXCL, int MODE>
__global__ void __launch_bounds__(kThreads)
scan_rows_kernel(const float* __restrict__ in, const float* __restrict__ aux_y,
                 const float* __restrict__ aux_x, float* __restrict__ out,
                 const int64_t* __restrict__ packed_info, long long n_rays, bool reverse,
                 bool normalize) {
    const int lane = lane_id();
    const long long warps_total = (long long)gridDim.x * kWarps;
    for (long long r = (long long)blockIdx.x * kWarps + warp_id(); r < n_rays; r += warps_total) {
        const RayChunk ch = load_chunk(packed_info, r);
        if (ch.cnt <= 0) continue;
        float carry = Op::identity();
        for (long long j0 = 0; j0 < ch.cnt; j0 += 32) {
            const long long j = j0 + lane;
            const bool ok = j < ch.cnt;
            const long long i = reverse ? (ch.start + ch.cnt - 1 - j) : (ch.start + j);
            float x = Op::identity();
            if (ok) {
                x = in[i];
                if (MODE == kProdBwd) x = __fmul_rn(x, aux_y[i]);
            }
            float t = x;
#pragma unroll
            for (int d = 1; d < 32; d <<= 1) {
                const float y = __shfl_up_sync(kFull, t, d);
                if (lane >= d) t = Op::apply(y, t);
            }
            const float incl = Op::apply(carry, t);
            float res = incl;
            if (EXCL) {
                float e = __shfl_up_sync(kFull, incl, 1);
                if (lane == 0) e = carry;
                res = e;
            }
            if (ok) {
                if (MODE == kProdBwd) res = __fdiv_rn(res, fmaxf(aux_x[i], 1e-10f));
                out[i] = res;
            }
            carry = __shfl_sync(kFull, incl, 31);
        }
        if (normalize) {  // utils_scan.cuh:102-110, 229-237 (the first exclusive item stays `init`)
            __syncwarp();
            const float den = fmaxf(carry, 1e-10f);
            for (long long j = lane + (EXCL ? 1 : 0); j < ch.cnt; j += 32) {
                const long long i = reverse ? (ch.start + ch.cnt - 1 - j) : (ch.start + j);
                out[i] = __fdiv_rn(out[i], den);
            }
        }
    }
}

static int prepare_workspace(void* ws, size_t ws_bytes, long long n, uint32_t& epoch, cudaStream_t s) {
    const size_t need = nfb_workspace_bytes(n);
    if (ws == nullptr || ws_bytes < need) {
        set_error("workspace too small: need %zu bytes, got %zu", need, ws_bytes);
        return NFB_ERR_WORKSPACE;
    }
    if (epoch == 0) {
        if (cudaMemsetAsync(ws, 0, need, s) != cudaSuccess) {
            set_error("workspace memset failed: %s", cudaGetErrorString(cudaGetLastError()));
            return NFB_ERR_CUDA;
        }
        epoch = kEpochMemset;
    } else if (epoch >= kEpochMemset) {
        set_error("epoch must be in [1, 2^30-2]");
        return NFB_ERR_INVALID_ARGUMENT;
    }
    return NFB_OK;
}

template <int MODE>
static int launch_scan(const float* in, const float* aux_y, const float* aux_x, float* out,
                       const int64_t* packed_info, int64_t n_rays, int64_t n, int op,
                       int exclusive, int reverse, int normalize, int canonical, void* ws,
                       size_t ws_bytes, uint32_t epoch, cudaStream_t s) {
    NFB_CHECK_ARG(n >= 0 && n_rays >= 0, "negative size");
    NFB_CHECK_ARG(op == NFB_OP_SUM || op == NFB_OP_PROD, "unknown op %d", op);
    if (n == 0 || n_rays == 0) return NFB_OK;
    NFB_CHECK_ARG(in && out && packed_info, "null pointer");
    NFB_CHECK_ARG(aligned16(packed_info), "packed_info must be 16-byte aligned");
    const bool flat = canonical && !normalize;
    if (flat) {
        int rc = prepare_workspace(ws, ws_bytes, n, epoch, s);
        if (rc != NFB_OK) return rc;
        const long long nt = num_tiles(n);
        const bool vec_ok = aligned16(in) && aligned16(out) && (MODE == kPlain || (aligned16(aux_y) && aligned16(aux_x)));
        auto* st = static_cast<unsigned long long*>(ws);
#define NFB_LAUNCH_FLAT(OP, REV, EXCL) \
        scan_flat_kernel<OP, REV, EXCL, MODE><<<(unsigned)nt, kThreads, 0, s>>>(in, aux_y, aux_x, out, packed_info, n_rays, n, nt, st, epoch, vec_ok)
        if (op == NFB_OP_SUM) {
            if (reverse) { if (exclusive) NFB_LAUNCH_FLAT(SumOp, true, true); else NFB_LAUNCH_FLAT(SumOp, true, false); }
            else { if (exclusive) NFB_LAUNCH_FLAT(SumOp, false, true); else NFB_LAUNCH_FLAT(SumOp, false, false); }
        } else {
            if (reverse) { if (exclusive) NFB_LAUNCH_FLAT(ProdOp, true, true); else NFB_LAUNCH_FLAT(ProdOp, true, false); }
            else { if (exclusive) NFB_LAUNCH_FLAT(ProdOp, false, true); else NFB_LAUNCH_FLAT(ProdOp, false, false); }
        }
#undef NFB_LAUNCH_FLAT
    } else {
        long long blocks = ceil_div<long long>(n_rays, kWarps);
        const long long cap = (long long)sm_count() * 8 * 4;
        if (blocks > cap) blocks = cap;
#define NFB_LAUNCH_ROWS(OP, EXCL) \
        scan_rows_kernel<OP, EXCL, MODE><<<(unsigned)blocks, kThreads, 0, s>>>(in, aux_y, aux_x, out, packed_info, n_rays, reverse != 0, normalize != 0)
        if (op == NFB_OP_SUM) { if (exclusive) NFB_LAUNCH_ROWS(SumOp, true); else NFB_LAUNCH_ROWS(SumOp, false); }
        else { if (exclusive) NFB_LAUNCH_ROWS(ProdOp, true); else NFB_LAUNCH_ROWS(ProdOp, false); }
#undef NFB_LAUNCH_ROWS
    }
    count_launch();
    NFB_CHECK_LAUNCH("scan kernel");
    return NFB_OK;
}

template <int MODE>
static int launch_scan_bits(const float* in, const float* aux_y, const float* aux_x, float* out,
                            const uint32_t* bits, int64_t n, int op, int exclusive, int reverse,
                            void* ws, size_t ws_bytes, uint32_t epoch, cudaStream_t s) {
    NFB_CHECK_ARG(n >= 0, "negative size");
    NFB_CHECK_ARG(op == NFB_OP_SUM || op == NFB_OP_PROD, "unknown op %d", op);
    if (n == 0) return NFB_OK;
    NFB_CHECK_ARG(in && out && bits, "null pointer");
    int rc = prepare_workspace(ws, ws_bytes, n, epoch, s);
    if (rc != NFB_OK) return rc;
    const long long n_wt = num_warp_tiles(n);
    const long long nt = ceil_div<long long>(n_wt, kWarps);
    const bool vec_ok = aligned16(in) && aligned16(out) && (MODE == kPlain || (aligned16(aux_y) && aligned16(aux_x)));
    auto* st = static_cast<unsigned long long*>(ws);
#define NFB_LAUNCH_BITS(OP, REV, EXCL) \
    scan_bits_kernel<OP, REV, EXCL, MODE><<<(unsigned)nt, kThreads, 0, s>>>(in, aux_y, aux_x, out, bits, n, n_wt, st, epoch, vec_ok)
    if (op == NFB_OP_SUM) {
        if (reverse) { if (exclusive) NFB_LAUNCH_BITS(SumOp, true, true); else NFB_LAUNCH_BITS(SumOp, true, false); }
        else { if (exclusive) NFB_LAUNCH_BITS(SumOp, false, true); else NFB_LAUNCH_BITS(SumOp, false, false); }
    } else {
        if (reverse) { if (exclusive) NFB_LAUNCH_BITS(ProdOp, true, true); else NFB_LAUNCH_BITS(ProdOp, true, false); }
        else { if (exclusive) NFB_LAUNCH_BITS(ProdOp, false, true); else NFB_LAUNCH_BITS(ProdOp, false, false); }
    }
#undef NFB_LAUNCH_BITS
    count_launch();
    NFB_CHECK_LAUNCH("scan_bits kernel");
    return NFB_OK;
}

}  // namespace nfb

extern "C" {

NFB_API size_t nfb_head_bits_bytes(int64_t n_elements) {
    return nfb::head_bits_words(n_elements > 0 ? n_elements : 0) * sizeof(uint32_t);
}

NFB_API int nfb_head_bits(const int64_t* packed_info, int64_t n_rays, int64_t n, uint32_t* bits,
                          nfb_stream_t stream) {
    using namespace nfb;
    NFB_CHECK_ARG(n >= 0 && n_rays >= 0, "negative size");
    NFB_CHECK_ARG(bits != nullptr && (packed_info != nullptr || n_rays == 0), "null pointer");
    NFB_CHECK_ARG(n_rays == 0 || aligned16(packed_info), "packed_info must be 16-byte aligned");
    auto s = static_cast<cudaStream_t>(stream);
    if (cudaMemsetAsync(bits, 0, nfb_head_bits_bytes(n), s) != cudaSuccess) {
        set_error("head bits memset failed: %s", cudaGetErrorString(cudaGetLastError()));
        return NFB_ERR_CUDA;
    }
    long long blocks = ceil_div<long long>(n_rays > 0 ? n_rays : 1, 256);
    const long long cap = (long long)sm_count() * 8;
    if (blocks > cap) blocks = cap;
    head_bits_kernel<<<(unsigned)blocks, 256, 0, s>>>(packed_info, n_rays, n, bits);
    count_launch();
    NFB_CHECK_LAUNCH("head_bits kernel");
    return NFB_OK;
}

NFB_API int nfb_scan_bits_f32(const float* in, float* out, const uint32_t* head_bits, int64_t n, int op,
                              int exclusive, int reverse, void* workspace, size_t workspace_bytes,
                              uint32_t epoch, nfb_stream_t stream) {
    return nfb::launch_scan_bits<nfb::kPlain>(in, nullptr, nullptr, out, head_bits, n, op, exclusive,
                                              reverse, workspace, workspace_bytes, epoch,
                                              static_cast<cudaStream_t>(stream));
}

NFB_API int nfb_prod_scan_bwd_bits_f32(const float* grad_out, const float* in, const float* out,
                                       float* grad_in, const uint32_t* head_bits, int64_t n,
                                       int exclusive, void* workspace, size_t workspace_bytes,
                                       uint32_t epoch, nfb_stream_t stream) {
    if (n > 0 && (in == nullptr || out == nullptr)) {
        nfb::set_error("null pointer");
        return NFB_ERR_INVALID_ARGUMENT;
    }
    return nfb::launch_scan_bits<nfb::kProdBwd>(grad_out, out, in, grad_in, head_bits, n, NFB_OP_SUM,
                                                exclusive, /*reverse=*/1, workspace, workspace_bytes,
                                                epoch, static_cast<cudaStream_t>(stream));
}


NFB_API size_t nfb_workspace_bytes(int64_t n_elements) {
    const long long nt = nfb::num_tiles(n_elements > 0 ? n_elements : 1);
    return (size_t)nt * nfb::kStatusPerTile * sizeof(unsigned long long) + 256;
}

NFB_API int nfb_scan_f32(const float* in, float* out, const int64_t* packed_info, int64_t n_rays,
                         int64_t n, int op, int exclusive, int reverse, int normalize,
                         int canonical, void* workspace, size_t workspace_bytes, uint32_t epoch,
                         nfb_stream_t stream) {
    return nfb::launch_scan<nfb::kPlain>(in, nullptr, nullptr, out, packed_info, n_rays, n, op,
                                         exclusive, reverse, normalize, canonical, workspace,
                                         workspace_bytes, epoch, static_cast<cudaStream_t>(stream));
}

NFB_API int nfb_prod_scan_bwd_f32(const float* grad_out, const float* in, const float* out,
                                  float* grad_in, const int64_t* packed_info, int64_t n_rays,
                                  int64_t n, int exclusive, int canonical, void* workspace,
                                  size_t workspace_bytes, uint32_t epoch, nfb_stream_t stream) {
    if (n > 0 && n_rays > 0 && (in == nullptr || out == nullptr)) {
        nfb::set_error("null pointer");
        return NFB_ERR_INVALID_ARGUMENT;
    }
    return nfb::launch_scan<nfb::kProdBwd>(grad_out, out, in, grad_in, packed_info, n_rays, n,
                                           NFB_OP_SUM, exclusive, /*reverse=*/1, /*normalize=*/0,
                                           canonical, workspace, workspace_bytes, epoch,
                                           static_cast<cudaStream_t>(stream));
}

}  // extern "C"
