// scan.cu — packed segmented scans (sum / prod, inclusive / exclusive, forward / reverse)
// and the fused product-scan backward.
//
// Replaces csrc/scan.cu + include/utils_scan.cuh of the reference (one 16-thread row per
// ray, Blelloch sweep in shared memory, ~11 block barriers per 32 elements).  Two kernels:
//   * scan_flat_kernel  — canonical packed_info: flat 4096-element tiles chained with
//     decoupled look-back (seg_tile.cuh); bandwidth-bound for any ray-length mix.
//   * scan_rows_kernel  — any non-overlapping packed_info (gaps, arbitrary order) and the
//     `normalize` option: one warp per ray, shuffle scan over 32-element strips.
#include "seg_tile.cuh"

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
// Warp tiles: every warp owns VEC*128 consecutive items and is on its own — no shared memory,
// no block barrier, no packed_info search (ray boundaries come as 1 bit per item, nfb_head_bits).
//
// A tile that does not start on a ray boundary needs the running value of the ray it continues.
// Waiting for the previous warp to publish it (look-back) puts one L2 round trip under load on
// every warp's critical path and halves the bytes in flight.  Instead the previous warp finishes
// that ray itself: every warp also loads the 128 items after its tile (the "overhang") and, if
// the next tile's first boundary lies among them, writes the items before that boundary — both
// warps read the same bits, so they agree on who writes what without talking.  Only tiles whose
// first boundary is further than 128 items away (rays of hundreds of samples) look back.
template <bool REV>
__device__ __forceinline__ void load_vec(const float* __restrict__ p, long long g, long long n, bool vec_ok,
                                         bool slow, float fill, float (&v)[4]) {
    const bool whole = vec_ok && g >= 0 && g + 3 < n;
#pragma unroll
    for (int c = 0; c < 4; ++c) v[c] = fill;
    if (REV) ldg_stream4_if(whole, p + g, v[3], v[2], v[1], v[0]);
    else ldg_stream4_if(whole, p + g, v[0], v[1], v[2], v[3]);
    if (slow) {
#pragma unroll
        for (int c = 0; c < 4; ++c) {
            const long long gi = g + (REV ? 3 - c : c);
            ldg_stream_if(!whole && gi >= 0 && gi < n, p + gi, v[c]);
        }
    }
}
// mask: bit c set <=> logical component c is written
template <bool REV>
__device__ __forceinline__ void store_vec(float* __restrict__ p, long long g, long long n, bool vec_ok,
                                          bool slow, const float (&v)[4], uint32_t mask) {
    const bool whole = vec_ok && g >= 0 && g + 3 < n && mask == 0xfu;
    if (REV) stg_stream4_if(whole, p + g, v[3], v[2], v[1], v[0]);
    else stg_stream4_if(whole, p + g, v[0], v[1], v[2], v[3]);
    if (slow || (mask != 0xfu && mask != 0u)) {
#pragma unroll
        for (int c = 0; c < 4; ++c) {
            const long long gi = g + (REV ? 3 - c : c);
            stg_stream_if(!whole && ((mask >> c) & 1u) && gi >= 0 && gi < n, p + gi, v[c]);
        }
    }
}
// position of the first head among a warp's 128 items (4 per lane, nibble flags); 128 = none
__device__ __forceinline__ int first_head_128(uint32_t nib) {
    const unsigned hm = __ballot_sync(kFull, nib != 0u);
    if (hm == 0u) return 128;
    const int l = __ffs(hm) - 1;
    return 4 * l + __ffs(__shfl_sync(kFull, nib, l)) - 1;
}

template <class Op, bool REV, bool EXCL, int MODE, int VEC>
__global__ void __launch_bounds__(kThreads, VEC == 4 ? 4 : 3)
scan_bits_kernel(const float* __restrict__ in, const float* __restrict__ aux_y,
                 const float* __restrict__ aux_x, float* __restrict__ out,
                 const uint32_t* __restrict__ bits, long long n, long long n_wt,
                 unsigned long long* status, uint32_t epoch, bool vec_ok) {
    constexpr int W = VEC * 128;
    const int lane = lane_id();
    const long long t = (long long)blockIdx.x * kWarps + warp_id();  // logical order == launch order
    if (t >= n_wt) return;
    const long long pt = REV ? n_wt - 1 - t : t;
    const long long base = pt * W;
    const bool slow = !vec_ok || base + W + 128 > n;
    const bool has_next = REV ? (pt > 0) : (base + W < n);
    const long long g0 = REV ? base + (long long)(VEC - 1) * 128 + (31 - lane) * 4 : base + lane * 4;
    const long long gstep = REV ? -128 : 128;  // physical start of logical vector k: g0 + k * gstep
    const long long go = has_next ? (REV ? base - 128 + (31 - lane) * 4 : base + W + lane * 4) : -8;

    float v[VEC][4], u[4];
    uint32_t f[VEC];
#pragma unroll
    for (int k = 0; k < VEC; ++k) load_vec<REV>(in, g0 + k * gstep, n, vec_ok, slow, Op::identity(), v[k]);
    load_vec<REV>(in, go, n, vec_ok, slow, Op::identity(), u);
    uint32_t nib0 = 0u;
#pragma unroll
    for (int k = 0; k < VEC; ++k) {
        const uint32_t nb = head_nibble<REV>(bits, g0 + k * gstep, n);
        if (k == 0) nib0 = nb;
        f[k] = nibble_to_flags(nb);
    }
    const uint32_t nibo = head_nibble<REV>(bits, go, n);
    if (MODE == kProdBwd) {
#pragma unroll
        for (int k = 0; k < VEC; ++k) {
            float y[4];
            load_vec<REV>(aux_y, g0 + k * gstep, n, vec_ok, slow, 0.0f, y);
#pragma unroll
            for (int c = 0; c < 4; ++c) v[k][c] = __fmul_rn(v[k][c], y[c]);
        }
        float y[4];
        load_vec<REV>(aux_y, go, n, vec_ok, slow, 0.0f, y);
#pragma unroll
        for (int c = 0; c < 4; ++c) u[c] = __fmul_rn(u[c], y[c]);
    }
    float E[VEC], wc;
    bool open_vec[VEC], warp_open;
    warp_segscan<Op, VEC>(v, f, E, open_vec, wc, warp_open);

    const int p_own = first_head_128(nib0);
    const bool by_pred = t > 0 && p_own > 0 && p_own < 128;  // my first items are the previous warp's overhang
    const bool need_x = t > 0 && p_own == 128;
    const bool final_now = (t == 0) || !warp_open;
    if (lane == 0) st_relaxed_u64(status + t, pack_status(wc, epoch, final_now ? 2u : 1u));
    float X = Op::identity();
    if (need_x) {
        X = warp_lookback_carry<Op>(status, t, epoch);
        if (!final_now && lane == 0) st_relaxed_u64(status + t, pack_status(Op::apply(X, wc), epoch, 2u));
    }
#pragma unroll
    for (int k = 0; k < VEC; ++k) {
        const float Ek = open_vec[k] ? Op::apply(X, E[k]) : E[k];
        float o[4];
        if (EXCL) finish_exclusive<Op>(v[k], f[k], Ek, o);
        else finish_inclusive<Op>(v[k], f[k], Ek, o);
        if (MODE == kProdBwd) {
            float x[4];
            load_vec<REV>(aux_x, g0 + k * gstep, n, vec_ok, slow, 1.0f, x);
#pragma unroll
            for (int c = 0; c < 4; ++c) o[c] = __fdiv_rn(o[c], fmaxf(x[c], 1e-10f));
        }
        uint32_t mask = 0xfu;
        if (k == 0 && by_pred) {
            const int d = p_own - 4 * lane;  // components below d belong to the previous warp
            mask = d <= 0 ? 0xfu : (d >= 4 ? 0u : (0xfu << d) & 0xfu);
        }
        store_vec<REV>(out, g0 + k * gstep, n, vec_ok, slow, o, mask);
    }
    if (!has_next) return;
    const int p_nx = first_head_128(nibo);
    if (p_nx == 0 || p_nx == 128) return;
    // finish the ray that runs over the end of my tile: items [0, p_nx) of the next tile
    const float carry = warp_open ? Op::apply(X, wc) : wc;  // inclusive value at my last item
#pragma unroll
    for (int c = 1; c < 4; ++c) u[c] = Op::apply(u[c - 1], u[c]);
    float tt = u[3];
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
        const float y = __shfl_up_sync(kFull, tt, d);
        if (lane >= d) tt = Op::apply(y, tt);
    }
    float e = __shfl_up_sync(kFull, tt, 1);
    e = (lane == 0) ? carry : Op::apply(carry, e);
    float o[4];
#pragma unroll
    for (int c = 0; c < 4; ++c) o[c] = EXCL ? (c == 0 ? e : Op::apply(e, u[c - 1])) : Op::apply(e, u[c]);
    if (MODE == kProdBwd) {
        float x[4];
        load_vec<REV>(aux_x, go, n, vec_ok, slow, 1.0f, x);
#pragma unroll
        for (int c = 0; c < 4; ++c) o[c] = __fdiv_rn(o[c], fmaxf(x[c], 1e-10f));
    }
    const int d = p_nx - 4 * lane;  // components below d are mine
    const uint32_t mask = d <= 0 ? 0u : (d >= 4 ? 0xfu : ~(0xfu << d) & 0xfu);
    store_vec<REV>(out, go, n, vec_ok, slow, o, mask);
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
    constexpr int kBitsVec = 4;
    const long long n_wt = ceil_div<long long>(n, kBitsVec * 128);
    const long long nt = ceil_div<long long>(n_wt, kWarps);
    const bool vec_ok = aligned16(in) && aligned16(out) && (MODE == kPlain || (aligned16(aux_y) && aligned16(aux_x)));
    auto* st = static_cast<unsigned long long*>(ws);
#define NFB_LAUNCH_BITS(OP, REV, EXCL) \
    scan_bits_kernel<OP, REV, EXCL, MODE, kBitsVec><<<(unsigned)nt, kThreads, 0, s>>>(in, aux_y, aux_x, out, bits, n, n_wt, st, epoch, vec_ok)
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
