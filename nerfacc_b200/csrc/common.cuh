// common.cuh — shared device/host helpers for libnerfacc_b200 (sm_100a only).
//
// Nothing here depends on libtorch: the library is plain CUDA behind a C ABI
// (include/nerfacc_b200.h).  Memory is owned by the caller; kernels launch on the
// caller's stream and never synchronise.
#pragma once

#include <cuda_runtime.h>
#include <stdint.h>

#include "../../include/nerfacc_b200.h"

#ifndef __CUDA_ARCH__
#define NFB_ARCH_OK 1
#elif __CUDA_ARCH__ >= 1000
#define NFB_ARCH_OK 1
#else
#error "libnerfacc_b200 is written for sm_100a (B200) only"
#endif

namespace nfb {

// ---------------------------------------------------------------- host side
void set_error(const char* fmt, ...);
void count_launch(int n = 1);

#define NFB_CHECK_ARG(cond, ...)            \
    do {                                    \
        if (!(cond)) {                      \
            ::nfb::set_error(__VA_ARGS__);  \
            return NFB_ERR_INVALID_ARGUMENT; \
        }                                   \
    } while (0)

#define NFB_CHECK_LAUNCH(what)                                                   \
    do {                                                                         \
        cudaError_t e__ = cudaGetLastError();                                    \
        if (e__ != cudaSuccess) {                                                \
            ::nfb::set_error("%s: %s", what, cudaGetErrorString(e__));           \
            return NFB_ERR_CUDA;                                                 \
        }                                                                        \
    } while (0)

int sm_count();

static inline bool aligned16(const void* p) { return (reinterpret_cast<uintptr_t>(p) & 15u) == 0; }

template <typename T>
static inline T ceil_div(T a, T b) { return (a + b - 1) / b; }

// ---------------------------------------------------------------- device side
constexpr int kWarp = 32;
constexpr unsigned kFull = 0xffffffffu;

__device__ __forceinline__ int lane_id() { return threadIdx.x & 31; }
__device__ __forceinline__ int warp_id() { return threadIdx.x >> 5; }

// Streaming 128-bit / 32-bit accesses: every array on this path is touched once
// per kernel, so keep it out of L1 (ld.global.nc.L1::no_allocate) and write with
// st.global.L1::no_allocate; L2 (126 MB) stays the only cache level in play.
__device__ __forceinline__ float4 ldg_stream4(const float* p) {
    float4 r;
    asm volatile("ld.global.nc.L1::no_allocate.v4.f32 {%0,%1,%2,%3}, [%4];"
                 : "=f"(r.x), "=f"(r.y), "=f"(r.z), "=f"(r.w) : "l"(p));
    return r;
}
__device__ __forceinline__ float ldg_stream(const float* p) {
    float r;
    asm volatile("ld.global.nc.L1::no_allocate.f32 %0, [%1];" : "=f"(r) : "l"(p));
    return r;
}
__device__ __forceinline__ void stg_stream4(float* p, float4 v) {
    asm volatile("st.global.L1::no_allocate.v4.f32 [%0], {%1,%2,%3,%4};"
                 :: "l"(p), "f"(v.x), "f"(v.y), "f"(v.z), "f"(v.w) : "memory");
}
__device__ __forceinline__ void stg_stream(float* p, float v) {
    asm volatile("st.global.L1::no_allocate.f32 [%0], %1;" :: "l"(p), "f"(v) : "memory");
}

// Predicated forms: the access happens only when `on`; the registers keep their value otherwise.
// The predicate lives inside the asm so that a tile-edge test never becomes a branch: a branch
// around a load makes ptxas merge the result through a temporary, and the MOV that follows the
// LDG then waits for the data — four dependent DRAM round trips per thread instead of four
// loads in flight.
__device__ __forceinline__ void ldg_stream4_if(bool on, const float* p, float& a, float& b, float& c, float& d) {
    asm volatile("{\n\t.reg .pred q;\n\tsetp.ne.s32 q, %5, 0;\n\t"
                 "@q ld.global.nc.L1::no_allocate.v4.f32 {%0,%1,%2,%3}, [%4];\n\t}"
                 : "+f"(a), "+f"(b), "+f"(c), "+f"(d) : "l"(p), "r"((int)on));
}
__device__ __forceinline__ void ldg_stream_if(bool on, const float* p, float& a) {
    asm volatile("{\n\t.reg .pred q;\n\tsetp.ne.s32 q, %2, 0;\n\t"
                 "@q ld.global.nc.L1::no_allocate.f32 %0, [%1];\n\t}" : "+f"(a) : "l"(p), "r"((int)on));
}
__device__ __forceinline__ void stg_stream4_if(bool on, float* p, float a, float b, float c, float d) {
    asm volatile("{\n\t.reg .pred q;\n\tsetp.ne.s32 q, %5, 0;\n\t"
                 "@q st.global.L1::no_allocate.v4.f32 [%0], {%1,%2,%3,%4};\n\t}"
                 :: "l"(p), "f"(a), "f"(b), "f"(c), "f"(d), "r"((int)on) : "memory");
}
__device__ __forceinline__ void stg_stream_if(bool on, float* p, float a) {
    asm volatile("{\n\t.reg .pred q;\n\tsetp.ne.s32 q, %2, 0;\n\t"
                 "@q st.global.L1::no_allocate.f32 [%0], %1;\n\t}" :: "l"(p), "f"(a), "r"((int)on) : "memory");
}

// 64-bit relaxed device-scope accesses for look-back status words.
__device__ __forceinline__ void st_relaxed_u64(unsigned long long* p, unsigned long long v) {
    asm volatile("st.relaxed.gpu.global.u64 [%0], %1;" :: "l"(p), "l"(v) : "memory");
}
__device__ __forceinline__ unsigned long long ld_relaxed_u64(const unsigned long long* p) {
    unsigned long long v;
    asm volatile("ld.relaxed.gpu.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
    return v;
}

// Decoupled look-back by one warp (all 32 lanes call it, tile > 0, the tile's own status word
// already published as "aggregate"): the sum of everything before `tile`.  A status word is
// value << 2 | state (0 not there yet, 1 the tile's aggregate, 2 its inclusive prefix); the array
// starts zeroed.  The warp reads 32 predecessors per round, so a launch whose tiles are all resident
// at once (a scan over rays: a few hundred tiles) resolves in a handful of dependent loads instead
// of one per predecessor.
__device__ __forceinline__ long long warp_lookback_u64(const unsigned long long* status, long long tile) {
    const int lane = threadIdx.x & 31;
    long long X = 0;
    long long k = tile - 1;  // lane 0 looks at the nearest predecessor
    while (true) {
        const long long idx = k - lane;
        // (before tile 0 there is nothing: an inclusive prefix of zero)
        const unsigned long long s = idx >= 0 ? ld_relaxed_u64(status + idx) : 2ull;
        const unsigned incl = __ballot_sync(0xffffffffu, (s & 3ull) == 2ull);
        const unsigned ready = __ballot_sync(0xffffffffu, (s & 3ull) != 0ull);
        // the lanes up to and including the first inclusive prefix must all be there
        const unsigned need = incl ? (0xffffffffu >> (32 - __ffs(incl))) : 0xffffffffu;
        if ((ready & need) != need) { __nanosleep(20); continue; }
        long long v = ((need >> lane) & 1u) ? (long long)(s >> 2) : 0;
#pragma unroll
        for (int d = 16; d > 0; d >>= 1) v += __shfl_xor_sync(0xffffffffu, v, d);
        X += v;
        if (incl) return X;
        k -= 32;
    }
}

// packed_info is int64 [n_rays, 2] = (start, count) interleaved: one 16-byte load.
struct RayChunk { long long start, cnt; };
__device__ __forceinline__ RayChunk load_chunk(const int64_t* packed_info, long long r) {
    const longlong2 v = __ldg(reinterpret_cast<const longlong2*>(packed_info) + r);
    return RayChunk{v.x, v.y};
}

// Warp-cooperative 32-ary search: first index r in [0, n) with pred(r) true
// (pred must be monotone false..true), n if none.  All 32 lanes must call it with
// the same arguments; every lane returns the same answer.  log32(n) dependent
// probes instead of log2(n): 4 rounds for 2^20 rays.
template <typename Pred>
__device__ __forceinline__ long long warp_search_first(long long n, Pred pred) {
    long long lo = 0, hi = n;  // answer in [lo, hi]
    const int lane = lane_id();
    while (hi - lo > 0) {
        const long long len = hi - lo;
        if (len <= 32) {
            const bool p = (lane < len) ? pred(lo + lane) : true;
            const unsigned m = __ballot_sync(kFull, p);
            if (m == 0u) return hi;
            const int f = __ffs(m) - 1;
            return (f < len) ? lo + f : hi;
        }
        const long long stride = (len + 31) >> 5;
        long long probe = lo + (long long)(lane + 1) * stride - 1;
        if (probe > hi - 1) probe = hi - 1;
        const bool p = pred(probe);
        const unsigned m = __ballot_sync(kFull, p);
        if (m == 0) return hi;  // even the last probe (hi-1) is false
        const int f = __ffs(m) - 1;
        long long nlo = lo + (long long)f * stride;
        long long nhi = lo + (long long)(f + 1) * stride - 1;
        if (nhi > hi - 1) nhi = hi - 1;
        // pred(nhi) is true, pred(nlo-1) is false: answer in [nlo, nhi]
        lo = nlo;
        hi = nhi;
        if (lo == hi) return lo;
    }
    return lo;
}

// ---- 1-D bulk copies (TMA) for the tile's input span: one instruction per array instead of a
// block-wide loop of 128-bit loads and stores; completion is counted in bytes on an mbarrier.
__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint64_t* bar, int count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" :: "r"(smem_u32(bar)), "r"(count));
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" :: "r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void bulk_g2s(void* dst, const void* src, uint32_t bytes, uint64_t* bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                 :: "r"(smem_u32(dst)), "l"(src), "r"(bytes), "r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
    asm volatile("{\n\t.reg .pred p;\n\tWAIT_%=:\n\t"
                 "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n\t"
                 "@!p bra WAIT_%=;\n\t}" :: "r"(smem_u32(bar)), "r"(parity) : "memory");
}

// q = upper_bound(c[0..n), u) (pdf.cu:44-63: the count of leading entries <= u) for a sorted row in
// shared memory, as a ladder of power-of-two steps without bounds checks: `top` is the largest power
// of two <= n (ltop its log2).  If c[top-1] <= u the answer lies in [top, n]: continue from
// base = n + 1 - top (<= top, so c[base-1] <= u as well) where the remaining steps top/2 .. 1 reach
// exactly up to n; else it lies in [0, top-1] and base = 0.  Every probe c[base + st - 1] is inside the
// row.  Three instructions per step: a load at a constant offset, a compare, a conditional add.
// The probes go through a 32-bit shared-memory address (a pointer would drag its 64-bit generic twin
// through every step).
__device__ __forceinline__ float lds_f32(uint32_t addr) {
    float v;
    asm volatile("ld.shared.f32 %0, [%1];" : "=f"(v) : "r"(addr));
    return v;
}
__device__ __forceinline__ int ladder_upper_bound(const float* c, int n, int top, int ltop, float u) {
    const uint32_t c0 = smem_u32(c);
    uint32_t q = !(lds_f32(c0 + 4u * (top - 1)) > u) ? c0 + 4u * (n + 1 - top) : c0;
#define NFB_LADDER(st) { const float v_ = lds_f32(q + 4u * ((st) - 1)); if (!(v_ > u)) q += 4u * (st); }
    switch (ltop) {
        case 19: NFB_LADDER(1 << 18) case 18: NFB_LADDER(1 << 17) case 17: NFB_LADDER(1 << 16)
        case 16: NFB_LADDER(1 << 15) case 15: NFB_LADDER(1 << 14) case 14: NFB_LADDER(1 << 13)
        case 13: NFB_LADDER(1 << 12) case 12: NFB_LADDER(1 << 11) case 11: NFB_LADDER(1 << 10)
        case 10: NFB_LADDER(1 << 9) case 9: NFB_LADDER(1 << 8) case 8: NFB_LADDER(1 << 7)
        case 7: NFB_LADDER(1 << 6) case 6: NFB_LADDER(1 << 5) case 5: NFB_LADDER(1 << 4)
        case 4: NFB_LADDER(1 << 3) case 3: NFB_LADDER(1 << 2) case 2: NFB_LADDER(1 << 1)
        case 1: NFB_LADDER(1)
        default: break;
    }
#undef NFB_LADDER
    return (int)((q - c0) >> 2);
}


}  // namespace nfb
