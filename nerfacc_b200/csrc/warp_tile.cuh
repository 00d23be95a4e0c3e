// warp_tile.cuh — segmented scans over a dense packed layout from head bits, one warp per tile.
//
// Every warp owns kWtVec*128 = 512 consecutive items and is on its own: no shared memory, no
// block barrier, no packed_info search (ray boundaries come as 1 bit per item, nfb_head_bits).
//
// A tile that does not start on a ray boundary needs the running value of the ray it continues.
// Waiting for the previous warp to publish it (look-back) puts one L2 round trip under load on
// every warp's critical path and halves the bytes in flight (measured: 113 us vs 58 us for a
// 36.4M-sample scan).  Instead the previous warp finishes that ray itself: every warp also loads
// the 128 items after its tile (the "overhang", vector index kWtVec) and, if the next tile's
// first boundary lies among them, produces the items before that boundary — both warps read the
// same bits, so they agree on who writes what without talking.  Only tiles whose first boundary
// is 128 or more items away (rays of hundreds of samples) look back, warp to warp.
//
// Logical order is the scan direction: REV walks the same physical tiles from the end of the
// array with (vector, lane, component) mirrored, so every access stays a coalesced, 16-byte
// aligned 512-byte warp transaction.
#pragma once

#include "seg_tile.cuh"

namespace nfb {

constexpr int kWtVec = 4;
constexpr int kWtItems = kWtVec * 128;
constexpr int kWtNV = kWtVec + 1;  // tile vectors + the overhang vector

static inline long long num_warp_tiles(long long n) { return (n + kWtItems - 1) / kWtItems; }

// position of the first head among a warp's 128 items (4 per lane, nibble flags); 128 = none
__device__ __forceinline__ int first_head_128(uint32_t nib) {
    const unsigned hm = __ballot_sync(kFull, nib != 0u);
    if (hm == 0u) return 128;
    const int l = __ffs(hm) - 1;
    return 4 * l + __ffs(__shfl_sync(kFull, nib, l)) - 1;
}

template <bool REV>
struct WarpTile {
    long long n, t, g0, gstep, go;
    int lane, p_own, p_nx;
    bool vec_ok, slow, has_next;
    bool by_pred;  // my first p_own items are written by the previous warp
    bool need_x;   // the ray I continue started too far back: look back for its running value
    bool over;     // I write items [0, p_nx) of the next tile
    uint32_t f[kWtNV];  // head flags, one byte per logical component (flag_at); f[kWtVec] == 0

    // false: this warp has no tile (past the end of the array)
    __device__ __forceinline__ bool init(const uint32_t* __restrict__ bits, long long n_, long long n_wt,
                                         bool vec_ok_) {
        n = n_;
        vec_ok = vec_ok_;
        lane = lane_id();
        t = (long long)blockIdx.x * (blockDim.x >> 5) + warp_id();  // logical order == launch order
        if (t >= n_wt) return false;
        const long long pt = REV ? n_wt - 1 - t : t;
        const long long base = pt * kWtItems;
        slow = !vec_ok || base + kWtItems + 128 > n;
        has_next = REV ? (pt > 0) : (base + kWtItems < n);
        g0 = REV ? base + (long long)(kWtVec - 1) * 128 + (31 - lane) * 4 : base + lane * 4;
        gstep = REV ? -128 : 128;
        go = has_next ? (REV ? base - 128 + (31 - lane) * 4 : base + kWtItems + lane * 4) : -8;
        uint32_t nib0 = 0u;
#pragma unroll
        for (int k = 0; k < kWtVec; ++k) {
            const uint32_t nb = head_nibble<REV>(bits, g0 + k * gstep, n);
            if (k == 0) nib0 = nb;
            f[k] = nibble_to_flags(nb);
        }
        f[kWtVec] = 0u;
        const uint32_t nibo = head_nibble<REV>(bits, go, n);
        p_own = first_head_128(nib0);
        by_pred = t > 0 && p_own > 0 && p_own < 128;
        need_x = t > 0 && p_own == 128;
        p_nx = has_next ? first_head_128(nibo) : 0;
        over = p_nx > 0 && p_nx < 128;
        return true;
    }
    // Pull the tile's (and overhang's) part of a stream into L2: lanes touch its 128-byte lines.
    __device__ __forceinline__ void prefetch_l2(const void* p, int bytes_per_item) const {
        const long long pt = (g0 - (REV ? (long long)(kWtVec - 1) * 128 + (31 - lane) * 4 : lane * 4));  // tile base
        const int ov = over ? p_nx : 0;  // overhang items this warp produces
        long long lo = REV ? pt - ov : pt, hi = REV ? pt + kWtItems : pt + kWtItems + ov;
        if (lo < 0) lo = 0;
        if (hi > n) hi = n;
        const char* b = reinterpret_cast<const char*>(p) + lo * bytes_per_item;
        const long long bytes = (hi - lo) * bytes_per_item;
        for (long long off = (long long)lane * 128; off < bytes; off += 32 * 128)
            asm volatile("prefetch.global.L2 [%0];" :: "l"(b + off));
    }
    // physical index of the first item of logical vector k (k == kWtVec: the overhang)
    __device__ __forceinline__ long long g(int k) const { return k < kWtVec ? g0 + k * gstep : go; }
    // bit c set <=> this warp writes logical component c of vector k
    __device__ __forceinline__ uint32_t mask(int k) const {
        if (k == 0 && by_pred) {
            const int d = p_own - 4 * lane;
            return d <= 0 ? 0xfu : (d >= 4 ? 0u : (0xfu << d) & 0xfu);
        }
        if (k == kWtVec) {
            const int d = over ? p_nx - 4 * lane : 0;
            return d <= 0 ? 0u : (d >= 4 ? 0xfu : ~(0xfu << d) & 0xfu);
        }
        return 0xfu;
    }
    // physical index of logical component c of vector k
    __device__ __forceinline__ long long item(int k, int c) const { return g(k) + (REV ? 3 - c : c); }

    __device__ __forceinline__ void load(const float* __restrict__ p, int k, float fill, float (&v)[4]) const {
        const long long gg = g(k);
        const bool whole = vec_ok && gg >= 0 && gg + 3 < n;
#pragma unroll
        for (int c = 0; c < 4; ++c) v[c] = fill;
        if (REV) ldg_stream4_if(whole, p + gg, v[3], v[2], v[1], v[0]);
        else ldg_stream4_if(whole, p + gg, v[0], v[1], v[2], v[3]);
        if (slow) {
#pragma unroll
            for (int c = 0; c < 4; ++c) {
                const long long gi = gg + (REV ? 3 - c : c);
                ldg_stream_if(!whole && gi >= 0 && gi < n, p + gi, v[c]);
            }
        }
    }
    // writes the components this warp owns (mask)
    __device__ __forceinline__ void store(float* __restrict__ p, int k, const float (&v)[4]) const {
        const long long gg = g(k);
        const uint32_t m = mask(k);
        const bool whole = vec_ok && gg >= 0 && gg + 3 < n && m == 0xfu;
        if (REV) stg_stream4_if(whole, p + gg, v[3], v[2], v[1], v[0]);
        else stg_stream4_if(whole, p + gg, v[0], v[1], v[2], v[3]);
        if (slow || (m != 0xfu && m != 0u)) {
#pragma unroll
            for (int c = 0; c < 4; ++c) {
                const long long gi = gg + (REV ? 3 - c : c);
                stg_stream_if(!whole && ((m >> c) & 1u) && gi >= 0 && gi < n, p + gi, v[c]);
            }
        }
    }

    // Segmented inclusive scan of the tile and its overhang.
    //   in : v[k][c] values (k == kWtVec: the overhang vector)
    //   out: v[k][c] thread-local scans, E[k] carries, exactly as tile_segscan leaves them: feed
    //        (v[k], f[k], E[k]) to finish_inclusive / finish_exclusive for k = 0..kWtVec-1 and,
    //        when `over`, for k == kWtVec.
    // status: n_slots words per warp tile; `slot` selects the one this scan uses.
    // ALWAYS: scan the overhang whether or not it is used (kernels with a long tail after the
    // scan: a branch on `over` here makes the compiler clone that tail for both outcomes).
    template <class Op, bool ALWAYS = false>
    __device__ __forceinline__ void scan(float (&v)[kWtNV][4], float (&E)[kWtNV], unsigned long long* status,
                                         int n_slots, int slot, uint32_t epoch) const {
        float wc;
        bool open_vec[kWtNV], warp_open;
        warp_segscan<Op, kWtVec, kWtNV>(v, f, E, open_vec, wc, warp_open);
        unsigned long long* mine = status + t * n_slots + slot;
        const bool final_now = (t == 0) || !warp_open;
        if (lane == 0) st_relaxed_u64(mine, pack_status(wc, epoch, final_now ? 2u : 1u));
        float X = Op::identity();
        if (need_x) {
            X = warp_lookback_carry<Op>(status + slot, n_slots, t, epoch);
            if (!final_now && lane == 0) st_relaxed_u64(mine, pack_status(Op::apply(X, wc), epoch, 2u));
        }
#pragma unroll
        for (int k = 0; k < kWtVec; ++k)
            if (open_vec[k]) E[k] = Op::apply(X, E[k]);
        if (!ALWAYS && !over) return;
        const float carry = warp_open ? Op::apply(X, wc) : wc;  // inclusive value at my last item
        float (&u)[4] = v[kWtVec];
#pragma unroll
        for (int c = 1; c < 4; ++c) u[c] = Op::apply(u[c - 1], u[c]);
        float tt = u[3];
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) {
            const float y = __shfl_up_sync(kFull, tt, d);
            if (lane >= d) tt = Op::apply(y, tt);
        }
        const float e = __shfl_up_sync(kFull, tt, 1);
        E[kWtVec] = (lane == 0) ? carry : Op::apply(carry, e);
    }
};

}  // namespace nfb
