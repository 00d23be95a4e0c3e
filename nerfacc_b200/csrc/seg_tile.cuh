// seg_tile.cuh — the flat segmented-scan tile shared by the scan / volume-rendering /
// accumulation kernels.
//
// The packed sample array is cut into 4096-element tiles, one CTA (256 threads) per
// tile, independent of how the samples are split into rays: a 1-sample ray and a
// 10^5-sample ray cost the same bytes.  Ray boundaries ("heads") come from packed_info
// (16 B per ray, not 8 B per sample): each CTA finds the rays that start inside its
// tile with a warp-cooperative 32-ary search and scatters 1-byte head flags into shared
// memory.  Inside the tile the scan is warp-shuffle based (one SHFL per step, head
// masks travel as ballots); tiles are chained with decoupled look-back, where a tile
// that contains a head publishes its carry-out immediately, so for typical NeRF rays
// (tens of samples) the look-back distance is exactly one tile.
//
// Layout of one tile (logical order): warp w owns items [512w, 512w+512); its lane l
// holds, for k = 0..3, the float4 at 512w + 128k + 4l — every global access is a fully
// coalesced 512-byte warp transaction.  A reverse scan (the autograd backward) walks
// the same physical tiles from the end of the array and mirrors (warp, k, lane,
// component), so loads stay aligned and coalesced.
#pragma once

#include "common.cuh"

namespace nfb {

constexpr int kThreads = 256;
constexpr int kWarps = kThreads / 32;
constexpr int kVec = 4;                   // float4 per thread
constexpr int kTile = kThreads * kVec * 4;  // 4096 items
constexpr int kWarpItems = kTile / kWarps;  // 512
constexpr int kStatusPerTile = 8;          // look-back words reserved per tile
constexpr uint32_t kEpochMemset = 0x3fffffffu;

struct SumOp {
    __device__ static __forceinline__ float identity() { return 0.0f; }
    __device__ static __forceinline__ float apply(float a, float b) { return __fadd_rn(a, b); }
};
struct ProdOp {
    __device__ static __forceinline__ float identity() { return 1.0f; }
    __device__ static __forceinline__ float apply(float a, float b) { return __fmul_rn(a, b); }
};

static inline long long num_tiles(long long n) { return (n + kTile - 1) / kTile; }

// ------------------------------------------------------------------ geometry
template <bool REV>
struct TileGeom {
    long long n;        // number of items in the array
    long long base;     // physical index of the first item of this CTA's tile
    __device__ __forceinline__ TileGeom(long long n_, long long n_tiles) : n(n_) {
        const long long pt = REV ? (n_tiles - 1 - (long long)blockIdx.x) : (long long)blockIdx.x;
        base = pt * kTile;
    }
    // physical tile-local offset of this thread's k-th vector
    __device__ __forceinline__ int voff(int k) const {
        const int w = warp_id(), l = lane_id();
        return REV ? ((kWarps - 1 - w) * kWarpItems + (kVec - 1 - k) * 128 + (31 - l) * 4)
                   : (w * kWarpItems + k * 128 + l * 4);
    }
    // logical tile-local index of physical tile-local item P
    __device__ __forceinline__ static int logical(int P) { return REV ? (kTile - 1 - P) : P; }

    // Load this thread's k-th vector (logical component order); items past the end of
    // the array read as `fill`.  Branch-free for whole vectors (see ldg_stream4_if); only a
    // tile that crosses the end of the array, or misaligned pointers, take the scalar path.
    __device__ __forceinline__ void load(const float* __restrict__ p, bool vec_ok, int k, float fill,
                                         float (&v)[4]) const {
        const long long g = base + voff(k);
        const bool whole = vec_ok && g + 3 < n;
#pragma unroll
        for (int c = 0; c < 4; ++c) v[c] = fill;
        if (REV) ldg_stream4_if(whole, p + g, v[3], v[2], v[1], v[0]);
        else ldg_stream4_if(whole, p + g, v[0], v[1], v[2], v[3]);
        if (!(vec_ok && base + kTile <= n)) {
#pragma unroll
            for (int c = 0; c < 4; ++c) {
                const long long gi = g + (REV ? 3 - c : c);
                ldg_stream_if(!whole && gi < n, p + gi, v[c]);
            }
        }
    }
    __device__ __forceinline__ void store(float* __restrict__ p, bool vec_ok, int k,
                                          const float (&v)[4]) const {
        const long long g = base + voff(k);
        const bool whole = vec_ok && g + 3 < n;
        if (REV) stg_stream4_if(whole, p + g, v[3], v[2], v[1], v[0]);
        else stg_stream4_if(whole, p + g, v[0], v[1], v[2], v[3]);
        if (!(vec_ok && base + kTile <= n)) {
#pragma unroll
            for (int c = 0; c < 4; ++c) {
                const long long gi = g + (REV ? 3 - c : c);
                stg_stream_if(!whole && gi < n, p + gi, v[c]);
            }
        }
    }
};

// ------------------------------------------------------------------ head flags
// Where the segments are.  Either a canonical packed_info (int64 [n_rays,2]) or, when
// packed_info == nullptr, uniform segments of `uniform_len` items (batched [n_rays, len]
// tensors need no packed_info at all).
struct SegSpec {
    const int64_t* packed_info;
    long long n_rays;
    long long uniform_len;
};

// Scatter one byte per segment head of this tile into s_flags (logical order).  Forward: a
// head sits at each non-empty ray's first sample; reverse: at its last sample.  With
// WITH_NEXT the byte at logical index kTile tells whether the item right after the tile
// starts a new segment (forward only).  Must be called by all 256 threads; ends with a
// barrier.  s_range receives [first ray with start>=a (fwd) / end>a (rev), first ray with
// start>=b / end>b].  With WITH_OWNED (forward only) s_range[2..3] additionally receive the
// range of rays whose end (start+count) lies in (a, b] — the rays that *finish* in this tile.
template <bool REV, bool WITH_NEXT = false, bool WITH_OWNED = false>
__device__ __forceinline__ void scatter_heads(const SegSpec& seg, long long base, long long n,
                                              uint32_t* s_flags /*[kTile/4 + 1]*/,
                                              long long* s_range /*[2] or [4]*/) {
    reinterpret_cast<uint4*>(s_flags)[threadIdx.x] = make_uint4(0, 0, 0, 0);
    if (threadIdx.x == 0) s_flags[kTile / 4] = 0;
    const long long a = base, b = base + kTile;
    uint8_t* fb = reinterpret_cast<uint8_t*>(s_flags);
    if (seg.packed_info == nullptr) {
        __syncthreads();
        const long long L = seg.uniform_len;
        // forward heads at multiples of L, reverse heads at L-1 (mod L)
        const long long off = REV ? (L - 1) : 0;
        long long h = ((a - off + L - 1) / L) * L + off;  // first head position >= a
        if (a - off < 0) h = off;
        for (h += (long long)threadIdx.x * L; h < b + (WITH_NEXT ? 1 : 0) && h < n + (WITH_NEXT ? 1 : 0); h += (long long)kThreads * L) {
            const long long P = h - a;
            if (P < kTile) fb[TileGeom<REV>::logical((int)P)] = 1;
            else fb[kTile] = 1;
        }
        __syncthreads();
        return;
    }
    const int64_t* __restrict__ packed_info = seg.packed_info;
    const long long n_rays = seg.n_rays;
    if (warp_id() < 2) {
        const long long x = (warp_id() == 0) ? a : b + (WITH_NEXT ? 1 : 0);
        long long r;
        if (!REV) r = warp_search_first(n_rays, [&](long long i) { return load_chunk(packed_info, i).start >= x; });
        else r = warp_search_first(n_rays, [&](long long i) { const RayChunk c = load_chunk(packed_info, i); return c.start + c.cnt > x; });
        if (lane_id() == 0) s_range[warp_id()] = r;
    } else if (WITH_OWNED && warp_id() < 4) {
        const long long x = (warp_id() == 2) ? a : b;
        const long long r = warp_search_first(n_rays, [&](long long i) { const RayChunk c = load_chunk(packed_info, i); return c.start + c.cnt > x; });
        if (lane_id() == 0) s_range[warp_id()] = (warp_id() == 2 && a == 0) ? 0 : r;
    }
    __syncthreads();
    const long long r_lo = s_range[0], r_hi = s_range[1];
    for (long long r = r_lo + threadIdx.x; r < r_hi; r += kThreads) {
        const RayChunk c = load_chunk(packed_info, r);
        if (c.cnt > 0) {
            const long long pos = REV ? (c.start + c.cnt - 1) : c.start;
            const long long P = pos - a;
            if (P >= 0 && P < kTile) fb[TileGeom<REV>::logical((int)P)] = 1;
            else if (WITH_NEXT && P == kTile) fb[kTile] = 1;
        }
    }
    __syncthreads();
}

// flags of this thread's k-th vector: byte c != 0 <=> logical component c is a head
__device__ __forceinline__ uint32_t vec_flags(const uint32_t* s_flags, int k) {
    return s_flags[warp_id() * 128 + k * 32 + lane_id()];
}
__device__ __forceinline__ bool flag_at(uint32_t f, int c) { return (f >> (8 * c)) & 0xffu; }

// ------------------------------------------------------------------ look-back
__device__ __forceinline__ unsigned long long pack_status(float v, uint32_t epoch, uint32_t state) {
    return ((unsigned long long)((epoch << 2) | state) << 32) | (unsigned long long)__float_as_uint(v);
}

// Carry into `tile` from its predecessors: walk back to the nearest tile whose carry-out
// is final (state 2), then fold the aggregates forward left-to-right — the same
// association whichever tile the walk stops at, so results are run-to-run deterministic.
template <class Op>
__device__ __forceinline__ float lookback_carry(unsigned long long* st, int slot, long long tile,
                                                uint32_t epoch) {
    long long k = tile - 1;
    unsigned long long w;
    while (true) {
        w = ld_relaxed_u64(st + k * kStatusPerTile + slot);
        const uint32_t meta = (uint32_t)(w >> 32);
        if ((meta >> 2) != epoch || (meta & 3u) == 0u) { __nanosleep(40); continue; }
        if ((meta & 3u) == 2u) break;
        --k;  // tile 0 always publishes state 2, so k never drops below 0
    }
    float acc = __uint_as_float((uint32_t)w);
    for (long long j = k + 1; j < tile; ++j) {
        w = ld_relaxed_u64(st + j * kStatusPerTile + slot);
        const float val = __uint_as_float((uint32_t)w);
        acc = (((uint32_t)(w >> 32)) & 3u) == 2u ? val : Op::apply(acc, val);
    }
    return acc;
}

// ------------------------------------------------------------------ head bits
// One bit per item instead of a search through packed_info: bit i is set iff item i is the
// first item of a ray, bit n is always set.  Because a canonical layout covers [0, n) without
// gaps, "item i is the last of its ray" is bit i+1, so the same array serves both scan
// directions.  ceil((n+1)/32)+2 words (the reverse read funnels two words).
static inline size_t head_bits_words(long long n) { return (size_t)((n + 1 + 31) / 32 + 2); }

// Head bits (bit c = logical component c) of the vector whose first physical item is g (a
// multiple of 4); 0 outside [0, n).
template <bool REV>
__device__ __forceinline__ uint32_t head_nibble(const uint32_t* __restrict__ bits, long long g, long long n) {
    if (g >= n || g < 0) return 0u;
    if (!REV) return (__ldg(bits + (g >> 5)) >> (unsigned)(g & 31)) & 0xfu;
    const long long h = g + 1;  // tails of [g, g+4) are heads of [g+1, g+5)
    const uint32_t lo = __ldg(bits + (h >> 5)), hi = __ldg(bits + (h >> 5) + 1);
    return __brev(__funnelshift_r(lo, hi, (unsigned)(h & 31)) & 0xfu) >> 28;  // logical c = 3 - physical
}
// bit c -> byte c (the form vec_flags / flag_at use)
__device__ __forceinline__ uint32_t nibble_to_flags(uint32_t b) { return (b * 0x00204081u) & 0x01010101u; }

// ------------------------------------------------------------------ warp look-back
// Decoupled look-back at warp granularity (one status word per 512 items, no block barrier
// anywhere): all 32 lanes inspect 32 predecessors per round.  Same fold order as
// lookback_carry: from the nearest final word forward, left to right.
template <class Op>
__device__ __noinline__ float warp_lookback_far(unsigned long long* st, int stride, long long tile, uint32_t epoch) {
    const int lane = lane_id();
    long long hi = tile - 1, k_final;
    while (true) {  // pass 1: nearest predecessor whose word is final, all nearer ones published
        const long long idx = hi - lane;
        uint32_t state = 2u;
        if (idx >= 0) {
            const uint32_t meta = (uint32_t)(ld_relaxed_u64(st + idx * stride) >> 32);
            state = ((meta >> 2) == epoch) ? (meta & 3u) : 0u;
        }
        const unsigned fin = __ballot_sync(kFull, state == 2u), none = __ballot_sync(kFull, state == 0u);
        if (fin) {
            const int f = __ffs(fin) - 1;
            if ((none & ((1u << f) - 1u)) == 0u) { k_final = hi - f; break; }
        } else if (none == 0u) {
            hi -= 32;
            continue;
        }
        __nanosleep(100);
    }
    float acc = Op::identity();
    for (long long b = k_final; b < tile; b += 32) {  // pass 2: fold forward
        const long long idx = b + lane;
        const unsigned long long w = (idx < tile) ? ld_relaxed_u64(st + idx * stride) : 0ull;
        const float val = __uint_as_float((uint32_t)w);
        const bool fin = (((uint32_t)(w >> 32)) & 3u) == 2u;
        const int cnt = (tile - b < 32) ? (int)(tile - b) : 32;
        for (int j = 0; j < cnt; ++j) {
            const float vj = __shfl_sync(kFull, val, j);
            const bool fj = __shfl_sync(kFull, (int)fin, j) != 0;
            acc = fj ? vj : Op::apply(acc, vj);
        }
    }
    return acc;
}

// The common case first: the word right before this tile is final (its warp held a ray
// boundary, or has finished its own look-back) — one broadcast load per poll.
template <class Op>
__device__ __forceinline__ float warp_lookback_carry(unsigned long long* st, int stride, long long tile,
                                                     uint32_t epoch) {
    while (true) {
        const unsigned long long w = ld_relaxed_u64(st + (tile - 1) * stride);
        const uint32_t meta = (uint32_t)(w >> 32);
        if ((meta >> 2) == epoch && (meta & 3u) != 0u) {
            if ((meta & 3u) == 2u) return __uint_as_float((uint32_t)w);
            break;
        }
        __nanosleep(100);
    }
    return warp_lookback_far<Op>(st, stride, tile, epoch);
}

// ------------------------------------------------------------------ tile scan
struct TileSmem {
    uint32_t flags[kTile / 4 + 2];
    long long range[4];
    float wagg[kStatusPerTile][kWarps];
    int whead[kWarps];
    float carry[kStatusPerTile];
};

// Segmented inclusive scan of one tile, in place.
//   in : v[k][c]  values, f[k] head flags (logical order)
//   out: v[k][c]  = scan restricted to [last head at or before the item within this
//                   thread's vector .. item]   (the thread-local part), and
//        E[k]     = value carried into the vector from everything before it (previous
//                   lanes, warps and tiles) up to the segment's head;
//        open[k]  bit c set <=> component c has no head at or before it inside the vector,
//                   i.e. its full result is Op(E[k], v[k][c]); otherwise it is v[k][c].
// `slot` selects the look-back word (kernels scanning several quantities call this once
// per quantity with different slots).  Contains two block barriers.
// The warp part: segmented inclusive scan of the 512 items one warp holds.  wc = aggregate
// of the items after the warp's last head (of all of them if warp_open), E[k] = value
// carried into vector k from the warp's earlier items, open_vec[k] <=> no head between the
// warp's first item and vector k.
template <class Op, int VEC, int NV = VEC>
__device__ __forceinline__ void warp_segscan(float (&v)[NV][4], const uint32_t (&f)[NV],
                                             float (&E)[NV], bool (&open_vec)[NV], float& wc,
                                             bool& warp_open) {
    const int lane = lane_id();
    const unsigned lt_mask = (1u << lane) - 1u;
    wc = Op::identity();  // carry inside the warp, relative to the warp's first item
    warp_open = true;     // no head so far in this warp
#pragma unroll
    for (int k = 0; k < VEC; ++k) {
        // thread-local segmented inclusive scan over the 4 components
#pragma unroll
        for (int c = 1; c < 4; ++c)
            if (!flag_at(f[k], c)) v[k][c] = Op::apply(v[k][c - 1], v[k][c]);
        const bool any = f[k] != 0u;
        const unsigned hmask = __ballot_sync(kFull, any);
        float t = v[k][3];  // after the last head of this lane (or the whole lane)
        // a partial from distance d is absorbed iff lanes (lane-d, lane] hold no head, i.e.
        // d <= lane - (nearest head lane at or before this one); without one, d <= lane
        const unsigned le = hmask & (lt_mask | (1u << lane));
        const int lim = le ? (lane - 31 + __clz(le)) : lane;
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) {
            const float y = __shfl_up_sync(kFull, t, d);
            if (d <= lim) t = Op::apply(y, t);
        }
        float e = __shfl_up_sync(kFull, t, 1);
        const bool no_head_before = (hmask & lt_mask) == 0u;
        if (lane == 0) e = wc;
        else if (no_head_before) e = Op::apply(wc, e);
        E[k] = e;
        open_vec[k] = warp_open && no_head_before;
        const float t31 = __shfl_sync(kFull, t, 31);
        wc = (hmask == 0u) ? Op::apply(wc, t31) : t31;
        warp_open = warp_open && (hmask == 0u);
    }
}

template <class Op, bool LOOKBACK = true>
__device__ __forceinline__ void tile_segscan(float (&v)[kVec][4], const uint32_t (&f)[kVec],
                                             float (&E)[kVec], unsigned long long* status,
                                             int slot, uint32_t epoch, TileSmem& sm) {
    const int lane = lane_id(), w = warp_id();
    float wc;
    bool warp_open;
    bool open_vec[kVec];
    warp_segscan<Op, kVec>(v, f, E, open_vec, wc, warp_open);
    if (lane == 0) { sm.wagg[slot][w] = wc; sm.whead[w] = warp_open ? 0 : 1; }
    __syncthreads();
    // carry into this warp from the earlier warps of the tile (8 entries, serial fold)
    float C = Op::identity();
    bool head_before = false;
    float A = Op::identity();
    bool tile_head = false;
#pragma unroll
    for (int j = 0; j < kWarps; ++j) {
        if (j == w) { C = A; head_before = tile_head; }
        const float aj = sm.wagg[slot][j];
        if (sm.whead[j]) { A = aj; tile_head = true; }
        else A = Op::apply(A, aj);
    }
    if (!LOOKBACK) {
        // tile-local scan: the tile start acts as a head, nothing is published
#pragma unroll
        for (int k = 0; k < kVec; ++k)
            if (open_vec[k]) E[k] = Op::apply(C, E[k]);
        return;
    }
    if (threadIdx.x == 0) {
        const long long tile = blockIdx.x;  // logical tile order == launch order
        unsigned long long* mine = status + tile * kStatusPerTile + slot;
        float X = Op::identity();
        if (tile == 0) {
            st_relaxed_u64(mine, pack_status(A, epoch, 2u));
        } else {
            st_relaxed_u64(mine, pack_status(A, epoch, tile_head ? 2u : 1u));
            X = lookback_carry<Op>(status, slot, tile, epoch);
            if (!tile_head) st_relaxed_u64(mine, pack_status(Op::apply(X, A), epoch, 2u));
        }
        sm.carry[slot] = X;
    }
    __syncthreads();
    const float Cw = head_before ? C : Op::apply(sm.carry[slot], C);
#pragma unroll
    for (int k = 0; k < kVec; ++k)
        if (open_vec[k]) E[k] = Op::apply(Cw, E[k]);
}

// Finish helpers: full inclusive / exclusive results of one vector.
template <class Op>
__device__ __forceinline__ void finish_inclusive(const float (&s)[4], uint32_t f, float E, float (&o)[4]) {
    bool open = true;
#pragma unroll
    for (int c = 0; c < 4; ++c) {
        open = open && !flag_at(f, c);
        o[c] = open ? Op::apply(E, s[c]) : s[c];
    }
}
template <class Op>
__device__ __forceinline__ void finish_exclusive(const float (&s)[4], uint32_t f, float E, float (&o)[4]) {
    bool open = true;
    float prev = E;  // inclusive result of the item before component c
#pragma unroll
    for (int c = 0; c < 4; ++c) {
        const bool h = flag_at(f, c);
        o[c] = h ? Op::identity() : prev;
        open = open && !h;
        prev = open ? Op::apply(E, s[c]) : s[c];
    }
}

}  // namespace nfb
