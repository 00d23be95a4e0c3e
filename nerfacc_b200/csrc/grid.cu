// grid.cu — ray_aabb_intersect, occupancy bit-packing and the two-pass grid traversal
// (count -> per-ray scan -> fill) behind the C ABI.
//
// Replaces csrc/grid.cu of the reference.  What changed structurally:
//  * ray/box intersection + event sort are fused into the traversal when the caller does
//    not pass precomputed t_sorted/t_indices/hits (the reference runs ray_aabb_intersect,
//    torch.cat and torch.sort with int64 indices first, grid.py:158-162);
//  * occupancy is a 1-bit/cell field (256 KiB per 128^3 level) plus an "any occupied" mip
//    that the marching kernels keep in shared memory, instead of 1-byte bools in HBM;
//  * marching runs on the warp-convergent producer/consumer engine of grid.cuh (persistent
//    threads, event queue, exact closed-form t += dt recurrence);
//  * per-ray counts go straight into the int64 packed_info layout and one look-back scan
//    produces starts and the grand totals (the reference: 2x torch.cumsum + 2x .item());
//  * with a constant dt pass 1 also records each ray's samples as a few arithmetic
//    progressions ("runs"), and pass 2 of OccGridEstimator.sampling is a pure,
//    warp-cooperative expansion of those runs — the grid is marched once, not twice.
#include <algorithm>
#include <type_traits>

#include "grid.cuh"

namespace nfb {

int launch_ray_scan(int64_t* packed_info, long long n_rays, long long* total_out, void* status,
                    cudaStream_t s, const unsigned long long* only_if = nullptr,
                    const unsigned long long* also_src = nullptr, long long* also_dst = nullptr);
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
    // optional work list: march rays ray_list[0..n_work) instead of 0..n_rays (the fill pass
    // re-marches only the rays whose runs overflowed the run buffer)
    const int* __restrict__ ray_list;
    long long n_work;
    // optional, read by the kernel when it starts (the chunked test-time loop keeps both on the
    // device so that an iteration needs no host round trip): the length of ray_list (n_work is then
    // only an upper bound that sizes the launch) and traverse_steps_limit
    const int* __restrict__ n_work_dev;
    const int* __restrict__ limit_dev;
};

// ---------------------------------------------------------------------------------- sinks
// Sink methods take the policy first: it is a kernel parameter (constant bank), so what a sink
// needs from it costs no registers.
struct CountSink {
    static constexpr bool kPieces = false;
    template <class P> __device__ __forceinline__ void run(const P&, float, float, int, bool, int, int) const {}
    template <class P> __device__ __forceinline__ void one(const P&, float, float, bool, int, int) const {}
};

// Run buffer written by the count pass when marching with a constant dt (cone_angle == 0,
// step_size > 0): per ray up to kRunSlots runs (t0, q, n, s0) — n samples whose edges are
// t0 + j*q, the first of them sample s0 of the ray — followed by one int32 run count per ray
// (kRunSlots + 1: the run pool was full and the ray is marched again by the fill pass).  Only the slots in use
// are ever touched, so the traffic is ~2 sectors per ray, not the reserved 128 bytes.
constexpr int kRunSlots = 8;
struct __align__(16) Run { float t0, q; int n, s0; };

// Runs beyond a ray's kRunSlots inline slots go to a global pool (bump-allocated, kPoolPerRay
// entries per ray on average), expanded by traverse_expand_pool_kernel; only when the pool is
// full is a ray flagged for a second march.
constexpr int kPoolPerRay = 2;
struct RunPool {
    Run* __restrict__ runs;
    int* __restrict__ ray;
    unsigned long long* __restrict__ cursor;
    unsigned long long cap;
};

struct RunSink {
    static constexpr bool kPieces = false;
    int idx, r;
    bool failed;
    // Every piece advance() hands over becomes one run: pieces are contiguous only at binade
    // crossings, so merging them here would save few slots and cost registers in the hot loop.
    template <class P>
    __device__ __forceinline__ void run(const P& pol, float t0, float q, int k, bool, int nS, int) {
        const int4 v = make_int4(__float_as_int(t0), __float_as_int(q), k, nS);
        if (idx < kRunSlots) {
            *reinterpret_cast<int4*>(pol.runs + (long long)r * kRunSlots + idx) = v;
        } else if (!failed) {
            const unsigned long long e = atomicAdd(pol.pool.cursor, 1ull);
            if (e < pol.pool.cap) {
                *reinterpret_cast<int4*>(pol.pool.runs + e) = v;
                pol.pool.ray[e] = r;
            } else {
                failed = true;
            }
        }
        ++idx;
    }
    template <class P> __device__ __forceinline__ void one(const P&, float, float, bool, int, int) { failed = true; }
};

// Piece log written by the count pass when marching with cone stepping (cone_angle > 0, step_size > 0):
// dt follows t, so a run of samples has no closed form — but it is still the pure recurrence
// t += calc_dt(t) from its first t.  The engine (grid.cuh, kPieces) cuts every run into pieces of at
// most kPieceLen samples and appends one record (t0, n, s0, ray) per piece: the first sample of the
// piece is sample s0 of its ray and starts at t0.  Pass 2 (traverse_expand_cone_kernel) replays the
// pieces, 32 at a time, instead of marching the grid again.
constexpr int kPiecesPerRay = 32;  // log capacity per ray on average; a full log sends pass 2 back to marching
struct PieceLog {
    int4* __restrict__ log;           // [n_chunks][kPieceChunk]
    int* __restrict__ chunk_fill;     // records in use per chunk (zeroed before the pass)
    unsigned long long* __restrict__ cursor;  // next free chunk
    unsigned long long* __restrict__ failed;  // set when the log ran out of chunks
    int n_chunks;
};
struct PieceSink {
    static constexpr bool kPieces = true;
    float p_t0;  // the open piece: starts at sample p_s0 of the ray, at t = p_t0
    int p_s0;
    template <class P> __device__ __forceinline__ void run(const P&, float, float, int, bool, int, int) const {}
    template <class P> __device__ __forceinline__ void one(const P&, float, float, bool, int, int) const {}
};

struct SamplingSink {  // OccGridEstimator.sampling outputs (occ_grid.py:174-177), one lane per ray
    static constexpr bool kPieces = false;
    int64_t* __restrict__ ray_indices;
    float* __restrict__ t_starts;
    float* __restrict__ t_ends;
    long long base, r;
    template <class P>
    __device__ __forceinline__ void one(const P&, float t_last, float t_next, bool, int nS, int) const {
        const long long i = base + nS;
        if (ray_indices) ray_indices[i] = r;
        t_starts[i] = t_last;
        t_ends[i] = t_next;
    }
    template <class P>
    __device__ __forceinline__ void run(const P& pol, float t0, float q, int k, bool, int nS, int) const {
        for (int j = 0; j < k; ++j)
            one(pol, __fmaf_rn((float)j, q, t0), __fmaf_rn((float)(j + 1), q, t0), true, nS + j, 0);
    }
};

struct FullSink {  // traverse_grids outputs (grid.cu:218-255)
    static constexpr bool kPieces = false;
    float* __restrict__ iv_vals; int64_t* __restrict__ iv_ray;
    uint8_t* __restrict__ iv_left; uint8_t* __restrict__ iv_right;
    float* __restrict__ sm_vals; int64_t* __restrict__ sm_ray; uint8_t* __restrict__ sm_valid;
    long long iv_base, sm_base, r;
    template <class P>
    __device__ __forceinline__ void one(const P&, float t_last, float t_next, bool continuous, int nS, int nI) const {
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
    template <class P>
    __device__ __forceinline__ void run(const P& pol, float t0, float q, int k, bool continuous, int nS, int nI) const {
        for (int j = 0; j < k; ++j) {
            one(pol, __fmaf_rn((float)j, q, t0), __fmaf_rn((float)(j + 1), q, t0), continuous, nS + j, nI);
            nI += continuous ? 1 : 2;
            continuous = true;
        }
    }
};

// ---------------------------------------------------------------------------------- policies
// Pass 1: counts (into packed_info[2r+1]), termination planes and, optionally, the run buffer.
// MODE: 0 counts only, 1 + the run buffer (constant dt), 2 + the piece log (cone stepping)
template <int MODE>
struct CountPolicy {
    static constexpr bool RUNS = MODE == 1;
    using Sink = typename std::conditional<MODE == 1, RunSink, typename std::conditional<MODE == 2, PieceSink, CountSink>::type>::type;
    int64_t* __restrict__ sample_info;
    int64_t* __restrict__ interval_info;
    float* __restrict__ term;
    Run* __restrict__ runs;
    int* __restrict__ n_runs;  // inline runs of the ray, or kRunSlots + 1: march it again
    int* __restrict__ overflow_list;
    unsigned long long* __restrict__ n_overflow;
    RunPool pool;
    PieceLog pieces;
    __device__ __forceinline__ bool accept(long long) const { return true; }
    __device__ __forceinline__ void skipped(long long) const {}
    __device__ __forceinline__ bool wants_term() const { return term != nullptr; }
    __device__ __forceinline__ Sink begin(long long r) const {
        Sink s;
        if constexpr (RUNS) { s.idx = 0; s.r = (int)r; s.failed = false; }
        if constexpr (MODE == 2) { s.p_t0 = 0.0f; s.p_s0 = 0; }
        return s;
    }
    __device__ __forceinline__ void end(long long r, Sink& s, int nS, int nI, float t_last) const {
        if (sample_info) sample_info[2 * r + 1] = nS;
        if (interval_info) interval_info[2 * r + 1] = nI;
        if (term) term[r] = t_last;
        if constexpr (RUNS) {
            n_runs[r] = s.failed ? kRunSlots + 1 : (s.idx < kRunSlots ? s.idx : kRunSlots);
            if (s.failed) overflow_list[atomicAdd(n_overflow, 1ull)] = (int)r;
        }
    }
};

template <bool FULL>
struct FillPolicy {
    using Sink = typename std::conditional<FULL, FullSink, SamplingSink>::type;
    const uint8_t* __restrict__ rays_mask;  // honoured in single-pass mode only (grid.cu:380)
    const float* __restrict__ near_arr; float near_s;
    // where each ray's data starts: info[r * stride]; packed_info (stride 2) or a plain array
    const int64_t* sm_starts; int sm_stride;
    const int64_t* iv_starts; int iv_stride;
    const int* slot_mul_dev;  // optional: sample starts are slot numbers, times this many samples per slot
    // two-pass: skip rays whose count (info[2r+1]) is zero; single pass: write the counts
    int64_t* sample_info; int64_t* interval_info; bool single_pass;
    float* term;
    Sink proto;
    __device__ __forceinline__ bool masked(long long r) const { return rays_mask != nullptr && !rays_mask[r]; }
    __device__ __forceinline__ bool wants_term() const { return term != nullptr; }
    __device__ __forceinline__ bool accept(long long r) const {
        if (masked(r)) return false;
        if (!single_pass && sample_info[2 * r + 1] == 0) return false;  // grid.cu:103-106
        return true;
    }
    __device__ __forceinline__ void skipped(long long r) const {
        if (single_pass && masked(r)) {
            if (sample_info) sample_info[2 * r + 1] = 0;
            if (interval_info) interval_info[2 * r + 1] = 0;
            if (term) term[r] = near_arr ? near_arr[r] : near_s;
        }
    }
    __device__ __forceinline__ Sink begin(long long r) const {
        Sink s = proto;
        s.r = r;
        if constexpr (FULL) { s.sm_base = sm_starts[r * sm_stride]; s.iv_base = iv_starts[r * iv_stride]; }
        else s.base = sm_starts[r * sm_stride] * (slot_mul_dev ? (long long)*slot_mul_dev : 1ll);
        return s;
    }
    __device__ __forceinline__ void end(long long r, Sink&, int nS, int nI, float t_last) const {
        if (single_pass) {
            if (sample_info) sample_info[2 * r + 1] = nS;
            if (interval_info) interval_info[2 * r + 1] = nI;
        }
        if (term) term[r] = t_last;
    }
};

// ---------------------------------------------------------------------------------- kernels
// Persistent CTAs: dynamic shared memory = [mip words][event queue kQ x blockDim floats]
// [prepared-ray pool: warps x kPoolWords x 32 words].
constexpr int kConeThreads = 512;
constexpr int kMarchThreads = 384;  // x 2 CTAs/SM: 85 registers per thread (512 x 2 = 64 spills ten loop invariants of the DDA)

// SINGLE: one grid level, known at compile time — everything the engine does for nested levels
// (event sorting, later segments and their DDA setup) folds away: a smaller kernel stalls less on
// instruction fetch.  Instantiated for the constant-dt count pass only (launch_march).
// Cone kernels (DIRECT) run 512 threads per CTA at 64 registers: a launch of 2^17 rays then has a lane per
// ray instead of two rays behind one another on some lanes (cfg4 sampling at 2^17 rays: 1.74 -> 1.64 ms;
// at 2^20 rays 10.03 -> 10.14 ms; 640 threads: 2.06 / 11.7 ms); the queue kernels lose 7 % at 512 and stay at 384.
template <bool DIRECT> constexpr int march_threads() { return DIRECT ? kConeThreads : kMarchThreads; }

// One thread per ray: the ray's first segment and its DDA state at the occupied box (prepare_ray), 64 bytes
// per ray, for a march_kernel launched right behind it with `prepped` set.
__global__ void __launch_bounds__(256)
march_prepare_kernel(const RayArgs A, const GridSpec Gp, bool clip, uint4* __restrict__ out) {
    const long long r = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (r >= A.n_rays) return;
    GridSpec G = Gp;
    G.n_grids = 1;  // (only launched for one level: the event arrays then unroll into registers)
    uint32_t w[16];
    prepare_ray(A.rays_o, A.rays_d, A.near_arr, A.near_s, A.far_arr, A.far_s, A.t_sorted, A.t_indices, A.hits, G, r, clip, w);
    out[4 * r] = make_uint4(w[0], w[1], w[2], w[3]);
    out[4 * r + 1] = make_uint4(w[4], w[5], w[6], w[7]);
    out[4 * r + 2] = make_uint4(w[8], w[9], w[10], w[11]);
    out[4 * r + 3] = make_uint4(w[12], w[13], w[14], w[15]);
}

template <bool MIP, bool DIRECT, class Policy, bool SINGLE = false, bool PREPPED = false>
__global__ void __launch_bounds__(march_threads<DIRECT>(), 2)
march_kernel(const RayArgs A, const GridSpec Gp, const MarchOpts opt, Policy P,
             unsigned long long* ray_counter, const uint4* __restrict__ prepped) {
    extern __shared__ uint32_t s_dyn[];
    GridSpec G = stage_mip(Gp, s_dyn);
    if (SINGLE) G.n_grids = 1;
    float* s_evt = reinterpret_cast<float*>(s_dyn + ((Gp.mip_words + 3) & ~3));
    uint32_t* s_pool = reinterpret_cast<uint32_t*>(s_evt + kQ * march_threads<DIRECT>());
    float* s_ck = reinterpret_cast<float*>(s_pool + kPoolWords * march_threads<DIRECT>());
    int* s_nck = reinterpret_cast<int*>(s_ck + kMaxCk);
    // constant dt from one near plane: all rays walk one t sequence, checkpointed once per CTA
    if (threadIdx.x == 0) {
        const bool common = !DIRECT && opt.step_size > 0.0f && opt.cone_angle == 0.0f && A.near_arr == nullptr;
        *s_nck = common ? make_checkpoints(A.near_s, opt.step_size, s_ck) : 0;
    }
    __syncthreads();
    MarchOpts o = opt;
    if (A.limit_dev) o.limit = *A.limit_dev;
    const long long n_work = A.n_work_dev ? (long long)*A.n_work_dev : (A.ray_list ? A.n_work : A.n_rays);
    march_engine<MIP, DIRECT, Policy, PREPPED>(A.rays_o, A.rays_d, A.near_arr, A.near_s, A.far_arr, A.far_s, A.t_sorted, A.t_indices,
                 A.hits, A.ray_list, n_work, G, o, P, s_evt, s_pool, s_ck, *s_nck,
                 ray_counter, prepped);
}

// Pass 2 of OccGridEstimator.sampling when pass 1 left a run buffer: no marching at all.
// One warp owns 32 consecutive rays, whose samples are one contiguous stretch of the packed
// output.  The lanes list their rays' runs in shared memory in sample order; then the warp
// walks the OUTPUT: each lane takes four consecutive samples of a 16-byte aligned group, finds
// their run by binary search in the list and writes ray_indices / t_starts / t_ends as 128-bit
// stores — full 512-byte warp transactions whatever the run lengths are.
// Runs beyond a ray's inline slots are expanded from the pool by traverse_expand_pool_kernel;
// rays flagged kRunSlots + 1 (pool full) are left to a march_kernel launch over a work list.
constexpr int kExpandWarps = 8;
constexpr int kExpandRuns = 32 * kRunSlots;  // inline runs of a warp's 32 rays
__global__ void __launch_bounds__(kExpandWarps * 32)
traverse_expand_kernel(long long n_rays, const int64_t* __restrict__ sample_info,
                       const Run* __restrict__ runs, const int* __restrict__ n_runs,
                       int64_t* __restrict__ ray_indices, float* __restrict__ t_starts,
                       float* __restrict__ t_ends, bool vec_ok, long long capacity) {
    // (capacity: the outputs hold that many samples — a speculative launch, made before the host knows
    // the total, must not write past them; the caller re-runs with exact buffers if the total is larger)
    // the warp's runs in sample order: first sample (relative to the warp's first), length, t0, q, ray
    __shared__ int s_start[kExpandWarps][kExpandRuns];
    __shared__ int s_n[kExpandWarps][kExpandRuns];
    __shared__ float s_t0[kExpandWarps][kExpandRuns];
    __shared__ float s_q[kExpandWarps][kExpandRuns];
    __shared__ uint8_t s_ray[kExpandWarps][kExpandRuns];
    const int lane = lane_id(), w = warp_id();
    const long long n_groups = (n_rays + 31) >> 5;
    for (long long g = (long long)blockIdx.x * kExpandWarps + w; g < n_groups;
         g += (long long)gridDim.x * kExpandWarps) {
        const long long r = g * 32 + lane;
        long long start = 0, cnt = 0;
        int nr = 0;
        if (r < n_rays) {
            const RayChunk c = load_chunk(sample_info, r);
            start = c.start; cnt = c.cnt;
            nr = __ldg(n_runs + r);
        }
        const long long base0 = __shfl_sync(kFull, start, 0);
        const int span = __reduce_max_sync(kFull, (int)(start + cnt - base0));  // the warp's samples: [base0, base0 + span)
        const int mine = nr <= kRunSlots ? nr : 0;
        int pos = mine;  // exclusive prefix of the run counts
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) {
            const int y = __shfl_up_sync(kFull, pos, d);
            if (lane >= d) pos += y;
        }
        const int total = __shfl_sync(kFull, pos, 31);
        pos -= mine;
        for (int s = 0; s < mine; ++s) {
            const int4 v = __ldg(reinterpret_cast<const int4*>(runs + r * kRunSlots + s));
            s_t0[w][pos + s] = __int_as_float(v.x);
            s_q[w][pos + s] = __int_as_float(v.y);
            s_n[w][pos + s] = v.z;
            s_start[w][pos + s] = (int)(start - base0) + v.w;
            s_ray[w][pos + s] = (uint8_t)lane;
        }
        __syncwarp();
        if (total > 0) {
            // Every lane takes four consecutive samples of a 16-byte aligned group and finds the run
            // its first one lies in by binary search; samples of rays whose runs are pooled or
            // re-marched lie in no listed run and are left to those passes.
            const long long first = base0 & ~3ll;
            int kbase = 0;  // last run starting at or before the chunk's first sample (warp-uniform)
            for (long long c0 = first; c0 < base0 + span; c0 += 128) {  // (warp-uniform trip count)
                const long long i0 = c0 + 4 * lane;
                const int rel = (int)(i0 - base0);  // >= -3
                const int e0 = rel < 0 ? 0 : rel;
                // runs are sorted by start: move kbase up to the chunk, 32 candidates at a time;
                // each lane then walks on from there (a chunk of 128 samples meets a handful of runs)
                const int cs = (int)(c0 - base0) < 0 ? 0 : (int)(c0 - base0);
                while (true) {
                    const int kk = kbase + 1 + lane;
                    const unsigned m = __ballot_sync(kFull, kk < total && s_start[w][kk] <= cs);
                    kbase += __popc(m);
                    if (m != kFull) break;
                }
                int k = s_start[w][kbase] <= e0 ? kbase : -1;
                if (i0 >= base0 + span) continue;
                long long ri[4];
                float ts[4], te[4];
                unsigned ok = 0u;
#pragma unroll
                for (int c = 0; c < 4; ++c) {
                    const int e = rel + c;
                    while (k + 1 < total && s_start[w][k + 1] <= e) ++k;
                    ri[c] = 0; ts[c] = te[c] = 0.0f;
                    if (e >= 0 && e < span && k >= 0) {
                        const int j = e - s_start[w][k];
                        if (j < s_n[w][k]) {
                            const float t0 = s_t0[w][k], q = s_q[w][k];
                            ri[c] = g * 32 + s_ray[w][k];
                            ts[c] = __fmaf_rn((float)j, q, t0);
                            te[c] = __fmaf_rn((float)(j + 1), q, t0);
                            ok |= 1u << c;
                        }
                    }
                }
                if (i0 + 4 > capacity) {
#pragma unroll
                    for (int c = 0; c < 4; ++c)
                        if (i0 + c >= capacity) ok &= ~(1u << c);
                }
                if (ok == 0xfu && vec_ok) {
                    *reinterpret_cast<longlong2*>(ray_indices + i0) = make_longlong2(ri[0], ri[1]);
                    *reinterpret_cast<longlong2*>(ray_indices + i0 + 2) = make_longlong2(ri[2], ri[3]);
                    stg_stream4(t_starts + i0, make_float4(ts[0], ts[1], ts[2], ts[3]));
                    stg_stream4(t_ends + i0, make_float4(te[0], te[1], te[2], te[3]));
                } else {
#pragma unroll
                    for (int c = 0; c < 4; ++c) {
                        if ((ok >> c) & 1u) {
                            ray_indices[i0 + c] = ri[c];
                            t_starts[i0 + c] = ts[c];
                            t_ends[i0 + c] = te[c];
                        }
                    }
                }
            }
        }
        __syncwarp();
    }
}

// The pooled runs (rays with more than kRunSlots runs): one warp per run.
__global__ void __launch_bounds__(kExpandWarps * 32)
traverse_expand_pool_kernel(const int64_t* __restrict__ sample_info, const RunPool pool,
                            int64_t* __restrict__ ray_indices, float* __restrict__ t_starts,
                            float* __restrict__ t_ends, long long capacity) {
    unsigned long long cnt = *pool.cursor;
    if (cnt > pool.cap) cnt = pool.cap;
    const int lane = lane_id();
    for (unsigned long long e = (unsigned long long)blockIdx.x * kExpandWarps + warp_id(); e < cnt;
         e += (unsigned long long)gridDim.x * kExpandWarps) {
        const int4 raw = __ldg(reinterpret_cast<const int4*>(pool.runs + e));
        const Run run = *reinterpret_cast<const Run*>(&raw);
        const long long rb = __ldg(pool.ray + e);
        long long i = load_chunk(sample_info, rb).start + run.s0 + lane;
        for (int j = lane; j < run.n && i < capacity; j += 32, i += 32) {
            ray_indices[i] = rb;
            t_starts[i] = __fmaf_rn((float)j, run.q, run.t0);
            t_ends[i] = __fmaf_rn((float)(j + 1), run.q, run.t0);
        }
    }
}

// Pass 2 of OccGridEstimator.sampling when pass 1 left a piece log (cone stepping): a warp takes 32
// records, every lane replays its piece's recurrence (<= kPieceLen steps, the arithmetic of the march:
// t_next = t + calc_dt(t)) into shared memory, and the warp then writes the 32 pieces row by row — each a
// contiguous stretch of ray_indices / t_starts / t_ends.  Chunks are read in whatever order the count
// pass's warps filled them; where a record lands depends only on its ray's offset and its s0.
constexpr int kConeExpandWarps = 8;
__global__ void __launch_bounds__(kConeExpandWarps * 32)
traverse_expand_cone_kernel(const int64_t* __restrict__ sample_info, const int4* __restrict__ log,
                            const int* __restrict__ chunk_fill, const unsigned long long* __restrict__ cursor,
                            int n_chunks, float step, float cone, int64_t* __restrict__ ray_indices,
                            float* __restrict__ t_starts, float* __restrict__ t_ends, long long capacity) {
    __shared__ float s_t[kConeExpandWarps][32][kPieceLen + 1];
    const int lane = lane_id(), w = warp_id();
    unsigned long long used = *cursor;
    if (used > (unsigned long long)n_chunks) used = (unsigned long long)n_chunks;
    constexpr int kGroups = kPieceChunk / 32;
    const long long n_groups = (long long)used * kGroups;
    // the records of group g (one per lane) and where their rays' samples start: two dependent global
    // loads, issued one group ahead of their use
    auto fetch = [&](long long g, int4& rec, long long& base, int& n) {
        n = 0; base = 0; rec = make_int4(0, 0, 0, 0);
        if (g >= n_groups) return;
        const int chunk = (int)(g / kGroups), first = (int)(g % kGroups) * 32;
        const int fill = __ldg(chunk_fill + chunk);
        if (first + lane < fill) {
            rec = __ldg(log + (long long)chunk * kPieceChunk + first + lane);
            n = rec.y;
            base = __ldg(reinterpret_cast<const long long*>(sample_info) + 2 * (long long)rec.w) + rec.z;
        }
    };
    const long long stride = (long long)gridDim.x * kConeExpandWarps;
    int4 rec_n; long long base_n; int n_n;
    long long g = (long long)blockIdx.x * kConeExpandWarps + w;
    fetch(g, rec_n, base_n, n_n);
    for (; g < n_groups; g += stride) {
        const int4 rec = rec_n;
        const long long base = base_n;
        const int n = n_n;
        fetch(g + stride, rec_n, base_n, n_n);
        if (__ballot_sync(kFull, n > 0) == 0u) continue;
        float t = __int_as_float(rec.x);
        const int rid = rec.w;
        int nmax = n;
#pragma unroll
        for (int d = 16; d > 0; d >>= 1) nmax = max(nmax, __shfl_xor_sync(kFull, nmax, d));
        float* row = s_t[w][lane];
        for (int j = 0; j <= nmax; ++j) {
            if (j <= n) row[j] = t;
            t = __fadd_rn(t, calc_dt(t, cone, step));
        }
        __syncwarp();
        for (int k = 0; k < 32; ++k) {
            const int nk = __shfl_sync(kFull, n, k);
            if (nk == 0) continue;
            const long long bk = __shfl_sync(kFull, base, k);
            const long long rk = __shfl_sync(kFull, rid, k);
            if (lane < nk && bk + lane < capacity) {
                ray_indices[bk + lane] = rk;
                t_starts[bk + lane] = s_t[w][k][lane];
                t_ends[bk + lane] = s_t[w][k][lane + 1];
            }
        }
        __syncwarp();
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

// "any occupied" mip: one thread per coarse cell ORs its (2^shift)^3 fine cells
__global__ void pack_mip_kernel(const uint8_t* __restrict__ binaries, int n_grids, int rx, int ry, int rz,
                                int shift, int mx, int my, int mz, uint32_t* __restrict__ mip) {
    const long long n_coarse = (long long)n_grids * mx * my * mz;
    const long long n_words = (n_coarse + 31) >> 5;
    const int lane = threadIdx.x & 31;
    const int B = 1 << shift;
    for (long long w = ((long long)blockIdx.x * blockDim.x + threadIdx.x) >> 5; w < n_words;
         w += ((long long)gridDim.x * blockDim.x) >> 5) {
        const long long c = w * 32 + lane;
        bool on = false;
        if (c < n_coarse) {
            const int cz = (int)(c % mz), cy = (int)((c / mz) % my), cx = (int)((c / ((long long)mz * my)) % mx);
            const int l = (int)(c / ((long long)mz * my * mx));
            for (int i = cx * B; i < min(cx * B + B, rx) && !on; ++i)
                for (int j = cy * B; j < min(cy * B + B, ry) && !on; ++j)
                    for (int k = cz * B; k < min(cz * B + B, rz); ++k)
                        if (binaries[(((long long)l * rx + i) * ry + j) * rz + k]) { on = true; break; }
        }
        const unsigned m = __ballot_sync(kFull, on);
        if (lane == 0) mip[w] = m;
    }
}

// Index box of the occupied cells of each level: bounds[l] = (x0,y0,z0,x1,y1,z1), from the bit
// field (one thread per 32-cell word, warp-reduced, then six atomics per warp that saw any).
__global__ void occ_bounds_init_kernel(int n_grids, int rx, int ry, int rz, int* __restrict__ bounds) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n_grids * 6) {
        const int k = i % 6;
        bounds[i] = k >= 3 ? -1 : (k == 0 ? rx : (k == 1 ? ry : rz));
    }
}
__global__ void occ_bounds_kernel(const uint32_t* __restrict__ bits, int n_grids, int rx, int ry, int rz,
                                  int* __restrict__ bounds) {
    const long long per_level = (long long)rx * ry * rz;
    const long long n_cells = per_level * n_grids;
    const long long n_words = (n_cells + 31) >> 5;
    for (long long w0 = ((long long)blockIdx.x * blockDim.x + threadIdx.x) & ~31ll; w0 < n_words;
         w0 += (long long)gridDim.x * blockDim.x) {
        const long long w = w0 + (threadIdx.x & 31);
        uint32_t word = w < n_words ? __ldg(bits + w) : 0u;
        // a warp's 1024 cells may straddle levels: take them one level at a time
        while (true) {
            int level = 0x7fffffff;
            if (word) {
                const long long c0 = w * 32 + (__ffs(word) - 1);
                if (c0 < n_cells) level = (int)(c0 / per_level);
            }
            level = __reduce_min_sync(kFull, level);
            if (level == 0x7fffffff) break;
            int lo[3] = {rx, ry, rz}, hi[3] = {-1, -1, -1};
            uint32_t rest = 0u;
            while (word) {
                const int b = __ffs(word) - 1;
                word &= word - 1u;
                const long long c = w * 32 + b;
                if (c >= n_cells) continue;
                if ((int)(c / per_level) != level) { rest |= 1u << b; continue; }
                const long long q = c - (long long)level * per_level;
                const int z = (int)(q % rz), y = (int)((q / rz) % ry), x = (int)(q / ((long long)rz * ry));
                lo[0] = min(lo[0], x); lo[1] = min(lo[1], y); lo[2] = min(lo[2], z);
                hi[0] = max(hi[0], x); hi[1] = max(hi[1], y); hi[2] = max(hi[2], z);
            }
            word = rest;
#pragma unroll
            for (int k = 0; k < 3; ++k) {
                lo[k] = __reduce_min_sync(kFull, lo[k]);
                hi[k] = __reduce_max_sync(kFull, hi[k]);
            }
            if ((threadIdx.x & 31) == 0 && hi[0] >= 0) {
#pragma unroll
                for (int k = 0; k < 3; ++k) {
                    atomicMin(bounds + 6 * level + k, lo[k]);
                    atomicMax(bounds + 6 * level + 3 + k, hi[k]);
                }
            }
        }
    }
}

// ---------------------------------------------------------------------------------- chunked loop
// The test-time loop of the reference's callers (examples/utils.py:252-425) as three launches per
// iteration around the user's radiance field, with the loop state (alive list, per-ray near
// planes, the step limit of the next chunk) kept on the device:
//   march_kernel (FillPolicy<false>, single pass)  the alive rays march up to k samples from their
//        termination planes into k-sample slots (k, the list and its length are device values);
//   chunk_compact_kernel    slots -> the dense, ray-ordered (ray_indices, t_starts, t_ends) the
//        callback sees (what the reference gets from three boolean-mask selects);
//   chunk_composite_kernel  weights with prefix_trans = 1 - opacity (volrend.py:312-362), the
//        alpha_thre filter, rgb / opacity / depth accumulated in place (accumulate_along_rays_ x3),
//        then the ray's fate: alive iff opacity <= 1 - early_stop_eps and it used its whole chunk
//        (examples/utils.py:404-411) -> next alive list, its slot number, or a zeroed count;
//   chunk_plan_kernel       k of the next chunk from the number of rays alive (utils.py:337).
struct ChunkState {
    int n_alive;   // length of the current alive list
    int limit;     // k: samples per ray of the current chunk
    int n_next;    // rays appended to the next list so far
    int pad;
};

__global__ void __launch_bounds__(256)
chunk_compact_kernel(const int* __restrict__ alive, const ChunkState* __restrict__ st,
                     const int64_t* __restrict__ sample_info, const float* __restrict__ slot_ts,
                     const float* __restrict__ slot_te, int64_t* __restrict__ ray_indices,
                     float* __restrict__ t_starts, float* __restrict__ t_ends) {
    const int lane = lane_id();
    const int n_alive = st->n_alive, k = st->limit;
    for (long long j = (long long)blockIdx.x * 8 + warp_id(); j < n_alive; j += (long long)gridDim.x * 8) {
        const int r = __ldg(alive + j);
        const RayChunk c = load_chunk(sample_info, r);
        const long long src = j * (long long)k;
        for (int i = lane; i < c.cnt; i += 32) {
            ray_indices[c.start + i] = r;
            t_starts[c.start + i] = __ldg(slot_ts + src + i);
            t_ends[c.start + i] = __ldg(slot_te + src + i);
        }
    }
}

struct ChunkCompositeArgs {
    const int* alive; int* alive_next; ChunkState* st;
    int64_t* sample_info; int64_t* slot_of;
    const float* t_starts; const float* t_ends; const float* sigmas; const float* rgbs;
    float* rgb; float* opacity; float* depth;
    float alpha_thre, opc_thre;
};

__global__ void __launch_bounds__(256) chunk_composite_kernel(const ChunkCompositeArgs p) {
    const int lane = lane_id();
    const int n_alive = p.st->n_alive, k = p.st->limit;
    for (long long j = (long long)blockIdx.x * 8 + warp_id(); j < n_alive; j += (long long)gridDim.x * 8) {
        const int r = __ldg(p.alive + j);
        const RayChunk c = load_chunk(p.sample_info, r);
        const float opac0 = p.opacity[r];
        const float prefix = __fsub_rn(1.0f, opac0);  // prefix_trans = 1 - opacity[ray_indices]
        float carry = 0.0f;  // sum of sigma * delta before this round of 32 samples
        float a_r = 0.0f, a_g = 0.0f, a_b = 0.0f, a_o = 0.0f, a_d = 0.0f;
        for (int i0 = 0; i0 < c.cnt; i0 += 32) {
            const int i = i0 + lane;
            const bool on = i < c.cnt;
            const long long s = c.start + i;
            const float ts = on ? __ldg(p.t_starts + s) : 0.0f, te = on ? __ldg(p.t_ends + s) : 0.0f;
            const float sd = on ? __fmul_rn(__ldg(p.sigmas + s), __fsub_rn(te, ts)) : 0.0f;
            float x = sd;
#pragma unroll
            for (int d = 1; d < 32; d <<= 1) {
                const float y = __shfl_up_sync(kFull, x, d);
                if (lane >= d) x = __fadd_rn(y, x);
            }
            const float incl = __fadd_rn(carry, x);
            float excl = __shfl_up_sync(kFull, incl, 1);
            if (lane == 0) excl = carry;
            carry = __shfl_sync(kFull, incl, 31);
            const float alpha = __fsub_rn(1.0f, expf(-sd));
            const float T = __fmul_rn(expf(-excl), prefix);
            float wgt = __fmul_rn(T, alpha);
            if (!on || (p.alpha_thre > 0.0f && !(alpha >= p.alpha_thre))) wgt = 0.0f;  // utils.py:382-390
            float v[5];
            v[0] = on ? __fmul_rn(wgt, __ldg(p.rgbs + 3 * s)) : 0.0f;
            v[1] = on ? __fmul_rn(wgt, __ldg(p.rgbs + 3 * s + 1)) : 0.0f;
            v[2] = on ? __fmul_rn(wgt, __ldg(p.rgbs + 3 * s + 2)) : 0.0f;
            v[3] = wgt;
            v[4] = __fmul_rn(wgt, __fmul_rn(__fadd_rn(ts, te), 0.5f));  // (t_starts + t_ends) / 2
#pragma unroll
            for (int q = 0; q < 5; ++q) {
#pragma unroll
                for (int d = 16; d > 0; d >>= 1) v[q] = __fadd_rn(v[q], __shfl_xor_sync(kFull, v[q], d));
            }
            a_r = __fadd_rn(a_r, v[0]); a_g = __fadd_rn(a_g, v[1]); a_b = __fadd_rn(a_b, v[2]);
            a_o = __fadd_rn(a_o, v[3]); a_d = __fadd_rn(a_d, v[4]);
        }
        if (lane == 0) {
            const float opac = __fadd_rn(opac0, a_o);
            p.rgb[3 * r] = __fadd_rn(p.rgb[3 * r], a_r);
            p.rgb[3 * r + 1] = __fadd_rn(p.rgb[3 * r + 1], a_g);
            p.rgb[3 * r + 2] = __fadd_rn(p.rgb[3 * r + 2], a_b);
            p.opacity[r] = opac;
            p.depth[r] = __fadd_rn(p.depth[r], a_d);
            const bool alive = (opac <= p.opc_thre) && (c.cnt == k);  // utils.py:404-411
            if (alive) {
                const int pos = atomicAdd(&p.st->n_next, 1);
                p.alive_next[pos] = r;
                p.slot_of[r] = pos;
            } else {
                p.sample_info[2 * r + 1] = 0;  // retired: contributes nothing to later chunks' scans
            }
        }
    }
}

__global__ void chunk_seed_kernel(ChunkState* st, int n_rays) {
    st->n_alive = 0; st->limit = 0; st->n_next = n_rays; st->pad = 0;
}

__global__ void chunk_plan_kernel(ChunkState* st, long long n_rays, int min_samples, int64_t* plan_host) {
    const int n_alive = st->n_next;
    int k = (int)(n_alive > 0 ? n_rays / n_alive : 64);
    k = k < 64 ? k : 64;
    k = k > min_samples ? k : min_samples;  // utils.py:337
    st->n_alive = n_alive;
    st->limit = k;
    st->n_next = 0;
    if (plan_host) { plan_host[0] = n_alive; plan_host[1] = k; }
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
constexpr long long kMaxMipBytes = 128 * 1024;
constexpr size_t kCounterBytes = 64;  // ray counter + overflow counter, at the end of `scratch`

// Persistent launch: one CTA per SM (the engine is register-heavy), dynamic shared memory
// for the mip and the event queue; zeroes the ray counter first.
template <class Policy>
static int launch_march(const RayArgs& A, const GridSpec& G, const MarchOpts& opt, const Policy& P,
                        unsigned long long* counter, cudaStream_t s, uint4* prep_buf = nullptr, bool clip = false) {
    // cone stepping (dt follows t) folds nothing: events are applied in place, without the queue
    // (measured: through the queue a cfg4 march takes 43 ms against 24 ms applied in place)
    const bool direct = opt.step_size > 0.0f && opt.cone_angle != 0.0f;
    const int threads = direct ? kConeThreads : kMarchThreads;
    const size_t smem = (size_t)((G.mip_words + 3) & ~3) * 4 + (size_t)kQ * threads * sizeof(float) +
                        (size_t)kPoolWords * threads * sizeof(uint32_t) + (kMaxCk + 4) * sizeof(float);
    auto kernel = G.mip != nullptr ? (direct ? march_kernel<true, true, Policy> : march_kernel<true, false, Policy>)
                                   : (direct ? march_kernel<false, true, Policy> : march_kernel<false, false, Policy>);
    if (G.n_grids == 1)
        kernel = G.mip != nullptr
                     ? (direct ? march_kernel<true, true, Policy, true> : march_kernel<true, false, Policy, true>)
                     : (direct ? march_kernel<false, true, Policy, true> : march_kernel<false, false, Policy, true>);
    if (cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem) != cudaSuccess) {
        set_error("cannot reserve %zu bytes of shared memory for the marching kernel", smem);
        return NFB_ERR_CUDA;
    }
    if (cudaMemsetAsync(counter, 0, sizeof(unsigned long long), s) != cudaSuccess) {
        set_error("ray counter memset failed");
        return NFB_ERR_CUDA;
    }
    // a warp takes 32 rays at a time: one CTA per 32 rays until the machine is full, so that
    // small launches (a re-march list, a test-time chunk) spread over all SMs
    long long b = ceil_div<long long>(A.ray_list ? A.n_work : A.n_rays, 32);
    const long long cap = (long long)sm_count() * (smem <= 112 * 1024 ? 2 : 1);
    if (b > cap) b = cap;
    if (b < 1) b = 1;
    // Rays are handed out dynamically, `batch` per warp at a time: 32 when there is work for
    // many rounds, fewer for small launches so that the last round does not leave most warps idle
    // behind the few that drew a second batch.
    MarchOpts o2 = opt;
    {
        const long long n_work = A.ray_list ? A.n_work : A.n_rays;
        const long long warps = b * (threads / 32);
        // (the cone loop ties a lane to a ray: one round that fills the lanes beats two half-empty ones)
        long long per = ceil_div<long long>(n_work, direct ? warps : warps * 2);
        o2.batch = (int)(per < 8 ? 8 : (per > 32 ? 32 : per));
    }
    const uint4* prepped = nullptr;
    bool use_prep = false;
    if constexpr (std::is_same<Policy, CountPolicy<1>>::value)
        use_prep = prep_buf != nullptr && !direct && G.n_grids == 1 && A.ray_list == nullptr && A.n_work_dev == nullptr &&
                   A.n_rays > 0;
    if (use_prep) {
        if constexpr (std::is_same<Policy, CountPolicy<1>>::value)
            kernel = G.mip != nullptr ? march_kernel<true, false, Policy, true, true> : march_kernel<false, false, Policy, true, true>;
        if (cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem) != cudaSuccess) {
            set_error("cannot reserve %zu bytes of shared memory for the marching kernel", smem);
            return NFB_ERR_CUDA;
        }
        march_prepare_kernel<<<(unsigned)ceil_div<long long>(A.n_rays, 256), 256, 0, s>>>(A, G, clip, prep_buf);
        count_launch();
        NFB_CHECK_LAUNCH("march_prepare_kernel");
        prepped = prep_buf;
    }
    kernel<<<(unsigned)b, threads, smem, s>>>(A, G, o2, P, counter, prepped);
    count_launch();
    NFB_CHECK_LAUNCH("march_kernel");
    return NFB_OK;
}

static int check_common(const nfb_rays_t* rays, const nfb_grid_t* grid, const nfb_march_t* march) {
    NFB_CHECK_ARG(rays && grid && march, "null descriptor");
    NFB_CHECK_ARG(rays->n_rays >= 0, "negative n_rays");
    NFB_CHECK_ARG(grid->n_grids >= 1 && grid->n_grids <= kMaxLevels, "n_grids must be in [1, %d]", kMaxLevels);
    NFB_CHECK_ARG(grid->res[0] > 0 && grid->res[1] > 0 && grid->res[2] > 0, "bad resolution");
    NFB_CHECK_ARG((long long)grid->n_grids * grid->res[0] * grid->res[1] * grid->res[2] < (1ll << 31),
                  "grids with 2^31 or more cells are not supported");
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
    a.ray_list = nullptr; a.n_work = 0; a.n_work_dev = nullptr; a.limit_dev = nullptr;
    return a;
}
static GridSpec to_grid(const nfb_grid_t* g) {
    GridSpec G;
    G.occ_bits = g->occ_bits; G.aabbs = g->aabbs; G.n_grids = g->n_grids;
    G.res[0] = g->res[0]; G.res[1] = g->res[1]; G.res[2] = g->res[2];
    G.cells_per_level = (long long)g->res[0] * g->res[1] * g->res[2];
    G.bounds = g->occ_bounds;
    G.mip = nullptr; G.mip_shift = 0; G.mres[0] = G.mres[1] = G.mres[2] = 0;
    G.mip_cells_per_level = 0; G.mip_words = 0;
    if (g->occ_mip != nullptr && g->mip_shift > 0) {
        const int sh = g->mip_shift;
        long long mc = 1;
        for (int k = 0; k < 3; ++k) { G.mres[k] = (g->res[k] + (1 << sh) - 1) >> sh; mc *= G.mres[k]; }
        const long long words = (mc * g->n_grids + 31) / 32;
        if (words * 4 <= kMaxMipBytes) {
            G.mip = g->occ_mip; G.mip_shift = sh; G.mip_cells_per_level = (int)mc; G.mip_words = (int)words;
        }
    }
    return G;
}
static MarchOpts to_opts(const nfb_march_t* m) { return MarchOpts{m->step_size, m->cone_angle, m->traverse_steps_limit, 32}; }

static size_t run_ints_bytes(size_t n) { return (2 * n * sizeof(int) + 15) & ~(size_t)15; }
// the pieces of a run buffer of n rays (see nfb_traverse_run_bytes); cursor = counters + 2
static RunPool pool_of(void* run_buf, long long n, unsigned long long* cursor) {
    char* base = static_cast<char*>(run_buf) + (size_t)n * kRunSlots * sizeof(Run) + run_ints_bytes((size_t)n);
    RunPool P;
    P.runs = reinterpret_cast<Run*>(base);
    P.cap = (unsigned long long)n * kPoolPerRay;
    P.ray = reinterpret_cast<int*>(base + P.cap * sizeof(Run));
    P.cursor = cursor;
    return P;
}

// the piece log of n rays (see nfb_traverse_piece_bytes): chunk_fill ints (padded to 16 B) | records
static int piece_chunks(long long n) {
    // kPiecesPerRay records per ray on average + one partly used chunk per marching warp
    const long long warps = (long long)sm_count() * 2 * (kConeThreads / 32);
    long long c = (n * kPiecesPerRay + kPieceChunk - 1) / kPieceChunk + (n / 32 + 1 < warps ? n / 32 + 1 : warps) + 1;
    return (int)(c < (1ll << 30) ? c : (1ll << 30));
}
static size_t piece_fill_bytes(int n_chunks) { return ((size_t)n_chunks * sizeof(int) + 15) & ~(size_t)15; }
static PieceLog pieces_of(void* buf, long long n, unsigned long long* counters) {
    PieceLog L;
    L.n_chunks = piece_chunks(n);
    L.chunk_fill = static_cast<int*>(buf);
    L.log = reinterpret_cast<int4*>(static_cast<char*>(buf) + piece_fill_bytes(L.n_chunks));
    L.failed = counters + 1;
    L.cursor = counters + 2;
    return L;
}

static uint4* prepared_of(void* scratch, long long n_rays) {
    const size_t one = (ray_scan_status_bytes(n_rays) + 63) & ~(size_t)63;
    return reinterpret_cast<uint4*>(static_cast<char*>(scratch) + 2 * one);
}

static unsigned long long* counters_of(void* scratch, size_t scratch_bytes, long long n_rays) {
    const size_t need = nfb_traverse_scratch_bytes(n_rays);
    if (scratch == nullptr || scratch_bytes < need) return nullptr;
    return reinterpret_cast<unsigned long long*>(static_cast<char*>(scratch) + need - kCounterBytes);
}

static int scan_both(int64_t* sample_info, int64_t* interval_info, long long n_rays, int64_t* totals,
                     void* scratch, cudaStream_t s, const unsigned long long* also_src = nullptr,
                     long long* also_dst = nullptr) {
    const size_t one = (ray_scan_status_bytes(n_rays) + 63) & ~(size_t)63;
    int rc = launch_ray_scan(sample_info, n_rays, reinterpret_cast<long long*>(totals), scratch, s, nullptr,
                             also_src, also_dst);
    if (rc != NFB_OK) return rc;
    if (interval_info != nullptr)
        rc = launch_ray_scan(interval_info, n_rays, totals ? reinterpret_cast<long long*>(totals) + 1 : nullptr,
                             static_cast<char*>(scratch) + one, s);
    return rc;
}
}  // namespace nfb

extern "C" {

NFB_API size_t nfb_traverse_scratch_bytes(int64_t n_rays) {
    // two ray-scan status arrays | prepared-ray records (64 B per ray) | counters
    const size_t one = (nfb::ray_scan_status_bytes(n_rays) + 63) & ~(size_t)63;
    return 2 * one + (size_t)(n_rays > 0 ? n_rays : 0) * 64 + nfb::kCounterBytes;
}

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

NFB_API int nfb_pack_occupancy(const uint8_t* binaries, int32_t n_grids, const int32_t* res_host,
                               uint32_t* bits, uint32_t* mip_bits, int32_t mip_shift,
                               nfb_stream_t stream) {
    using namespace nfb;
    NFB_CHECK_ARG(n_grids >= 0 && res_host != nullptr, "bad grid shape");
    const long long n_cells = (long long)n_grids * res_host[0] * res_host[1] * res_host[2];
    if (n_cells == 0) return NFB_OK;
    NFB_CHECK_ARG(binaries && bits, "null pointer");
    cudaStream_t s = static_cast<cudaStream_t>(stream);
    pack_bits_kernel<<<grid_for(n_cells, 256), 256, 0, s>>>(binaries, n_cells, bits);
    count_launch();
    if (mip_bits != nullptr) {
        NFB_CHECK_ARG(mip_shift >= 1 && mip_shift <= 8, "mip_shift must be in [1, 8]");
        const int B = 1 << mip_shift;
        const int mx = (res_host[0] + B - 1) / B, my = (res_host[1] + B - 1) / B, mz = (res_host[2] + B - 1) / B;
        pack_mip_kernel<<<grid_for((long long)n_grids * mx * my * mz, 256), 256, 0, s>>>(
            binaries, n_grids, res_host[0], res_host[1], res_host[2], mip_shift, mx, my, mz, mip_bits);
        count_launch();
    }
    NFB_CHECK_LAUNCH("pack_occupancy");
    return NFB_OK;
}

NFB_API int nfb_occupancy_bounds(const uint32_t* bits, int32_t n_grids, const int32_t* res_host,
                                 int32_t* bounds, nfb_stream_t stream) {
    using namespace nfb;
    NFB_CHECK_ARG(n_grids >= 0 && res_host != nullptr, "bad grid shape");
    if (n_grids == 0) return NFB_OK;
    NFB_CHECK_ARG(bits && bounds, "null pointer");
    cudaStream_t s = static_cast<cudaStream_t>(stream);
    const long long n_cells = (long long)n_grids * res_host[0] * res_host[1] * res_host[2];
    occ_bounds_init_kernel<<<1, 64, 0, s>>>(n_grids, res_host[0], res_host[1], res_host[2], bounds);
    if (n_cells > 0)
        occ_bounds_kernel<<<grid_for((n_cells + 31) / 32, 256), 256, 0, s>>>(bits, n_grids, res_host[0], res_host[1],
                                                                             res_host[2], bounds);
    count_launch(2);
    NFB_CHECK_LAUNCH("occupancy_bounds");
    return NFB_OK;
}

NFB_API size_t nfb_traverse_run_bytes(int64_t n_rays) {
    const size_t n = (size_t)(n_rays > 0 ? n_rays : 0);
    // inline runs | n_runs, overflow list (ints, padded to 16 B) | pooled runs | their ray ids
    return n * nfb::kRunSlots * sizeof(nfb::Run) + nfb::run_ints_bytes(n) +
           n * nfb::kPoolPerRay * (sizeof(nfb::Run) + sizeof(int)) + 16;
}

NFB_API size_t nfb_traverse_piece_bytes(int64_t n_rays) {
    const long long n = n_rays > 0 ? n_rays : 0;
    const int c = nfb::piece_chunks(n);
    return nfb::piece_fill_bytes(c) + (size_t)c * nfb::kPieceChunk * sizeof(int4) + 16;
}

NFB_API int nfb_traverse_count(const nfb_rays_t* rays, const nfb_grid_t* grid,
                               const nfb_march_t* march, int64_t* sample_info,
                               int64_t* interval_info, float* termination_planes,
                               int64_t* totals, void* scratch, size_t scratch_bytes,
                               void* run_buf, size_t run_buf_bytes, nfb_stream_t stream) {
    using namespace nfb;
    int rc = check_common(rays, grid, march);
    if (rc != NFB_OK) return rc;
    cudaStream_t s = static_cast<cudaStream_t>(stream);
    if (rays->n_rays == 0) {
        if (totals) cudaMemsetAsync(totals, 0, 3 * sizeof(int64_t), s);
        return NFB_OK;
    }
    NFB_CHECK_ARG(sample_info != nullptr, "null sample_info");
    unsigned long long* counters = counters_of(scratch, scratch_bytes, rays->n_rays);
    NFB_CHECK_ARG(counters != nullptr, "scratch too small: need %zu bytes", nfb_traverse_scratch_bytes(rays->n_rays));
    RayArgs A = to_args(rays);
    A.rays_mask = nullptr;  // the two-pass path ignores rays_mask (grid.cu:418,450)
    const GridSpec G = to_grid(grid);
    const bool want_runs = run_buf != nullptr && march->step_size > 0.0f && march->cone_angle == 0.0f;
    const bool want_pieces = run_buf != nullptr && march->step_size > 0.0f && march->cone_angle != 0.0f;
    if (want_pieces) {
        NFB_CHECK_ARG(run_buf_bytes >= nfb_traverse_piece_bytes(rays->n_rays), "piece log too small");
        NFB_CHECK_ARG(aligned16(run_buf), "piece log must be 16-byte aligned");
        CountPolicy<2> P;
        P.sample_info = sample_info; P.interval_info = interval_info; P.term = termination_planes;
        P.runs = nullptr; P.n_runs = nullptr; P.overflow_list = nullptr; P.n_overflow = nullptr; P.pool = RunPool{};
        P.pieces = pieces_of(run_buf, rays->n_rays, counters);
        cudaMemsetAsync(counters + 1, 0, 2 * sizeof(unsigned long long), s);
        cudaMemsetAsync(P.pieces.chunk_fill, 0, piece_fill_bytes(P.pieces.n_chunks), s);
        rc = launch_march(A, G, to_opts(march), P, counters, s);
    } else if (want_runs) {
        NFB_CHECK_ARG(run_buf_bytes >= nfb_traverse_run_bytes(rays->n_rays), "run buffer too small");
        NFB_CHECK_ARG(aligned16(run_buf), "run buffer must be 16-byte aligned");
        CountPolicy<1> P;
        P.sample_info = sample_info; P.interval_info = interval_info; P.term = termination_planes;
        P.runs = static_cast<Run*>(run_buf);
        P.n_runs = reinterpret_cast<int*>(P.runs + rays->n_rays * kRunSlots);
        P.overflow_list = P.n_runs + rays->n_rays;
        P.n_overflow = counters + 1;
        P.pool = pool_of(run_buf, rays->n_rays, counters + 2);
        P.pieces = PieceLog{};
        cudaMemsetAsync(counters + 1, 0, 2 * sizeof(unsigned long long), s);
        // (the engine's `clip`: constant dt folds empty cells; without termination planes the walk may stop
        // at the occupied box)
        rc = launch_march(A, G, to_opts(march), P, counters, s,
                          aligned16(scratch) ? prepared_of(scratch, rays->n_rays) : nullptr,
                          /*clip=*/termination_planes == nullptr);
    } else {
        CountPolicy<0> P;
        P.sample_info = sample_info; P.interval_info = interval_info; P.term = termination_planes;
        P.runs = nullptr; P.n_runs = nullptr; P.overflow_list = nullptr; P.n_overflow = nullptr; P.pool = RunPool{};
        P.pieces = PieceLog{};
        cudaMemsetAsync(counters + 1, 0, sizeof(unsigned long long), s);
        rc = launch_march(A, G, to_opts(march), P, counters, s);
    }
    if (rc != NFB_OK) return rc;
    // totals[2] (rays beyond the run buffer) goes out with the scan's own total, no copy on the stream
    return scan_both(sample_info, interval_info, rays->n_rays, totals, scratch, s,
                     totals != nullptr ? counters + 1 : nullptr,
                     totals != nullptr ? reinterpret_cast<long long*>(totals) + 2 : nullptr);
}

// the expansion of the run buffer: traverse_expand_kernel + the pooled runs
static int launch_expand(int64_t n_rays, const int64_t* sample_info, const void* run_buf, int64_t* ray_indices,
                         float* t_starts, float* t_ends, long long capacity, void* scratch, size_t scratch_bytes,
                         cudaStream_t s) {
    using namespace nfb;
    NFB_CHECK_ARG(aligned16(run_buf) && aligned16(sample_info), "run buffer / sample_info must be 16-byte aligned");
    const Run* runs = static_cast<const Run*>(run_buf);
    const int* n_runs = reinterpret_cast<const int*>(runs + n_rays * kRunSlots);
    const long long groups = (n_rays + 31) / 32;
    traverse_expand_kernel<<<grid_for(groups, kExpandWarps, 32), kExpandWarps * 32, 0, s>>>(
        n_rays, sample_info, runs, n_runs, ray_indices, t_starts, t_ends,
        aligned16(ray_indices) && aligned16(t_starts) && aligned16(t_ends), capacity);
    count_launch();
    NFB_CHECK_LAUNCH("traverse_expand");
    // pooled runs; the count stays on the device (the kernel reads the cursor)
    unsigned long long* ctr = counters_of(scratch, scratch_bytes, n_rays);
    NFB_CHECK_ARG(ctr != nullptr, "scratch too small: need %zu bytes", nfb_traverse_scratch_bytes(n_rays));
    const RunPool pool = pool_of(const_cast<void*>(run_buf), n_rays, ctr + 2);
    traverse_expand_pool_kernel<<<sm_count() * 4, kExpandWarps * 32, 0, s>>>(sample_info, pool, ray_indices, t_starts,
                                                                             t_ends, capacity);
    count_launch();
    NFB_CHECK_LAUNCH("traverse_expand_pool");
    return NFB_OK;
}

static int launch_expand_cone(int64_t n_rays, const int64_t* sample_info, const void* piece_buf, float step, float cone,
                              int64_t* ray_indices, float* t_starts, float* t_ends, long long capacity, void* scratch,
                              size_t scratch_bytes, cudaStream_t s) {
    using namespace nfb;
    NFB_CHECK_ARG(aligned16(piece_buf) && aligned16(sample_info), "piece log / sample_info must be 16-byte aligned");
    unsigned long long* ctr = counters_of(scratch, scratch_bytes, n_rays);
    NFB_CHECK_ARG(ctr != nullptr, "scratch too small: need %zu bytes", nfb_traverse_scratch_bytes(n_rays));
    const PieceLog L = pieces_of(const_cast<void*>(piece_buf), n_rays, ctr);
    // sized for the whole log; the kernel reads how many chunks were handed out
    const long long groups = (long long)L.n_chunks * (kPieceChunk / 32);
    traverse_expand_cone_kernel<<<grid_for(groups, kConeExpandWarps, 6), kConeExpandWarps * 32, 0, s>>>(
        sample_info, L.log, L.chunk_fill, L.cursor, L.n_chunks, step, cone, ray_indices, t_starts, t_ends, capacity);
    count_launch();
    NFB_CHECK_LAUNCH("traverse_expand_cone");
    return NFB_OK;
}

NFB_API int nfb_traverse_expand_cone(int64_t n_rays, const int64_t* sample_info, const void* piece_buf,
                                     float step_size, float cone_angle, int64_t capacity, int64_t* ray_indices,
                                     float* t_starts, float* t_ends, void* scratch, size_t scratch_bytes,
                                     nfb_stream_t stream) {
    using namespace nfb;
    NFB_CHECK_ARG(n_rays >= 0 && capacity >= 0, "negative size");
    if (n_rays == 0 || capacity == 0) return NFB_OK;
    NFB_CHECK_ARG(sample_info && piece_buf && ray_indices && t_starts && t_ends, "null pointer");
    NFB_CHECK_ARG(step_size > 0.0f && cone_angle != 0.0f, "the piece log belongs to cone stepping");
    return launch_expand_cone(n_rays, sample_info, piece_buf, step_size, cone_angle, ray_indices, t_starts, t_ends,
                              capacity, scratch, scratch_bytes, static_cast<cudaStream_t>(stream));
}

NFB_API int nfb_traverse_expand(int64_t n_rays, const int64_t* sample_info, const void* run_buf,
                                int64_t capacity, int64_t* ray_indices, float* t_starts, float* t_ends,
                                void* scratch, size_t scratch_bytes, nfb_stream_t stream) {
    using namespace nfb;
    NFB_CHECK_ARG(n_rays >= 0 && capacity >= 0, "negative size");
    if (n_rays == 0 || capacity == 0) return NFB_OK;
    NFB_CHECK_ARG(sample_info && run_buf && ray_indices && t_starts && t_ends, "null pointer");
    return launch_expand(n_rays, sample_info, run_buf, ray_indices, t_starts, t_ends, capacity, scratch, scratch_bytes,
                         static_cast<cudaStream_t>(stream));
}

NFB_API int nfb_traverse_fill_sampling(const nfb_rays_t* rays, const nfb_grid_t* grid,
                                       const nfb_march_t* march, const int64_t* sample_info,
                                       int64_t* ray_indices, float* t_starts, float* t_ends,
                                       const void* run_buf, int64_t n_overflowed, void* scratch,
                                       size_t scratch_bytes, nfb_stream_t stream) {
    using namespace nfb;
    int rc = check_common(rays, grid, march);
    if (rc != NFB_OK) return rc;
    if (rays->n_rays == 0) return NFB_OK;
    NFB_CHECK_ARG(sample_info != nullptr, "null sample_info");
    cudaStream_t s = static_cast<cudaStream_t>(stream);
    RayArgs A = to_args(rays);
    A.rays_mask = nullptr;
    if (run_buf != nullptr && march->step_size > 0.0f && march->cone_angle != 0.0f) {
        // cone stepping: the count pass left a complete piece log (a caller whose log ran full —
        // totals[2] != 0 — passes no run_buf and gets the march below)
        NFB_CHECK_ARG(n_overflowed == 0 || n_overflowed == -1, "a full piece log cannot be expanded: pass run_buf = NULL");
        if (n_overflowed < 0) return NFB_OK;  // nfb_traverse_expand_cone has run already
        return launch_expand_cone(rays->n_rays, sample_info, run_buf, march->step_size, march->cone_angle, ray_indices,
                                  t_starts, t_ends, (long long)1 << 62, scratch, scratch_bytes, s);
    }
    const bool have_runs = run_buf != nullptr && march->step_size > 0.0f && march->cone_angle == 0.0f;
    if (have_runs) {
        // n_overflowed < 0: nfb_traverse_expand has expanded the runs already; march only the
        // -n_overflowed - 1 rays that overflowed
        const bool expanded = n_overflowed < 0;
        if (expanded) n_overflowed = -n_overflowed - 1;
        else {
            rc = launch_expand(rays->n_rays, sample_info, run_buf, ray_indices, t_starts, t_ends,
                               (long long)1 << 62, scratch, scratch_bytes, s);
            if (rc != NFB_OK) return rc;
        }
        if (n_overflowed == 0) return NFB_OK;
        // the few rays with more than kRunSlots runs and no room in the pool: re-march exactly
        // those (the count pass left their ids in the overflow list)
        NFB_CHECK_ARG(n_overflowed > 0 && n_overflowed <= rays->n_rays, "bad n_overflowed");
        const Run* runs = static_cast<const Run*>(run_buf);
        A.ray_list = reinterpret_cast<const int*>(runs + rays->n_rays * kRunSlots) + rays->n_rays;
        A.n_work = n_overflowed;
    }
    unsigned long long* counters = counters_of(scratch, scratch_bytes, rays->n_rays);
    NFB_CHECK_ARG(counters != nullptr, "scratch too small: need %zu bytes", nfb_traverse_scratch_bytes(rays->n_rays));
    FillPolicy<false> P{};
    P.rays_mask = nullptr; P.near_arr = A.near_arr; P.near_s = A.near_s;
    P.sm_starts = sample_info; P.sm_stride = 2;
    P.sample_info = const_cast<int64_t*>(sample_info); P.interval_info = nullptr; P.single_pass = false;
    P.term = nullptr;
    P.proto.ray_indices = ray_indices; P.proto.t_starts = t_starts; P.proto.t_ends = t_ends;
    return launch_march(A, to_grid(grid), to_opts(march), P, counters, s);
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
    unsigned long long* counters = counters_of(scratch, scratch_bytes, rays->n_rays);
    NFB_CHECK_ARG(counters != nullptr, "scratch too small: need %zu bytes", nfb_traverse_scratch_bytes(rays->n_rays));
    RayArgs A = to_args(rays);
    FillPolicy<true> P{};
    P.near_arr = A.near_arr; P.near_s = A.near_s;
    P.sample_info = sample_info; P.interval_info = interval_info; P.term = termination_planes;
    P.proto.iv_vals = iv_vals; P.proto.iv_ray = iv_ray_indices; P.proto.iv_left = iv_is_left; P.proto.iv_right = iv_is_right;
    P.proto.sm_vals = sm_vals; P.proto.sm_ray = sm_ray_indices; P.proto.sm_valid = sm_is_valid;
    if (over_allocate) {
        NFB_CHECK_ARG(march->traverse_steps_limit > 0, "traverse_steps_limit must be > 0 when over_allocate is true");
        NFB_CHECK_ARG(sample_starts && interval_starts, "over_allocate needs the allocated starts");
        P.rays_mask = A.rays_mask;
        P.single_pass = true;
        P.sm_starts = sample_starts; P.sm_stride = 1;
        P.iv_starts = interval_starts; P.iv_stride = 1;
    } else {
        P.rays_mask = nullptr;
        P.single_pass = false;
        P.term = nullptr;  // already written by nfb_traverse_count for every ray
        P.sm_starts = sample_info; P.sm_stride = 2;
        P.iv_starts = interval_info; P.iv_stride = 2;
    }
    A.rays_mask = nullptr;
    rc = launch_march(A, to_grid(grid), to_opts(march), P, counters, s);
    if (rc != NFB_OK) return rc;
    if (over_allocate) {
        rc = scan_both(sample_info, interval_info, rays->n_rays, totals, scratch, s);
        if (rc == NFB_OK && totals != nullptr) cudaMemsetAsync(totals + 2, 0, sizeof(int64_t), s);
    }
    return rc;
}

/* ---- the chunked test-time loop (examples/utils.py:252-425), see the kernels above */
NFB_API size_t nfb_chunk_state_bytes(void) { return sizeof(nfb::ChunkState); }

NFB_API int nfb_chunk_init(void* state, int64_t n_rays, int32_t min_samples, int64_t* plan_host,
                           nfb_stream_t stream) {
    using namespace nfb;
    NFB_CHECK_ARG(state != nullptr && n_rays >= 0 && n_rays < (1ll << 31), "bad state / n_rays");
    cudaStream_t s = static_cast<cudaStream_t>(stream);
    // "next" = all rays; the plan kernel turns that into the first chunk
    chunk_seed_kernel<<<1, 1, 0, s>>>(static_cast<ChunkState*>(state), (int)n_rays);
    chunk_plan_kernel<<<1, 1, 0, s>>>(static_cast<ChunkState*>(state), n_rays, min_samples, plan_host);
    count_launch(2);
    NFB_CHECK_LAUNCH("chunk_init");
    return NFB_OK;
}

NFB_API int nfb_chunk_march(const nfb_rays_t* rays, const nfb_grid_t* grid, const nfb_march_t* march,
                            const int32_t* alive, int64_t n_alive_max, const void* state,
                            const int64_t* slot_of, int64_t* sample_info, float* slot_t_starts,
                            float* slot_t_ends, int64_t* ray_indices, float* t_starts, float* t_ends,
                            int64_t* totals, void* scratch, size_t scratch_bytes, nfb_stream_t stream) {
    using namespace nfb;
    int rc = check_common(rays, grid, march);
    if (rc != NFB_OK) return rc;
    cudaStream_t s = static_cast<cudaStream_t>(stream);
    if (rays->n_rays == 0 || n_alive_max <= 0) {
        if (totals) cudaMemsetAsync(totals, 0, 3 * sizeof(int64_t), s);
        return NFB_OK;
    }
    NFB_CHECK_ARG(alive && state && slot_of && sample_info && slot_t_starts && slot_t_ends && ray_indices && t_starts && t_ends,
                  "null pointer");
    NFB_CHECK_ARG(rays->near_planes != nullptr, "the chunked loop keeps per-ray near planes (termination planes) on the device");
    NFB_CHECK_ARG(n_alive_max <= rays->n_rays, "n_alive_max > n_rays");
    unsigned long long* counters = counters_of(scratch, scratch_bytes, rays->n_rays);
    NFB_CHECK_ARG(counters != nullptr, "scratch too small: need %zu bytes", nfb_traverse_scratch_bytes(rays->n_rays));
    const ChunkState* st = static_cast<const ChunkState*>(state);
    RayArgs A = to_args(rays);
    A.rays_mask = nullptr;
    A.ray_list = alive; A.n_work = n_alive_max; A.n_work_dev = &st->n_alive; A.limit_dev = &st->limit;
    FillPolicy<false> P{};
    P.rays_mask = nullptr; P.near_arr = A.near_arr; P.near_s = A.near_s;
    P.sm_starts = slot_of; P.sm_stride = 1; P.slot_mul_dev = &st->limit;
    P.sample_info = sample_info; P.interval_info = nullptr; P.single_pass = true;
    P.term = const_cast<float*>(A.near_arr);  // termination planes become the next chunk's near planes, in place
    P.proto.ray_indices = nullptr; P.proto.t_starts = slot_t_starts; P.proto.t_ends = slot_t_ends;
    nfb_march_t m2 = *march;
    m2.traverse_steps_limit = 64;  // (placeholder > 0: the kernel reads the chunk's k from the state)
    rc = launch_march(A, to_grid(grid), to_opts(&m2), P, counters, s);
    if (rc != NFB_OK) return rc;
    rc = launch_ray_scan(sample_info, rays->n_rays, reinterpret_cast<long long*>(totals), scratch, s);
    if (rc != NFB_OK) return rc;
    long long b = ceil_div<long long>(n_alive_max, 8);
    const long long cap = (long long)sm_count() * 16;
    if (b > cap) b = cap;
    chunk_compact_kernel<<<(unsigned)b, 256, 0, s>>>(alive, st, sample_info, slot_t_starts, slot_t_ends, ray_indices,
                                                    t_starts, t_ends);
    count_launch();
    NFB_CHECK_LAUNCH("chunk_march");
    return NFB_OK;
}

NFB_API int nfb_chunk_composite(const int32_t* alive, int32_t* alive_next, int64_t n_alive_max, void* state,
                                int64_t* sample_info, int64_t* slot_of, const float* t_starts,
                                const float* t_ends, const float* sigmas, const float* rgbs, int64_t n_rays,
                                float* rgb, float* opacity, float* depth, float alpha_thre,
                                float opacity_thre, int32_t min_samples, int64_t* plan_host,
                                nfb_stream_t stream) {
    using namespace nfb;
    NFB_CHECK_ARG(n_alive_max >= 0 && n_rays >= 0, "negative size");
    NFB_CHECK_ARG(state != nullptr, "null state");
    cudaStream_t s = static_cast<cudaStream_t>(stream);
    if (n_alive_max > 0) {
        NFB_CHECK_ARG(alive && alive_next && sample_info && slot_of && t_starts && t_ends && sigmas && rgbs && rgb &&
                      opacity && depth, "null pointer");
        ChunkCompositeArgs a{};
        a.alive = alive; a.alive_next = alive_next; a.st = static_cast<ChunkState*>(state);
        a.sample_info = sample_info; a.slot_of = slot_of; a.t_starts = t_starts; a.t_ends = t_ends;
        a.sigmas = sigmas; a.rgbs = rgbs; a.rgb = rgb; a.opacity = opacity; a.depth = depth;
        a.alpha_thre = alpha_thre; a.opc_thre = opacity_thre;
        long long b = ceil_div<long long>(n_alive_max, 8);
        const long long cap = (long long)sm_count() * 16;
        if (b > cap) b = cap;
        chunk_composite_kernel<<<(unsigned)b, 256, 0, s>>>(a);
        count_launch();
    }
    chunk_plan_kernel<<<1, 1, 0, s>>>(static_cast<ChunkState*>(state), n_rays, min_samples, plan_host);
    count_launch();
    NFB_CHECK_LAUNCH("chunk_composite");
    return NFB_OK;
}

}  // extern "C"
