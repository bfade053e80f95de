// Device code for the GPU accounting path (sm_100a).  HBM-bound integer work:
// no tensor cores; the design rules are coalesced 128-bit loads, shared-memory
// staging of prefix sums, warp shuffles, and few, spread-out atomics.
//
// Reference semantics (paths relative to /root/reference):
//   quantity_value   vendor/k8s.io/apimachinery/pkg/api/resource/amount.go:95-101,
//                    math.go:144-203 (and the dropped `ok`, quantity.go:713-714)
//   pods kernel      cmd/arena/commands/gpu.go:39-79, pod_helper.go:15-31,
//                    top_node.go:88-97, 232-242
//   (same kernel)    gpu.go:18-37, top_node.go:115-118 (node capacities, cluster total)
//   job kernels      job_info.go:51-70, trainer_mpi.go:241-282,
//                    trainer_standalone.go:198-221, trainer.go:25-37
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

namespace arena {

typedef unsigned long long u64;
typedef long long i64;

constexpr unsigned kNone = 0xFFFFFFFFu;
constexpr int kPhaseSucceeded = 3;
constexpr int kPhaseFailed = 4;

constexpr int kThreads = 256;              // 8 warps per CTA
constexpr int kWarps = kThreads / 32;
constexpr int kPodsPerLane = 4;            // one 128-bit load per u32 column
constexpr int kWarpPods = 32 * kPodsPerLane;   // 128 pods per warp tile
constexpr int kRowCtrs = 64;               // containers per row: 2 per lane, one 128-bit load
constexpr int kRows = 4;                   // rows per chunk held in registers
constexpr int kChunkCtrs = kRowCtrs * kRows;   // 256 containers -> 2 KB of prefix sums per warp
constexpr int kMaxPartials = 4096;         // upper bound on CTAs per launch

__device__ __constant__ i64 kPow10[19] = {
    1ll, 10ll, 100ll, 1000ll, 10000ll, 100000ll, 1000000ll, 10000000ll, 100000000ll,
    1000000000ll, 10000000000ll, 100000000000ll, 1000000000000ll, 10000000000000ll,
    100000000000000ll, 1000000000000000ll, 10000000000000000ll, 100000000000000000ll,
    1000000000000000000ll};

// Quantity.Value() of int64Amount{v, s}.
//   s < 0: divide by 10^-s, rounding AWAY from zero (negativeScaleInt64: the
//          digit-by-digit loop is equivalent to one truncating division plus a
//          +-1 when any remainder was dropped, including the early exit at 0).
//   s > 0: multiply by 10^s.  The caller drops `ok` and keeps the value, and the
//          value on overflow depends on the scale (math.go:144-168): scales
//          1, 2, 3, 6, 9 are ONE checked multiply that returns the WRAPPED
//          product (math.go:118-140); every other scale is a loop of x10 steps
//          that returns 0 once a step overflows.  INT64_MIN gives 0 for any s >= 1.
__device__ __noinline__ i64 quantity_value_slow(i64 v, int s)
{
    if (s < 0) {
        int k = -s;
        if (k > 18) return v > 0 ? 1 : (v < 0 ? -1 : 0);
        i64 d = kPow10[k];
        i64 q, r;
        if (v >= -2147483647ll && v <= 2147483647ll && k <= 9) {   // cheap 32-bit divide
            int vi = (int)v, di = (int)d;
            int qi = vi / di;
            q = qi;
            r = vi - qi * di;
        } else {
            q = v / d;
            r = v - q * d;
        }
        if (r != 0) q += v > 0 ? 1 : -1;
        return q;
    }
    if (s > 18) return 0;
    if (v == (i64)0x8000000000000000ull) return 0;
    i64 m = kPow10[s];
    i64 lo = (i64)((u64)v * (u64)m);
    if (s == 1 || s == 2 || s == 3 || s == 6 || s == 9) return lo;
    i64 hi = __mul64hi(v, m);
    return hi == (lo >> 63) ? lo : 0;
}

// floor(2^64 / 10^k) + 1: with a < 2^32, umul64hi(a, kRecip10[k]) == a / 10^k exactly
// (the error term a * (M*d - 2^64) / (d * 2^64) is below 2^-32 < 1/d).
__device__ __constant__ u64 kRecip10[10] = {
    0ull, 1844674407370955162ull, 184467440737095517ull, 18446744073709552ull, 1844674407370956ull,
    184467440737096ull, 18446744073710ull, 1844674407371ull, 184467440738ull, 18446744074ull};

// scale != 0.  The common shape ("500m": scale in [-9,-1], |value| < 2^32) is
// branch-light: one multiply-high by a reciprocal instead of a 64-bit divide.
__device__ __forceinline__ i64 quantity_value_nz(i64 v, int s)
{
    const u64 a = v < 0 ? 0ull - (u64)v : (u64)v;
    if (s < 0 && s >= -9 && (a >> 32) == 0) {
        const unsigned k = (unsigned)(-s);
        const unsigned d = (unsigned)kPow10[k];
        unsigned q = (unsigned)__umul64hi(a, kRecip10[k]);
        const unsigned r = (unsigned)a - q * d;
        q += (r != 0u);                       // round away from zero
        return v < 0 ? -(i64)q : (i64)q;
    }
    return quantity_value_slow(v, s);
}

__device__ __forceinline__ i64 quantity_value(i64 v, int s)
{
    if (s == 0) return v;
    return quantity_value_nz(v, s);
}

__device__ __forceinline__ bool phase_active(unsigned ph)
{
    return ph != kPhaseSucceeded && ph != kPhaseFailed;
}

__device__ __forceinline__ u64 shfl_u64(u64 v, int src)
{
    return __shfl_sync(0xffffffffu, v, src);
}

// Inclusive warp scans (wrapping adds).  shfl.up's predicate output says
// whether the source lane exists, so each step is SHFL + predicated ADD.
__device__ __forceinline__ unsigned warp_inclusive_scan_u32(unsigned x)
{
    // one asm statement for the five steps: the compiler then guards the whole
    // sequence with a single convergence check instead of one per shuffle
    asm volatile("{\n\t.reg .u32 y;\n\t.reg .pred p;\n\t"
                 "shfl.sync.up.b32 y|p, %0, 1, 0, 0xffffffff;\n\t@p add.u32 %0, %0, y;\n\t"
                 "shfl.sync.up.b32 y|p, %0, 2, 0, 0xffffffff;\n\t@p add.u32 %0, %0, y;\n\t"
                 "shfl.sync.up.b32 y|p, %0, 4, 0, 0xffffffff;\n\t@p add.u32 %0, %0, y;\n\t"
                 "shfl.sync.up.b32 y|p, %0, 8, 0, 0xffffffff;\n\t@p add.u32 %0, %0, y;\n\t"
                 "shfl.sync.up.b32 y|p, %0, 16, 0, 0xffffffff;\n\t@p add.u32 %0, %0, y;\n\t}"
                 : "+r"(x));
    return x;
}

__device__ __forceinline__ u64 warp_inclusive_scan(u64 x, int lane)
{
    unsigned lo = (unsigned)x, hi = (unsigned)(x >> 32);
#pragma unroll
    for (int d = 1; d < 32; d <<= 1)
        asm volatile("{\n\t.reg .u32 ylo, yhi;\n\t.reg .pred p;\n\t"
                     "shfl.sync.up.b32 ylo|p, %0, %2, 0, 0xffffffff;\n\t"
                     "shfl.sync.up.b32 yhi, %1, %2, 0, 0xffffffff;\n\t"
                     "@p add.cc.u32 %0, %0, ylo;\n\t"
                     "@p addc.u32 %1, %1, yhi;\n\t}"
                     : "+r"(lo), "+r"(hi) : "r"(d));
    (void)lane;
    return ((u64)hi << 32) | lo;
}

// Scan of values that are usually tiny: when every lane's value is below 2^26
// the 32-lane sum fits 32 bits and the cheap scan is exact; otherwise (huge or
// negative quantities) fall back to the 64-bit one.  Warp-uniform branch.
__device__ __forceinline__ u64 warp_inclusive_scan_auto(u64 x, int lane)
{
    if (!__any_sync(0xffffffffu, (x >> 26) != 0ull)) return (u64)warp_inclusive_scan_u32((unsigned)x);
    return warp_inclusive_scan(x, lane);
}

__device__ __forceinline__ u64 warp_reduce_sum(u64 x)
{
#pragma unroll
    for (int d = 16; d > 0; d >>= 1) x += __shfl_xor_sync(0xffffffffu, x, d);
    return x;
}

__device__ __forceinline__ void red_add(u64 *p, u64 v)
{
    atomicAdd(p, v);   // result unused -> RED.E.ADD.64
}

// Streaming loads: read-only path, do not pollute L1 with data touched once.
__device__ __forceinline__ uint4 ldg_stream(const uint4 *p)
{
    uint4 r;
    asm volatile("ld.global.nc.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];"
                 : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w) : "l"(p));
    return r;
}
__device__ __forceinline__ ulonglong2 ldg_stream(const ulonglong2 *p)
{
    ulonglong2 r;
    asm volatile("ld.global.nc.L1::no_allocate.v2.u64 {%0,%1}, [%2];"
                 : "=l"(r.x), "=l"(r.y) : "l"(p));
    return r;
}

// Scratch shared by the kernels of one device context.  Tickets are left at 0
// by the CTA that finishes last, so no launch needs a memset first.
// Each ticket id owns its own half of the partial arrays (kMaxPartials / 2 CTAs per launch), so a
// `top node` query (ticket 1) and a `top job` query (ticket 2) on the same handle do not share them.
struct Scratch {
    u64 partial_a[kMaxPartials];   // per-CTA partial of the first total
    u64 partial_b[kMaxPartials];   // per-CTA partial of the second total
    unsigned ticket[4];
    unsigned next_round[4];        // per ticket id: CTA-rounds handed out from the pool of the running launch
};

// ---- totals across GPUs without a collective launch -------------------------------------------
// Every rank owns a mailbox in its own HBM that all ranks (itself included) can write through
// NVLink peer mappings (CUDA IPC).  The last CTA of a rank's kernel stores the rank's two totals
// and then the step's epoch into slot [rank] of EVERY mailbox, waits until its own mailbox
// shows the epoch in all slots, and adds them up: the exchange rides in the tail of the kernel
// that produced the numbers.
//
// Two ways to collect.  Synchronous (lag 0): the launch that posted epoch e waits for every
// peer's epoch e, so each step ends at the pace of the slowest rank of THAT step.  Deferred
// (lag 1): the launch posts epoch e and collects epoch e-1, which every peer posted a whole
// step ago, so a rank that is late by less than a step stalls nobody; the totals of step e are
// delivered by step e+1's launch, or by peer_collect_kernel at the end of a sequence.
//
// Slots are a ring of kPeerRing epochs.  A rank cannot finish the tail of step e before every
// peer has posted e-1 (deferred) or e (synchronous), so a writer is never more than three epochs
// ahead of the oldest epoch a peer may still be reading (e+1 written while e-2 is read): four
// slots keep them apart.
constexpr int kMaxPeers = 16;
constexpr int kPeerRing = 4;
// A cell is {value, epoch} written with ONE aligned 16-byte store and polled with one 16-byte
// load: the value is valid as soon as the epoch beside it is the expected one, so no fence (a round
// trip over NVLink) has to separate data from flag.
struct __align__(16) PeerCell { unsigned long long value, epoch; };
struct PeerMailbox {
    PeerCell cell[kPeerRing][kMaxPeers][2];          // [slot][writer][first total, second total]
};

__device__ __forceinline__ void peer_cell_store(PeerCell *p, u64 value, u64 epoch)
{
    asm volatile("st.volatile.global.v2.u64 [%0], {%1, %2};" :: "l"(p), "l"(value), "l"(epoch) : "memory");
}
__device__ __forceinline__ PeerCell peer_cell_load(const PeerCell *p)
{
    PeerCell c;
    asm volatile("ld.volatile.global.v2.u64 {%0, %1}, [%2];" : "=l"(c.value), "=l"(c.epoch) : "l"(p) : "memory");
    return c;
}
struct PeerExchange {
    int rank, world;
    unsigned timed_out;                          // set when a peer did not show up within the spin budget
    unsigned long long epoch;                    // steps posted so far: kept on the DEVICE and advanced by the launch
                                                 // itself, so a captured launch (CUDA graph) can be replayed
    PeerMailbox *box[kMaxPeers];                 // rank i's mailbox as mapped into this process
};

// Last-CTA-done reduction of per-CTA partial sums into *out (and *out_b).
// Every CTA calls this once with its partial(s) in thread 0.
// Thread i < world: wait until rank i's totals of `epoch` have arrived in this rank's mailbox and
// return them (bounded spin; a rank that never shows up sets px->timed_out).
__device__ __forceinline__ void peer_collect(PeerExchange *px, u64 epoch, u64 &ga, u64 &gb)
{
    PeerMailbox *mine = px->box[px->rank];
    const int slot = (int)(epoch % (u64)kPeerRing);
    const long long t0 = clock64();
    for (;;) {
        const PeerCell a = peer_cell_load(&mine->cell[slot][threadIdx.x][0]);
        const PeerCell b = peer_cell_load(&mine->cell[slot][threadIdx.x][1]);
        if (a.epoch == epoch && b.epoch == epoch) { ga = a.value; gb = b.value; return; }
        __nanosleep(64);
        if (clock64() - t0 > 4000000000ll) { px->timed_out = 1u; return; }      // ~2 s: a peer is gone
    }
}

__device__ __forceinline__ void finish_totals(u64 block_a, u64 block_b, bool two,
                                              Scratch *sc, int ticket_id, i64 *out_a, i64 *out_b,
                                              PeerExchange *px = nullptr, unsigned lag = 0)
{
    __shared__ bool s_last;
    u64 *const part_a = sc->partial_a + (ticket_id == 1 ? 0 : kMaxPartials / 2);
    u64 *const part_b = sc->partial_b + (ticket_id == 1 ? 0 : kMaxPartials / 2);
    if (threadIdx.x == 0) {
        part_a[blockIdx.x] = block_a;
        if (two) part_b[blockIdx.x] = block_b;
        __threadfence();
        unsigned t = atomicAdd(&sc->ticket[ticket_id], 1u);
        s_last = (t == gridDim.x - 1);
    }
    __syncthreads();
    if (!s_last) return;
    __threadfence();
    u64 a = 0, b = 0;
    for (unsigned i = threadIdx.x; i < gridDim.x; i += blockDim.x) {
        a += *((volatile u64 *)&part_a[i]);
        if (two) b += *((volatile u64 *)&part_b[i]);
    }
    __shared__ u64 s_a[32], s_b[32];            // (any CTA size)
    a = warp_reduce_sum(a);
    b = warp_reduce_sum(b);
    if ((threadIdx.x & 31) == 0) { s_a[threadIdx.x >> 5] = a; s_b[threadIdx.x >> 5] = b; }
    __syncthreads();
    __shared__ u64 s_tot[2];
    if (threadIdx.x == 0) {
        u64 ta = 0, tb = 0;
        for (int w = 0; w < (int)(blockDim.x >> 5); w++) { ta += s_a[w]; tb += s_b[w]; }
        s_tot[0] = ta; s_tot[1] = tb;
        if (!px) {
            *out_a = (i64)ta;
            if (two) *out_b = (i64)tb;
        }
        sc->ticket[ticket_id] = 0;      // ready for the next launch
        sc->next_round[ticket_id] = 0;
        if (px) px->epoch += 1;         // this CTA is the only writer; the next launch is ordered behind it
    }
    if (!px) return;
    __syncthreads();
    const u64 epoch = *((volatile unsigned long long *)&px->epoch);
    // post to every rank, then collect from every rank: thread i talks to rank i
    const int slot = (int)(epoch % (u64)kPeerRing);
    u64 ga = 0, gb = 0;
    if ((int)threadIdx.x < px->world) {
        PeerMailbox *theirs = px->box[threadIdx.x];
        peer_cell_store(&theirs->cell[slot][px->rank][0], s_tot[0], epoch);
        peer_cell_store(&theirs->cell[slot][px->rank][1], s_tot[1], epoch);
        if (epoch > (u64)lag) peer_collect(px, epoch - lag, ga, gb);
    }
    if (epoch <= (u64)lag) return;     // deferred collection, first step: nothing older to deliver
    if (threadIdx.x < 32) {            // world <= 16: all in the first warp
        ga = warp_reduce_sum(ga);
        gb = warp_reduce_sum(gb);
        const bool timed_out = __any_sync(0xffffffffu, *((volatile unsigned *)&px->timed_out) != 0u);
        if (threadIdx.x == 0) {
            // a peer that never posted: leave no half-sum behind that could be taken for a total
            *out_a = timed_out ? (i64)0x8000000000000000ull : (i64)ga;
            if (two) *out_b = timed_out ? (i64)0x8000000000000000ull : (i64)gb;
        }
    }
}

// ---------------------------------------------------------------------------
// Pods kernel.  One warp owns a tile of 128 consecutive pods:
//
//  1. 128-bit loads of the pod columns (4 pods per lane).
//  2. The tile's containers are a contiguous range of the container columns
//     (CSR).  They are read 64 per row -- one 128-bit load of two int64 values
//     and one 16-bit load of two scales per lane, fully coalesced -- decoded
//     with quantity_value, prefix-summed across the warp with shuffles, and the
//     running EXCLUSIVE prefix sums are parked in the warp's 2 KB of shared
//     memory.  gpuInPod(p) is then prefix[end(p)] - prefix[begin(p)]: two
//     shared-memory reads per pod, no per-pod loop, any containers-per-pod.
//  3. Per-key (node or job) sums use the same telescoping trick one level up:
//     with S the inclusive prefix sum of the masked pod values over the tile,
//     every position where the key changes adds +S to the key on the left and
//     -S to the key on the right (and the tile end adds +S to the last key).
//     Unsigned arithmetic wraps exactly like Go's int64 `+=`, so the result is
//     bit-exact whatever the order of the atomics, and it is CORRECT for any
//     pod order -- grouped input merely makes key changes (hence atomics) rare.
//
// kJob == false: key = pod_node, one sum (allocated GPUs), phase-masked.
// kJob == true : key = pod_job, two sums (requested: unmasked; allocated:
//                phase-masked); ARENA_ROLE_LATEST pods are left out of the sums
//                and instead bid for "newest pod of the job" with atomicMax.
// ---------------------------------------------------------------------------
struct PodsArgs {
    u64 P, C, K;                    // pods, containers, number of keys (N or J)
    const unsigned *pod_key;        // pod_node or pod_job
    const unsigned char *pod_phase;
    const unsigned *pod_ctr_begin;
    const i64 *ctr_val;
    const signed char *ctr_scale;
    const unsigned char *pod_role;  // kJob only
    const i64 *pod_created;         // kJob only
    u64 *key_alloc;                 // [K] pre-zeroed
    u64 *key_req;                   // [K] pre-zeroed (kJob only)
    i64 *key_maxcreated;            // [K] pre-set to INT64_MIN (kJob only)
    i64 *pod_gpu;                   // [P] or nullptr
    i64 *total_alloc;               // sum over keys of key_alloc
    i64 *total_req;                 // kJob only
    Scratch *scratch;
    // top node only: the per-node capacities, decoded by the same launch
    u64 n_nodes;
    const i64 *cap_val;
    const signed char *cap_scale;
    i64 *node_total;                // [N] Value(capacity)
    i64 *total_capacity;            // sum of node_total
    // multi-GPU: exchange the two totals with the peer GPUs inside this launch (or nullptr)
    PeerExchange *px;
    unsigned px_lag;                // 0: collect this step's totals; 1: the previous step's (deferred)
    // Tiles below n_static (a multiple of the grid's warp count) are taken grid-stride; the rest form a pool
    // that CTAs draw from, one run of kPodWarps consecutive tiles at a time (top_node_pods_kernel).
    unsigned n_static;
};

// Rare path (a row that runs past the end of the container columns) kept out
// of line so the hot loop stays small.  Everything by value: no stack traffic.
struct RowTail { u64 x, y; unsigned sc; };
__device__ __noinline__ RowTail load_row_tail(const i64 *ctr_val, const signed char *ctr_scale, u64 C, u64 idx)
{
    RowTail t;
    t.x = 0; t.y = 0; t.sc = 0;
    if (idx < C) {
        t.x = (u64)__ldg(ctr_val + idx);
        t.sc = (unsigned char)__ldg(ctr_scale + idx);
    }
    if (idx + 1 < C) {
        t.y = (u64)__ldg(ctr_val + idx + 1);
        t.sc |= (unsigned)(unsigned char)__ldg(ctr_scale + idx + 1) << 8;
    }
    return t;
}

struct TileSums { u64 alloc, req; };   // tile totals, valid in lane 31

// Generic tile: any containers-per-pod (multi-chunk), any int64 value, any
// scale, ragged last tile.  64-bit arithmetic throughout.  This is the whole
// job kernel; the top-node kernel only falls back to it for tiles its 32-bit
// fast path cannot take.
template <bool kJob>
__device__ __forceinline__ TileSums process_tile_generic(const PodsArgs &a, const unsigned tile, const bool full_tile,
                                                         const int lane, u64 *prefix)
{
    const unsigned FULL = 0xffffffffu;
    const u64 q = (u64)tile * kWarpPods + (u64)lane * kPodsPerLane;

    // ---- 1. pod columns ------------------------------------------------
    unsigned cb[kPodsPerLane + 1], key[kPodsPerLane];
    unsigned ph4, r4 = 0;
    if (full_tile) {                                     // vector loads
        uint4 c4 = ldg_stream(reinterpret_cast<const uint4 *>(a.pod_ctr_begin + q));
        uint4 k4 = ldg_stream(reinterpret_cast<const uint4 *>(a.pod_key + q));
        ph4 = __ldg(reinterpret_cast<const unsigned *>(a.pod_phase + q));
        unsigned last = 0;
        if (lane == 31) last = __ldg(a.pod_ctr_begin + q + kPodsPerLane);
        if (kJob) r4 = __ldg(reinterpret_cast<const unsigned *>(a.pod_role + q));
        cb[0] = c4.x; cb[1] = c4.y; cb[2] = c4.z; cb[3] = c4.w;
        key[0] = k4.x; key[1] = k4.y; key[2] = k4.z; key[3] = k4.w;
        unsigned nxt = __shfl_down_sync(FULL, c4.x, 1);
        cb[4] = lane == 31 ? last : nxt;
    } else {                                             // ragged last tile
        ph4 = 0;
#pragma unroll
        for (int j = 0; j <= kPodsPerLane; j++) {
            const u64 i = q + j;
            cb[j] = __ldg(a.pod_ctr_begin + (i < a.P ? i : a.P));
        }
#pragma unroll
        for (int j = 0; j < kPodsPerLane; j++) {
            const u64 i = q + j;
            const bool in = i < a.P;
            key[j] = in ? __ldg(a.pod_key + i) : kNone;
            ph4 |= (in ? (unsigned)__ldg(a.pod_phase + i) : 0u) << (8 * j);
            if (kJob) r4 |= (in ? (unsigned)__ldg(a.pod_role + i) : 0u) << (8 * j);
        }
    }

    // ---- 2. container prefix sums --------------------------------------
    const unsigned wc0 = __shfl_sync(FULL, cb[0], 0);
    const unsigned wc1 = __shfl_sync(FULL, cb[kPodsPerLane], 31);
    u64 B[kPodsPerLane + 1];
#pragma unroll
    for (int j = 0; j <= kPodsPerLane; j++) B[j] = 0;

    u64 carry = 0;
    // (64-bit loop variable: with C close to 2^32 a 32-bit one would wrap past wc1 and never leave)
    for (u64 chunk_lo = wc0 & ~(unsigned)(kRowCtrs - 1); chunk_lo < wc1; chunk_lo += kChunkCtrs) {
        ulonglong2 v[kRows];
        unsigned sc[kRows];
        // issue every load of the chunk before using any of them
#pragma unroll
        for (int r = 0; r < kRows; r++) {
            const u64 rs = chunk_lo + r * kRowCtrs;
            const u64 idx = rs + 2u * lane;
            v[r].x = 0; v[r].y = 0; sc[r] = 0;
            if (rs < wc1) {
                if (rs + kRowCtrs <= a.C) {
                    v[r] = ldg_stream(reinterpret_cast<const ulonglong2 *>(a.ctr_val + idx));
                    sc[r] = __ldg(reinterpret_cast<const unsigned short *>(a.ctr_scale + idx));
                } else {
                    const RowTail t = load_row_tail(a.ctr_val, a.ctr_scale, a.C, idx);
                    v[r].x = t.x; v[r].y = t.y; sc[r] = t.sc;
                }
            }
        }
        unsigned n_valid = 0;
#pragma unroll
        for (int r = 0; r < kRows; r++) {
            const u64 rs = chunk_lo + r * kRowCtrs;
            if (rs < wc1) {                                // warp-uniform
                u64 x0 = v[r].x, x1 = v[r].y;
                if (__any_sync(FULL, sc[r] != 0u)) {       // some quantity in the row has a scale
                    const int s0 = (int)(signed char)(sc[r] & 0xffu);
                    const int s1 = (int)(signed char)(sc[r] >> 8);
                    if (s0 != 0) x0 = (u64)quantity_value_nz((i64)x0, s0);
                    if (s1 != 0) x1 = (u64)quantity_value_nz((i64)x1, s1);
                }
                const u64 s = x0 + x1;
                const u64 inc = warp_inclusive_scan_auto(s, lane);
                const u64 e0 = carry + inc - s;
                ulonglong2 e;
                e.x = e0;
                e.y = e0 + x0;
                *reinterpret_cast<ulonglong2 *>(prefix + r * kRowCtrs + 2 * lane) = e;
                carry += shfl_u64(inc, 31);
                if (lane == 31) prefix[(r + 1) * kRowCtrs] = carry;   // inclusive end of the row
                n_valid = (r + 1) * kRowCtrs;
            }
        }
        __syncwarp();
#pragma unroll
        for (int j = 0; j <= kPodsPerLane; j++) {
            const unsigned off = cb[j] - (unsigned)chunk_lo;   // wraps when cb[j] < chunk_lo
            if (off <= n_valid) B[j] = prefix[off];
        }
        __syncwarp();
    }

    // ---- 3. pod values, masks, telescoping sums per key -----------------
    u64 g[kPodsPerLane];
#pragma unroll
    for (int j = 0; j < kPodsPerLane; j++) g[j] = B[j + 1] - B[j];

    if (a.pod_gpu != nullptr) {
        if (full_tile) {
            ulonglong2 o0, o1;
            o0.x = g[0]; o0.y = g[1]; o1.x = g[2]; o1.y = g[3];
            ulonglong2 *dst = reinterpret_cast<ulonglong2 *>(a.pod_gpu + q);
            dst[0] = o0;
            dst[1] = o1;
        } else {
#pragma unroll
            for (int j = 0; j < kPodsPerLane; j++)
                if (q + j < a.P) a.pod_gpu[q + j] = (i64)g[j];
        }
    }

    u64 S[kPodsPerLane], R[kPodsPerLane];
    u64 run_a = 0, run_r = 0;
#pragma unroll
    for (int j = 0; j < kPodsPerLane; j++) {
        const unsigned ph = (ph4 >> (8 * j)) & 0xffu;
        bool counts = key[j] < a.K;
        if (kJob) {
            const unsigned role = (r4 >> (8 * j)) & 0xffu;
            if (counts && role == 1) {                    // bids for "latest pod of the job"
                counts = false;
                atomicMax(a.key_maxcreated + key[j], __ldg(a.pod_created + q + j));
            }
        }
        const u64 req = counts ? g[j] : 0;
        const u64 alc = phase_active(ph) ? req : 0;
        run_a += alc;
        S[j] = run_a;
        if (kJob) { run_r += req; R[j] = run_r; }
    }
    TileSums out;
    out.alloc = 0; out.req = 0;
    {
        const u64 inc = warp_inclusive_scan_auto(run_a, lane);
        const u64 off = inc - run_a;
#pragma unroll
        for (int j = 0; j < kPodsPerLane; j++) S[j] += off;
        out.alloc = inc;
    }
    if (kJob) {
        const u64 inc = warp_inclusive_scan_auto(run_r, lane);
        const u64 off = inc - run_r;
#pragma unroll
        for (int j = 0; j < kPodsPerLane; j++) R[j] += off;
        out.req = inc;
    }
    unsigned key_next = __shfl_down_sync(FULL, key[0], 1);
    if (lane == 31) key_next = kNone;
    // a lane has work here only where the key changes (rare on grouped input)
    const bool any_edge = (key[0] != key[1]) | (key[1] != key[2]) | (key[2] != key[3]) |
                          (key[3] != key_next) | (lane == 31);
    if (any_edge) {
#pragma unroll
        for (int j = 0; j < kPodsPerLane; j++) {
            const unsigned kl = key[j];
            const unsigned kr = j + 1 < kPodsPerLane ? key[j + 1] : key_next;
            const bool tile_end = (lane == 31 && j == kPodsPerLane - 1);
            if (kl != kr || tile_end) {
                if (kl < a.K) {
                    red_add(a.key_alloc + kl, S[j]);
                    if (kJob) red_add(a.key_req + kl, R[j]);
                }
                if (!tile_end && kr < a.K) {
                    red_add(a.key_alloc + kr, 0ull - S[j]);
                    if (kJob) red_add(a.key_req + kr, 0ull - R[j]);
                }
            }
        }
    }
    return out;
}

// Out-of-line copy for the top-node kernel's fallback.  The kernel parameter is
// __grid_constant__, so its address can be passed without a local copy.
__device__ __noinline__ TileSums process_tile_fallback(const PodsArgs *a, unsigned tile, bool full_tile, int lane, u64 *prefix)
{
    return process_tile_generic<false>(*a, tile, full_tile, lane, prefix);
}

__device__ __noinline__ TileSums process_tile_fallback_job(const PodsArgs *a, unsigned tile, bool full_tile, int lane, u64 *prefix)
{
    return process_tile_generic<true>(*a, tile, full_tile, lane, prefix);
}

// Job kernel: generic tiles only.
__global__ void __launch_bounds__(kThreads) job_pods_kernel(PodsArgs a)
{
    __shared__ __align__(16) u64 s_prefix[kWarps][kChunkCtrs + 2];
    __shared__ u64 s_wa[kWarps], s_wb[kWarps];
    const int lane = threadIdx.x & 31;
    const int warp = threadIdx.x >> 5;
    u64 acc_alloc = 0, acc_req = 0;
    const unsigned n_tiles = (unsigned)((a.P + kWarpPods - 1) / kWarpPods);
    const unsigned n_full = (unsigned)(a.P / kWarpPods);
    for (unsigned tile = blockIdx.x * kWarps + warp; tile < n_tiles; tile += gridDim.x * kWarps) {
        const TileSums t = process_tile_generic<true>(a, tile, tile < n_full, lane, s_prefix[warp]);
        acc_alloc += t.alloc;
        acc_req += t.req;
    }
    if (lane == 31) { s_wa[warp] = acc_alloc; s_wb[warp] = acc_req; }
    __syncthreads();
    u64 block_a = 0, block_b = 0;
    if (threadIdx.x == 0)
        for (int w = 0; w < kWarps; w++) { block_a += s_wa[w]; block_b += s_wb[w]; }
    finish_totals(block_a, block_b, true, a.scratch, 2, a.total_alloc, a.total_req);
}

// ---------------------------------------------------------------------------
// Top-node pods kernel: 32-bit fast path.
//
// A tile takes the fast path when (warp-uniform tests)
//   * it is a full tile of 128 pods,
//   * its containers fit one chunk: at most 256 container slots counted from
//     the multiple of 16 at or below its first container (128-byte aligned
//     value rows), all inside the columns,
//   * every quantity it touches is small and plain: 0 <= value < 2^23 and
//     scale in [-9, 0].
// Then every prefix sum fits 32 bits (256 * 2^23 = 2^31), so the whole tile is
// 32-bit adds, 32-bit shuffles and 32-bit shared-memory words.
//
// Row geometry: a row is 128 containers = two halves of 64.  Lane l owns the
// pair {2l, 2l+1} of each half, so every value load is one fully coalesced
// 128-bit load per lane (512 contiguous bytes per warp) and every scale load a
// coalesced 16-bit load.  The two halves need two prefix scans; when the lane
// sums are below 2^11 (any realistic GPU count) they are packed into the two
// 16-bit fields of ONE 32-bit scan.  A typical tile is 2 rows = 2 scans.
// Quantities with a negative scale ("500m") are rounded up with a
// multiply-by-reciprocal table (kCeilDiv10, exact on this range) in a loop that
// only runs for the lanes that hold one.  Anything else -- negative or huge
// values, positive scales, pods with hundreds of containers, the ragged last
// tile -- is handed to the 64-bit generic tile, so the result is always exact.
// ---------------------------------------------------------------------------
constexpr int kHalf = 64;                  // containers per half row
constexpr int kFastRow = 2 * kHalf;        // containers per fast-path row
constexpr unsigned kSmallBits = 23;

// k = -scale in [0, 9]:  ceil(x / 10^k) == ((x + 10^k - 1) * m) >> sh exactly for
// x < 2^23 (sh = 31 + ceil(log2 10^k), m = floor(2^sh / 10^k) + 1 < 2^32;
// t = x + 10^k - 1 < 2^31 so t * 10^k < 2^sh, which bounds the error term below
// 1/10^k).  Checked exhaustively by tests/test_tables.py.
struct CeilDiv10 { unsigned m, sh, dm1, pad; };
__device__ __constant__ CeilDiv10 kCeilDiv10[10] = {
    {1u, 0u, 0u, 0u},
    {3435973837u, 35u, 9u, 0u},
    {2748779070u, 38u, 99u, 0u},
    {2199023256u, 41u, 999u, 0u},
    {3518437209u, 45u, 9999u, 0u},
    {2814749768u, 48u, 99999u, 0u},
    {2251799814u, 51u, 999999u, 0u},
    {3602879702u, 55u, 9999999u, 0u},
    {2882303762u, 58u, 99999999u, 0u},
    {2305843010u, 61u, 999999999u, 0u},
};

// {m, 10^k - 1} of kCeilDiv10, padded to 16 entries so that any nibble is a
// valid index (entries 0 and 10..15 are never used for valid scales).
__device__ __constant__ uint2 kCeilDiv10Packed[16] = {
    {0u, 0u}, {3435973837u, 9u}, {2748779070u, 99u}, {2199023256u, 999u}, {3518437209u, 9999u},
    {2814749768u, 99999u}, {2251799814u, 999999u}, {3602879702u, 9999999u}, {2882303762u, 99999999u},
    {2305843010u, 999999999u}, {0u, 0u}, {0u, 0u}, {0u, 0u}, {0u, 0u}, {0u, 0u}, {0u, 0u}};

__device__ __forceinline__ void red_add_u64(u64 *p, u64 v)
{
    asm volatile("red.global.add.u64 [%0], %1;" :: "l"(p), "l"(v) : "memory");
}

// Decode one scaled quantity: sb = scale byte, already validated to be 0xF7..0xFF
// (-9..-1), so k = -scale is the low nibble of -sb.  ceil(x / 10^k) is the HIGH
// word of (x + 10^k - 1) * m shifted by sh - 32 (every table shift is >= 35, and
// sh - 32 == (k * 3402) >> 10 for k = 1..9); m and 10^k - 1 come from constant
// memory with one 64-bit load (kCeilDiv10Packed, indexable by any nibble).
__device__ __forceinline__ unsigned decode_one(unsigned x, unsigned sb)
{
    const unsigned k = (0u - sb) & 0xfu;
    const uint2 t = kCeilDiv10Packed[k];                      // {m, 10^k - 1}
    return __umulhi(x + t.y, t.x) >> ((k * 3402u) >> 10);
}

// One row: decode, prefix-sum, park exclusive prefixes in pre32[0..128], return
// the running total after the row.  (xa0,xa1) = lane's pair in the first half,
// (xb0,xb1) = its pair in the second half.
__device__ __forceinline__ unsigned fast_row(unsigned xa0, unsigned xa1, unsigned xb0, unsigned xb1, unsigned sbytes,
                                             unsigned carry, unsigned *pre32, int lane, bool &bad)
{
    const unsigned FULL = 0xffffffffu;
    // scaled quantities are rare: each of the four positions is decoded only by
    // the lanes that hold one there (divergent, skipped when no lane does)
    if (sbytes != 0u) {
        // every byte must be 0 or in [0xF7, 0xFF] (scale 0 or -9..-1); anything else sends the
        // tile to the generic path (the values decoded below are then thrown away)
        bad |= (__vcmpgeu4(sbytes, 0xF7F7F7F7u) | __vcmpeq4(sbytes, 0u)) != 0xFFFFFFFFu;
        if (sbytes & 0x000000ffu) xa0 = decode_one(xa0, sbytes & 0xffu);
        if (sbytes & 0x0000ff00u) xa1 = decode_one(xa1, (sbytes >> 8) & 0xffu);
        if (sbytes & 0x00ff0000u) xb0 = decode_one(xb0, (sbytes >> 16) & 0xffu);
        if (sbytes & 0xff000000u) xb1 = decode_one(xb1, sbytes >> 24);
    }
    const unsigned sA = xa0 + xa1, sB = xb0 + xb1;
    unsigned exA, exB, tA, tB;
    if (!__any_sync(FULL, (sA | sB) >= 2048u)) {             // both scans in one: 16-bit fields
        const unsigned w = sA | (sB << 16);
        const unsigned inc = warp_inclusive_scan_u32(w);
        const unsigned tw = __shfl_sync(FULL, inc, 31);
        const unsigned ex = inc - w;
        exA = ex & 0xffffu; exB = ex >> 16;
        tA = tw & 0xffffu; tB = tw >> 16;
    } else {
        const unsigned incA = warp_inclusive_scan_u32(sA);
        const unsigned incB = warp_inclusive_scan_u32(sB);
        exA = incA - sA; exB = incB - sB;
        tA = __shfl_sync(FULL, incA, 31);
        tB = __shfl_sync(FULL, incB, 31);
    }
    exA += carry;
    exB += carry + tA;
    *reinterpret_cast<uint2 *>(pre32 + 2 * lane) = make_uint2(exA, exA + xa0);
    *reinterpret_cast<uint2 *>(pre32 + kHalf + 2 * lane) = make_uint2(exB, exB + xb0);
    const unsigned total = carry + tA + tB;
    if (lane == 31) pre32[kFastRow] = total;                  // inclusive end of the row
    return total;
}

// top_node_pods_kernel comes in two CTA shapes with the same 48 warps per SM: 8 warps x 6 CTAs, and 16 x 3 for
// the summary kernel on long launches -- fewer CTAs for the SM to play favourites with (1e8 pods: 437 -> 432 us;
// at 1.25e7 pods per GPU with the peer exchange the small shape is the quicker one by 1.5 us: tools/ab_shard2.sh).
constexpr int kPodWarps = 8, kPodThreads = kPodWarps * 32;
constexpr int kPodWarpsLong = 16, kPodThreadsLong = kPodWarpsLong * 32;

// kDirect: pods are NOT grouped by node (almost every position would be an edge and telescoping
// would cost two atomics per pod): each pod's own value is added to its node instead, one atomic
// per pod that contributes anything.  Both instantiations are exact for any order; the host picks
// one from a sample of the pod_node column (probe_grouping / arena_snapshot_upload).
// kDetails: also write gpuInPod of every pod (the details view); kept out of the summary kernel.
// kJob: the same tiles keyed by pod_job for `top job` (job_info.go:51-70): requested = every pod,
// allocated = active pods, ARENA_ROLE_LATEST pods only bid with their creation time; per-pod adds
// (pods of one job are scattered over the node-grouped rows), two totals.
#ifdef ARENA_TIMELINE
// Debug builds only (tools/variants.sh timeline "-DARENA_TIMELINE"): thread 0 of every CTA notes the global
// timer at six points of a launch; arena_debug_timeline() copies the table out (tools/timeline.py).
__device__ unsigned long long g_timeline[2048 * 6];
__device__ __forceinline__ void timeline_mark(int k)
{
    if (threadIdx.x == 0 && blockIdx.x < 2048) {
        unsigned long long t;
        asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
        g_timeline[blockIdx.x * 6 + k] = t;
        if (k == 0) {                                   // where the CTA runs: the SM id rides in the top bits of mark 0
            unsigned smid;
            asm volatile("mov.u32 %0, %%smid;" : "=r"(smid));
            g_timeline[blockIdx.x * 6] = (t & 0x0000FFFFFFFFFFFFull) | ((unsigned long long)smid << 48);
        }
    }
}
#define ARENA_MARK(k) timeline_mark(k)
#else
#define ARENA_MARK(k)
#endif

template <bool kDirect, bool kDetails, bool kJob = false, int kPW = kPodWarps>
__global__ void __launch_bounds__(kPW * 32, 48 / kPW) top_node_pods_kernel(const __grid_constant__ PodsArgs a)
{
    ARENA_MARK(0);
    __shared__ __align__(16) u64 s_prefix[kPW][kChunkCtrs + 2];
    __shared__ u64 s_wa[kPW];

    const int lane = threadIdx.x & 31;
    const int warp = threadIdx.x >> 5;
    const unsigned FULL = 0xffffffffu;
    unsigned *pre32 = reinterpret_cast<unsigned *>(s_prefix[warp]);

    u64 acc_alloc = 0;
    u64 job_req = 0, job_alloc = 0;             // kJob: every lane's own partial totals
    const unsigned n_tiles = (unsigned)((a.P + kWarpPods - 1) / kWarpPods);
    const unsigned n_full = (unsigned)(a.P / kWarpPods);
    const unsigned K32 = (unsigned)(a.K < 0xFFFFFFFFull ? a.K : 0xFFFFFFFFull);
    const unsigned tile_step = gridDim.x * kPW;

    // Tiles the 32-bit fast path cannot take (huge or negative values, positive scales, pods with hundreds of
    // containers, the ragged last tile) are not summed where they are met: the call into the generic 64-bit
    // tile would sit inside the hot loop and its live registers would be spilled around it on every trip.
    // A warp notes up to kDeferred such tiles and sums them after its loop; a warp that meets more leaves the
    // hot loop for good and takes ALL its remaining tiles generically (an input like that has no fast tiles).
    //
    // Who takes which tile.  The warps of an SM do not advance at one pace: its CTAs finish equal shares up to
    // 10 us apart (the scheduler favours some of them for the whole launch: profiles/r02_iterations.md), and a
    // launch is over when the slowest is.  So only the first n_static tiles are shared out in advance, grid-stride;
    // the last few rounds are a pool from which a CTA draws one run of kPW consecutive tiles at a time (one
    // atomic per CTA and run, asked for a run ahead; one barrier per run).  CTAs that are ahead take more of them.
    constexpr int kDeferred = 8;
    __shared__ unsigned s_deferred[kPW][kDeferred];
    __shared__ unsigned s_claim[2];
    unsigned n_deferred = 0;
    bool generic = false;                                  // this warp has left the fast code for good
    unsigned *const pool = &a.scratch->next_round[kJob ? 2 : 1];
    const unsigned n_static = a.n_static < n_tiles ? a.n_static : n_tiles;
    unsigned pending = 0, parity = 0;
    if (threadIdx.x == 0) pending = atomicAdd(pool, 1u);   // this CTA's first run from the pool: asked for now, read after its share
    unsigned tile = blockIdx.x * kPW + warp, limit = n_static;
    for (;;) {
    if (!generic)
    for (; tile < limit; tile += tile_step) {
        bool fast = tile < n_full;
        u64 tile_alloc = 0;
        if (fast) {
            const u64 q = (u64)tile * kWarpPods + (u64)lane * kPodsPerLane;
            // ---- pod columns -------------------------------------------------
            const uint4 c4 = ldg_stream(reinterpret_cast<const uint4 *>(a.pod_ctr_begin + q));
            const uint4 k4 = ldg_stream(reinterpret_cast<const uint4 *>(a.pod_key + q));
            const unsigned ph4 = __ldg(reinterpret_cast<const unsigned *>(a.pod_phase + q));
            unsigned cb4 = 0;
            if (lane == 31) cb4 = __ldg(a.pod_ctr_begin + q + kPodsPerLane);
            {
                const unsigned nxt = __shfl_down_sync(FULL, c4.x, 1);
                if (lane != 31) cb4 = nxt;
            }
            const unsigned wc0 = __shfl_sync(FULL, c4.x, 0);
            const unsigned wc1 = __shfl_sync(FULL, cb4, 31);
            const unsigned base = wc0 & ~15u;                 // 16 values = 128 bytes
            const unsigned span = wc1 - base;
            fast = span <= (unsigned)kChunkCtrs && (u64)base + kChunkCtrs <= a.C;
            if (fast) {
                // ---- container rows: every load first -------------------------------
                const unsigned i0 = base + 2u * lane;
                const bool two = span > (unsigned)kFastRow;       // warp-uniform
                const ulonglong2 vA0 = ldg_stream(reinterpret_cast<const ulonglong2 *>(a.ctr_val + i0));
                const ulonglong2 vB0 = ldg_stream(reinterpret_cast<const ulonglong2 *>(a.ctr_val + i0 + kHalf));
                unsigned s0 = __ldg(reinterpret_cast<const unsigned short *>(a.ctr_scale + i0));
                s0 |= (unsigned)__ldg(reinterpret_cast<const unsigned short *>(a.ctr_scale + i0 + kHalf)) << 16;
                ulonglong2 vA1, vB1;
                unsigned s1 = 0;
                vA1.x = vA1.y = vB1.x = vB1.y = 0;
                if (two) {
                    vA1 = ldg_stream(reinterpret_cast<const ulonglong2 *>(a.ctr_val + i0 + kFastRow));
                    vB1 = ldg_stream(reinterpret_cast<const ulonglong2 *>(a.ctr_val + i0 + kFastRow + kHalf));
                    s1 = __ldg(reinterpret_cast<const unsigned short *>(a.ctr_scale + i0 + kFastRow));
                    s1 |= (unsigned)__ldg(reinterpret_cast<const unsigned short *>(a.ctr_scale + i0 + kFastRow + kHalf)) << 16;
                }
                // ---- eligibility: small non-negative values, scales in [-9, 0] ----
                bool bad = ((vA0.x | vA0.y | vB0.x | vB0.y | vA1.x | vA1.y | vB1.x | vB1.y) >> kSmallBits) != 0ull;
                // offsets must be non-decreasing through the tile (lane chains to lane via cb4), which
                // keeps every shared-memory index below inside [0, span]: a corrupt CSR column is
                // summed by the generic path, whose lookups are range-checked, instead of faulting
                bad |= (c4.x > c4.y) | (c4.y > c4.z) | (c4.z > c4.w) | (c4.w > cb4);
                fast = !__any_sync(FULL, bad);
                if (fast) {
                    unsigned total = fast_row((unsigned)vA0.x, (unsigned)vA0.y, (unsigned)vB0.x, (unsigned)vB0.y, s0,
                                              0u, pre32, lane, bad);
                    if (two)
                        total = fast_row((unsigned)vA1.x, (unsigned)vA1.y, (unsigned)vB1.x, (unsigned)vB1.y, s1,
                                         total, pre32 + kFastRow, lane, bad);
                    // a scale outside [-9, 0] turned up while decoding: redo the tile generically
                    // (nothing has been published yet: shared memory is scratch, atomics come later)
                    fast = !__any_sync(FULL, bad);
                }
                if (fast) {
                    __syncwarp();
                    // ---- gpuInPod = prefix[end] - prefix[begin] -------------------------
                    const unsigned b0 = pre32[c4.x - base], b1 = pre32[c4.y - base], b2 = pre32[c4.z - base],
                                   b3 = pre32[c4.w - base], b4 = pre32[cb4 - base];
                    __syncwarp();
                    const unsigned g0 = b1 - b0, g1 = b2 - b1, g2 = b3 - b2, g3 = b4 - b3;
                    if (kDetails) {
                        ulonglong2 *dst = reinterpret_cast<ulonglong2 *>(a.pod_gpu + q);
                        ulonglong2 o0, o1;
                        o0.x = g0; o0.y = g1; o1.x = g2; o1.y = g3;
                        dst[0] = o0;
                        dst[1] = o1;
                    }
                    if (kJob) {
                        const unsigned r4 = __ldg(reinterpret_cast<const unsigned *>(a.pod_role + q));
                        const unsigned ks[4] = {k4.x, k4.y, k4.z, k4.w}, gs[4] = {g0, g1, g2, g3};
#pragma unroll
                        for (int j = 0; j < 4; j++) {
                            if (ks[j] >= K32) continue;
                            if (((r4 >> (8 * j)) & 0xffu) == 1u) {            // bids for "latest pod of the job"
                                atomicMax(a.key_maxcreated + ks[j], __ldg(a.pod_created + q + j));
                                continue;
                            }
                            const u64 g = gs[j];
                            if (g) red_add_u64(a.key_req + ks[j], g);
                            job_req += g;
                            if (phase_active((ph4 >> (8 * j)) & 0xffu)) {
                                if (g) red_add_u64(a.key_alloc + ks[j], g);
                                job_alloc += g;
                            }
                        }
                    } else {
                    // ---- masks: node exists, phase active ------------------------------
                    const unsigned a0 = (k4.x < K32 && phase_active(ph4 & 0xffu)) ? g0 : 0u;
                    const unsigned a1 = (k4.y < K32 && phase_active((ph4 >> 8) & 0xffu)) ? g1 : 0u;
                    const unsigned a2 = (k4.z < K32 && phase_active((ph4 >> 16) & 0xffu)) ? g2 : 0u;
                    const unsigned a3 = (k4.w < K32 && phase_active(ph4 >> 24)) ? g3 : 0u;
                    const unsigned S0 = a0, S1 = S0 + a1, S2 = S1 + a2, S3 = S2 + a3;
                    const unsigned inc = warp_inclusive_scan_u32(S3);
                    const unsigned off = inc - S3;
                    tile_alloc = inc;
                    // ---- telescoping per-node sums: work only where the node changes ----
                    unsigned key_next = __shfl_down_sync(FULL, k4.x, 1);
                    if (lane == 31) key_next = kNone;
                    const bool e0 = k4.x != k4.y, e1 = k4.y != k4.z, e2 = k4.z != k4.w;
                    // the tile's last run is closed by lane 31 on its own, so that the block below runs
                    // only in tiles where the node really changes (7 of 10 at 100 pods per node)
                    const bool e3 = (k4.w != key_next) & (lane != 31);
                    if (!kDirect && lane == 31 && k4.w < K32) red_add_u64(a.key_alloc + k4.w, (u64)(off + S3));
                    if (kDirect) {
                        if (a0) red_add_u64(a.key_alloc + k4.x, a0);
                        if (a1) red_add_u64(a.key_alloc + k4.y, a1);
                        if (a2) red_add_u64(a.key_alloc + k4.z, a2);
                        if (a3) red_add_u64(a.key_alloc + k4.w, a3);
                    } else if (e0 | e1 | e2 | e3) {
                        if (e0) {
                            const u64 v = off + S0;
                            if (k4.x < K32) red_add_u64(a.key_alloc + k4.x, v);
                            if (k4.y < K32) red_add_u64(a.key_alloc + k4.y, 0ull - v);
                        }
                        if (e1) {
                            const u64 v = off + S1;
                            if (k4.y < K32) red_add_u64(a.key_alloc + k4.y, v);
                            if (k4.z < K32) red_add_u64(a.key_alloc + k4.z, 0ull - v);
                        }
                        if (e2) {
                            const u64 v = off + S2;
                            if (k4.z < K32) red_add_u64(a.key_alloc + k4.z, v);
                            if (k4.w < K32) red_add_u64(a.key_alloc + k4.w, 0ull - v);
                        }
                        if (e3) {
                            const u64 v = off + S3;
                            if (k4.w < K32) red_add_u64(a.key_alloc + k4.w, v);
                            if (key_next < K32) red_add_u64(a.key_alloc + key_next, 0ull - v);
                        }
                    }
                    }
                }
            }
        }
        if (!fast) {                                   // warp-uniform
            if (n_deferred == (unsigned)kDeferred) { generic = true; break; }      // (`tile` is still to be done)
            if (lane == 0) s_deferred[warp][n_deferred] = tile;
            n_deferred++;
        }
        if (!kJob && lane == 31) acc_alloc += tile_alloc;
        ARENA_MARK(1);
    }
    if (generic)
        for (; tile < limit; tile += tile_step) {
            if (kJob) {
                const TileSums ts = process_tile_fallback_job(&a, tile, tile < n_full, lane, s_prefix[warp]);
                if (lane == 31) { job_req += ts.req; job_alloc += ts.alloc; }
            } else {
                const u64 t = process_tile_fallback(&a, tile, tile < n_full, lane, s_prefix[warp]).alloc;
                if (lane == 31) acc_alloc += t;
            }
        }
    // ---- the CTA's next run from the pool -----------------------------------------------------
    if (threadIdx.x == 0) s_claim[parity] = pending;
    __syncthreads();
    const unsigned run0 = n_static + s_claim[parity] * (unsigned)kPW;
    parity ^= 1u;
    if (run0 >= n_tiles) break;                            // the same for the whole CTA
    if (threadIdx.x == 0) pending = atomicAdd(pool, 1u);   // the run after this one
    tile = run0 + warp;
    limit = tile + 1 < n_tiles ? tile + 1 : n_tiles;       // exactly this tile (none when the last run is short)
    }
    __syncwarp();
    ARENA_MARK(2);
    for (unsigned k = 0; k < n_deferred; k++) {
        const unsigned tile = s_deferred[warp][k];
        if (kJob) {
            const TileSums ts = process_tile_fallback_job(&a, tile, tile < n_full, lane, s_prefix[warp]);
            if (lane == 31) { job_req += ts.req; job_alloc += ts.alloc; }
        } else {
            const u64 t = process_tile_fallback(&a, tile, tile < n_full, lane, s_prefix[warp]).alloc;
            if (lane == 31) acc_alloc += t;
        }
    }
    if (kJob) {
        __shared__ u64 s_wr[kPW];
        job_req = warp_reduce_sum(job_req);
        job_alloc = warp_reduce_sum(job_alloc);
        if (lane == 0) { s_wr[warp] = job_req; s_wa[warp] = job_alloc; }
        __syncthreads();
        u64 block_r = 0, block_al = 0;
        if (threadIdx.x == 0)
            for (int w = 0; w < kPW; w++) { block_r += s_wr[w]; block_al += s_wa[w]; }
        finish_totals(block_al, block_r, true, a.scratch, 2, a.total_alloc, a.total_req);
        return;
    }

    // ---- node capacities (gpuInNode, gpu.go:18-37): 9N bytes in, 8N out.  Done by the same
    // launch, after the pod tiles, where it fills the tail of the grid instead of costing a
    // second launch (node_alloc was zeroed by a memset before this kernel).
    u64 cap_sum = 0;
    for (u64 n = (u64)blockIdx.x * blockDim.x + threadIdx.x; n < a.n_nodes; n += (u64)gridDim.x * blockDim.x) {
        const i64 t = quantity_value(__ldg(a.cap_val + n), (int)__ldg(a.cap_scale + n));
        a.node_total[n] = t;
        cap_sum += (u64)t;
    }
    cap_sum = warp_reduce_sum(cap_sum);
    ARENA_MARK(3);

    __shared__ u64 s_wc[kPW];
    if (lane == 31) s_wa[warp] = acc_alloc;
    if (lane == 0) s_wc[warp] = cap_sum;
    __syncthreads();
    u64 block_a = 0, block_c = 0;
    if (threadIdx.x == 0)
        for (int w = 0; w < kPW; w++) { block_a += s_wa[w]; block_c += s_wc[w]; }
    ARENA_MARK(4);
    finish_totals(block_a, block_c, true, a.scratch, 1, a.total_alloc, a.total_capacity, a.px, a.px_lag);
    ARENA_MARK(5);
}

// ---------------------------------------------------------------------------
// The same summary kernel with the three POD columns staged by the bulk-copy engine (TMA), one
// iteration ahead.  The eight tiles a CTA works on in one iteration are 1024 consecutive pods, so
// their offsets, node ids and phases are three contiguous spans (4 KB + 4 KB + 1 KB): one thread
// requests the NEXT iteration's spans with three cp.async.bulk copies (no LSU request, no register)
// while the warps work on the current ones out of shared memory.  What remains per tile is ONE
// dependent global phase, the container rows.  Two stages, an mbarrier pair per stage: `full`
// (transaction bytes of the copies) and `empty` (one arrival per warp when it has read its tile).
// Grouped input, summary only (no pod_gpu, no per-pod adds): the launcher picks it for that case.
// ---------------------------------------------------------------------------
constexpr int kStagePods = kWarps * kWarpPods;                 // 1024 pods per CTA iteration
constexpr int kStageBytes = kStagePods * 9;                    // offsets + node ids + phases

__device__ __forceinline__ unsigned smem_u32(const void *p) { return (unsigned)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(u64 *bar, unsigned count)
{
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" :: "r"(smem_u32(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_expect_tx(u64 *bar, unsigned bytes)
{
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" :: "r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_arrive(u64 *bar)
{
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" :: "r"(smem_u32(bar)) : "memory");
}
// bounded: a copy that never lands must not hang the GPU (a trap fails the launch instead)
__device__ __forceinline__ void mbar_wait(u64 *bar, unsigned parity)
{
    const unsigned addr = smem_u32(bar);
    for (unsigned spins = 0;; spins++) {
        unsigned done;
        asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}"
                     : "=r"(done) : "r"(addr), "r"(parity) : "memory");
        if (done) return;
        if (spins > (1u << 26)) __trap();
    }
}
__device__ __forceinline__ void bulk_load(void *smem_dst, const void *gmem_src, unsigned bytes, u64 *bar)
{
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                 :: "r"(smem_u32(smem_dst)), "l"(gmem_src), "r"(bytes), "r"(smem_u32(bar)) : "memory");
}

__global__ void __launch_bounds__(kThreads, 6) top_node_pods_staged_kernel(const __grid_constant__ PodsArgs a)
{
    __shared__ __align__(16) u64 s_prefix[kWarps][kChunkCtrs + 2];
    __shared__ __align__(128) unsigned char s_stage[2][kStageBytes];
    __shared__ __align__(8) u64 s_full[2], s_empty[2];
    __shared__ u64 s_wa[kWarps];

    const int lane = threadIdx.x & 31;
    const int warp = threadIdx.x >> 5;
    const unsigned FULL = 0xffffffffu;
    unsigned *pre32 = reinterpret_cast<unsigned *>(s_prefix[warp]);

    const unsigned n_tiles = (unsigned)((a.P + kWarpPods - 1) / kWarpPods);
    const unsigned n_full = (unsigned)(a.P / kWarpPods);
    const unsigned K32 = (unsigned)(a.K < 0xFFFFFFFFull ? a.K : 0xFFFFFFFFull);
    const unsigned n_blocks = (n_tiles + kWarps - 1) / kWarps;     // CTA iterations in the whole launch

    if (threadIdx.x == 0) {
        mbar_init(&s_full[0], 1); mbar_init(&s_full[1], 1);
        mbar_init(&s_empty[0], kWarps); mbar_init(&s_empty[1], kWarps);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncthreads();
    // the full tiles of block `blk`, as bytes of each column, into stage `st`
    auto request = [&](unsigned blk, int st) {
        const u64 first = (u64)blk * kStagePods;
        const u64 lim = (u64)n_full * kWarpPods;
        if (first >= lim) return;                                  // no full tile in this block: nobody will wait
        const unsigned pods = (unsigned)(lim - first < (u64)kStagePods ? lim - first : (u64)kStagePods);
        unsigned char *dst = s_stage[st];
        mbar_expect_tx(&s_full[st], pods * 9u);
        bulk_load(dst, a.pod_ctr_begin + first, pods * 4u, &s_full[st]);
        bulk_load(dst + kStagePods * 4, a.pod_key + first, pods * 4u, &s_full[st]);
        bulk_load(dst + kStagePods * 8, a.pod_phase + first, pods, &s_full[st]);
    };
    if (threadIdx.x == 0 && blockIdx.x < n_blocks) request(blockIdx.x, 0);

    u64 acc_alloc = 0;
    unsigned it = 0;
    for (unsigned blk = blockIdx.x; blk < n_blocks; blk += gridDim.x, it++) {
        const int st = (int)(it & 1u);
        if (threadIdx.x == 0 && blk + gridDim.x < n_blocks) {      // the next iteration's spans
            if (it >= 1) mbar_wait(&s_empty[st ^ 1], ((it - 1) >> 1) & 1u);
            request(blk + gridDim.x, st ^ 1);
        }
        const unsigned tile = blk * kWarps + warp;
        u64 tile_alloc = 0;
        if (tile < n_tiles) {
            bool fast = tile < n_full;
            if (fast) {
                const u64 q = (u64)tile * kWarpPods + (u64)lane * kPodsPerLane;
                unsigned cb4 = 0;
                if (lane == 31) cb4 = __ldg(a.pod_ctr_begin + q + kPodsPerLane);
                mbar_wait(&s_full[st], (it >> 1) & 1u);
                const unsigned char *sg = s_stage[st];
                const uint4 c4 = *reinterpret_cast<const uint4 *>(sg + (warp * kWarpPods + lane * kPodsPerLane) * 4);
                {
                    const unsigned nxt = __shfl_down_sync(FULL, c4.x, 1);
                    if (lane != 31) cb4 = nxt;
                }
                const unsigned wc0 = __shfl_sync(FULL, c4.x, 0);
                const unsigned wc1 = __shfl_sync(FULL, cb4, 31);
                const unsigned base = wc0 & ~15u;
                const unsigned span = wc1 - base;
                fast = span <= (unsigned)kChunkCtrs && (u64)base + kChunkCtrs <= a.C;
                if (fast) {
                    const unsigned i0 = base + 2u * lane;
                    const bool two = span > (unsigned)kFastRow;
                    const ulonglong2 vA0 = ldg_stream(reinterpret_cast<const ulonglong2 *>(a.ctr_val + i0));
                    const ulonglong2 vB0 = ldg_stream(reinterpret_cast<const ulonglong2 *>(a.ctr_val + i0 + kHalf));
                    unsigned s0 = __ldg(reinterpret_cast<const unsigned short *>(a.ctr_scale + i0));
                    s0 |= (unsigned)__ldg(reinterpret_cast<const unsigned short *>(a.ctr_scale + i0 + kHalf)) << 16;
                    ulonglong2 vA1, vB1;
                    unsigned s1 = 0;
                    vA1.x = vA1.y = vB1.x = vB1.y = 0;
                    if (two) {
                        vA1 = ldg_stream(reinterpret_cast<const ulonglong2 *>(a.ctr_val + i0 + kFastRow));
                        vB1 = ldg_stream(reinterpret_cast<const ulonglong2 *>(a.ctr_val + i0 + kFastRow + kHalf));
                        s1 = __ldg(reinterpret_cast<const unsigned short *>(a.ctr_scale + i0 + kFastRow));
                        s1 |= (unsigned)__ldg(reinterpret_cast<const unsigned short *>(a.ctr_scale + i0 + kFastRow + kHalf)) << 16;
                    }
                    bool bad = ((vA0.x | vA0.y | vB0.x | vB0.y | vA1.x | vA1.y | vB1.x | vB1.y) >> kSmallBits) != 0ull;
                    bad |= (c4.x > c4.y) | (c4.y > c4.z) | (c4.z > c4.w) | (c4.w > cb4);
                    fast = !__any_sync(FULL, bad);
                    if (fast) {
                        unsigned total = fast_row((unsigned)vA0.x, (unsigned)vA0.y, (unsigned)vB0.x, (unsigned)vB0.y, s0,
                                                  0u, pre32, lane, bad);
                        if (two)
                            total = fast_row((unsigned)vA1.x, (unsigned)vA1.y, (unsigned)vB1.x, (unsigned)vB1.y, s1,
                                             total, pre32 + kFastRow, lane, bad);
                        fast = !__any_sync(FULL, bad);
                    }
                    if (fast) {
                        __syncwarp();
                        const unsigned b0 = pre32[c4.x - base], b1 = pre32[c4.y - base], b2 = pre32[c4.z - base],
                                       b3 = pre32[c4.w - base], b4 = pre32[cb4 - base];
                        __syncwarp();
                        const unsigned g0 = b1 - b0, g1 = b2 - b1, g2 = b3 - b2, g3 = b4 - b3;
                        const uint4 k4 = *reinterpret_cast<const uint4 *>(sg + kStagePods * 4 + (warp * kWarpPods + lane * kPodsPerLane) * 4);
                        const unsigned ph4 = *reinterpret_cast<const unsigned *>(sg + kStagePods * 8 + warp * kWarpPods + lane * kPodsPerLane);
                        const unsigned a0 = (k4.x < K32 && phase_active(ph4 & 0xffu)) ? g0 : 0u;
                        const unsigned a1 = (k4.y < K32 && phase_active((ph4 >> 8) & 0xffu)) ? g1 : 0u;
                        const unsigned a2 = (k4.z < K32 && phase_active((ph4 >> 16) & 0xffu)) ? g2 : 0u;
                        const unsigned a3 = (k4.w < K32 && phase_active(ph4 >> 24)) ? g3 : 0u;
                        const unsigned S0 = a0, S1 = S0 + a1, S2 = S1 + a2, S3 = S2 + a3;
                        const unsigned inc = warp_inclusive_scan_u32(S3);
                        const unsigned off = inc - S3;
                        tile_alloc = inc;
                        unsigned key_next = __shfl_down_sync(FULL, k4.x, 1);
                        if (lane == 31) key_next = kNone;
                        const bool e0 = k4.x != k4.y, e1 = k4.y != k4.z, e2 = k4.z != k4.w;
                        const bool e3 = (k4.w != key_next) & (lane != 31);
                        if (lane == 31 && k4.w < K32) red_add_u64(a.key_alloc + k4.w, (u64)(off + S3));
                        if (e0 | e1 | e2 | e3) {
                            if (e0) {
                                const u64 v = off + S0;
                                if (k4.x < K32) red_add_u64(a.key_alloc + k4.x, v);
                                if (k4.y < K32) red_add_u64(a.key_alloc + k4.y, 0ull - v);
                            }
                            if (e1) {
                                const u64 v = off + S1;
                                if (k4.y < K32) red_add_u64(a.key_alloc + k4.y, v);
                                if (k4.z < K32) red_add_u64(a.key_alloc + k4.z, 0ull - v);
                            }
                            if (e2) {
                                const u64 v = off + S2;
                                if (k4.z < K32) red_add_u64(a.key_alloc + k4.z, v);
                                if (k4.w < K32) red_add_u64(a.key_alloc + k4.w, 0ull - v);
                            }
                            if (e3) {
                                const u64 v = off + S3;
                                if (k4.w < K32) red_add_u64(a.key_alloc + k4.w, v);
                                if (key_next < K32) red_add_u64(a.key_alloc + key_next, 0ull - v);
                            }
                        }
                    }
                }
            }
            if (!fast) tile_alloc = process_tile_fallback(&a, tile, tile < n_full, lane, s_prefix[warp]).alloc;
            if (lane == 31) acc_alloc += tile_alloc;
        }
        __syncwarp();
        if (lane == 0) mbar_arrive(&s_empty[st]);                  // this warp is done with the stage
    }

    u64 cap_sum = 0;
    for (u64 n = (u64)blockIdx.x * blockDim.x + threadIdx.x; n < a.n_nodes; n += (u64)gridDim.x * blockDim.x) {
        const i64 t = quantity_value(__ldg(a.cap_val + n), (int)__ldg(a.cap_scale + n));
        a.node_total[n] = t;
        cap_sum += (u64)t;
    }
    cap_sum = warp_reduce_sum(cap_sum);
    __shared__ u64 s_wc[kWarps];
    if (lane == 31) s_wa[warp] = acc_alloc;
    if (lane == 0) s_wc[warp] = cap_sum;
    __syncthreads();
    u64 block_a = 0, block_c = 0;
    if (threadIdx.x == 0)
        for (int w = 0; w < kWarps; w++) { block_a += s_wa[w]; block_c += s_wc[w]; }
    finish_totals(block_a, block_c, true, a.scratch, 1, a.total_alloc, a.total_capacity, a.px, a.px_lag);
}

// The totals of the most recent step (posted by the launches of every rank) summed into out = {total capacity,
// allocated} (the mailbox carries them in the kernel's order: allocated first): the end of a
// sequence of deferred-collection steps.  One CTA of 32 threads.
__global__ void peer_collect_kernel(PeerExchange *px, i64 *out)
{
    u64 ga = 0, gb = 0;
    const u64 epoch = *((volatile unsigned long long *)&px->epoch);
    if (epoch == 0) {                             // nothing posted yet
        if (threadIdx.x == 0) { out[0] = 0; out[1] = 0; }
        return;
    }
    if ((int)threadIdx.x < px->world) peer_collect(px, epoch, ga, gb);
    ga = warp_reduce_sum(ga);
    gb = warp_reduce_sum(gb);
    const bool timed_out = *((volatile unsigned *)&px->timed_out) != 0u;
    if (threadIdx.x == 0) {
        out[1] = timed_out ? (i64)0x8000000000000000ull : (i64)ga;
        out[0] = timed_out ? (i64)0x8000000000000000ull : (i64)gb;
    }
}

// ---------------------------------------------------------------------------
// top job helpers
// ---------------------------------------------------------------------------
struct JobInitArgs {
    u64 J;
    u64 *job_req, *job_alloc;
    i64 *job_maxcreated;
    unsigned *job_latest;
};
__global__ void job_init_kernel(JobInitArgs a)
{
    for (u64 j = (u64)blockIdx.x * blockDim.x + threadIdx.x; j < a.J; j += (u64)gridDim.x * blockDim.x) {
        a.job_req[j] = 0;
        a.job_alloc[j] = 0;
        a.job_maxcreated[j] = (i64)0x8000000000000000ull;
        a.job_latest[j] = kNone;
    }
}

// Among ARENA_ROLE_LATEST pods whose timestamp equals the job's maximum, the
// FIRST in list order wins (`latest.Before(current)` is strict,
// trainer_mpi.go:271): atomicMin on the pod index.
struct JobLatestArgs {
    u64 P, J;
    const unsigned *pod_job;
    const unsigned char *pod_role;
    const i64 *pod_created;
    const i64 *job_maxcreated;
    unsigned *job_latest;
};
__global__ void job_latest_kernel(JobLatestArgs a)
{
    for (u64 p = (u64)blockIdx.x * blockDim.x + threadIdx.x; p < a.P; p += (u64)gridDim.x * blockDim.x) {
        if (a.pod_role[p] != 1) continue;
        unsigned j = a.pod_job[p];
        if (j >= a.J) continue;
        if (a.pod_created[p] == a.job_maxcreated[j]) atomicMin(a.job_latest + j, (unsigned)p);
    }
}

// Add the winning pod to its job, emit sort keys, and fold the winners into
// the totals the pods kernel produced.
struct JobFinishArgs {
    u64 J, Jpad;
    const unsigned *job_latest;
    const unsigned char *pod_phase;
    const unsigned *pod_ctr_begin;
    const i64 *ctr_val;
    const signed char *ctr_scale;
    u64 *job_req, *job_alloc;
    u64 *sort_key;          // [Jpad]
    unsigned *sort_idx;     // [Jpad]
    i64 *totals;            // {req, alloc}: in = sums of ROLE_COUNTS pods, out = full totals
    Scratch *scratch;
    struct SortPlan *plan;  // or nullptr: OR / AND of all sort keys (which key bits vary at all)
};
struct SortPlan { u64 or_acc, and_acc; };
__global__ void __launch_bounds__(kThreads) job_finish_kernel(JobFinishArgs a)
{
    u64 add_req = 0, add_alloc = 0;
    u64 key_or = 0, key_and = ~0ull;
    for (u64 j = (u64)blockIdx.x * blockDim.x + threadIdx.x; j < a.Jpad; j += (u64)gridDim.x * blockDim.x) {
        if (j >= a.J) {                      // padding sorts last
            a.sort_key[j] = ~0ull;
            a.sort_idx[j] = kNone;
            continue;
        }
        u64 req = a.job_req[j], alc = a.job_alloc[j];
        unsigned p = a.job_latest[j];
        if (p != kNone) {
            u64 g = 0;
            for (unsigned c = a.pod_ctr_begin[p]; c < a.pod_ctr_begin[p + 1]; c++)
                g += (u64)quantity_value(a.ctr_val[c], a.ctr_scale[c]);
            req += g;
            add_req += g;
            if (phase_active(a.pod_phase[p])) { alc += g; add_alloc += g; }
            a.job_req[j] = req;
            a.job_alloc[j] = alc;
        }
        // requested DESC as an ascending unsigned key: flip sign bit, invert
        const u64 key = ~(req ^ 0x8000000000000000ull);
        a.sort_key[j] = key;
        a.sort_idx[j] = (unsigned)j;
        key_or |= key;
        key_and &= key;
    }
    if (a.plan) {                            // one pair of atomics per warp
        const unsigned FULL = 0xffffffffu;
        const unsigned olo = __reduce_or_sync(FULL, (unsigned)key_or), ohi = __reduce_or_sync(FULL, (unsigned)(key_or >> 32));
        const unsigned alo = __reduce_and_sync(FULL, (unsigned)key_and), ahi = __reduce_and_sync(FULL, (unsigned)(key_and >> 32));
        if ((threadIdx.x & 31) == 0) {
            atomicOr(&a.plan->or_acc, ((u64)ohi << 32) | olo);
            atomicAnd(&a.plan->and_acc, ((u64)ahi << 32) | alo);
        }
    }
    __shared__ u64 s_a[kWarps], s_b[kWarps];
    add_req = warp_reduce_sum(add_req);
    add_alloc = warp_reduce_sum(add_alloc);
    if ((threadIdx.x & 31) == 0) { s_a[threadIdx.x >> 5] = add_req; s_b[threadIdx.x >> 5] = add_alloc; }
    __syncthreads();
    if (threadIdx.x == 0) {
        u64 r = 0, al = 0;
        for (int w = 0; w < kWarps; w++) { r += s_a[w]; al += s_b[w]; }
        if (r) red_add((u64 *)&a.totals[0], r);
        if (al) red_add((u64 *)&a.totals[1], al);
    }
}

// Bitonic sort of (key, idx) pairs, ascending by key then idx: a total order,
// so the result does not depend on the network being unstable.
__device__ __forceinline__ bool pair_less(u64 ka, unsigned ia, u64 kb, unsigned ib)
{
    return ka < kb || (ka == kb && ia < ib);
}

constexpr int kSortTile = 2048;    // pairs per CTA in shared memory (24 KB): the no-alloc entry point's one tile
constexpr int kBigSortTile = 4096; // the general path: 48 KB, fewer passes through global memory

// Runs every compare-exchange step with distance < kSortTile of stage `k`
// (or, when full_sort, all stages k = 2..kSortTile) inside shared memory.
template <int kSortTile>
__global__ void __launch_bounds__(1024) bitonic_tile_kernel(u64 *key, unsigned *idx, u64 n, u64 k_stage, int full_sort)
{
    __shared__ u64 s_key[kSortTile];
    __shared__ unsigned s_idx[kSortTile];
    const u64 base = (u64)blockIdx.x * kSortTile;
    for (int t = threadIdx.x; t < kSortTile; t += blockDim.x) {
        s_key[t] = key[base + t];
        s_idx[t] = idx[base + t];
    }
    __syncthreads();
    u64 k_first = full_sort ? 2 : k_stage;
    u64 k_last = full_sort ? (u64)kSortTile : k_stage;
    for (u64 k = k_first; k <= k_last; k <<= 1) {
        u64 j0 = k >> 1;
        if (j0 >= (u64)kSortTile) j0 = kSortTile >> 1;
        for (u64 j = j0; j > 0; j >>= 1) {
            for (int t = threadIdx.x; t < kSortTile / 2; t += blockDim.x) {
                // t-th compare pair of distance j
                unsigned lo = (unsigned)(((unsigned)t / (unsigned)j) * 2u * (unsigned)j + ((unsigned)t % (unsigned)j));
                unsigned hi = lo + (unsigned)j;
                bool up = (((base + lo) & k) == 0);
                u64 ka = s_key[lo], kb = s_key[hi];
                unsigned ia = s_idx[lo], ib = s_idx[hi];
                bool swap = up ? pair_less(kb, ib, ka, ia) : pair_less(ka, ia, kb, ib);
                if (swap) { s_key[lo] = kb; s_key[hi] = ka; s_idx[lo] = ib; s_idx[hi] = ia; }
            }
            __syncthreads();
        }
    }
    for (int t = threadIdx.x; t < kSortTile; t += blockDim.x) {
        key[base + t] = s_key[t];
        idx[base + t] = s_idx[t];
    }
}

// One compare-exchange step of distance j >= kSortTile in global memory.
__global__ void bitonic_global_kernel(u64 *key, unsigned *idx, u64 n, u64 k, u64 j)
{
    for (u64 t = (u64)blockIdx.x * blockDim.x + threadIdx.x; t < n / 2; t += (u64)gridDim.x * blockDim.x) {
        u64 lo = (t / j) * 2 * j + (t % j);
        u64 hi = lo + j;
        bool up = ((lo & k) == 0);
        u64 ka = key[lo], kb = key[hi];
        unsigned ia = idx[lo], ib = idx[hi];
        bool swap = up ? pair_less(kb, ib, ka, ia) : pair_less(ka, ia, kb, ib);
        if (swap) { key[lo] = kb; key[hi] = ka; idx[lo] = ib; idx[hi] = ia; }
    }
}

// ---- ordering of many jobs: LSD radix sort on the key bits that vary ----------------------------
// The sort key is `requested` as an ascending unsigned number (sign flipped, inverted: requested desc).
// Real clusters ask for a handful of GPUs per job, so of the key's 64 bits a few low ones vary and the
// rest is the same in every job.  The first `top job` query on a snapshot sorts with the bitonic
// network below and publishes WHICH bits vary (OR of all keys xor AND of all keys) to page-locked host
// memory; later queries on the same (immutable) snapshot radix-sort on exactly those bits, eight per
// pass -- one pass for any realistic snapshot.  LSD radix is stable, and the input is in id order, so
// the documented tie-break (ascending job id) needs no id in the key.
struct SortHost { u64 varmask; unsigned ready, pad; };

__global__ void sort_plan_kernel(SortPlan *plan, SortHost *host)
{
    host->varmask = plan->or_acc ^ plan->and_acc;
    __threadfence_system();
    *((volatile unsigned *)&host->ready) = 1u;
    plan->or_acc = 0;                        // ready for the next snapshot query
    plan->and_acc = ~0ull;
}

constexpr int kRadixTile = 2048;           // keys per CTA of 256 threads: 8 per thread

// this pass's digit: the key bits selected by `mask` (at most eight), packed together
__device__ __forceinline__ unsigned radix_digit(u64 key, u64 mask, int shift, int contiguous)
{
    if (contiguous) return (unsigned)((key & mask) >> shift);
    unsigned d = 0, i = 0;
    while (mask) {
        const int b = __ffsll((long long)mask) - 1;
        d |= (unsigned)((key >> b) & 1ull) << i++;
        mask &= mask - 1;
    }
    return d;
}

__global__ void __launch_bounds__(256) radix_hist_kernel(const u64 *__restrict__ key, u64 n, u64 mask, int shift, int contiguous,
                                                         unsigned *__restrict__ block_hist)
{
    __shared__ unsigned hist[256];
    hist[threadIdx.x] = 0;
    __syncthreads();
    const u64 base = (u64)blockIdx.x * kRadixTile;
#pragma unroll
    for (int k = 0; k < kRadixTile / 256; k++) {
        const u64 i = base + (u64)k * 256 + threadIdx.x;
        if (i < n) atomicAdd(&hist[radix_digit(key[i], mask, shift, contiguous)], 1u);
    }
    __syncthreads();
    block_hist[(u64)blockIdx.x * 256 + threadIdx.x] = hist[threadIdx.x];
}

// block_hist[b][d] becomes where block b's keys of digit d start in the output.  One CTA of 1024
// threads: four threads per digit walk a quarter of the blocks each, then the quarters and the digits
// are chained.
__global__ void __launch_bounds__(1024) radix_scan_kernel(unsigned *__restrict__ block_hist, unsigned n_blocks)
{
    __shared__ unsigned part[4][256];
    __shared__ unsigned digit_base[256];
    const unsigned d = threadIdx.x & 255, quarter = threadIdx.x >> 8;
    const unsigned per = (n_blocks + 3) / 4, b0 = quarter * per, b1 = b0 + per < n_blocks ? b0 + per : n_blocks;
    unsigned sum = 0;
    for (unsigned b = b0; b < b1; b++) sum += block_hist[(u64)b * 256 + d];
    part[quarter][d] = sum;
    __syncthreads();
    if (threadIdx.x < 256) {
        unsigned total = part[0][d] + part[1][d] + part[2][d] + part[3][d];
        digit_base[d] = total;
    }
    __syncthreads();
    if (threadIdx.x == 0) {                  // 256 values: a serial exclusive scan is a microsecond
        unsigned run = 0;
        for (int i = 0; i < 256; i++) { const unsigned t = digit_base[i]; digit_base[i] = run; run += t; }
    }
    __syncthreads();
    unsigned run = digit_base[d];
    for (unsigned qd = 0; qd < quarter; qd++) run += part[qd][d];
    for (unsigned b = b0; b < b1; b++) {
        const unsigned t = block_hist[(u64)b * 256 + d];
        block_hist[(u64)b * 256 + d] = run;
        run += t;
    }
}

// Stable scatter of one tile: warp w owns the tile's keys [256 w, 256 w + 256) and walks them 32 at a
// time in order; inside a round, equal digits are ranked by lane (match_any + population count of the
// lower lanes), across rounds by a per-warp running count, across warps by a prefix over the warps.
__global__ void __launch_bounds__(256) radix_scatter_kernel(const u64 *__restrict__ key_in, const unsigned *__restrict__ idx_in, u64 n,
                                                            u64 mask, int shift, int contiguous,
                                                            const unsigned *__restrict__ block_off, u64 *__restrict__ key_out,
                                                            unsigned *__restrict__ idx_out)
{
    __shared__ unsigned warp_cnt[8][256];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    for (int i = threadIdx.x; i < 8 * 256; i += 256) (&warp_cnt[0][0])[i] = 0;
    __syncthreads();
    const u64 base = (u64)blockIdx.x * kRadixTile + (u64)warp * 256;
    u64 k[8];
    unsigned id[8], dg[8], rank[8];
#pragma unroll
    for (int r = 0; r < 8; r++) {
        const u64 i = base + (u64)r * 32 + lane;
        const bool in = i < n;
        k[r] = in ? key_in[i] : 0ull;
        id[r] = in ? idx_in[i] : 0u;
        dg[r] = in ? radix_digit(k[r], mask, shift, contiguous) : 0xffffffffu;
    }
#pragma unroll
    for (int r = 0; r < 8; r++) {
        const unsigned peers = __match_any_sync(0xffffffffu, dg[r]);
        const unsigned below = __popc(peers & ((1u << lane) - 1u));
        unsigned start = 0;
        const int leader = __ffs(peers) - 1;
        if (dg[r] != 0xffffffffu && lane == leader) {
            start = warp_cnt[warp][dg[r]];
            warp_cnt[warp][dg[r]] = start + __popc(peers);
        }
        start = __shfl_sync(0xffffffffu, start, leader);
        rank[r] = start + below;
        __syncwarp();
    }
    __syncthreads();
    {   // exclusive prefix over the warps, per digit, plus this block's start for the digit
        const unsigned d = threadIdx.x;
        unsigned run = block_off[(u64)blockIdx.x * 256 + d];
#pragma unroll
        for (int w = 0; w < 8; w++) { const unsigned t = warp_cnt[w][d]; warp_cnt[w][d] = run; run += t; }
    }
    __syncthreads();
#pragma unroll
    for (int r = 0; r < 8; r++) {
        if (dg[r] == 0xffffffffu) continue;
        const unsigned pos = warp_cnt[warp][dg[r]] + rank[r];
        key_out[pos] = k[r];
        idx_out[pos] = id[r];
    }
}

// Samples adjacent pairs of the key column: *out = number of sampled pairs whose keys differ.
__global__ void probe_grouping_kernel(const unsigned *key, u64 P, unsigned n_samples, unsigned *out)
{
    unsigned diff = 0;
    for (unsigned i = blockIdx.x * blockDim.x + threadIdx.x; i < n_samples; i += gridDim.x * blockDim.x) {
        const u64 p = (u64)i * (P - 1) / n_samples;
        diff += key[p] != key[p + 1];
    }
    diff = __reduce_add_sync(0xffffffffu, diff);
    if ((threadIdx.x & 31) == 0 && diff) atomicAdd(out, diff);
}

// Calibration: how fast can this GPU merely READ the columns the pods kernel
// reads (no arithmetic to speak of)?  One 128-bit load per lane, grid-stride,
// over the five columns back to back.  bench.py reports it beside the roofline.
struct StreamReadArgs {
    const uint4 *col[5];
    u64 n16[5];               // 16-byte vectors per column
    u64 *sink;
};
__global__ void __launch_bounds__(kThreads) stream_read_kernel(StreamReadArgs a)
{
    u64 acc = 0;
    const u64 stride = (u64)gridDim.x * blockDim.x;
#pragma unroll 1
    for (int c = 0; c < 5; c++) {
        const uint4 *p = a.col[c];
        const u64 n = a.n16[c];
        u64 i = (u64)blockIdx.x * blockDim.x + threadIdx.x;
        for (; i + 3 * stride < n; i += 4 * stride) {          // 4 loads in flight per thread
            const uint4 v0 = ldg_stream(p + i), v1 = ldg_stream(p + i + stride), v2 = ldg_stream(p + i + 2 * stride),
                        v3 = ldg_stream(p + i + 3 * stride);
            acc += v0.x ^ v0.w ^ v1.y ^ v1.z ^ v2.x ^ v2.w ^ v3.y ^ v3.z;
        }
        for (; i < n; i += stride) { const uint4 v = ldg_stream(p + i); acc += v.x ^ v.w; }
    }
    if (acc == 0x9E3779B97F4A7C15ull) *a.sink = acc;            // never true in practice; defeats dead-code removal
}

// Calibration 2: the pods kernel's access PATTERN without its arithmetic or its
// dependent loads: per warp tile, 512 B of offsets + 512 B of node ids + 128 B of
// phases + the tile's share of the value and scale columns (rows placed by
// proportion, 2 KB + 256 B), five streams at once in grid-stride order.
__global__ void __launch_bounds__(kThreads) stream_pattern_kernel(PodsArgs a, u64 *sink)
{
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const unsigned n_full = (unsigned)(a.P / kWarpPods);
    const unsigned step = gridDim.x * kWarps;
    const u64 c_per_tile = n_full ? a.C / n_full : 0;
    u64 acc = 0;
    for (unsigned tile = blockIdx.x * kWarps + warp; tile < n_full; tile += step) {
        const u64 q = (u64)tile * kWarpPods + 4u * lane;
        const uint4 c4 = ldg_stream(reinterpret_cast<const uint4 *>(a.pod_ctr_begin + q));
        const uint4 k4 = ldg_stream(reinterpret_cast<const uint4 *>(a.pod_key + q));
        const unsigned ph = __ldg(reinterpret_cast<const unsigned *>(a.pod_phase + q));
        u64 base = ((u64)tile * c_per_tile) & ~15ull;
        if (base + kChunkCtrs > a.C) base = 0;
        const u64 i0 = base + 2u * lane;
        const ulonglong2 v0 = ldg_stream(reinterpret_cast<const ulonglong2 *>(a.ctr_val + i0));
        const ulonglong2 v1 = ldg_stream(reinterpret_cast<const ulonglong2 *>(a.ctr_val + i0 + 64));
        const ulonglong2 v2 = ldg_stream(reinterpret_cast<const ulonglong2 *>(a.ctr_val + i0 + 128));
        const ulonglong2 v3 = ldg_stream(reinterpret_cast<const ulonglong2 *>(a.ctr_val + i0 + 192));
        const unsigned s0 = __ldg(reinterpret_cast<const unsigned short *>(a.ctr_scale + i0));
        const unsigned s1 = __ldg(reinterpret_cast<const unsigned short *>(a.ctr_scale + i0 + 64));
        const unsigned s2 = __ldg(reinterpret_cast<const unsigned short *>(a.ctr_scale + i0 + 128));
        const unsigned s3 = __ldg(reinterpret_cast<const unsigned short *>(a.ctr_scale + i0 + 192));
        acc += c4.x ^ c4.w ^ k4.y ^ k4.z ^ ph ^ v0.x ^ v1.y ^ v2.x ^ v3.y ^ s0 ^ s1 ^ s2 ^ s3;
    }
    if (acc == 0x9E3779B97F4A7C15ull) *sink = acc;
}

// Quantity.Value() probe.
__global__ void quantity_values_kernel(const i64 *val, const signed char *scale, u64 n, i64 *out)
{
    for (u64 i = (u64)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (u64)gridDim.x * blockDim.x)
        out[i] = quantity_value(val[i], scale[i]);
}

}  // namespace arena

// ---- packed host snapshots: expansion into the wide layout (arena_top_node_host_packed) ---------
// pod_meta: low nibble phase, high nibble container count.  Tile = 1024 pods = one CTA of 256
// threads, four pods (one 32-bit word) per thread.
namespace arena {

constexpr int kPackTile = 1024;

__device__ __forceinline__ unsigned packed_counts(unsigned meta4) { return (meta4 >> 4) & 0x0f0f0f0fu; }

__device__ __forceinline__ unsigned cta256_exclusive_sum(unsigned mine, unsigned &total)
{
    __shared__ unsigned warp_tot[8];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    unsigned inc = mine;
#pragma unroll
    for (int off = 1; off < 32; off <<= 1) {
        const unsigned up = __shfl_up_sync(0xffffffffu, inc, off);
        if (lane >= off) inc += up;
    }
    if (lane == 31) warp_tot[warp] = inc;
    __syncthreads();
    unsigned before = 0, all = 0;
#pragma unroll
    for (int w = 0; w < 8; w++) {
        if (w < warp) before += warp_tot[w];
        all += warp_tot[w];
    }
    total = all;
    __syncthreads();
    return before + inc - mine;
}

// containers per tile
__global__ void __launch_bounds__(256) packed_tile_counts_kernel(const unsigned *__restrict__ meta4, u64 n_words,
                                                                 unsigned *__restrict__ tile_sum)
{
    const u64 w = (u64)blockIdx.x * 256 + threadIdx.x;
    const unsigned c = w < n_words ? packed_counts(__ldg(meta4 + w)) : 0u;
    unsigned total;
    cta256_exclusive_sum((c * 0x01010101u) >> 24, total);
    if (threadIdx.x == 0) tile_sum[blockIdx.x] = total;
}

// exclusive scan of the tile sums, in place: one CTA of 1024 threads
__global__ void __launch_bounds__(1024) packed_tile_scan_kernel(unsigned *__restrict__ tile_sum, u64 n_tiles)
{
    __shared__ unsigned part[1024];
    const u64 per = (n_tiles + 1023) / 1024;
    const u64 lo = (u64)threadIdx.x * per, hi = lo + per < n_tiles ? lo + per : n_tiles;
    unsigned s = 0;
    for (u64 t = lo; t < hi; t++) s += tile_sum[t];
    part[threadIdx.x] = s;
    __syncthreads();
    for (int off = 1; off < 1024; off <<= 1) {                 // Hillis-Steele: 10 rounds
        const unsigned v = threadIdx.x >= (unsigned)off ? part[threadIdx.x - off] : 0u;
        __syncthreads();
        part[threadIdx.x] += v;
        __syncthreads();
    }
    unsigned run = part[threadIdx.x] - s;
    for (u64 t = lo; t < hi; t++) { const unsigned v = tile_sum[t]; tile_sum[t] = run; run += v; }
}

// pod_ctr_begin and pod_phase of four pods per thread
__global__ void __launch_bounds__(256) packed_expand_pods_kernel(const unsigned *__restrict__ meta4, u64 n_words, u64 P, u64 C,
                                                                 const unsigned *__restrict__ tile_base,
                                                                 unsigned *__restrict__ pod_ctr_begin,
                                                                 unsigned *__restrict__ pod_phase4)
{
    const u64 w = (u64)blockIdx.x * 256 + threadIdx.x;
    const unsigned m = w < n_words ? __ldg(meta4 + w) : 0u;
    const unsigned c = packed_counts(m);
    unsigned total;
    const unsigned b0 = tile_base[blockIdx.x] + cta256_exclusive_sum((c * 0x01010101u) >> 24, total);
    if (w < n_words) {
        const unsigned b1 = b0 + (c & 0xffu), b2 = b1 + ((c >> 8) & 0xffu), b3 = b2 + ((c >> 16) & 0xffu);
        reinterpret_cast<uint4 *>(pod_ctr_begin)[w] = make_uint4(b0, b1, b2, b3);
        unsigned ph = m & 0x0f0f0f0fu;                           // nibble 15 = "some other phase" = 255
        ph |= __vcmpeq4(ph, 0x0f0f0f0fu) & 0xf0f0f0f0u;
        pod_phase4[w] = ph;
    }
    if (w == 0) pod_ctr_begin[P] = (unsigned)C;                 // (written after the vector store of a tiny snapshot too)
}

// ctr_val (int64, scale 0): four containers per thread from bytes, eight from nibbles
// (n_pairs = ceil(C / 2): the column has room for whole pairs, not for whole words of them)
template <int kBits>
__global__ void __launch_bounds__(256) packed_expand_ctrs_kernel(const unsigned *__restrict__ gpu4, u64 n_words, u64 n_pairs,
                                                                 longlong2 *__restrict__ ctr_val2)
{
    const u64 w = (u64)blockIdx.x * 256 + threadIdx.x;
    if (w >= n_words) return;
    const unsigned g = __ldg(gpu4 + w);
    if (kBits == 8) {
        if (2 * w < n_pairs) ctr_val2[2 * w] = make_longlong2((long long)(g & 0xffu), (long long)((g >> 8) & 0xffu));
        if (2 * w + 1 < n_pairs) ctr_val2[2 * w + 1] = make_longlong2((long long)((g >> 16) & 0xffu), (long long)(g >> 24));
    } else {
#pragma unroll
        for (int k = 0; k < 4; k++)
            if (4 * w + k < n_pairs)
                ctr_val2[4 * w + k] = make_longlong2((long long)((g >> (8 * k)) & 0xfu), (long long)((g >> (8 * k + 4)) & 0xfu));
    }
}

// pod_node: one warp per node fills its rows; warp N fills the rows that name no node
__global__ void __launch_bounds__(256) packed_expand_nodes_kernel(const unsigned *__restrict__ node_pod_begin, u64 N, u64 P,
                                                                  unsigned *__restrict__ pod_node)
{
    const u64 n = (u64)blockIdx.x * 8 + (threadIdx.x >> 5);
    if (n > N) return;
    const u64 lo = node_pod_begin[n], hi = n < N ? (u64)node_pod_begin[n + 1] : P;
    const unsigned id = n < N ? (unsigned)n : kNone;
    for (u64 r = lo + (threadIdx.x & 31); r < hi; r += 32) pod_node[r] = id;
}

}  // namespace arena

// ---- `top node` summary straight from the packed columns (no expansion) ---------------------------
// One HALF-warp per node: its pods are the rows [b[n], b[n+1]) and their containers the values from
// node_ctr_begin[n] on, so the join is a pair of offsets and nothing is shared between half-warps: no
// atomics, no pod_node column.  128 pods per trip: eight meta bytes per lane, a 16-lane scan of the
// container counts places every lane in the value column, each lane adds up its own pods' values in one
// pass (4-bit values) or pod by pod (8-bit values, lanes with dozens of containers).  Nodes
// above kBigNodePods rows (power-law snapshots) are cut by the host into pieces of at most that many
// rows, each with its own container offset; pieces are spread over the grid and meet in node_alloc
// with one atomic each.
namespace arena {

constexpr unsigned kBigNodePods = 2048;

struct PackedArgs {
    u64 N, n_pieces;                                    // this launch covers nodes [n_lo, N) and pieces [piece_lo, n_pieces)
    u64 n_lo, piece_lo;
    const unsigned *node_pod_begin, *node_ctr_begin;    // [N+1]
    const unsigned char *meta, *gpu;
    const uint4 *pieces;                                // {node, first row, end row, first container}
    const i64 *cap_val;
    const signed char *cap_scale;
    u64 *node_alloc;                                    // [N] zeroed before the launch
    i64 *node_total;
    i64 *total_alloc, *total_capacity;
    Scratch *scratch;
    PeerExchange *px;
};

// sum of the `cnt` values that start at value index i: eight nibbles (four bytes) per round, taken
// from two aligned words with a funnel shift and added up inside the register
template <int kBits>
__device__ __forceinline__ unsigned packed_values_sum(const unsigned char *gpu, unsigned i, unsigned cnt)
{
    constexpr unsigned kPer = kBits == 4 ? 8u : 4u;
    const unsigned *words = reinterpret_cast<const unsigned *>(gpu);
    unsigned s = 0;
    while (cnt) {
        const unsigned *w = words + i / kPer;
        unsigned x = __funnelshift_r(__ldg(w), __ldg(w + 1), (i % kPer) * kBits);
        const unsigned take = cnt < kPer ? cnt : kPer;
        if (take < kPer) x &= (1u << (take * kBits)) - 1u;
        if (kBits == 4) {
            x = (x & 0x0f0f0f0fu) + ((x >> 4) & 0x0f0f0f0fu);
            s += (x * 0x01010101u) >> 24;
        } else {
            x = (x & 0x00ff00ffu) + ((x >> 8) & 0x00ff00ffu);
            s += (x * 0x00010001u) >> 16;
        }
        i += take;
        cnt -= take;
    }
    return s;
}

// ---- the 4-bit fast lane: eight pods, at most sixteen values, no per-pod loop -----------------------
// The pods of a lane own consecutive runs of the value column, so "which values count" is a run-length
// code: a TOGGLE bit at the value index where the activity changes (pod j toggles at its first value
// when its activity differs from pod j-1's; pods without containers toggle on the same index and
// cancel), and the mask is the running parity of the toggles.  One shift per pod builds the toggle
// word, a 256-entry table in shared memory turns eight toggle bits into eight mask nibbles, and the
// sixteen values are added up inside two registers.
__device__ __forceinline__ unsigned shl_clamp(unsigned x, unsigned n)      // 0 when n >= 32 (PTX shl), not n mod 32
{
    unsigned r;
    asm("shl.b32 %0, %1, %2;" : "=r"(r) : "r"(x), "r"(n));
    return r;
}
// 1 in every byte whose phase nibble is counted (not Succeeded = 3, not Failed = 4)
__device__ __forceinline__ unsigned packed_active_bytes(unsigned meta4)
{
    const unsigned x = (meta4 & 0x0f0f0f0fu) + 0x05050505u;              // 3, 4 -> 8, 9: the only ones with (x & 0x1e) == 8
    const unsigned y = (x & 0x1e1e1e1eu) ^ 0x08080808u;                  // zero byte <=> not counted
    return ((y + 0x7f7f7f7fu) >> 7) & 0x01010101u;
}
// table entry b: nibble i = 0xF when an odd number of bits 0..i of b are set
__device__ __forceinline__ unsigned packed_parity_nibbles(unsigned b)
{
    unsigned m = 0, par = 0;
#pragma unroll
    for (int i = 0; i < 8; i++) {
        par ^= (b >> i) & 1u;
        m |= (0u - par) & (0xFu << (4 * i));
    }
    return m;
}
// cl / ch: container counts of pods 0..3 / 4..7, one per byte; al / ah: their activity bytes; the
// values start at nibble index `start`, and there are `total` <= 24 of them (the third word is only
// touched by the lanes that have more than 16)
__device__ __forceinline__ unsigned packed_lane_sum24(const unsigned *words, const unsigned *lut, unsigned cl, unsigned ch,
                                                      unsigned al, unsigned ah, unsigned start, unsigned total)
{
    const unsigned pre_lo = cl * 0x01010101u;
    const unsigned pre_hi = ch * 0x01010101u + (pre_lo >> 24) * 0x01010101u;
    const unsigned ex_lo = pre_lo - cl, ex_hi = pre_hi - ch;             // byte j = first value of pod j (0..16)
    const unsigned t_lo = al ^ (al << 8), t_hi = ah ^ __funnelshift_l(al, ah, 8);     // byte j = activity changes at pod j
    // shift counts: the pod's first value, or 32 + that (= no bit at all) when it does not toggle
    const unsigned sh_lo = ex_lo | ((t_lo ^ 0x01010101u) << 5), sh_hi = ex_hi | ((t_hi ^ 0x01010101u) << 5);
    unsigned d = shl_clamp(ah >> 24, pre_hi >> 24);                      // the last pod's run ends here
    d ^= shl_clamp(1u, __byte_perm(sh_lo, 0u, 0x4440u)) ^ shl_clamp(1u, __byte_perm(sh_lo, 0u, 0x4441u));
    d ^= shl_clamp(1u, __byte_perm(sh_lo, 0u, 0x4442u)) ^ shl_clamp(1u, __byte_perm(sh_lo, 0u, 0x4443u));
    d ^= shl_clamp(1u, __byte_perm(sh_hi, 0u, 0x4440u)) ^ shl_clamp(1u, __byte_perm(sh_hi, 0u, 0x4441u));
    d ^= shl_clamp(1u, __byte_perm(sh_hi, 0u, 0x4442u)) ^ shl_clamp(1u, __byte_perm(sh_hi, 0u, 0x4443u));
    const unsigned m_lo = lut[d & 0xffu];
    const unsigned m_hi = lut[(d >> 8) & 0xffu] ^ (0u - (__popc(d & 0xffu) & 1u));
    const unsigned *w = words + (start >> 3);
    const unsigned x0 = __ldg(w), x1 = __ldg(w + 1), x2 = __ldg(w + 2), sh = (start & 7u) * 4u;
    const unsigned v_lo = __funnelshift_r(x0, x1, sh) & m_lo, v_hi = __funnelshift_r(x1, x2, sh) & m_hi;
    unsigned y = (v_lo & 0x0f0f0f0fu) + ((v_lo >> 4) & 0x0f0f0f0fu) + (v_hi & 0x0f0f0f0fu) + ((v_hi >> 4) & 0x0f0f0f0fu);
    if (total > 16u) {
        const unsigned m_3 = lut[(d >> 16) & 0xffu] ^ (0u - (__popc(d & 0xffffu) & 1u));
        const unsigned v_3 = __funnelshift_r(x2, __ldg(w + 3), sh) & m_3;
        y += (v_3 & 0x0f0f0f0fu) + ((v_3 >> 4) & 0x0f0f0f0fu);          // bytes <= 90
    }
    y = (y & 0x00ff00ffu) + ((y >> 8) & 0x00ff00ffu);
    return (y * 0x00010001u) >> 16;                                      // <= 360
}

// Masked GPU sum of the rows [lo, hi) whose containers start at value index c0, by one HALF-WARP
// (16 lanes): the same value in every lane of the half.  Both halves of the warp are in here together
// and stay together (the trip count is the larger of the two; a half without rows left works on
// nothing), so every shuffle is a full-warp instruction of width 16.  At most kBigNodePods rows.  A lane owns EIGHT
// consecutive pods per trip, so a half-warp covers 128 rows at once and an ordinary node (about a
// hundred pods) is one trip; the two halves of a warp work on two different nodes, which doubles
// the number of independent load chains in flight (the kernel is a chain of dependent loads per
// node: offsets -> meta bytes -> container values).
//   meta bytes   two funnel-shifted 32-bit words (the rows start at any byte), rows past hi zeroed
//   counts       high nibbles; a multiply by 0x01010101 turns four count bytes into their running sums
//   placement    a 16-lane scan of the lane totals places every lane in the value column
//   values       each pod's few nibbles (bytes): packed_values_sum, skipped for pods without containers
template <int kBits>
__device__ __forceinline__ unsigned packed_rows_sum(const unsigned char *meta, const unsigned char *gpu, unsigned lo, unsigned hi,
                                                    unsigned c0, int l16, const unsigned *lut)
{
    const unsigned hmask = 0xffffffffu;
    const unsigned *mw = reinterpret_cast<const unsigned *>(meta);
    unsigned acc = 0;
    const unsigned rows = hi - lo;
    for (unsigned done = 0; __any_sync(hmask, done < rows); done += 128) {
        const unsigned q = lo + done + 8u * (unsigned)l16;
        unsigned m_lo = 0, m_hi = 0;
        if (q < hi) {
            const unsigned *w = mw + (q >> 2);
            const unsigned w0 = __ldg(w), w1 = __ldg(w + 1), w2 = __ldg(w + 2), sh = (q & 3u) * 8u;
            m_lo = __funnelshift_r(w0, w1, sh);
            m_hi = __funnelshift_r(w1, w2, sh);
        }
        {   // keep the rows of this lane that are inside the range: min(hi - q, 8) bytes (none when q >= hi)
            const int nv = max((int)(hi - q), 0), n4 = min(nv, 4), n8 = min(nv, 8);
            m_lo &= shl_clamp(1u, 8u * (unsigned)n4) - 1u;
            m_hi &= shl_clamp(1u, 8u * (unsigned)(n8 - n4)) - 1u;
        }
        const unsigned cnt_lo = (m_lo >> 4) & 0x0f0f0f0fu, cnt_hi = (m_hi >> 4) & 0x0f0f0f0fu;
        const unsigned pre_lo = cnt_lo * 0x01010101u;               // byte k = counts 0..k (at most 60: no carry)
        const unsigned pre_hi = cnt_hi * 0x01010101u + (pre_lo >> 24) * 0x01010101u;     // at most 120
        const unsigned total = pre_hi >> 24;
        unsigned inc = total;
#pragma unroll
        for (int d = 1; d < 16; d <<= 1) {
            const unsigned up = __shfl_up_sync(hmask, inc, d, 16);
            if (l16 >= d) inc += up;
        }
        const unsigned start = c0 + inc - total;
        c0 += __shfl_sync(hmask, inc, 15, 16);
        if (kBits == 4) {
            // whole lane at once when it has at most 24 values (three pods out of four have one or two
            // containers); two rounds of four pods when only the halves fit; the per-pod loop below for the
            // rest (a pod with more than a dozen containers)
            const unsigned tot_lo = pre_lo >> 24;
            unsigned rounds = 1;
            if (total > 24u) rounds = (tot_lo <= 24u && total - tot_lo <= 24u) ? 2u : 0u;
            if (rounds) {
                const unsigned a_lo = packed_active_bytes(m_lo), a_hi = packed_active_bytes(m_hi);
                const unsigned *words = reinterpret_cast<const unsigned *>(gpu);
                if (rounds == 1) acc += packed_lane_sum24(words, lut, cnt_lo, cnt_hi, a_lo, a_hi, start, total);
                else {
                    acc += packed_lane_sum24(words, lut, cnt_lo, 0u, a_lo, 0u, start, tot_lo);
                    acc += packed_lane_sum24(words, lut, cnt_hi, 0u, a_hi, 0u, start + tot_lo, total - tot_lo);
                }
                continue;
            }
        }
        const unsigned ex_lo = pre_lo - cnt_lo, ex_hi = pre_hi - cnt_hi;   // byte k = counts before pod k
#pragma unroll
        for (int j = 0; j < 8; j++) {
            const unsigned sh = 8u * (unsigned)(j & 3);
            const unsigned cnt = ((j < 4 ? cnt_lo : cnt_hi) >> sh) & 0xffu;
            if (cnt == 0) continue;                                  // (also every row past hi)
            const unsigned ph = ((j < 4 ? m_lo : m_hi) >> sh) & 0xfu;
            const unsigned off = ((j < 4 ? ex_lo : ex_hi) >> sh) & 0xffu;
            const unsigned v = packed_values_sum<kBits>(gpu, start + off, cnt);
            if (ph != kPhaseSucceeded && ph != kPhaseFailed) acc += v;
        }
    }
#pragma unroll
    for (int d = 8; d > 0; d >>= 1) acc += __shfl_xor_sync(hmask, acc, d, 16);      // at most 2048 * 15 * 255
    return acc;
}

template <int kBits>
__global__ void __launch_bounds__(kThreads) top_node_packed_kernel(const __grid_constant__ PackedArgs a)
{
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int l16 = lane & 15;
    const unsigned half = (unsigned)lane >> 4;
    __shared__ unsigned s_lut[256];
    if (kBits == 4) {
        s_lut[threadIdx.x] = packed_parity_nibbles(threadIdx.x);       // kThreads == 256
        __syncthreads();
    }
    // One unit of work per HALF-warp; the loops run on the warp's pair of units so that both halves make
    // the same trips (the odd one out at the end of the range gets an empty node).
    // (node and piece numbers fit 32 bits with room for one stride: check_packed)
    const unsigned n_units = gridDim.x * kWarps * 2, w0 = (blockIdx.x * kWarps + warp) * 2;
    const unsigned N = (unsigned)a.N, n_pieces = (unsigned)a.n_pieces;
    u64 alloc_sum = 0;                                  // lane 0 of every half
    unsigned lo_n = 0, hi_n = 0, c_n = 0;               // the next node's offsets, fetched one node ahead
    {
        const unsigned first = (unsigned)a.n_lo + w0 + half;
        if (first < N) { lo_n = __ldg(a.node_pod_begin + first); hi_n = __ldg(a.node_pod_begin + first + 1); c_n = __ldg(a.node_ctr_begin + first); }
    }
    for (unsigned nb = (unsigned)a.n_lo + w0; nb < N; nb += n_units) {
        const unsigned n = nb + half, nn = n + n_units;
        const unsigned lo = lo_n, c0 = c_n;
        unsigned hi = hi_n;
        lo_n = hi_n = c_n = 0;
        if (nn < N) {
            lo_n = __ldg(a.node_pod_begin + nn);
            hi_n = __ldg(a.node_pod_begin + nn + 1);
            c_n = __ldg(a.node_ctr_begin + nn);
        }
        const bool own = n < N && hi - lo <= kBigNodePods;              // larger nodes come as pieces, below
        if (!own) hi = lo;
        const unsigned s = packed_rows_sum<kBits>(a.meta, a.gpu, lo, hi, c0, l16, s_lut);
        if (l16 == 0 && own) { a.node_alloc[n] = (u64)s; alloc_sum += s; }
    }
    for (unsigned kb = (unsigned)a.piece_lo + w0; kb < n_pieces; kb += n_units) {
        const unsigned k = kb + half;
        uint4 pc = make_uint4(0u, 0u, 0u, 0u);
        if (k < n_pieces) pc = __ldg(a.pieces + k);
        const unsigned s = packed_rows_sum<kBits>(a.meta, a.gpu, pc.y, pc.z, pc.w, l16, s_lut);
        if (l16 == 0 && s) { red_add_u64(a.node_alloc + pc.x, (u64)s); alloc_sum += s; }
    }
    // node capacities (gpuInNode): one thread per node, after the rows
    u64 cap_sum = 0;
    for (u64 n = a.n_lo + (u64)blockIdx.x * blockDim.x + threadIdx.x; n < a.N; n += (u64)gridDim.x * blockDim.x) {
        const i64 t = quantity_value(__ldg(a.cap_val + n), (int)__ldg(a.cap_scale + n));
        a.node_total[n] = t;
        cap_sum += (u64)t;
    }
    cap_sum = warp_reduce_sum(cap_sum);
    alloc_sum += __shfl_down_sync(0xffffffffu, alloc_sum, 16);
    __shared__ u64 s_a[kWarps], s_c[kWarps];
    if (lane == 0) { s_a[warp] = alloc_sum; s_c[warp] = cap_sum; }
    __syncthreads();
    u64 block_a = 0, block_c = 0;
    if (threadIdx.x == 0)
        for (int w = 0; w < kWarps; w++) { block_a += s_a[w]; block_c += s_c[w]; }
    finish_totals(block_a, block_c, true, a.scratch, 2, a.total_alloc, a.total_capacity, a.px);
}

}  // namespace arena

// ---- `top job` for small clusters in ONE launch ------------------------------------------------
// BASELINE configs[1] (1e4 pods, 1e3 jobs) is latency, not bandwidth: the general path is seven
// launches of a few microseconds each.  Up to kSmallJobPods pods and kSmallJobs jobs everything --
// per-job sums, the latest-pod rule, totals, the ordering -- fits one CTA and its shared memory.
// Same arithmetic and the same tie rules as the general kernels above.
namespace arena {

constexpr unsigned kSmallJobs = 2048, kSmallJobPods = 65536;
constexpr size_t kSmallJobSmem = (size_t)kSmallJobs * (4 * 8 + 2 * 4);

struct JobSmallArgs {
    u64 P, J;
    const unsigned *pod_job;
    const unsigned char *pod_phase, *pod_role;
    const unsigned *pod_ctr_begin;
    const i64 *ctr_val;
    const signed char *ctr_scale;
    const i64 *pod_created;
    i64 *out_req, *out_alloc;
    unsigned *out_order, *out_latest;      // out_latest may be nullptr
    i64 *out_totals;
};

__device__ __forceinline__ u64 small_gpu_in_pod(const JobSmallArgs &a, u64 p)
{
    u64 g = 0;
    for (unsigned c = a.pod_ctr_begin[p]; c < a.pod_ctr_begin[p + 1]; c++)
        g += (u64)quantity_value(a.ctr_val[c], a.ctr_scale[c]);
    return g;
}

__global__ void __launch_bounds__(1024) job_small_kernel(const __grid_constant__ JobSmallArgs a)
{
    extern __shared__ __align__(16) unsigned char smem[];
    u64 *req = reinterpret_cast<u64 *>(smem), *alc = req + kSmallJobs, *key = alc + kSmallJobs;
    i64 *maxc = reinterpret_cast<i64 *>(key + kSmallJobs);
    unsigned *latest = reinterpret_cast<unsigned *>(maxc + kSmallJobs), *idx = latest + kSmallJobs;
    const unsigned tid = threadIdx.x;
    for (unsigned j = tid; j < kSmallJobs; j += 1024) {
        req[j] = 0; alc[j] = 0; maxc[j] = (i64)0x8000000000000000ull; latest[j] = kNone;
    }
    __syncthreads();
    u64 tot_req = 0, tot_alloc = 0;
    for (u64 p = tid; p < a.P; p += 1024) {
        const unsigned j = a.pod_job[p];
        if (j >= a.J) continue;
        if (a.pod_role[p] == 1) { atomicMax(reinterpret_cast<long long *>(maxc + j), (long long)a.pod_created[p]); continue; }
        const u64 g = small_gpu_in_pod(a, p);
        if (g) atomicAdd(reinterpret_cast<unsigned long long *>(req + j), g);
        tot_req += g;
        if (phase_active(a.pod_phase[p])) {
            if (g) atomicAdd(reinterpret_cast<unsigned long long *>(alc + j), g);
            tot_alloc += g;
        }
    }
    __syncthreads();
    for (u64 p = tid; p < a.P; p += 1024) {            // first pod in list order among the newest
        const unsigned j = a.pod_job[p];
        if (j < a.J && a.pod_role[p] == 1 && a.pod_created[p] == maxc[j]) atomicMin(latest + j, (unsigned)p);
    }
    __syncthreads();
    for (unsigned j = tid; j < kSmallJobs; j += 1024) {
        if (j >= a.J) { key[j] = ~0ull; idx[j] = kNone; continue; }
        u64 r = req[j], al = alc[j];
        const unsigned p = latest[j];
        if (p != kNone) {
            const u64 g = small_gpu_in_pod(a, p);
            r += g; tot_req += g;
            if (phase_active(a.pod_phase[p])) { al += g; tot_alloc += g; }
        }
        a.out_req[j] = (i64)r;
        a.out_alloc[j] = (i64)al;
        if (a.out_latest) a.out_latest[j] = p;
        key[j] = ~(r ^ 0x8000000000000000ull);          // requested DESC as an ascending key
        idx[j] = j;
    }
    // totals
    __shared__ u64 s_r[32], s_a[32];
    tot_req = warp_reduce_sum(tot_req);
    tot_alloc = warp_reduce_sum(tot_alloc);
    if ((tid & 31) == 0) { s_r[tid >> 5] = tot_req; s_a[tid >> 5] = tot_alloc; }
    __syncthreads();
    if (tid == 0) {
        u64 r = 0, al = 0;
        for (int w = 0; w < 32; w++) { r += s_r[w]; al += s_a[w]; }
        a.out_totals[0] = (i64)r;
        a.out_totals[1] = (i64)al;
    }
    // bitonic sort of the 2048 (key, id) pairs: ascending key, then id
    for (unsigned k = 2; k <= kSmallJobs; k <<= 1)
        for (unsigned d = k >> 1; d > 0; d >>= 1) {
            for (unsigned t = tid; t < kSmallJobs; t += 1024) {
                const unsigned u = t ^ d;
                if (u > t) {
                    const bool up = (t & k) == 0;
                    const u64 ka = key[t], kb = key[u];
                    const unsigned ia = idx[t], ib = idx[u];
                    const bool a_less = ka < kb || (ka == kb && ia < ib);
                    if (a_less != up) { key[t] = kb; key[u] = ka; idx[t] = ib; idx[u] = ia; }
                }
            }
            __syncthreads();
        }
    for (unsigned j = tid; j < a.J; j += 1024) a.out_order[j] = idx[j];
}

}  // namespace arena
