// PodList JSON bytes -> device-resident `top node` columns, parsed ON the GPU
// (the wire-format step in front of the accounting kernels).
//
//   H2D  the document, padded with spaces to a multiple of 64 bytes
//   K1   classify_words    one thread per 64-byte word: bitmasks of unescaped quotes and of
//                          brackets, plus the word's effect on (inside-string, depth) for both
//                          possible start states; per-CTA composition of those effects
//   K2   scan_ctas         one CTA: exclusive composition over the per-CTA effects
//   K3   word_states       (inside-string, depth) at every word start; the few containers that
//                          open at depth 2 (the root object's members) are reported to the host
//   host                   parses the ENVELOPE with the ordinary parser (json.cc), asking the
//                          device result where the "items" array ends -> [A, B)
//   K4   count_items       per word: brackets that open/close a depth-3 value inside (A, B)
//   scan                   exclusive prefix sums -> item index of every such bracket
//   K5   scatter_items     item k's byte range
//   K6   scan_items        ONE THREAD PER POD: pod_scan.cuh validates the item and extracts
//                          nodeName -> node id (hash join), phase, GPU limits per container
//   scan + K7 gather       CSR container columns in the final layout
//
// Anything the per-item scanner does not handle itself (escaped keys, duplicate keys, exotic
// quantities ...) is FLAGGED, and a flagged document is loaded by the host loader instead
// (arena_load_json + upload): results are identical either way, stats->fell_back says which.
// The kernels are byte/integer work bound by HBM (and, end to end, by the PCIe copy of the
// document): no tensor cores, 128-bit loads, grids sized to the data.
#include "../../include/arena_top.h"
#include "gpuacct_internal.h"
#include "host_snapshot.h"
#include "json.h"
#include "pod_scan.cuh"

#include <algorithm>
#include <chrono>
#include <cstdlib>
#include <cstring>
#include <mutex>
#include <string>
#include <thread>
#include <vector>

using arena_host::fail;
using namespace arena_scan;

namespace {

typedef unsigned long long u64;
typedef unsigned int u32;

#define CUDA_TRY(expr)                                                                   \
    do {                                                                                 \
        cudaError_t e_ = (expr);                                                         \
        if (e_ != cudaSuccess)                                                           \
            return fail(ARENA_E_CUDA, "%s failed: %s (%s:%d)", #expr, cudaGetErrorString(e_), \
                        __FILE__, __LINE__);                                             \
    } while (0)

constexpr int kWordBytes = 64;
constexpr int kCta = 256;                 // threads (= words) per CTA in the word kernels
constexpr u32 kMaxRootContainers = 1024;  // depth-2 containers reported to the host

// Effect of a span of text on the scan state: starting inside a string or not (s = 1 / 0),
// the span flips that state iff `parity`, and changes the bracket depth by d[s].
struct Effect {
    int parity, d0, d1;
};

__host__ __device__ inline Effect compose(const Effect &a, const Effect &b)        // a, then b
{
    Effect r;
    r.parity = a.parity ^ b.parity;
    r.d0 = a.d0 + (a.parity ? b.d1 : b.d0);
    r.d1 = a.d1 + (a.parity ? b.d0 : b.d1);
    return r;
}

__device__ inline u32 pack_effect(const Effect &e)
{
    return (u32)(e.parity & 1) | ((u32)(e.d0 & 0xff) << 8) | ((u32)(e.d1 & 0xff) << 16);
}
__device__ inline Effect unpack_effect(u32 w)
{
    Effect e;
    e.parity = (int)(w & 1u);
    e.d0 = (int)(signed char)((w >> 8) & 0xffu);
    e.d1 = (int)(signed char)((w >> 16) & 0xffu);
    return e;
}

__device__ inline Effect shfl_up_effect(const Effect &e, int off)
{
    Effect r;
    r.parity = __shfl_up_sync(0xffffffffu, e.parity, off);
    r.d0 = __shfl_up_sync(0xffffffffu, e.d0, off);
    r.d1 = __shfl_up_sync(0xffffffffu, e.d1, off);
    return r;
}

// Inclusive ordered scan of one Effect per thread over a CTA of kCta threads; returns the
// EXCLUSIVE prefix of the calling thread and, in `total`, the composition of the whole CTA.
__device__ inline Effect cta_exclusive_effects(Effect mine, Effect &total)
{
    __shared__ Effect warp_tot[kCta / 32];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    Effect inc = mine;
    for (int off = 1; off < 32; off <<= 1) {
        const Effect up = shfl_up_effect(inc, off);
        if (lane >= off) inc = compose(up, inc);
    }
    if (lane == 31) warp_tot[warp] = inc;
    __syncthreads();
    Effect before{0, 0, 0};
    for (int w = 0; w < warp; w++) before = compose(before, warp_tot[w]);
    Effect all = before;
    for (int w = warp; w < kCta / 32; w++) all = compose(all, warp_tot[w]);
    total = all;
    Effect prev = shfl_up_effect(inc, 1);
    if (lane == 0) prev = Effect{0, 0, 0};
    __syncthreads();
    return compose(before, prev);
}

__device__ inline u32 mask4(u32 eq) { return ((eq & 0x08040201u) * 0x01010101u) >> 24; }

__device__ inline u64 prefix_xor(u64 x)
{
    x ^= x << 1; x ^= x << 2; x ^= x << 4; x ^= x << 8; x ^= x << 16; x ^= x << 32;
    return x;
}

struct ScanResult {                       // small device -> host record
    u32 error;                            // structural scan could not cope (host loader decides)
    u32 n_root_open, n_root_close;
    int final_depth, final_in_string;
    u32 any_flag;                         // some item was flagged by the per-item scanner
    u32 first_flag_item, first_flag_reason;
    u64 n_opens, n_closes, n_containers;
    u32 root_open[kMaxRootContainers], root_close[kMaxRootContainers];
};

enum : u32 { kErrBackslashRun = 1, kErrRootOverflow = 2 };

// ---- K1 ----------------------------------------------------------------------------------
__global__ void __launch_bounds__(kCta) classify_words_kernel(const uint4 *__restrict__ text, u64 n_words,
                                                             u64 *__restrict__ quote_mask, u64 *__restrict__ open_mask,
                                                             u64 *__restrict__ close_mask, u32 *__restrict__ word_effect,
                                                             Effect *__restrict__ cta_effect, ScanResult *res)
{
    const u64 w = (u64)blockIdx.x * kCta + threadIdx.x;
    Effect mine{0, 0, 0};
    if (w < n_words) {
        u64 quote = 0, bslash = 0, open = 0, close = 0;
#pragma unroll
        for (int q = 0; q < 4; q++) {
            const uint4 v = __ldg(text + w * 4 + q);
            const u32 x[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
            for (int k = 0; k < 4; k++) {
                const int sh = 16 * q + 4 * k;
                const u32 lower = x[k] | 0x20202020u;                 // '[' -> '{', ']' -> '}'
                quote |= (u64)mask4(__vcmpeq4(x[k], 0x22222222u)) << sh;
                bslash |= (u64)mask4(__vcmpeq4(x[k], 0x5c5c5c5cu)) << sh;
                open |= (u64)mask4(__vcmpeq4(lower, 0x7b7b7b7bu)) << sh;
                close |= (u64)mask4(__vcmpeq4(lower, 0x7d7d7d7du)) << sh;
            }
        }
        // a quote is escaped when an ODD run of backslashes ends right before it; the run may
        // start in an earlier word
        bool carry = false;
        if (w > 0) {
            const unsigned char *bytes = reinterpret_cast<const unsigned char *>(text);
            u64 p = w * kWordBytes;
            int run = 0;
            while (p > 0 && bytes[p - 1] == '\\' && run < 4096) { p--; run++; }
            if (run == 4096) atomicOr(&res->error, kErrBackslashRun);
            carry = run & 1;
        }
        u64 escaped = 0;
        if (bslash | (u64)carry) {
            bool e = carry;
            for (int k = 0; k < 64; k++) {
                if (e) { escaped |= 1ull << k; e = false; }
                else if ((bslash >> k) & 1ull) e = true;
            }
        }
        quote &= ~escaped;
        const u64 inside = prefix_xor(quote);                         // assuming the word starts outside a string
        mine.parity = __popcll(quote) & 1;
        mine.d0 = __popcll(open & ~inside) - __popcll(close & ~inside);
        mine.d1 = __popcll(open & inside) - __popcll(close & inside);
        quote_mask[w] = quote;
        open_mask[w] = open;
        close_mask[w] = close;
        word_effect[w] = pack_effect(mine);
    }
    Effect total;
    cta_exclusive_effects(mine, total);
    if (threadIdx.x == 0) cta_effect[blockIdx.x] = total;
}

// ---- K2: one CTA ------------------------------------------------------------------------------
__global__ void __launch_bounds__(kCta) scan_ctas_kernel(const Effect *__restrict__ cta_effect, u64 n_ctas,
                                                        int2 *__restrict__ cta_state, ScanResult *res)
{
    const u64 per = (n_ctas + kCta - 1) / kCta;
    const u64 lo = (u64)threadIdx.x * per, hi = lo + per < n_ctas ? lo + per : n_ctas;
    Effect mine{0, 0, 0};
    for (u64 b = lo; b < hi; b++) mine = compose(mine, cta_effect[b]);
    Effect total;
    const Effect before = cta_exclusive_effects(mine, total);
    int in = before.parity, depth = before.d0;                       // the document starts outside, at depth 0
    for (u64 b = lo; b < hi; b++) {
        cta_state[b] = make_int2(in, depth);
        const Effect e = cta_effect[b];
        depth += in ? e.d1 : e.d0;
        in ^= e.parity;
    }
    if (threadIdx.x == 0) { res->final_in_string = total.parity; res->final_depth = total.d0; }
}

// Walks the structural brackets of one word.  f(bit, is_open, depth_after_open_or_before_close)
template <class F>
__device__ inline void walk_brackets(u64 quote, u64 open, u64 close, int in, int depth, F f)
{
    const u64 inside = prefix_xor(quote) ^ (in ? ~0ull : 0ull);
    u64 st = (open | close) & ~inside;
    while (st) {
        const int k = __ffsll((long long)st) - 1;
        st &= st - 1;
        if ((open >> k) & 1ull) { depth++; f(k, true, depth); }
        else { f(k, false, depth); depth--; }
    }
}

// ---- K3 --------------------------------------------------------------------------------------
__global__ void __launch_bounds__(kCta) word_states_kernel(const u64 *__restrict__ quote_mask, const u64 *__restrict__ open_mask,
                                                          const u64 *__restrict__ close_mask, const u32 *__restrict__ word_effect,
                                                          const int2 *__restrict__ cta_state, u64 n_words,
                                                          int *__restrict__ word_state, ScanResult *res)
{
    const u64 w = (u64)blockIdx.x * kCta + threadIdx.x;
    Effect mine{0, 0, 0};
    if (w < n_words) mine = unpack_effect(word_effect[w]);
    Effect total;
    const Effect before = cta_exclusive_effects(mine, total);
    if (w >= n_words) return;
    const int2 start = cta_state[blockIdx.x];
    const int in = start.x ^ before.parity;
    const int depth = start.y + (start.x ? before.d1 : before.d0);
    word_state[w] = depth * 2 + in;
    // the root object's container members: open at depth 2
    walk_brackets(quote_mask[w], open_mask[w], close_mask[w], in, depth, [&](int k, bool is_open, int d) {
            if (d != 2) return;
            u32 *count = is_open ? &res->n_root_open : &res->n_root_close;
            u32 *list = is_open ? res->root_open : res->root_close;
            const u32 slot = atomicAdd(count, 1u);
            if (slot < kMaxRootContainers) list[slot] = (u32)(w * kWordBytes + k);
        else atomicOr(&res->error, kErrRootOverflow);
    });
}

// ---- K4 / K5 -----------------------------------------------------------------------------------
// Items are the values that open at depth 3 strictly inside (A, B): the elements of the root
// object's "items" array, whose '[' sits at A and whose ']' at B - 1.
__global__ void __launch_bounds__(kCta) count_items_kernel(const u64 *__restrict__ quote_mask, const u64 *__restrict__ open_mask,
                                                          const u64 *__restrict__ close_mask, const int *__restrict__ word_state,
                                                          u64 first_word, u64 n_words_in_range, u64 A, u64 B,
                                                          u64 *__restrict__ counts)
{
    const u64 r = (u64)blockIdx.x * kCta + threadIdx.x;
    if (r >= n_words_in_range) return;
    const u64 w = first_word + r;
    const int st = word_state[w];
    u32 opens = 0, closes = 0;
    walk_brackets(quote_mask[w], open_mask[w], close_mask[w], st & 1, st >> 1, [&](int k, bool is_open, int d) {
        const u64 pos = w * kWordBytes + k;
        if (d != 3 || pos <= A || pos >= B - 1) return;
        if (is_open) opens++; else closes++;
    });
    counts[r] = (u64)opens | ((u64)closes << 32);
}

__global__ void __launch_bounds__(kCta) scatter_items_kernel(const u64 *__restrict__ quote_mask, const u64 *__restrict__ open_mask,
                                                            const u64 *__restrict__ close_mask, const int *__restrict__ word_state,
                                                            u64 first_word, u64 n_words_in_range, u64 A, u64 B,
                                                            const u64 *__restrict__ offsets, u32 *__restrict__ item_begin,
                                                            u32 *__restrict__ item_end)
{
    const u64 r = (u64)blockIdx.x * kCta + threadIdx.x;
    if (r >= n_words_in_range) return;
    const u64 w = first_word + r;
    const int st = word_state[w];
    const u64 off = offsets[r];
    u32 o = (u32)off, c = (u32)(off >> 32);
    walk_brackets(quote_mask[w], open_mask[w], close_mask[w], st & 1, st >> 1, [&](int k, bool is_open, int d) {
        const u64 pos = w * kWordBytes + k;
        if (d != 3 || pos <= A || pos >= B - 1) return;
        if (is_open) item_begin[o++] = (u32)pos;
        else item_end[c++] = (u32)pos + 1u;
    });
}

// ---- exclusive prefix sum over u64 (also used with two packed u32 counters) -------------------
constexpr int kScanPerThread = 4;
constexpr int kScanTile = kCta * kScanPerThread;

__device__ inline u64 cta_exclusive_sum(u64 mine, u64 &total)
{
    __shared__ u64 warp_tot[kCta / 32];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    u64 inc = mine;
    for (int off = 1; off < 32; off <<= 1) {
        const u64 up = __shfl_up_sync(0xffffffffu, inc, off);
        if (lane >= off) inc += up;
    }
    if (lane == 31) warp_tot[warp] = inc;
    __syncthreads();
    u64 before = 0, all = 0;
    for (int w = 0; w < kCta / 32; w++) {
        if (w < warp) before += warp_tot[w];
        all += warp_tot[w];
    }
    total = all;
    __syncthreads();
    return before + inc - mine;
}

__global__ void __launch_bounds__(kCta) scan_tiles_kernel(const u64 *__restrict__ in, u64 n, u64 *__restrict__ tile_sum)
{
    const u64 base = (u64)blockIdx.x * kScanTile + (u64)threadIdx.x * kScanPerThread;
    u64 s = 0;
    for (int k = 0; k < kScanPerThread; k++)
        if (base + k < n) s += in[base + k];
    u64 total;
    cta_exclusive_sum(s, total);
    if (threadIdx.x == 0) tile_sum[blockIdx.x] = total;
}

__global__ void __launch_bounds__(kCta) scan_tile_sums_kernel(u64 *__restrict__ tile_sum, u64 n_tiles, u64 *__restrict__ grand_total)
{
    const u64 per = (n_tiles + kCta - 1) / kCta;
    const u64 lo = (u64)threadIdx.x * per, hi = lo + per < n_tiles ? lo + per : n_tiles;
    u64 s = 0;
    for (u64 b = lo; b < hi; b++) s += tile_sum[b];
    u64 total;
    u64 run = cta_exclusive_sum(s, total);
    for (u64 b = lo; b < hi; b++) { const u64 v = tile_sum[b]; tile_sum[b] = run; run += v; }
    if (threadIdx.x == 0) *grand_total = total;
}

__global__ void __launch_bounds__(kCta) scan_apply_kernel(const u64 *__restrict__ in, u64 n, const u64 *__restrict__ tile_sum,
                                                         u64 *__restrict__ out)
{
    const u64 base = (u64)blockIdx.x * kScanTile + (u64)threadIdx.x * kScanPerThread;
    u64 v[kScanPerThread], s = 0;
    for (int k = 0; k < kScanPerThread; k++) { v[k] = base + k < n ? in[base + k] : 0; s += v[k]; }
    u64 total;
    u64 run = tile_sum[blockIdx.x] + cta_exclusive_sum(s, total);
    for (int k = 0; k < kScanPerThread; k++)
        if (base + k < n) { out[base + k] = run; run += v[k]; }
}

// ---- K6: one thread per pod ---------------------------------------------------------------------
struct ItemArgs {
    const unsigned char *text;
    const u32 *item_begin, *item_end;
    u64 n_items, A, B;
    unsigned lanes;                // pods per warp
    NodeTable nodes;
    u32 *pod_node;
    unsigned char *pod_phase, *pod_flag;
    u64 *pod_nctr;                 // u64: feeds the prefix sum
    long long *ctr_tmp;            // [kInlineCtrs][n_items]
    signed char *scale_tmp;        // [kInlineCtrs][n_items]
    ScanResult *res;
    PodMeta *meta;                 // [n_items], kMeta launches only
};

template <bool kMeta>
__global__ void __launch_bounds__(128) scan_items_kernel(const ItemArgs a, const uint4 *__restrict__ fused_table)
{
    // every lane indexes the table with its own state: shared memory, not constant
    __shared__ __align__(16) uint32_t s_fused[kFusedEntries];
    __shared__ __align__(128) unsigned char s_ring[128 * 64];          // [thread][4 windows of 16 bytes]
    const uint32_t fused_base = (uint32_t)__cvta_generic_to_shared(s_fused);
    for (int k = threadIdx.x; k < kFusedEntries / 4; k += blockDim.x) {
        uint4 v = __ldg(fused_table + k);                              // row offsets -> shared-memory addresses
        v.x += fused_base; v.y += fused_base; v.z += fused_base; v.w += fused_base;
        reinterpret_cast<uint4 *>(s_fused)[k] = v;
    }
    __syncthreads();
    // With few pods only the first `lanes` lanes of every warp take one: the walk is a long
    // dependent chain per lane, so a small job finishes sooner spread over many thin warps
    // (more warps per scheduler to switch between, fewer lanes dragging each other through
    // their rare paths) than packed into a few full ones.
    const unsigned lane = threadIdx.x & 31u;
    if (lane >= a.lanes) return;
    const u64 p = ((u64)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5)) * a.lanes + lane;
    if (p >= a.n_items) return;
    const u64 begin = a.item_begin[p], end = a.item_end[p];
    PodOut out;
    out.flag = kOk;
    const u64 prev_end = p == 0 ? a.A + 1 : (u64)a.item_end[p - 1];
    if (end <= begin || prev_end > begin || !gap_ok(a.text, prev_end, begin, p == 0)) out.flag = kFlagSeparator;
    else if (p + 1 == a.n_items && !gap_ok(a.text, end, a.B - 1, true)) out.flag = kFlagSeparator;
    else scan_pod_item<SharedLex, RingFeed, kMeta>(a.text, begin, end, a.nodes, SharedLex{fused_base},
                                                   RingFeed(a.text, s_ring + threadIdx.x * 64), out, kMeta ? a.meta + p : nullptr);
    if (out.flag) {
        atomicOr(&a.res->any_flag, 1u);
        const u32 old = atomicMin(&a.res->first_flag_item, (u32)p);
        (void)old;
        out.n_ctrs = 0;
        out.node = kNoNode;
    }
    a.pod_flag[p] = (unsigned char)out.flag;
    a.pod_node[p] = out.node;
    a.pod_phase[p] = out.phase;
    const u32 n = out.n_ctrs <= (u32)kInlineCtrs ? out.n_ctrs : 0u;
    a.pod_nctr[p] = n;
    for (u32 k = 0; k < n; k++) {
        a.ctr_tmp[(u64)k * a.n_items + p] = out.ctr_val[k];
        a.scale_tmp[(u64)k * a.n_items + p] = out.ctr_scale[k];
    }
}

// ---- K7 ----------------------------------------------------------------------------------------
__global__ void __launch_bounds__(kCta) gather_columns_kernel(u64 n_items, const u64 *__restrict__ ctr_begin64,
                                                             const u64 *__restrict__ pod_nctr, const long long *__restrict__ ctr_tmp,
                                                             const signed char *__restrict__ scale_tmp,
                                                             u64 n_containers, u32 *__restrict__ pod_ctr_begin,
                                                             long long *__restrict__ ctr_val, signed char *__restrict__ ctr_scale)
{
    const u64 p = (u64)blockIdx.x * kCta + threadIdx.x;
    if (p == 0) pod_ctr_begin[n_items] = (u32)n_containers;
    if (p >= n_items) return;
    const u64 b = ctr_begin64[p];
    pod_ctr_begin[p] = (u32)b;
    const u32 n = (u32)pod_nctr[p];
    // (an item the host parsed may hold more containers than the inline slots: those are patched in afterwards)
    for (u32 k = 0; k < n && k < (u32)kInlineCtrs; k++) {
        ctr_val[b + k] = ctr_tmp[(u64)k * n_items + p];
        ctr_scale[b + k] = scale_tmp[(u64)k * n_items + p];
    }
}

// ---- items handed back to the host: their rows patched into the device arrays ----------------------
struct PodPatch { u32 item, node, n_ctrs, phase; };
__global__ void apply_pod_patches_kernel(const PodPatch *__restrict__ patch, u32 n, u32 *__restrict__ pod_node,
                                         unsigned char *__restrict__ pod_phase, u64 *__restrict__ pod_nctr)
{
    const u32 k = blockIdx.x * blockDim.x + threadIdx.x;
    if (k >= n) return;
    const PodPatch p = patch[k];
    pod_node[p.item] = p.node;
    pod_phase[p.item] = (unsigned char)p.phase;
    pod_nctr[p.item] = p.n_ctrs;
}
struct CtrPatch { u32 item, k; long long val; long long scale; };
__global__ void apply_ctr_patches_kernel(const CtrPatch *__restrict__ patch, u32 n, const u32 *__restrict__ pod_ctr_begin,
                                         long long *__restrict__ ctr_val, signed char *__restrict__ ctr_scale)
{
    const u32 i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const CtrPatch p = patch[i];
    const u64 at = (u64)pod_ctr_begin[p.item] + p.k;
    ctr_val[at] = p.val;
    ctr_scale[at] = (signed char)p.scale;
}

// =================================== host side ====================================================

size_t pad16(size_t n) { return (n + 15) & ~(size_t)15; }
size_t pad256(size_t n) { return (n + 255) & ~(size_t)255; }

// Device workspace of the loader: a bump allocator over chunks that stay allocated between
// calls (cudaMalloc / cudaFree cost more than every kernel here).  A call that needed more
// than one chunk merges them afterwards, so a steady workload runs out of one allocation.
struct Chunk {
    char *base;
    size_t cap;
};
std::vector<Chunk> g_chunks;             // guarded by g_load_mu
std::mutex g_load_mu;

struct DeviceBuffers {
    size_t chunk = 0, used = 0, total = 0;
    template <class T>
    cudaError_t alloc(T **out, size_t bytes)
    {
        bytes = (bytes + 255) & ~(size_t)255;
        if (bytes == 0) bytes = 256;
        total += bytes;
        while (chunk < g_chunks.size() && g_chunks[chunk].cap - used < bytes) { chunk++; used = 0; }
        if (chunk == g_chunks.size()) {
            size_t cap = bytes > (size_t)(8u << 20) ? bytes : (size_t)(8u << 20);
            if (!g_chunks.empty() && g_chunks.back().cap * 2 > cap) cap = g_chunks.back().cap * 2;
            void *p = nullptr;
            cudaError_t e = cudaMalloc(&p, cap);
            if (e != cudaSuccess && cap > bytes) { cap = bytes; e = cudaMalloc(&p, cap); }
            if (e != cudaSuccess) { *out = nullptr; return e; }
            g_chunks.push_back(Chunk{(char *)p, cap});
            used = 0;
        }
        *out = (T *)(g_chunks[chunk].base + used);
        used += bytes;
        return cudaSuccess;
    }
    ~DeviceBuffers()
    {
        if (g_chunks.size() <= 1) return;
        for (Chunk &c : g_chunks) cudaFree(c.base);           // cudaFree waits for the device: nothing is in flight
        g_chunks.clear();
        void *p = nullptr;
        const size_t cap = total + total / 8;
        if (cudaMalloc(&p, cap) == cudaSuccess) g_chunks.push_back(Chunk{(char *)p, cap});
    }
};

// ---- pageable documents: staged upload ---------------------------------------------------------
// A caller's buffer is usually ordinary (pageable) memory -- a Go []byte, a Python bytes -- which the
// driver copies through its own small staging area at a fraction of the PCIe rate.  A large pageable
// document is therefore cut into 4 MiB chunks that kStageThreads host threads copy into page-locked
// slots (two per thread, kept between calls) and send from there, each on its own stream: the
// memcpys of several cores overlap one another and the DMA.
constexpr int kStageThreads = 4;
constexpr size_t kStageChunk = 4u << 20;
struct StageLane {
    void *slot[2] = {nullptr, nullptr};
    cudaEvent_t sent[2] = {nullptr, nullptr};
    cudaStream_t stream = nullptr;
};
constexpr int kStageDevices = 16;                 // lanes (streams, page-locked slots) are per device
StageLane g_stage_all[kStageDevices][kStageThreads];
std::mutex g_stage_mu_all[kStageDevices];

bool is_pageable(const void *p)
{
    cudaPointerAttributes at{};
    if (cudaPointerGetAttributes(&at, p) != cudaSuccess) { cudaGetLastError(); return true; }
    return at.type == cudaMemoryTypeUnregistered;
}

int staged_upload(unsigned char *d_dst, const char *src, size_t len, cudaStream_t st, int device)
{
    if (device < 0 || device >= kStageDevices) return fail(ARENA_E_INVALID, "device ordinal %d beyond the staging lanes", device);
    std::lock_guard<std::mutex> lock(g_stage_mu_all[device]);
    StageLane *g_stage = g_stage_all[device];
    for (int t = 0; t < kStageThreads; t++) {
        StageLane &l = g_stage[t];
        if (l.stream) continue;
        CUDA_TRY(cudaStreamCreateWithFlags(&l.stream, cudaStreamNonBlocking));
        for (int k = 0; k < 2; k++) {
            CUDA_TRY(cudaHostAlloc(&l.slot[k], kStageChunk, cudaHostAllocDefault));
            CUDA_TRY(cudaEventCreateWithFlags(&l.sent[k], cudaEventDisableTiming));
        }
    }
    const size_t n_chunks = (len + kStageChunk - 1) / kStageChunk;
    cudaError_t err[kStageThreads];
    std::vector<std::thread> pool;
    auto work = [&](int t) {
        err[t] = cudaSetDevice(device);
        StageLane &l = g_stage[t];
        int k = 0;
        for (size_t c = t; c < n_chunks && err[t] == cudaSuccess; c += kStageThreads, k ^= 1) {
            const size_t off = c * kStageChunk, n = std::min(kStageChunk, len - off);
            err[t] = cudaEventSynchronize(l.sent[k]);            // the slot's previous chunk has left
            if (err[t] != cudaSuccess) break;
            std::memcpy(l.slot[k], src + off, n);
            err[t] = cudaMemcpyAsync(d_dst + off, l.slot[k], n, cudaMemcpyHostToDevice, l.stream);
            if (err[t] == cudaSuccess) err[t] = cudaEventRecord(l.sent[k], l.stream);
        }
    };
    for (int t = 1; t < kStageThreads; t++) pool.emplace_back(work, t);
    work(0);
    for (std::thread &th : pool) th.join();
    for (int t = 0; t < kStageThreads; t++) {
        if (err[t] != cudaSuccess) return fail(ARENA_E_CUDA, "staged upload: %s", cudaGetErrorString(err[t]));
        for (int k = 0; k < 2; k++) CUDA_TRY(cudaStreamWaitEvent(st, g_stage[t].sent[k], 0));   // kernels wait for every chunk
    }
    return ARENA_OK;
}

int exclusive_scan(const u64 *d_in, u64 n, u64 *d_out, u64 *d_tile_sums, u64 *d_total, cudaStream_t st)
{
    const u64 tiles = (n + kScanTile - 1) / kScanTile;
    if (n == 0) { CUDA_TRY(cudaMemsetAsync(d_total, 0, sizeof(u64), st)); return ARENA_OK; }
    scan_tiles_kernel<<<(unsigned)tiles, kCta, 0, st>>>(d_in, n, d_tile_sums);
    scan_tile_sums_kernel<<<1, kCta, 0, st>>>(d_tile_sums, tiles, d_total);
    scan_apply_kernel<<<(unsigned)tiles, kCta, 0, st>>>(d_in, n, d_tile_sums, d_out);
    arena_internal::add_launches(3);
    CUDA_TRY(cudaGetLastError());
    return ARENA_OK;
}

}  // namespace

namespace arena_internal {

int copy_to_device(void *d_dst, const void *src, size_t bytes, cudaStream_t st)
{
    if (!bytes) return ARENA_OK;
    if (bytes >= 4 * kStageChunk && is_pageable(src)) {
        int device = 0;
        CUDA_TRY(cudaGetDevice(&device));
        CUDA_TRY(cudaStreamSynchronize(st));                     // earlier work on `st` that touches the target is done
        return staged_upload((unsigned char *)d_dst, (const char *)src, bytes, st, device);
    }
    CUDA_TRY(cudaMemcpyAsync(d_dst, src, bytes, cudaMemcpyHostToDevice, st));
    return ARENA_OK;
}

void free_loader_workspace()
{
    std::lock_guard<std::mutex> lock(g_load_mu);
    for (Chunk &c : g_chunks) cudaFree(c.base);
    g_chunks.clear();
    for (int d = 0; d < kStageDevices; d++) {
        std::lock_guard<std::mutex> stage_lock(g_stage_mu_all[d]);
        for (StageLane &l : g_stage_all[d]) {
            for (int k = 0; k < 2; k++) {
                if (l.slot[k]) cudaFreeHost(l.slot[k]);
                if (l.sent[k]) cudaEventDestroy(l.sent[k]);
            }
            if (l.stream) cudaStreamDestroy(l.stream);
            l = StageLane{};
        }
    }
}

}  // namespace arena_internal

namespace arena_scan_host {

// Open-addressing table for find_node(): at most half full.
void build_node_table(const std::vector<std::string> &names, std::vector<NodeSlot> &slots, std::vector<unsigned char> &blob)
{
    size_t cap = 2;
    while (cap < names.size() * 2 + 2) cap <<= 1;
    slots.assign(cap, NodeSlot{0, 0, 0, 0});
    blob.clear();
    for (size_t n = 0; n < names.size(); n++) {
        const std::string &s = names[n];
        const uint64_t h = hash_bytes((const unsigned char *)s.data(), s.size());
        size_t idx = h & (cap - 1);
        while (slots[idx].hash != 0) idx = (idx + 1) & (cap - 1);
        slots[idx] = NodeSlot{h, (uint32_t)n, (uint32_t)s.size(), (uint64_t)blob.size()};
        blob.insert(blob.end(), s.begin(), s.end());
    }
    if (blob.empty()) blob.push_back(0);
}

}  // namespace arena_scan_host

namespace {

struct RootContainers {
    std::vector<u32> open, close;            // sorted; k-th open pairs with k-th close
};

bool root_container_end(void *user, size_t open, size_t *end)
{
    const RootContainers *rc = (const RootContainers *)user;
    auto it = std::lower_bound(rc->open.begin(), rc->open.end(), (u32)open);
    if (it == rc->open.end() || *it != (u32)open) return false;
    const size_t k = (size_t)(it - rc->open.begin());
    if (k >= rc->close.size() || rc->close[k] <= open) return false;
    *end = (size_t)rc->close[k] + 1;
    return true;
}

double ms_since(std::chrono::steady_clock::time_point t0)
{
    return std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0).count();
}

// One pod record from what the scanner reported: the fields loader.cc extract_pod reads from its tree.
void rec_from_scan(const char *text, uint32_t node, uint8_t phase, uint32_t n_ctrs, const int64_t *ctr_val, size_t val_stride,
                   const int8_t *ctr_scale, const PodMeta &m, arena_host::PodRec &r)
{
    auto str = [&](int f) { return std::string(text + m.off[f], m.len[f]); };
    auto has = [&](int f) { return (m.present >> f) & 1u; };
    r.name = str(kMetaName);
    r.ns = str(kMetaNamespace);
    r.host_ip = str(kMetaHostIP);
    r.created = has(kMetaCreated) ? arena_host::canonical_time_unix(text + m.off[kMetaCreated]) : arena_host::kZeroTimeUnix;
    r.owner_is_job = m.owner_is_job != 0;
    r.labels.has = 0;
    if (has(kMetaRelease)) { r.labels.has |= arena_host::kHasRelease; r.labels.release = str(kMetaRelease); }
    if (has(kMetaApp)) { r.labels.has |= arena_host::kHasApp; r.labels.app = str(kMetaApp); }
    if (has(kMetaGroup)) { r.labels.has |= arena_host::kHasGroup; r.labels.group_name = str(kMetaGroup); }
    if (has(kMetaReplicaType)) { r.labels.has |= arena_host::kHasReplicaType; r.labels.replica_type = str(kMetaReplicaType); }
    if (has(kMetaReplicaIndex)) { r.labels.has |= arena_host::kHasReplicaIndex; r.labels.replica_index = str(kMetaReplicaIndex); }
    r.node = node;
    r.phase = phase;
    r.ctr_val.resize(n_ctrs);
    r.ctr_scale.resize(n_ctrs);
    for (uint32_t k = 0; k < n_ctrs; k++) { r.ctr_val[k] = ctr_val[k * val_stride]; r.ctr_scale[k] = ctr_scale[k * val_stride]; }
}

// Returns ARENA_OK with *handled == false when the document must go through the host loader.
// pods_out == nullptr: the `top node` columns are left on the device (*snap_out; `nodes` holds the capacities).
// pods_out != nullptr: every pod comes back as a host record with the job / printer fields too (nothing stays
// on the device; `nodes` and snap_out are not used).
int load_on_device(const char *podlist, size_t len, const std::vector<std::string> &node_names, const arena_host_snapshot *nodes,
                   arena_snapshot **snap_out, arena_load_device_stats *stats, bool *handled,
                   std::vector<arena_host::PodRec> *pods_out)
{
    *handled = false;
    cudaStream_t st;
    int sm_count = 0;
    int rc = arena_internal::ready(&st, &sm_count);
    if (rc) return rc;
    if (len == 0 || len >= 0xFFFFFF00ull) return ARENA_OK;              // byte offsets are u32 on the device
    const auto t_start = std::chrono::steady_clock::now();

    std::lock_guard<std::mutex> lock(g_load_mu);
    DeviceBuffers bufs;
    const u64 n_words = (len + kWordBytes - 1) / kWordBytes;
    const u64 n_ctas = (n_words + kCta - 1) / kCta;
    unsigned char *d_text;
    u64 *d_quote, *d_open, *d_close;
    u32 *d_effect;
    Effect *d_cta_effect;
    int2 *d_cta_state;
    int *d_word_state;
    ScanResult *d_res;
    CUDA_TRY(bufs.alloc(&d_text, n_words * kWordBytes + 512));      // the per-pod kernel reads and prefetches past the end
    CUDA_TRY(bufs.alloc(&d_quote, n_words * 8));
    CUDA_TRY(bufs.alloc(&d_open, n_words * 8));
    CUDA_TRY(bufs.alloc(&d_close, n_words * 8));
    CUDA_TRY(bufs.alloc(&d_effect, n_words * 4));
    CUDA_TRY(bufs.alloc(&d_word_state, n_words * 4));
    CUDA_TRY(bufs.alloc(&d_cta_effect, n_ctas * sizeof(Effect)));
    CUDA_TRY(bufs.alloc(&d_cta_state, n_ctas * sizeof(int2)));
    CUDA_TRY(bufs.alloc(&d_res, sizeof(ScanResult)));

    CUDA_TRY(cudaMemsetAsync(d_res, 0, offsetof(ScanResult, root_open), st));
    CUDA_TRY(cudaMemsetAsync(d_text + (n_words - 1) * kWordBytes, ' ', kWordBytes + 64, st));
    rc = arena_internal::copy_to_device(d_text, podlist, len, st);   // (after the padding memset)
    if (rc) return rc;
    if (stats) { CUDA_TRY(cudaStreamSynchronize(st)); stats->ms_h2d = ms_since(t_start); }

    classify_words_kernel<<<(unsigned)n_ctas, kCta, 0, st>>>((const uint4 *)d_text, n_words, d_quote, d_open, d_close, d_effect,
                                                            d_cta_effect, d_res);
    scan_ctas_kernel<<<1, kCta, 0, st>>>(d_cta_effect, n_ctas, d_cta_state, d_res);
    word_states_kernel<<<(unsigned)n_ctas, kCta, 0, st>>>(d_quote, d_open, d_close, d_effect, d_cta_state, n_words, d_word_state,
                                                         d_res);
    arena_internal::add_launches(3);
    CUDA_TRY(cudaGetLastError());

    std::vector<unsigned char> res_bytes(sizeof(ScanResult));
    ScanResult *res = (ScanResult *)res_bytes.data();
    CUDA_TRY(cudaMemcpyAsync(res, d_res, sizeof(ScanResult), cudaMemcpyDeviceToHost, st));
    CUDA_TRY(cudaStreamSynchronize(st));
    if (stats) stats->ms_structure = ms_since(t_start);
    if (res->error || res->final_depth != 0 || res->final_in_string != 0 || res->n_root_open != res->n_root_close)
        return ARENA_OK;                                                 // damaged or unusual: the host loader reports it

    // ---- envelope on the host -----------------------------------------------------------------
    RootContainers roots;
    roots.open.assign(res->root_open, res->root_open + res->n_root_open);
    roots.close.assign(res->root_close, res->root_close + res->n_root_close);
    std::sort(roots.open.begin(), roots.open.end());
    std::sort(roots.close.begin(), roots.close.end());
    arena_json::Tree envelope;
    arena_json::ParseError perr;
    arena_json::ParseOptions opt;
    std::vector<arena_json::Range> items_range;
    bool split = false;
    opt.filter = [](std::string_view, std::string_view key) { return key == "items"; };
    opt.split_key = "items";
    opt.split_out = &items_range;
    opt.split_done = &split;
    opt.split_hook = root_container_end;
    opt.split_user = &roots;
    if (!arena_json::parse(std::string_view(podlist, len), envelope, perr, opt)) return ARENA_OK;   // host reports the error
    u64 A = 0, B = 0;
    bool have_items = false;
    if (split && items_range.size() == 1) { A = items_range[0].begin; B = items_range[0].end; have_items = true; }
    else {
        // no "items", or "items": null  => zero pods; anything else (scalar items, duplicates) => host
        const arena_json::Value *items = envelope.root().is_object() ? envelope.root().get("items") : nullptr;
        if (!envelope.root().is_object() || (items && !items->is_null())) return ARENA_OK;
    }

    // ---- item ranges ------------------------------------------------------------------------------
    u64 P = 0;
    u32 *d_item_begin = nullptr, *d_item_end = nullptr;
    if (have_items) {
        const u64 first_word = A / kWordBytes, last_word = (B - 1) / kWordBytes;
        const u64 nw = last_word - first_word + 1;
        const u64 tiles = (nw + kScanTile - 1) / kScanTile;
        u64 *d_counts, *d_offsets, *d_tile_sums;
        CUDA_TRY(bufs.alloc(&d_counts, nw * 8));
        CUDA_TRY(bufs.alloc(&d_offsets, nw * 8));
        CUDA_TRY(bufs.alloc(&d_tile_sums, tiles * 8));
        const unsigned grid = (unsigned)((nw + kCta - 1) / kCta);
        count_items_kernel<<<grid, kCta, 0, st>>>(d_quote, d_open, d_close, d_word_state, first_word, nw, A, B, d_counts);
        arena_internal::add_launches(1);
        rc = exclusive_scan(d_counts, nw, d_offsets, d_tile_sums, &d_res->n_opens, st);
        if (rc) return rc;
        u64 packed = 0;
        CUDA_TRY(cudaMemcpyAsync(&packed, &d_res->n_opens, 8, cudaMemcpyDeviceToHost, st));
        CUDA_TRY(cudaStreamSynchronize(st));
        P = packed & 0xFFFFFFFFull;
        if ((packed >> 32) != P) return ARENA_OK;
        if (P) {
            CUDA_TRY(bufs.alloc(&d_item_begin, P * 4));
            CUDA_TRY(bufs.alloc(&d_item_end, P * 4));
            scatter_items_kernel<<<grid, kCta, 0, st>>>(d_quote, d_open, d_close, d_word_state, first_word, nw, A, B, d_offsets,
                                                        d_item_begin, d_item_end);
            arena_internal::add_launches(1);
            CUDA_TRY(cudaGetLastError());
        } else {
            // "items": [ ]  -- only whitespace may sit between the brackets
            for (u64 k = A + 1; k + 1 < B; k++)
                if (!detail::is_ws((unsigned char)podlist[k])) return ARENA_OK;
        }
    }

    // ---- node table ---------------------------------------------------------------------------------
    std::vector<NodeSlot> slots;
    std::vector<unsigned char> blob;
    arena_scan_host::build_node_table(node_names, slots, blob);
    NodeSlot *d_slots;
    unsigned char *d_blob;
    uint32_t *d_fused;
    std::vector<uint32_t> fused(kFusedEntries);
    build_fused(fused.data());
    CUDA_TRY(bufs.alloc(&d_fused, fused.size() * 4));
    CUDA_TRY(cudaMemcpyAsync(d_fused, fused.data(), fused.size() * 4, cudaMemcpyHostToDevice, st));
    CUDA_TRY(bufs.alloc(&d_slots, slots.size() * sizeof(NodeSlot)));
    CUDA_TRY(bufs.alloc(&d_blob, blob.size()));
    CUDA_TRY(cudaMemcpyAsync(d_slots, slots.data(), slots.size() * sizeof(NodeSlot), cudaMemcpyHostToDevice, st));
    CUDA_TRY(cudaMemcpyAsync(d_blob, blob.data(), blob.size(), cudaMemcpyHostToDevice, st));

    // ---- one thread per pod -------------------------------------------------------------------------
    u32 *d_pod_node_tmp = nullptr;
    unsigned char *d_pod_phase_tmp = nullptr, *d_pod_flag = nullptr;
    u64 *d_nctr = nullptr, *d_ctr_begin64 = nullptr, *d_tile_sums2 = nullptr;
    long long *d_ctr_tmp = nullptr;
    signed char *d_scale_tmp = nullptr;
    CtrPatch *d_ctr_patches = nullptr;
    u32 n_ctr_patches = 0;
    u64 C = 0;
    PodMeta *d_meta = nullptr;
    if (P && pods_out) CUDA_TRY(bufs.alloc(&d_meta, P * sizeof(PodMeta)));
    if (P) {
        CUDA_TRY(bufs.alloc(&d_pod_node_tmp, P * 4));
        CUDA_TRY(bufs.alloc(&d_pod_phase_tmp, P));
        CUDA_TRY(bufs.alloc(&d_pod_flag, P));
        CUDA_TRY(bufs.alloc(&d_nctr, P * 8));
        CUDA_TRY(bufs.alloc(&d_ctr_begin64, P * 8));
        CUDA_TRY(bufs.alloc(&d_tile_sums2, ((P + kScanTile - 1) / kScanTile) * 8));
        CUDA_TRY(bufs.alloc(&d_ctr_tmp, P * 8 * kInlineCtrs));
        CUDA_TRY(bufs.alloc(&d_scale_tmp, P * kInlineCtrs));
        const u32 none = 0xFFFFFFFFu;
        CUDA_TRY(cudaMemcpyAsync(&d_res->first_flag_item, &none, 4, cudaMemcpyHostToDevice, st));
        ItemArgs ia;
        ia.text = d_text; ia.item_begin = d_item_begin; ia.item_end = d_item_end;
        ia.n_items = P; ia.A = A; ia.B = B;
        ia.nodes = NodeTable{d_slots, (uint64_t)slots.size() - 1, d_blob};
        ia.pod_node = d_pod_node_tmp; ia.pod_phase = d_pod_phase_tmp; ia.pod_flag = d_pod_flag;
        ia.pod_nctr = d_nctr; ia.ctr_tmp = d_ctr_tmp; ia.scale_tmp = d_scale_tmp; ia.res = d_res;
        ia.meta = d_meta;
        // aim at >= 4 warps per SM sub-partition before filling the warps up
        const u64 warps_wanted = (u64)sm_count * 16;
        unsigned lanes = 32;
        while (lanes > 4 && (P + lanes - 1) / lanes < warps_wanted) lanes >>= 1;
        if (const char *env = std::getenv("ARENA_LOADER_LANES")) {     // tuning aid
            const long v = std::strtol(env, nullptr, 10);
            if (v == 1 || v == 2 || v == 4 || v == 8 || v == 16 || v == 32) lanes = (unsigned)v;
        }
        ia.lanes = lanes;
        const u64 n_warps = (P + lanes - 1) / lanes;
        if (pods_out) scan_items_kernel<true><<<(unsigned)((n_warps + 3) / 4), 128, 0, st>>>(ia, (const uint4 *)d_fused);
        else scan_items_kernel<false><<<(unsigned)((n_warps + 3) / 4), 128, 0, st>>>(ia, (const uint4 *)d_fused);
        arena_internal::add_launches(1);
        CUDA_TRY(cudaGetLastError());
        if (pods_out) {
            // ---- everything comes home: one record per pod, flagged items through the host's item code ----
            std::vector<unsigned char> flags(P), phase(P);
            std::vector<u32> node(P), ib, ie;
            std::vector<u64> nctr(P);
            std::vector<long long> vals(P * kInlineCtrs);
            std::vector<signed char> scales(P * kInlineCtrs);
            std::vector<PodMeta> meta(P);
            CUDA_TRY(cudaMemcpyAsync(res, d_res, offsetof(ScanResult, root_open), cudaMemcpyDeviceToHost, st));
            CUDA_TRY(cudaMemcpyAsync(flags.data(), d_pod_flag, P, cudaMemcpyDeviceToHost, st));
            CUDA_TRY(cudaMemcpyAsync(phase.data(), d_pod_phase_tmp, P, cudaMemcpyDeviceToHost, st));
            CUDA_TRY(cudaMemcpyAsync(node.data(), d_pod_node_tmp, P * 4, cudaMemcpyDeviceToHost, st));
            CUDA_TRY(cudaMemcpyAsync(nctr.data(), d_nctr, P * 8, cudaMemcpyDeviceToHost, st));
            CUDA_TRY(cudaMemcpyAsync(vals.data(), d_ctr_tmp, P * 8 * kInlineCtrs, cudaMemcpyDeviceToHost, st));
            CUDA_TRY(cudaMemcpyAsync(scales.data(), d_scale_tmp, P * kInlineCtrs, cudaMemcpyDeviceToHost, st));
            CUDA_TRY(cudaMemcpyAsync(meta.data(), d_meta, P * sizeof(PodMeta), cudaMemcpyDeviceToHost, st));
            CUDA_TRY(cudaStreamSynchronize(st));
            if (stats) { stats->n_items = P; stats->ms_items = ms_since(t_start); }
            if (res->any_flag) {
                ib.resize(P); ie.resize(P);
                CUDA_TRY(cudaMemcpyAsync(ib.data(), d_item_begin, P * 4, cudaMemcpyDeviceToHost, st));
                CUDA_TRY(cudaMemcpyAsync(ie.data(), d_item_end, P * 4, cudaMemcpyDeviceToHost, st));
                CUDA_TRY(cudaStreamSynchronize(st));
                if (stats) { stats->flagged_item = res->first_flag_item; stats->flag_reason = flags[res->first_flag_item]; }
            }
            arena_host::PodItemParser parser(node_names);
            const u64 max_handed = std::max<u64>(64, P / 8);
            u64 handed = 0;
            pods_out->assign(P, arena_host::PodRec());
            for (u64 k = 0; k < P; k++) {                       // the few flagged items first: any of them may end the attempt
                if (!flags[k]) continue;
                if (flags[k] == kFlagSeparator || ++handed > max_handed) return ARENA_OK;          // not handled
                if (!parser.parse_full(podlist, len, ib[k], ie[k], (*pods_out)[k])) return ARENA_OK;
            }
            // the rest is copying a dozen short strings per pod: spread over a few threads when there are many
            auto build = [&](u64 lo, u64 hi) {
                for (u64 k = lo; k < hi; k++)
                    if (!flags[k])
                        rec_from_scan(podlist, node[k], phase[k], (uint32_t)nctr[k], (const int64_t *)vals.data() + k, P,
                                      (const int8_t *)scales.data() + k, meta[k], (*pods_out)[k]);
            };
            const unsigned hw = std::thread::hardware_concurrency();
            const u64 n_thr = P < 4096 ? 1 : std::min<u64>(hw ? hw : 1, 8);
            if (n_thr <= 1) build(0, P);
            else {
                std::vector<std::thread> pool;
                const u64 per = (P + n_thr - 1) / n_thr;
                for (u64 t = 0; t < n_thr; t++) pool.emplace_back(build, std::min(P, t * per), std::min(P, (t + 1) * per));
                for (auto &th : pool) th.join();
            }
            if (stats) { stats->n_handed_back = handed; stats->ms_device = ms_since(t_start); }
            *handled = true;
            return ARENA_OK;
        }
        rc = exclusive_scan(d_nctr, P, d_ctr_begin64, d_tile_sums2, &d_res->n_containers, st);
        if (rc) return rc;
        CUDA_TRY(cudaMemcpyAsync(res, d_res, offsetof(ScanResult, root_open), cudaMemcpyDeviceToHost, st));
        CUDA_TRY(cudaStreamSynchronize(st));
        if (stats) { stats->n_items = P; stats->ms_items = ms_since(t_start); }
        if (res->any_flag) {
            // Only the flagged items go to the host: each is parsed by the host loader's own item code
            // (json.cc parse_range + loader.cc extract_pod) and its row is patched into the device arrays.
            // Should one of them fail there -- or should there be very many -- the whole document takes
            // the host path, which then reports the error exactly as a sequential decode would.
            std::vector<unsigned char> flags(P);
            std::vector<u32> ib(P), ie(P);
            CUDA_TRY(cudaMemcpyAsync(flags.data(), d_pod_flag, P, cudaMemcpyDeviceToHost, st));
            CUDA_TRY(cudaMemcpyAsync(ib.data(), d_item_begin, P * 4, cudaMemcpyDeviceToHost, st));
            CUDA_TRY(cudaMemcpyAsync(ie.data(), d_item_end, P * 4, cudaMemcpyDeviceToHost, st));
            CUDA_TRY(cudaStreamSynchronize(st));
            if (stats) { stats->flagged_item = res->first_flag_item; stats->flag_reason = flags[res->first_flag_item]; }
            std::vector<PodPatch> pod_patches;
            std::vector<CtrPatch> ctr_patches;
            arena_host::PodItemParser parser(nodes->node_names);
            std::vector<int64_t> pv;
            std::vector<int8_t> ps;
            const u64 max_handed = std::max<u64>(64, P / 8);
            bool give_up = false;
            for (u64 k = 0; k < P && !give_up; k++) {
                if (!flags[k]) continue;
                if (flags[k] == kFlagSeparator || pod_patches.size() >= max_handed) { give_up = true; break; }
                uint32_t node = 0;
                uint8_t phase = 0;
                if (!parser.parse(podlist, len, ib[k], ie[k], node, phase, pv, ps)) { give_up = true; break; }
                pod_patches.push_back(PodPatch{(u32)k, node, (u32)pv.size(), phase});
                for (size_t c = 0; c < pv.size(); c++) ctr_patches.push_back(CtrPatch{(u32)k, (u32)c, (long long)pv[c], (long long)ps[c]});
            }
            if (give_up) return ARENA_OK;                        // handled stays false: host loader + upload
            PodPatch *d_pp = nullptr;
            CUDA_TRY(bufs.alloc(&d_pp, pod_patches.size() * sizeof(PodPatch)));
            CUDA_TRY(cudaMemcpyAsync(d_pp, pod_patches.data(), pod_patches.size() * sizeof(PodPatch), cudaMemcpyHostToDevice, st));
            apply_pod_patches_kernel<<<(unsigned)((pod_patches.size() + 255) / 256), 256, 0, st>>>(d_pp, (u32)pod_patches.size(), d_pod_node_tmp,
                                                                                               d_pod_phase_tmp, d_nctr);
            arena_internal::add_launches(1);
            rc = exclusive_scan(d_nctr, P, d_ctr_begin64, d_tile_sums2, &d_res->n_containers, st);
            if (rc) return rc;
            CUDA_TRY(cudaMemcpyAsync(res, d_res, offsetof(ScanResult, root_open), cudaMemcpyDeviceToHost, st));
            CUDA_TRY(cudaStreamSynchronize(st));
            if (!ctr_patches.empty()) {
                CUDA_TRY(bufs.alloc(&d_ctr_patches, ctr_patches.size() * sizeof(CtrPatch)));
                CUDA_TRY(cudaMemcpyAsync(d_ctr_patches, ctr_patches.data(), ctr_patches.size() * sizeof(CtrPatch), cudaMemcpyHostToDevice, st));
                CUDA_TRY(cudaStreamSynchronize(st));             // the vector goes out of scope below
                n_ctr_patches = (u32)ctr_patches.size();
            }
            if (stats) stats->n_handed_back = pod_patches.size();
        }
        C = res->n_containers;
    }

    if (pods_out) {                                                // (no pods at all)
        pods_out->clear();
        if (stats) stats->ms_device = ms_since(t_start);
        *handled = true;
        return ARENA_OK;
    }

    // ---- final block: the layout arena_snapshot_upload() uses -----------------------------------------
    const u64 N = nodes->v.n_nodes;
    const size_t o_node = 0, o_phase = o_node + pad16(P * 4), o_cb = o_phase + pad16(P), o_val = o_cb + pad16((P + 1) * 4),
                 o_scale = o_val + pad16(C * 8), o_capv = o_scale + pad16(C), o_caps = o_capv + pad16(N * 8),
                 total = o_caps + pad16(N);
    char *block = nullptr;
    const size_t block_bytes = pad256(total ? total : 16);
    CUDA_TRY(arena_internal::device_alloc((void **)&block, block_bytes));
    cudaError_t e = cudaSuccess;
    auto step = [&](cudaError_t x) { if (e == cudaSuccess) e = x; };
    step(cudaMemsetAsync(block + o_cb, 0, pad16((P + 1) * 4), st));
    if (P) {
        step(cudaMemcpyAsync(block + o_node, d_pod_node_tmp, P * 4, cudaMemcpyDeviceToDevice, st));
        step(cudaMemcpyAsync(block + o_phase, d_pod_phase_tmp, P, cudaMemcpyDeviceToDevice, st));
        gather_columns_kernel<<<(unsigned)((P + kCta - 1) / kCta), kCta, 0, st>>>(P, d_ctr_begin64, d_nctr, d_ctr_tmp, d_scale_tmp, C,
                                                                                 (u32 *)(block + o_cb), (long long *)(block + o_val),
                                                                                 (signed char *)(block + o_scale));
        if (n_ctr_patches)
            apply_ctr_patches_kernel<<<(n_ctr_patches + 255) / 256, 256, 0, st>>>(d_ctr_patches, n_ctr_patches, (const u32 *)(block + o_cb),
                                                                               (long long *)(block + o_val), (signed char *)(block + o_scale));
        arena_internal::add_launches(1 + (n_ctr_patches ? 1 : 0));
        step(cudaGetLastError());
    }
    if (N) {
        step(cudaMemcpyAsync(block + o_capv, nodes->v.node_cap_val, N * 8, cudaMemcpyHostToDevice, st));
        step(cudaMemcpyAsync(block + o_caps, nodes->v.node_cap_scale, N, cudaMemcpyHostToDevice, st));
    }
    step(cudaStreamSynchronize(st));
    if (e != cudaSuccess) { arena_internal::device_release(block, block_bytes); return fail(ARENA_E_CUDA, "device loader: %s", cudaGetErrorString(e)); }

    arena_snapshot_view v{};
    v.n_pods = P; v.n_containers = C; v.n_nodes = N; v.n_jobs = 0;
    v.pod_node = (const uint32_t *)(block + o_node);
    v.pod_phase = (const uint8_t *)(block + o_phase);
    v.pod_ctr_begin = (const uint32_t *)(block + o_cb);
    v.ctr_val = (const int64_t *)(block + o_val);
    v.ctr_scale = (const int8_t *)(block + o_scale);
    v.node_cap_val = (const int64_t *)(block + o_capv);
    v.node_cap_scale = (const int8_t *)(block + o_caps);
    rc = arena_internal::adopt_device_block(block, block_bytes, v, snap_out);
    if (rc) return rc;
    if (stats) stats->ms_device = ms_since(t_start);
    *handled = true;
    return ARENA_OK;
}

}  // namespace

extern "C" {

int arena_pinned_alloc(size_t bytes, void **out)
{
    if (!out) return fail(ARENA_E_INVALID, "out is NULL");
    *out = nullptr;
    int rc = arena_internal::ready(nullptr, nullptr);
    if (rc) return rc;
    CUDA_TRY(cudaHostAlloc(out, bytes ? bytes : 1, cudaHostAllocDefault));
    return ARENA_OK;
}

void arena_pinned_free(void *p)
{
    if (p) cudaFreeHost(p);
}

int arena_load_json_device(const char *podlist_json, size_t podlist_len, const char *nodelist_json, size_t nodelist_len,
                           arena_host_snapshot **nodes_out, arena_snapshot **snapshot_out, arena_load_device_stats *stats)
{
    if (!nodes_out || !snapshot_out) return fail(ARENA_E_INVALID, "nodes_out / snapshot_out is NULL");
    *nodes_out = nullptr;
    *snapshot_out = nullptr;
    if (stats) std::memset(stats, 0, sizeof *stats);
    if (!podlist_json || !nodelist_json) return fail(ARENA_E_INVALID, "PodList and NodeList are required");
    int rc = arena_internal::ready(nullptr, nullptr);
    if (rc) return rc;                                                     // no CUDA, no loader: there is no CPU stand-in here

    // nodes: host loader with an empty PodList (N is small next to P; names and addresses stay on the host)
    static const char kNoPods[] = "{\"items\":[]}";
    arena_host_snapshot *nodes = nullptr;
    rc = arena_load_json(kNoPods, sizeof kNoPods - 1, nodelist_json, nodelist_len, nullptr, 0, nullptr, 0, nullptr, 0, 0, &nodes);
    if (rc) {
        // a sequential decode would have met a PodList syntax error first: let the host loader pick the message
        arena_host_snapshot *full = nullptr;
        int rc2 = arena_load_json(podlist_json, podlist_len, nodelist_json, nodelist_len, nullptr, 0, nullptr, 0, nullptr, 0, 0, &full);
        if (!rc2) { arena_host_snapshot_free(full); return fail(ARENA_E_PARSE, "NodeList failed to load"); }
        return rc2;
    }
    bool handled = false;
    rc = load_on_device(podlist_json, podlist_len, nodes->node_names, nodes, snapshot_out, stats, &handled, nullptr);
    if (rc) { arena_host_snapshot_free(nodes); return rc; }
    if (handled) { *nodes_out = nodes; return ARENA_OK; }

    // ---- host loader + upload: same columns, grouped by node -------------------------------------
    arena_host_snapshot_free(nodes);
    arena_host_snapshot *full = nullptr;
    rc = arena_load_json(podlist_json, podlist_len, nodelist_json, nodelist_len, nullptr, 0, nullptr, 0, nullptr, 0, 0, &full);
    if (rc) return rc;
    arena_snapshot_view hv;
    arena_host_snapshot_view(full, &hv);
    rc = arena_snapshot_upload(&hv, snapshot_out);
    if (rc) { arena_host_snapshot_free(full); return rc; }
    if (stats) stats->fell_back = 1;
    *nodes_out = full;
    return ARENA_OK;
}

int arena_load_json_gpu(const char *podlist_json, size_t podlist_len, const char *nodelist_json, size_t nodelist_len,
                        const char *joblist_json, size_t joblist_len, const char *tfjoblist_json, size_t tfjoblist_len,
                        const char *helm_list_text, size_t helm_len, int group_by_job, arena_host_snapshot **out,
                        arena_load_device_stats *stats)
{
    if (!out) return fail(ARENA_E_INVALID, "out is NULL");
    *out = nullptr;
    if (stats) std::memset(stats, 0, sizeof *stats);
    if (!podlist_json || !nodelist_json) return fail(ARENA_E_INVALID, "PodList and NodeList are required");
    int rc = arena_internal::ready(nullptr, nullptr);
    if (rc) return rc;                                                     // no CUDA, no loader: there is no CPU stand-in here
    struct DeviceSource : arena_host::PodSource {
        const char *text; size_t len; arena_load_device_stats *stats; int err = ARENA_OK;
        bool pods(const std::vector<std::string> &node_names, std::vector<arena_host::PodRec> &pods_out) override
        {
            bool handled = false;
            err = load_on_device(text, len, node_names, nullptr, nullptr, stats, &handled, &pods_out);
            return !err && handled;
        }
    } src;
    src.text = podlist_json; src.len = podlist_len; src.stats = stats;
    bool handled = false;
    rc = arena_host::load_with_pods(src, nodelist_json, nodelist_len, joblist_json, joblist_len, tfjoblist_json, tfjoblist_len,
                                    helm_list_text, helm_len, group_by_job != 0, out, &handled);
    if (src.err) return src.err;
    if (rc) return rc;
    if (handled) return ARENA_OK;
    if (stats) stats->fell_back = 1;
    return arena_load_json(podlist_json, podlist_len, nodelist_json, nodelist_len, joblist_json, joblist_len, tfjoblist_json,
                           tfjoblist_len, helm_list_text, helm_len, group_by_job, out);
}

// Diagnostic twin of arena_load_json_gpu: the same scanner source compiled for the host walks the items
// (no CUDA call), so the CPU test-suite can compare whole snapshots with arena_load_json.
int arena_load_json_hostscan(const char *podlist_json, size_t podlist_len, const char *nodelist_json, size_t nodelist_len,
                             const char *joblist_json, size_t joblist_len, const char *tfjoblist_json, size_t tfjoblist_len,
                             const char *helm_list_text, size_t helm_len, int group_by_job, arena_host_snapshot **out,
                             uint64_t *n_handed_back)
{
    if (!out) return fail(ARENA_E_INVALID, "out is NULL");
    *out = nullptr;
    if (n_handed_back) *n_handed_back = 0;
    if (!podlist_json || !nodelist_json) return fail(ARENA_E_INVALID, "PodList and NodeList are required");
    struct HostScanSource : arena_host::PodSource {
        const char *text; size_t len; uint64_t handed = 0;
        bool pods(const std::vector<std::string> &node_names, std::vector<arena_host::PodRec> &pods_out) override
        {
            std::vector<arena_json::Range> ranges;
            const std::string_view doc(text, len);
            size_t open = doc.find("\"items\"");
            if (open == std::string_view::npos) return false;
            open = doc.find('[', open);
            size_t end = 0;
            if (open == std::string_view::npos || !arena_json::split_array(doc, open, ranges, end)) return false;
            std::vector<NodeSlot> slots;
            std::vector<unsigned char> blob;
            arena_scan_host::build_node_table(node_names, slots, blob);
            const NodeTable nt{slots.data(), (uint64_t)slots.size() - 1, blob.data()};
            std::vector<uint32_t> fused(kFusedEntries);
            build_fused(fused.data());
            arena_host::PodItemParser parser(node_names);
            pods_out.assign(ranges.size(), arena_host::PodRec());
            for (size_t k = 0; k < ranges.size(); k++) {
                PodOut o = PodOut();
                PodMeta m;
                scan_pod_item<DirectLex, DirectFeed, true>((const unsigned char *)text, ranges[k].begin, ranges[k].end, nt,
                                                           DirectLex{fused.data()}, DirectFeed{(const unsigned char *)text}, o, &m);
                if (o.flag) {
                    handed++;
                    if (!parser.parse_full(text, len, ranges[k].begin, ranges[k].end, pods_out[k])) return false;
                    continue;
                }
                rec_from_scan(text, o.node, o.phase, o.n_ctrs, o.ctr_val, 1, o.ctr_scale, m, pods_out[k]);
            }
            return true;
        }
    } src;
    src.text = podlist_json; src.len = podlist_len;
    bool handled = false;
    int rc = arena_host::load_with_pods(src, nodelist_json, nodelist_len, joblist_json, joblist_len, tfjoblist_json, tfjoblist_len,
                                        helm_list_text, helm_len, group_by_job != 0, out, &handled);
    if (rc) return rc;
    if (n_handed_back) *n_handed_back = src.handed;
    if (handled) return ARENA_OK;
    return fail(ARENA_E_PARSE, "the scanner path did not take this document (arena_load_json reports why)");
}

int arena_scan_podlist_host(const char *podlist_json, size_t podlist_len, const arena_host_snapshot *nodes, uint64_t capacity,
                            uint64_t *n_items_out, uint32_t *pod_node, uint8_t *pod_phase, uint32_t *pod_n_ctrs,
                            int64_t *ctr_val, int8_t *ctr_scale, uint8_t *pod_flag)
{
    if (!podlist_json || !nodes || !n_items_out) return fail(ARENA_E_INVALID, "NULL argument");
    std::vector<arena_json::Range> ranges;
    const std::string_view text(podlist_json, podlist_len);
    size_t open = text.find("\"items\"");
    if (open == std::string_view::npos) return fail(ARENA_E_INVALID, "no \"items\" member");
    open = text.find('[', open);
    size_t end = 0;
    if (open == std::string_view::npos || !arena_json::split_array(text, open, ranges, end))
        return fail(ARENA_E_PARSE, "items array cannot be split");
    *n_items_out = ranges.size();
    if (ranges.size() > capacity) return fail(ARENA_E_INVALID, "capacity %llu < %zu items", (unsigned long long)capacity, ranges.size());
    std::vector<NodeSlot> slots;
    std::vector<unsigned char> blob;
    arena_scan_host::build_node_table(nodes->node_names, slots, blob);
    const NodeTable nt{slots.data(), (uint64_t)slots.size() - 1, blob.data()};
    std::vector<uint32_t> fused(kFusedEntries);
    build_fused(fused.data());
    for (size_t k = 0; k < ranges.size(); k++) {
        PodOut out = PodOut();
        scan_pod_item((const unsigned char *)podlist_json, ranges[k].begin, ranges[k].end, nt, DirectLex{fused.data()},
                      DirectFeed{(const unsigned char *)podlist_json}, out);
        if (pod_flag) pod_flag[k] = (uint8_t)out.flag;
        if (pod_node) pod_node[k] = out.node;
        if (pod_phase) pod_phase[k] = out.phase;
        if (pod_n_ctrs) pod_n_ctrs[k] = out.n_ctrs;
        if (ctr_val)
            for (int c = 0; c < kInlineCtrs; c++) ctr_val[k * kInlineCtrs + c] = (uint32_t)c < out.n_ctrs ? out.ctr_val[c] : 0;
        if (ctr_scale)
            for (int c = 0; c < kInlineCtrs; c++) ctr_scale[k * kInlineCtrs + c] = (uint32_t)c < out.n_ctrs ? out.ctr_scale[c] : 0;
    }
    return ARENA_OK;
}

}  // extern "C"
