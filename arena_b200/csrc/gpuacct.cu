// C ABI of libarena_gpuacct.so (see include/arena_gpuacct.h): device context,
// snapshot handles and kernel launches.  No torch types, no CPU fallback: if
// CUDA is unavailable every entry point fails with ARENA_E_CUDA / NOT_INIT.
#include "../../include/arena_gpuacct.h"
#include "kernels.cuh"
#include "host_snapshot.h"
#include "gpuacct_internal.h"

#include <algorithm>
#include <atomic>
#include <cstdarg>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <mutex>
#include <new>
#include <vector>

using namespace arena;

namespace {

using arena_host::fail;

#define CUDA_TRY(expr)                                                                   \
    do {                                                                                 \
        cudaError_t e_ = (expr);                                                         \
        if (e_ != cudaSuccess)                                                           \
            return fail(ARENA_E_CUDA, "%s failed: %s (%s:%d)", #expr, cudaGetErrorString(e_), \
                        __FILE__, __LINE__);                                             \
    } while (0)

struct Context {
    bool ready = false;
    int device = -1;
    int sm_count = 0;
    int pods_ctas_per_sm = 0, pods_long_ctas_per_sm = 0;
    int staged_ctas_per_sm = 0;      // > 0: the bulk-copy staged top-node kernel is used for grouped summaries
    int jobs_ctas_per_sm = 0;
    int packed_ctas_per_sm = 0;
    cudaStream_t stream = nullptr;
    // one-shot host calls: staging snapshot and packed staging block, kept between calls
    arena_snapshot *host_snap = nullptr;
    void *packed_block = nullptr;
    size_t packed_bytes = 0;
    bool pipe_ready = false;                     // streams / events of the pipelined packed one-shot call
    cudaStream_t pipe_stream[2] = {nullptr, nullptr};
    cudaEvent_t pipe_event[17] = {};
    // multi-GPU totals exchange: this device's mailbox, the peers' mappings, the descriptor
    PeerMailbox *my_box = nullptr;
    PeerExchange *px = nullptr;                  // device copy of the descriptor
    PeerExchange px_host{};
    void *peer_mapped[kMaxPeers] = {};           // CUDA IPC mappings to close (none for in-process peers)
};

// One context per GPU this process drives.  Slot 0 is the device of arena_acct_init(); the others
// exist after arena_acct_init_devices().  Every entry point works on the CALLING THREAD's current
// slot (0 unless the multi-device driver selected another), so the single-device API is unchanged.
constexpr int kMaxSlots = 16;
Context g_ctxs[kMaxSlots];
thread_local int t_slot = 0;
inline Context &ctx() { return g_ctxs[t_slot]; }
std::mutex g_mu, g_host_mu;
std::atomic<uint64_t> g_launches{0};

inline void count_launch() { g_launches.fetch_add(1, std::memory_order_relaxed); }

// Optional per-kernel timing for bench.py's roofline: CUDA events recorded on
// the launch stream around the nodes kernel and the pods kernel of every
// arena_top_node*() call between arena_profile_begin() and arena_profile_end().
struct Profile {
    bool on = false;
    unsigned every = 1, seen = 0;    // record every `every`-th call only (the three event records cost ~10 us per call)
    std::vector<cudaEvent_t> ev;     // 3 per call: before nodes, between, after pods
    size_t used = 0;
};
Profile g_prof;

size_t pad16(size_t n) { return (n + 15) & ~(size_t)15; }

}  // namespace

// Device-resident snapshot.  `owned` columns were allocated by upload().
struct arena_snapshot {
    arena_snapshot_view v{};     // device pointers
    bool owns = false;
    int slot = t_slot;           // the device context the handle lives on (the creating thread's current slot)
    void *block = nullptr;       // single allocation holding all owned columns
    size_t block_bytes = 0;
    Scratch *scratch = nullptr;  // per-handle: distinct handles may run concurrently
    bool direct = false;         // pods not grouped by node: per-pod adds instead of telescoping
    // top job work arrays (allocated on first use)
    u64 *job_req = nullptr, *job_alloc = nullptr;
    i64 *job_maxcreated = nullptr;
    unsigned *job_latest = nullptr;
    u64 *sort_key = nullptr;
    unsigned *sort_idx = nullptr;
    u64 jpad = 0;
    // radix sort of the job ordering (kernels.cuh): second key / id buffers, per-tile digit counts, the
    // OR / AND accumulators, and the page-locked word the first query publishes the varying bits to
    u64 *alt_key = nullptr;
    unsigned *alt_idx = nullptr, *block_hist = nullptr;
    SortPlan *sort_plan = nullptr;
    SortHost *sort_host = nullptr;
    // the `top node` step (memset + launch) as a CUDA graph, replayed while the arguments stay the same
    struct StepGraph {
        cudaGraphExec_t exec = nullptr;
        void *d_total = nullptr, *d_alloc = nullptr, *d_cluster = nullptr, *d_pod_gpu = nullptr, *px = nullptr;
        int exchange = 0;
        bool direct = false;
        arena_snapshot_view view{};  // the columns the captured launch reads (a staging snapshot is refilled between calls)
        void *scratch = nullptr;
    } step_graph;
    bool graph_broken = false;       // capture failed once on this handle: plain launches from then on
    // device output staging for the host-output entry points
    i64 *d_out = nullptr;
    size_t d_out_bytes = 0;
};

// multi-GPU totals exchange (arena_peer_exchange_*): this rank's mailbox, the peers' mappings
// A few freed device blocks are kept for the next snapshot of about the same size: a polling caller
// (load, query, free, again) otherwise pays cudaMalloc + cudaFree, which cost more than its kernels.
struct FreeBlock { void *p; size_t bytes; cudaEvent_t idle; int device; };   // idle: recorded on the library stream at release
static FreeBlock g_free_blocks[8] = {};
static std::mutex g_pool_mu;

extern "C" {

uint64_t arena_kernel_launches(void) { return g_launches.load(); }

// Binds the calling thread's current slot to `device` (g_mu held).
static int init_current_slot(int device)
{
    if (ctx().ready) {
        if (ctx().device == device) return ARENA_OK;
        return fail(ARENA_E_INVALID, "already initialised on device %d", ctx().device);
    }
    int n = 0;
    CUDA_TRY(cudaGetDeviceCount(&n));
    if (device < 0 || device >= n) return fail(ARENA_E_INVALID, "device %d out of range (have %d)", device, n);
    CUDA_TRY(cudaSetDevice(device));
    cudaDeviceProp prop;
    CUDA_TRY(cudaGetDeviceProperties(&prop, device));
    // the library carries one sm_100a cubin and no PTX: arch-specific code loads on exactly that chip
    if (prop.major != 10 || prop.minor != 0)
        return fail(ARENA_E_CUDA, "device %d is sm_%d%d; this library is built for sm_100a (B200) only and carries no "
                    "code for other architectures", device, prop.major, prop.minor);
    ctx().device = device;
    ctx().sm_count = prop.multiProcessorCount;
    CUDA_TRY(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&ctx().pods_ctas_per_sm, top_node_pods_kernel<false, false>, kPodThreads, 0));
    CUDA_TRY(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&ctx().pods_long_ctas_per_sm,
                                                           top_node_pods_kernel<false, false, false, kPodWarpsLong>, kPodThreadsLong, 0));
    {
        const char *env = getenv("ARENA_STAGED");                    // A/B switch while the variant is evaluated
        if (env && env[0] == '1')
            CUDA_TRY(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&ctx().staged_ctas_per_sm, top_node_pods_staged_kernel, kThreads, 0));
    }
    CUDA_TRY(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&ctx().jobs_ctas_per_sm, job_pods_kernel, kThreads, 0));
    CUDA_TRY(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&ctx().packed_ctas_per_sm, top_node_packed_kernel<4>, kThreads, 0));
    if (ctx().pods_ctas_per_sm < 1 || ctx().jobs_ctas_per_sm < 1 || ctx().packed_ctas_per_sm < 1)
        return fail(ARENA_E_CUDA, "kernel does not fit on an SM");
    CUDA_TRY(cudaStreamCreateWithFlags(&ctx().stream, cudaStreamNonBlocking));
    ctx().ready = true;
    return ARENA_OK;
}

int arena_acct_init(int device)
{
    std::lock_guard<std::mutex> lk(g_mu);
    arena_internal::UseSlot first(0);
    return init_current_slot(device);
}

void arena_acct_shutdown(void)
{
    {
    std::lock_guard<std::mutex> lk(g_mu);
    for (int slot = kMaxSlots - 1; slot >= 0; slot--) {
        arena_internal::UseSlot use(slot);
        if (!ctx().ready) continue;
        cudaSetDevice(ctx().device);
        cudaStreamSynchronize(ctx().stream);
        if (ctx().host_snap) { arena_snapshot_free(ctx().host_snap); ctx().host_snap = nullptr; }
        if (ctx().packed_block) { cudaFree(ctx().packed_block); ctx().packed_block = nullptr; ctx().packed_bytes = 0; }
        arena_peer_exchange_destroy();
        if (slot == 0) {
            arena_internal::device_pool_clear();
            arena_internal::free_loader_workspace();
        }
        if (ctx().pipe_ready) {
            cudaStreamDestroy(ctx().pipe_stream[0]);
            cudaStreamDestroy(ctx().pipe_stream[1]);
            for (cudaEvent_t e : ctx().pipe_event) if (e) cudaEventDestroy(e);
        }
        cudaStreamDestroy(ctx().stream);
        ctx() = Context{};
    }
    }
    arena_internal::multi_reset();           // (after g_mu is released: arena_acct_init_devices takes the two locks the other way round)
}

int arena_acct_device(void) { return g_ctxs[0].ready ? g_ctxs[0].device : (int)ARENA_E_NOT_INIT; }
int arena_acct_sm_count(void) { return g_ctxs[0].ready ? g_ctxs[0].sm_count : (int)ARENA_E_NOT_INIT; }

}  // extern "C"

namespace {

int check_ready()
{
    if (!ctx().ready) return fail(ARENA_E_NOT_INIT, "arena_acct_init() has not been called");
    cudaError_t e = cudaSetDevice(ctx().device);
    if (e != cudaSuccess) return fail(ARENA_E_CUDA, "cudaSetDevice: %s", cudaGetErrorString(e));
    return ARENA_OK;
}

// need_nodes: the `top node` columns (pod_node, node capacities) must be there; `top job` alone
// (arena_top_job_device_noalloc) reads neither.
int check_view(const arena_snapshot_view *v, bool device_ptrs, bool need_nodes = true)
{
    if (!v) return fail(ARENA_E_INVALID, "snapshot view is NULL");
    if (v->n_pods >= 0xFFFFFFFFull) return fail(ARENA_E_TOO_LARGE, "n_pods %llu >= 2^32-1", (unsigned long long)v->n_pods);
    if (v->n_containers > 0xFFFFFFFFull) return fail(ARENA_E_TOO_LARGE, "n_containers exceeds u32 offsets");
    if (v->n_nodes >= 0xFFFFFFFFull || v->n_jobs >= 0xFFFFFFFFull) return fail(ARENA_E_TOO_LARGE, "n_nodes / n_jobs >= 2^32-1");
    if (!v->pod_ctr_begin) return fail(ARENA_E_INVALID, "pod_ctr_begin is NULL (needs P+1 entries even when P == 0)");
    if (v->n_pods && ((need_nodes && !v->pod_node) || !v->pod_phase)) return fail(ARENA_E_INVALID, "pod columns are NULL");
    if (v->n_containers && (!v->ctr_val || !v->ctr_scale)) return fail(ARENA_E_INVALID, "container columns are NULL");
    if (need_nodes && v->n_nodes && (!v->node_cap_val || !v->node_cap_scale)) return fail(ARENA_E_INVALID, "node columns are NULL");
    if (v->n_jobs && v->n_pods && (!v->pod_job || !v->pod_role || !v->pod_created))
        return fail(ARENA_E_INVALID, "n_jobs > 0 but the job columns are NULL");
    const void *cols[] = {v->pod_node, v->pod_phase, v->pod_ctr_begin, v->ctr_val, v->ctr_scale,
                          v->node_cap_val, v->node_cap_scale, v->pod_job, v->pod_role, v->pod_created};
    if (device_ptrs)       // the kernels issue 128-bit loads on these; upload() re-aligns host columns
        for (const void *p : cols)
            if (((uintptr_t)p & 15) != 0) return fail(ARENA_E_INVALID, "a device column is not 16-byte aligned");
    return ARENA_OK;
}

int alloc_common(arena_snapshot *s)
{
    CUDA_TRY(arena_internal::device_alloc((void **)&s->scratch, sizeof(Scratch)));
    CUDA_TRY(cudaMemsetAsync(s->scratch, 0, sizeof(Scratch), ctx().stream));
    return ARENA_OK;
}

// Are the pods grouped by node?  Sample 4096 adjacent pairs of the DEVICE pod_node column; when
// more than a quarter of them change node the snapshot uses the per-pod-add kernel.  Either
// kernel is exact for any order, so this only picks the faster one.  Synchronises the stream.
int probe_grouping(arena_snapshot *s)
{
    s->direct = false;
    const u64 P = s->v.n_pods;
    if (P < 2 || !s->scratch) return ARENA_OK;
    const unsigned n = P - 1 < 4096 ? (unsigned)(P - 1) : 4096u;
    unsigned *flag = &s->scratch->ticket[3];
    CUDA_TRY(cudaMemsetAsync(flag, 0, sizeof(unsigned), ctx().stream));
    probe_grouping_kernel<<<16, 256, 0, ctx().stream>>>(s->v.pod_node, P, n, flag);
    count_launch();
    unsigned diff = 0;
    CUDA_TRY(cudaMemcpyAsync(&diff, flag, sizeof(unsigned), cudaMemcpyDeviceToHost, ctx().stream));
    CUDA_TRY(cudaStreamSynchronize(ctx().stream));
    CUDA_TRY(cudaMemsetAsync(flag, 0, sizeof(unsigned), ctx().stream));
    s->direct = diff * 4u > n;
    return ARENA_OK;
}

int grid_for(u64 work_items, u64 items_per_cta, int ctas_per_sm)
{
    u64 want = (work_items + items_per_cta - 1) / items_per_cta;
    u64 cap = (u64)ctx().sm_count * (u64)ctas_per_sm;      // one resident wave: persistent CTAs
    if (cap > (u64)kMaxPartials / 2) cap = kMaxPartials / 2;   // finish_totals: half of the partials per ticket id
    if (want < 1) want = 1;
    return (int)(want < cap ? want : cap);
}

int ensure_out(arena_snapshot *s, size_t bytes)
{
    if (s->d_out_bytes >= bytes) return ARENA_OK;
    if (s->d_out) cudaFree(s->d_out);
    s->d_out = nullptr;
    s->d_out_bytes = 0;
    CUDA_TRY(cudaMalloc((void **)&s->d_out, bytes));
    s->d_out_bytes = bytes;
    return ARENA_OK;
}

int ensure_job_arrays(arena_snapshot *s)
{
    if (s->job_req) return ARENA_OK;
    u64 J = s->v.n_jobs;
    u64 jpad = kSortTile;
    while (jpad < J) jpad <<= 1;
    const u64 n_tiles = (J + kRadixTile - 1) / kRadixTile;
    size_t bytes = pad16(J * 8) * 3 + pad16(J * 4) + pad16(jpad * 8) + pad16(jpad * 4) + pad16(J * 8) + pad16(J * 4) +
                   pad16(n_tiles * 256 * 4) + 16;
    char *base = nullptr;
    CUDA_TRY(cudaMalloc((void **)&base, bytes ? bytes : 16));
    CUDA_TRY(cudaHostAlloc((void **)&s->sort_host, sizeof(SortHost), cudaHostAllocMapped));
    *s->sort_host = SortHost{0, 0, 0};
    char *p = base;
    s->job_req = (u64 *)p; p += pad16(J * 8);
    s->job_alloc = (u64 *)p; p += pad16(J * 8);
    s->job_maxcreated = (i64 *)p; p += pad16(J * 8);
    s->sort_key = (u64 *)p; p += pad16(jpad * 8);
    s->job_latest = (unsigned *)p; p += pad16(J * 4);
    s->sort_idx = (unsigned *)p; p += pad16(jpad * 4);
    s->alt_key = (u64 *)p; p += pad16(J * 8);
    s->alt_idx = (unsigned *)p; p += pad16(J * 4);
    s->block_hist = (unsigned *)p; p += pad16(n_tiles * 256 * 4);
    s->sort_plan = (SortPlan *)p;
    const SortPlan fresh{0ull, ~0ull};
    CUDA_TRY(cudaMemcpy(s->sort_plan, &fresh, sizeof fresh, cudaMemcpyHostToDevice));
    s->jpad = jpad;
    return ARENA_OK;
}

// PodsArgs::n_static: how many tiles a launch shares out grid-stride before its CTAs turn to the pool: all but
// the last four rounds when the launch has at least 32 of them.  Shorter launches keep everything grid-stride: a
// pool round is a third slower than a fixed one (a barrier per run), which costs a 14-round launch more than the
// shorter tail gives back (1.27e7 pods: 67.8 -> 69.8 us; 1e8 pods: 431.6 -> 424.1 us; profiles/r02_iterations.md).
// ARENA_DYN_ROUNDS=r forces the last r rounds into the pool (0: none).
constexpr u64 kLongLaunchRounds = 32;
unsigned static_tiles(u64 n_pods, int grid, int warps_per_cta)
{
    const u64 W = (u64)grid * (u64)warps_per_cta, n_tiles = (n_pods + kWarpPods - 1) / kWarpPods;
    const u64 rounds = n_tiles / W;
    static const long forced = [] { const char *e = getenv("ARENA_DYN_ROUNDS"); return e ? std::strtol(e, nullptr, 10) : -1L; }();
    const u64 pool_rounds = forced >= 0 ? (u64)forced : (rounds >= kLongLaunchRounds ? 4 : 0);
    if (pool_rounds == 0) return 0xFFFFFFFFu;
    const u64 static_rounds = rounds > pool_rounds ? rounds - pool_rounds : 1;
    return (unsigned)std::min<u64>(static_rounds * W, 0xFFFFFFFFull);
}

template <typename K>
cudaError_t launch_pods(K kernel, int grid, PodsArgs pa, cudaStream_t st, int threads = kPodThreads)
{
    pa.n_static = static_tiles(pa.P, grid, threads / 32);
    kernel<<<grid, threads, 0, st>>>(pa);
    return cudaGetLastError();
}

// exchange: 0 none, 1 in-kernel totals exchange collected by the same launch, 2 collected one step late
int launch_top_node_plain(const arena_snapshot *s, i64 *d_total, i64 *d_alloc, i64 *d_cluster, i64 *d_pod_gpu,
                          cudaStream_t st, int exchange, bool prof);

// A step is a memset and a launch with the same arguments every time a caller polls one snapshot into
// the same buffers.  The pair is captured once per (buffers, stream kind, exchange mode) into a CUDA
// graph kept on the handle and replayed afterwards: one submission instead of two, no gap between the
// memset and the kernel (1.25e7 pods: 65.4 -> 63.7 us per step, tools/kbench_step.py).  Profiled calls,
// a stream that is itself being captured, and the legacy default stream (which cannot be captured)
// take the plain launches.  ARENA_NO_GRAPH=1 switches the replay off.
int launch_top_node(const arena_snapshot *cs, i64 *d_total, i64 *d_alloc, i64 *d_cluster, i64 *d_pod_gpu,
                    cudaStream_t st, int exchange = 0)
{
    arena_snapshot *s = const_cast<arena_snapshot *>(cs);
    const bool prof = g_prof.on && g_prof.used + 3 <= g_prof.ev.size() && (g_prof.seen++ % g_prof.every) == 0;
    static const bool no_graph = [] { const char *e = getenv("ARENA_NO_GRAPH"); return e && e[0] == '1'; }();
    bool replay = !prof && !no_graph && !s->graph_broken && st != cudaStreamLegacy && st != nullptr;
    if (replay) {
        cudaStreamCaptureStatus cap = cudaStreamCaptureStatusNone;
        if (cudaStreamIsCapturing(st, &cap) != cudaSuccess || cap != cudaStreamCaptureStatusNone) { cudaGetLastError(); replay = false; }
    }
    if (!replay) return launch_top_node_plain(s, d_total, d_alloc, d_cluster, d_pod_gpu, st, exchange, prof);
    arena_snapshot::StepGraph &g = s->step_graph;
    const bool same = g.exec && g.d_total == d_total && g.d_alloc == d_alloc && g.d_cluster == d_cluster && g.d_pod_gpu == d_pod_gpu &&
                      g.exchange == exchange && g.direct == s->direct && g.px == (exchange ? (void *)ctx().px : nullptr) &&
                      g.scratch == (void *)s->scratch && std::memcmp(&g.view, &s->v, sizeof(arena_snapshot_view)) == 0;
    if (!same) {
        if (g.exec) { cudaGraphExecDestroy(g.exec); g.exec = nullptr; }
        cudaGraph_t graph = nullptr;
        if (cudaStreamBeginCapture(st, cudaStreamCaptureModeThreadLocal) != cudaSuccess) {
            cudaGetLastError();
            s->graph_broken = true;
            return launch_top_node_plain(s, d_total, d_alloc, d_cluster, d_pod_gpu, st, exchange, false);
        }
        const uint64_t launches_before = g_launches.load();
        const int rc = launch_top_node_plain(s, d_total, d_alloc, d_cluster, d_pod_gpu, st, exchange, false);
        g_launches.store(launches_before);                      // nothing ran yet: the replay below counts
        const cudaError_t ce = cudaStreamEndCapture(st, &graph);
        if (rc || ce != cudaSuccess || !graph || cudaGraphInstantiate(&g.exec, graph, 0) != cudaSuccess) {
            if (graph) cudaGraphDestroy(graph);
            cudaGetLastError();
            g.exec = nullptr;
            s->graph_broken = true;
            return launch_top_node_plain(s, d_total, d_alloc, d_cluster, d_pod_gpu, st, exchange, false);
        }
        cudaGraphDestroy(graph);
        g.d_total = d_total; g.d_alloc = d_alloc; g.d_cluster = d_cluster; g.d_pod_gpu = d_pod_gpu;
        g.exchange = exchange; g.direct = s->direct; g.px = exchange ? (void *)ctx().px : nullptr;
        g.view = s->v; g.scratch = (void *)s->scratch;
    }
    CUDA_TRY(cudaGraphLaunch(g.exec, st));
    count_launch();
    return ARENA_OK;
}

int launch_top_node_plain(const arena_snapshot *s, i64 *d_total, i64 *d_alloc, i64 *d_cluster, i64 *d_pod_gpu,
                          cudaStream_t st, int exchange, bool prof)
{
    const arena_snapshot_view &v = s->v;
    cudaEvent_t *pe = prof ? &g_prof.ev[g_prof.used] : nullptr;
    if (prof) { g_prof.used += 3; CUDA_TRY(cudaEventRecord(pe[0], st)); }
    // node_alloc is accumulated with atomics: it starts from zero.  (A launch that clears it itself --
    // cooperative launch, every CTA zeroes a slice, one grid-wide barrier -- was measured and dropped:
    // the barrier costs what the memset costs, profiles/r02_iterations.md.)
    int gp = grid_for(v.n_pods, (u64)kPodWarps * kWarpPods, ctx().pods_ctas_per_sm);
    const int gn = grid_for(v.n_nodes, (u64)kPodThreads * 4, ctx().pods_ctas_per_sm);
    if (gn > gp) gp = gn;
    if (v.n_nodes) CUDA_TRY(cudaMemsetAsync(d_alloc, 0, v.n_nodes * 8, st));
    if (prof) CUDA_TRY(cudaEventRecord(pe[1], st));

    PodsArgs pa{};
    pa.P = v.n_pods; pa.C = v.n_containers; pa.K = v.n_nodes;
    pa.pod_key = v.pod_node;
    pa.pod_phase = v.pod_phase;
    pa.pod_ctr_begin = v.pod_ctr_begin;
    pa.ctr_val = (const i64 *)v.ctr_val;
    pa.ctr_scale = (const signed char *)v.ctr_scale;
    pa.key_alloc = (u64 *)d_alloc;
    pa.pod_gpu = d_pod_gpu;
    pa.total_alloc = d_cluster + 1;
    pa.scratch = s->scratch;
    pa.n_nodes = v.n_nodes;
    pa.cap_val = (const i64 *)v.node_cap_val;
    pa.cap_scale = (const signed char *)v.node_cap_scale;
    pa.node_total = d_total;
    pa.total_capacity = d_cluster;
    if (exchange) {
        pa.px = ctx().px;
        pa.px_lag = exchange == 2 ? 1u : 0u;
    }
    cudaError_t e;
    if (d_pod_gpu) {
        if (s->direct) e = launch_pods(top_node_pods_kernel<true, true>, gp, pa, st);
        else e = launch_pods(top_node_pods_kernel<false, true>, gp, pa, st);
    } else {
        if (s->direct) e = launch_pods(top_node_pods_kernel<true, false>, gp, pa, st);
        else if (ctx().staged_ctas_per_sm > 0) {
            // the pod columns staged one iteration ahead by the bulk-copy engine (kernels.cuh)
            int gs = grid_for(v.n_pods, (u64)kWarps * kWarpPods, ctx().staged_ctas_per_sm);
            if (gn > gs) gs = std::min(gn, ctx().sm_count * ctx().staged_ctas_per_sm);
            e = launch_pods(top_node_pods_staged_kernel, gs, pa, st, kThreads);
        } else {
            // a long launch (32 rounds and more: 3e7 pods on a B200) takes the 16-warp shape and the pool
            const u64 n_tiles = (v.n_pods + kWarpPods - 1) / kWarpPods;
            const u64 warps = (u64)ctx().sm_count * (u64)ctx().pods_long_ctas_per_sm * kPodWarpsLong;
            if (ctx().pods_long_ctas_per_sm > 0 && n_tiles / warps >= kLongLaunchRounds) {
                int gl = grid_for(v.n_pods, (u64)kPodWarpsLong * kWarpPods, ctx().pods_long_ctas_per_sm);
                e = launch_pods(top_node_pods_kernel<false, false, false, kPodWarpsLong>, gl, pa, st, kPodThreadsLong);
            } else e = launch_pods(top_node_pods_kernel<false, false>, gp, pa, st);
        }
    }
    if (e != cudaSuccess) return fail(ARENA_E_CUDA, "top-node kernel launch: %s", cudaGetErrorString(e));
    count_launch();
    if (prof) CUDA_TRY(cudaEventRecord(pe[2], st));
    return ARENA_OK;
}

int launch_top_job(arena_snapshot *s, i64 *d_req, i64 *d_alloc, unsigned *d_order, unsigned *d_latest,
                   i64 *d_totals, cudaStream_t st)
{
    const arena_snapshot_view &v = s->v;
    const u64 J = v.n_jobs;
    if (J == 0) {                                               // no job at all (an empty job range of a multi-GPU cut)
        CUDA_TRY(cudaMemsetAsync(d_totals, 0, 16, st));
        return ARENA_OK;
    }
    if (v.n_pods <= kSmallJobPods && J <= kSmallJobs) {         // a small cluster: one launch does it all
        static std::atomic<bool> smem_ok{false};            // (setting the attribute twice is harmless)
        if (!smem_ok.load(std::memory_order_acquire)) {
            CUDA_TRY(cudaFuncSetAttribute(job_small_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kSmallJobSmem));
            smem_ok.store(true, std::memory_order_release);
        }
        JobSmallArgs sa{};
        sa.P = v.n_pods; sa.J = J;
        sa.pod_job = v.pod_job; sa.pod_phase = v.pod_phase; sa.pod_role = v.pod_role;
        sa.pod_ctr_begin = v.pod_ctr_begin;
        sa.ctr_val = (const i64 *)v.ctr_val; sa.ctr_scale = (const signed char *)v.ctr_scale;
        sa.pod_created = (const i64 *)v.pod_created;
        sa.out_req = d_req; sa.out_alloc = d_alloc; sa.out_order = d_order; sa.out_latest = d_latest;
        sa.out_totals = d_totals;
        job_small_kernel<<<1, 1024, kSmallJobSmem, st>>>(sa);
        count_launch();
        CUDA_TRY(cudaGetLastError());
        return ARENA_OK;
    }
    int rc = ensure_job_arrays(s);
    if (rc) return rc;

    JobInitArgs ia{J, s->job_req, s->job_alloc, s->job_maxcreated, s->job_latest};
    job_init_kernel<<<grid_for(J, kThreads * 4, 8), kThreads, 0, st>>>(ia);
    count_launch();

    PodsArgs pa{};
    pa.P = v.n_pods; pa.C = v.n_containers; pa.K = J;
    pa.pod_key = v.pod_job;
    pa.pod_phase = v.pod_phase;
    pa.pod_ctr_begin = v.pod_ctr_begin;
    pa.ctr_val = (const i64 *)v.ctr_val;
    pa.ctr_scale = (const signed char *)v.ctr_scale;
    pa.pod_role = v.pod_role;
    pa.pod_created = (const i64 *)v.pod_created;
    pa.key_alloc = s->job_alloc;
    pa.key_req = s->job_req;
    pa.key_maxcreated = s->job_maxcreated;
    pa.total_req = d_totals;
    pa.total_alloc = d_totals + 1;
    pa.scratch = s->scratch;
    // the top-node kernel's 32-bit fast tiles, keyed by job (generic tiles where a tile does not qualify)
    if (launch_pods(top_node_pods_kernel<true, false, true>, grid_for(v.n_pods, (u64)kPodWarps * kWarpPods, ctx().pods_ctas_per_sm), pa, st) != cudaSuccess)
        return fail(ARENA_E_CUDA, "top-job pods kernel launch failed");
    count_launch();

    if (v.n_pods) {
        JobLatestArgs la{v.n_pods, J, v.pod_job, v.pod_role, (const i64 *)v.pod_created, s->job_maxcreated,
                         s->job_latest};
        job_latest_kernel<<<grid_for(v.n_pods, kThreads * 8, 8), kThreads, 0, st>>>(la);
        count_launch();
    }

    JobFinishArgs fa;
    fa.J = J; fa.Jpad = s->jpad;
    fa.job_latest = s->job_latest;
    fa.pod_phase = v.pod_phase;
    fa.pod_ctr_begin = v.pod_ctr_begin;
    fa.ctr_val = (const i64 *)v.ctr_val;
    fa.ctr_scale = (const signed char *)v.ctr_scale;
    fa.job_req = s->job_req; fa.job_alloc = s->job_alloc;
    fa.sort_key = s->sort_key; fa.sort_idx = s->sort_idx;
    fa.totals = d_totals;
    fa.scratch = s->scratch;
    // Which key bits vary is a property of the snapshot.  A snapshot the library uploaded cannot change,
    // so once its first query has published them, later queries radix-sort on exactly those bits; columns
    // the caller merely lent (arena_snapshot_wrap_device) may change under us and keep the bitonic network.
    const bool radix = s->owns && *((volatile unsigned *)&s->sort_host->ready) != 0u;
    fa.plan = (s->owns && !radix) ? s->sort_plan : nullptr;
    job_finish_kernel<<<grid_for(s->jpad, kThreads, 8), kThreads, 0, st>>>(fa);
    count_launch();

    if (radix) {
        u64 var = s->sort_host->varmask;
        const u64 *k_in = s->sort_key;
        const unsigned *i_in = s->sort_idx;
        u64 *k_out = s->alt_key;
        unsigned *i_out = s->alt_idx;
        const unsigned tiles = (unsigned)((J + kRadixTile - 1) / kRadixTile);
        while (var) {                                    // least significant eight varying bits first
            u64 mask = 0;
            for (int b = 0; b < 8 && var; b++) { mask |= var & (0 - var); var &= var - 1; }
            const int shift = __builtin_ctzll(mask);
            const int contiguous = ((mask >> shift) & ((mask >> shift) + 1)) == 0;
            radix_hist_kernel<<<tiles, 256, 0, st>>>(k_in, J, mask, shift, contiguous, s->block_hist);
            radix_scan_kernel<<<1, 1024, 0, st>>>(s->block_hist, tiles);
            radix_scatter_kernel<<<tiles, 256, 0, st>>>(k_in, i_in, J, mask, shift, contiguous, s->block_hist, k_out, i_out);
            g_launches.fetch_add(3, std::memory_order_relaxed);
            std::swap(const_cast<u64 *&>(k_in), k_out);
            std::swap(const_cast<unsigned *&>(i_in), i_out);
        }
        CUDA_TRY(cudaGetLastError());
        CUDA_TRY(cudaMemcpyAsync(d_req, s->job_req, J * 8, cudaMemcpyDeviceToDevice, st));
        CUDA_TRY(cudaMemcpyAsync(d_alloc, s->job_alloc, J * 8, cudaMemcpyDeviceToDevice, st));
        CUDA_TRY(cudaMemcpyAsync(d_order, i_in, J * 4, cudaMemcpyDeviceToDevice, st));
        if (d_latest) CUDA_TRY(cudaMemcpyAsync(d_latest, s->job_latest, J * 4, cudaMemcpyDeviceToDevice, st));
        return ARENA_OK;
    }

    // sort (requested desc, id asc)
    const u64 n = s->jpad;
    // large sorts: 4096-pair tiles (fewer passes through global memory); small ones: 2048 (more CTAs)
    const bool big = n >= (1ull << 18);
    const u64 tile = big ? kBigSortTile : kSortTile;
    const int tiles = (int)(n / tile);
    auto tile_pass = [&](u64 k, int full) {
        if (big) bitonic_tile_kernel<kBigSortTile><<<tiles, 1024, 0, st>>>(s->sort_key, s->sort_idx, n, k, full);
        else bitonic_tile_kernel<kSortTile><<<tiles, 1024, 0, st>>>(s->sort_key, s->sort_idx, n, k, full);
        count_launch();
    };
    tile_pass(0, 1);
    for (u64 k = 2 * tile; k <= n; k <<= 1) {
        for (u64 j = k >> 1; j >= tile; j >>= 1) {
            bitonic_global_kernel<<<grid_for(n / 2, kThreads, 8), kThreads, 0, st>>>(s->sort_key, s->sort_idx, n, k, j);
            count_launch();
        }
        tile_pass(k, 0);
    }
    if (fa.plan) {                                       // publish the varying bits for the next query on this handle
        SortHost *mapped = nullptr;
        CUDA_TRY(cudaHostGetDevicePointer((void **)&mapped, s->sort_host, 0));
        sort_plan_kernel<<<1, 1, 0, st>>>(s->sort_plan, mapped);
        count_launch();
    }
    CUDA_TRY(cudaGetLastError());
    if (J) {
        CUDA_TRY(cudaMemcpyAsync(d_req, s->job_req, J * 8, cudaMemcpyDeviceToDevice, st));
        CUDA_TRY(cudaMemcpyAsync(d_alloc, s->job_alloc, J * 8, cudaMemcpyDeviceToDevice, st));
        CUDA_TRY(cudaMemcpyAsync(d_order, s->sort_idx, J * 4, cudaMemcpyDeviceToDevice, st));
        if (d_latest) CUDA_TRY(cudaMemcpyAsync(d_latest, s->job_latest, J * 4, cudaMemcpyDeviceToDevice, st));
    }
    return ARENA_OK;
}

}  // namespace

extern "C" {

// Copy host columns into s->block (grown when too small; one allocation, every
// column padded to 16 bytes plus a spare vector so aligned 128-bit reads stay
// inside it) and point s->v at the device copies.  Asynchronous on the
// library stream.
// Device block of `s` sized for (P, C, N) wide columns; with `h` the host columns are copied in
// (async, library stream), without it they are left for the caller to fill.
static int fill_snapshot(arena_snapshot *s, const arena_snapshot_view *h, bool jobs, const u64 *pcn = nullptr)
{
    const u64 P = h ? h->n_pods : pcn[0], C = h ? h->n_containers : pcn[1], N = h ? h->n_nodes : pcn[2];
    size_t sizes[10] = {P * 4, P, (P + 1) * 4, C * 8, C, N * 8, N, jobs ? P * 4 : 0, jobs ? P : 0, jobs ? P * 8 : 0};
    const void *src[10] = {nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr};
    if (h) {
        const void *from[10] = {h->pod_node, h->pod_phase, h->pod_ctr_begin, h->ctr_val, h->ctr_scale,
                                h->node_cap_val, h->node_cap_scale, h->pod_job, h->pod_role, h->pod_created};
        for (int i = 0; i < 10; i++) src[i] = from[i];
    }
    size_t total = 0;
    for (size_t z : sizes) total += pad16(z) + 16;
    if (total < 16) total = 16;
    if (s->block_bytes < total) {
        if (s->block) arena_internal::device_release(s->block, s->block_bytes);
        s->block = nullptr;
        s->block_bytes = 0;
        cudaError_t e = arena_internal::device_alloc(&s->block, total);
        if (e != cudaSuccess)
            return fail(e == cudaErrorMemoryAllocation ? ARENA_E_NOMEM : ARENA_E_CUDA, "cudaMalloc(%zu): %s", total,
                        cudaGetErrorString(e));
        s->block_bytes = total;
    }
    s->owns = true;
    char *p = (char *)s->block;
    const void *dst[10];
    for (int i = 0; i < 10; i++) {
        dst[i] = p;
        if (sizes[i] && src[i]) {
            const int crc = arena_internal::copy_to_device(p, src[i], sizes[i], ctx().stream);
            if (crc) return crc;
        }
        p += pad16(sizes[i]) + 16;
    }
    if (h) s->v = *h;
    else { s->v = arena_snapshot_view{}; s->v.n_pods = P; s->v.n_containers = C; s->v.n_nodes = N; }
    s->v.pod_node = (const uint32_t *)dst[0];
    s->v.pod_phase = (const uint8_t *)dst[1];
    s->v.pod_ctr_begin = (const uint32_t *)dst[2];
    s->v.ctr_val = (const int64_t *)dst[3];
    s->v.ctr_scale = (const int8_t *)dst[4];
    s->v.node_cap_val = (const int64_t *)dst[5];
    s->v.node_cap_scale = (const int8_t *)dst[6];
    s->v.pod_job = jobs ? (const uint32_t *)dst[7] : nullptr;
    s->v.pod_role = jobs ? (const uint8_t *)dst[8] : nullptr;
    s->v.pod_created = jobs ? (const int64_t *)dst[9] : nullptr;
    if (!jobs) s->v.n_jobs = 0;
    return ARENA_OK;
}

int arena_snapshot_upload(const arena_snapshot_view *h, arena_snapshot **out)
{
    int rc = check_ready();
    if (rc) return rc;
    if (!out) return fail(ARENA_E_INVALID, "out is NULL");
    *out = nullptr;
    rc = check_view(h, false);
    if (rc) return rc;
    arena_snapshot *s = new (std::nothrow) arena_snapshot();
    if (!s) return fail(ARENA_E_NOMEM, "out of host memory");
    rc = fill_snapshot(s, h, h->n_jobs > 0 && h->pod_job);
    if (!rc) rc = alloc_common(s);
    if (!rc) rc = probe_grouping(s);
    if (rc) { arena_snapshot_free(s); return rc; }
    cudaError_t e = cudaStreamSynchronize(ctx().stream);     // host buffers may be reused on return
    if (e != cudaSuccess) { arena_snapshot_free(s); return fail(ARENA_E_CUDA, "upload: %s", cudaGetErrorString(e)); }
    *out = s;
    return ARENA_OK;
}

int arena_snapshot_wrap_device(const arena_snapshot_view *d, arena_snapshot **out)
{
    int rc = check_ready();
    if (rc) return rc;
    if (!out) return fail(ARENA_E_INVALID, "out is NULL");
    *out = nullptr;
    rc = check_view(d, true);
    if (rc) return rc;
    arena_snapshot *s = new (std::nothrow) arena_snapshot();
    if (!s) return fail(ARENA_E_NOMEM, "out of host memory");
    s->v = *d;
    if (!d->pod_job) s->v.n_jobs = 0;
    rc = alloc_common(s);
    if (!rc) rc = probe_grouping(s);
    if (rc) { arena_snapshot_free(s); return rc; }
    cudaError_t e = cudaStreamSynchronize(ctx().stream);
    if (e != cudaSuccess) { arena_snapshot_free(s); return fail(ARENA_E_CUDA, "wrap: %s", cudaGetErrorString(e)); }
    *out = s;
    return ARENA_OK;
}

}  // extern "C"

namespace arena_internal {

int current_slot() { return t_slot; }
void set_slot(int slot) { t_slot = slot; }

int init_slot(int device)
{
    if (t_slot < 0 || t_slot >= kMaxSlots) return fail(ARENA_E_INVALID, "at most %d devices per process", kMaxSlots);
    std::lock_guard<std::mutex> lk(g_mu);
    return init_current_slot(device);
}

// Mailboxes of slots 0..n-1 wired to one another through peer access: a device pointer of slot j is
// valid on slot i once cudaDeviceEnablePeerAccess(j) has been called with i current (one address
// space per process).  Returns ARENA_E_INVALID (and leaves nothing behind) when some pair of devices
// cannot reach each other; the caller then adds the per-device totals on the host.
int peer_exchange_in_process(int n)
{
    if (n < 1 || n > kMaxPeers || n > kMaxSlots) return fail(ARENA_E_INVALID, "%d devices (at most %d)", n, kMaxPeers);
    for (int i = 0; i < n; i++) {
        if (!g_ctxs[i].ready) return fail(ARENA_E_NOT_INIT, "slot %d is not initialised", i);
        if (g_ctxs[i].my_box) return fail(ARENA_E_INVALID, "peer exchange already created");
    }
    for (int i = 0; i < n; i++)
        for (int j = 0; j < n; j++) {
            if (i == j || g_ctxs[i].device == g_ctxs[j].device) continue;
            int can = 0;
            CUDA_TRY(cudaDeviceCanAccessPeer(&can, g_ctxs[i].device, g_ctxs[j].device));
            if (!can) return fail(ARENA_E_INVALID, "device %d cannot access device %d", g_ctxs[i].device, g_ctxs[j].device);
        }
    for (int i = 0; i < n; i++) {
        CUDA_TRY(cudaSetDevice(g_ctxs[i].device));
        for (int j = 0; j < n; j++) {
            if (i == j || g_ctxs[i].device == g_ctxs[j].device) continue;
            cudaError_t e = cudaDeviceEnablePeerAccess(g_ctxs[j].device, 0);
            if (e == cudaErrorPeerAccessAlreadyEnabled) cudaGetLastError();
            else if (e != cudaSuccess) return fail(ARENA_E_CUDA, "cudaDeviceEnablePeerAccess: %s", cudaGetErrorString(e));
        }
        CUDA_TRY(cudaMalloc((void **)&g_ctxs[i].my_box, sizeof(PeerMailbox)));
        CUDA_TRY(cudaMemset(g_ctxs[i].my_box, 0, sizeof(PeerMailbox)));
    }
    for (int i = 0; i < n; i++) {
        CUDA_TRY(cudaSetDevice(g_ctxs[i].device));
        g_ctxs[i].px_host = PeerExchange{};
        g_ctxs[i].px_host.rank = i;
        g_ctxs[i].px_host.world = n;
        for (int j = 0; j < n; j++) g_ctxs[i].px_host.box[j] = g_ctxs[j].my_box;
        CUDA_TRY(cudaMalloc((void **)&g_ctxs[i].px, sizeof(PeerExchange)));
        CUDA_TRY(cudaMemcpy(g_ctxs[i].px, &g_ctxs[i].px_host, sizeof(PeerExchange), cudaMemcpyHostToDevice));
    }
    return ARENA_OK;
}

cudaError_t device_alloc(void **out, size_t bytes)
{
    if (bytes < 256) bytes = 256;
    {
        std::lock_guard<std::mutex> lk(g_pool_mu);
        int best = -1;
        for (int i = 0; i < 8; i++)
            if (g_free_blocks[i].p && g_free_blocks[i].device == ctx().device && g_free_blocks[i].bytes >= bytes &&
                g_free_blocks[i].bytes <= 2 * bytes + (1u << 20) &&
                (best < 0 || g_free_blocks[i].bytes < g_free_blocks[best].bytes))
                best = i;
        if (best >= 0) {
            *out = g_free_blocks[best].p;
            // whatever the library stream still had to do with the block comes first
            if (g_free_blocks[best].idle) {
                cudaStreamWaitEvent(ctx().stream, g_free_blocks[best].idle, 0);
                cudaEventDestroy(g_free_blocks[best].idle);
            }
            g_free_blocks[best] = FreeBlock{nullptr, 0, nullptr, -1};
            return cudaSuccess;
        }
    }
    cudaError_t e = cudaMalloc(out, bytes);
    if (e == cudaErrorMemoryAllocation) {          // give the cache back and try once more
        device_pool_clear();
        cudaGetLastError();
        e = cudaMalloc(out, bytes);
    }
    return e;
}

// `bytes` = what the block was allocated with (device_alloc's argument).  Stream-ordered: an event
// on the library stream marks the point after which the block is idle, and whoever takes the block
// out of the cache makes the library stream wait for it -- no other stream of the process is stalled.
// (Work a caller enqueued on a stream of its own must be complete before it frees the handle:
// arena_snapshot_free says so.)  A block that is really freed goes through cudaFree, which waits.
void device_release(void *p, size_t bytes)
{
    if (!p) return;
    if (bytes < 256) bytes = 256;
    cudaEvent_t idle = nullptr;
    if (ctx().ready && cudaEventCreateWithFlags(&idle, cudaEventDisableTiming) == cudaSuccess) {
        if (cudaEventRecord(idle, ctx().stream) != cudaSuccess) { cudaEventDestroy(idle); idle = nullptr; }
    }
    if (!idle) cudaDeviceSynchronize();            // no event to order by: fall back to waiting here
    std::lock_guard<std::mutex> lk(g_pool_mu);
    int slot = -1;
    for (int i = 0; i < 8; i++)
        if (!g_free_blocks[i].p) { slot = i; break; }
    if (slot < 0) {                                 // full: drop the smallest
        slot = 0;
        for (int i = 1; i < 8; i++)
            if (g_free_blocks[i].bytes < g_free_blocks[slot].bytes) slot = i;
        if (g_free_blocks[slot].bytes >= bytes) { if (idle) cudaEventDestroy(idle); cudaFree(p); return; }
        if (g_free_blocks[slot].idle) cudaEventDestroy(g_free_blocks[slot].idle);
        cudaFree(g_free_blocks[slot].p);
    }
    g_free_blocks[slot] = FreeBlock{p, bytes, idle, ctx().device};
}

void device_pool_clear()
{
    std::lock_guard<std::mutex> lk(g_pool_mu);
    for (FreeBlock &b : g_free_blocks) {
        if (b.idle) cudaEventDestroy(b.idle);
        if (b.p) cudaFree(b.p);
        b = FreeBlock{nullptr, 0, nullptr, -1};
    }
}

int ready(cudaStream_t *stream, int *sm_count)
{
    int rc = check_ready();
    if (rc) return rc;
    if (stream) *stream = ctx().stream;
    if (sm_count) *sm_count = ctx().sm_count;
    return ARENA_OK;
}

void add_launches(unsigned long long n) { g_launches.fetch_add(n, std::memory_order_relaxed); }

int adopt_device_block(void *block, size_t block_bytes, const arena_snapshot_view &device_view, arena_snapshot **out)
{
    *out = nullptr;
    int rc = check_ready();
    if (!rc) rc = check_view(&device_view, true);
    arena_snapshot *s = rc ? nullptr : new (std::nothrow) arena_snapshot();
    if (!rc && !s) rc = fail(ARENA_E_NOMEM, "out of host memory");
    if (rc) { device_release(block, block_bytes); return rc; }
    s->v = device_view;
    s->v.n_jobs = 0;
    s->owns = true;
    s->block = block;
    s->block_bytes = block_bytes;
    rc = alloc_common(s);
    if (!rc) rc = probe_grouping(s);
    if (rc) { arena_snapshot_free(s); return rc; }
    *out = s;
    return ARENA_OK;
}

}  // namespace arena_internal

extern "C" {

int arena_snapshot_counts(const arena_snapshot *s, uint64_t counts_out[4])
{
    if (!s || !counts_out) return fail(ARENA_E_INVALID, "snapshot or counts_out is NULL");
    counts_out[0] = s->v.n_pods; counts_out[1] = s->v.n_containers;
    counts_out[2] = s->v.n_nodes; counts_out[3] = s->v.n_jobs;
    return ARENA_OK;
}

int arena_snapshot_download(const arena_snapshot *s, const arena_snapshot_view *o)
{
    arena_internal::UseSlot use(s ? s->slot : arena_internal::current_slot());
    int rc = check_ready();
    if (rc) return rc;
    if (!s || !o) return fail(ARENA_E_INVALID, "snapshot or host_out is NULL");
    const arena_snapshot_view &v = s->v;
    const u64 P = v.n_pods, C = v.n_containers, N = v.n_nodes;
    struct Col { const void *src; const void *dst; size_t bytes; };
    const Col cols[] = {
        {v.pod_node, o->pod_node, P * 4}, {v.pod_phase, o->pod_phase, P}, {v.pod_ctr_begin, o->pod_ctr_begin, (P + 1) * 4},
        {v.ctr_val, o->ctr_val, C * 8}, {v.ctr_scale, o->ctr_scale, C}, {v.node_cap_val, o->node_cap_val, N * 8},
        {v.node_cap_scale, o->node_cap_scale, N}, {v.pod_job, o->pod_job, P * 4}, {v.pod_role, o->pod_role, P},
        {v.pod_created, o->pod_created, P * 8}};
    for (const Col &c : cols)
        if (c.src && c.dst && c.bytes)
            CUDA_TRY(cudaMemcpyAsync(const_cast<void *>(c.dst), c.src, c.bytes, cudaMemcpyDeviceToHost, ctx().stream));
    CUDA_TRY(cudaStreamSynchronize(ctx().stream));
    return ARENA_OK;
}

void arena_snapshot_free(arena_snapshot *s)
{
    if (!s) return;
    arena_internal::UseSlot use(s->slot);
    if (ctx().ready) cudaSetDevice(ctx().device);
    if (s->owns && s->block) arena_internal::device_release(s->block, s->block_bytes);
    if (s->scratch) arena_internal::device_release(s->scratch, sizeof(Scratch));
    if (s->step_graph.exec) cudaGraphExecDestroy(s->step_graph.exec);
    if (s->job_req) cudaFree(s->job_req);
    if (s->sort_host) cudaFreeHost(s->sort_host);
    if (s->d_out) cudaFree(s->d_out);
    delete s;
}

int arena_top_node_device(const arena_snapshot *s, int64_t *d_node_total, int64_t *d_node_alloc,
                          int64_t *d_cluster, int64_t *d_pod_gpu, void *stream)
{
    arena_internal::UseSlot use(s ? s->slot : arena_internal::current_slot());
    int rc = check_ready();
    if (rc) return rc;
    if (!s || !d_cluster) return fail(ARENA_E_INVALID, "snapshot or d_cluster is NULL");
    if (s->v.n_nodes && (!d_node_total || !d_node_alloc)) return fail(ARENA_E_INVALID, "node output buffers are NULL");
    if (((uintptr_t)d_pod_gpu & 15) != 0) return fail(ARENA_E_INVALID, "d_pod_gpu is not 16-byte aligned");
    cudaStream_t st = stream ? (cudaStream_t)stream : ctx().stream;
    return launch_top_node(s, (i64 *)d_node_total, (i64 *)d_node_alloc, (i64 *)d_cluster, (i64 *)d_pod_gpu, st);
}

int arena_top_node(const arena_snapshot *cs, int64_t *node_total, int64_t *node_alloc, int64_t cluster[2],
                   int64_t *pod_gpu)
{
    arena_internal::UseSlot use(cs ? cs->slot : arena_internal::current_slot());
    int rc = check_ready();
    if (rc) return rc;
    if (!cs || !cluster) return fail(ARENA_E_INVALID, "snapshot or cluster is NULL");
    arena_snapshot *s = const_cast<arena_snapshot *>(cs);
    const u64 N = s->v.n_nodes, P = s->v.n_pods;
    if (N && (!node_total || !node_alloc)) return fail(ARENA_E_INVALID, "node output buffers are NULL");
    size_t bytes = 16 + pad16(N * 8) * 2 + (pod_gpu ? pad16(P * 8) : 0);
    rc = ensure_out(s, bytes);
    if (rc) return rc;
    char *base = (char *)s->d_out;
    i64 *d_cluster = (i64 *)base;
    i64 *d_total = (i64 *)(base + 16);
    i64 *d_alloc = (i64 *)(base + 16 + pad16(N * 8));
    i64 *d_pg = pod_gpu ? (i64 *)(base + 16 + 2 * pad16(N * 8)) : nullptr;
    cudaStream_t st = ctx().stream;
    rc = launch_top_node(s, d_total, d_alloc, d_cluster, d_pg, st);
    if (rc) return rc;
    CUDA_TRY(cudaMemcpyAsync(cluster, d_cluster, 16, cudaMemcpyDeviceToHost, st));
    if (N) {
        CUDA_TRY(cudaMemcpyAsync(node_total, d_total, N * 8, cudaMemcpyDeviceToHost, st));
        CUDA_TRY(cudaMemcpyAsync(node_alloc, d_alloc, N * 8, cudaMemcpyDeviceToHost, st));
    }
    if (pod_gpu && P) CUDA_TRY(cudaMemcpyAsync(pod_gpu, d_pg, P * 8, cudaMemcpyDeviceToHost, st));
    CUDA_TRY(cudaStreamSynchronize(st));
    return ARENA_OK;
}

int arena_top_node_host(const arena_snapshot_view *host, int64_t *node_total, int64_t *node_alloc,
                        int64_t cluster[2], int64_t *pod_gpu)
{
    int rc = check_ready();
    if (rc) return rc;
    rc = check_view(host, false);
    if (rc) return rc;
    // One staging snapshot per process, kept between calls: a polling caller
    // pays for cudaMalloc only when its snapshot grows.  Calls are serialised.
    std::lock_guard<std::mutex> lk(g_host_mu);
    if (!ctx().host_snap) {
        ctx().host_snap = new (std::nothrow) arena_snapshot();
        if (!ctx().host_snap) return fail(ARENA_E_NOMEM, "out of host memory");
        rc = alloc_common(ctx().host_snap);
        if (rc) { arena_snapshot_free(ctx().host_snap); ctx().host_snap = nullptr; return rc; }
    }
    rc = fill_snapshot(ctx().host_snap, host, false);            // async H2D on the library stream ...
    if (rc) return rc;
    {   // grouping probe on the HOST column (no extra device round trip on this path)
        const u64 P = host->n_pods;
        unsigned diff = 0, n = 0;
        if (P >= 2) {
            n = P - 1 < 4096 ? (unsigned)(P - 1) : 4096u;
            for (unsigned i = 0; i < n; i++) {
                const u64 p = (u64)i * (P - 1) / n;
                diff += host->pod_node[p] != host->pod_node[p + 1];
            }
        }
        ctx().host_snap->direct = diff * 4u > n;
    }
    return arena_top_node(ctx().host_snap, node_total, node_alloc, cluster, pod_gpu);   // ... kernels + D2H behind it
}

// ---- the packed columns on the device ------------------------------------------------------------
// One block: [node_pod_begin | node_ctr_begin | pod_meta | ctr_gpu | pieces | cap_val | cap_scale |
// node_total | node_alloc | cluster pairs], every column padded so that the kernel's whole-word reads
// past a column's end stay inside the block.
constexpr int kPackedChunks = 8;                 // pipeline depth of the one-shot call
struct PackedLayout {
    u64 P, C, N, K, gpu_bytes;
    unsigned bits;
    size_t o_begin, o_cbegin, o_meta, o_gpu, o_pieces, o_capv, o_caps, o_total, o_alloc, o_cluster, need;
};
struct arena_packed_device {     // device-resident packed snapshot (arena_snapshot_upload_packed)
    PackedLayout L{};
    void *block = nullptr;
    Scratch *scratch = nullptr;
    int slot = t_slot;
};
static PackedLayout packed_layout(const arena_packed_view *h)
{
    PackedLayout L{};
    L.P = h->n_pods; L.C = h->n_containers; L.N = h->n_nodes; L.K = h->n_pieces; L.bits = h->ctr_bits;
    L.gpu_bytes = h->ctr_bits == 4 ? (L.C + 1) / 2 : L.C;
    L.o_begin = 0;
    L.o_cbegin = L.o_begin + pad16((L.N + 1) * 4);
    L.o_meta = L.o_cbegin + pad16((L.N + 1) * 4);
    L.o_gpu = L.o_meta + pad16(L.P + 16);
    L.o_pieces = L.o_gpu + pad16(L.gpu_bytes + 16);
    L.o_capv = L.o_pieces + pad16(L.K * 16);
    L.o_caps = L.o_capv + pad16(L.N * 8);
    L.o_total = L.o_caps + pad16(L.N);
    L.o_alloc = L.o_total + pad16(L.N * 8);
    L.o_cluster = L.o_alloc + pad16(L.N * 8);
    L.need = L.o_cluster + 16 * kPackedChunks;
    return L;
}
static PackedArgs packed_args(const PackedLayout &L, char *pb, Scratch *scratch)
{
    PackedArgs a{};
    a.N = L.N; a.n_pieces = L.K;
    a.node_pod_begin = (const unsigned *)(pb + L.o_begin);
    a.node_ctr_begin = (const unsigned *)(pb + L.o_cbegin);
    a.meta = (const unsigned char *)(pb + L.o_meta);
    a.gpu = (const unsigned char *)(pb + L.o_gpu);
    a.pieces = (const uint4 *)(pb + L.o_pieces);
    a.cap_val = (const i64 *)(pb + L.o_capv);
    a.cap_scale = (const signed char *)(pb + L.o_caps);
    a.node_total = (i64 *)(pb + L.o_total);
    a.node_alloc = (u64 *)(pb + L.o_alloc);
    a.total_capacity = (i64 *)(pb + L.o_cluster);
    a.total_alloc = (i64 *)(pb + L.o_cluster) + 1;
    a.scratch = scratch;
    return a;
}
static int packed_launch(const PackedArgs &a, unsigned bits, cudaStream_t st)
{
    const int grid = grid_for((a.N - a.n_lo) + (a.n_pieces - a.piece_lo), kWarps * 2, ctx().packed_ctas_per_sm);
    if (bits == 4) top_node_packed_kernel<4><<<grid, kThreads, 0, st>>>(a);
    else top_node_packed_kernel<8><<<grid, kThreads, 0, st>>>(a);
    count_launch();
    CUDA_TRY(cudaGetLastError());
    return ARENA_OK;
}
static int check_packed(const arena_packed_view *h)
{
    if (!h) return fail(ARENA_E_INVALID, "packed view is NULL");
    const u64 P = h->n_pods, C = h->n_containers, N = h->n_nodes;
    // (the kernel counts nodes and pieces in 32 bits and steps past the end by one grid stride)
    if (P >= 0xFFFFFFFFull || C > 0xFFFFFFFFull || N >= 0xFFF00000ull || h->n_pieces >= 0xFFF00000ull)
        return fail(ARENA_E_TOO_LARGE, "snapshot exceeds 32-bit offsets");
    if (!h->node_pod_begin || (P && !h->pod_meta) || (C && !h->ctr_gpu) || (N && (!h->node_cap_val || !h->node_cap_scale)))
        return fail(ARENA_E_INVALID, "packed columns are NULL");
    if (h->ctr_bits != 4 && h->ctr_bits != 8) return fail(ARENA_E_INVALID, "ctr_bits must be 4 or 8, not %u", h->ctr_bits);
    if (!h->node_ctr_begin || (h->n_pieces && !h->pieces)) return fail(ARENA_E_INVALID, "node_ctr_begin / pieces are NULL");
    return ARENA_OK;
}

// Summary straight from the packed columns, pipelined: the node range is cut into kPackedChunks
// pieces of about equal pod count; chunk k's columns go up on the copy stream while chunk k-1 is
// summed on the compute stream and chunk k-2's slice of the table comes down on the download
// stream (PCIe is full duplex), so the kernel and the download hide behind the upload.  Every
// chunk leaves its own pair of totals; the host adds them (wrapping, like Go's +=).
static int top_node_packed_native(const arena_packed_view *h, int64_t *node_total, int64_t *node_alloc, int64_t cluster[2])
{
    const PackedLayout L = packed_layout(h);
    const u64 P = L.P, N = L.N, K = L.K;
    if (ctx().packed_bytes < L.need) {
        if (ctx().packed_block) cudaFree(ctx().packed_block);
        ctx().packed_block = nullptr;
        ctx().packed_bytes = 0;
        CUDA_TRY(cudaMalloc(&ctx().packed_block, L.need));
        ctx().packed_bytes = L.need;
    }
    char *pb = (char *)ctx().packed_block;
    if (!ctx().pipe_ready) {
        CUDA_TRY(cudaStreamCreateWithFlags(&ctx().pipe_stream[0], cudaStreamNonBlocking));
        CUDA_TRY(cudaStreamCreateWithFlags(&ctx().pipe_stream[1], cudaStreamNonBlocking));
        for (int k = 0; k < 2 * kPackedChunks + 1; k++) CUDA_TRY(cudaEventCreateWithFlags(&ctx().pipe_event[k], cudaEventDisableTiming));
        ctx().pipe_ready = true;
    }
    cudaStream_t up = ctx().stream, run = ctx().pipe_stream[0], down = ctx().pipe_stream[1];
    // chunk boundaries in nodes: equal shares of the rows (a node is never cut)
    // about 8 Mi rows per chunk, eight chunks at most: below that the per-copy overheads of a chunk
    // (five uploads, a launch, three downloads) outweigh what the overlap hides
    u64 chunk_rows = 8u << 20;
    if (const char *env = getenv("ARENA_PACKED_CHUNK_PODS")) {           // tuning aid (and the tests' way to pipeline small inputs)
        const long long v = atoll(env);
        if (v > 0) chunk_rows = (u64)v;
    }
    int chunks = (int)std::min<u64>((u64)kPackedChunks, P / chunk_rows);
    if (chunks < 1 || N < (u64)kPackedChunks) chunks = 1;
    u64 cut[kPackedChunks + 1];
    cut[0] = 0;
    for (int k = 1; k < chunks; k++) {
        const uint32_t target = (uint32_t)(P * (u64)k / (u64)chunks);
        u64 n = std::upper_bound(h->node_pod_begin, h->node_pod_begin + N + 1, target) - h->node_pod_begin;
        n = n ? n - 1 : 0;
        cut[k] = std::max(cut[k - 1], std::min(n, N));
    }
    cut[chunks] = N;
    const uint4 *pieces = (const uint4 *)h->pieces;
    int64_t part[kPackedChunks][2];
    int crc;
    if (K) CUDA_TRY(cudaMemcpyAsync(pb + L.o_pieces, h->pieces, K * 16, cudaMemcpyHostToDevice, up));
    CUDA_TRY(cudaEventRecord(ctx().pipe_event[2 * kPackedChunks], up));      // what earlier calls left on the library stream
    CUDA_TRY(cudaStreamWaitEvent(run, ctx().pipe_event[2 * kPackedChunks], 0));
    u64 piece_lo = 0;
    for (int k = 0; k < chunks; k++) {
        const u64 n0 = cut[k], n1 = cut[k + 1];
        // rows and containers of the chunk; the rows that name no node (after node N - 1) are never read
        const u64 r0 = h->node_pod_begin[n0], r1 = h->node_pod_begin[n1];
        const u64 c0 = h->node_ctr_begin[n0], c1 = h->node_ctr_begin[n1];
        const u64 g0 = L.bits == 4 ? c0 / 2 : c0, g1 = L.bits == 4 ? (c1 + 1) / 2 : c1;
        // (one entry past the chunk: the kernel reads begin[n + 1])
        CUDA_TRY(cudaMemcpyAsync(pb + L.o_begin + n0 * 4, h->node_pod_begin + n0, (n1 - n0 + 1) * 4, cudaMemcpyHostToDevice, up));
        CUDA_TRY(cudaMemcpyAsync(pb + L.o_cbegin + n0 * 4, h->node_ctr_begin + n0, (n1 - n0 + 1) * 4, cudaMemcpyHostToDevice, up));
        if (r1 > r0 && (crc = arena_internal::copy_to_device(pb + L.o_meta + r0, h->pod_meta + r0, r1 - r0, up))) return crc;
        if (g1 > g0 && (crc = arena_internal::copy_to_device(pb + L.o_gpu + g0, h->ctr_gpu + g0, g1 - g0, up))) return crc;
        if (n1 > n0) {
            CUDA_TRY(cudaMemcpyAsync(pb + L.o_capv + n0 * 8, h->node_cap_val + n0, (n1 - n0) * 8, cudaMemcpyHostToDevice, up));
            CUDA_TRY(cudaMemcpyAsync(pb + L.o_caps + n0, h->node_cap_scale + n0, n1 - n0, cudaMemcpyHostToDevice, up));
        }
        CUDA_TRY(cudaEventRecord(ctx().pipe_event[2 * k], up));
        // compute
        CUDA_TRY(cudaStreamWaitEvent(run, ctx().pipe_event[2 * k], 0));
        u64 piece_hi = piece_lo;
        while (piece_hi < K && pieces[piece_hi].x < n1) piece_hi++;
        if (piece_hi > piece_lo) CUDA_TRY(cudaMemsetAsync(pb + L.o_alloc + n0 * 8, 0, (n1 - n0) * 8, run));   // big nodes are summed with atomics
        PackedArgs a = packed_args(L, pb, ctx().host_snap->scratch);
        a.n_lo = n0; a.N = n1; a.piece_lo = piece_lo; a.n_pieces = piece_hi;
        a.total_capacity = (i64 *)(pb + L.o_cluster) + 2 * k;
        a.total_alloc = a.total_capacity + 1;
        if ((crc = packed_launch(a, L.bits, run))) return crc;
        piece_lo = piece_hi;
        CUDA_TRY(cudaEventRecord(ctx().pipe_event[2 * k + 1], run));
    }
    // downloads, issued after every upload and launch is queued: a copy into pageable memory blocks the host
    for (int k = 0; k < chunks; k++) {
        const u64 n0 = cut[k], n1 = cut[k + 1];
        CUDA_TRY(cudaStreamWaitEvent(down, ctx().pipe_event[2 * k + 1], 0));
        CUDA_TRY(cudaMemcpyAsync(part[k], pb + L.o_cluster + 16 * k, 16, cudaMemcpyDeviceToHost, down));
        if (n1 > n0) {
            CUDA_TRY(cudaMemcpyAsync(node_total + n0, pb + L.o_total + n0 * 8, (n1 - n0) * 8, cudaMemcpyDeviceToHost, down));
            CUDA_TRY(cudaMemcpyAsync(node_alloc + n0, pb + L.o_alloc + n0 * 8, (n1 - n0) * 8, cudaMemcpyDeviceToHost, down));
        }
    }
    CUDA_TRY(cudaStreamSynchronize(down));
    CUDA_TRY(cudaStreamSynchronize(up));
    u64 tc = 0, ta = 0;
    for (int k = 0; k < chunks; k++) { tc += (u64)part[k][0]; ta += (u64)part[k][1]; }
    cluster[0] = (int64_t)tc;
    cluster[1] = (int64_t)ta;
    return ARENA_OK;
}

int arena_snapshot_upload_packed(const arena_packed_view *h, arena_packed_device **out)
{
    int rc = check_ready();
    if (rc) return rc;
    if (!out) return fail(ARENA_E_INVALID, "out is NULL");
    *out = nullptr;
    rc = check_packed(h);
    if (rc) return rc;
    arena_packed_device *p = new (std::nothrow) arena_packed_device();
    if (!p) return fail(ARENA_E_NOMEM, "out of host memory");
    p->L = packed_layout(h);
    const PackedLayout &L = p->L;
    cudaError_t e = arena_internal::device_alloc(&p->block, L.need);
    if (e == cudaSuccess) e = arena_internal::device_alloc((void **)&p->scratch, sizeof(Scratch));
    if (e != cudaSuccess) {
        arena_packed_device_free(p);
        return fail(e == cudaErrorMemoryAllocation ? ARENA_E_NOMEM : ARENA_E_CUDA, "cudaMalloc(%zu): %s", L.need, cudaGetErrorString(e));
    }
    cudaStream_t st = ctx().stream;
    char *pb = (char *)p->block;
    rc = ARENA_OK;
    auto up = [&](size_t off, const void *src, size_t bytes) {
        if (!rc && bytes) rc = arena_internal::copy_to_device(pb + off, src, bytes, st);
    };
    if (cudaMemsetAsync(p->scratch, 0, sizeof(Scratch), st) != cudaSuccess) rc = fail(ARENA_E_CUDA, "cudaMemsetAsync failed");
    up(L.o_begin, h->node_pod_begin, (L.N + 1) * 4);
    up(L.o_cbegin, h->node_ctr_begin, (L.N + 1) * 4);
    up(L.o_meta, h->pod_meta, L.P);
    up(L.o_gpu, h->ctr_gpu, L.gpu_bytes);
    up(L.o_pieces, h->pieces, L.K * 16);
    up(L.o_capv, h->node_cap_val, L.N * 8);
    up(L.o_caps, h->node_cap_scale, L.N);
    if (!rc && cudaStreamSynchronize(st) != cudaSuccess) rc = fail(ARENA_E_CUDA, "packed upload failed");
    if (rc) { arena_packed_device_free(p); return rc; }
    *out = p;
    return ARENA_OK;
}

void arena_packed_device_free(arena_packed_device *p)
{
    if (!p) return;
    arena_internal::UseSlot use(p->slot);
    if (ctx().ready) cudaSetDevice(ctx().device);
    if (p->block) arena_internal::device_release(p->block, p->L.need);
    if (p->scratch) arena_internal::device_release(p->scratch, sizeof(Scratch));
    delete p;
}

int arena_packed_device_bytes(const arena_packed_device *p, uint64_t *read_write_bytes)
{
    if (!p || !read_write_bytes) return fail(ARENA_E_INVALID, "NULL argument");
    const PackedLayout &L = p->L;
    // what one summary launch reads and writes: both offset columns, one byte per pod, the value
    // column, the pieces, the capacities; the two table columns and the totals out
    *read_write_bytes = 8 * (L.N + 1) + L.P + L.gpu_bytes + 16 * L.K + 9 * L.N + 16 * L.N + 16;
    return ARENA_OK;
}

int arena_top_node_packed_device(const arena_packed_device *p, int64_t *d_node_total, int64_t *d_node_alloc,
                                 int64_t *d_cluster, void *stream)
{
    arena_internal::UseSlot use(p ? p->slot : arena_internal::current_slot());
    int rc = check_ready();
    if (rc) return rc;
    if (!p || !d_cluster) return fail(ARENA_E_INVALID, "packed snapshot or d_cluster is NULL");
    if (p->L.N && (!d_node_total || !d_node_alloc)) return fail(ARENA_E_INVALID, "node output buffers are NULL");
    cudaStream_t st = stream ? (cudaStream_t)stream : ctx().stream;
    if (p->L.K) CUDA_TRY(cudaMemsetAsync(d_node_alloc, 0, p->L.N * 8, st));      // nodes cut into pieces meet in atomics
    PackedArgs a = packed_args(p->L, (char *)p->block, p->scratch);
    a.node_total = (i64 *)d_node_total;
    a.node_alloc = (u64 *)d_node_alloc;
    a.total_capacity = (i64 *)d_cluster;
    a.total_alloc = (i64 *)d_cluster + 1;
    return packed_launch(a, p->L.bits, st);
}

int arena_top_node_host_packed(const arena_packed_view *h, int64_t *node_total, int64_t *node_alloc, int64_t cluster[2],
                               int64_t *pod_gpu)
{
    int rc = check_ready();
    if (rc) return rc;
    rc = check_packed(h);
    if (rc) return rc;
    const u64 P = h->n_pods, C = h->n_containers, N = h->n_nodes;
    std::lock_guard<std::mutex> lk(g_host_mu);
    if (!ctx().host_snap) {
        ctx().host_snap = new (std::nothrow) arena_snapshot();
        if (!ctx().host_snap) return fail(ARENA_E_NOMEM, "out of host memory");
        rc = alloc_common(ctx().host_snap);
        if (rc) { arena_snapshot_free(ctx().host_snap); ctx().host_snap = nullptr; return rc; }
    }
    if (h->ctr_bits != 4 && h->ctr_bits != 8) return fail(ARENA_E_INVALID, "ctr_bits must be 4 or 8, not %u", h->ctr_bits);
    if (N && (!node_total || !node_alloc)) return fail(ARENA_E_INVALID, "node output buffers are NULL");
    if (!cluster) return fail(ARENA_E_INVALID, "cluster is NULL");
    if (!pod_gpu) return top_node_packed_native(h, node_total, node_alloc, cluster);
    const u64 pcn[3] = {P, C, N};
    rc = fill_snapshot(ctx().host_snap, nullptr, false, pcn);       // wide columns: reserved, filled below on the device
    if (rc) return rc;
    const arena_snapshot_view &w = ctx().host_snap->v;
    cudaStream_t st = ctx().stream;

    // packed staging: [node_pod_begin | pod_meta (word padded) | ctr_gpu (word padded) | tile sums]
    if (h->ctr_bits != 4 && h->ctr_bits != 8) return fail(ARENA_E_INVALID, "ctr_bits must be 4 or 8, not %u", h->ctr_bits);
    const u64 gpu_bytes = h->ctr_bits == 4 ? (C + 1) / 2 : C;
    const u64 meta_words = (P + 3) / 4, gpu_words = (gpu_bytes + 3) / 4, tiles = (meta_words + 255) / 256;
    const size_t o_begin = 0, o_meta = o_begin + pad16((N + 1) * 4), o_gpu = o_meta + pad16(meta_words * 4 + 4),
                 o_tiles = o_gpu + pad16(gpu_words * 4 + 4), need = o_tiles + pad16((tiles + 1) * 4);
    if (ctx().packed_bytes < need) {
        if (ctx().packed_block) cudaFree(ctx().packed_block);
        ctx().packed_block = nullptr;
        ctx().packed_bytes = 0;
        CUDA_TRY(cudaMalloc(&ctx().packed_block, need));
        ctx().packed_bytes = need;
    }
    char *pb = (char *)ctx().packed_block;
    unsigned *d_begin = (unsigned *)(pb + o_begin), *d_meta = (unsigned *)(pb + o_meta), *d_gpu = (unsigned *)(pb + o_gpu),
             *d_tiles = (unsigned *)(pb + o_tiles);
    if (P & 3) CUDA_TRY(cudaMemsetAsync(d_meta + (meta_words - 1), 0, 4, st));       // the pods past P count nothing
    if (gpu_bytes & 3) CUDA_TRY(cudaMemsetAsync(d_gpu + (gpu_words - 1), 0, 4, st));
    CUDA_TRY(cudaMemcpyAsync(d_begin, h->node_pod_begin, (N + 1) * 4, cudaMemcpyHostToDevice, st));
    if (P) CUDA_TRY(cudaMemcpyAsync(d_meta, h->pod_meta, P, cudaMemcpyHostToDevice, st));
    if (C) CUDA_TRY(cudaMemcpyAsync(d_gpu, h->ctr_gpu, gpu_bytes, cudaMemcpyHostToDevice, st));
    if (N) {
        CUDA_TRY(cudaMemcpyAsync(const_cast<int64_t *>(w.node_cap_val), h->node_cap_val, N * 8, cudaMemcpyHostToDevice, st));
        CUDA_TRY(cudaMemcpyAsync(const_cast<int8_t *>(w.node_cap_scale), h->node_cap_scale, N, cudaMemcpyHostToDevice, st));
    }
    // expansion into the wide layout
    if (C) CUDA_TRY(cudaMemsetAsync(const_cast<int8_t *>(w.ctr_scale), 0, C, st));   // ctr_gpu is a resolved Value(): scale 0
    if (P) {
        packed_tile_counts_kernel<<<(unsigned)tiles, 256, 0, st>>>(d_meta, meta_words, d_tiles);
        packed_tile_scan_kernel<<<1, 1024, 0, st>>>(d_tiles, tiles);
        packed_expand_pods_kernel<<<(unsigned)tiles, 256, 0, st>>>(d_meta, meta_words, P, C, d_tiles,
                                                                  const_cast<unsigned *>(w.pod_ctr_begin),
                                                                  (unsigned *)const_cast<uint8_t *>(w.pod_phase));
        packed_expand_nodes_kernel<<<(unsigned)((N + 1 + 7) / 8), 256, 0, st>>>(d_begin, N, P, const_cast<unsigned *>(w.pod_node));
        g_launches.fetch_add(4, std::memory_order_relaxed);
    } else {
        CUDA_TRY(cudaMemsetAsync(const_cast<unsigned *>(w.pod_ctr_begin), 0, 4, st));
    }
    if (C) {
        longlong2 *dst = (longlong2 *)const_cast<int64_t *>(w.ctr_val);
        const unsigned grid = (unsigned)((gpu_words + 255) / 256);
        if (h->ctr_bits == 4) packed_expand_ctrs_kernel<4><<<grid, 256, 0, st>>>(d_gpu, gpu_words, (C + 1) / 2, dst);
        else packed_expand_ctrs_kernel<8><<<grid, 256, 0, st>>>(d_gpu, gpu_words, (C + 1) / 2, dst);
        count_launch();
    }
    CUDA_TRY(cudaGetLastError());
    ctx().host_snap->direct = false;                                // grouped by construction
    return arena_top_node(ctx().host_snap, node_total, node_alloc, cluster, pod_gpu);
}

// ---- multi-GPU totals exchange inside the kernel (see PeerExchange in kernels.cuh) ----------
int arena_peer_exchange_create(int rank, int world, unsigned char handle_out[64])
{
    int rc = check_ready();
    if (rc) return rc;
    if (world < 1 || world > kMaxPeers || rank < 0 || rank >= world || !handle_out)
        return fail(ARENA_E_INVALID, "rank %d / world %d (at most %d ranks)", rank, world, kMaxPeers);
    if (ctx().my_box) return fail(ARENA_E_INVALID, "peer exchange already created");
    static_assert(sizeof(cudaIpcMemHandle_t) == 64, "IPC handle size");
    CUDA_TRY(cudaMalloc((void **)&ctx().my_box, sizeof(PeerMailbox)));
    CUDA_TRY(cudaMemset(ctx().my_box, 0, sizeof(PeerMailbox)));
    cudaIpcMemHandle_t h;
    CUDA_TRY(cudaIpcGetMemHandle(&h, ctx().my_box));
    std::memcpy(handle_out, &h, 64);
    ctx().px_host = PeerExchange{};
    ctx().px_host.rank = rank;
    ctx().px_host.world = world;
    // (the step counter lives in the device copy of the descriptor and starts from 0 on every rank)
    return ARENA_OK;
}

int arena_peer_exchange_connect(const unsigned char *handles)
{
    int rc = check_ready();
    if (rc) return rc;
    if (!ctx().my_box || !handles) return fail(ARENA_E_INVALID, "arena_peer_exchange_create() first");
    for (int i = 0; i < ctx().px_host.world; i++) {
        if (i == ctx().px_host.rank) { ctx().px_host.box[i] = ctx().my_box; continue; }
        cudaIpcMemHandle_t h;
        std::memcpy(&h, handles + 64 * i, 64);
        void *p = nullptr;
        CUDA_TRY(cudaIpcOpenMemHandle(&p, h, cudaIpcMemLazyEnablePeerAccess));
        ctx().peer_mapped[i] = p;
        ctx().px_host.box[i] = (PeerMailbox *)p;
    }
    CUDA_TRY(cudaMalloc((void **)&ctx().px, sizeof(PeerExchange)));
    CUDA_TRY(cudaMemcpy(ctx().px, &ctx().px_host, sizeof(PeerExchange), cudaMemcpyHostToDevice));
    return ARENA_OK;
}

void arena_peer_exchange_destroy(void)
{
    if (ctx().ready) cudaSetDevice(ctx().device);
    for (int i = 0; i < kMaxPeers; i++)
        if (ctx().peer_mapped[i]) { cudaIpcCloseMemHandle(ctx().peer_mapped[i]); ctx().peer_mapped[i] = nullptr; }
    if (ctx().px) { cudaFree(ctx().px); ctx().px = nullptr; }
    if (ctx().my_box) { cudaFree(ctx().my_box); ctx().my_box = nullptr; }
    ctx().px_host = PeerExchange{};
}

int arena_top_node_device_fused(const arena_snapshot *s, int64_t *d_node_total, int64_t *d_node_alloc,
                                int64_t *d_cluster, void *stream)
{
    arena_internal::UseSlot use(s ? s->slot : arena_internal::current_slot());
    int rc = check_ready();
    if (rc) return rc;
    if (!ctx().px) return fail(ARENA_E_INVALID, "arena_peer_exchange_connect() has not been called");
    if (!s || !d_cluster) return fail(ARENA_E_INVALID, "snapshot or d_cluster is NULL");
    if (s->v.n_nodes && (!d_node_total || !d_node_alloc)) return fail(ARENA_E_INVALID, "node output buffers are NULL");
    cudaStream_t st = stream ? (cudaStream_t)stream : ctx().stream;
    return launch_top_node(s, (i64 *)d_node_total, (i64 *)d_node_alloc, (i64 *)d_cluster, nullptr, st, 1);
}

int arena_top_node_device_posted(const arena_snapshot *s, int64_t *d_node_total, int64_t *d_node_alloc,
                                 int64_t *d_cluster_prev, void *stream)
{
    arena_internal::UseSlot use(s ? s->slot : arena_internal::current_slot());
    int rc = check_ready();
    if (rc) return rc;
    if (!ctx().px) return fail(ARENA_E_INVALID, "arena_peer_exchange_connect() has not been called");
    if (!s || !d_cluster_prev) return fail(ARENA_E_INVALID, "snapshot or d_cluster_prev is NULL");
    if (s->v.n_nodes && (!d_node_total || !d_node_alloc)) return fail(ARENA_E_INVALID, "node output buffers are NULL");
    cudaStream_t st = stream ? (cudaStream_t)stream : ctx().stream;
    return launch_top_node(s, (i64 *)d_node_total, (i64 *)d_node_alloc, (i64 *)d_cluster_prev, nullptr, st, 2);
}

int arena_peer_exchange_collect(int64_t *d_cluster, void *stream)
{
    int rc = check_ready();
    if (rc) return rc;
    if (!ctx().px) return fail(ARENA_E_INVALID, "arena_peer_exchange_connect() has not been called");
    if (!d_cluster) return fail(ARENA_E_INVALID, "d_cluster is NULL");
    cudaStream_t st = stream ? (cudaStream_t)stream : ctx().stream;
    peer_collect_kernel<<<1, 32, 0, st>>>(ctx().px, (i64 *)d_cluster);
    count_launch();
    CUDA_TRY(cudaGetLastError());
    return ARENA_OK;
}

int arena_peer_exchange_timed_out(void)
{
    if (!ctx().px) return 0;
    PeerExchange h{};
    if (cudaMemcpy(&h, ctx().px, sizeof h, cudaMemcpyDeviceToHost) != cudaSuccess) return -1;
    return (int)h.timed_out;
}

int arena_top_job_device(const arena_snapshot *cs, int64_t *d_job_req, int64_t *d_job_alloc,
                         uint32_t *d_job_order, uint32_t *d_job_latest, int64_t *d_totals, void *stream)
{
    arena_internal::UseSlot use(cs ? cs->slot : arena_internal::current_slot());
    int rc = check_ready();
    if (rc) return rc;
    if (!cs || !d_totals) return fail(ARENA_E_INVALID, "snapshot or d_totals is NULL");
    arena_snapshot *s = const_cast<arena_snapshot *>(cs);
    if (s->v.n_pods && s->v.n_jobs && !s->v.pod_job) return fail(ARENA_E_INVALID, "snapshot has no job columns");
    if (s->v.n_jobs && (!d_job_req || !d_job_alloc || !d_job_order)) return fail(ARENA_E_INVALID, "job output buffers are NULL");
    cudaStream_t st = stream ? (cudaStream_t)stream : ctx().stream;
    return launch_top_job(s, (i64 *)d_job_req, (i64 *)d_job_alloc, d_job_order, d_job_latest, (i64 *)d_totals, st);
}

int arena_top_job(const arena_snapshot *cs, int64_t *job_req, int64_t *job_alloc, uint32_t *job_order,
                  uint32_t *job_latest, int64_t totals[2])
{
    arena_internal::UseSlot use(cs ? cs->slot : arena_internal::current_slot());
    int rc = check_ready();
    if (rc) return rc;
    if (!cs || !totals) return fail(ARENA_E_INVALID, "snapshot or totals is NULL");
    arena_snapshot *s = const_cast<arena_snapshot *>(cs);
    const u64 J = s->v.n_jobs;
    if (s->v.n_pods && J && !s->v.pod_job) return fail(ARENA_E_INVALID, "snapshot has no job columns");
    if (J && (!job_req || !job_alloc || !job_order)) return fail(ARENA_E_INVALID, "job output buffers are NULL");
    size_t bytes = 16 + pad16(J * 8) * 2 + pad16(J * 4) * 2;
    rc = ensure_out(s, bytes);
    if (rc) return rc;
    char *base = (char *)s->d_out;
    i64 *d_totals = (i64 *)base;
    i64 *d_req = (i64 *)(base + 16);
    i64 *d_alloc = (i64 *)(base + 16 + pad16(J * 8));
    unsigned *d_order = (unsigned *)(base + 16 + 2 * pad16(J * 8));
    unsigned *d_latest = (unsigned *)(base + 16 + 2 * pad16(J * 8) + pad16(J * 4));
    cudaStream_t st = ctx().stream;
    rc = launch_top_job(s, d_req, d_alloc, d_order, d_latest, d_totals, st);
    if (rc) return rc;
    CUDA_TRY(cudaMemcpyAsync(totals, d_totals, 16, cudaMemcpyDeviceToHost, st));
    if (J) {
        CUDA_TRY(cudaMemcpyAsync(job_req, d_req, J * 8, cudaMemcpyDeviceToHost, st));
        CUDA_TRY(cudaMemcpyAsync(job_alloc, d_alloc, J * 8, cudaMemcpyDeviceToHost, st));
        CUDA_TRY(cudaMemcpyAsync(job_order, d_order, J * 4, cudaMemcpyDeviceToHost, st));
        if (job_latest) CUDA_TRY(cudaMemcpyAsync(job_latest, d_latest, J * 4, cudaMemcpyDeviceToHost, st));
    }
    CUDA_TRY(cudaStreamSynchronize(st));
    return ARENA_OK;
}

int arena_calibrate_stream_read(const arena_snapshot *s, int iters, double *gbs_out, double *ms_out)
{
    arena_internal::UseSlot use(s ? s->slot : arena_internal::current_slot());
    int rc = check_ready();
    if (rc) return rc;
    if (!s || iters < 1 || !gbs_out) return fail(ARENA_E_INVALID, "bad argument");
    const arena_snapshot_view &v = s->v;
    StreamReadArgs a;
    const void *cols[5] = {v.pod_ctr_begin, v.pod_node, v.pod_phase, v.ctr_val, v.ctr_scale};
    const u64 bytes[5] = {v.n_pods * 4, v.n_pods * 4, v.n_pods, v.n_containers * 8, v.n_containers};
    u64 total = 0;
    for (int i = 0; i < 5; i++) { a.col[i] = (const uint4 *)cols[i]; a.n16[i] = bytes[i] / 16; total += a.n16[i] * 16; }
    a.sink = (u64 *)s->scratch;
    cudaEvent_t e0, e1;
    CUDA_TRY(cudaEventCreate(&e0));
    CUDA_TRY(cudaEventCreate(&e1));
    const int grid = ctx().sm_count * 8;
    cudaStream_t st = ctx().stream;
    stream_read_kernel<<<grid, kThreads, 0, st>>>(a);           // warm-up
    CUDA_TRY(cudaEventRecord(e0, st));
    for (int i = 0; i < iters; i++) { stream_read_kernel<<<grid, kThreads, 0, st>>>(a); count_launch(); }
    CUDA_TRY(cudaEventRecord(e1, st));
    CUDA_TRY(cudaEventSynchronize(e1));
    float ms = 0;
    CUDA_TRY(cudaEventElapsedTime(&ms, e0, e1));
    cudaEventDestroy(e0);
    cudaEventDestroy(e1);
    *gbs_out = (double)total * iters / (ms / 1e3) / 1e9;
    if (ms_out) *ms_out = ms / iters;
    return ARENA_OK;
}

int arena_calibrate_stream_pattern(const arena_snapshot *s, int iters, double *gbs_out, double *ms_out)
{
    arena_internal::UseSlot use(s ? s->slot : arena_internal::current_slot());
    int rc = check_ready();
    if (rc) return rc;
    if (!s || iters < 1 || !gbs_out) return fail(ARENA_E_INVALID, "bad argument");
    const arena_snapshot_view &v = s->v;
    PodsArgs pa{};
    pa.P = v.n_pods; pa.C = v.n_containers; pa.K = v.n_nodes;
    pa.pod_key = v.pod_node; pa.pod_phase = v.pod_phase; pa.pod_ctr_begin = v.pod_ctr_begin;
    pa.ctr_val = (const i64 *)v.ctr_val; pa.ctr_scale = (const signed char *)v.ctr_scale;
    const u64 n_full = v.n_pods / kWarpPods;
    const double total = (double)n_full * (kWarpPods * 9.0 + kChunkCtrs * 9.0);     // bytes requested per pass
    cudaEvent_t e0, e1;
    CUDA_TRY(cudaEventCreate(&e0));
    CUDA_TRY(cudaEventCreate(&e1));
    const int grid = grid_for(v.n_pods, (u64)kWarps * kWarpPods, ctx().pods_ctas_per_sm);
    cudaStream_t st = ctx().stream;
    stream_pattern_kernel<<<grid, kThreads, 0, st>>>(pa, (u64 *)s->scratch);
    CUDA_TRY(cudaEventRecord(e0, st));
    for (int i = 0; i < iters; i++) { stream_pattern_kernel<<<grid, kThreads, 0, st>>>(pa, (u64 *)s->scratch); count_launch(); }
    CUDA_TRY(cudaEventRecord(e1, st));
    CUDA_TRY(cudaEventSynchronize(e1));
    float ms = 0;
    CUDA_TRY(cudaEventElapsedTime(&ms, e0, e1));
    cudaEventDestroy(e0);
    cudaEventDestroy(e1);
    *gbs_out = total * iters / (ms / 1e3) / 1e9;
    if (ms_out) *ms_out = ms / iters;
    return ARENA_OK;
}

int arena_profile_begin(int max_calls)
{
    int rc = check_ready();
    if (rc) return rc;
    if (max_calls < 1 || max_calls > (1 << 20)) return fail(ARENA_E_INVALID, "max_calls out of range");
    for (cudaEvent_t e : g_prof.ev) cudaEventDestroy(e);
    g_prof.ev.assign((size_t)max_calls * 3, nullptr);
    for (auto &e : g_prof.ev) CUDA_TRY(cudaEventCreate(&e));
    g_prof.used = 0;
    g_prof.seen = 0;
    g_prof.every = 1;
    g_prof.on = true;
    return ARENA_OK;
}

int arena_profile_begin_sampled(int max_calls, int every)
{
    if (every < 1) return fail(ARENA_E_INVALID, "every must be >= 1");
    int rc = arena_profile_begin(max_calls);
    if (rc) return rc;
    g_prof.every = (unsigned)every;
    return ARENA_OK;
}

int arena_profile_end(double *nodes_ms_sum, double *pods_ms_sum, uint64_t *calls)
{
    int rc = check_ready();
    if (rc) return rc;
    g_prof.on = false;
    double a = 0, b = 0;
    uint64_t n = g_prof.used / 3;
    for (uint64_t i = 0; i < n; i++) {
        float ms = 0;
        CUDA_TRY(cudaEventSynchronize(g_prof.ev[3 * i + 2]));
        CUDA_TRY(cudaEventElapsedTime(&ms, g_prof.ev[3 * i], g_prof.ev[3 * i + 1]));
        a += ms;
        CUDA_TRY(cudaEventElapsedTime(&ms, g_prof.ev[3 * i + 1], g_prof.ev[3 * i + 2]));
        b += ms;
    }
    for (cudaEvent_t e : g_prof.ev) cudaEventDestroy(e);
    g_prof.ev.clear();
    g_prof.used = 0;
    if (nodes_ms_sum) *nodes_ms_sum = a;
    if (pods_ms_sum) *pods_ms_sum = b;
    if (calls) *calls = n;
    return ARENA_OK;
}

}  // extern "C" (reopened below)

// ---------------------------------------------------------------------------
// arena_top_job_device_noalloc: `top job` for callers that must not allocate.
//
// Some hosts of this library may not call cudaMalloc on the query path: a
// monitoring agent that polls `top job` on a fixed device-memory budget next to
// training jobs, a caller whose CUDA context has allocation hooked or forbidden
// after start-up, a caller in the middle of building a launch sequence where an
// allocation would synchronise the device.  The regular entry points allocate
// per-handle work arrays (per-job accumulators, sort buffers, the CTA-partials
// block) with cudaMalloc; this one keeps the same work arrays in ONE file-scope
// device array that is part of the loaded module, so nothing is allocated or
// freed here -- not by us, not by a library on our behalf.
//
//   * Size: kNoAllocMaxJobs jobs at most (the array is ~105 KiB and every GPU
//     that loads the library carries it until the process exits).  More jobs
//     are refused with ARENA_E_TOO_LARGE; use arena_top_job_device for those.
//   * The array's name is only a host-side handle for the symbol, not a device
//     address, so the device address is looked up with cudaGetSymbolAddress
//     once per device, with that device current.  The FIRST call on a device
//     performs that look-up: make it before starting any stream capture.
//   * Every call uses the same array and finds the previous call's contents in
//     it.  What has to start from a known value (the last-CTA tickets, the
//     per-job maxima, the sort padding) is re-initialised on the stream by
//     each call.  Two calls that overlap on the GPU (different streams, or
//     different host threads) would race on it: serialise them.
// ---------------------------------------------------------------------------
namespace {

constexpr unsigned kNoAllocMaxJobs = 2048;     // == kSortTile: one shared-memory sort tile

struct NoAllocWork {
    Scratch scratch;                            // CTA partials + tickets
    i64 maxcreated[kNoAllocMaxJobs];
    u64 sort_key[kNoAllocMaxJobs];
    unsigned sort_idx[kNoAllocMaxJobs];
};

static __device__ __align__(256) unsigned char g_noalloc_work[sizeof(NoAllocWork)];

constexpr int kMaxDevices = 64;
std::atomic<NoAllocWork *> g_noalloc_ptr[kMaxDevices] = {};   // device address of g_noalloc_work, per device

// One-time, per device: resolve the symbol to a device address.  Passing
// g_noalloc_work itself to a kernel from host code would hand it the host
// shadow of the symbol.
int noalloc_work(NoAllocWork **out)
{
    const int dev = ctx().device;
    if (dev < 0 || dev >= kMaxDevices) return fail(ARENA_E_INVALID, "device ordinal %d not supported by the no-alloc path", dev);
    NoAllocWork *w = g_noalloc_ptr[dev].load(std::memory_order_acquire);
    if (!w) {                                    // two threads racing here resolve the same address
        void *p = nullptr;
        cudaError_t e = cudaGetSymbolAddress(&p, g_noalloc_work);     // current device == dev (check_ready)
        if (e != cudaSuccess || !p)
            return fail(ARENA_E_CUDA, "cudaGetSymbolAddress(g_noalloc_work): %s", cudaGetErrorString(e));
        w = static_cast<NoAllocWork *>(p);
        g_noalloc_ptr[dev].store(w, std::memory_order_release);
    }
    *out = w;
    return ARENA_OK;
}

}  // namespace

extern "C" {

int arena_top_job_device_noalloc(const arena_snapshot_view *dev, int64_t *d_job_req, int64_t *d_job_alloc,
                                 uint32_t *d_job_order, uint32_t *d_job_latest, int64_t *d_totals, void *stream)
{
    int rc = check_ready();
    if (rc) return rc;
    rc = check_view(dev, true, false);           // `top job` reads neither pod_node nor the node capacities
    if (rc) return rc;
    const u64 J = dev->n_jobs, P = dev->n_pods;
    if (J > kNoAllocMaxJobs)
        return fail(ARENA_E_TOO_LARGE,
                    "arena_top_job_device_noalloc handles at most %u jobs (got %llu): its work arrays live in a "
                    "fixed device array; use arena_top_job_device for larger inputs",
                    kNoAllocMaxJobs, (unsigned long long)J);
    if (P && !dev->pod_job) return fail(ARENA_E_INVALID, "snapshot view has no job columns");
    if (!d_totals) return fail(ARENA_E_INVALID, "d_totals is NULL");
    if (J && (!d_job_req || !d_job_alloc || !d_job_order || !d_job_latest))
        return fail(ARENA_E_INVALID, "job output buffers are NULL (d_job_latest is required here: it doubles as a work array)");
    NoAllocWork *w = nullptr;
    rc = noalloc_work(&w);
    if (rc) return rc;
    cudaStream_t st = stream ? (cudaStream_t)stream : ctx().stream;

    // leftovers of an earlier (possibly interrupted) call must not leak into this one
    CUDA_TRY(cudaMemsetAsync(w->scratch.ticket, 0, sizeof(w->scratch.ticket) + sizeof(w->scratch.next_round), st));   // (adjacent)

    // the per-job accumulators are the caller's output buffers themselves
    JobInitArgs ia{J, (u64 *)d_job_req, (u64 *)d_job_alloc, w->maxcreated, d_job_latest};
    job_init_kernel<<<grid_for(J, kThreads * 4, 8), kThreads, 0, st>>>(ia);
    count_launch();

    PodsArgs pa{};
    pa.P = P; pa.C = dev->n_containers; pa.K = J;
    pa.pod_key = dev->pod_job;
    pa.pod_phase = dev->pod_phase;
    pa.pod_ctr_begin = dev->pod_ctr_begin;
    pa.ctr_val = (const i64 *)dev->ctr_val;
    pa.ctr_scale = (const signed char *)dev->ctr_scale;
    pa.pod_role = dev->pod_role;
    pa.pod_created = (const i64 *)dev->pod_created;
    pa.key_alloc = (u64 *)d_job_alloc;
    pa.key_req = (u64 *)d_job_req;
    pa.key_maxcreated = w->maxcreated;
    pa.total_req = (i64 *)d_totals;
    pa.total_alloc = (i64 *)d_totals + 1;
    pa.scratch = &w->scratch;
    job_pods_kernel<<<grid_for(P, (u64)kWarps * kWarpPods, ctx().jobs_ctas_per_sm), kThreads, 0, st>>>(pa);
    count_launch();

    if (P) {
        JobLatestArgs la{P, J, dev->pod_job, dev->pod_role, (const i64 *)dev->pod_created, w->maxcreated, d_job_latest};
        job_latest_kernel<<<grid_for(P, kThreads * 8, 8), kThreads, 0, st>>>(la);
        count_launch();
    }

    JobFinishArgs fa;
    fa.J = J; fa.Jpad = kNoAllocMaxJobs;
    fa.job_latest = d_job_latest;
    fa.pod_phase = dev->pod_phase;
    fa.pod_ctr_begin = dev->pod_ctr_begin;
    fa.ctr_val = (const i64 *)dev->ctr_val;
    fa.ctr_scale = (const signed char *)dev->ctr_scale;
    fa.job_req = (u64 *)d_job_req; fa.job_alloc = (u64 *)d_job_alloc;
    fa.sort_key = w->sort_key; fa.sort_idx = w->sort_idx;
    fa.totals = (i64 *)d_totals;
    fa.scratch = &w->scratch;
    fa.plan = nullptr;
    job_finish_kernel<<<grid_for(kNoAllocMaxJobs, kThreads, 8), kThreads, 0, st>>>(fa);
    count_launch();

    bitonic_tile_kernel<kSortTile><<<1, 1024, 0, st>>>(w->sort_key, w->sort_idx, kNoAllocMaxJobs, 0, 1);
    count_launch();
    CUDA_TRY(cudaGetLastError());
    if (J) CUDA_TRY(cudaMemcpyAsync(d_job_order, w->sort_idx, J * 4, cudaMemcpyDeviceToDevice, st));
    return ARENA_OK;
}

int arena_quantity_values(const int64_t *val, const int8_t *scale, uint64_t n, int64_t *out)
{
    int rc = check_ready();
    if (rc) return rc;
    if (n == 0) return ARENA_OK;
    if (!val || !scale || !out) return fail(ARENA_E_INVALID, "NULL argument");
    char *d = nullptr;
    CUDA_TRY(cudaMalloc((void **)&d, pad16(n * 8) * 2 + pad16(n)));
    i64 *dv = (i64 *)d;
    i64 *dout = (i64 *)(d + pad16(n * 8));
    signed char *ds = (signed char *)(d + 2 * pad16(n * 8));
    cudaStream_t st = ctx().stream;
    cudaError_t e = cudaMemcpyAsync(dv, val, n * 8, cudaMemcpyHostToDevice, st);
    if (e == cudaSuccess) e = cudaMemcpyAsync(ds, scale, n, cudaMemcpyHostToDevice, st);
    if (e == cudaSuccess) {
        quantity_values_kernel<<<grid_for(n, kThreads, 8), kThreads, 0, st>>>(dv, ds, n, dout);
        count_launch();
        e = cudaGetLastError();
    }
    if (e == cudaSuccess) e = cudaMemcpyAsync(out, dout, n * 8, cudaMemcpyDeviceToHost, st);
    if (e == cudaSuccess) e = cudaStreamSynchronize(st);
    cudaFree(d);
    if (e != cudaSuccess) return fail(ARENA_E_CUDA, "arena_quantity_values: %s", cudaGetErrorString(e));
    return ARENA_OK;
}

}  // extern "C"

#ifdef ARENA_TIMELINE
extern "C" int arena_debug_timeline(unsigned long long *out, int n)
{
    CUDA_TRY(cudaDeviceSynchronize());
    CUDA_TRY(cudaMemcpyFromSymbol(out, arena::g_timeline, sizeof(unsigned long long) * (size_t)n));
    return ARENA_OK;
}
#endif
