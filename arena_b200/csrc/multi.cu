// One process, several GPUs: `top node` over node-range shards and `top job` over job-id range
// shards (include/arena_gpuacct.h, "single process, several GPUs").  The reference CLI is ONE
// process (cmd/arena/main.go:12-28) whose `top node` loops over all nodes (top_node.go:109-118)
// and whose `top job` loops over all releases (top_job.go:53-66); this file is what lets a caller
// of that shape -- a cgo shim inside the Go binary, the arena-top CLI -- reach every GPU of the box
// without becoming N processes.
//
// Everything here goes through the library's own entry points, once per device, with that device's
// context ("slot", gpuacct_internal.h) selected for the calling thread.  Launches are asynchronous,
// so one host thread keeps all devices busy; uploads use one host thread per device because staging
// a pageable column is host work.  The cluster totals are traded inside the kernels over peer
// memory when every pair of devices can reach each other, and added up on the host otherwise.
#include "../../include/arena_gpuacct.h"
#include "../../include/arena_top.h"
#include "gpuacct_internal.h"
#include "host_snapshot.h"

#include <algorithm>
#include <cstring>
#include <mutex>
#include <queue>
#include <string>
#include <thread>
#include <vector>

using arena_host::fail;
using arena_internal::UseSlot;

namespace {

constexpr int kMaxDevices = 16;
std::mutex g_multi_mu;
int g_n_dev = 0;                 // devices bound by arena_acct_init_devices (slots 0..g_n_dev-1)
bool g_peer_totals = false;      // in-kernel totals exchange available between them

struct DeviceOut {               // per-device output staging, grown on demand
    void *p = nullptr;
    size_t bytes = 0;
};

size_t pad16(size_t n) { return (n + 15) & ~(size_t)15; }

}  // namespace

struct arena_multi_snapshot {
    int n = 0;
    bool by_job = false;
    uint64_t n_pods = 0, n_nodes = 0, n_jobs = 0;
    std::vector<uint64_t> key_cut, pod_cut;      // [n + 1]
    arena_snapshot *dev[kMaxDevices] = {};
    DeviceOut out[kMaxDevices];
};

namespace {

int ensure_out(arena_multi_snapshot *m, int i, size_t bytes)
{
    if (m->out[i].bytes >= bytes) return ARENA_OK;
    if (m->out[i].p) cudaFree(m->out[i].p);
    m->out[i] = DeviceOut{};
    cudaError_t e = cudaMalloc(&m->out[i].p, bytes);
    if (e != cudaSuccess) return fail(e == cudaErrorMemoryAllocation ? ARENA_E_NOMEM : ARENA_E_CUDA, "cudaMalloc(%zu): %s", bytes, cudaGetErrorString(e));
    m->out[i].bytes = bytes;
    return ARENA_OK;
}

#define CUDA_TRY(expr)                                                                                  \
    do {                                                                                                \
        cudaError_t e_ = (expr);                                                                        \
        if (e_ != cudaSuccess) return fail(ARENA_E_CUDA, "%s failed: %s", #expr, cudaGetErrorString(e_)); \
    } while (0)

}  // namespace

namespace arena_internal {
void multi_reset()
{
    std::lock_guard<std::mutex> lk(g_multi_mu);
    g_n_dev = 0;
    g_peer_totals = false;
}
}  // namespace arena_internal

extern "C" {

int arena_acct_init_devices(int n_dev, const int *dev_ids)
{
    if (n_dev < 1 || n_dev > kMaxDevices || !dev_ids) return fail(ARENA_E_INVALID, "n_dev %d (1..%d) / dev_ids", n_dev, kMaxDevices);
    for (int i = 0; i < n_dev; i++)
        for (int j = 0; j < i; j++)
            if (dev_ids[i] == dev_ids[j]) return fail(ARENA_E_INVALID, "device %d is listed twice", dev_ids[i]);
    std::lock_guard<std::mutex> lk(g_multi_mu);
    if (g_n_dev) {
        if (g_n_dev == n_dev) return ARENA_OK;                    // idempotent, like arena_acct_init
        return fail(ARENA_E_INVALID, "already initialised on %d devices", g_n_dev);
    }
    for (int i = 0; i < n_dev; i++) {
        UseSlot use(i);
        int rc = arena_internal::init_slot(dev_ids[i]);
        if (rc) return rc;
    }
    g_peer_totals = n_dev > 1 && arena_internal::peer_exchange_in_process(n_dev) == ARENA_OK;
    g_n_dev = n_dev;
    return ARENA_OK;
}

int arena_acct_device_count(void) { return g_n_dev; }

void arena_multi_snapshot_free(arena_multi_snapshot *m)
{
    if (!m) return;
    for (int i = 0; i < m->n; i++) {
        if (m->dev[i]) arena_snapshot_free(m->dev[i]);           // selects the handle's own slot
        if (m->out[i].p) cudaFree(m->out[i].p);
    }
    delete m;
}

int arena_multi_snapshot_upload(const arena_snapshot_view *host, int by_job, arena_multi_snapshot **out)
{
    if (!out) return fail(ARENA_E_INVALID, "out is NULL");
    *out = nullptr;
    if (!host) return fail(ARENA_E_INVALID, "snapshot view is NULL");
    const int n = g_n_dev;
    if (n < 1) return fail(ARENA_E_NOT_INIT, "arena_acct_init_devices() has not been called");
    if (by_job && host->n_pods && !host->pod_job) return fail(ARENA_E_INVALID, "snapshot has no job columns");
    arena_multi_snapshot *m = new (std::nothrow) arena_multi_snapshot();
    if (!m) return fail(ARENA_E_NOMEM, "out of host memory");
    m->n = n;
    m->by_job = by_job != 0;
    m->n_pods = host->n_pods; m->n_nodes = host->n_nodes; m->n_jobs = host->n_jobs;
    m->key_cut.assign(n + 1, 0);
    m->pod_cut.assign(n + 1, 0);
    int rc = arena_shard_plan(host, n, by_job, m->key_cut.data(), m->pod_cut.data());
    if (rc) { delete m; return rc; }
    // one host thread per device: cutting a shard and staging its pageable columns is host work
    std::vector<int> codes(n, ARENA_OK);
    std::vector<std::string> texts(n);
    auto work = [&](int i) {
        UseSlot use(i);
        arena_host_snapshot *sh = nullptr;
        int c = arena_shard_extract(host, by_job, m->key_cut[i], m->key_cut[i + 1], m->pod_cut[i], m->pod_cut[i + 1], &sh);
        if (!c) {
            arena_snapshot_view v;
            arena_host_snapshot_view(sh, &v);
            c = arena_snapshot_upload(&v, &m->dev[i]);           // returns once the copies are done: the shard may go
        }
        if (c) texts[i] = arena_last_error();
        if (sh) arena_host_snapshot_free(sh);
        codes[i] = c;
    };
    std::vector<std::thread> pool;
    for (int i = 1; i < n; i++) pool.emplace_back(work, i);
    work(0);
    for (std::thread &t : pool) t.join();
    for (int i = 0; i < n; i++)
        if (codes[i]) {
            rc = fail(codes[i], "device slot %d: %s", i, texts[i].c_str());
            arena_multi_snapshot_free(m);
            return rc;
        }
    *out = m;
    return ARENA_OK;
}

int arena_multi_snapshot_shards(const arena_multi_snapshot *m, int *n_shards, uint64_t *key_cut, uint64_t *pod_cut)
{
    if (!m) return fail(ARENA_E_INVALID, "multi snapshot is NULL");
    if (n_shards) *n_shards = m->n;
    for (int i = 0; i <= m->n; i++) {
        if (key_cut) key_cut[i] = m->key_cut[i];
        if (pod_cut) pod_cut[i] = m->pod_cut[i];
    }
    return ARENA_OK;
}

int arena_top_node_multi(const arena_multi_snapshot *cm, int64_t *node_total, int64_t *node_alloc, int64_t cluster[2],
                         int64_t *pod_gpu)
{
    arena_multi_snapshot *m = const_cast<arena_multi_snapshot *>(cm);
    if (!m || !cluster) return fail(ARENA_E_INVALID, "multi snapshot or cluster is NULL");
    if (m->by_job) return fail(ARENA_E_INVALID, "this snapshot was cut by job range: use arena_top_job_multi");
    if (m->n_nodes && (!node_total || !node_alloc)) return fail(ARENA_E_INVALID, "node output buffers are NULL");
    const int n = m->n;
    const bool peer = g_peer_totals && n > 1 && !pod_gpu;        // the details view goes through the plain launch
    cudaStream_t st[kMaxDevices];
    int64_t *d_cluster[kMaxDevices], *d_total[kMaxDevices], *d_alloc[kMaxDevices], *d_pg[kMaxDevices];
    // 1. every device computes its node range (asynchronous launches: they all run at once)
    for (int i = 0; i < n; i++) {
        UseSlot use(i);
        int rc = arena_internal::ready(&st[i], nullptr);
        if (rc) return rc;
        const uint64_t Ni = m->key_cut[i + 1] - m->key_cut[i], Pi = m->pod_cut[i + 1] - m->pod_cut[i];
        rc = ensure_out(m, i, 16 + 2 * pad16(Ni * 8) + (pod_gpu ? pad16(Pi * 8) : 0));
        if (rc) return rc;
        char *base = (char *)m->out[i].p;
        d_cluster[i] = (int64_t *)base;
        d_total[i] = (int64_t *)(base + 16);
        d_alloc[i] = (int64_t *)(base + 16 + pad16(Ni * 8));
        d_pg[i] = pod_gpu ? (int64_t *)(base + 16 + 2 * pad16(Ni * 8)) : nullptr;
        rc = peer ? arena_top_node_device_fused(m->dev[i], d_total[i], d_alloc[i], d_cluster[i], st[i])
                  : arena_top_node_device(m->dev[i], d_total[i], d_alloc[i], d_cluster[i], d_pg[i], st[i]);
        if (rc) return rc;
    }
    // 2. the slices come home: disjoint node ranges, so they land side by side in the caller's table
    int64_t part[kMaxDevices][2];
    for (int i = 0; i < n; i++) {
        UseSlot use(i);
        arena_internal::ready(nullptr, nullptr);                 // makes the device current
        const uint64_t lo = m->key_cut[i], Ni = m->key_cut[i + 1] - lo, Pi = m->pod_cut[i + 1] - m->pod_cut[i];
        CUDA_TRY(cudaMemcpyAsync(part[i], d_cluster[i], 16, cudaMemcpyDeviceToHost, st[i]));
        if (Ni) {
            CUDA_TRY(cudaMemcpyAsync(node_total + lo, d_total[i], Ni * 8, cudaMemcpyDeviceToHost, st[i]));
            CUDA_TRY(cudaMemcpyAsync(node_alloc + lo, d_alloc[i], Ni * 8, cudaMemcpyDeviceToHost, st[i]));
        }
        if (pod_gpu && Pi) CUDA_TRY(cudaMemcpyAsync(pod_gpu + m->pod_cut[i], d_pg[i], Pi * 8, cudaMemcpyDeviceToHost, st[i]));
    }
    for (int i = 0; i < n; i++) {
        UseSlot use(i);
        arena_internal::ready(nullptr, nullptr);
        CUDA_TRY(cudaStreamSynchronize(st[i]));
    }
    // 3. cluster totals (top_node.go:109-118): every device already holds the global pair when the kernels
    // traded them over peer memory; otherwise the host adds the per-device pairs (wrapping, like Go's +=)
    if (peer) {
        if (arena_peer_exchange_timed_out())
            return fail(ARENA_E_CUDA, "a device did not post its totals in time (in-kernel exchange)");
        cluster[0] = part[0][0];
        cluster[1] = part[0][1];
        for (int i = 1; i < n; i++)
            if (part[i][0] != part[0][0] || part[i][1] != part[0][1])
                return fail(ARENA_E_CUDA, "devices disagree on the exchanged cluster totals");
    } else {
        uint64_t a = 0, b = 0;
        for (int i = 0; i < n; i++) { a += (uint64_t)part[i][0]; b += (uint64_t)part[i][1]; }
        cluster[0] = (int64_t)a;
        cluster[1] = (int64_t)b;
    }
    return ARENA_OK;
}

int arena_top_job_multi(const arena_multi_snapshot *cm, int64_t *job_req, int64_t *job_alloc, uint32_t *job_order,
                        uint32_t *job_latest, int64_t totals[2])
{
    arena_multi_snapshot *m = const_cast<arena_multi_snapshot *>(cm);
    if (!m || !totals) return fail(ARENA_E_INVALID, "multi snapshot or totals is NULL");
    if (!m->by_job) return fail(ARENA_E_INVALID, "this snapshot was cut by node range: upload it with by_job = 1 for `top job`");
    const uint64_t J = m->n_jobs;
    if (J && (!job_req || !job_alloc || !job_order)) return fail(ARENA_E_INVALID, "job output buffers are NULL");
    const int n = m->n;
    cudaStream_t st[kMaxDevices];
    int64_t *d_totals[kMaxDevices], *d_req[kMaxDevices], *d_alloc[kMaxDevices];
    uint32_t *d_order[kMaxDevices], *d_latest[kMaxDevices];
    for (int i = 0; i < n; i++) {
        UseSlot use(i);
        int rc = arena_internal::ready(&st[i], nullptr);
        if (rc) return rc;
        const uint64_t Ji = m->key_cut[i + 1] - m->key_cut[i];
        rc = ensure_out(m, i, 16 + 2 * pad16(Ji * 8) + 2 * pad16(Ji * 4));
        if (rc) return rc;
        char *base = (char *)m->out[i].p;
        d_totals[i] = (int64_t *)base;
        d_req[i] = (int64_t *)(base + 16);
        d_alloc[i] = (int64_t *)(base + 16 + pad16(Ji * 8));
        d_order[i] = (uint32_t *)(base + 16 + 2 * pad16(Ji * 8));
        d_latest[i] = (uint32_t *)(base + 16 + 2 * pad16(Ji * 8) + pad16(Ji * 4));
        rc = arena_top_job_device(m->dev[i], d_req[i], d_alloc[i], d_order[i], d_latest[i], d_totals[i], st[i]);
        if (rc) return rc;
    }
    std::vector<uint32_t> local_order(J);                         // shard-local ids, shard after shard
    int64_t part[kMaxDevices][2];
    for (int i = 0; i < n; i++) {
        UseSlot use(i);
        arena_internal::ready(nullptr, nullptr);
        const uint64_t lo = m->key_cut[i], Ji = m->key_cut[i + 1] - lo;
        CUDA_TRY(cudaMemcpyAsync(part[i], d_totals[i], 16, cudaMemcpyDeviceToHost, st[i]));
        if (!Ji) continue;
        CUDA_TRY(cudaMemcpyAsync(job_req + lo, d_req[i], Ji * 8, cudaMemcpyDeviceToHost, st[i]));
        CUDA_TRY(cudaMemcpyAsync(job_alloc + lo, d_alloc[i], Ji * 8, cudaMemcpyDeviceToHost, st[i]));
        CUDA_TRY(cudaMemcpyAsync(local_order.data() + lo, d_order[i], Ji * 4, cudaMemcpyDeviceToHost, st[i]));
        if (job_latest) CUDA_TRY(cudaMemcpyAsync(job_latest + lo, d_latest[i], Ji * 4, cudaMemcpyDeviceToHost, st[i]));
    }
    for (int i = 0; i < n; i++) {
        UseSlot use(i);
        arena_internal::ready(nullptr, nullptr);
        CUDA_TRY(cudaStreamSynchronize(st[i]));
    }
    uint64_t a = 0, b = 0;                                        // list.go:102-103, wrapping like Go's +=
    for (int i = 0; i < n; i++) { a += (uint64_t)part[i][0]; b += (uint64_t)part[i][1]; }
    totals[0] = (int64_t)a;
    totals[1] = (int64_t)b;
    // winning rows are shard-local: back to rows of the snapshot that was uploaded
    if (job_latest)
        for (int i = 0; i < n; i++)
            for (uint64_t j = m->key_cut[i]; j < m->key_cut[i + 1]; j++)
                if (job_latest[j] != ARENA_NONE) job_latest[j] += (uint32_t)m->pod_cut[i];
    // ordering (trainer.go:25-37 with the documented tie-break): every device has its job range in
    // (requested desc, id asc) order already, and the ranges ascend in id, so the global order is a
    // k-way merge that prefers the lower shard on equal requested values
    struct Head { int64_t req; int shard; };
    auto after = [](const Head &x, const Head &y) { return x.req != y.req ? x.req < y.req : x.shard > y.shard; };
    std::priority_queue<Head, std::vector<Head>, decltype(after)> heap(after);
    std::vector<uint64_t> pos(n);
    for (int i = 0; i < n; i++) {
        pos[i] = m->key_cut[i];
        if (pos[i] < m->key_cut[i + 1]) heap.push(Head{job_req[m->key_cut[i] + local_order[pos[i]]], i});
    }
    for (uint64_t k = 0; k < J; k++) {
        const Head h = heap.top();
        heap.pop();
        const int i = h.shard;
        job_order[k] = (uint32_t)(m->key_cut[i] + local_order[pos[i]]);
        if (++pos[i] < m->key_cut[i + 1]) heap.push(Head{job_req[m->key_cut[i] + local_order[pos[i]]], i});
    }
    return ARENA_OK;
}

}  // extern "C"
