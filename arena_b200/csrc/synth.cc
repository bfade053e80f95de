// Seeded synthetic columnar snapshots ("Synthetic
// inputs"), host column storage, and the node-range / job-range sharder.
//
// Everything is counter-based (value = hash(seed, stream, index)), so the
// result is independent of the number of worker threads.
#include "host_snapshot.h"

#include <cuda_runtime.h>

#include <algorithm>
#include <cmath>
#include <cstdlib>
#include <cstring>
#include <numeric>
#include <thread>

namespace arena_host {

Column::~Column()
{
    if (!ptr) return;
    if (pinned) cudaFreeHost(ptr);
    else free(ptr);
}

int Column::alloc(size_t nbytes, bool want_pinned)
{
    size_t n = (nbytes + 63) & ~(size_t)63;
    if (n == 0) n = 64;
    if (want_pinned) {
        cudaError_t e = cudaHostAlloc(&ptr, n, cudaHostAllocDefault);
        if (e != cudaSuccess) {
            ptr = nullptr;
            return fail(ARENA_E_CUDA, "cudaHostAlloc(%zu): %s", n, cudaGetErrorString(e));
        }
        pinned = true;
    } else {
        if (posix_memalign(&ptr, 64, n) != 0) {
            ptr = nullptr;
            return fail(ARENA_E_NOMEM, "posix_memalign(%zu) failed", n);
        }
    }
    bytes = nbytes;
    return ARENA_OK;
}

namespace {

inline uint64_t mix64(uint64_t z)
{
    z += 0x9E3779B97F4A7C15ull;
    z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
    z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
    return z ^ (z >> 31);
}

inline uint64_t draw(uint64_t seed, uint64_t stream, uint64_t index)
{
    return mix64(mix64(seed ^ (stream * 0xD1B54A32D192ED03ull)) + index * 0x9E3779B97F4A7C15ull);
}

template <class F> void parallel_for(uint64_t n, int threads, F f)
{
    if (threads <= 1 || n < 65536) { f(0, n, 0); return; }
    std::vector<std::thread> ts;
    uint64_t per = (n + threads - 1) / threads;
    for (int t = 0; t < threads; t++) {
        uint64_t lo = std::min<uint64_t>(n, per * t), hi = std::min<uint64_t>(n, per * (t + 1));
        ts.emplace_back([=] { f(lo, hi, t); });
    }
    for (auto &t : ts) t.join();
}

// phases: Running 70 / Pending 10 / Succeeded 12 / Failed 6 / Unknown 2 (%)
inline uint8_t synth_phase(uint64_t r)
{
    unsigned x = (unsigned)(r % 100);
    if (x < 70) return ARENA_PHASE_RUNNING;
    if (x < 80) return ARENA_PHASE_PENDING;
    if (x < 92) return ARENA_PHASE_SUCCEEDED;
    if (x < 98) return ARENA_PHASE_FAILED;
    return ARENA_PHASE_UNKNOWN;
}

// containers per pod: {1:.60, 2:.30, 3:.08, 4:.02}
inline unsigned synth_nctr(uint64_t r)
{
    unsigned x = (unsigned)(r % 100);
    if (x < 60) return 1;
    if (x < 90) return 2;
    if (x < 98) return 3;
    return 4;
}

// per-container limit: {0:.50, 1:.25, 2:.10, 4:.08, 8:.05, "500m":.02}
inline void synth_limit(uint64_t r, int64_t *val, int8_t *scale)
{
    unsigned x = (unsigned)(r % 100);
    *scale = 0;
    if (x < 50) *val = 0;
    else if (x < 75) *val = 1;
    else if (x < 85) *val = 2;
    else if (x < 93) *val = 4;
    else if (x < 98) *val = 8;
    else { *val = 500; *scale = -3; }
}

uint64_t gcd64(uint64_t a, uint64_t b) { while (b) { uint64_t t = a % b; a = b; b = t; } return a; }

}  // namespace
}  // namespace arena_host

using namespace arena_host;

extern "C" {

void arena_host_snapshot_free(arena_host_snapshot *h) { delete h; }

int arena_host_snapshot_view(const arena_host_snapshot *h, arena_snapshot_view *out)
{
    if (!h || !out) return fail(ARENA_E_INVALID, "NULL argument");
    *out = h->v;
    return ARENA_OK;
}

const uint32_t *arena_host_pod_order(const arena_host_snapshot *h)
{
    return (h && !h->pod_order.empty()) ? h->pod_order.data() : nullptr;
}

int arena_synth_snapshot(const arena_synth_params *sp, arena_host_snapshot **out)
{
    if (!sp || !out) return fail(ARENA_E_INVALID, "NULL argument");
    *out = nullptr;
    const uint64_t P = sp->n_pods, N = sp->n_nodes, J = sp->n_jobs, seed = sp->seed;
    if (P >= 0xFFFFFFFFull || N >= 0xFFFFFFFFull || J >= 0xFFFFFFFFull) return fail(ARENA_E_TOO_LARGE, "counts must be < 2^32-1");
    if (P > 0 && N == 0) return fail(ARENA_E_INVALID, "pods need at least one node");
    int threads = sp->threads > 0 ? sp->threads : (int)std::thread::hardware_concurrency();
    if (threads < 1) threads = 1;
    if (threads > 32) threads = 32;
    const bool pin = sp->pinned != 0;

    arena_host_snapshot *h = new arena_host_snapshot();
    int rc = 0;
    auto bail = [&](int code) { delete h; return code; };
    if ((rc = h->c_pod_node.alloc(P * 4, pin))) return bail(rc);
    if ((rc = h->c_pod_phase.alloc(P, pin))) return bail(rc);
    if ((rc = h->c_pod_ctr_begin.alloc((P + 1) * 4, pin))) return bail(rc);
    if ((rc = h->c_node_cap_val.alloc(N * 8, pin))) return bail(rc);
    if ((rc = h->c_node_cap_scale.alloc(N, pin))) return bail(rc);
    uint32_t *pod_node = h->c_pod_node.as<uint32_t>();
    uint8_t *pod_phase = h->c_pod_phase.as<uint8_t>();
    uint32_t *ctr_begin = h->c_pod_ctr_begin.as<uint32_t>();
    int64_t *cap_val = h->c_node_cap_val.as<int64_t>();
    int8_t *cap_scale = h->c_node_cap_scale.as<int8_t>();

    // ---- node capacity {0:.2, 4:.2, 8:.6}; GPU-sharing flag (skewed only) ----
    std::vector<uint8_t> sharing(sp->skewed ? N : 0);
    for (uint64_t n = 0; n < N; n++) {
        unsigned x = (unsigned)(draw(seed, 5, n) % 10);
        cap_val[n] = x < 2 ? 0 : (x < 4 ? 4 : 8);
        cap_scale[n] = 0;
        if (sp->skewed) sharing[n] = (draw(seed, 6, n) % 10) == 0;
    }

    // ---- pods per node -----------------------------------------------------
    // stream 1 decides scheduled / unscheduled (3%) and, for the uniform case,
    // the node of each pod.
    std::vector<uint64_t> count(N + 1, 0);       // count[N] = unscheduled
    if (!sp->skewed) {
        std::vector<std::vector<uint32_t>> local(threads);
        parallel_for(P, threads, [&](uint64_t lo, uint64_t hi, int t) {
            std::vector<uint32_t> &c = local[t];
            c.assign(N + 1, 0);
            for (uint64_t p = lo; p < hi; p++) {
                uint64_t r = draw(seed, 1, p);
                if (r % 100 < 3) { c[N]++; if (!sp->grouped) pod_node[p] = ARENA_NONE; }
                else {
                    uint32_t n = (uint32_t)((r >> 24) % N);
                    c[n]++;
                    if (!sp->grouped) pod_node[p] = n;
                }
            }
        });
        for (auto &c : local)
            if (!c.empty())
                for (uint64_t n = 0; n <= N; n++) count[n] += c[n];
    } else {
        uint64_t unsched = P * 3 / 100;
        uint64_t sched = P - unsched;
        count[N] = unsched;
        const uint64_t cap = 1000000;
        std::vector<double> w(N);
        for (uint64_t r = 0; r < N; r++) w[r] = std::pow((double)(r + 1), -1.1);
        std::vector<uint64_t> by_rank(N, 0);
        // water-filling: ranks whose share exceeds the cap are pinned at the cap
        uint64_t remaining = sched;
        uint64_t first_free = 0;
        for (int iter = 0; iter < 64 && first_free < N; iter++) {
            double W = 0;
            for (uint64_t r = first_free; r < N; r++) W += w[r];
            uint64_t newly = 0;
            uint64_t r = first_free;
            while (r < N && (double)remaining * w[r] / W > (double)cap) { by_rank[r] = cap; newly++; r++; }
            if (!newly) {
                uint64_t used = 0;
                for (uint64_t q = first_free; q < N; q++) {
                    by_rank[q] = (uint64_t)std::floor((double)remaining * w[q] / W);
                    used += by_rank[q];
                }
                uint64_t left = remaining - used;           // < N - first_free
                for (uint64_t q = first_free; left > 0 && q < N; q++) { by_rank[q]++; left--; }
                remaining = 0;
                break;
            }
            if (newly * cap > remaining) {                   // tiny inputs: just fill in order
                for (uint64_t q = first_free; q < first_free + newly; q++) {
                    by_rank[q] = std::min<uint64_t>(cap, remaining);
                    remaining -= by_rank[q];
                }
                break;
            }
            remaining -= newly * cap;
            first_free += newly;
        }
        count[N] += remaining;                               // cannot be placed under the cap
        // rank -> node id through an affine bijection
        uint64_t A = (uint64_t)((double)N * 0.6180339887) | 1ull;
        while (gcd64(A, N) != 1) A += 2;
        for (uint64_t r = 0; r < N; r++) count[(r * A + 7) % N] = by_rank[r];
        if (!sp->grouped) { delete h; return fail(ARENA_E_INVALID, "skewed snapshots are generated grouped only"); }
    }

    if (sp->grouped) {
        std::vector<uint64_t> begin(N + 2, 0);
        for (uint64_t n = 0; n <= N; n++) begin[n + 1] = begin[n] + count[n];
        parallel_for(N + 1, threads, [&](uint64_t lo, uint64_t hi, int) {
            for (uint64_t n = lo; n < hi; n++) {
                uint32_t id = n == N ? ARENA_NONE : (uint32_t)n;
                for (uint64_t p = begin[n]; p < begin[n + 1]; p++) pod_node[p] = id;
            }
        });
    }

    // ---- phases and container counts ------------------------------------------
    std::vector<uint64_t> chunk_ctrs(threads + 1, 0);
    parallel_for(P, threads, [&](uint64_t lo, uint64_t hi, int t) {
        uint64_t s = 0;
        for (uint64_t p = lo; p < hi; p++) {
            pod_phase[p] = synth_phase(draw(seed, 2, p));
            s += synth_nctr(draw(seed, 3, p));
        }
        chunk_ctrs[t + 1] = s;
    });
    for (int t = 0; t < threads; t++) chunk_ctrs[t + 1] += chunk_ctrs[t];
    uint64_t C = 0;
    if (P < 65536 || threads <= 1) C = chunk_ctrs[1];
    else C = chunk_ctrs[threads];
    if (C > 0xFFFFFFFFull) { delete h; return fail(ARENA_E_TOO_LARGE, "more than 2^32-1 containers"); }
    if ((rc = h->c_ctr_val.alloc(C * 8, pin))) return bail(rc);
    if ((rc = h->c_ctr_scale.alloc(C, pin))) return bail(rc);
    int64_t *ctr_val = h->c_ctr_val.as<int64_t>();
    int8_t *ctr_scale = h->c_ctr_scale.as<int8_t>();
    parallel_for(P, threads, [&](uint64_t lo, uint64_t hi, int t) {
        uint64_t c = chunk_ctrs[t];
        for (uint64_t p = lo; p < hi; p++) {
            ctr_begin[p] = (uint32_t)c;
            unsigned k = synth_nctr(draw(seed, 3, p));
            bool share = sp->skewed && pod_node[p] != ARENA_NONE && sharing[pod_node[p]];
            for (unsigned j = 0; j < k; j++) {
                uint64_t r = draw(seed, 4, p * 8 + j);
                if (share) { ctr_val[c] = (r & 1) ? 500 : 250; ctr_scale[c] = -3; }
                else synth_limit(r, &ctr_val[c], &ctr_scale[c]);
                c++;
            }
        }
    });
    ctr_begin[P] = (uint32_t)C;

    // ---- job columns --------------------------------------------------------------
    if (J > 0) {
        if ((rc = h->c_pod_job.alloc(P * 4, pin))) return bail(rc);
        if ((rc = h->c_pod_role.alloc(P, pin))) return bail(rc);
        if ((rc = h->c_pod_created.alloc(P * 8, pin))) return bail(rc);
        uint32_t *pod_job = h->c_pod_job.as<uint32_t>();
        uint8_t *pod_role = h->c_pod_role.as<uint8_t>();
        int64_t *pod_created = h->c_pod_created.as<int64_t>();
        parallel_for(P, threads, [&](uint64_t lo, uint64_t hi, int) {
            for (uint64_t p = lo; p < hi; p++) {
                uint64_t r = draw(seed, 7, p);
                pod_job[p] = (r % 50 == 0) ? ARENA_NONE : (uint32_t)((r >> 20) % J);
                pod_role[p] = ((r >> 8) % 10 == 0) ? ARENA_ROLE_LATEST : ARENA_ROLE_COUNTS;
                pod_created[p] = 1532304000ll + (int64_t)((r >> 40) % 600);   // second resolution: ties happen
            }
        });
    }

    h->v.n_pods = P; h->v.n_containers = C; h->v.n_nodes = N; h->v.n_jobs = J;
    h->v.pod_node = pod_node; h->v.pod_phase = pod_phase; h->v.pod_ctr_begin = ctr_begin;
    h->v.ctr_val = ctr_val; h->v.ctr_scale = ctr_scale;
    h->v.node_cap_val = cap_val; h->v.node_cap_scale = cap_scale;
    if (J > 0) {
        h->v.pod_job = h->c_pod_job.as<uint32_t>();
        h->v.pod_role = h->c_pod_role.as<uint8_t>();
        h->v.pod_created = h->c_pod_created.as<int64_t>();
    }
    *out = h;
    return ARENA_OK;
}

int arena_shard_plan(const arena_snapshot_view *v, int n_shards, int by_job, uint64_t *key_cut, uint64_t *pod_cut)
{
    if (!v || !key_cut || !pod_cut || n_shards < 1) return fail(ARENA_E_INVALID, "bad argument");
    const uint32_t *key = by_job ? v->pod_job : v->pod_node;
    const uint64_t K = by_job ? v->n_jobs : v->n_nodes;
    const uint64_t P = v->n_pods;
    if (P && !key) return fail(ARENA_E_INVALID, "key column is NULL");
    for (uint64_t p = 1; p < P; p++)
        if (key[p] < key[p - 1]) return fail(ARENA_E_INVALID, "snapshot is not grouped: key column must be non-decreasing to shard by range");
    // number of pods with a real key (the rest sort last)
    uint64_t Pk = std::lower_bound(key, key + P, (uint32_t)std::min<uint64_t>(K, 0xFFFFFFFFull)) - key;
    key_cut[0] = 0; pod_cut[0] = 0;
    for (int s = 1; s < n_shards; s++) {
        // equal shares of ALL pods: the unmatched ones sit at the end, so they
        // count towards the last shard's share
        uint64_t target = P * (uint64_t)s / (uint64_t)n_shards;
        uint64_t k;
        if (target >= Pk) k = K;
        else {
            // cut at the boundary of the key holding pod `target` that is nearer to it
            uint32_t at = key[target];
            uint64_t lo = std::lower_bound(key, key + Pk, at) - key;
            uint64_t hi = std::upper_bound(key, key + Pk, at) - key;
            k = (target - lo <= hi - target) ? at : (uint64_t)at + 1;
        }
        if (k < key_cut[s - 1]) k = key_cut[s - 1];
        if (k > K) k = K;
        key_cut[s] = k;
        pod_cut[s] = std::lower_bound(key, key + Pk, (uint32_t)std::min<uint64_t>(k, 0xFFFFFFFFull)) - key;
    }
    key_cut[n_shards] = K;
    pod_cut[n_shards] = P;      // unmatched pods ride with the last shard
    return ARENA_OK;
}

int arena_shard_extract(const arena_snapshot_view *v, int by_job, uint64_t key_lo, uint64_t key_hi,
                        uint64_t pod_lo, uint64_t pod_hi, arena_host_snapshot **out)
{
    if (!v || !out) return fail(ARENA_E_INVALID, "NULL argument");
    *out = nullptr;
    if (pod_lo > pod_hi || pod_hi > v->n_pods || key_lo > key_hi) return fail(ARENA_E_INVALID, "bad shard bounds");
    const uint64_t K = by_job ? v->n_jobs : v->n_nodes;
    if (key_hi > K) return fail(ARENA_E_INVALID, "key_hi beyond key count");
    const uint64_t Ps = pod_hi - pod_lo;
    arena_host_snapshot *h = new arena_host_snapshot();
    int rc;
    auto bail = [&](int code) { delete h; return code; };
    if ((rc = h->c_pod_ctr_begin.alloc((Ps + 1) * 4, false))) return bail(rc);
    Column &kc = by_job ? h->c_pod_job : h->c_pod_node;
    if ((rc = kc.alloc(Ps * 4, false))) return bail(rc);
    uint32_t *cb = h->c_pod_ctr_begin.as<uint32_t>();
    uint32_t *keys = kc.as<uint32_t>();
    const uint32_t *src_key = by_job ? v->pod_job : v->pod_node;
    const uint32_t c0 = v->pod_ctr_begin[pod_lo];
    for (uint64_t p = 0; p <= Ps; p++) cb[p] = v->pod_ctr_begin[pod_lo + p] - c0;
    for (uint64_t p = 0; p < Ps; p++) {
        uint32_t k = src_key[pod_lo + p];
        keys[p] = (k >= key_lo && k < key_hi) ? (uint32_t)(k - key_lo) : ARENA_NONE;
    }
    h->v = *v;
    h->v.n_pods = Ps;
    h->v.n_containers = cb[Ps];
    h->v.pod_ctr_begin = cb;
    h->v.pod_phase = v->pod_phase + pod_lo;
    h->v.ctr_val = v->ctr_val + c0;
    h->v.ctr_scale = v->ctr_scale + c0;
    if (by_job) {
        h->v.n_jobs = key_hi - key_lo;
        h->v.pod_job = keys;
        h->v.pod_node = v->pod_node + pod_lo;      // node ids stay global
        h->v.pod_role = v->pod_role + pod_lo;
        h->v.pod_created = v->pod_created + pod_lo;
    } else {
        h->v.n_nodes = key_hi - key_lo;
        h->v.pod_node = keys;
        h->v.node_cap_val = v->node_cap_val + key_lo;
        h->v.node_cap_scale = v->node_cap_scale + key_lo;
        h->v.n_jobs = 0; h->v.pod_job = nullptr; h->v.pod_role = nullptr; h->v.pod_created = nullptr;
    }
    *out = h;
    return ARENA_OK;
}

// Rows regrouped by key (stable counting sort: rows of one key keep their order, which is what the
// latest-pod tie rule needs); rows whose key names no node / job go last.  Owns all its pod and
// container columns; the node columns are copied so that the result outlives its source.
int arena_host_snapshot_regroup(const arena_snapshot_view *v, int by_job, arena_host_snapshot **out)
{
    if (!v || !out) return fail(ARENA_E_INVALID, "NULL argument");
    *out = nullptr;
    const uint32_t *key = by_job ? v->pod_job : v->pod_node;
    const uint64_t K = by_job ? v->n_jobs : v->n_nodes, P = v->n_pods, C = v->n_containers, N = v->n_nodes;
    if (P && (!key || !v->pod_node || !v->pod_phase || !v->pod_ctr_begin)) return fail(ARENA_E_INVALID, "pod columns are NULL");
    const bool jobs = v->pod_job && v->pod_role && v->pod_created;
    if (by_job && !jobs) return fail(ARENA_E_INVALID, "snapshot has no job columns to group by");
    arena_host_snapshot *h = new arena_host_snapshot();
    int rc;
    auto bail = [&](int code) { delete h; return code; };
    if ((rc = h->c_pod_node.alloc(P * 4, false)) || (rc = h->c_pod_phase.alloc(P, false)) ||
        (rc = h->c_pod_ctr_begin.alloc((P + 1) * 4, false)) || (rc = h->c_ctr_val.alloc(C * 8, false)) ||
        (rc = h->c_ctr_scale.alloc(C, false)) || (rc = h->c_node_cap_val.alloc(N * 8, false)) ||
        (rc = h->c_node_cap_scale.alloc(N, false)))
        return bail(rc);
    if (jobs && ((rc = h->c_pod_job.alloc(P * 4, false)) || (rc = h->c_pod_role.alloc(P, false)) ||
                 (rc = h->c_pod_created.alloc(P * 8, false))))
        return bail(rc);
    std::vector<uint64_t> start(K + 2, 0);                       // bucket K: rows with no key
    for (uint64_t p = 0; p < P; p++) start[std::min<uint64_t>(key[p], K) + 1]++;
    for (uint64_t k = 0; k <= K; k++) start[k + 1] += start[k];
    std::vector<uint32_t> row_of(P);                             // old row -> new row
    for (uint64_t p = 0; p < P; p++) row_of[p] = (uint32_t)start[std::min<uint64_t>(key[p], K)]++;
    uint32_t *cb = h->c_pod_ctr_begin.as<uint32_t>();
    for (uint64_t p = 0; p < P; p++) cb[row_of[p] + 1] = v->pod_ctr_begin[p + 1] - v->pod_ctr_begin[p];
    if (P + 1) cb[0] = 0;
    for (uint64_t r = 0; r < P; r++) cb[r + 1] += cb[r];
    uint32_t *pn = h->c_pod_node.as<uint32_t>();
    uint8_t *ph = h->c_pod_phase.as<uint8_t>();
    int64_t *cv = h->c_ctr_val.as<int64_t>();
    int8_t *cs = h->c_ctr_scale.as<int8_t>();
    uint32_t *pj = h->c_pod_job.as<uint32_t>();
    uint8_t *pr = h->c_pod_role.as<uint8_t>();
    int64_t *pc = h->c_pod_created.as<int64_t>();
    for (uint64_t p = 0; p < P; p++) {
        const uint32_t r = row_of[p];
        pn[r] = v->pod_node[p];
        ph[r] = v->pod_phase[p];
        if (jobs) { pj[r] = v->pod_job[p]; pr[r] = v->pod_role[p]; pc[r] = v->pod_created[p]; }
        const uint32_t n = v->pod_ctr_begin[p + 1] - v->pod_ctr_begin[p];
        if (n) {
            std::memcpy(cv + cb[r], v->ctr_val + v->pod_ctr_begin[p], (size_t)n * 8);
            std::memcpy(cs + cb[r], v->ctr_scale + v->pod_ctr_begin[p], n);
        }
    }
    if (N) {
        std::memcpy(h->c_node_cap_val.ptr, v->node_cap_val, N * 8);
        std::memcpy(h->c_node_cap_scale.ptr, v->node_cap_scale, N);
    }
    h->v = *v;
    h->v.pod_node = pn; h->v.pod_phase = ph; h->v.pod_ctr_begin = cb; h->v.ctr_val = cv; h->v.ctr_scale = cs;
    h->v.node_cap_val = h->c_node_cap_val.as<int64_t>(); h->v.node_cap_scale = h->c_node_cap_scale.as<int8_t>();
    if (jobs) { h->v.pod_job = pj; h->v.pod_role = pr; h->v.pod_created = pc; }
    else { h->v.pod_job = nullptr; h->v.pod_role = nullptr; h->v.pod_created = nullptr; h->v.n_jobs = 0; }
    h->pod_order = std::move(row_of);                            // (row -> source row needs the inverse)
    std::vector<uint32_t> inv(P);
    for (uint64_t p = 0; p < P; p++) inv[h->pod_order[p]] = (uint32_t)p;
    h->pod_order = std::move(inv);
    *out = h;
    return ARENA_OK;
}

}  // extern "C"
