/*
 * ORACLE -- test infrastructure only.  Not part of the product, never linked
 * into libarena_gpuacct.so.  Only tests/, __graft_entry__.smoke() and
 * bench.py's cpu_baseline / --impl reference legs may load this library.
 *
 * Plain-C restatement of the reference's GPU accounting arithmetic
 * (paths relative to /root/reference):
 *
 *   oracle_quantity_value     vendor/k8s.io/apimachinery/pkg/api/resource/
 *                             amount.go:95-101 (AsScaledInt64 with scale 0),
 *                             math.go:92-140 (int64MultiplyScale*),
 *                             math.go:144-168 (positiveScaleInt64),
 *                             math.go:173-203 (negativeScaleInt64);
 *                             the dropped `ok` of quantity.go:713-714
 *   oracle_top_node_columnar  cmd/arena/commands/gpu.go:39-46 (gpuInPod),
 *                             gpu.go:18-37 (gpuInNode),
 *                             pod_helper.go:15-31 (active-pod selector:
 *                             phase != Succeeded && phase != Failed),
 *                             top_node.go:69-97 (join pod.nodeName == node.name),
 *                             top_node.go:232-242 (caculateNodeGPU),
 *                             top_node.go:109-118 (cluster totals)
 *   oracle_top_node_aos_*     the same over array-of-structs inputs with string
 *                             node names and keyed limits (gpu.go:61-79 lookup
 *                             order: primary key if PRESENT, else deprecated key)
 *   oracle_top_job_columnar   job_info.go:51-70 / trainer_tensorflow.go:135-155
 *                             (RequestedGPU = sum gpuInPod over the job's pods,
 *                             AllocatedGPU = sum gpuInActivePod, gpu.go:48-59),
 *                             trainer_mpi.go:241-282 and
 *                             trainer_standalone.go:198-221 (latest Job-owned /
 *                             latest pod wins, strict Before => first in list
 *                             wins ties), list.go:102-103 (totals)
 *
 * Parity status: the reference has no tests or golden vectors for this path
 * and cannot be compiled here (Go, no toolchain): "parity unpinned" by the
 * reference; pinned by the hand-derived vectors under tests/golden/ and, for
 * quantity_value, by the worked examples of the reference's own doc comments
 * (tests/test_reference_examples.py).
 *
 * All int64 accumulation wraps like Go's `+=`: done in uint64_t here because
 * signed overflow is undefined in C.
 */
#include <stdint.h>
#include <stddef.h>
#include <stdlib.h>
#include <string.h>

#define PHASE_SUCCEEDED 3
#define PHASE_FAILED    4
#define NODE_NONE       0xFFFFFFFFu

/* ---- math.go:92-140 ---------------------------------------------------- */
static int mul_scale(int64_t a, int64_t b, int64_t *out)
{
    if (a == 0 || a == 1) { *out = a * b; return 1; }
    if (a == INT64_MIN && b != 1) { *out = 0; return 0; }
    int64_t c = (int64_t)((uint64_t)a * (uint64_t)b);
    *out = c;
    return c / b == a;
}

/* ---- math.go:144-168 ---------------------------------------------------- */
static int positive_scale_int64(int64_t base, int32_t scale, int64_t *out)
{
    switch (scale) {
    case 0: *out = base; return 1;
    case 1: return mul_scale(base, 10, out);
    case 2: return mul_scale(base, 100, out);
    case 3: return mul_scale(base, 1000, out);
    case 6: return mul_scale(base, 1000000, out);
    case 9: return mul_scale(base, 1000000000, out);
    default: {
        int64_t value = base;
        for (int32_t i = 0; i < scale; i++) {
            if (!mul_scale(value, 10, &value)) { *out = 0; return 0; }
        }
        *out = value;
        return 1;
    }
    }
}

/* ---- math.go:173-203 ---------------------------------------------------- */
static int64_t negative_scale_int64(int64_t base, int32_t scale)
{
    if (scale == 0) return base;
    int64_t value = base;
    int fraction = 0;
    for (int32_t i = 0; i < scale; i++) {
        if (!fraction && value % 10 != 0) fraction = 1;
        value = value / 10;
        if (value == 0) {
            if (fraction) return base > 0 ? 1 : -1;
            return 0;
        }
    }
    if (fraction) value += base > 0 ? 1 : -1;
    return value;
}

/* ---- amount.go:95-101 with scale == 0; quantity.go:713-714 drops `ok` ---- */
int64_t oracle_quantity_value(int64_t value, int32_t scale)
{
    if (scale < 0) return negative_scale_int64(value, -scale);
    int64_t r;
    /* `i, _ := q.i.AsScaledInt64(scale)`: the result is used whatever `ok` says.
     * The looped scales return (0,false) on overflow, but the single-multiply
     * scales 1,2,3,6,9 return the WRAPPED product with ok=false (math.go:118-140). */
    (void)positive_scale_int64(value, scale, &r);
    return r;
}

static inline int phase_active(uint8_t ph)
{
    /* pod_helper.go:18 "status.phase!=Succeeded,status.phase!=Failed";
     * fields/selector.go:118-120 is plain string inequality, so every other
     * phase string (including "") is active. */
    return ph != PHASE_SUCCEEDED && ph != PHASE_FAILED;
}

/* gpu.go:39-46 over the columnar container columns */
static inline uint64_t gpu_in_pod(const uint32_t *ctr_begin, const int64_t *ctr_val,
                                  const int8_t *ctr_scale, uint64_t p)
{
    uint64_t s = 0;
    for (uint32_t c = ctr_begin[p]; c < ctr_begin[p + 1]; c++)
        s += (uint64_t)oracle_quantity_value(ctr_val[c], ctr_scale[c]);
    return s;
}

/*
 * B3: one pass over the columnar snapshot.  Does NOT assume the pods are
 * grouped by node: it is the scatter form of the join of top_node.go:88-97
 * (dictionary id equality == string equality of the names the ids stand for).
 * pod_gpu (may be NULL) receives gpuInPod for EVERY pod regardless of phase
 * (details view calls gpuInPod on already-filtered pods; the caller masks).
 */
void oracle_top_node_columnar(uint64_t P, uint64_t N,
                              const uint32_t *pod_node, const uint8_t *pod_phase,
                              const uint32_t *pod_ctr_begin,
                              const int64_t *ctr_val, const int8_t *ctr_scale,
                              const int64_t *node_cap_val, const int8_t *node_cap_scale,
                              int64_t *node_total, int64_t *node_alloc,
                              int64_t cluster[2], int64_t *pod_gpu)
{
    uint64_t tot = 0, alloc = 0;
    for (uint64_t n = 0; n < N; n++) {
        node_total[n] = oracle_quantity_value(node_cap_val[n], node_cap_scale[n]);
        node_alloc[n] = 0;
    }
    for (uint64_t p = 0; p < P; p++) {
        uint64_t g = gpu_in_pod(pod_ctr_begin, ctr_val, ctr_scale, p);
        if (pod_gpu) pod_gpu[p] = (int64_t)g;
        if (!phase_active(pod_phase[p])) continue;      /* filtered server-side in the reference */
        uint32_t n = pod_node[p];
        if (n == NODE_NONE || n >= N) continue;         /* matches no node row */
        node_alloc[n] = (int64_t)((uint64_t)node_alloc[n] + g);
    }
    for (uint64_t n = 0; n < N; n++) {                   /* top_node.go:115-118 */
        tot += (uint64_t)node_total[n];
        alloc += (uint64_t)node_alloc[n];
    }
    cluster[0] = (int64_t)tot;
    cluster[1] = (int64_t)alloc;
}

/*
 * top job roll-up over the columnar snapshot.
 *   pod_job[p]       job id or 0xFFFFFFFF (pod belongs to no listed release)
 *   pod_role[p]      0 = always counts (tfjob pods; horovod non-Job-owned pods)
 *                    1 = "latest wins" candidate (horovod Job-owned pods;
 *                        every pod of a standalone job)
 *   pod_created[p]   creationTimestamp, unix seconds
 * Pods are visited in list order, so `latest.Before(current)` strictness makes
 * the FIRST pod with the maximum timestamp win (trainer_mpi.go:263-278).
 * job_latest (may be NULL) receives the winning pod index or 0xFFFFFFFF.
 */
void oracle_top_job_columnar(uint64_t P, uint64_t J,
                             const uint32_t *pod_job, const uint8_t *pod_role,
                             const int64_t *pod_created, const uint8_t *pod_phase,
                             const uint32_t *pod_ctr_begin,
                             const int64_t *ctr_val, const int8_t *ctr_scale,
                             int64_t *job_req, int64_t *job_alloc,
                             uint32_t *job_latest, int64_t totals[2])
{
    uint32_t *latest = (uint32_t *)malloc((J ? J : 1) * sizeof(uint32_t));
    for (uint64_t j = 0; j < J; j++) { job_req[j] = 0; job_alloc[j] = 0; latest[j] = NODE_NONE; }
    for (uint64_t p = 0; p < P; p++) {
        uint32_t j = pod_job[p];
        if (j == NODE_NONE || j >= J) continue;
        if (pod_role[p] == 1) {
            if (latest[j] == NODE_NONE || pod_created[latest[j]] < pod_created[p])
                latest[j] = (uint32_t)p;
            continue;
        }
        uint64_t g = gpu_in_pod(pod_ctr_begin, ctr_val, ctr_scale, p);
        job_req[j] = (int64_t)((uint64_t)job_req[j] + g);                 /* gpu.go:39-46 */
        if (phase_active(pod_phase[p]))                                    /* gpu.go:48-59 */
            job_alloc[j] = (int64_t)((uint64_t)job_alloc[j] + g);
    }
    uint64_t treq = 0, talloc = 0;
    for (uint64_t j = 0; j < J; j++) {
        if (latest[j] != NODE_NONE) {
            uint64_t p = latest[j];
            uint64_t g = gpu_in_pod(pod_ctr_begin, ctr_val, ctr_scale, p);
            job_req[j] = (int64_t)((uint64_t)job_req[j] + g);
            if (phase_active(pod_phase[p]))
                job_alloc[j] = (int64_t)((uint64_t)job_alloc[j] + g);
        }
        if (job_latest) job_latest[j] = latest[j];
        treq += (uint64_t)job_req[j];                                      /* list.go:102-103 */
        talloc += (uint64_t)job_alloc[j];
    }
    totals[0] = (int64_t)treq;
    totals[1] = (int64_t)talloc;
    free(latest);
}

/* ------------------------------------------------------------------------
 * Array-of-structs restatement: what the reference actually walks.  A pod
 * carries its node NAME, phase STRING and containers whose limits are a small
 * keyed list (stand-in for map[ResourceName]Quantity).
 * ---------------------------------------------------------------------- */
typedef struct {
    const char *key;          /* resource name */
    int64_t value;            /* int64Amount.value */
    int32_t scale;            /* int64Amount.scale */
} oracle_limit;

typedef struct {
    uint32_t n_limits;
    const oracle_limit *limits;
} oracle_container;

typedef struct {
    const char *node_name;    /* spec.nodeName ("" when unscheduled) */
    const char *phase;        /* status.phase */
    uint32_t n_containers;
    const oracle_container *containers;
} oracle_pod;

typedef struct {
    const char *name;
    uint32_t n_capacity;
    const oracle_limit *capacity;
} oracle_node;

static const char KEY_PRIMARY[] = "nvidia.com/gpu";                    /* const.go:7 */
static const char KEY_DEPRECATED[] = "alpha.kubernetes.io/nvidia-gpu"; /* const.go:9 */

static const oracle_limit *lookup(const oracle_limit *l, uint32_t n, const char *key)
{
    const oracle_limit *hit = NULL;
    for (uint32_t i = 0; i < n; i++)
        if (strcmp(l[i].key, key) == 0) hit = &l[i];   /* duplicate JSON keys: last wins */
    return hit;
}

/* gpu.go:61-79 and gpu.go:18-37: presence, not positivity, picks the key */
static int64_t gpu_in_list(const oracle_limit *l, uint32_t n)
{
    const oracle_limit *v = lookup(l, n, KEY_PRIMARY);
    if (!v) v = lookup(l, n, KEY_DEPRECATED);
    if (!v) return 0;
    return oracle_quantity_value(v->value, v->scale);
}

static uint64_t gpu_in_pod_aos(const oracle_pod *pod)
{
    uint64_t s = 0;
    for (uint32_t c = 0; c < pod->n_containers; c++)
        s += (uint64_t)gpu_in_list(pod->containers[c].limits, pod->containers[c].n_limits);
    return s;
}

static int pod_active_aos(const oracle_pod *pod)
{
    return strcmp(pod->phase, "Succeeded") != 0 && strcmp(pod->phase, "Failed") != 0;
}

/* B1: the faithful nested loop, O(N*P) string compares (top_node.go:77-97). */
void oracle_top_node_aos_nested(uint64_t P, const oracle_pod *pods,
                                uint64_t N, const oracle_node *nodes,
                                int64_t *node_total, int64_t *node_alloc, int64_t cluster[2])
{
    uint64_t tot = 0, alloc = 0;
    for (uint64_t n = 0; n < N; n++) {
        uint64_t a = 0;
        for (uint64_t p = 0; p < P; p++) {
            if (!pod_active_aos(&pods[p])) continue;            /* server-side selector */
            if (strcmp(pods[p].node_name, nodes[n].name) == 0)  /* top_node.go:91 */
                a += gpu_in_pod_aos(&pods[p]);
        }
        node_total[n] = gpu_in_list(nodes[n].capacity, nodes[n].n_capacity);
        node_alloc[n] = (int64_t)a;
        tot += (uint64_t)node_total[n];
        alloc += a;
    }
    cluster[0] = (int64_t)tot;
    cluster[1] = (int64_t)alloc;
}

/* B2: same semantics, the join done through a string hash table: O(N+P). */
static uint64_t fnv1a(const char *s)
{
    uint64_t h = 1469598103934665603ull;
    for (; *s; s++) { h ^= (uint8_t)*s; h *= 1099511628211ull; }
    return h;
}

void oracle_top_node_aos_hash(uint64_t P, const oracle_pod *pods,
                              uint64_t N, const oracle_node *nodes,
                              int64_t *node_total, int64_t *node_alloc, int64_t cluster[2])
{
    uint64_t cap = 16;
    while (cap < 2 * N + 1) cap <<= 1;
    /* slot -> LAST node index with that name is not what the reference does:
     * with duplicate node names every duplicate row gets the pods.  Keep a
     * chain so duplicates all receive the sum. */
    int64_t *slot = (int64_t *)malloc(cap * sizeof(int64_t));
    int64_t *next = (int64_t *)malloc((N ? N : 1) * sizeof(int64_t));
    for (uint64_t i = 0; i < cap; i++) slot[i] = -1;
    for (uint64_t n = 0; n < N; n++) {
        node_total[n] = gpu_in_list(nodes[n].capacity, nodes[n].n_capacity);
        node_alloc[n] = 0;
        uint64_t h = fnv1a(nodes[n].name) & (cap - 1);
        /* open addressing on the first node of each distinct name */
        for (;;) {
            if (slot[h] < 0) { slot[h] = (int64_t)n; next[n] = -1; break; }
            if (strcmp(nodes[slot[h]].name, nodes[n].name) == 0) {
                next[n] = next[slot[h]]; next[slot[h]] = (int64_t)n; break;
            }
            h = (h + 1) & (cap - 1);
        }
    }
    for (uint64_t p = 0; p < P; p++) {
        if (!pod_active_aos(&pods[p])) continue;
        uint64_t h = fnv1a(pods[p].node_name) & (cap - 1);
        for (;;) {
            if (slot[h] < 0) break;
            if (strcmp(nodes[slot[h]].name, pods[p].node_name) == 0) {
                uint64_t g = gpu_in_pod_aos(&pods[p]);
                for (int64_t n = slot[h]; n >= 0; n = next[n])
                    node_alloc[n] = (int64_t)((uint64_t)node_alloc[n] + g);
                break;
            }
            h = (h + 1) & (cap - 1);
        }
    }
    uint64_t tot = 0, alloc = 0;
    for (uint64_t n = 0; n < N; n++) { tot += (uint64_t)node_total[n]; alloc += (uint64_t)node_alloc[n]; }
    cluster[0] = (int64_t)tot;
    cluster[1] = (int64_t)alloc;
    free(slot);
    free(next);
}

/* ------------------------------------------------------------------------
 * Helpers used by tests / the CPU baseline to build AoS inputs from a columnar
 * snapshot (node names "node-%07u"), so B1/B2 can be checked against B3 and
 * timed on the same workload.  Everything is allocated in one arena that the
 * caller frees with oracle_aos_free().
 * ---------------------------------------------------------------------- */
typedef struct {
    oracle_pod *pods;
    oracle_node *nodes;
    oracle_container *containers;
    oracle_limit *limits;
    char *names;
    uint64_t P, N;
} oracle_aos;

static const char *PHASE_STR[6] = {"", "Pending", "Running", "Succeeded", "Failed", "Unknown"};

oracle_aos *oracle_aos_from_columnar(uint64_t P, uint64_t N,
                                     const uint32_t *pod_node, const uint8_t *pod_phase,
                                     const uint32_t *pod_ctr_begin,
                                     const int64_t *ctr_val, const int8_t *ctr_scale,
                                     const int64_t *node_cap_val, const int8_t *node_cap_scale)
{
    uint64_t C = pod_ctr_begin[P];
    oracle_aos *a = (oracle_aos *)calloc(1, sizeof(*a));
    a->P = P; a->N = N;
    a->pods = (oracle_pod *)calloc(P ? P : 1, sizeof(oracle_pod));
    a->nodes = (oracle_node *)calloc(N ? N : 1, sizeof(oracle_node));
    a->containers = (oracle_container *)calloc(C ? C : 1, sizeof(oracle_container));
    /* up to two limits per container (see below), one per node */
    a->limits = (oracle_limit *)calloc((2 * C + N) ? (2 * C + N) : 1, sizeof(oracle_limit));
    a->names = (char *)malloc((N + 2) * 16);
    for (uint64_t n = 0; n < N; n++) {
        char *nm = a->names + n * 16;
        /* "node-0000000" */
        memcpy(nm, "node-", 5);
        uint64_t v = n;
        for (int d = 11; d >= 5; d--) { nm[d] = (char)('0' + v % 10); v /= 10; }
        nm[12] = 0;
        a->nodes[n].name = nm;
        a->nodes[n].n_capacity = 1;
        /* every third node advertises its GPUs under the deprecated key only (gpu.go:28-37) */
        a->limits[2 * C + n].key = n % 3 == 2 ? KEY_DEPRECATED : KEY_PRIMARY;
        a->limits[2 * C + n].value = node_cap_val[n];
        a->limits[2 * C + n].scale = node_cap_scale[n];
        a->nodes[n].capacity = &a->limits[2 * C + n];
    }
    char *empty = a->names + N * 16; empty[0] = 0;
    char *ghost = a->names + (N + 1) * 16; memcpy(ghost, "ghost-node", 11);
    /* The columnar form carries the limit of the SELECTED key.  The structs spread it over the spellings
     * the selection rule has to see through (gpu.go:61-79: presence of the primary key decides):
     *   c % 8 == 5   only the deprecated key carries the limit
     *   c % 8 == 6   both keys: the primary one carries the limit, the deprecated one a decoy (value + 1)
     *   c % 8 == 7   an unrelated key ("cpu") in front of the primary one
     *   otherwise    the primary key alone */
    static const char KEY_CPU[] = "cpu";
    for (uint64_t c = 0; c < C; c++) {
        oracle_limit *l = &a->limits[2 * c];
        const unsigned shape = (unsigned)(c % 8);
        l[0].value = ctr_val[c]; l[0].scale = ctr_scale[c];
        a->containers[c].limits = l;
        a->containers[c].n_limits = 1;
        if (shape == 5) {
            l[0].key = KEY_DEPRECATED;
        } else if (shape == 6) {
            l[0].key = KEY_DEPRECATED; l[0].value = ctr_val[c] + 1;
            l[1].key = KEY_PRIMARY; l[1].value = ctr_val[c]; l[1].scale = ctr_scale[c];
            a->containers[c].n_limits = 2;
        } else if (shape == 7) {
            l[0].key = KEY_CPU; l[0].value = 64; l[0].scale = 0;
            l[1].key = KEY_PRIMARY; l[1].value = ctr_val[c]; l[1].scale = ctr_scale[c];
            a->containers[c].n_limits = 2;
        } else {
            l[0].key = KEY_PRIMARY;
        }
    }
    for (uint64_t p = 0; p < P; p++) {
        uint32_t n = pod_node[p];
        a->pods[p].node_name = n == NODE_NONE ? empty : (n < N ? a->nodes[n].name : ghost);
        a->pods[p].phase = pod_phase[p] < 6 ? PHASE_STR[pod_phase[p]] : "Other";
        a->pods[p].n_containers = pod_ctr_begin[p + 1] - pod_ctr_begin[p];
        a->pods[p].containers = &a->containers[pod_ctr_begin[p]];
    }
    return a;
}

void oracle_aos_run(const oracle_aos *a, int nested,
                    int64_t *node_total, int64_t *node_alloc, int64_t cluster[2])
{
    if (nested) oracle_top_node_aos_nested(a->P, a->pods, a->N, a->nodes, node_total, node_alloc, cluster);
    else oracle_top_node_aos_hash(a->P, a->pods, a->N, a->nodes, node_total, node_alloc, cluster);
}

void oracle_aos_free(oracle_aos *a)
{
    if (!a) return;
    free(a->pods); free(a->nodes); free(a->containers); free(a->limits); free(a->names);
    free(a);
}

/* ------------------------------------------------------------------------
 * All-cores variants for the CPU baseline (bench.py --impl reference and the
 * cpu_baseline object).  The reference itself is single-goroutine
 * (top_node.go:31-50); these give it every host thread so that the GPU is
 * compared against the strongest CPU form of the same algorithm.  Pod ranges
 * are disjoint per thread; a node whose pods straddle two ranges is updated
 * with an atomic add, so the result is identical to the serial one.
 * ---------------------------------------------------------------------- */
#include <pthread.h>

typedef struct {
    uint64_t p_lo, p_hi, N;
    const uint32_t *pod_node; const uint8_t *pod_phase; const uint32_t *pod_ctr_begin;
    const int64_t *ctr_val; const int8_t *ctr_scale;
    int64_t *node_alloc;
} col_job;

static void *col_worker(void *arg)
{
    col_job *j = (col_job *)arg;
    for (uint64_t p = j->p_lo; p < j->p_hi; p++) {
        if (!phase_active(j->pod_phase[p])) continue;
        uint32_t n = j->pod_node[p];
        if (n == NODE_NONE || n >= j->N) continue;
        uint64_t g = gpu_in_pod(j->pod_ctr_begin, j->ctr_val, j->ctr_scale, p);
        __atomic_fetch_add((uint64_t *)&j->node_alloc[n], g, __ATOMIC_RELAXED);
    }
    return NULL;
}

void oracle_threads_top_node_columnar(int threads, uint64_t P, uint64_t N,
                                      const uint32_t *pod_node, const uint8_t *pod_phase,
                                      const uint32_t *pod_ctr_begin,
                                      const int64_t *ctr_val, const int8_t *ctr_scale,
                                      const int64_t *node_cap_val, const int8_t *node_cap_scale,
                                      int64_t *node_total, int64_t *node_alloc, int64_t cluster[2])
{
    if (threads < 1) threads = 1;
    if (threads > 256) threads = 256;
    for (uint64_t n = 0; n < N; n++) {
        node_total[n] = oracle_quantity_value(node_cap_val[n], node_cap_scale[n]);
        node_alloc[n] = 0;
    }
    pthread_t th[256];
    col_job jobs[256];
    for (int t = 0; t < threads; t++) {
        jobs[t] = (col_job){P * (uint64_t)t / threads, P * (uint64_t)(t + 1) / threads, N, pod_node, pod_phase,
                            pod_ctr_begin, ctr_val, ctr_scale, node_alloc};
        pthread_create(&th[t], NULL, col_worker, &jobs[t]);
    }
    for (int t = 0; t < threads; t++) pthread_join(th[t], NULL);
    uint64_t tot = 0, alloc = 0;
    for (uint64_t n = 0; n < N; n++) { tot += (uint64_t)node_total[n]; alloc += (uint64_t)node_alloc[n]; }
    cluster[0] = (int64_t)tot;
    cluster[1] = (int64_t)alloc;
}

typedef struct {
    uint64_t p_lo, p_hi, cap;
    const oracle_aos *a;
    const int64_t *slot, *next;
    int64_t *node_alloc;
} aos_job;

static void *aos_worker(void *arg)
{
    aos_job *j = (aos_job *)arg;
    const oracle_pod *pods = j->a->pods;
    const oracle_node *nodes = j->a->nodes;
    for (uint64_t p = j->p_lo; p < j->p_hi; p++) {
        if (!pod_active_aos(&pods[p])) continue;
        uint64_t h = fnv1a(pods[p].node_name) & (j->cap - 1);
        for (;;) {
            if (j->slot[h] < 0) break;
            if (strcmp(nodes[j->slot[h]].name, pods[p].node_name) == 0) {
                uint64_t g = gpu_in_pod_aos(&pods[p]);
                for (int64_t n = j->slot[h]; n >= 0; n = j->next[n])
                    __atomic_fetch_add((uint64_t *)&j->node_alloc[n], g, __ATOMIC_RELAXED);
                break;
            }
            h = (h + 1) & (j->cap - 1);
        }
    }
    return NULL;
}

void oracle_aos_run_threads(const oracle_aos *a, int threads,
                            int64_t *node_total, int64_t *node_alloc, int64_t cluster[2])
{
    if (threads < 1) threads = 1;
    if (threads > 256) threads = 256;
    const uint64_t N = a->N, P = a->P;
    uint64_t cap = 16;
    while (cap < 2 * N + 1) cap <<= 1;
    int64_t *slot = (int64_t *)malloc(cap * sizeof(int64_t));
    int64_t *next = (int64_t *)malloc((N ? N : 1) * sizeof(int64_t));
    for (uint64_t i = 0; i < cap; i++) slot[i] = -1;
    for (uint64_t n = 0; n < N; n++) {
        node_total[n] = gpu_in_list(a->nodes[n].capacity, a->nodes[n].n_capacity);
        node_alloc[n] = 0;
        uint64_t h = fnv1a(a->nodes[n].name) & (cap - 1);
        for (;;) {
            if (slot[h] < 0) { slot[h] = (int64_t)n; next[n] = -1; break; }
            if (strcmp(a->nodes[slot[h]].name, a->nodes[n].name) == 0) {
                next[n] = next[slot[h]]; next[slot[h]] = (int64_t)n; break;
            }
            h = (h + 1) & (cap - 1);
        }
    }
    pthread_t th[256];
    aos_job jobs[256];
    for (int t = 0; t < threads; t++) {
        jobs[t] = (aos_job){P * (uint64_t)t / threads, P * (uint64_t)(t + 1) / threads, cap, a, slot, next, node_alloc};
        pthread_create(&th[t], NULL, aos_worker, &jobs[t]);
    }
    for (int t = 0; t < threads; t++) pthread_join(th[t], NULL);
    uint64_t tot = 0, alloc = 0;
    for (uint64_t n = 0; n < N; n++) { tot += (uint64_t)node_total[n]; alloc += (uint64_t)node_alloc[n]; }
    cluster[0] = (int64_t)tot;
    cluster[1] = (int64_t)alloc;
    free(slot);
    free(next);
}
