/*
 * arena_gpuacct.h -- C ABI of libarena_gpuacct.so: the GPU resource accounting
 * behind `arena top node` / `arena top job`, on B200 (sm_100a).
 *
 * The reference (a Go CLI, read-only at /root/reference) has no FFI for this
 * path; its seams are unexported Go functions in package `commands`.  Each
 * entry point below names the Go seam it stands in for (paths relative to
 * /root/reference).  INTEGRATION.md shows the cgo binding a maintainer adds.
 *
 * Conventions
 *   - plain C types only; every function returns 0 or a negative arena_status;
 *     arena_last_error() gives the message for the calling thread.
 *   - the caller owns every host buffer it passes; the library keeps no host
 *     pointer after a call returns (cgo pointer-passing rule).
 *   - arena_acct_init binds the process to one GPU (one rank per GPU under
 *     torchrun / one goroutine locked to an OS thread); a process that wants
 *     every GPU of the box uses arena_acct_init_devices and the *_multi entry
 *     points instead.  Every entry point makes its device current itself.
 *   - one query in flight per snapshot handle.
 *   - int64 sums wrap modulo 2^64 exactly like Go's `+=`.
 *   - all columns must be 16-byte aligned (cudaMalloc / malloc'd arrays are).
 */
#ifndef ARENA_GPUACCT_H
#define ARENA_GPUACCT_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef enum arena_status {
    ARENA_OK = 0,
    ARENA_E_INVALID = -1,    /* bad argument (NULL, misaligned, inconsistent counts) */
    ARENA_E_CUDA = -2,       /* CUDA runtime error; text in arena_last_error() */
    ARENA_E_NOT_INIT = -3,   /* arena_acct_init() not called for this process */
    ARENA_E_NOMEM = -4,
    ARENA_E_TOO_LARGE = -5,  /* a size exceeds what the entry point supports */
    ARENA_E_PARSE = -6       /* loader: malformed JSON / quantity (arena_top.h) */
} arena_status;

/* spec.nodeName / job id that matches no row ("" or a node absent from the
 * NodeList: such pods add to no row and to no total, top_node.go:88-97). */
#define ARENA_NONE 0xFFFFFFFFu

/* status.phase codes (vendor/k8s.io/api/core/v1/types.go:2339-2358).  Active
 * <=> code != SUCCEEDED && code != FAILED, the selector of pod_helper.go:18. */
enum {
    ARENA_PHASE_EMPTY = 0, ARENA_PHASE_PENDING = 1, ARENA_PHASE_RUNNING = 2,
    ARENA_PHASE_SUCCEEDED = 3, ARENA_PHASE_FAILED = 4, ARENA_PHASE_UNKNOWN = 5,
    ARENA_PHASE_OTHER = 255
};

/* pod_role for `top job` (trainer_mpi.go:241-282, trainer_standalone.go:198-221,
 * trainer_tensorflow.go:331-344) */
enum {
    ARENA_ROLE_COUNTS = 0,   /* always summed (tfjob pods; horovod pods not owned by a Job) */
    ARENA_ROLE_LATEST = 1    /* only the newest such pod of the job is summed */
};

/*
 * Columnar snapshot (struct of arrays).  P pods, C containers, N nodes, J jobs.
 * A limit / capacity is the reference's int64Amount{value, scale}
 * (resource/amount.go:57-60) of the SELECTED key: "nvidia.com/gpu" when that
 * key is present, else "alpha.kubernetes.io/nvidia-gpu", else {0,0}
 * (gpu.go:18-37, 61-79).  Slow-path quantities (inf.Dec) are resolved by the
 * loader to {Value(), 0}.  Scale is in [-18, 18] after the loader canonicalises
 * (scale > 18 can only yield 0).
 *
 * Fast when pods that share a node (job) are adjacent; correct in any order.
 */
typedef struct arena_snapshot_view {
    uint64_t n_pods;                 /* P */
    uint64_t n_containers;           /* C == pod_ctr_begin[P] */
    uint64_t n_nodes;                /* N */
    uint64_t n_jobs;                 /* J (0 when the job columns are NULL) */
    const uint32_t *pod_node;        /* [P]   dictionary id of spec.nodeName, or ARENA_NONE */
    const uint8_t  *pod_phase;       /* [P]   ARENA_PHASE_* */
    const uint32_t *pod_ctr_begin;   /* [P+1] CSR offsets into ctr_* (spec.containers only) */
    const int64_t  *ctr_val;         /* [C]   int64Amount.value */
    const int8_t   *ctr_scale;       /* [C]   int64Amount.scale */
    const int64_t  *node_cap_val;    /* [N]   status.capacity, same encoding */
    const int8_t   *node_cap_scale;  /* [N] */
    const uint32_t *pod_job;         /* [P]   job id or ARENA_NONE   (top job; may be NULL) */
    const uint8_t  *pod_role;        /* [P]   ARENA_ROLE_*           (top job; may be NULL) */
    const int64_t  *pod_created;     /* [P]   creationTimestamp, unix seconds (top job) */
} arena_snapshot_view;

typedef struct arena_snapshot arena_snapshot;   /* opaque; device-resident columns */

/* Process set-up / tear-down.  Stands in for setupKubeconfig+initKubeClient
 * (cmd/arena/commands/common.go:29-66): the one-time acquisition of the
 * backend a `top` command talks to.  Idempotent; `device` is a CUDA ordinal. */
int  arena_acct_init(int device);
void arena_acct_shutdown(void);
int  arena_acct_device(void);                     /* ordinal, or ARENA_E_NOT_INIT */
int  arena_acct_sm_count(void);

/* Thread-local, static storage; never NULL. */
const char *arena_last_error(void);

/* Replaces the two List calls that materialise []v1.Pod / []v1.Node
 * (pod_helper.go:15-43, top_node.go:71): copy host columns to HBM. */
int  arena_snapshot_upload(const arena_snapshot_view *host, arena_snapshot **out);
/* Same, but `dev` already holds DEVICE pointers (e.g. torch tensors); nothing
 * is copied and the caller keeps ownership of the columns. */
int  arena_snapshot_wrap_device(const arena_snapshot_view *dev, arena_snapshot **out);
/* Stream-ordered on the library's own stream (freed blocks are cached for the next snapshot and no
 * other stream of the process is stalled).  Queries the caller enqueued on a stream of ITS OWN
 * (the `stream` argument of the *_device entry points) must have completed before the handle is
 * freed. */
void arena_snapshot_free(arena_snapshot *s);
/* counts_out = { n_pods, n_containers, n_nodes, n_jobs } of a device snapshot. */
int  arena_snapshot_counts(const arena_snapshot *s, uint64_t counts_out[4]);
/* Copies the device columns back into the HOST buffers `host_out` points to (sized from
 * arena_snapshot_counts; NULL pointers are skipped).  For inspection and tests: no
 * reference counterpart. */
int  arena_snapshot_download(const arena_snapshot *s, const arena_snapshot_view *host_out);

/*
 * `arena top node`: NodeDescriber.getAllNodeInfos + caculateNodeGPU + the
 * cluster roll-up (top_node.go:69-97, 232-242, 109-118), with gpuInNode /
 * gpuInPod / gpuInContainer (gpu.go:18-79) and Quantity.Value()
 * (resource/quantity.go:699-718) inside.
 *   node_total[N], node_alloc[N]  per-node table (HOST buffers)
 *   cluster[2]                    {sum node_total, sum node_alloc}
 *   pod_gpu[P] or NULL            gpuInPod of every pod (details view,
 *                                 top_node.go:186-195; unmasked by phase)
 */
int  arena_top_node(const arena_snapshot *s, int64_t *node_total, int64_t *node_alloc,
                    int64_t cluster[2], int64_t *pod_gpu);

/* Same computation, DEVICE output buffers, enqueued on `stream` (a
 * cudaStream_t; NULL = the library's own stream, so pass cudaStreamLegacy to
 * mean the legacy default stream) without synchronising.  Used
 * to time the kernels alone and to feed d_cluster to an NCCL all-reduce. */
int  arena_top_node_device(const arena_snapshot *s, int64_t *d_node_total, int64_t *d_node_alloc,
                           int64_t *d_cluster, int64_t *d_pod_gpu, void *stream);

/* One-shot form for callers holding HOST columns: upload, compute, download.
 * This is the end-to-end call bench.py times as `e2e`. */
int  arena_top_node_host(const arena_snapshot_view *host, int64_t *node_total,
                         int64_t *node_alloc, int64_t cluster[2], int64_t *pod_gpu);

/*
 * Multi-GPU `top node` without a collective launch (one process per GPU, one node range per
 * process).  Each rank creates a mailbox in its own HBM and publishes its 64-byte CUDA IPC
 * handle (arena_peer_exchange_create); the ranks trade handles by any means they have (the
 * torchrun driver gathers them as Python objects) and map each other's mailboxes
 * (arena_peer_exchange_connect: `handles` = world x 64 bytes in rank order).  After that
 * arena_top_node_device_fused is arena_top_node_device whose kernel also trades the two
 * cluster totals with every peer over NVLink in its last CTA -- peer stores, an epoch flag,
 * a bounded spin (~2 s; after that arena_peer_exchange_timed_out() != 0, the call itself has long
 * returned ARENA_OK, and both entries of d_cluster hold INT64_MIN instead of a partial sum) -- so
 * d_cluster receives the GLOBAL {total, allocated} and no all-reduce is launched.
 * Every rank must make the same sequence of fused calls.  At most 16 ranks.  Replaces the
 * single-process loop over all nodes of top_node.go:109-118; no reference counterpart for
 * the exchange itself.
 */
int  arena_peer_exchange_create(int rank, int world, unsigned char ipc_handle_out[64]);
int  arena_peer_exchange_connect(const unsigned char *handles);
void arena_peer_exchange_destroy(void);
int  arena_peer_exchange_timed_out(void);
int  arena_top_node_device_fused(const arena_snapshot *s, int64_t *d_node_total, int64_t *d_node_alloc,
                                 int64_t *d_cluster, void *stream);
/*
 * The same launch with the collection deferred by one step, for a caller that issues a SEQUENCE of
 * queries (a poller, a benchmark): the kernel posts this step's totals to every peer and collects
 * the PREVIOUS step's, which every peer posted a whole step ago -- a rank that runs late by less
 * than a step stalls nobody.  d_cluster_prev receives the global totals of the previous posted step
 * (untouched by the first call after arena_peer_exchange_create); arena_peer_exchange_collect
 * enqueues a one-warp kernel that delivers the totals of the most recent step, at the end of the
 * sequence or whenever the caller wants them now.  Same rules as above: every rank makes the same
 * sequence of calls; the per-node slices are complete when the launch is.
 */
int  arena_top_node_device_posted(const arena_snapshot *s, int64_t *d_node_total, int64_t *d_node_alloc,
                                  int64_t *d_cluster_prev, void *stream);
int  arena_peer_exchange_collect(int64_t *d_cluster, void *stream);

/*
 * The same one-shot call for a PACKED host snapshot: the narrowest exact encoding of what
 * `top node` reads, for callers on the far side of PCIe (the copy is the whole cost of
 * arena_top_node_host: 9 bytes per pod + 9 per container; packed it is 1 + 1).
 *
 *   node_pod_begin[N+1]  rows of node n are [b[n], b[n+1]): pods are grouped by node, in
 *                        node order; rows from b[N] on name no node of the NodeList
 *   pod_meta[P]          low nibble: phase (ARENA_PHASE_EMPTY..UNKNOWN = 0..5, 15 = any other
 *                        phase string); high nibble: number of containers (0..15)
 *   ctr_gpu              Quantity.Value() of the container's selected GPU limit (gpu.go:61-79,
 *                        quantity.go:699-718) in pod order: ctr_bits == 8: one byte each, 0..255;
 *                        ctr_bits == 4: two per byte, 0..15, container 2k in the low nibble
 *   node_cap_val/scale   as in arena_snapshot_view
 *
 * A snapshot with a pod of more than 15 containers, a container above 255 GPUs or pods that
 * are not grouped has no packed form: use the wide one (arena_host_snapshot_pack says so).
 * The summary (pod_gpu == NULL) is computed straight from the packed columns, one half-warp per
 * node, and pipelined: a large node range (from 16 Mi pods) is cut into up to eight chunks whose uploads, kernels and
 * downloads overlap (page-locked buffers, arena_pinned_alloc, make the copies asynchronous); with pod_gpu (per row, as in arena_top_node) the columns are first expanded into the
 * wide layout on the device and the ordinary kernels run on that.  No reference counterpart.
 */
typedef struct arena_packed_view {
    uint64_t n_pods, n_containers, n_nodes;
    const uint32_t *node_pod_begin;
    const uint8_t  *pod_meta;
    const uint8_t  *ctr_gpu;
    const int64_t  *node_cap_val;
    const int8_t   *node_cap_scale;
    uint32_t ctr_bits;                /* 8 or 4 */
    /* where each node's containers start in ctr_gpu ([N+1]; [N] = those of the rows that name no
     * node), and the rows of every node above 2048 pods cut into pieces of at most 2048:
     * pieces[4k..4k+3] = { node, first row, end row, first container } */
    const uint32_t *node_ctr_begin;
    const uint32_t *pieces;
    uint64_t n_pieces;
} arena_packed_view;

int  arena_top_node_host_packed(const arena_packed_view *host, int64_t *node_total,
                                int64_t *node_alloc, int64_t cluster[2], int64_t *pod_gpu);

/* The packed columns RESIDENT on the device, for a caller that queries one snapshot many times and
 * can afford to hold it in its narrow form only (0.2 GB instead of 2.3 GB per 1e8 pods): upload once,
 * then arena_top_node_packed_device enqueues the summary kernel on `stream` (NULL = the library's)
 * with DEVICE outputs, like arena_top_node_device.  arena_packed_device_bytes reports the bytes one
 * such launch reads and writes (its own roofline denominator: it reads other columns than the wide
 * kernel does).  No reference counterpart. */
typedef struct arena_packed_device arena_packed_device;
int  arena_snapshot_upload_packed(const arena_packed_view *host, arena_packed_device **out);
void arena_packed_device_free(arena_packed_device *p);
int  arena_packed_device_bytes(const arena_packed_device *p, uint64_t *read_write_bytes);
int  arena_top_node_packed_device(const arena_packed_device *p, int64_t *d_node_total, int64_t *d_node_alloc,
                                  int64_t *d_cluster, void *stream);

/*
 * Single process, several GPUs.  The reference CLI is one process (cmd/arena/main.go:12-28); its
 * `top node` loops over all nodes (top_node.go:109-118) and its `top job` over all releases
 * (top_job.go:53-66).  These entry points split those loops over the devices of one process:
 *
 *   arena_acct_init_devices      binds dev_ids[0..n_dev-1]; dev_ids[0] also becomes the device of the
 *                                single-device API above (as if arena_acct_init(dev_ids[0]) had been called)
 *   arena_multi_snapshot_upload  cuts a HOST snapshot into n_dev contiguous key ranges of about equal
 *                                pod count (arena_shard_plan: node ranges for `top node`, by_job = 0; job-id
 *                                ranges for `top job`, by_job = 1 -- rows must be grouped by that key,
 *                                arena_host_snapshot_regroup) and uploads one range per device
 *   arena_top_node_multi         every device computes its node range at once; the per-node slices are
 *                                disjoint and land side by side in node_total / node_alloc (HOST, [N]); the
 *                                two cluster totals are traded inside the kernels over peer memory (peer
 *                                access, no IPC) when every pair of devices can reach each other, else added
 *                                on the host.  pod_gpu (HOST, [P], rows as uploaded) or NULL
 *   arena_top_job_multi          every device answers `top job` for its job range; per-job tables are
 *                                disjoint slices, totals are added, the per-device orderings are merged
 *                                (requested desc, id asc); job_latest holds rows of the uploaded snapshot
 * Results are bit-identical to the single-device entry points for any number of devices (integer sums).
 * Calls on one multi snapshot must not overlap.  No reference counterpart beyond the loops named above.
 */
typedef struct arena_multi_snapshot arena_multi_snapshot;
int  arena_acct_init_devices(int n_dev, const int *dev_ids);
int  arena_acct_device_count(void);                       /* devices bound by arena_acct_init_devices, else 0 */
int  arena_multi_snapshot_upload(const arena_snapshot_view *host, int by_job, arena_multi_snapshot **out);
void arena_multi_snapshot_free(arena_multi_snapshot *m);
/* n_shards, and the key / pod boundaries of the shards ([n_shards + 1] each; NULL pointers are skipped) */
int  arena_multi_snapshot_shards(const arena_multi_snapshot *m, int *n_shards, uint64_t *key_cut, uint64_t *pod_cut);
int  arena_top_node_multi(const arena_multi_snapshot *m, int64_t *node_total, int64_t *node_alloc,
                          int64_t cluster[2], int64_t *pod_gpu);
int  arena_top_job_multi(const arena_multi_snapshot *m, int64_t *job_req, int64_t *job_alloc,
                         uint32_t *job_order, uint32_t *job_latest, int64_t totals[2]);

/*
 * `arena top job`: TrainingJob.RequestedGPU / AllocatedGPU (job_info.go:51-70,
 * trainer_tensorflow.go:135-155) for every job, the "latest pod wins" rule of
 * trainer_mpi.go:263-278 / trainer_standalone.go:198-221, the ordering of
 * makeTrainingJobOrderdByGPUCount (trainer.go:25-37, 68-75; ties, which
 * sort.Sort leaves unspecified, broken by ascending job id) and the totals of
 * list.go:102-103.
 *   job_req[J], job_alloc[J]   per-job sums (HOST)
 *   job_order[J]               job ids, requested desc then id asc
 *   job_latest[J] or NULL      index of the winning ARENA_ROLE_LATEST pod, or ARENA_NONE
 *   totals[2]                  {sum job_req, sum job_alloc}
 * Row order and the latest-pod rule: among ARENA_ROLE_LATEST pods of a job with the same newest
 * pod_created the one with the LOWEST ROW INDEX wins.  The reference takes the first in LIST order
 * (`latest.Before(current)` is strict), so a caller that builds the columns itself either keeps
 * such pods in list order or settles the tie beforehand, as arena_load_json does when it regroups
 * rows by node (it moves the later pods of a tie one second back).  One query in flight per handle
 * (a `top node` and a `top job` query on the same handle may overlap; two of the same kind may not).
 */
int  arena_top_job(const arena_snapshot *s, int64_t *job_req, int64_t *job_alloc,
                   uint32_t *job_order, uint32_t *job_latest, int64_t totals[2]);
int  arena_top_job_device(const arena_snapshot *s, int64_t *d_job_req, int64_t *d_job_alloc,
                          uint32_t *d_job_order, uint32_t *d_job_latest, int64_t *d_totals,
                          void *stream);

/*
 * `top job` without any allocation.  (No counterpart in the reference: a Go
 * program never manages device memory.)  For callers that cannot call
 * cudaMalloc on the query path -- an agent polling next to training jobs on a
 * fixed device-memory budget, a host whose CUDA context forbids or hooks
 * allocation after start-up, a caller for whom an allocation's implicit device
 * synchronisation is unacceptable.  Same results as arena_top_job_device, but:
 *   - `dev` holds DEVICE column pointers (no snapshot handle: making one allocates); only the
 *     columns `top job` reads are required (pod_node and the node capacities may be NULL), rows
 *     in the order described at arena_top_job;
 *   - the work arrays live in a fixed ~105 KiB device array inside the library,
 *     so at most ARENA_NOALLOC_MAX_JOBS jobs; more -> ARENA_E_TOO_LARGE;
 *   - d_job_latest is required (it doubles as a work array);
 *   - the first call on a device resolves the array's device address: make
 *     that call before beginning any stream capture;
 *   - all calls share the one array: calls that overlap on the GPU race.
 */
#define ARENA_NOALLOC_MAX_JOBS 2048
int  arena_top_job_device_noalloc(const arena_snapshot_view *dev, int64_t *d_job_req, int64_t *d_job_alloc,
                                  uint32_t *d_job_order, uint32_t *d_job_latest, int64_t *d_totals,
                                  void *stream);

/* Quantity.Value() of n int64Amount pairs on the GPU (parity probe for
 * resource/amount.go:95-101 + math.go:144-203); HOST in, HOST out. */
int  arena_quantity_values(const int64_t *val, const int8_t *scale, uint64_t n, int64_t *out);

/* Per-kernel device timing for the roofline report (no counterpart in the
 * reference; its only profiler is --pprof, cmd/arena/main.go:14-22).  Between
 * begin and end, every arena_top_node*() call records CUDA events on its
 * launch stream around the node_alloc memset and the top-node kernel (at most
 * max_calls calls are recorded).  end() waits for them and returns the summed
 * durations in milliseconds. */
int  arena_profile_begin(int max_calls);
/* The same, recording every `every`-th call only: the three event records of a profiled call cost
 * about 10 us of stream time, which is most of what a 65 us step of a small shard can spare. */
int  arena_profile_begin_sampled(int max_calls, int every);
int  arena_profile_end(double *memset_ms_sum, double *kernel_ms_sum, uint64_t *calls);

/* Calibration (no counterpart in the reference): read the five columns the
 * pods kernel reads with plain coalesced 128-bit loads and no arithmetic,
 * `iters` times; *gbs_out = bytes read / time.  The ceiling a read-only
 * streaming kernel reaches on this GPU, reported next to the roofline. */
int  arena_calibrate_stream_read(const arena_snapshot *s, int iters, double *gbs_out, double *ms_out);
/* Same, but with the pods kernel's access pattern (five columns at once, per
 * 128-pod tile, grid-stride) and no dependent loads: what that pattern can
 * reach before any arithmetic.  *gbs_out counts the bytes requested. */
int  arena_calibrate_stream_pattern(const arena_snapshot *s, int iters, double *gbs_out, double *ms_out);

/* Number of kernels this library has launched in this process (bench.py's
 * gpu_launches claim). */
uint64_t arena_kernel_launches(void);

#ifdef __cplusplus
}
#endif
#endif /* ARENA_GPUACCT_H */
