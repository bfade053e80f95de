/*
 * arena_top.h -- host side of libarena_gpuacct.so: everything around the
 * kernels that the reference does in Go on the CPU for `arena top node` /
 * `arena top job`, rebuilt as C++ behind a C ABI (no Go toolchain exists in
 * the build image; INTEGRATION.md shows the cgo stubs).
 *
 *   loader     PodList / NodeList / JobList / TFJobList JSON + `helm list`
 *              text  ->  columnar snapshot.  Replaces client-go's List + JSON
 *              decode (cmd/arena/commands/pod_helper.go:15-55, top_node.go:71,
 *              util/helm/helm.go:164-195) and Quantity.UnmarshalJSON
 *              (vendor/k8s.io/apimachinery/pkg/api/resource/quantity.go:648-667).
 *   printers   displayTopNodeSummary / displayTopNodeDetails
 *              (top_node.go:107-229) and displayTrainingJobList (list.go:82-133)
 *              through a text/tabwriter work-alike.
 *   synth      seeded columnar generators for the large configurations.
 *   shard      node-range (or job-range) partition for multi-GPU runs.
 *
 * Same conventions as arena_gpuacct.h: 0 or a negative arena_status,
 * arena_last_error() for the text, caller-owned host buffers.
 */
#ifndef ARENA_TOP_H
#define ARENA_TOP_H

#include "arena_gpuacct.h"

#ifdef __cplusplus
extern "C" {
#endif

/* Host-resident snapshot: the columns plus the names the printers need. */
typedef struct arena_host_snapshot arena_host_snapshot;

void arena_host_snapshot_free(arena_host_snapshot *h);
/* Fills `out` with HOST pointers into `h` (valid until h is freed). */
int  arena_host_snapshot_view(const arena_host_snapshot *h, arena_snapshot_view *out);

/* ---- loader ------------------------------------------------------------- */
/*
 * podlist_json / nodelist_json: a v1.PodList / v1.NodeList as the API server
 * returns them (vendor/k8s.io/api/core/v1/types.go:3075, 4030).  Required.
 * joblist_json (batch/v1 JobList), tfjoblist_json (kubeflow.org/v1alpha2
 * TFJobList) and helm_list_text (stdout of `helm list`) are needed for
 * `top job` only and may be NULL / 0.
 *
 * Pods are re-ordered so that pods of one node are adjacent (stable: list
 * order is kept inside a node; pods that match no node go last), which is the
 * grouping the kernels are fastest on; arena_host_pod_order() gives the
 * original list index of each row.  With group_by_job != 0 (use it for `top
 * job`) the rows are grouped by job instead, again stably, so that "first pod
 * in list order wins a timestamp tie" (trainer_mpi.go:263-278) is "lowest row
 * wins" inside every job.  A malformed quantity fails the whole load
 * with ARENA_E_PARSE, as a failed List decode does in the reference
 * (top_node.go:39-42).
 */
int  arena_load_json(const char *podlist_json, size_t podlist_len,
                     const char *nodelist_json, size_t nodelist_len,
                     const char *joblist_json, size_t joblist_len,
                     const char *tfjoblist_json, size_t tfjoblist_len,
                     const char *helm_list_text, size_t helm_len,
                     int group_by_job, arena_host_snapshot **out);

/*
 * The same wire-format step with the PodList parsed ON THE GPU (`top node` columns only).
 * The document is copied to the device; a structural scan there (quote / bracket bitmasks,
 * 64 bytes per thread, prefix-composed across the grid) finds every element of .items, and
 * one thread per pod validates its item and extracts spec.nodeName (joined to the NodeList
 * by an exact hash lookup), status.phase and the GPU limit of every container.  Replaces the
 * same reference code as arena_load_json (pod_helper.go:15-43, json.go:171,
 * quantity.go:648-667) for the fields A1 of SURVEY.md section 8 names.
 *
 *   nodes_out      host snapshot of the NodeList (names, addresses, capacities: what the
 *                  printers need); it has ZERO pods unless the host loader had to take over
 *   snapshot_out   device snapshot, rows in PodList order (the accounting kernels are exact
 *                  for any order and pick the ungrouped variant when it is faster)
 *
 * A pod item that holds something the device scanner does not decide itself -- escaped or
 * duplicate keys, a GPU quantity off the reference's fast path (resource/quantity.go:276-343),
 * a timestamp that is not in the API server's canonical form, more than 8 containers -- is
 * handed back ALONE: the host loader's item code parses it and its row is patched into the
 * device columns (stats->n_handed_back).  Malformed text, an item the host rejects, or more
 * than an eighth of the items flagged make the whole document go through arena_load_json +
 * arena_snapshot_upload instead: the columns, the results and the error messages are those
 * of the host loader either way, and stats->fell_back says which path ran.  Needs arena_acct_init(); there is no CPU stand-in
 * for the device (ARENA_E_NOT_INIT / ARENA_E_CUDA).  PodLists of 4 GiB and more take the
 * host path.
 */
typedef struct arena_load_device_stats {
    uint64_t n_items;         /* pods found by the device scan */
    uint64_t flagged_item;    /* first item the device handed back (when fell_back) */
    uint32_t flag_reason;     /* arena_scan::Flag of that item (diagnostic) */
    uint32_t fell_back;       /* 1 = the host loader produced the snapshot */
    /* host clock since the call started, at the points where the host waits for the device */
    double   ms_h2d;          /* document copied to the device */
    double   ms_structure;    /* + structural scan (3 kernels) */
    double   ms_items;        /* + envelope on the host, item ranges, one thread per pod, CSR offsets */
    double   ms_device;       /* + final columns: the whole device path */
    uint64_t n_handed_back;   /* items the device scanner flagged and the host parsed one by one (device path kept) */
} arena_load_device_stats;

/* Page-locked host memory for documents: reading the API server's response straight into
 * such a buffer lets the copy to the device run at full PCIe speed (a pageable buffer is
 * staged by the driver at roughly a third of it). */
int  arena_pinned_alloc(size_t bytes, void **out);
void arena_pinned_free(void *p);

int  arena_load_json_device(const char *podlist_json, size_t podlist_len,
                            const char *nodelist_json, size_t nodelist_len,
                            arena_host_snapshot **nodes_out, arena_snapshot **snapshot_out,
                            arena_load_device_stats *stats /* may be NULL */);

/*
 * arena_load_json with the PodList decoded on the GPU: the SAME host snapshot (columns, names, jobs,
 * row order) as arena_load_json gives for the same five inputs, so `top job`, `top node -d` and
 * `arena list` can use the device loader too.  The device scanner reports, per pod, the accounting
 * columns and the PLACES of metadata.name / namespace / creationTimestamp, the five labels the trainers
 * read (trainer_mpi.go:315-332, trainer_standalone.go:257-273, trainer_tensorflow.go:394-418),
 * ownerReferences[].kind == "Job" and status.hostIP; the host copies those bytes out and runs the job
 * membership rules and the grouping as arena_load_json does.  Replaces pod_helper.go:15-43 +
 * json.go:171 for all three commands.  Items the scanner does not decide (escaped strings, exotic
 * quantities ...) are parsed one by one by the host's item code; a document it cannot take at all goes
 * through arena_load_json as a whole (stats->fell_back), which also owns every error message.
 */
int  arena_load_json_gpu(const char *podlist_json, size_t podlist_len,
                         const char *nodelist_json, size_t nodelist_len,
                         const char *joblist_json, size_t joblist_len,       /* may be NULL */
                         const char *tfjoblist_json, size_t tfjoblist_len,   /* may be NULL */
                         const char *helm_list_text, size_t helm_len,        /* may be NULL */
                         int group_by_job, arena_host_snapshot **out,
                         arena_load_device_stats *stats /* may be NULL */);

/* Diagnostic twin: the same scanner source compiled for the HOST walks the items (no CUDA call), then
 * the same record building and job rules run.  For the CPU test-suite (whole snapshots against
 * arena_load_json without a GPU); not a fallback of anything.  n_handed_back: items that went through
 * the host's item code.  A document the scanner path does not take is an error here. */
int  arena_load_json_hostscan(const char *podlist_json, size_t podlist_len,
                              const char *nodelist_json, size_t nodelist_len,
                              const char *joblist_json, size_t joblist_len,
                              const char *tfjoblist_json, size_t tfjoblist_len,
                              const char *helm_list_text, size_t helm_len,
                              int group_by_job, arena_host_snapshot **out, uint64_t *n_handed_back);

/*
 * Diagnostic: the per-item scanner of arena_load_json_device (csrc/pod_scan.cuh, the same
 * source compiled for the host) over every element of .items, in PodList order.  `nodes`
 * is a snapshot loaded from the NodeList.  Arrays hold `capacity` pods (ctr_val / ctr_scale:
 * 8 entries per pod, the int64Amount{value, scale} of each container's selected GPU limit); pod_flag[k] != 0 means item k would be handed to the host loader.  Used by the
 * CPU test-suite to compare the device code with the host loader without a GPU.
 */
int  arena_scan_podlist_host(const char *podlist_json, size_t podlist_len, const arena_host_snapshot *nodes,
                             uint64_t capacity, uint64_t *n_items_out, uint32_t *pod_node, uint8_t *pod_phase,
                             uint32_t *pod_n_ctrs, int64_t *ctr_val, int8_t *ctr_scale, uint8_t *pod_flag);

/*
 * Packs a host snapshot (arena_snapshot_view with HOST pointers, rows grouped by node as the
 * loader and the generators produce them) into the form arena_top_node_host_packed takes.
 * ARENA_E_INVALID, with the reason in arena_last_error(), when the snapshot has no packed
 * form.  pinned != 0 puts the columns in page-locked memory.
 */
typedef struct arena_packed_snapshot arena_packed_snapshot;
int  arena_host_snapshot_pack(const arena_snapshot_view *host, int pinned, arena_packed_snapshot **out);
int  arena_packed_snapshot_view(const arena_packed_snapshot *p, arena_packed_view *out);
void arena_packed_snapshot_free(arena_packed_snapshot *p);

/* Row -> original PodList index ([P]); job id -> name / trainer type; node id and
 * row -> metadata.name ("" for synthetic snapshots, which carry no names). */
const uint32_t *arena_host_pod_order(const arena_host_snapshot *h);
const char *arena_host_job_name(const arena_host_snapshot *h, uint32_t job);
const char *arena_host_job_trainer(const arena_host_snapshot *h, uint32_t job);
const char *arena_host_node_name(const arena_host_snapshot *h, uint32_t node);
const char *arena_host_pod_name(const arena_host_snapshot *h, uint32_t row);

/* Quantity.UnmarshalJSON + Value() of one raw JSON token (quotes included for
 * strings), on the host.  Returns ARENA_E_PARSE for what the reference rejects. */
int  arena_quantity_parse(const char *raw, size_t len, int64_t *value_out,
                          int64_t *amount_val_out, int32_t *amount_scale_out, int *slow_path_out);

/* ---- printers ------------------------------------------------------------ */
/* Text is malloc'd; release with arena_free_text().  `details` selects
 * displayTopNodeDetails (needs pod_gpu[P]); otherwise pod_gpu may be NULL. */
int  arena_format_top_node(const arena_host_snapshot *h, const int64_t *node_total,
                           const int64_t *node_alloc, const int64_t *pod_gpu, int details,
                           char **text_out, size_t *len_out);
/* AGE / STATUS / NODE are wall-clock and job-object dependent; they are
 * rendered from the loaded Job / TFJob objects with `now_unix` as the clock. */
int  arena_format_top_job(const arena_host_snapshot *h, const int64_t *job_req,
                          const int64_t *job_alloc, const uint32_t *job_order,
                          const uint32_t *job_latest, int64_t now_unix,
                          char **text_out, size_t *len_out);
/* `arena list` (cmd/arena/commands/list.go:16-80): the other user of the same printer.
 * arena_job_order_by_age is makeTrainingJobOrderdByAge (trainer.go:39-66): jobs without
 * status.startTime first, then the newest start first (ties, which sort.Sort leaves open, by
 * ascending job id) -- host work, the keys are job metadata.  arena_format_job_list is
 * displayTrainingJobList(jobs, displayGPU) (list.go:82-133): display_gpu != 0 is the `top job`
 * text above; display_gpu == 0 prints NAME STATUS TRAINER AGE NODE only (job_req / job_alloc may be
 * NULL), but STATUS and NODE still need job_latest -- the latest-pod rule the `top job` kernels
 * evaluate -- for mpijob and standalonejob rows. */
int  arena_job_order_by_age(const arena_host_snapshot *h, uint32_t *job_order);
int  arena_format_job_list(const arena_host_snapshot *h, const int64_t *job_req,
                           const int64_t *job_alloc, const uint32_t *job_order,
                           const uint32_t *job_latest, int64_t now_unix, int display_gpu,
                           char **text_out, size_t *len_out);
void arena_free_text(char *text);

/* ---- synthetic snapshots --------------------------------------------------- */
typedef struct arena_synth_params {
    uint64_t n_pods;          /* P */
    uint64_t n_nodes;         /* N */
    uint64_t n_jobs;          /* J; 0 = no job columns */
    uint64_t seed;
    int      skewed;          /* 0: pods -> nodes uniform (configs 3, 4)
                                 1: Zipf(1.1) pods per node, truncated at 1e6,
                                    10% GPU-sharing nodes (config 5) */
    int      grouped;         /* 1: pods of a node adjacent (canonical); 0: shuffled list order */
    int      pinned;          /* 1: allocate the columns in page-locked memory (needs CUDA) */
    int      threads;         /* worker threads; 0 = hardware concurrency */
} arena_synth_params;

int  arena_synth_snapshot(const arena_synth_params *p, arena_host_snapshot **out);

/* ---- sharding --------------------------------------------------------------- */
/*
 * Cut a GROUPED snapshot into n_shards contiguous node ranges holding about
 * equal pod counts (cuts only at node boundaries; pods that match no node,
 * which sort last, go to the last shard).  node_cut[n_shards+1] and
 * pod_cut[n_shards+1] receive the boundaries.  by_job != 0 cuts on pod_job /
 * n_jobs instead (for `top job`, where jobs span nodes).
 */
int  arena_shard_plan(const arena_snapshot_view *host, int n_shards, int by_job,
                      uint64_t *key_cut, uint64_t *pod_cut);
/* A copy of `host` with its rows regrouped by node id (by_job == 0) or by job id (by_job != 0):
 * stable, so rows of one key keep their order; rows that name no node / job go last.  What
 * arena_shard_plan needs before it can cut by that key (e.g. job ranges for a multi-GPU `top job`
 * from a node-grouped snapshot).  arena_host_pod_order() of the result maps a row to its source
 * row.  The result owns its columns.  No reference counterpart. */
int  arena_host_snapshot_regroup(const arena_snapshot_view *host, int by_job, arena_host_snapshot **out);
/* Shard [key_lo,key_hi) x [pod_lo,pod_hi) as a host snapshot of its own: keys
 * and container offsets are rebased to the shard (keys outside the range
 * become ARENA_NONE); the other columns alias the parent's memory, which must
 * outlive the shard. */
int  arena_shard_extract(const arena_snapshot_view *host, int by_job,
                         uint64_t key_lo, uint64_t key_hi, uint64_t pod_lo, uint64_t pod_hi,
                         arena_host_snapshot **out);

#ifdef __cplusplus
}
#endif
#endif /* ARENA_TOP_H */
