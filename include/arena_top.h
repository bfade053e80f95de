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
