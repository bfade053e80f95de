// Host-resident snapshot shared by the loader, the synthetic generators, the
// sharder and the printers.
#pragma once
#include "../../include/arena_top.h"

#include <cstdint>
#include <string>
#include <vector>

namespace arena_host {

int fail(int code, const char *fmt, ...);   // sets arena_last_error() text (gpuacct.cu owns the buffer)

// 64-byte aligned (optionally page-locked) raw column storage.
struct Column {
    void *ptr = nullptr;
    size_t bytes = 0;
    bool pinned = false;
    Column() = default;
    Column(const Column &) = delete;
    Column &operator=(const Column &) = delete;
    ~Column();
    int alloc(size_t nbytes, bool want_pinned);
    template <class T> T *as() { return static_cast<T *>(ptr); }
};

struct JobMeta {
    std::string name;          // helm release name
    std::string ns;            // namespace from `helm list`
    std::string trainer;       // "mpijob" | "standalonejob" | "tfjob"
    // batch/v1 Job (mpijob, standalonejob): status.active / succeeded / failed
    int64_t active = 0, succeeded = 0, failed = 0;
    // TFJob: GetStatus() resolved from status.conditions at load time
    bool tf_found = false;
    std::string tf_status;     // "SUCCEEDED" | "FAILED" | "PENDING" | "RUNNING"
    uint32_t tf_chief_row = 0xFFFFFFFFu;   // row of the chief pod (worker index 0), or none
    bool start_present = false;   // status.startTime is not nil (what makeTrainingJobOrderdByAge tests)
    bool has_start = false;    // status.startTime present and not the zero time
    int64_t start_unix = 0;
};

// One pod as the loader holds it between the decode and the columns (loader.cc).  The GPU loader
// fills the same record from the places its scanner reports (pod_scan.cuh PodMeta).
enum : uint8_t { kHasRelease = 1, kHasApp = 2, kHasGroup = 4, kHasReplicaType = 8, kHasReplicaIndex = 16 };
struct PodLabels {             // the labels the job-membership rules read
    uint8_t has = 0;
    std::string release, app, group_name, replica_type, replica_index;
    bool is(uint8_t bit, const std::string &field, const char *want) const { return (has & bit) && field == want; }
};
constexpr int64_t kZeroTimeUnix = -62135596800ll;      // Go's time.Time{} (0001-01-01T00:00:00Z)
struct PodRec {
    std::string name, ns, node_name, host_ip;
    uint8_t phase = 0;
    std::vector<int64_t> ctr_val;
    std::vector<int8_t> ctr_scale;
    int64_t created = kZeroTimeUnix;
    bool owner_is_job = false;
    PodLabels labels;
    uint32_t node = 0xFFFFFFFFu, job = 0xFFFFFFFFu;      // ARENA_NONE
    uint8_t role = 0;                                    // ARENA_ROLE_COUNTS
};
// "YYYY-MM-DDTHH:MM:SSZ", already checked for that exact form -> seconds since the epoch
int64_t canonical_time_unix(const char *p);

// Where load_with_pods() gets the decoded pods from instead of parsing the PodList itself: called once,
// after the nodes are known; returns false when it cannot deliver (the caller of load_with_pods then
// takes the plain host path from the start).
struct PodSource {
    virtual ~PodSource() = default;
    virtual bool pods(const std::vector<std::string> &node_names, std::vector<PodRec> &out) = 0;
};
// arena_load_json with the PodList decode replaced by `src`; same snapshot, same errors for every
// other document.  *handled = false: nothing was produced, run arena_load_json.
int load_with_pods(PodSource &src, const char *nodelist, size_t nodelen, const char *joblist, size_t joblen,
                   const char *tfjoblist, size_t tflen, const char *helm, size_t helmlen, bool group_by_job,
                   arena_host_snapshot **out, bool *handled);

// One PodList item through the HOST parser (json.cc + loader.cc extract_pod): what the GPU loader
// calls for the few items its per-pod scanner hands back.  parse() returns false on ANY failure --
// the caller then runs the whole document through the host loader, which owns every error message.
struct PodItemParser {
    explicit PodItemParser(const std::vector<std::string> &node_names);
    ~PodItemParser();
    bool parse(const char *text, size_t text_len, size_t begin, size_t end, uint32_t &node, uint8_t &phase,
               std::vector<int64_t> &ctr_val, std::vector<int8_t> &ctr_scale);
    bool parse_full(const char *text, size_t text_len, size_t begin, size_t end, PodRec &rec);   // every field
    struct Impl;
    Impl *impl;
};

}  // namespace arena_host

struct arena_host_snapshot {
    arena_snapshot_view v{};
    // owned columns (any may be empty when the view borrows from a parent)
    arena_host::Column c_pod_node, c_pod_phase, c_pod_ctr_begin, c_ctr_val, c_ctr_scale, c_node_cap_val,
        c_node_cap_scale, c_pod_job, c_pod_role, c_pod_created;
    // names for the printers; empty for synthetic snapshots ("node-%07u")
    std::vector<std::string> node_names, node_internal_ip, node_first_addr;
    std::vector<uint8_t> node_has_addr, node_is_master;
    std::vector<std::string> pod_names, pod_namespaces, pod_host_ip;   // in ROW order
    std::vector<uint32_t> pod_order;                      // row -> PodList index
    std::vector<arena_host::JobMeta> jobs;
};
