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

// One PodList item through the HOST parser (json.cc + loader.cc extract_pod): what the GPU loader
// calls for the few items its per-pod scanner hands back.  parse() returns false on ANY failure --
// the caller then runs the whole document through the host loader, which owns every error message.
struct PodItemParser {
    explicit PodItemParser(const std::vector<std::string> &node_names);
    ~PodItemParser();
    bool parse(const char *text, size_t text_len, size_t begin, size_t end, uint32_t &node, uint8_t &phase,
               std::vector<int64_t> &ctr_val, std::vector<int8_t> &ctr_scale);
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
