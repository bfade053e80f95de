// Text output of `arena top node` / `arena top job`, byte-compatible with the
// reference printers (paths relative to /root/reference):
//   displayTopNodeSummary   cmd/arena/commands/top_node.go:107-157
//   displayTopNodeDetails   cmd/arena/commands/top_node.go:159-229
//   displayTrainingJobList  cmd/arena/commands/list.go:82-133 (displayGPU = true)
//   ShortHumanDuration      util/duration.go:10-27
// through a work-alike of Go's text/tabwriter configured as at top_node.go:108:
// NewWriter(os.Stdout, minwidth 0, tabwidth 0, padding 2, padchar ' ', flags 0).
// (text/tabwriter is Go standard library, not in the reference tree: its layout
// rule is restated from the package documentation and pinned by the sample
// outputs in docs/userguide/1-tfjob-standalone.md:7-17, 73-98.)
#include "host_snapshot.h"

#include <algorithm>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <vector>

using namespace arena_host;

namespace {

// ---- tabwriter ------------------------------------------------------------------
// Cells are tab-TERMINATED; the trailing cell of a line is not part of any
// column.  A column block is a run of consecutive lines that have a cell in
// that column; its width is the widest cell + padding.
class TabWriter {
public:
    explicit TabWriter(int padding) : padding_(padding) { lines_.emplace_back(); lines_.back().emplace_back(); }

    void write(const std::string &s)
    {
        for (char ch : s) {
            if (ch == '\t') lines_.back().emplace_back();
            else if (ch == '\n') { lines_.emplace_back(); lines_.back().emplace_back(); }
            else lines_.back().back().push_back(ch);
        }
    }

    std::string flush()
    {
        out_.clear();
        widths_.clear();
        // the last line is the unterminated remainder: printed as is (usually empty)
        const size_t n = lines_.size();
        format(0, n - 1);
        const auto &tail = lines_[n - 1];
        for (size_t j = 0; j < tail.size(); j++) out_ += tail[j];   // no newline was written after it
        return out_;
    }

private:
    static size_t rune_count(const std::string &s)
    {
        size_t n = 0;
        for (unsigned char c : s) n += (c & 0xC0) != 0x80;
        return n;
    }

    void write_lines(size_t line0, size_t line1)
    {
        for (size_t i = line0; i < line1; i++) {
            const auto &line = lines_[i];
            for (size_t j = 0; j < line.size(); j++) {
                out_ += line[j];
                if (j < widths_.size()) {
                    const size_t w = rune_count(line[j]);
                    if (widths_[j] > w) out_.append(widths_[j] - w, ' ');
                }
            }
            out_.push_back('\n');
        }
    }

    void format(size_t line0, size_t line1)
    {
        const size_t column = widths_.size();
        for (size_t cur = line0; cur < line1; cur++) {
            if (column + 1 >= lines_[cur].size()) continue;      // no tab-terminated cell in this column
            write_lines(line0, cur);
            line0 = cur;
            size_t width = 0;                                    // minwidth 0
            for (; cur < line1; cur++) {
                const auto &line = lines_[cur];
                if (column + 1 >= line.size()) break;
                const size_t w = rune_count(line[column]) + (size_t)padding_;
                if (w > width) width = w;
            }
            widths_.push_back(width);
            format(line0, cur);
            widths_.pop_back();
            line0 = cur;      // lines_[cur] (if any) has no cell in this column: the loop step skips it
        }
        write_lines(line0, line1);
    }

    int padding_;
    std::vector<std::vector<std::string>> lines_;
    std::vector<size_t> widths_;
    std::string out_;
};

std::string i64s(int64_t v) { return std::to_string(v); }

// int64(float64(a) / float64(b) * 100), top_node.go:146-153 (compiled with -ffp-contract=off)
int64_t percent(int64_t alloc, int64_t total)
{
    double usage = 0;
    if (total > 0) {
        volatile double q = (double)alloc / (double)total;
        usage = q * 100.0;
    }
    return (int64_t)usage;
}

const char kRule[] = "-----------------------------------------------------------------------------------------\n";

std::string node_name(const arena_host_snapshot *h, size_t n)
{
    if (n < h->node_names.size()) return h->node_names[n];
    char buf[32];
    snprintf(buf, sizeof(buf), "node-%07zu", n);                 // synthetic snapshots carry no names
    return buf;
}

int emit(const std::string &text, char **text_out, size_t *len_out)
{
    char *p = (char *)malloc(text.size() + 1);
    if (!p) return fail(ARENA_E_NOMEM, "out of host memory");
    std::memcpy(p, text.data(), text.size());
    p[text.size()] = 0;
    *text_out = p;
    if (len_out) *len_out = text.size();
    return ARENA_OK;
}

// util/duration.go:10-27
std::string short_human_duration(int64_t seconds)
{
    if (seconds < -1) return "<invalid>";
    if (seconds < 0) return "0s";
    if (seconds < 60) return i64s(seconds) + "s";
    const int64_t minutes = seconds / 60;
    if (minutes < 60) return i64s(minutes) + "m";
    const int64_t hours = seconds / 3600;
    if (hours < 24) return i64s(hours) + "h";
    if (hours < 24 * 365) return i64s(hours / 24) + "d";
    return i64s(hours / 24 / 365) + "y";
}

std::string upper(std::string s)
{
    for (char &c : s)
        if (c >= 'a' && c <= 'z') c = (char)(c - 32);
    return s;
}

}  // namespace

extern "C" {

void arena_free_text(char *text) { free(text); }

int arena_format_top_node(const arena_host_snapshot *h, const int64_t *node_total, const int64_t *node_alloc,
                          const int64_t *pod_gpu, int details, char **text_out, size_t *len_out)
{
    if (!h || !text_out) return fail(ARENA_E_INVALID, "NULL argument");
    const size_t N = h->v.n_nodes, P = h->v.n_pods;
    if (N && (!node_total || !node_alloc)) return fail(ARENA_E_INVALID, "node table is NULL");
    if (details && P && !pod_gpu) return fail(ARENA_E_INVALID, "details view needs pod_gpu");
    TabWriter w(2);
    uint64_t total_c = 0, alloc_c = 0;                           // Go int64 += wraps
    if (!details) {
        w.write("NAME\tIPADDRESS\tROLE\tGPU(Total)\tGPU(Allocated)\n");
        for (size_t n = 0; n < N; n++) {
            total_c += (uint64_t)node_total[n];
            alloc_c += (uint64_t)node_alloc[n];
            const std::string addr = n < h->node_internal_ip.size() ? h->node_internal_ip[n] : "unknown";
            const bool master = n < h->node_is_master.size() && h->node_is_master[n];
            w.write(node_name(h, n) + "\t" + addr + "\t" + (master ? "master" : "worker") + "\t" + i64s(node_total[n]) +
                    "\t" + i64s(node_alloc[n]) + "\n");
        }
        w.write(kRule);
        w.write("Allocated/Total GPUs In Cluster:\n");
        w.write(i64s((int64_t)alloc_c) + "/" + i64s((int64_t)total_c) + " (" +
                i64s(percent((int64_t)alloc_c, (int64_t)total_c)) + "%)\t\n");
        return emit(w.flush(), text_out, len_out);
    }

    // details: pods of node n, in list order, with gpuInPod > 0 (gpu.go:8-15).  The
    // pods the reference sees are the ACTIVE ones (server-side selector).
    std::vector<std::vector<uint32_t>> rows_of_node(N);
    {
        std::vector<std::pair<uint32_t, uint32_t>> tmp;          // (list index, row)
        for (size_t row = 0; row < P; row++) {
            const uint32_t n = h->v.pod_node[row];
            if (n >= N) continue;
            const uint8_t ph = h->v.pod_phase[row];
            if (ph == ARENA_PHASE_SUCCEEDED || ph == ARENA_PHASE_FAILED) continue;
            if (pod_gpu[row] <= 0) continue;
            rows_of_node[n].push_back((uint32_t)row);
        }
        if (!h->pod_order.empty())
            for (auto &v : rows_of_node)
                std::sort(v.begin(), v.end(), [&](uint32_t a, uint32_t b) { return h->pod_order[a] < h->pod_order[b]; });
    }
    w.write("\n");
    for (size_t n = 0; n < N; n++) {
        total_c += (uint64_t)node_total[n];
        alloc_c += (uint64_t)node_alloc[n];
        std::string addr = "unknown";
        if (n < h->node_has_addr.size() && h->node_has_addr[n]) addr = h->node_first_addr[n];   // top_node.go:172-175
        const bool master = n < h->node_is_master.size() && h->node_is_master[n];
        const std::string name = node_name(h, n);
        w.write("\n");
        w.write("NAME:\t" + name + "\n");
        w.write("IPADDRESS:\t" + addr + "\n");
        w.write(std::string("ROLE:\t") + (master ? "master" : "worker") + "\n");
        if (!rows_of_node[n].empty()) {
            w.write("\n");
            w.write("NAMESPACE\tNAME\tGPU REQUESTS\tGPU LIMITS\n");
            for (uint32_t row : rows_of_node[n]) {
                const std::string ns = row < h->pod_namespaces.size() ? h->pod_namespaces[row] : "";
                const std::string pn = row < h->pod_names.size() ? h->pod_names[row] : "pod-" + std::to_string(row);
                w.write(ns + "\t" + pn + "\t" + i64s(pod_gpu[row]) + "\t" + i64s(pod_gpu[row]) + "\n");
            }
            w.write("\n");
        }
        if (!(node_total[n] > 0)) w.write("\n");                 // top_node.go:199-203
        w.write("Total GPUs In Node " + name + ":\t" + i64s(node_total[n]) + " \t\n");
        w.write("Allocated GPUs In Node " + name + ":\t" + i64s(node_alloc[n]) + " (" +
                i64s(percent(node_alloc[n], node_total[n])) + "%)\t\n");
        w.write(kRule);
    }
    w.write("\n");
    w.write("\n");
    w.write("Allocated/Total GPUs In Cluster:\t");
    w.write(i64s((int64_t)alloc_c) + "/" + i64s((int64_t)total_c) + " (" + i64s(percent((int64_t)alloc_c, (int64_t)total_c)) +
            "%)\t\n");
    return emit(w.flush(), text_out, len_out);
}

// makeTrainingJobOrderdByAge (trainer.go:39-66): Less(i, j) = StartTime(i) == nil, else false when
// StartTime(j) == nil, else StartTime(i).After(StartTime(j)) -- jobs without a start time first, then
// the newest first.  sort.Sort leaves ties open (and two nil times compare "less" both ways): they
// are broken by ascending job id here, as for `top job`.  Host work: the keys are job metadata.
int arena_job_order_by_age(const arena_host_snapshot *h, uint32_t *job_order)
{
    if (!h) return fail(ARENA_E_INVALID, "host snapshot is NULL");
    const size_t J = h->jobs.size();
    if (J && !job_order) return fail(ARENA_E_INVALID, "job_order is NULL");
    for (size_t j = 0; j < J; j++) job_order[j] = (uint32_t)j;
    std::stable_sort(job_order, job_order + J, [&](uint32_t a, uint32_t b) {
        const JobMeta &x = h->jobs[a], &y = h->jobs[b];
        if (x.start_present != y.start_present) return !x.start_present;
        if (!x.start_present) return false;
        return x.start_unix > y.start_unix;
    });
    return ARENA_OK;
}

int arena_format_top_job(const arena_host_snapshot *h, const int64_t *job_req, const int64_t *job_alloc,
                         const uint32_t *job_order, const uint32_t *job_latest, int64_t now_unix, char **text_out,
                         size_t *len_out)
{
    return arena_format_job_list(h, job_req, job_alloc, job_order, job_latest, now_unix, 1, text_out, len_out);
}

int arena_format_job_list(const arena_host_snapshot *h, const int64_t *job_req, const int64_t *job_alloc,
                          const uint32_t *job_order, const uint32_t *job_latest, int64_t now_unix, int display_gpu,
                          char **text_out, size_t *len_out)
{
    if (!h || !text_out) return fail(ARENA_E_INVALID, "NULL argument");
    const size_t J = h->jobs.size();
    if (J && (!job_order || !job_latest)) return fail(ARENA_E_INVALID, "job table is NULL");
    if (J && display_gpu && (!job_req || !job_alloc)) return fail(ARENA_E_INVALID, "job table is NULL");
    TabWriter w(2);
    uint64_t total_alloc = 0, total_req = 0;
    if (display_gpu) w.write("NAME\tSTATUS\tTRAINER\tAGE\tNODE\tGPU(Requests)\tGPU(Allocated)\n");
    else w.write("NAME\tSTATUS\tTRAINER\tAGE\tNODE\n");                       // list.go:92-96
    for (size_t k = 0; k < J; k++) {
        const uint32_t j = job_order[k];
        if (j >= J) return fail(ARENA_E_INVALID, "job_order[%zu] = %u out of range", k, j);
        const JobMeta &jm = h->jobs[j];
        std::string status, host_ip = "N/A";
        if (jm.trainer == "tfjob") {                              // trainer_tensorflow.go:69-86, 158-165
            status = jm.tf_found ? jm.tf_status : "UNKNOWN";
            if (status == "RUNNING")
                host_ip = jm.tf_chief_row < h->pod_host_ip.size() ? h->pod_host_ip[jm.tf_chief_row] : "";
        } else {                                                  // job_info.go:88-108, 41-48
            if (jm.active > 0) status = "RUNNING";
            else if (jm.succeeded > 0) status = "SUCCEEDED";
            else if (jm.failed > 0) status = "FAILED";
            const uint32_t row = job_latest[j];
            const bool have = row < h->pod_host_ip.size();
            const std::string pod_ip = have ? h->pod_host_ip[row] : "";
            if (status == "RUNNING") {
                if (pod_ip.empty()) status = "PENDING";
                else if (have && h->v.pod_phase[row] == ARENA_PHASE_PENDING) status = "PENDING";
            }
            if (status == "RUNNING") host_ip = pod_ip;
        }
        const std::string age = jm.has_start ? short_human_duration(now_unix - jm.start_unix) : "0s";
        if (!display_gpu) {                                       // list.go:112-118
            w.write(jm.name + "\t" + status + "\t" + upper(jm.trainer) + "\t" + age + "\t" + host_ip + "\n");
            continue;
        }
        total_alloc += (uint64_t)job_alloc[j];
        total_req += (uint64_t)job_req[j];
        w.write(jm.name + "\t" + status + "\t" + upper(jm.trainer) + "\t" + age + "\t" + host_ip + "\t" + i64s(job_req[j]) +
                "\t" + i64s(job_alloc[j]) + "\n");
    }
    if (!display_gpu) return emit(w.flush(), text_out, len_out);
    w.write("\n");
    w.write("\n");
    w.write("Total Allocated GPUs of Training Job:\n");
    w.write(i64s((int64_t)total_alloc) + " \t\n");
    w.write("\n");
    w.write("Total Requested GPUs of Training Job:\n");
    w.write(i64s((int64_t)total_req) + " \t\n");
    return emit(w.flush(), text_out, len_out);
}

}  // extern "C"
