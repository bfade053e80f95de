// JSON fixtures -> columnar snapshot.  Replaces, for fixture-driven runs, the
// List calls and the client-side decoding the reference does before its
// accounting (paths relative to /root/reference):
//   pods / nodes     cmd/arena/commands/pod_helper.go:15-43, top_node.go:71
//   jobs / tfjobs    pod_helper.go:45-55, trainer_tensorflow.go:176-208
//   helm releases    util/helm/helm.go:164-195
//   job membership   trainer_mpi.go:91-119, 222-332; trainer_standalone.go:71-99,
//                    182-273; trainer_tensorflow.go:215-245, 315-418; trainer.go:10-23
//   key selection    gpu.go:18-37, 61-79 (presence decides, not positivity)
//   schema           vendor/k8s.io/api/core/v1/types.go (field names only)
#include "host_snapshot.h"
#include "json.h"
#include "quantity.h"

#include <algorithm>
#include <atomic>
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <limits>
#include <map>
#include <numeric>
#include <thread>
#include <unordered_map>

using arena_json::Value;
using namespace arena_host;

namespace {

const char kPrimary[] = "nvidia.com/gpu";                        // const.go:7
const char kDeprecated[] = "alpha.kubernetes.io/nvidia-gpu";     // const.go:9
const char kMasterLabel[] = "node-role.kubernetes.io/master";    // const.go:11

struct LoadError {
    std::string msg;
};

[[noreturn]] void bad(const std::string &m) { throw LoadError{m}; }

const Value *child(const Value *v, const char *key)
{
    if (!v || v->is_null()) return nullptr;
    if (!v->is_object()) bad(std::string("expected an object around \"") + key + "\"");
    const Value *c = v->get(key);
    if (c && c->is_null()) return nullptr;
    return c;
}

std::string str_field(const Value *v, const char *key)
{
    const Value *c = child(v, key);
    if (!c) return std::string();
    if (!c->is_string()) bad(std::string("field \"") + key + "\" is not a string");
    return c->str();
}

int64_t int_field(const Value *v, const char *key)
{
    const Value *c = child(v, key);
    if (!c) return 0;
    if (c->type != arena_json::Type::Number) bad(std::string("field \"") + key + "\" is not a number");
    int64_t x = 0;
    bool neg = false;
    for (char ch : c->raw) {
        if (ch == '-') { neg = true; continue; }
        if (ch < '0' || ch > '9') bad(std::string("field \"") + key + "\" is not an integer");
        x = x * 10 + (ch - '0');
    }
    return neg ? -x : x;
}

// map[string]string lookup: (found, value)
bool label(const Value *labels, const char *key, std::string &out)
{
    if (!labels) return false;
    if (!labels->is_object()) bad("labels is not an object");
    const Value *c = labels->get(key);
    if (!c) return false;
    if (c->is_null()) { out.clear(); return true; }       // null decodes to "" with the key present
    if (!c->is_string()) bad("label value is not a string");
    out = c->str();
    return true;
}

// Decode every entry of a ResourceList (a bad quantity anywhere fails the List
// decode in the reference) and return the selected GPU quantity.
void resource_list(const Value *rl, const char *what, bool select, int64_t &val, int8_t &scale)
{
    val = 0;
    scale = 0;
    if (!rl) return;
    if (!rl->is_object()) bad(std::string(what) + " is not an object");
    const Value *primary = nullptr, *deprecated = nullptr;
    for (const Value &m : rl->children()) {
        ParsedQuantity q;
        if (!unmarshal_quantity_json(m.raw, q))
            bad(std::string("invalid quantity ") + std::string(m.raw) + " for \"" + std::string(m.key) + "\" in " + what);
        if (m.key == kPrimary) primary = &m;                    // duplicates: last wins
        else if (m.key == kDeprecated) deprecated = &m;
    }
    if (!select) return;
    const Value *pick = primary ? primary : deprecated;          // gpu.go:61-79: key PRESENCE
    if (!pick) return;
    ParsedQuantity q;
    unmarshal_quantity_json(pick->raw, q);
    column_encoding(q, val, scale);
}

int64_t days_from_civil(int64_t y, unsigned m, unsigned d)
{
    y -= m <= 2;
    const int64_t era = (y >= 0 ? y : y - 399) / 400;
    const unsigned yoe = (unsigned)(y - era * 400);
    const unsigned doy = (153 * (m + (m > 2 ? -3 : 9)) + 2) / 5 + d - 1;
    const unsigned doe = yoe * 365 + yoe / 4 - yoe / 100 + doy;
    return era * 146097 + (int64_t)doe - 719468;
}

using arena_host::kZeroTimeUnix;                   // Go's time.Time{} (0001-01-01T00:00:00Z)

// metav1.Time.UnmarshalJSON: null -> zero time, else time.Parse(RFC3339)
// (vendor/k8s.io/apimachinery/pkg/apis/meta/v1/time.go:105-121).  Sub-second
// digits are accepted and dropped (the API server writes whole seconds).
int64_t parse_time(const Value *v, const char *what)
{
    if (!v) return kZeroTimeUnix;
    if (!v->is_string()) bad(std::string(what) + " is not a string");
    const std::string s = v->str();
    auto num = [&](size_t at, size_t n) -> int {
        if (at + n > s.size()) bad(std::string("bad RFC3339 time \"") + s + "\"");
        int x = 0;
        for (size_t i = at; i < at + n; i++) {
            if (s[i] < '0' || s[i] > '9') bad(std::string("bad RFC3339 time \"") + s + "\"");
            x = x * 10 + (s[i] - '0');
        }
        return x;
    };
    if (s.size() < 20 || s[4] != '-' || s[7] != '-' || (s[10] != 'T' && s[10] != 't') || s[13] != ':' || s[16] != ':')
        bad(std::string("bad RFC3339 time \"") + s + "\"");
    const int Y = num(0, 4), M = num(5, 2), D = num(8, 2), h = num(11, 2), mi = num(14, 2), sec = num(17, 2);
    if (M < 1 || M > 12 || D < 1 || D > 31 || h > 23 || mi > 59 || sec > 59) bad(std::string("bad RFC3339 time \"") + s + "\"");
    size_t i = 19;
    if (i < s.size() && s[i] == '.') {
        i++;
        size_t st = i;
        while (i < s.size() && s[i] >= '0' && s[i] <= '9') i++;
        if (i == st) bad(std::string("bad RFC3339 time \"") + s + "\"");
    }
    int64_t off = 0;
    if (i < s.size() && (s[i] == 'Z' || s[i] == 'z')) i++;
    else if (i < s.size() && (s[i] == '+' || s[i] == '-')) {
        const int sign = s[i] == '-' ? -1 : 1;
        if (i + 6 > s.size() || s[i + 3] != ':') bad(std::string("bad RFC3339 time \"") + s + "\"");
        off = sign * (num(i + 1, 2) * 3600 + num(i + 4, 2) * 60);
        i += 6;
    } else bad(std::string("bad RFC3339 time \"") + s + "\"");
    if (i != s.size()) bad(std::string("bad RFC3339 time \"") + s + "\"");
    return days_from_civil(Y, (unsigned)M, (unsigned)D) * 86400 + h * 3600 + mi * 60 + sec - off;
}

uint8_t phase_code(const std::string &p)
{
    if (p.empty()) return ARENA_PHASE_EMPTY;
    if (p == "Pending") return ARENA_PHASE_PENDING;
    if (p == "Running") return ARENA_PHASE_RUNNING;
    if (p == "Succeeded") return ARENA_PHASE_SUCCEEDED;
    if (p == "Failed") return ARENA_PHASE_FAILED;
    if (p == "Unknown") return ARENA_PHASE_UNKNOWN;
    return ARENA_PHASE_OTHER;
}

// (PodRec / PodLabels: host_snapshot.h.  The labels are copied out so a pod record does not point
// into its per-item, short-lived JSON tree.)
using arena_host::PodLabels;
using arena_host::PodRec;
using arena_host::kHasRelease;
using arena_host::kHasApp;
using arena_host::kHasGroup;
using arena_host::kHasReplicaType;
using arena_host::kHasReplicaIndex;
static_assert(ARENA_NONE == 0xFFFFFFFFu && ARENA_ROLE_COUNTS == 0, "PodRec defaults");

struct NodeRec {
    std::string name, internal_ip, first_addr;
    uint8_t has_addr = 0, is_master = 0;
    int64_t cap_val = 0;
    int8_t cap_scale = 0;
};

struct ObjRec {                 // batch/v1 Job or TFJob: what membership and the printer need
    std::string ns;
    const Value *labels = nullptr;
    const Value *status = nullptr;
};

const Value *items_of(const Value &doc, const char *what)
{
    if (!doc.is_object()) bad(std::string(what) + " is not a JSON object");
    const Value *items = doc.get("items");
    if (!items || items->is_null()) return nullptr;
    if (!items->is_array()) bad(std::string(what) + ".items is not an array");
    return items;
}

// Which members the loader ever reads.  Everything else (a real PodList is mostly
// managedFields, volumes, env, containerStatuses ...) is validated and dropped by the
// parser without building nodes.  Inside the string->string / string->Quantity maps
// every key is kept: a bad quantity under ANY resource name fails the decode.
bool is_map_parent(std::string_view k)
{
    return k == "labels" || k == "limits" || k == "requests" || k == "capacity" || k == "allocatable";
}

bool one_of(std::string_view key, std::initializer_list<std::string_view> set)
{
    for (std::string_view s : set)
        if (key == s) return true;
    return false;
}

bool keep_core(std::string_view parent, std::string_view key)      // PodList, NodeList
{
    if (is_map_parent(parent)) return true;
    switch (key.size()) {
    case 4: return one_of(key, {"spec", "name", "kind", "type"});
    case 5: return one_of(key, {"items", "phase"});
    case 6: return one_of(key, {"status", "labels", "limits", "hostIP"});
    case 7: return one_of(key, {"address"});
    case 8: return one_of(key, {"metadata", "nodeName", "requests", "capacity"});
    case 9: return one_of(key, {"namespace", "resources", "addresses"});
    case 10: return one_of(key, {"containers"});
    case 11: return one_of(key, {"allocatable"});
    case 14: return one_of(key, {"initContainers"});
    case 15: return one_of(key, {"ownerReferences"});
    case 17: return one_of(key, {"creationTimestamp"});
    default: return false;
    }
}

bool keep_job(std::string_view parent, std::string_view key)       // JobList, TFJobList
{
    if (is_map_parent(parent)) return true;
    switch (key.size()) {
    case 4: return one_of(key, {"type"});
    case 5: return one_of(key, {"items"});
    case 6: return one_of(key, {"status", "labels", "active", "failed"});
    case 8: return one_of(key, {"metadata"});
    case 9: return one_of(key, {"namespace", "succeeded", "startTime"});
    case 10: return one_of(key, {"conditions"});
    default: return false;
    }
}

struct Doc {
    arena_json::Tree tree;
    std::vector<arena_json::Range> item_ranges; // elements of .items when `split`
    bool split = false;
};

struct SyntaxError {
    size_t offset;
    std::string message;
};

[[noreturn]] void bad_syntax(const char *what, size_t offset, const std::string &message)
{
    bad(std::string(what) + ": JSON syntax error at byte " + std::to_string(offset) + ": " + message);
}

// Parses the list envelope.  With `split` the elements of the root "items" array are
// left as byte ranges (doc.split says whether that happened) for for_each_item.
void parse_doc(const char *text, size_t len, const char *what, Doc &doc, arena_json::KeyFilter filter, bool split)
{
    arena_json::ParseError e;
    arena_json::ParseOptions opt;
    opt.filter = filter;
    if (split) {
        opt.split_key = "items";
        opt.split_out = &doc.item_ranges;
        opt.split_done = &doc.split;
    }
    if (!arena_json::parse(std::string_view(text, len), doc.tree, e, opt)) bad_syntax(what, e.offset, e.message);
}

// ARENA_LOADER_TRACE=1: per-phase wall times on stderr (tools/loader_timing.py reads them)
struct PhaseTrace {
    bool on = std::getenv("ARENA_LOADER_TRACE") != nullptr;
    std::chrono::steady_clock::time_point t = std::chrono::steady_clock::now();
    void mark(const char *phase)
    {
        if (!on) return;
        const auto now = std::chrono::steady_clock::now();
        std::fprintf(stderr, "[arena loader] %-22s %8.2f ms\n", phase, std::chrono::duration<double, std::milli>(now - t).count());
        t = now;
    }
};

unsigned loader_threads(size_t n_items)
{
    unsigned t = std::thread::hardware_concurrency();
    if (const char *env = std::getenv("ARENA_LOADER_THREADS")) {
        const long v = std::strtol(env, nullptr, 10);
        if (v > 0) t = (unsigned)v;
    }
    if (t < 1) t = 1;
    if (t > 64) t = 64;
    const size_t by_work = n_items / 256 + 1;                    // a thread is not worth starting for fewer items
    return (unsigned)std::min<size_t>(t, by_work);
}

// Runs extract(item, rec) over every element of doc's "items", in parallel over the
// byte ranges when the envelope was split.  Error precedence is that of a sequential
// pass: the syntax error at the lowest byte offset, else the extraction error of the
// lowest item index.
template <class Rec, class Extract>
void for_each_item(const char *text, size_t len, const char *what, Doc &doc, arena_json::KeyFilter filter,
                   std::vector<Rec> &recs, Extract extract)
{
    const Value *items = items_of(doc.tree.root(), what);
    if (!items) { recs.clear(); return; }
    if (!doc.split) {
        recs.resize(items->size());
        size_t k = 0;
        for (const Value &item : items->children()) extract(item, recs[k++]);
        return;
    }
    const std::vector<arena_json::Range> &ranges = doc.item_ranges;
    const size_t n = ranges.size();
    recs.resize(n);
    const unsigned T = loader_threads(n);
    struct Outcome {
        bool syntax = false, extraction = false, nomem = false;
        size_t syntax_offset = 0, item = 0;
        std::string syntax_message, message;
    };
    std::vector<Outcome> outcomes(T);
    std::atomic<size_t> next{0};
    const size_t kChunk = 64;
    const std::string_view doc_text(text, len);
    auto work = [&](unsigned tid) {
        Outcome &o = outcomes[tid];
        arena_json::ParseOptions opt;
        opt.filter = filter;
        arena_json::Tree item;                                   // reused: its node pool stops growing after a few items
        for (;;) {
            const size_t lo = next.fetch_add(kChunk);
            if (lo >= n) break;
            const size_t hi = std::min(n, lo + kChunk);
            for (size_t k = lo; k < hi; k++) {
                try {
                    arena_json::ParseError e;
                    if (!arena_json::parse_range(doc_text, ranges[k], 2, "items", item, e, opt)) {
                        if (!o.syntax || e.offset < o.syntax_offset) {
                            o.syntax = true; o.syntax_offset = e.offset; o.syntax_message = e.message;
                        }
                        continue;
                    }
                    extract(item.root(), recs[k]);
                } catch (const LoadError &e) {
                    if (!o.extraction || k < o.item) { o.extraction = true; o.item = k; o.message = e.msg; }
                } catch (const std::bad_alloc &) {
                    o.nomem = true;
                    return;
                }
            }
        }
    };
    if (T <= 1) {
        work(0);
    } else {
        std::vector<std::thread> pool;
        pool.reserve(T - 1);
        for (unsigned t = 1; t < T; t++) pool.emplace_back(work, t);
        work(0);
        for (std::thread &th : pool) th.join();
    }
    const Outcome *first_syntax = nullptr, *first_extract = nullptr;
    for (const Outcome &o : outcomes) {
        if (o.nomem) throw std::bad_alloc();
        if (o.syntax && (!first_syntax || o.syntax_offset < first_syntax->syntax_offset)) first_syntax = &o;
        if (o.extraction && (!first_extract || o.item < first_extract->item)) first_extract = &o;
    }
    if (first_syntax) bad_syntax(what, first_syntax->syntax_offset, first_syntax->syntax_message);
    if (first_extract) bad(first_extract->message);
}

void pod_labels(const Value *labels, PodLabels &out)
{
    if (!labels) return;
    if (label(labels, "release", out.release)) out.has |= kHasRelease;
    if (label(labels, "app", out.app)) out.has |= kHasApp;
    if (label(labels, "group_name", out.group_name)) out.has |= kHasGroup;
    if (label(labels, "tf-replica-type", out.replica_type)) out.has |= kHasReplicaType;
    if (label(labels, "tf-replica-index", out.replica_index)) out.has |= kHasReplicaIndex;
}

void extract_node(const Value &node, NodeRec &r)
{
    if (!node.is_object()) bad("NodeList item is not an object");
    const Value *meta = child(&node, "metadata");
    const Value *status = child(&node, "status");
    r.name = str_field(meta, "name");
    const Value *labels = child(meta, "labels");
    std::string tmp;
    r.is_master = label(labels, kMasterLabel, tmp) ? 1 : 0;               // top_node.go:244-250: key presence
    int64_t v; int8_t s;
    resource_list(child(status, "allocatable"), "node status.allocatable", false, v, s);
    resource_list(child(status, "capacity"), "node status.capacity", true, r.cap_val, r.cap_scale);
    r.internal_ip = "unknown";                                            // top_node.go:120-129
    const Value *addrs = child(status, "addresses");
    if (!addrs) return;
    if (!addrs->is_array()) bad("node status.addresses is not an array");
    bool have_internal = false;
    size_t i = 0;
    for (const Value &a : addrs->children()) {
        std::string type = str_field(&a, "type"), addr = str_field(&a, "address");
        if (!have_internal && type == "InternalIP") { r.internal_ip = addr; have_internal = true; }
        if (i++ == 0) { r.first_addr = std::move(addr); r.has_addr = 1; }   // top_node.go:173-175
    }
}

void extract_pod(const Value &pod, PodRec &r, const std::unordered_map<std::string, uint32_t> &node_id)
{
    if (!pod.is_object()) bad("PodList item is not an object");
    const Value *meta = child(&pod, "metadata");
    const Value *spec = child(&pod, "spec");
    const Value *status = child(&pod, "status");
    r.name = str_field(meta, "name");
    r.ns = str_field(meta, "namespace");
    pod_labels(child(meta, "labels"), r.labels);
    r.created = parse_time(child(meta, "creationTimestamp"), "metadata.creationTimestamp");
    if (const Value *owners = child(meta, "ownerReferences")) {
        if (!owners->is_array()) bad("metadata.ownerReferences is not an array");
        for (const Value &o : owners->children())
            if (str_field(&o, "kind") == "Job") { r.owner_is_job = true; break; }
    }
    r.node_name = str_field(spec, "nodeName");
    r.phase = phase_code(str_field(status, "phase"));
    r.host_ip = str_field(status, "hostIP");
    for (const char *list : {"initContainers", "containers"}) {
        const Value *cs = child(spec, list);
        if (!cs) continue;
        if (!cs->is_array()) bad(std::string("spec.") + list + " is not an array");
        const bool counted = std::strcmp(list, "containers") == 0;             // gpu.go:40: spec.containers only
        for (const Value &c : cs->children()) {
            if (!c.is_object()) bad("container is not an object");
            const Value *res = child(&c, "resources");
            int64_t v; int8_t s;
            resource_list(child(res, "requests"), "resources.requests", false, v, s);
            resource_list(child(res, "limits"), "resources.limits", counted, v, s);   // gpu.go:61-69: LIMITS
            if (counted) { r.ctr_val.push_back(v); r.ctr_scale.push_back(s); }
        }
    }
    if (!r.node_name.empty()) {
        auto it = node_id.find(r.node_name);
        if (it != node_id.end()) r.node = it->second;
    }
}

// Stable grouping of rows by key (node or job id; ARENA_NONE last): a counting sort.
std::vector<uint32_t> group_rows(const std::vector<PodRec> &pods, size_t n_keys, bool by_job)
{
    auto key_of = [&](const PodRec &r) -> size_t {
        const uint32_t k = by_job ? r.job : r.node;
        return k == ARENA_NONE ? n_keys : (size_t)k;
    };
    std::vector<size_t> start(n_keys + 2, 0);
    for (const PodRec &r : pods) start[key_of(r) + 1]++;
    for (size_t k = 1; k < start.size(); k++) start[k] += start[k - 1];
    std::vector<uint32_t> order(pods.size());
    for (size_t p = 0; p < pods.size(); p++) order[start[key_of(pods[p])]++] = (uint32_t)p;
    return order;
}

// hasCondition / isSucceeded / isFailed / isPending: trainer_tensorflow.go:420-439
bool tf_condition(const Value *status, const char *type)
{
    const Value *conds = child(status, "conditions");
    if (!conds) return false;
    if (!conds->is_array()) bad("tfjob status.conditions is not an array");
    for (const Value &c : conds->children())
        if (str_field(&c, "type") == type && str_field(&c, "status") == "True") return true;
    return false;
}

struct PodsNotDelivered {};     // thrown by load() when its PodSource gives up: nothing was produced

// `src` != nullptr: the pods come from there (the GPU loader) and `podlist` is not looked at.
int load(const char *podlist, size_t podlen, const char *nodelist, size_t nodelen, const char *joblist, size_t joblen,
         const char *tfjoblist, size_t tflen, const char *helm, size_t helmlen, bool group_by_job,
         arena_host_snapshot *h, arena_host::PodSource *src = nullptr)
{
    PhaseTrace trace;
    Doc pod_doc, node_doc, job_doc, tf_doc;
    parse_doc(nodelist, nodelen, "NodeList", node_doc, keep_core, true);
    if (!src) parse_doc(podlist, podlen, "PodList", pod_doc, keep_core, true);
    trace.mark("envelopes + item scan");

    // ---- nodes -------------------------------------------------------------------
    std::vector<NodeRec> nodes;
    try {
        for_each_item(nodelist, nodelen, "NodeList", node_doc, keep_core, nodes, extract_node);
    } catch (const LoadError &e) {
        // a sequential decode reads BOTH documents before looking at any field, so a syntax
        // error inside a pod outranks a bad field in a node
        if (src) throw PodsNotDelivered{};          // the plain path sorts out which document's error comes first
        if (pod_doc.split && e.msg.find("JSON syntax error") == std::string::npos) {
            Doc whole;
            parse_doc(podlist, podlen, "PodList", whole, [](std::string_view, std::string_view) { return false; }, false);
        }
        throw;
    }
    const size_t N = nodes.size();
    std::unordered_map<std::string, uint32_t> node_id;
    node_id.reserve(N * 2);
    h->node_names.resize(N); h->node_internal_ip.resize(N); h->node_first_addr.resize(N);
    h->node_has_addr.assign(N, 0); h->node_is_master.assign(N, 0);
    std::vector<int64_t> cap_val(N);
    std::vector<int8_t> cap_scale(N);
    std::string dup_node;
    for (size_t n = 0; n < N; n++) {
        NodeRec &r = nodes[n];
        if (!node_id.emplace(r.name, (uint32_t)n).second && dup_node.empty()) dup_node = r.name;
        h->node_names[n] = std::move(r.name);
        h->node_internal_ip[n] = std::move(r.internal_ip);
        h->node_first_addr[n] = std::move(r.first_addr);
        h->node_has_addr[n] = r.has_addr;
        h->node_is_master[n] = r.is_master;
        cap_val[n] = r.cap_val;
        cap_scale[n] = r.cap_scale;
    }
    if (!dup_node.empty()) {
        if (src) throw PodsNotDelivered{};
        bad("duplicate node name \"" + dup_node + "\" (the API server guarantees unique node names)");
    }

    trace.mark("nodes");

    // ---- pods --------------------------------------------------------------------
    std::vector<PodRec> pods;
    if (src) {
        if (!src->pods(h->node_names, pods)) throw PodsNotDelivered{};
    } else {
        for_each_item(podlist, podlen, "PodList", pod_doc, keep_core, pods,
                      [&](const Value &item, PodRec &r) { extract_pod(item, r, node_id); });
    }
    const size_t P = pods.size();
    trace.mark("pods");

    // ---- jobs (only when a `helm list` text is given) -------------------------------
    if (helm) {
        // util/helm/helm.go:180-192: name = first field, namespace = last field,
        // lines containing "NAME" are skipped, later lines overwrite earlier ones.
        std::map<std::string, std::string> releases;                  // sorted by name => job ids
        std::string_view text(helm, helmlen);
        size_t pos = 0;
        while (pos <= text.size()) {
            size_t nl = text.find('\n', pos);
            std::string_view line = text.substr(pos, nl == std::string_view::npos ? std::string_view::npos : nl - pos);
            pos = nl == std::string_view::npos ? text.size() + 1 : nl + 1;
            if (line.find("NAME") != std::string_view::npos) continue;
            std::vector<std::string_view> cols;
            size_t i = 0;
            auto is_sp = [](char c) { return c == ' ' || c == '\t' || c == '\r' || c == '\v' || c == '\f'; };
            while (i < line.size()) {
                while (i < line.size() && is_sp(line[i])) i++;
                size_t st = i;
                while (i < line.size() && !is_sp(line[i])) i++;
                if (i > st) cols.push_back(line.substr(st, i - st));
            }
            if (cols.size() > 1) releases[std::string(cols.front())] = std::string(cols.back());
        }

        std::vector<ObjRec> jobs, tfjobs;
        if (joblist) {
            parse_doc(joblist, joblen, "JobList", job_doc, keep_job, false);
            if (const Value *it = items_of(job_doc.tree.root(), "JobList"))
                for (const Value &j : it->children()) {
                    const Value *meta = child(&j, "metadata");
                    jobs.push_back({str_field(meta, "namespace"), child(meta, "labels"), child(&j, "status")});
                }
        }
        // NewTensorFlowJobTrainer lists TFJobs only when allPods is non-empty
        // (trainer_tensorflow.go:188-202); without a TFJobList the trainer is disabled.
        const bool tf_enabled = tfjoblist != nullptr;
        if (tfjoblist && P > 0) {
            parse_doc(tfjoblist, tflen, "TFJobList", tf_doc, keep_job, false);
            if (const Value *it = items_of(tf_doc.tree.root(), "TFJobList"))
                for (const Value &j : it->children()) {
                    const Value *meta = child(&j, "metadata");
                    tfjobs.push_back({str_field(meta, "namespace"), child(meta, "labels"), child(&j, "status")});
                }
        }
        // first object in list order per (release, app, namespace): the reference scans the whole
        // list for every release (O(R x objects)); an index gives the same answer in O(R + objects)
        auto index_objs = [&](const std::vector<ObjRec> &objs) {
            std::unordered_map<std::string, const ObjRec *> idx;
            for (const ObjRec &o : objs) {
                std::string rel, app;
                if (!label(o.labels, "release", rel) || !label(o.labels, "app", app)) continue;
                std::string key = rel;
                key.push_back('\0'); key += app; key.push_back('\0'); key += o.ns;
                idx.emplace(std::move(key), &o);                         // keeps the FIRST
            }
            return idx;
        };
        const auto job_idx = index_objs(jobs), tf_idx = index_objs(tfjobs);
        auto find_obj = [&](const std::unordered_map<std::string, const ObjRec *> &idx, const std::string &name,
                            const std::string &ns, const char *app) -> const ObjRec * {
            std::string key = name;
            key.push_back('\0'); key += app; key.push_back('\0'); key += ns;
            auto it = idx.find(key);
            return it == idx.end() ? nullptr : it->second;
        };
        std::unordered_map<std::string, uint32_t> job_of_release;
        for (const auto &rel : releases) {
            const std::string &name = rel.first, &ns = rel.second;
            JobMeta jm;
            jm.name = name;
            jm.ns = ns;
            const ObjRec *obj = nullptr;
            // trainer precedence: MPI, Standalone, TensorFlow (trainer.go:10-23); with an
            // empty Job list the reference would fall back to API calls: not supported here.
            if ((obj = find_obj(job_idx, name, ns, "tf-horovod"))) jm.trainer = "mpijob";
            else if ((obj = find_obj(job_idx, name, ns, "training"))) jm.trainer = "standalonejob";
            else if (tf_enabled && (obj = find_obj(tf_idx, name, ns, "tfjob"))) jm.trainer = "tfjob";
            else continue;                                               // "Unkown chart": no row
            const Value *st = obj->status;
            if (jm.trainer == "tfjob") {
                jm.tf_found = true;
                if (tf_condition(st, "Succeeded")) jm.tf_status = "SUCCEEDED";
                else if (tf_condition(st, "Failed")) jm.tf_status = "FAILED";
                else if (tf_condition(st, "Created") || tf_condition(st, "Restarting")) jm.tf_status = "PENDING";
                else jm.tf_status = "RUNNING";
            } else {
                jm.active = int_field(st, "active");
                jm.succeeded = int_field(st, "succeeded");
                jm.failed = int_field(st, "failed");
            }
            if (const Value *t = child(st, "startTime")) {
                jm.start_unix = parse_time(t, "status.startTime");
                jm.start_present = true;
                jm.has_start = jm.start_unix != kZeroTimeUnix;
            }
            job_of_release.emplace(name, (uint32_t)h->jobs.size());
            h->jobs.push_back(std::move(jm));
        }
        // pod -> job: the release label names the job, the trainer's own predicate decides
        for (PodRec &r : pods) {
            if (!(r.labels.has & kHasRelease)) continue;
            auto it = job_of_release.find(r.labels.release);
            if (it == job_of_release.end()) continue;
            const JobMeta &jm = h->jobs[it->second];
            if (r.ns != jm.ns) continue;
            if (jm.trainer == "mpijob") {                                // trainer_mpi.go:315-332, 247-279
                if (!r.labels.is(kHasApp, r.labels.app, "tf-horovod")) continue;
                r.role = r.owner_is_job ? ARENA_ROLE_LATEST : ARENA_ROLE_COUNTS;
            } else if (jm.trainer == "standalonejob") {                  // trainer_standalone.go:198-221, 257-273
                if (!r.labels.is(kHasApp, r.labels.app, "training")) continue;
                r.role = ARENA_ROLE_LATEST;
            } else {                                                     // trainer_tensorflow.go:394-418
                if (!r.labels.is(kHasApp, r.labels.app, "tfjob") || !r.labels.is(kHasGroup, r.labels.group_name, "kubeflow.org")) continue;
                r.role = ARENA_ROLE_COUNTS;
            }
            r.job = it->second;
        }
        // "latest pod of the job": `latest.Before(current)` is strict (trainer_mpi.go:263-278,
        // trainer_standalone.go:198-221), so among ARENA_ROLE_LATEST pods with the same newest
        // second the FIRST IN LIST ORDER wins.  The kernels break that tie by ROW index, and
        // rows are about to be regrouped by node: settle it here, while list order is still
        // known, by moving the later ones of such a tie one second back (they lose either
        // way; pod_created serves nothing but this rule).
        std::vector<int64_t> newest(h->jobs.size(), std::numeric_limits<int64_t>::min());
        for (const PodRec &r : pods)
            if (r.job != ARENA_NONE && r.role == ARENA_ROLE_LATEST && r.created > newest[r.job]) newest[r.job] = r.created;
        std::vector<uint8_t> seen(h->jobs.size(), 0);
        for (PodRec &r : pods) {
            if (r.job == ARENA_NONE || r.role != ARENA_ROLE_LATEST || r.created != newest[r.job]) continue;
            if (seen[r.job] && r.created > std::numeric_limits<int64_t>::min()) r.created -= 1;
            seen[r.job] = 1;
        }
    }
    const size_t J = h->jobs.size();
    trace.mark("jobs");

    // ---- row order: stable grouping by node (or by job) --------------------------------
    const std::vector<uint32_t> order = group_rows(pods, group_by_job ? J : N, group_by_job);
    h->pod_order = order;

    size_t C = 0;
    for (const PodRec &r : pods) C += r.ctr_val.size();
    if (C > 0xFFFFFFFFull || P >= 0xFFFFFFFFull) return fail(ARENA_E_TOO_LARGE, "snapshot exceeds 32-bit offsets");
    int rc;
    if ((rc = h->c_pod_node.alloc(P * 4, false))) return rc;
    if ((rc = h->c_pod_phase.alloc(P, false))) return rc;
    if ((rc = h->c_pod_ctr_begin.alloc((P + 1) * 4, false))) return rc;
    if ((rc = h->c_ctr_val.alloc(C * 8, false))) return rc;
    if ((rc = h->c_ctr_scale.alloc(C, false))) return rc;
    if ((rc = h->c_node_cap_val.alloc(N * 8, false))) return rc;
    if ((rc = h->c_node_cap_scale.alloc(N, false))) return rc;
    if ((rc = h->c_pod_job.alloc(P * 4, false))) return rc;
    if ((rc = h->c_pod_role.alloc(P, false))) return rc;
    if ((rc = h->c_pod_created.alloc(P * 8, false))) return rc;
    uint32_t *pod_node = h->c_pod_node.as<uint32_t>();
    uint8_t *pod_phase = h->c_pod_phase.as<uint8_t>();
    uint32_t *cb = h->c_pod_ctr_begin.as<uint32_t>();
    int64_t *ctr_val = h->c_ctr_val.as<int64_t>();
    int8_t *ctr_scale = h->c_ctr_scale.as<int8_t>();
    uint32_t *pod_job = h->c_pod_job.as<uint32_t>();
    uint8_t *pod_role = h->c_pod_role.as<uint8_t>();
    int64_t *pod_created = h->c_pod_created.as<int64_t>();
    if (N) {                                 // (an empty vector's data() may be null: not a memcpy argument)
        std::memcpy(h->c_node_cap_val.ptr, cap_val.data(), N * 8);
        std::memcpy(h->c_node_cap_scale.ptr, cap_scale.data(), N);
    }
    h->pod_names.resize(P); h->pod_namespaces.resize(P); h->pod_host_ip.resize(P);
    std::vector<uint32_t> row_of(P);
    {   // CSR offsets first (a running sum in row order), then the rows themselves, which no longer depend on
        // each other, on a few threads: a row is a dozen scattered reads of its pod record
        size_t c = 0;
        for (size_t row = 0; row < P; row++) { cb[row] = (uint32_t)c; c += pods[order[row]].ctr_val.size(); }
        cb[P] = (uint32_t)c;
    }
    auto fill = [&](size_t lo, size_t hi) {
        for (size_t row = lo; row < hi; row++) {
            PodRec &r = pods[order[row]];
            row_of[order[row]] = (uint32_t)row;
            pod_node[row] = r.node;
            pod_phase[row] = r.phase;
            size_t c = cb[row];
            for (size_t k = 0; k < r.ctr_val.size(); k++, c++) { ctr_val[c] = r.ctr_val[k]; ctr_scale[c] = r.ctr_scale[k]; }
            pod_job[row] = r.job;
            pod_role[row] = r.role;
            pod_created[row] = r.created;
            h->pod_names[row] = std::move(r.name);
            h->pod_namespaces[row] = std::move(r.ns);
            h->pod_host_ip[row] = std::move(r.host_ip);
        }
    };
    const unsigned fill_threads = std::min(loader_threads(P / 8), 8u);     // (from ~4000 rows up)
    if (fill_threads <= 1) fill(0, P);
    else {
        std::vector<std::thread> pool;
        for (unsigned t = 1; t < fill_threads; t++) pool.emplace_back(fill, P * t / fill_threads, P * (t + 1) / fill_threads);
        fill(0, P / fill_threads);
        for (std::thread &th : pool) th.join();
    }

    // chief pod of a tfjob: worker index 0; the LAST match in list order is kept
    // (trainer_tensorflow.go:336-338, 357-372)
    for (size_t p = 0; p < P; p++) {
        const PodRec &r = pods[p];
        if (r.job == ARENA_NONE || h->jobs[r.job].trainer != "tfjob") continue;
        if (r.labels.is(kHasReplicaType, r.labels.replica_type, "worker") && r.labels.is(kHasReplicaIndex, r.labels.replica_index, "0"))
            h->jobs[r.job].tf_chief_row = row_of[p];
    }

    trace.mark("columns");
    arena_snapshot_view &v = h->v;
    v.n_pods = P; v.n_containers = C; v.n_nodes = N; v.n_jobs = J;
    v.pod_node = pod_node; v.pod_phase = pod_phase; v.pod_ctr_begin = cb;
    v.ctr_val = ctr_val; v.ctr_scale = ctr_scale;
    v.node_cap_val = h->c_node_cap_val.as<int64_t>(); v.node_cap_scale = h->c_node_cap_scale.as<int8_t>();
    v.pod_job = pod_job; v.pod_role = pod_role; v.pod_created = pod_created;
    return ARENA_OK;
}

}  // namespace

namespace arena_host {

struct PodItemParser::Impl {
    std::unordered_map<std::string, uint32_t> node_id;
    arena_json::Tree item;
};

PodItemParser::PodItemParser(const std::vector<std::string> &node_names) : impl(new Impl())
{
    impl->node_id.reserve(node_names.size() * 2);
    for (size_t n = 0; n < node_names.size(); n++) impl->node_id.emplace(node_names[n], (uint32_t)n);   // first wins, as in load()
}

PodItemParser::~PodItemParser() { delete impl; }

bool PodItemParser::parse_full(const char *text, size_t text_len, size_t begin, size_t end, PodRec &rec)
{
    try {
        arena_json::ParseOptions opt;
        opt.filter = keep_core;
        arena_json::ParseError e;
        arena_json::Range r;
        r.begin = begin;
        r.end = end;
        if (!arena_json::parse_range(std::string_view(text, text_len), r, 2, "items", impl->item, e, opt)) return false;
        rec = PodRec();
        extract_pod(impl->item.root(), rec, impl->node_id);
        return true;
    } catch (const LoadError &) {
        return false;
    } catch (const std::bad_alloc &) {
        return false;
    }
}

bool PodItemParser::parse(const char *text, size_t text_len, size_t begin, size_t end, uint32_t &node, uint8_t &phase,
                          std::vector<int64_t> &ctr_val, std::vector<int8_t> &ctr_scale)
{
    PodRec rec;
    if (!parse_full(text, text_len, begin, end, rec)) return false;
    node = rec.node;
    phase = rec.phase;
    ctr_val = std::move(rec.ctr_val);
    ctr_scale = std::move(rec.ctr_scale);
    return true;
}

int64_t canonical_time_unix(const char *p)
{
    auto two = [&](int at) { return (p[at] - '0') * 10 + (p[at + 1] - '0'); };
    const int64_t Y = two(0) * 100 + two(2);
    return days_from_civil(Y, (unsigned)two(5), (unsigned)two(8)) * 86400 + two(11) * 3600 + two(14) * 60 + two(17);
}

int load_with_pods(PodSource &src, const char *nodelist, size_t nodelen, const char *joblist, size_t joblen,
                   const char *tfjoblist, size_t tflen, const char *helm, size_t helmlen, bool group_by_job,
                   arena_host_snapshot **out, bool *handled)
{
    *out = nullptr;
    *handled = false;
    arena_host_snapshot *h = new arena_host_snapshot();
    int rc = ARENA_OK;
    try {
        rc = load(nullptr, 0, nodelist, nodelen, joblist, joblen, tfjoblist, tflen, helm, helmlen, group_by_job, h, &src);
        if (!rc) { *out = h; *handled = true; return ARENA_OK; }
    } catch (const PodsNotDelivered &) {
        rc = ARENA_OK;                                  // not handled: the caller runs the plain path
    } catch (const LoadError &) {
        rc = ARENA_OK;                                  // the plain path owns every error message (and their order)
    } catch (const std::bad_alloc &) {
        rc = fail(ARENA_E_NOMEM, "out of host memory while loading");
    } catch (const std::exception &) {
        rc = ARENA_OK;                                  // (a helper thread that could not start, ...): the plain path
    }
    delete h;
    return rc;
}

}  // namespace arena_host

extern "C" {

int arena_load_json(const char *podlist_json, size_t podlist_len, const char *nodelist_json, size_t nodelist_len,
                    const char *joblist_json, size_t joblist_len, const char *tfjoblist_json, size_t tfjoblist_len,
                    const char *helm_list_text, size_t helm_len, int group_by_job, arena_host_snapshot **out)
{
    if (!out) return fail(ARENA_E_INVALID, "out is NULL");
    *out = nullptr;
    if (!podlist_json || !nodelist_json) return fail(ARENA_E_INVALID, "PodList and NodeList are required");
    arena_host_snapshot *h = new arena_host_snapshot();
    int rc;
    try {
        rc = load(podlist_json, podlist_len, nodelist_json, nodelist_len, joblist_json, joblist_len, tfjoblist_json,
                  tfjoblist_len, helm_list_text, helm_len, group_by_job != 0, h);
    } catch (const LoadError &e) {
        rc = fail(ARENA_E_PARSE, "%s", e.msg.c_str());
    } catch (const std::bad_alloc &) {
        rc = fail(ARENA_E_NOMEM, "out of host memory while loading");
    }
    if (rc) { delete h; return rc; }
    *out = h;
    return ARENA_OK;
}

const char *arena_host_job_name(const arena_host_snapshot *h, uint32_t job)
{
    return (h && job < h->jobs.size()) ? h->jobs[job].name.c_str() : "";
}
const char *arena_host_job_trainer(const arena_host_snapshot *h, uint32_t job)
{
    return (h && job < h->jobs.size()) ? h->jobs[job].trainer.c_str() : "";
}
const char *arena_host_node_name(const arena_host_snapshot *h, uint32_t node)
{
    return (h && node < h->node_names.size()) ? h->node_names[node].c_str() : "";
}
const char *arena_host_pod_name(const arena_host_snapshot *h, uint32_t row)
{
    return (h && row < h->pod_names.size()) ? h->pod_names[row].c_str() : "";
}

int arena_quantity_parse(const char *raw, size_t len, int64_t *value_out, int64_t *amount_val_out,
                         int32_t *amount_scale_out, int *slow_path_out)
{
    if (!raw) return fail(ARENA_E_INVALID, "raw is NULL");
    ParsedQuantity q;
    if (!unmarshal_quantity_json(std::string_view(raw, len), q))
        return fail(ARENA_E_PARSE, "invalid quantity: %.*s", (int)(len > 200 ? 200 : len), raw);
    if (value_out) *value_out = q.value;
    if (amount_val_out) *amount_val_out = q.amount;
    if (amount_scale_out) *amount_scale_out = q.scale;
    if (slow_path_out) *slow_path_out = q.slow ? 1 : 0;
    return ARENA_OK;
}

}  // extern "C"
