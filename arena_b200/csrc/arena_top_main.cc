// arena-top: the `top` command surface of the reference CLI
// (cmd/arena/commands/top.go:17-32, top_node.go:26-55, top_job.go:13-84) over
// fixture files instead of an API server, computing on the GPU through the C ABI.
//
//   arena-top node [-d|--details] --pods PODLIST.json --nodes NODELIST.json
//   arena-top job  --pods PODLIST.json --nodes NODELIST.json --jobs JOBLIST.json
//                  [--tfjobs TFJOBLIST.json] --releases HELM_LIST.txt [--now UNIX_SECONDS]
//   arena-top list (same flags as job): `arena list` (cmd/arena/commands/list.go:16-80), the other
//                  user of the job printer: ordered by age, no GPU columns
// Common: [--device N] [--gpus N] [--time] [--loader host|device]
//   --gpus N (N > 1): this one process drives devices 0..N-1 (arena_acct_init_devices): node ranges for
//   `node`, job-id ranges for `job`.  Same output for any N.
//   --loader device (the default): the PodList is parsed on the GPU -- `node` without -d keeps the columns
//   there (arena_load_json_device); -d, `job`, `list` and --gpus N get the whole host snapshot with pod
//   names, labels and the job table from arena_load_json_gpu.  --loader host: the C++ loader + upload.
// Errors are printed to stdout and the exit status is 1, as in the reference
// (`fmt.Println(err); os.Exit(1)`, top_node.go:34-48).
#include "../../include/arena_top.h"

#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <ctime>
#include <fstream>
#include <sstream>
#include <string>
#include <vector>

static const char kTopLong[] =
    "Display Resource (GPU) usage.\n\nAvailable Commands:\n  node        Display Resource (GPU) usage of nodes\n"
    "  job         Display Resource (GPU) usage of pods\n    \n\nUsage:\n  arena-top [command]\n";

static bool slurp(const char *path, std::string &out)
{
    std::ifstream f(path, std::ios::binary);
    if (!f) return false;
    std::ostringstream ss;
    ss << f.rdbuf();
    out = ss.str();
    return true;
}

static int die(const std::string &msg)
{
    printf("%s\n", msg.c_str());
    return 1;
}

static int die_arena() { return die(arena_last_error()); }

int main(int argc, char **argv)
{
    if (argc < 2 || (strcmp(argv[1], "node") != 0 && strcmp(argv[1], "job") != 0 && strcmp(argv[1], "list") != 0)) {
        fputs(kTopLong, stdout);
        return argc < 2 ? 0 : 1;
    }
    const bool is_node = strcmp(argv[1], "node") == 0, is_list = strcmp(argv[1], "list") == 0;
    bool details = false, timing = false;
    const char *loader = nullptr;
    int device = 0, gpus = 1;
    long long now = (long long)time(nullptr);
    const char *pods_path = nullptr, *nodes_path = nullptr, *jobs_path = nullptr, *tfjobs_path = nullptr,
               *releases_path = nullptr;
    for (int i = 2; i < argc; i++) {
        const std::string a = argv[i];
        auto need = [&](const char **dst) { if (i + 1 >= argc) return false; *dst = argv[++i]; return true; };
        const char *tmp = nullptr;
        if (a == "-d" || a == "--details") details = true;
        else if (a == "--time") timing = true;
        else if (a == "--pods") { if (!need(&pods_path)) return die("flag needs an argument: --pods"); }
        else if (a == "--nodes") { if (!need(&nodes_path)) return die("flag needs an argument: --nodes"); }
        else if (a == "--jobs") { if (!need(&jobs_path)) return die("flag needs an argument: --jobs"); }
        else if (a == "--tfjobs") { if (!need(&tfjobs_path)) return die("flag needs an argument: --tfjobs"); }
        else if (a == "--releases") { if (!need(&releases_path)) return die("flag needs an argument: --releases"); }
        else if (a == "--device") { if (!need(&tmp)) return die("flag needs an argument: --device"); device = atoi(tmp); }
        else if (a == "--gpus") { if (!need(&tmp)) return die("flag needs an argument: --gpus"); gpus = atoi(tmp); if (gpus < 1 || gpus > 16) return die("--gpus must be 1..16"); }
        else if (a == "--loader") { if (!need(&loader)) return die("flag needs an argument: --loader"); }
        else if (a == "--now") { if (!need(&tmp)) return die("flag needs an argument: --now"); now = atoll(tmp); }
        else return die("unknown flag: " + a);
    }
    if (!pods_path || !nodes_path) return die("--pods and --nodes are required");
    if (!is_node && !releases_path) return die(is_list ? "list needs --releases (the output of `helm list`)" : "top job needs --releases (the output of `helm list`)");
    std::string pods, nodes, jobs, tfjobs, releases;
    if (!slurp(pods_path, pods)) return die(std::string("open ") + pods_path + ": no such file or directory");
    if (!slurp(nodes_path, nodes)) return die(std::string("open ") + nodes_path + ": no such file or directory");
    if (jobs_path && !slurp(jobs_path, jobs)) return die(std::string("open ") + jobs_path + ": no such file or directory");
    if (tfjobs_path && !slurp(tfjobs_path, tfjobs)) return die(std::string("open ") + tfjobs_path + ": no such file or directory");
    if (releases_path && !slurp(releases_path, releases)) return die(std::string("open ") + releases_path + ": no such file or directory");

    if (loader && strcmp(loader, "host") != 0 && strcmp(loader, "device") != 0) return die("--loader must be host or device");
    const bool device_loader = is_node && !details && gpus == 1 && !(loader && strcmp(loader, "host") == 0);
    const bool gpu_decode = !(loader && strcmp(loader, "host") == 0);      // every command decodes the PodList on the GPU by default
    if (device_loader) {
        if (arena_acct_init(device)) return die_arena();
        auto t0 = std::chrono::steady_clock::now();
        arena_host_snapshot *node_names = nullptr;
        arena_snapshot *dev = nullptr;
        arena_load_device_stats st;
        if (arena_load_json_device(pods.data(), pods.size(), nodes.data(), nodes.size(), &node_names, &dev, &st)) return die_arena();
        auto t1 = std::chrono::steady_clock::now();
        uint64_t counts[4];
        arena_snapshot_counts(dev, counts);
        std::vector<int64_t> total(counts[2] + 1), alloc(counts[2] + 1);
        int64_t cluster[2];
        char *text = nullptr;
        size_t len = 0;
        if (arena_top_node(dev, total.data(), alloc.data(), cluster, nullptr)) return die_arena();
        if (arena_format_top_node(node_names, total.data(), alloc.data(), nullptr, 0, &text, &len)) return die_arena();
        auto t2 = std::chrono::steady_clock::now();
        fwrite(text, 1, len, stdout);
        arena_free_text(text);
        if (timing) {
            auto ms = [](auto a, auto b) { return std::chrono::duration<double, std::milli>(b - a).count(); };
            char how[96];
            if (st.fell_back) snprintf(how, sizeof how, "host loader took over");
            else if (st.n_handed_back) snprintf(how, sizeof how, "parsed on the GPU, %llu items handed back to the host", (unsigned long long)st.n_handed_back);
            else snprintf(how, sizeof how, "parsed on the GPU");
            fprintf(stderr, "[arena-top] pods=%llu nodes=%llu  device load %.2f ms (%s; copy %.2f ms), query+format %.2f ms\n",
                    (unsigned long long)counts[0], (unsigned long long)counts[2], ms(t0, t1), how, st.ms_h2d, ms(t1, t2));
        }
        arena_snapshot_free(dev);
        arena_host_snapshot_free(node_names);
        arena_acct_shutdown();
        return 0;
    }

    // `-d`, `top job`, `list` (and --gpus N): the whole host snapshot -- names, job table -- with the PodList
    // decoded on the GPU (arena_load_json_gpu) unless --loader host
    if (gpus > 1) {
        std::vector<int> ids(gpus);
        for (int i = 0; i < gpus; i++) ids[i] = i;
        if (arena_acct_init_devices(gpus, ids.data())) return die_arena();
    } else if (arena_acct_init(device)) return die_arena();
    auto t0 = std::chrono::steady_clock::now();
    arena_host_snapshot *host = nullptr;
    arena_load_device_stats lst;
    memset(&lst, 0, sizeof lst);
    lst.fell_back = 1;
    const char *jobs_p = jobs_path ? jobs.data() : nullptr, *tf_p = tfjobs_path ? tfjobs.data() : nullptr,
               *rel_p = (!is_node && releases_path) ? releases.data() : nullptr;
    if (gpu_decode ? arena_load_json_gpu(pods.data(), pods.size(), nodes.data(), nodes.size(), jobs_p, jobs.size(), tf_p, tfjobs.size(),
                                         rel_p, releases.size(), is_node ? 0 : 1, &host, &lst)
                   : arena_load_json(pods.data(), pods.size(), nodes.data(), nodes.size(), jobs_p, jobs.size(), tf_p, tfjobs.size(),
                                     rel_p, releases.size(), is_node ? 0 : 1, &host))
        return die_arena();
    auto t1 = std::chrono::steady_clock::now();
    char how[96];
    if (lst.fell_back) snprintf(how, sizeof how, "host loader");
    else if (lst.n_handed_back) snprintf(how, sizeof how, "PodList parsed on the GPU, %llu items handed back to the host", (unsigned long long)lst.n_handed_back);
    else snprintf(how, sizeof how, "PodList parsed on the GPU");
    arena_snapshot_view view;
    arena_host_snapshot_view(host, &view);
    if (gpus > 1) {
        // one process, several GPUs: the loops of top_node.go:109-118 / top_job.go:53-66 split by key range
        arena_multi_snapshot *ms = nullptr;
        if (arena_multi_snapshot_upload(&view, is_node ? 0 : 1, &ms)) return die_arena();
        auto t2 = std::chrono::steady_clock::now();
        char *text = nullptr;
        size_t len = 0;
        if (is_node) {
            std::vector<int64_t> total(view.n_nodes + 1), alloc(view.n_nodes + 1), pod_gpu(details ? view.n_pods + 1 : 1);
            int64_t cluster[2];
            if (arena_top_node_multi(ms, total.data(), alloc.data(), cluster, details ? pod_gpu.data() : nullptr)) return die_arena();
            if (arena_format_top_node(host, total.data(), alloc.data(), details ? pod_gpu.data() : nullptr, details ? 1 : 0, &text, &len))
                return die_arena();
        } else {
            const size_t J = view.n_jobs;
            std::vector<int64_t> req(J + 1), alloc(J + 1);
            std::vector<uint32_t> order(J + 1), latest(J + 1);
            int64_t totals[2];
            if (arena_top_job_multi(ms, req.data(), alloc.data(), order.data(), latest.data(), totals)) return die_arena();
            if (is_list && arena_job_order_by_age(host, order.data())) return die_arena();
            if (arena_format_job_list(host, req.data(), alloc.data(), order.data(), latest.data(), now, is_list ? 0 : 1, &text, &len)) return die_arena();
        }
        auto t3 = std::chrono::steady_clock::now();
        fwrite(text, 1, len, stdout);
        arena_free_text(text);
        if (timing) {
            auto ms_ = [](auto a, auto b) { return std::chrono::duration<double, std::milli>(b - a).count(); };
            fprintf(stderr, "[arena-top] gpus=%d pods=%llu nodes=%llu jobs=%llu  load %.2f ms (%s), upload %.2f ms, query+format %.2f ms\n",
                    gpus, (unsigned long long)view.n_pods, (unsigned long long)view.n_nodes, (unsigned long long)view.n_jobs,
                    ms_(t0, t1), how, ms_(t1, t2), ms_(t2, t3));
        }
        arena_multi_snapshot_free(ms);
        arena_host_snapshot_free(host);
        arena_acct_shutdown();
        return 0;
    }
    arena_snapshot *dev = nullptr;
    if (arena_snapshot_upload(&view, &dev)) return die_arena();
    auto t2 = std::chrono::steady_clock::now();

    char *text = nullptr;
    size_t len = 0;
    if (is_node) {
        std::vector<int64_t> total(view.n_nodes + 1), alloc(view.n_nodes + 1), pod_gpu(details ? view.n_pods + 1 : 1);
        int64_t cluster[2];
        if (arena_top_node(dev, total.data(), alloc.data(), cluster, details ? pod_gpu.data() : nullptr)) return die_arena();
        if (arena_format_top_node(host, total.data(), alloc.data(), details ? pod_gpu.data() : nullptr, details ? 1 : 0, &text, &len))
            return die_arena();
    } else {
        const size_t J = view.n_jobs;
        std::vector<int64_t> req(J + 1), alloc(J + 1);
        std::vector<uint32_t> order(J + 1), latest(J + 1);
        int64_t totals[2];
        if (arena_top_job(dev, req.data(), alloc.data(), order.data(), latest.data(), totals)) return die_arena();
        if (is_list && arena_job_order_by_age(host, order.data())) return die_arena();
        if (arena_format_job_list(host, req.data(), alloc.data(), order.data(), latest.data(), now, is_list ? 0 : 1, &text, &len)) return die_arena();
    }
    auto t3 = std::chrono::steady_clock::now();
    fwrite(text, 1, len, stdout);
    arena_free_text(text);
    if (timing) {
        auto ms = [](auto a, auto b) { return std::chrono::duration<double, std::milli>(b - a).count(); };
        fprintf(stderr, "[arena-top] pods=%llu nodes=%llu jobs=%llu  load %.2f ms (%s), upload %.2f ms, query+format %.2f ms\n",
                (unsigned long long)view.n_pods, (unsigned long long)view.n_nodes, (unsigned long long)view.n_jobs,
                ms(t0, t1), how, ms(t1, t2), ms(t2, t3));
    }
    arena_snapshot_free(dev);
    arena_host_snapshot_free(host);
    arena_acct_shutdown();
    return 0;
}
