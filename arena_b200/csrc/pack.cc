// Wide host snapshot -> packed host snapshot (include/arena_gpuacct.h: arena_packed_view):
// one byte per pod, one per container, one row range per node.  The container byte is the
// RESOLVED Quantity.Value() (amount and scale folded on the host, quantity.cc), so what the
// device expands and sums is what the wide path decodes and sums.
#include "host_snapshot.h"
#include "quantity.h"

#include <algorithm>
#include <cstdio>
#include <cstring>
#include <new>
#include <thread>
#include <vector>

using namespace arena_host;

struct arena_packed_snapshot {
    arena_packed_view v{};
    Column c_node_pod_begin, c_pod_meta, c_ctr_gpu, c_node_cap_val, c_node_cap_scale, c_node_ctr_begin, c_pieces;
};

extern "C" {

int arena_host_snapshot_pack(const arena_snapshot_view *h, int pinned, arena_packed_snapshot **out)
{
    if (!h || !out) return fail(ARENA_E_INVALID, "snapshot or out is NULL");
    *out = nullptr;
    const uint64_t P = h->n_pods, C = h->n_containers, N = h->n_nodes;
    if (P >= 0xFFFFFFFFull || C > 0xFFFFFFFFull || N >= 0xFFFFFFFFull) return fail(ARENA_E_TOO_LARGE, "snapshot exceeds 32-bit offsets");
    if (!h->pod_ctr_begin || (P && (!h->pod_node || !h->pod_phase)) || (C && (!h->ctr_val || !h->ctr_scale)) ||
        (N && (!h->node_cap_val || !h->node_cap_scale)))
        return fail(ARENA_E_INVALID, "snapshot columns are NULL");
    arena_packed_snapshot *p = new (std::nothrow) arena_packed_snapshot();
    if (!p) return fail(ARENA_E_NOMEM, "out of host memory");
    int rc;
    // sizes are rounded up so that the device can read whole 32-bit words past the end
    // (the value column is allocated once the widest value is known: a nibble or a byte each)
    if ((rc = p->c_node_pod_begin.alloc((N + 1) * 4, pinned != 0)) || (rc = p->c_pod_meta.alloc(P + 4, pinned != 0)) ||
        (rc = p->c_node_cap_val.alloc(N * 8, pinned != 0)) || (rc = p->c_node_cap_scale.alloc(N, pinned != 0))) {
        delete p;
        return rc;
    }
    uint32_t *begin = p->c_node_pod_begin.as<uint32_t>();
    uint8_t *meta = p->c_pod_meta.as<uint8_t>();
    for (uint64_t k = 0; k < 4; k++) meta[P + k] = 0;

    // One pass over the rows and one over the containers (the wide columns are read ONCE: the packing
    // is a memory-bound transform of 2.3 GB into 0.2 GB per 1e8 pods), each cut into one contiguous
    // range per host thread; a short second pass folds the container bytes into nibbles when every value
    // fits one.  A failure is reported for the LOWEST offending row / container, as a sequential pass would.
    struct Err { uint64_t at = ~0ull; int code = 0; char text[160] = ""; };
    unsigned T = std::thread::hardware_concurrency();
    if (T == 0) T = 1;
    if (T > 32) T = 32;
    if (P + C < (1u << 20)) T = 1;
    std::vector<Err> errs(T);
    auto run = [&](auto &&fn) {
        std::vector<std::thread> pool;
        for (unsigned t = 1; t < T; t++) pool.emplace_back(fn, t);
        fn(0u);
        for (std::thread &th : pool) th.join();
    };
    auto first_error = [&]() -> int {
        const Err *worst = nullptr;
        for (const Err &e : errs)
            if (e.code && (!worst || e.at < worst->at)) worst = &e;
        if (!worst) return 0;
        return fail(worst->code, "%s", worst->text);
    };
#define PACK_FAIL(T_, AT_, ...)                                                 \
    do {                                                                        \
        Err &e_ = errs[T_];                                                     \
        if (!e_.code) { e_.code = ARENA_E_INVALID; e_.at = (AT_); snprintf(e_.text, sizeof e_.text, __VA_ARGS__); } \
        return;                                                                 \
    } while (0)

    // rows must be grouped by node, in node order, the pods of no node last
    auto key_of = [&](uint64_t r) { const uint32_t n = h->pod_node[r]; return n == ARENA_NONE ? N : (uint64_t)n; };
    run([&](unsigned t) {
        const uint64_t r0 = P * t / T, r1 = P * (t + 1) / T;
        // nodes whose first row falls into this range: every node after the previous row's, up to each row's own
        uint64_t node = r0 == 0 ? 0 : std::min<uint64_t>(key_of(r0 - 1), N) + 1, prev_key = r0 == 0 ? 0 : key_of(r0 - 1);
        for (uint64_t r = r0; r < r1; r++) {
            const uint32_t n = h->pod_node[r];
            const uint64_t key = n == ARENA_NONE ? N : (uint64_t)n;
            if (n != ARENA_NONE && n >= N) PACK_FAIL(t, r, "pod row %llu names node %u of %llu", (unsigned long long)r, n, (unsigned long long)N);
            if (key < prev_key) PACK_FAIL(t, r, "no packed form: rows are not grouped by node (row %llu)", (unsigned long long)r);
            prev_key = key;
            while (node <= key) begin[node++] = (uint32_t)r;
            const uint32_t cb = h->pod_ctr_begin[r], ce = h->pod_ctr_begin[r + 1];
            if (ce < cb || ce > C) PACK_FAIL(t, r, "pod_ctr_begin is not a monotonic CSR at row %llu", (unsigned long long)r);
            if (ce - cb > 15) PACK_FAIL(t, r, "no packed form: pod row %llu has %u containers (15 at most)", (unsigned long long)r, ce - cb);
            const uint8_t ph = h->pod_phase[r];
            uint8_t code;
            if (ph <= ARENA_PHASE_UNKNOWN) code = ph;
            else if (ph == ARENA_PHASE_OTHER) code = 15;
            else PACK_FAIL(t, r, "no packed form: unknown phase code %u at row %llu", ph, (unsigned long long)r);
            meta[r] = (uint8_t)(code | ((ce - cb) << 4));
        }
    });
    if ((rc = first_error())) { delete p; return rc; }
    {   // nodes after the last row's
        uint64_t node = P ? std::min<uint64_t>(key_of(P - 1), N) + 1 : 0;
        while (node <= N) begin[node++] = (uint32_t)P;
    }
    if (P && h->pod_ctr_begin[0] != 0) { delete p; return fail(ARENA_E_INVALID, "pod_ctr_begin[0] != 0"); }
    if (h->pod_ctr_begin[P] != C) { delete p; return fail(ARENA_E_INVALID, "pod_ctr_begin[P] != n_containers"); }
    // containers: Quantity.Value() of each, one byte each into a scratch column; the widest value then
    // decides between a byte and a nibble each
    Column bytes;                                        // (plain memory, freed on return)
    if ((rc = bytes.alloc(C + 8, false))) { delete p; return rc; }
    uint8_t *wide = bytes.as<uint8_t>();
    std::vector<uint8_t> tops(T, 0);
    run([&](unsigned t) {
        const uint64_t c0 = C * t / T, c1 = C * (t + 1) / T;
        uint8_t top = 0;
        for (uint64_t c = c0; c < c1; c++) {
            const int8_t sc = h->ctr_scale[c];
            const int64_t v = sc == 0 ? h->ctr_val[c] : amount_value(h->ctr_val[c], sc);       // Quantity.Value()
            if (v < 0 || v > 255) PACK_FAIL(t, c, "no packed form: container %llu holds %lld GPUs (0..255)", (unsigned long long)c, (long long)v);
            wide[c] = (uint8_t)v;
            top |= (uint8_t)v;
        }
        tops[t] = top;
    });
    if ((rc = first_error())) { delete p; return rc; }
    uint8_t top = 0;
    for (uint8_t x : tops) top |= x;
    p->v.ctr_bits = top < 16 ? 4 : 8;
    // sizes are rounded up so that the device can read whole 32-bit words past the end
    const uint64_t pairs = (C + 1) / 2, gpu_bytes = top < 16 ? pairs : C;
    if ((rc = p->c_ctr_gpu.alloc(gpu_bytes + 8, pinned != 0))) { delete p; return rc; }
    uint8_t *gpu = p->c_ctr_gpu.as<uint8_t>();
    for (uint64_t k = gpu_bytes; k < gpu_bytes + 8; k++) gpu[k] = 0;
    if (C & 1) wide[C] = 0;                               // the odd container's partner
    if (top < 16) {                          // every container fits a nibble: two per byte, container 2k in the low nibble
        run([&](unsigned t) {
            const uint64_t k0 = pairs * t / T, k1 = pairs * (t + 1) / T;
            for (uint64_t k = k0; k < k1; k++) gpu[k] = (uint8_t)(wide[2 * k] | (wide[2 * k + 1] << 4));
        });
    } else {
        run([&](unsigned t) {
            const uint64_t c0 = C * t / T, c1 = C * (t + 1) / T;
            if (c1 > c0) std::memcpy(gpu + c0, wide + c0, c1 - c0);
        });
    }
#undef PACK_FAIL
    for (uint64_t n = 0; n < N; n++) {
        p->c_node_cap_val.as<int64_t>()[n] = h->node_cap_val[n];
        p->c_node_cap_scale.as<int8_t>()[n] = h->node_cap_scale[n];
    }
    // container offsets per node, and the big nodes cut into pieces
    const uint32_t kPiece = 2048;
    uint64_t n_pieces = 0;
    for (uint64_t n = 0; n < N; n++) {
        const uint64_t rows = begin[n + 1] - begin[n];
        if (rows > kPiece) n_pieces += (rows + kPiece - 1) / kPiece;
    }
    if ((rc = p->c_node_ctr_begin.alloc((N + 1) * 4, pinned != 0)) || (rc = p->c_pieces.alloc(n_pieces * 16, pinned != 0))) {
        delete p;
        return rc;
    }
    uint32_t *ncb = p->c_node_ctr_begin.as<uint32_t>(), *pc = p->c_pieces.as<uint32_t>();
    for (uint64_t n = 0; n <= N; n++) ncb[n] = h->pod_ctr_begin[begin[n]];
    for (uint64_t n = 0, k = 0; n < N; n++) {
        if (begin[n + 1] - begin[n] <= kPiece) continue;
        for (uint32_t r = begin[n]; r < begin[n + 1]; r += kPiece, k++) {
            pc[4 * k] = (uint32_t)n;
            pc[4 * k + 1] = r;
            pc[4 * k + 2] = begin[n + 1] - r > kPiece ? r + kPiece : begin[n + 1];
            pc[4 * k + 3] = h->pod_ctr_begin[r];
        }
    }
    p->v.node_ctr_begin = ncb;
    p->v.pieces = pc;
    p->v.n_pieces = n_pieces;
    p->v.n_pods = P; p->v.n_containers = C; p->v.n_nodes = N;
    p->v.node_pod_begin = begin; p->v.pod_meta = meta; p->v.ctr_gpu = gpu;
    p->v.node_cap_val = p->c_node_cap_val.as<int64_t>();
    p->v.node_cap_scale = p->c_node_cap_scale.as<int8_t>();
    *out = p;
    return ARENA_OK;
}

int arena_packed_snapshot_view(const arena_packed_snapshot *p, arena_packed_view *out)
{
    if (!p || !out) return fail(ARENA_E_INVALID, "packed snapshot or out is NULL");
    *out = p->v;
    return ARENA_OK;
}

void arena_packed_snapshot_free(arena_packed_snapshot *p) { delete p; }

}  // extern "C"
