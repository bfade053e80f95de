// One PodList item -> the `top node` columns, as straight-line code that runs the
// same on the host and on the device (one GPU thread per pod).
//
// This is the device half of the wire-format step: it replaces,
// for the fields the accounting reads, the json-iterator decode of a v1.Pod
// (/root/reference: cmd/arena/commands/pod_helper.go:15-43,
// vendor/k8s.io/apimachinery/pkg/runtime/serializer/json/json.go:104-113, 171) and
// Quantity.UnmarshalJSON (vendor/k8s.io/apimachinery/pkg/api/resource/quantity.go:648-667).
//
// Contract: the scanner VALIDATES the item's JSON and extracts
//     spec.nodeName, status.phase, spec.containers[*].resources.limits[GPU key]
// exactly as the host loader (loader.cc extract_pod) does, or it FLAGS the item.
// A flag means "not handled here", never "invalid": the caller then runs the host
// loader, which is the authority for every error message and every exotic but legal
// spelling (escaped keys, duplicate keys, quantities with suffixes on a GPU key,
// RFC 3339 times with offsets, more than kInlineCtrs containers ...).  So the only
// thing this file must guarantee is: flag == 0  =>  the outputs equal the host's.
#pragma once
#include <cstddef>
#include <cstdint>
#include <type_traits>

#if defined(__CUDACC__)
#define ARENA_HD __host__ __device__
#else
#define ARENA_HD
#endif

namespace arena_scan {

constexpr int kInlineCtrs = 8;        // containers per pod handled without the host
constexpr int kMaxLevels = 40;        // nesting inside one item handled without the host
constexpr uint32_t kNoNode = 0xFFFFFFFFu;

// why an item was handed back to the host (diagnostics only; any non-zero value means the same)
enum Flag : uint32_t {
    kOk = 0,
    kFlagSyntax = 1,          // not well-formed here: the host reports the exact error
    kFlagShape = 2,           // a field has a type the host rejects or treats specially
    kFlagEscapedKey = 3,      // a key with a backslash in a place where key names matter
    kFlagDuplicateKey = 4,    // last-wins is left to the host
    kFlagQuantity = 5,        // a quantity outside the plain grammar, or a GPU quantity off the reference's fast path
    kFlagTime = 6,            // creationTimestamp not in the API server's canonical form
    kFlagTooDeep = 7,
    kFlagTooManyContainers = 8,
    kFlagEscapedValue = 9,    // nodeName / phase with a backslash
    kFlagSeparator = 10,      // something other than one comma between two items
    kFlagKeyHash = 11,        // a key with the length and hash of a known one but other bytes
    kFlagTooManyFields = 12,  // more deferred checks than the per-item list holds
};

// ---- node name -> id: open addressing, exact (bytes are compared) ----------------------
struct NodeSlot {
    uint64_t hash;            // 0 = empty
    uint32_t id;
    uint32_t name_len;
    uint64_t name_off;        // into NodeTable::names
};

struct NodeTable {
    const NodeSlot *slots;
    uint64_t mask;            // slot count - 1 (a power of two)
    const unsigned char *names;
};

ARENA_HD inline uint64_t hash_bytes(const unsigned char *p, size_t n)
{
    uint64_t h = 0xcbf29ce484222325ull;                 // FNV-1a
    for (size_t k = 0; k < n; k++) { h ^= p[k]; h *= 0x100000001b3ull; }
    return h | 0x8000000000000000ull;                   // never 0
}

ARENA_HD inline uint32_t find_node(const NodeTable &nt, const unsigned char *p, size_t n)
{
    const uint64_t h = hash_bytes(p, n);
    for (uint64_t idx = h & nt.mask;; idx = (idx + 1) & nt.mask) {
        const NodeSlot s = nt.slots[idx];
        if (s.hash == 0) return kNoNode;
        if (s.hash == h && s.name_len == n) {
            const unsigned char *q = nt.names + s.name_off;
            size_t k = 0;
            while (k < n && q[k] == p[k]) k++;
            if (k == n) return s.id;
        }
    }
}

struct PodOut {
    uint32_t flag;
    uint32_t node;                    // node id or kNoNode
    uint32_t n_ctrs;
    uint8_t phase;                    // ARENA_PHASE_*
    int64_t ctr_val[kInlineCtrs];     // column encoding: int64Amount.value ...
    int8_t ctr_scale[kInlineCtrs];    // ... and .scale (resource/amount.go:57-60)
};

// What `top job`, `top node -d` and `arena list` read besides the accounting columns, as places in
// the document: the host copies the bytes out (loader.cc extract_pod reads the same fields from its
// tree).  Filled only when the caller asks for it (scan_pod_item<..., true>).
enum MetaField : int {
    kMetaName, kMetaNamespace, kMetaHostIP, kMetaCreated,
    kMetaRelease, kMetaApp, kMetaGroup, kMetaReplicaType, kMetaReplicaIndex, kMetaFields
};
struct PodMeta {
    uint32_t off[kMetaFields];        // absolute byte offset of the string body
    uint16_t len[kMetaFields];
    uint16_t present;                 // bit f: the key is there (a label whose value is null counts: it reads as "")
    uint8_t owner_is_job;             // some ownerReference has kind "Job" (trainer_mpi.go:247-279)
};

namespace detail {

enum Role : uint8_t {
    R_SKIP_OBJ, R_SKIP_ARR, R_ITEM, R_META, R_LABELS, R_OWNERS, R_OWNER, R_SPEC, R_CTRS, R_CTR, R_RES, R_LIMITS,
    R_QMAP, R_ICTRS, R_ICTR, R_IRES, R_STATUS, R_COUNT
};

ARENA_HD inline bool role_is_array(uint8_t r) { return r == R_SKIP_ARR || r == R_OWNERS || r == R_CTRS || r == R_ICTRS; }

// key ids: bit positions in the per-role "seen" mask (0 = a key nobody looks at)
enum Key : uint8_t {
    K_OTHER = 0, K_METADATA, K_SPEC, K_STATUS,                                  // item
    K_NAME, K_NAMESPACE, K_LABELS, K_CREATED, K_OWNERS,                         // metadata
    K_L_RELEASE, K_L_APP, K_L_GROUP, K_L_RTYPE, K_L_RINDEX,                     // labels
    K_KIND,                                                                     // ownerReference
    K_NODENAME, K_CONTAINERS, K_INITCONTAINERS,                                 // spec
    K_RESOURCES,                                                                // container
    K_LIMITS, K_REQUESTS,                                                       // resources
    K_GPU, K_GPU_DEPRECATED, K_QUANTITY,                                        // resource maps
    K_PHASE, K_HOSTIP,                                                          // status
    K_ELEMENT                                                                   // array element
};

ARENA_HD inline bool bytes_are(const unsigned char *p, size_t n, const char *lit)
{
    size_t k = 0;
    for (; k < n; k++)
        if (lit[k] == 0 || (unsigned char)lit[k] != p[k]) return false;
    return lit[k] == 0;
}

// Keys are recognised by (length, FNV-1a of the bytes), both gathered while the lexer walks
// the string, so the hot loop never goes back to memory; every recognised key is then
// compared byte for byte (key_text) before its value is trusted.
ARENA_HD constexpr uint32_t fnv32(const char *s)
{
    uint32_t h = 2166136261u;
    for (int k = 0; s[k]; k++) h = (h ^ (uint8_t)s[k]) * 16777619u;
    return h;
}
ARENA_HD constexpr uint32_t lit_len(const char *s)
{
    uint32_t n = 0;
    while (s[n]) n++;
    return n;
}

ARENA_HD inline const char *key_text(uint8_t key)
{
    switch (key) {
    case K_METADATA: return "metadata";
    case K_SPEC: return "spec";
    case K_STATUS: return "status";
    case K_NAME: return "name";
    case K_NAMESPACE: return "namespace";
    case K_LABELS: return "labels";
    case K_CREATED: return "creationTimestamp";
    case K_OWNERS: return "ownerReferences";
    case K_L_RELEASE: return "release";
    case K_L_APP: return "app";
    case K_L_GROUP: return "group_name";
    case K_L_RTYPE: return "tf-replica-type";
    case K_L_RINDEX: return "tf-replica-index";
    case K_KIND: return "kind";
    case K_NODENAME: return "nodeName";
    case K_CONTAINERS: return "containers";
    case K_INITCONTAINERS: return "initContainers";
    case K_RESOURCES: return "resources";
    case K_LIMITS: return "limits";
    case K_REQUESTS: return "requests";
    case K_GPU: return "nvidia.com/gpu";                               // const.go:7
    case K_GPU_DEPRECATED: return "alpha.kubernetes.io/nvidia-gpu";    // const.go:9
    case K_PHASE: return "phase";
    case K_HOSTIP: return "hostIP";
    default: return "";
    }
}

#define ARENA_KEY_IS(lit) \
    (n == std::integral_constant<uint32_t, lit_len(lit)>::value && h == std::integral_constant<uint32_t, fnv32(lit)>::value)
ARENA_HD inline uint8_t key_id_hashed(uint8_t role, uint32_t n, uint32_t h)
{
    switch (role) {
    case R_ITEM:
        if (ARENA_KEY_IS("metadata")) return K_METADATA;
        if (ARENA_KEY_IS("spec")) return K_SPEC;
        if (ARENA_KEY_IS("status")) return K_STATUS;
        return K_OTHER;
    case R_META:
        if (ARENA_KEY_IS("name")) return K_NAME;
        if (ARENA_KEY_IS("namespace")) return K_NAMESPACE;
        if (ARENA_KEY_IS("labels")) return K_LABELS;
        if (ARENA_KEY_IS("creationTimestamp")) return K_CREATED;
        if (ARENA_KEY_IS("ownerReferences")) return K_OWNERS;
        return K_OTHER;
    case R_LABELS:
        if (ARENA_KEY_IS("release")) return K_L_RELEASE;
        if (ARENA_KEY_IS("app")) return K_L_APP;
        if (ARENA_KEY_IS("group_name")) return K_L_GROUP;
        if (ARENA_KEY_IS("tf-replica-type")) return K_L_RTYPE;
        if (ARENA_KEY_IS("tf-replica-index")) return K_L_RINDEX;
        return K_OTHER;
    case R_OWNER:
        return ARENA_KEY_IS("kind") ? K_KIND : K_OTHER;
    case R_SPEC:
        if (ARENA_KEY_IS("nodeName")) return K_NODENAME;
        if (ARENA_KEY_IS("containers")) return K_CONTAINERS;
        if (ARENA_KEY_IS("initContainers")) return K_INITCONTAINERS;
        return K_OTHER;
    case R_CTR:
    case R_ICTR:
        return ARENA_KEY_IS("resources") ? K_RESOURCES : K_OTHER;
    case R_RES:
    case R_IRES:
        if (ARENA_KEY_IS("limits")) return K_LIMITS;
        if (ARENA_KEY_IS("requests")) return K_REQUESTS;
        return K_OTHER;
    case R_LIMITS:
        if (ARENA_KEY_IS("nvidia.com/gpu")) return K_GPU;
        if (ARENA_KEY_IS("alpha.kubernetes.io/nvidia-gpu")) return K_GPU_DEPRECATED;
        return K_QUANTITY;
    case R_QMAP:
        return K_QUANTITY;
    case R_STATUS:
        if (ARENA_KEY_IS("phase")) return K_PHASE;
        if (ARENA_KEY_IS("hostIP")) return K_HOSTIP;
        return K_OTHER;
    default:
        return K_OTHER;
    }
}
#undef ARENA_KEY_IS

ARENA_HD inline bool is_digit(unsigned char c) { return c >= '0' && c <= '9'; }
ARENA_HD inline bool is_ws(unsigned char c) { return c == ' ' || c == '\t' || c == '\n' || c == '\r'; }

// A SUFFICIENT condition for ParseQuantity (quantity.go:159-390) to succeed:
//   [+-]? digits+ ( '.' digits* )? ( "" | n u m k M G T P E | Ki Mi Gi Ti Pi Ei | [eE][+-]?digits{1,9} )
ARENA_HD inline bool quantity_plain(const unsigned char *p, size_t n)
{
    if (n == 0 || n > 64) return false;
    size_t i = 0;
    if (p[0] == '+' || p[0] == '-') i++;
    size_t d0 = i;
    while (i < n && is_digit(p[i])) i++;
    if (i == d0) return false;
    if (i < n && p[i] == '.') {
        i++;
        while (i < n && is_digit(p[i])) i++;
    }
    const size_t left = n - i;
    if (left == 0) return true;
    const unsigned char a = p[i];
    if (left == 1) return a == 'n' || a == 'u' || a == 'm' || a == 'k' || a == 'M' || a == 'G' || a == 'T' || a == 'P' || a == 'E';
    if (left == 2 && p[i + 1] == 'i') return a == 'K' || a == 'M' || a == 'G' || a == 'T' || a == 'P' || a == 'E';
    if (a != 'e' && a != 'E') return false;
    size_t j = i + 1;
    if (p[j] == '+' || p[j] == '-') j++;
    const size_t e0 = j;
    while (j < n && is_digit(p[j])) j++;
    return j == n && j > e0 && j - e0 <= 9;
}

// A GPU quantity on the reference's FAST path (resource/quantity.go:276-343: sign, digits, one
// decimal point, a suffix from suffix.go:88-106 or an e/E exponent; at most 18 significant digits;
// scale >= -9; mantissa product inside int64) -> int64Amount{value, scale}, the column encoding the
// host loader produces (quantity.cc: parse_quantity + column_encoding).  Anything else -- the inf.Dec
// slow path, scales beyond +-18, malformed text -- returns false and the item is handed to the host.
ARENA_HD inline bool quantity_fast(const unsigned char *s, uint32_t n, int64_t &amount, int8_t &scale_out)
{
    amount = 0;
    scale_out = 0;
    if (n == 0 || n > 40) return false;
    uint32_t pos = 0;
    bool positive = true;
    if (s[0] == '-') { positive = false; pos = 1; }
    else if (s[0] == '+') pos = 1;
    if (pos >= n) return false;                                // a bare sign: the host decides
    const uint32_t zeros_from = pos;
    while (pos < n && s[pos] == '0') pos++;                    // leading zeros
    if (pos >= n) return true;                                 // "0", "000", "-0": zero
    const bool stripped = pos > zeros_from;
    uint64_t v = 0;
    uint32_t num_len = 0, denom_len = 0;
    while (pos < n && is_digit(s[pos])) { if (num_len < 19) v = v * 10 + (uint64_t)(s[pos] - '0'); num_len++; pos++; }
    if (pos < n && s[pos] == '.') {
        pos++;
        while (pos < n && is_digit(s[pos])) { if (num_len + denom_len < 19) v = v * 10 + (uint64_t)(s[pos] - '0'); denom_len++; pos++; }
    }
    if (!stripped && num_len == 0 && denom_len == 0) return false;   // no digit at all ("m", "."): the host decides
    const uint32_t digits = (num_len ? num_len : 1u) + denom_len;   // an empty numerator reads as "0"
    // suffix
    int32_t exponent = 0;
    bool binary = false;
    const uint32_t left = n - pos;
    if (left == 1) {
        switch (s[pos]) {
        case 'n': exponent = -9; break;
        case 'u': exponent = -6; break;
        case 'm': exponent = -3; break;
        case 'k': exponent = 3; break;
        case 'M': exponent = 6; break;
        case 'G': exponent = 9; break;
        case 'T': exponent = 12; break;
        case 'P': exponent = 15; break;
        case 'E': exponent = 18; break;
        default: return false;
        }
    } else if (left == 2 && s[pos + 1] == 'i') {
        binary = true;
        switch (s[pos]) {
        case 'K': exponent = 10; break;
        case 'M': exponent = 20; break;
        case 'G': exponent = 30; break;
        case 'T': exponent = 40; break;
        case 'P': exponent = 50; break;
        case 'E': exponent = 60; break;
        default: return false;
        }
    } else if (left >= 2) {
        if (s[pos] != 'e' && s[pos] != 'E') return false;
        uint32_t j = pos + 1;
        bool eneg = false;
        if (s[j] == '+' || s[j] == '-') { eneg = s[j] == '-'; j++; }
        if (j >= n || n - j > 3) return false;                // up to three exponent digits: enough for every fast-path scale
        int32_t e = 0;
        for (; j < n; j++) {
            if (!is_digit(s[j])) return false;
            e = e * 10 + (int32_t)(s[j] - '0');
        }
        exponent = eneg ? -e : e;
    }
    int64_t scale = 0;
    uint64_t mantissa = 1;
    int32_t precision;
    if (binary) {
        if (denom_len) return false;                           // fractional binary-SI: slow path
        mantissa = 1ull << exponent;
        precision = 15 - (int32_t)(num_len ? num_len : 1u) - (exponent * 3) / 10 - 1;
    } else {
        scale = exponent;
        precision = 18 - (int32_t)digits;
    }
    if (precision < 0) return false;
    scale -= (int64_t)denom_len;
    if (scale < -9 || scale > 18) return false;
    // int64Multiply(v, mantissa): v < 10^18 < 2^60
    if (v != 0 && mantissa != 1 && v > 0x7fffffffffffffffull / mantissa) return false;
    const uint64_t r = v * mantissa;
    amount = positive ? (int64_t)r : -(int64_t)r;
    scale_out = (int8_t)scale;
    return true;
}

// The API server's creationTimestamp: "YYYY-MM-DDTHH:MM:SSZ" with the field ranges
// time.Parse(RFC3339) checks first; everything else goes to the host.
ARENA_HD inline bool time_canonical(const unsigned char *p, size_t n)
{
    if (n != 20) return false;
    if (p[4] != '-' || p[7] != '-' || p[10] != 'T' || p[13] != ':' || p[16] != ':' || p[19] != 'Z') return false;
    const int at[] = {0, 1, 2, 3, 5, 6, 8, 9, 11, 12, 14, 15, 17, 18};
    for (int k : at)
        if (!is_digit(p[k])) return false;
    const int Y = (p[0] - '0') * 1000 + (p[1] - '0') * 100 + (p[2] - '0') * 10 + (p[3] - '0');
    const int M = (p[5] - '0') * 10 + (p[6] - '0'), D = (p[8] - '0') * 10 + (p[9] - '0');
    const int h = (p[11] - '0') * 10 + (p[12] - '0'), mi = (p[14] - '0') * 10 + (p[15] - '0');
    const int s = (p[17] - '0') * 10 + (p[18] - '0');
    if (M < 1 || M > 12 || D < 1 || h > 23 || mi > 59 || s > 59) return false;
    const bool leap = (Y % 4 == 0 && Y % 100 != 0) || Y % 400 == 0;
    const int dim = M == 2 ? (leap ? 29 : 28) : (M == 4 || M == 6 || M == 9 || M == 11) ? 30 : 31;
    return D <= dim;                                       // Go rejects "day out of range": leave those to the host
}

ARENA_HD inline uint8_t phase_code(const unsigned char *p, size_t n)
{
    if (n == 0) return 0;                                  // ARENA_PHASE_EMPTY
    if (bytes_are(p, n, "Pending")) return 1;
    if (bytes_are(p, n, "Running")) return 2;
    if (bytes_are(p, n, "Succeeded")) return 3;
    if (bytes_are(p, n, "Failed")) return 4;
    if (bytes_are(p, n, "Unknown")) return 5;
    return 255;                                            // ARENA_PHASE_OTHER
}

// ---- the lexer: a table-driven automaton, one byte per step ---------------------------------
// Every thread of a warp runs the SAME few instructions per byte (class lookup, transition
// lookup), whatever token it is in; only token boundaries take a (short) divergent path.
// Acceptance is that of json.cc's recursive parser.
enum Class : uint8_t {
    C_SPACE, C_WHITE, C_LCURB, C_RCURB, C_LSQRB, C_RSQRB, C_COLON, C_COMMA, C_QUOTE, C_BACKS, C_SLASH, C_PLUS, C_MINUS,
    C_POINT, C_ZERO, C_DIGIT, C_A, C_B, C_LOW_E, C_F, C_L, C_N, C_R, C_S, C_T, C_U, C_HEXLET, C_E, C_ETC, C_CTRL, C_COUNT
};

enum LexState : uint8_t {
    S_VAL, S_ARR0, S_OBJ0, S_KEY, S_COLON, S_AFTER, S_STR, S_ESC, S_U1, S_U2, S_U3, S_U4, S_MINUS, S_ZERO, S_INT,
    S_FRAC0, S_FRAC, S_EXP0, S_EXPS, S_EXP, S_T1, S_T2, S_T3, S_F1, S_F2, S_F3, S_F4, S_N1, S_N2, S_N3,
    S_KSTR, S_KESC, S_KU1, S_KU2, S_KU3, S_KU4,                          // a KEY string: same rules as S_STR, other exits
    S_COUNT
};

enum Action : uint8_t {
    A_NONE, A_ERR, A_OPEN, A_CLOSE, A_COMMA, A_KEY_BEGIN, A_KEY_END, A_STR_BEGIN, A_STR_END, A_ESCAPE, A_NUM_BEGIN, A_LIT_END,
    A_NUM_ENDED = 0x80                   // flag on the action of the byte that follows a number
};

struct LexTables {
    uint8_t cls[256];
    uint16_t next[S_COUNT * C_COUNT];          // action << 8 | next state
};

constexpr uint16_t lex_entry(uint8_t action, uint8_t state) { return (uint16_t)((uint16_t)action << 8 | state); }

constexpr LexTables make_lex_tables()
{
    LexTables t{};
    for (int c = 0; c < 256; c++) t.cls[c] = c < 0x20 ? C_CTRL : C_ETC;
    t.cls[' '] = C_SPACE; t.cls['\t'] = C_WHITE; t.cls['\n'] = C_WHITE; t.cls['\r'] = C_WHITE;
    t.cls['{'] = C_LCURB; t.cls['}'] = C_RCURB; t.cls['['] = C_LSQRB; t.cls[']'] = C_RSQRB;
    t.cls[':'] = C_COLON; t.cls[','] = C_COMMA; t.cls['"'] = C_QUOTE; t.cls['\\'] = C_BACKS; t.cls['/'] = C_SLASH;
    t.cls['+'] = C_PLUS; t.cls['-'] = C_MINUS; t.cls['.'] = C_POINT; t.cls['0'] = C_ZERO;
    for (int c = '1'; c <= '9'; c++) t.cls[c] = C_DIGIT;
    t.cls['b'] = C_B; t.cls['f'] = C_F; t.cls['n'] = C_N; t.cls['r'] = C_R; t.cls['t'] = C_T; t.cls['u'] = C_U;
    t.cls['a'] = C_A; t.cls['c'] = C_HEXLET; t.cls['d'] = C_HEXLET; t.cls['l'] = C_L; t.cls['s'] = C_S;
    t.cls['A'] = C_HEXLET; t.cls['B'] = C_HEXLET; t.cls['C'] = C_HEXLET; t.cls['D'] = C_HEXLET; t.cls['F'] = C_HEXLET;
    t.cls['e'] = C_LOW_E; t.cls['E'] = C_E;

    for (int k = 0; k < S_COUNT * C_COUNT; k++) t.next[k] = lex_entry(A_ERR, S_AFTER);
    auto at = [&](int st, int c) -> uint16_t & { return t.next[st * C_COUNT + c]; };
    // between tokens
    const int gaps[] = {S_VAL, S_ARR0, S_OBJ0, S_KEY, S_COLON, S_AFTER};
    for (int st : gaps) { at(st, C_SPACE) = lex_entry(A_NONE, (uint8_t)st); at(st, C_WHITE) = lex_entry(A_NONE, (uint8_t)st); }
    const int value_starts[] = {S_VAL, S_ARR0};
    for (int st : value_starts) {
        at(st, C_LCURB) = lex_entry(A_OPEN, S_OBJ0);
        at(st, C_LSQRB) = lex_entry(A_OPEN, S_ARR0);
        at(st, C_QUOTE) = lex_entry(A_STR_BEGIN, S_STR);
        at(st, C_MINUS) = lex_entry(A_NUM_BEGIN, S_MINUS);
        at(st, C_ZERO) = lex_entry(A_NUM_BEGIN, S_ZERO);
        at(st, C_DIGIT) = lex_entry(A_NUM_BEGIN, S_INT);
        at(st, C_T) = lex_entry(A_NONE, S_T1);
        at(st, C_F) = lex_entry(A_NONE, S_F1);
        at(st, C_N) = lex_entry(A_NONE, S_N1);
    }
    // true / false / null: the last letter ends the value
    at(S_T1, C_R) = lex_entry(A_NONE, S_T2); at(S_T2, C_U) = lex_entry(A_NONE, S_T3); at(S_T3, C_LOW_E) = lex_entry(A_LIT_END, S_AFTER);
    at(S_F1, C_A) = lex_entry(A_NONE, S_F2); at(S_F2, C_L) = lex_entry(A_NONE, S_F3); at(S_F3, C_S) = lex_entry(A_NONE, S_F4);
    at(S_F4, C_LOW_E) = lex_entry(A_LIT_END, S_AFTER);
    at(S_N1, C_U) = lex_entry(A_NONE, S_N2); at(S_N2, C_L) = lex_entry(A_NONE, S_N3); at(S_N3, C_L) = lex_entry(A_LIT_END, S_AFTER);
    at(S_ARR0, C_RSQRB) = lex_entry(A_CLOSE, S_AFTER);
    at(S_OBJ0, C_RCURB) = lex_entry(A_CLOSE, S_AFTER);
    at(S_OBJ0, C_QUOTE) = lex_entry(A_KEY_BEGIN, S_KSTR);
    at(S_KEY, C_QUOTE) = lex_entry(A_KEY_BEGIN, S_KSTR);
    at(S_COLON, C_COLON) = lex_entry(A_NONE, S_VAL);
    at(S_AFTER, C_COMMA) = lex_entry(A_COMMA, S_VAL);
    at(S_AFTER, C_RCURB) = lex_entry(A_CLOSE, S_AFTER);
    at(S_AFTER, C_RSQRB) = lex_entry(A_CLOSE, S_AFTER);
    // strings (value strings: S_STR..S_U4, key strings: S_KSTR..S_KU4)
    for (int fam = 0; fam < 2; fam++) {
        const int str = fam ? S_KSTR : S_STR, esc = fam ? S_KESC : S_ESC, u1 = fam ? S_KU1 : S_U1;
        for (int c = 0; c < C_COUNT; c++) at(str, c) = lex_entry(A_NONE, (uint8_t)str);
        at(str, C_WHITE) = lex_entry(A_ERR, S_AFTER);                  // \t \n \r are control characters
        at(str, C_CTRL) = lex_entry(A_ERR, S_AFTER);
        at(str, C_QUOTE) = fam ? lex_entry(A_KEY_END, S_COLON) : lex_entry(A_STR_END, S_AFTER);
        at(str, C_BACKS) = lex_entry(A_ESCAPE, (uint8_t)esc);
        const int escapes[] = {C_QUOTE, C_BACKS, C_SLASH, C_B, C_F, C_N, C_R, C_T};
        for (int c : escapes) at(esc, c) = lex_entry(A_NONE, (uint8_t)str);
        at(esc, C_U) = lex_entry(A_NONE, (uint8_t)u1);
        const int hex[] = {C_ZERO, C_DIGIT, C_A, C_B, C_LOW_E, C_F, C_HEXLET, C_E};
        for (int c : hex) {
            at(u1, c) = lex_entry(A_NONE, (uint8_t)(u1 + 1));
            at(u1 + 1, c) = lex_entry(A_NONE, (uint8_t)(u1 + 2));
            at(u1 + 2, c) = lex_entry(A_NONE, (uint8_t)(u1 + 3));
            at(u1 + 3, c) = lex_entry(A_NONE, (uint8_t)str);
        }
    }
    // numbers: -? (0 | [1-9][0-9]*) (\.[0-9]+)? ([eE][+-]?[0-9]+)?
    at(S_MINUS, C_ZERO) = lex_entry(A_NONE, S_ZERO);
    at(S_MINUS, C_DIGIT) = lex_entry(A_NONE, S_INT);
    at(S_INT, C_ZERO) = lex_entry(A_NONE, S_INT);
    at(S_INT, C_DIGIT) = lex_entry(A_NONE, S_INT);
    const int int_ends[] = {S_ZERO, S_INT};
    for (int st : int_ends) {
        at(st, C_POINT) = lex_entry(A_NONE, S_FRAC0);
        at(st, C_E) = lex_entry(A_NONE, S_EXP0);
        at(st, C_LOW_E) = lex_entry(A_NONE, S_EXP0);
    }
    at(S_FRAC0, C_ZERO) = lex_entry(A_NONE, S_FRAC);
    at(S_FRAC0, C_DIGIT) = lex_entry(A_NONE, S_FRAC);
    at(S_FRAC, C_ZERO) = lex_entry(A_NONE, S_FRAC);
    at(S_FRAC, C_DIGIT) = lex_entry(A_NONE, S_FRAC);
    at(S_FRAC, C_E) = lex_entry(A_NONE, S_EXP0);
    at(S_FRAC, C_LOW_E) = lex_entry(A_NONE, S_EXP0);
    at(S_EXP0, C_PLUS) = lex_entry(A_NONE, S_EXPS);
    at(S_EXP0, C_MINUS) = lex_entry(A_NONE, S_EXPS);
    const int exp_digits[] = {S_EXP0, S_EXPS, S_EXP};
    for (int st : exp_digits) { at(st, C_ZERO) = lex_entry(A_NONE, S_EXP); at(st, C_DIGIT) = lex_entry(A_NONE, S_EXP); }
    // a number ends at the first byte that can follow a value; that byte is then handled as in S_AFTER
    const int num_ends[] = {S_ZERO, S_INT, S_FRAC, S_EXP};
    const int followers[] = {C_SPACE, C_WHITE, C_COMMA, C_RCURB, C_RSQRB};
    for (int st : num_ends)
        for (int c : followers) at(st, c) = (uint16_t)(at(S_AFTER, c) | (A_NUM_ENDED << 8));
    return t;
}

}  // namespace detail

// The tables as data.  The scanner indexes ONE fused table of 32-bit entries, [state][byte]:
//     bits 24..31  action (0 on the common path)
//     bits  0..23  where the next state's row starts -- a byte offset from the table's base here;
//                  the device copy adds its shared-memory address, so that there the next
//                  lookup address is just  entry + 4 * byte.
static constexpr detail::LexTables kLexTables = detail::make_lex_tables();
constexpr int kFusedEntries = detail::S_COUNT * 256;
constexpr uint32_t kRowBytes = 256 * 4;

inline void build_fused(uint32_t *out)
{
    for (int k = 0; k < kFusedEntries; k++) {
        const uint16_t e = kLexTables.next[(k >> 8) * detail::C_COUNT + kLexTables.cls[k & 255]];
        out[k] = (uint32_t)(e >> 8) << 24 | (uint32_t)(e & 0xff) * kRowBytes;
    }
}

struct DirectLex {                       // host
    const uint32_t *table;
    ARENA_HD uint32_t row(unsigned state) const { return state * kRowBytes; }
    ARENA_HD uint32_t next(uint32_t row, unsigned c) const { return table[(row >> 2) + c]; }
};

#if defined(__CUDACC__)
struct SharedLex {                       // device: rows are shared-memory addresses
    uint32_t base;
    __device__ uint32_t row(unsigned state) const { return base + state * kRowBytes; }
    __device__ uint32_t next(uint32_t row, unsigned c) const
    {
        uint32_t e;
        asm("ld.shared.u32 %0, [%1];" : "=r"(e) : "r"(row + (c << 2)));
        return e;
    }
};
#endif

// ---- where the walk gets its bytes from ------------------------------------------------------
struct DirectFeed {                      // host: straight from the document
    const unsigned char *t;
    ARENA_HD unsigned get(uint32_t pos) const { return t[pos]; }
    ARENA_HD void start(uint32_t) {}
    ARENA_HD void entered_window(uint32_t) {}
};

#if defined(__CUDACC__)
// Device: every lane owns a 64-byte ring in shared memory (4 windows of 16 bytes) that
// cp.async fills two windows ahead of the walk.  The hot loop then depends on shared-memory
// loads only: a global load in it would stall all 32 lanes of the warp whenever any one of
// them refills, because a scoreboard covers a register for the whole warp.
struct RingFeed {
    const unsigned char *t;              // 16-byte aligned, readable 64 bytes past the item
    uint32_t ring_shared;                // this lane's 64 bytes (64-byte aligned), as a shared-window address
    __device__ RingFeed(const unsigned char *text, unsigned char *lane_ring)
        : t(text), ring_shared((uint32_t)__cvta_generic_to_shared(lane_ring)) {}
    __device__ void request(uint32_t window) const
    {
        asm volatile("cp.async.cg.shared.global [%0], [%1], 16;\n"
                     "cp.async.commit_group;" ::"r"(ring_shared + (window & 3u) * 16u), "l"(t + (size_t)window * 16));
    }
    __device__ void start(uint32_t pos) const
    {
        const uint32_t w = pos >> 4;
        request(w); request(w + 1); request(w + 2);
        asm volatile("cp.async.wait_group 2;" ::: "memory");            // window w has landed
    }
    // the walk stepped onto the first byte of a window: ask for the one two ahead
    __device__ void entered_window(uint32_t pos) const
    {
        request((pos >> 4) + 2);
        asm volatile("cp.async.wait_group 2;" ::: "memory");            // all but the last two requests: this window is in
    }
    __device__ unsigned get(uint32_t pos) const
    {
        unsigned c;
        asm volatile("ld.shared.u8 %0, [%1];" : "=r"(c) : "r"(ring_shared | (pos & 63u)));
        return c;
    }
};
#endif

// Scans the item in t[begin, end) (which must be exactly one JSON value).  See the contract
// at the top.  `lex` looks up the [state][byte] table, `feed` hands out the bytes (DirectLex /
// DirectFeed on the host, SharedLex / RingFeed on the device).
//
// Shape of the walk (what keeps a warp of 32 different pods together):
//   * the FAST loop eats bytes with one table lookup each.  Token boundaries inside a subtree
//     nobody reads ("quiet": managedFields, env, volumes, conditions ... most of a pod) are
//     settled right there with a bit stack of container types -- no role stack, no switch;
//   * a boundary that matters (a key or a value under metadata / spec / status / containers /
//     resources / limits) ends the fast loop.  The lanes of a warp meet again after it, and
//     pods of one schema reach their k-th such boundary together, so the SLOW handler runs
//     once per boundary, not once per lane;
//   * the walk never re-reads the text: keys are recognised by length + hash, and whatever
//     needs the bytes of a token (confirming a recognised key, checking a quantity or a
//     timestamp, decoding the phase, the GPU digits, the node name) goes on a short list
//     worked off after the walk, again side by side.
constexpr int kMaxTodo = 96;
constexpr int kRoleLevels = 8;           // ITEM > SPEC > CTRS > CTR > RES > LIMITS is the deepest path with a role
constexpr int kMaxBitDepth = 62;         // container types live in a 64-bit stack

template <class Lex, class Feed, bool kMeta = false>
ARENA_HD inline void scan_pod_item(const unsigned char *t, size_t begin_, size_t end_, const NodeTable &nodes,
                                   const Lex lex, Feed feed, PodOut &out, PodMeta *meta_out = nullptr)
{
    using namespace detail;
    enum What : uint8_t { W_KEY, W_QUANTITY, W_GPU, W_TIME, W_PHASE, W_KIND };
    PodMeta meta;
    if (kMeta) {
        for (int f = 0; f < kMetaFields; f++) { meta.off[f] = 0; meta.len[f] = 0; }
        meta.present = 0;
        meta.owner_is_job = 0;
    }
    struct Todo {
        uint32_t pos;                    // relative to `begin`
        uint16_t len;
        uint8_t what, arg;               // W_KEY: key id; W_GPU: index into gpu_value[]
    };
    Todo todo[kMaxTodo];
    int n_todo = 0;
    uint8_t stack[kRoleLevels];          // role of the open containers that have one (never the quiet ones)
    uint32_t seen[kRoleLevels];          // key ids already met in the object at that level
    stack[0] = R_SKIP_ARR;
    seen[0] = 0;
    const uint32_t begin = (uint32_t)begin_, end = (uint32_t)end_;       // the caller keeps documents below 4 GiB
    uint32_t flag = kOk;

    uint64_t bits = 0;                   // bit 0 = the innermost open container is an array
    int depth = 0;
    constexpr int kNotQuiet = 1 << 30;
    int quiet_base = kNotQuiet;          // quiet <=> depth > quiet_base

    // GPU quantities wait on the list; a container points at the one that counts for it
    constexpr int16_t kAbsent = -1, kNull = -2;                         // no GPU key / "key": null => 0
    int16_t ctr_ref[kInlineCtrs];
    int64_t gpu_value[2 * kInlineCtrs + 2];                              // evaluated W_GPU entries: amount ...
    int8_t gpu_scale[2 * kInlineCtrs + 2];                               // ... and scale
    int n_gpu = 0;
    int16_t cur_ref = kAbsent, gpu_ref = kAbsent, dep_ref = kAbsent;
    bool gpu_present = false, dep_present = false;

    uint32_t node_off = 0, node_len = 0;
    int phase_todo = -1;                 // W_PHASE entry of the status that counts, or -1 (=> "")
    uint32_t n_ctrs = 0;

    uint8_t key = K_ELEMENT;             // key id of the value being read, in its parent
    uint32_t tok = 0;                    // start of the current string body / number
    bool escaped = false, done = false;
    uint32_t row = lex.row(S_VAL);       // the lexer state, as the start of its table row
    uint32_t hash = 0;                   // FNV-1a of the bytes since the last key's opening quote

    auto defer = [&](uint8_t what, uint8_t arg, uint32_t body, uint32_t len) -> int {
        if (n_todo >= kMaxTodo || len > 0xFFFFu) { flag = kFlagTooManyFields; return -1; }
        todo[n_todo] = Todo{body - begin, (uint16_t)len, what, arg};
        return n_todo++;
    };

    // a string (or null) field the job rules and the printers read: remember where it is
    auto remember = [&](int field, bool is_string, uint32_t body, uint32_t len) {
        if (is_string && (escaped || len > 0xFFFFu)) { flag = escaped ? kFlagEscapedValue : kFlagTooManyFields; return; }
        meta.off[field] = is_string ? body : 0u;
        meta.len[field] = (uint16_t)(is_string ? len : 0u);
        meta.present |= (uint16_t)(1u << field);
    };

    // one finished scalar in context (parent role, key): the host loader's field rules
    auto scalar = [&](bool is_string, bool is_number, bool is_null, uint32_t body, uint32_t len) {
        if (depth == 0) { flag = kFlagShape; return; }                 // "PodList item is not an object"
        const uint8_t parent = stack[depth - 1];
        switch (key) {
        case K_OTHER: break;
        case K_ELEMENT:
            if (parent == R_CTRS || parent == R_ICTRS) flag = kFlagShape;          // "container is not an object"
            else if (parent == R_OWNERS && !is_null) flag = kFlagShape;
            break;
        case K_METADATA: case K_LABELS: case K_OWNERS: case K_INITCONTAINERS: case K_REQUESTS:
            if (!is_null) flag = kFlagShape;
            break;
        case K_SPEC: if (!is_null) flag = kFlagShape; node_len = 0; n_ctrs = 0; break;
        case K_STATUS: if (!is_null) flag = kFlagShape; phase_todo = -1; break;
        case K_CONTAINERS: if (!is_null) flag = kFlagShape; n_ctrs = 0; break;
        case K_RESOURCES: if (!is_null) flag = kFlagShape; cur_ref = kAbsent; break;
        case K_LIMITS: if (!is_null) flag = kFlagShape; cur_ref = kAbsent; break;
        case K_NAME: case K_NAMESPACE: case K_HOSTIP: case K_KIND:
        case K_L_RELEASE: case K_L_APP: case K_L_GROUP: case K_L_RTYPE: case K_L_RINDEX:
            if (!is_string && !is_null) { flag = kFlagShape; break; }
            if (!kMeta) break;
            switch (key) {
            case K_NAME: remember(kMetaName, is_string, body, len); break;
            case K_NAMESPACE: remember(kMetaNamespace, is_string, body, len); break;
            case K_HOSTIP: remember(kMetaHostIP, is_string, body, len); break;
            case K_L_RELEASE: remember(kMetaRelease, is_string, body, len); break;
            case K_L_APP: remember(kMetaApp, is_string, body, len); break;
            case K_L_GROUP: remember(kMetaGroup, is_string, body, len); break;
            case K_L_RTYPE: remember(kMetaReplicaType, is_string, body, len); break;
            case K_L_RINDEX: remember(kMetaReplicaIndex, is_string, body, len); break;
            default:                                                     // K_KIND: compared with "Job" after the walk
                if (is_string) {
                    if (escaped) flag = kFlagEscapedValue;
                    else defer(W_KIND, 0, body, len);
                }
                break;
            }
            break;
        case K_CREATED:
            if (is_null) break;
            if (!is_string) flag = kFlagShape;
            else if (escaped) flag = kFlagTime;
            else {
                defer(W_TIME, 0, body, len);
                if (kMeta) remember(kMetaCreated, true, body, len);
            }
            break;
        case K_NODENAME:
            node_len = 0;
            if (is_null) break;
            if (!is_string) flag = kFlagShape;
            else if (escaped) flag = kFlagEscapedValue;
            else { node_off = body; node_len = len; }
            break;
        case K_PHASE:
            phase_todo = -1;
            if (is_null) break;
            if (!is_string) flag = kFlagShape;
            else if (escaped) flag = kFlagEscapedValue;
            else phase_todo = defer(W_PHASE, 0, body, len);
            break;
        case K_QUANTITY:
            // Quantity.UnmarshalJSON: null, or the raw token with its quotes dropped
            if (is_null) break;
            if (!(is_string || is_number)) flag = kFlagQuantity;
            else defer(W_QUANTITY, 0, body, len);
            break;
        case K_GPU: case K_GPU_DEPRECATED: {
            int16_t ref = kNull;
            if (!is_null) {
                if (!(is_string || is_number) || n_gpu >= 2 * kInlineCtrs + 2) { flag = kFlagQuantity; break; }
                if (defer(W_GPU, (uint8_t)n_gpu, body, len) < 0) break;
                ref = (int16_t)n_gpu++;
            }
            if (key == K_GPU) { gpu_present = true; gpu_ref = ref; }
            else { dep_present = true; dep_ref = ref; }
            break;
        }
        default: break;
        }
    };

    // a boundary where roles matter (never called while quiet)
    auto slow = [&](unsigned act, uint32_t ns, const unsigned c, const uint32_t i) {
        if (act & A_NUM_ENDED) {                                       // this byte follows a number
            scalar(false, true, false, tok, i - tok);
            act &= 0x7fu;
        }
        switch (act) {
        case A_NONE: break;
        case A_OPEN: {
            const bool obj = c == '{';
            uint8_t role = obj ? R_SKIP_OBJ : R_SKIP_ARR;
            const uint8_t parent = depth ? stack[depth - 1] : (uint8_t)R_SKIP_ARR;
            if (depth == 0) {
                if (!obj) flag = kFlagShape;                           // "PodList item is not an object"
                role = R_ITEM;
            } else if (obj) {
                switch (key) {
                case K_METADATA: role = R_META; break;
                case K_SPEC: role = R_SPEC; node_len = 0; n_ctrs = 0; break;
                case K_STATUS: role = R_STATUS; phase_todo = -1; break;
                case K_LABELS: role = R_LABELS; break;
                case K_RESOURCES: role = parent == R_CTR ? R_RES : R_IRES; cur_ref = kAbsent; break;
                case K_LIMITS:
                    if (parent == R_RES) { role = R_LIMITS; cur_ref = kAbsent; gpu_present = dep_present = false; }
                    else role = R_QMAP;
                    break;
                case K_REQUESTS: role = R_QMAP; break;
                case K_ELEMENT:
                    if (parent == R_CTRS) { role = R_CTR; cur_ref = kAbsent; }
                    else if (parent == R_ICTRS) role = R_ICTR;
                    else if (parent == R_OWNERS) role = R_OWNER;
                    break;
                case K_OTHER: break;
                default: flag = kFlagShape; break;                     // an object where the host wants a scalar or a list
                }
            } else {
                switch (key) {
                case K_OWNERS: role = R_OWNERS; break;
                case K_CONTAINERS: role = R_CTRS; n_ctrs = 0; break;
                case K_INITCONTAINERS: role = R_ICTRS; break;
                case K_ELEMENT:
                    if (parent == R_CTRS || parent == R_ICTRS || parent == R_OWNERS) flag = kFlagShape;
                    break;
                case K_OTHER: break;
                default: flag = kFlagShape; break;                     // a list where the host wants an object or a scalar
                }
            }
            if (role == R_SKIP_OBJ || role == R_SKIP_ARR) quiet_base = depth;      // nobody reads this subtree
            else if (depth >= kRoleLevels) { flag = kFlagTooDeep; break; }
            else { stack[depth] = role; seen[depth] = 0; }
            bits = (bits << 1) | (obj ? 0u : 1u);
            depth++;
            key = K_ELEMENT;                                           // (arrays; objects set it at every key)
            break;
        }
        case A_CLOSE: {
            if (depth == 0) { flag = kFlagSyntax; break; }
            const uint8_t role = stack[depth - 1];
            if ((bool)(bits & 1) != (c == ']')) { flag = kFlagSyntax; break; }
            if (role == R_LIMITS) cur_ref = gpu_present ? gpu_ref : dep_present ? dep_ref : kAbsent;   // gpu.go:61-79: key presence
            else if (role == R_CTR) {
                if (n_ctrs < (uint32_t)kInlineCtrs) ctr_ref[n_ctrs] = cur_ref;
                else flag = kFlagTooManyContainers;
                n_ctrs++;
            }
            bits >>= 1;
            depth--;
            if (depth == 0) done = true;
            break;
        }
        case A_COMMA:
            if (depth == 0) { flag = kFlagSyntax; break; }
            if (bits & 1) { ns = lex.row(S_VAL); key = K_ELEMENT; }
            else ns = lex.row(S_KEY);
            break;
        case A_KEY_BEGIN: tok = i + 1; escaped = false; hash = 2166136261u; break;
        case A_STR_BEGIN: tok = i + 1; escaped = false; break;
        case A_ESCAPE: escaped = true; break;
        case A_KEY_END: {
            const uint8_t role = stack[depth - 1];
            if (escaped) flag = kFlagEscapedKey;
            key = key_id_hashed(role, i - tok, hash);
            if (key != K_OTHER && key != K_QUANTITY) {
                const uint32_t bit = 1u << key;
                if (seen[depth - 1] & bit) flag = kFlagDuplicateKey;
                seen[depth - 1] |= bit;
                defer(W_KEY, key, tok, i - tok);                       // confirmed byte for byte after the walk
            }
            break;
        }
        case A_STR_END: scalar(true, false, false, tok, i - tok); break;
        case A_NUM_BEGIN: tok = i; break;
        case A_LIT_END: scalar(false, false, c == 'l', i, 1); break;   // ...l = null
        default: flag = kFlagSyntax; break;                            // A_ERR
        }
        row = ns;
    };

    // ---- the walk ------------------------------------------------------------------------------
    // (ONE loop, one byte per trip, short ifs inside: each `if` ends where the lanes of a warp
    // meet again.  Nested loops here make the compiler put that meeting point after the whole
    // nest, and 32 lanes then drift apart and run one at a time.)
    uint32_t pos = begin;
    feed.start(pos);
    while (pos < end && !flag && !done) {
        unsigned e = 0, c = 0;
        bool pending = false;
        // FAST: until a boundary that needs the slow handler
        while (pos < end) {
            c = feed.get(pos);
            e = lex.next(row, c);
            if (e >> 24) {                                             // some action
                if (depth <= quiet_base) { pending = true; break; }    // roles matter here
                // quiet: container types are all that is needed
                const unsigned act = (e >> 24) & 0x7fu;
                e &= 0xFFFFFFu;
                if (act == A_OPEN) {
                    if (depth >= kMaxBitDepth) { flag = kFlagTooDeep; break; }
                    bits = (bits << 1) | (c == '[' ? 1u : 0u);
                    depth++;
                } else if (act == A_CLOSE) {
                    if ((bool)(bits & 1) != (c == ']')) { flag = kFlagSyntax; break; }
                    bits >>= 1;
                    depth--;
                    if (depth == quiet_base) quiet_base = kNotQuiet;   // the subtree is over; its value needed no handling
                } else if (act == A_COMMA) {
                    e = lex.row((bits & 1) ? S_VAL : S_KEY);
                } else if (act == A_ERR) {
                    flag = kFlagSyntax;
                    break;
                }
            } else {
                hash = (hash ^ c) * 16777619u;
            }
            row = e;
            pos++;
            if ((pos & 15u) == 0) feed.entered_window(pos);
        }
        if (!pending) continue;                                        // out of bytes, or flagged while quiet
        // SLOW: the lanes of a warp arrive here together
        slow(e >> 24, e & 0xFFFFFFu, c, pos);
        pos++;
        if ((pos & 15u) == 0) feed.entered_window(pos);
    }
    if (!flag && !done) flag = kFlagSyntax;                            // ran out of bytes inside the item
    if (!flag)
        for (uint32_t i = pos; i < end; i++)
            if (!is_ws(t[i])) { flag = kFlagSyntax; break; }

    // ---- the deferred checks: the only place that goes back to the text ---------------------
    uint8_t phase = 0;
    for (int k = 0; k < n_todo && !flag; k++) {
        const Todo d = todo[k];
        const unsigned char *p = t + begin + d.pos;
        switch (d.what) {
        case W_KEY:
            if (!bytes_are(p, d.len, key_text(d.arg))) flag = kFlagKeyHash;
            break;
        case W_QUANTITY:
            if (!quantity_plain(p, d.len)) flag = kFlagQuantity;
            break;
        case W_GPU: {
            // the fast-path grammar ("2", "500m", "1.5", "1e3", "1Ki"); the rest is the host's business
            int64_t v = 0;
            int8_t sc = 0;
            if (!quantity_fast(p, d.len, v, sc)) flag = kFlagQuantity;
            gpu_value[d.arg] = v;
            gpu_scale[d.arg] = sc;
            break;
        }
        case W_TIME:
            if (!time_canonical(p, d.len)) flag = kFlagTime;
            break;
        case W_KIND:
            if (kMeta && bytes_are(p, d.len, "Job")) meta.owner_is_job = 1;
            break;
        default:                                                       // W_PHASE
            if (k == phase_todo) phase = phase_code(p, d.len);
            break;
        }
    }
    if (!flag)
        for (uint32_t k = 0; k < n_ctrs && k < (uint32_t)kInlineCtrs; k++)
        {
            out.ctr_val[k] = ctr_ref[k] >= 0 ? gpu_value[ctr_ref[k]] : 0;
            out.ctr_scale[k] = ctr_ref[k] >= 0 ? gpu_scale[ctr_ref[k]] : (int8_t)0;
        }
    out.flag = flag;
    out.phase = phase;
    out.n_ctrs = n_ctrs;
    out.node = (!flag && node_len) ? find_node(nodes, t + node_off, node_len) : kNoNode;     // loader.cc: "" joins nothing
    if (kMeta && meta_out) *meta_out = meta;
}

// What must sit between two items (or between '[' and the first one): whitespace and,
// unless `first`, exactly one comma.
ARENA_HD inline bool gap_ok(const unsigned char *t, size_t from, size_t to, bool first)
{
    int commas = 0;
    for (size_t k = from; k < to; k++) {
        if (t[k] == ',') commas++;
        else if (!detail::is_ws(t[k])) return false;
    }
    return commas == (first ? 0 : 1);
}

}  // namespace arena_scan
