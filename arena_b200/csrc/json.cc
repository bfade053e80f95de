#include "json.h"

#include <cstdint>
#include <cstring>
#if defined(__SSE2__)
#include <emmintrin.h>
#endif

namespace arena_json {

namespace {

struct Parser {
    std::string_view t;
    size_t i = 0;
    ParseError *err;
    const ParseOptions *opt;
    KeyFilter filter = nullptr;
    Tree *tree = nullptr;
    int depth = 0;
    bool split_seen = false, split_conflict = false;

    bool fail(const char *msg)
    {
        err->offset = i;
        err->message = msg;
        return false;
    }
    void ws()
    {
        while (i < t.size() && (unsigned char)t[i] <= ' ' && (t[i] == ' ' || t[i] == '\t' || t[i] == '\n' || t[i] == '\r')) i++;
    }
    bool lit(const char *w, size_t n)
    {
        if (t.size() - i < n || t.compare(i, n, w) != 0) return fail("invalid literal");
        i += n;
        return true;
    }
    bool string_token(std::string_view &raw)
    {
        size_t start = i;
        i++;  // opening quote
        while (i < t.size()) {
#if defined(__SSE2__)
            // skip 16 bytes at a time while none of them is '"', '\\' or a control character
            while (t.size() - i >= 16) {
                const __m128i v = _mm_loadu_si128(reinterpret_cast<const __m128i *>(t.data() + i));
                const __m128i hit = _mm_or_si128(_mm_or_si128(_mm_cmpeq_epi8(v, _mm_set1_epi8('"')),
                                                              _mm_cmpeq_epi8(v, _mm_set1_epi8('\\'))),
                                                 _mm_cmpeq_epi8(_mm_min_epu8(v, _mm_set1_epi8(0x1f)), v));
                const int m = _mm_movemask_epi8(hit);
                if (m) { i += (size_t)__builtin_ctz((unsigned)m); break; }
                i += 16;
            }
            if (i >= t.size()) break;
#endif
            unsigned char c = (unsigned char)t[i];
            if (c == '"') {
                i++;
                raw = t.substr(start, i - start);
                return true;
            }
            if (c == '\\') {
                i++;
                if (i >= t.size()) break;
                char e = t[i];
                if (e == 'u') {
                    if (t.size() - i < 5) return fail("short \\u escape");
                    for (int k = 1; k <= 4; k++) {
                        char h = t[i + k];
                        bool ok = (h >= '0' && h <= '9') || (h >= 'a' && h <= 'f') || (h >= 'A' && h <= 'F');
                        if (!ok) return fail("bad \\u escape");
                    }
                    i += 5;
                    continue;
                }
                if (e != '"' && e != '\\' && e != '/' && e != 'b' && e != 'f' && e != 'n' && e != 'r' && e != 't')
                    return fail("bad escape");
                i++;
                continue;
            }
            if (c < 0x20) return fail("control character in string");
            i++;
        }
        return fail("unterminated string");
    }
    bool number_token(std::string_view &raw)
    {
        size_t start = i;
        if (i < t.size() && t[i] == '-') i++;
        if (i >= t.size()) return fail("bad number");
        if (t[i] == '0') i++;
        else if (t[i] >= '1' && t[i] <= '9') while (i < t.size() && t[i] >= '0' && t[i] <= '9') i++;
        else return fail("bad number");
        if (i < t.size() && t[i] == '.') {
            i++;
            if (i >= t.size() || t[i] < '0' || t[i] > '9') return fail("bad fraction");
            while (i < t.size() && t[i] >= '0' && t[i] <= '9') i++;
        }
        if (i < t.size() && (t[i] == 'e' || t[i] == 'E')) {
            i++;
            if (i < t.size() && (t[i] == '+' || t[i] == '-')) i++;
            if (i >= t.size() || t[i] < '0' || t[i] > '9') return fail("bad exponent");
            while (i < t.size() && t[i] >= '0' && t[i] <= '9') i++;
        }
        raw = t.substr(start, i - start);
        return true;
    }
    static constexpr uint32_t kNoNode = 0xFFFFFFFFu;

    uint32_t new_child(uint32_t parent, uint32_t &tail, std::string_view key)
    {
        std::vector<Value> &nodes = tree->nodes;
        const uint32_t c = (uint32_t)nodes.size();
        nodes.emplace_back();
        nodes[c].key = key;
        if (tail == kNoNode) nodes[parent].first = (int32_t)(c - parent);
        else nodes[tail].next = (int32_t)(c - tail);
        nodes[parent].n++;
        tail = c;
        return c;
    }
    // at == kNoNode: validate only, build nothing.  `parent` = key of the enclosing member.
    bool value(uint32_t at, std::string_view parent)
    {
        ws();
        if (i >= t.size()) return fail("unexpected end of input");
        if (++depth > 512) return fail("nesting too deep");
        bool ok = value_inner(at, parent);
        depth--;
        return ok;
    }
    // root-level list member: hand the elements out as byte ranges instead of descending
    bool try_split(uint32_t obj, uint32_t &tail, std::string_view key)
    {
        size_t j = i;
        while (j < t.size() && (t[j] == ' ' || t[j] == '\t' || t[j] == '\n' || t[j] == '\r')) j++;
        if (j >= t.size() || t[j] != '[') return false;
        if (split_seen) { split_conflict = true; return false; }      // duplicate key: parse() starts over without splitting
        size_t end = 0;
        std::vector<Range> ranges;
        if (opt->split_hook) {
            if (!opt->split_hook(opt->split_user, j, &end) || end <= j + 1 || end > t.size() || t[end - 1] != ']') return false;
            ranges.push_back({j, end});
        } else if (!split_array(t, j, ranges, end)) return false;
        split_seen = true;
        *opt->split_out = std::move(ranges);
        if (opt->split_done) *opt->split_done = true;
        const uint32_t a = new_child(obj, tail, key);
        tree->nodes[a].type = Type::Array;
        tree->nodes[a].raw = t.substr(j, end - j);
        i = end;
        return true;
    }
    std::string_view key_view(std::string_view kraw)
    {
        std::string_view body = kraw.substr(1, kraw.size() - 2);
        if (body.find('\\') == std::string_view::npos) return body;
        Value ktmp;
        ktmp.type = Type::String;
        ktmp.raw = kraw;
        tree->keys.push_back(ktmp.str());
        return tree->keys.back();
    }
    void set(uint32_t at, Type type, std::string_view raw)
    {
        if (at == kNoNode) return;
        tree->nodes[at].type = type;
        tree->nodes[at].raw = raw;
    }
    bool value_inner(uint32_t at, std::string_view parent)
    {
        const bool keep = at != kNoNode;
        char c = t[i];
        size_t start = i;
        if (c == '{') {
            i++;
            ws();
            if (i < t.size() && t[i] == '}') { i++; set(at, Type::Object, t.substr(start, i - start)); return true; }
            uint32_t tail = kNoNode;
            for (;;) {
                ws();
                if (i >= t.size() || t[i] != '"') return fail("expected object key");
                std::string_view kraw;
                if (!string_token(kraw)) return false;
                ws();
                if (i >= t.size() || t[i] != ':') return fail("expected ':'");
                i++;
                bool want = keep;
                std::string_view key;
                if (keep) {
                    key = key_view(kraw);
                    if (filter && !filter(parent, key)) want = false;
                }
                if (want && depth == 1 && opt->split_out && key == opt->split_key && try_split(at, tail, key)) {
                    // elements recorded for parse_range
                } else if (want) {
                    if (!value(new_child(at, tail, key), key)) return false;
                } else {
                    if (!value(kNoNode, std::string_view())) return false;
                }
                ws();
                if (i < t.size() && t[i] == ',') { i++; continue; }
                if (i < t.size() && t[i] == '}') { i++; break; }
                return fail("expected ',' or '}'");
            }
            set(at, Type::Object, t.substr(start, i - start));
            return true;
        }
        if (c == '[') {
            i++;
            ws();
            if (i < t.size() && t[i] == ']') { i++; set(at, Type::Array, t.substr(start, i - start)); return true; }
            uint32_t tail = kNoNode;
            for (;;) {
                // elements inherit the array's key as their parent key
                if (!value(keep ? new_child(at, tail, std::string_view()) : kNoNode, parent)) return false;
                ws();
                if (i < t.size() && t[i] == ',') { i++; continue; }
                if (i < t.size() && t[i] == ']') { i++; break; }
                return fail("expected ',' or ']'");
            }
            set(at, Type::Array, t.substr(start, i - start));
            return true;
        }
        std::string_view raw;
        if (c == '"') { if (!string_token(raw)) return false; set(at, Type::String, raw); return true; }
        if (c == 't') { if (!lit("true", 4)) return false; set(at, Type::Bool, t.substr(start, 4)); return true; }
        if (c == 'f') { if (!lit("false", 5)) return false; set(at, Type::Bool, t.substr(start, 5)); return true; }
        if (c == 'n') { if (!lit("null", 4)) return false; set(at, Type::Null, t.substr(start, 4)); return true; }
        if (c == '-' || (c >= '0' && c <= '9')) { if (!number_token(raw)) return false; set(at, Type::Number, raw); return true; }
        return fail("unexpected character");
    }
};

void append_utf8(std::string &out, uint32_t cp)
{
    if (cp < 0x80) out.push_back((char)cp);
    else if (cp < 0x800) { out.push_back((char)(0xC0 | (cp >> 6))); out.push_back((char)(0x80 | (cp & 0x3F))); }
    else if (cp < 0x10000) {
        out.push_back((char)(0xE0 | (cp >> 12)));
        out.push_back((char)(0x80 | ((cp >> 6) & 0x3F)));
        out.push_back((char)(0x80 | (cp & 0x3F)));
    } else {
        out.push_back((char)(0xF0 | (cp >> 18)));
        out.push_back((char)(0x80 | ((cp >> 12) & 0x3F)));
        out.push_back((char)(0x80 | ((cp >> 6) & 0x3F)));
        out.push_back((char)(0x80 | (cp & 0x3F)));
    }
}

uint32_t hex4(std::string_view s, size_t at)
{
    uint32_t v = 0;
    for (int k = 0; k < 4; k++) {
        char h = s[at + k];
        v = v * 16 + (uint32_t)(h <= '9' ? h - '0' : (h | 0x20) - 'a' + 10);
    }
    return v;
}

}  // namespace

const Value *Value::get(std::string_view k) const
{
    const Value *hit = nullptr;
    for (const Value *m = first_child(); m; m = m->next_sibling())
        if (m->key == k) hit = m;                  // duplicates: last wins
    return hit;
}

std::string Value::str() const
{
    std::string out;
    if (type != Type::String || raw.size() < 2) return out;
    std::string_view s = raw.substr(1, raw.size() - 2);
    out.reserve(s.size());
    for (size_t i = 0; i < s.size(); i++) {
        char c = s[i];
        if (c != '\\') { out.push_back(c); continue; }
        char e = s[++i];
        switch (e) {
        case 'b': out.push_back('\b'); break;
        case 'f': out.push_back('\f'); break;
        case 'n': out.push_back('\n'); break;
        case 'r': out.push_back('\r'); break;
        case 't': out.push_back('\t'); break;
        case 'u': {
            uint32_t cp = hex4(s, i + 1);
            i += 4;
            if (cp >= 0xD800 && cp <= 0xDBFF && i + 6 < s.size() + 0 && s[i + 1] == '\\' && s[i + 2] == 'u') {
                uint32_t lo = hex4(s, i + 3);
                if (lo >= 0xDC00 && lo <= 0xDFFF) { cp = 0x10000 + ((cp - 0xD800) << 10) + (lo - 0xDC00); i += 6; }
            }
            append_utf8(out, cp);
            break;
        }
        default: out.push_back(e); break;     // '"', '\\', '/'
        }
    }
    return out;
}

bool parse(std::string_view text, Tree &out, ParseError &err, const ParseOptions &opt)
{
    if (opt.split_done) *opt.split_done = false;
    out.clear();
    out.nodes.emplace_back();
    Parser p{text, 0, &err, &opt, opt.filter, &out};
    bool ok = p.value(0, std::string_view());
    if (p.split_conflict) {
        ParseOptions plain = opt;
        plain.split_out = nullptr;
        plain.split_done = nullptr;
        if (opt.split_done) *opt.split_done = false;
        opt.split_out->clear();
        return parse(text, out, err, plain);
    }
    if (!ok) return false;
    p.ws();
    if (p.i != text.size()) return p.fail("trailing characters after document");
    return true;
}

bool parse_range(std::string_view text, Range r, int depth, std::string_view parent_key, Tree &out, ParseError &err,
                 const ParseOptions &opt)
{
    ParseOptions o = opt;
    o.split_out = nullptr;
    out.clear();
    out.nodes.emplace_back();
    Parser p{text.substr(0, r.end), r.begin, &err, &o, o.filter, &out};
    p.depth = depth;
    if (!p.value(0, parent_key)) return false;
    p.ws();
    if (p.i != r.end) return p.fail("trailing characters after document");
    return true;
}

namespace {

// One 64-byte step of the structural scan: bit k of each mask describes byte k.
struct Block {
    uint64_t quote, backslash, open, close;
};

inline Block classify(const char *p)
{
    Block b{0, 0, 0, 0};
#if defined(__SSE2__)
    for (int k = 0; k < 4; k++) {
        const __m128i v = _mm_loadu_si128(reinterpret_cast<const __m128i *>(p + 16 * k));
        const __m128i lower = _mm_or_si128(v, _mm_set1_epi8(0x20));            // '[' -> '{', ']' -> '}'
        b.quote |= (uint64_t)(unsigned)_mm_movemask_epi8(_mm_cmpeq_epi8(v, _mm_set1_epi8('"'))) << (16 * k);
        b.backslash |= (uint64_t)(unsigned)_mm_movemask_epi8(_mm_cmpeq_epi8(v, _mm_set1_epi8('\\'))) << (16 * k);
        b.open |= (uint64_t)(unsigned)_mm_movemask_epi8(_mm_cmpeq_epi8(lower, _mm_set1_epi8('{'))) << (16 * k);
        b.close |= (uint64_t)(unsigned)_mm_movemask_epi8(_mm_cmpeq_epi8(lower, _mm_set1_epi8('}'))) << (16 * k);
    }
#else
    for (int k = 0; k < 64; k++) {
        const char c = p[k];
        const uint64_t bit = 1ull << k;
        if (c == '"') b.quote |= bit;
        else if (c == '\\') b.backslash |= bit;
        else if (c == '{' || c == '[') b.open |= bit;
        else if (c == '}' || c == ']') b.close |= bit;
    }
#endif
    return b;
}

// Index of the bracket that closes the one at text[at], or npos.  Strings are skipped
// (a quote preceded by an odd run of backslashes does not end one); '{' and '[' are
// not told apart here: the full parser checks that when it reads the range.
size_t find_close(std::string_view t, size_t at)
{
    uint64_t in_string = 0;          // all ones while inside a string at the block boundary
    bool next_escaped = false;       // the previous block ended in an unpaired backslash
    long depth = 0;
    size_t pos = at;
    char tail[64];
    while (pos < t.size()) {
        const char *p = t.data() + pos;
        const size_t left = t.size() - pos;
        if (left < 64) {
            std::memset(tail, ' ', sizeof tail);
            std::memcpy(tail, p, left);
            p = tail;
        }
        const Block b = classify(p);
        uint64_t escaped = 0;
        if (b.backslash || next_escaped) {
            bool e = next_escaped;
            for (int k = 0; k < 64; k++) {
                if (e) { escaped |= 1ull << k; e = false; }
                else if ((b.backslash >> k) & 1) e = true;
            }
            next_escaped = e;
        }
        uint64_t q = b.quote & ~escaped;
        q ^= q << 1; q ^= q << 2; q ^= q << 4; q ^= q << 8; q ^= q << 16; q ^= q << 32;     // prefix xor: inside-string mask
        q ^= in_string;
        in_string = (uint64_t)((int64_t)q >> 63);
        uint64_t structural = (b.open | b.close) & ~q;
        while (structural) {
            const int k = __builtin_ctzll(structural);
            structural &= structural - 1;
            if ((b.open >> k) & 1) depth++;
            else if (--depth == 0) return pos + (size_t)k;
        }
        pos += 64;
    }
    return std::string_view::npos;
}

inline bool is_ws(char c) { return c == ' ' || c == '\t' || c == '\n' || c == '\r'; }

}  // namespace

bool split_array(std::string_view t, size_t open, std::vector<Range> &out, size_t &end)
{
    out.clear();
    if (open >= t.size() || t[open] != '[') return false;
    size_t pos = open + 1;
    while (pos < t.size() && is_ws(t[pos])) pos++;
    if (pos < t.size() && t[pos] == ']') { end = pos + 1; return true; }
    for (;;) {
        while (pos < t.size() && is_ws(t[pos])) pos++;
        if (pos >= t.size() || (t[pos] != '{' && t[pos] != '[')) return false;        // scalar element or damage
        const size_t close = find_close(t, pos);
        if (close == std::string_view::npos) return false;
        out.push_back({pos, close + 1});
        pos = close + 1;
        while (pos < t.size() && is_ws(t[pos])) pos++;
        if (pos >= t.size()) return false;
        if (t[pos] == ',') { pos++; continue; }
        if (t[pos] == ']') { end = pos + 1; return true; }
        return false;
    }
}

}  // namespace arena_json
