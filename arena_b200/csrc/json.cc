#include "json.h"

#include <cstdint>

namespace arena_json {

namespace {

struct Parser {
    std::string_view t;
    size_t i = 0;
    ParseError *err;
    int depth = 0;

    bool fail(const char *msg)
    {
        err->offset = i;
        err->message = msg;
        return false;
    }
    void ws()
    {
        while (i < t.size() && (t[i] == ' ' || t[i] == '\t' || t[i] == '\n' || t[i] == '\r')) i++;
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
    bool value(Value &v)
    {
        ws();
        if (i >= t.size()) return fail("unexpected end of input");
        if (++depth > 512) return fail("nesting too deep");
        bool ok = value_inner(v);
        depth--;
        return ok;
    }
    bool value_inner(Value &v)
    {
        char c = t[i];
        size_t start = i;
        if (c == '{') {
            v.type = Type::Object;
            i++;
            ws();
            if (i < t.size() && t[i] == '}') { i++; v.raw = t.substr(start, i - start); return true; }
            for (;;) {
                ws();
                if (i >= t.size() || t[i] != '"') return fail("expected object key");
                std::string_view kraw;
                if (!string_token(kraw)) return false;
                Value ktmp;
                ktmp.type = Type::String;
                ktmp.raw = kraw;
                ws();
                if (i >= t.size() || t[i] != ':') return fail("expected ':'");
                i++;
                v.members.emplace_back(ktmp.str(), Value());
                if (!value(v.members.back().second)) return false;
                ws();
                if (i < t.size() && t[i] == ',') { i++; continue; }
                if (i < t.size() && t[i] == '}') { i++; break; }
                return fail("expected ',' or '}'");
            }
            v.raw = t.substr(start, i - start);
            return true;
        }
        if (c == '[') {
            v.type = Type::Array;
            i++;
            ws();
            if (i < t.size() && t[i] == ']') { i++; v.raw = t.substr(start, i - start); return true; }
            for (;;) {
                v.items.emplace_back();
                if (!value(v.items.back())) return false;
                ws();
                if (i < t.size() && t[i] == ',') { i++; continue; }
                if (i < t.size() && t[i] == ']') { i++; break; }
                return fail("expected ',' or ']'");
            }
            v.raw = t.substr(start, i - start);
            return true;
        }
        if (c == '"') { v.type = Type::String; return string_token(v.raw); }
        if (c == 't') { v.type = Type::Bool; if (!lit("true", 4)) return false; v.raw = t.substr(start, 4); return true; }
        if (c == 'f') { v.type = Type::Bool; if (!lit("false", 5)) return false; v.raw = t.substr(start, 5); return true; }
        if (c == 'n') { v.type = Type::Null; if (!lit("null", 4)) return false; v.raw = t.substr(start, 4); return true; }
        if (c == '-' || (c >= '0' && c <= '9')) { v.type = Type::Number; return number_token(v.raw); }
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

const Value *Value::get(std::string_view key) const
{
    const Value *hit = nullptr;
    for (const auto &m : members)
        if (m.first == key) hit = &m.second;       // duplicates: last wins
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

bool parse(std::string_view text, Value &out, ParseError &err)
{
    Parser p{text, 0, &err};
    if (!p.value(out)) return false;
    p.ws();
    if (p.i != text.size()) return p.fail("trailing characters after document");
    return true;
}

}  // namespace arena_json
