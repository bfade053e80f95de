#include "quantity.h"

#include <cstring>
#include <string>
#include <vector>

namespace arena_host {

namespace {

constexpr int64_t kI64Max = INT64_MAX;
constexpr int64_t kI64Min = INT64_MIN;

inline bool is_digit(char c) { return c >= '0' && c <= '9'; }

// ---- arbitrary-size non-negative decimal integer (slow path only) -------------
// Stored as decimal digits, most significant first, no leading zeros ("0" = empty).
struct BigDec {
    std::string d;

    static BigDec from_digits(std::string_view s)
    {
        BigDec b;
        size_t i = 0;
        while (i < s.size() && s[i] == '0') i++;
        b.d.assign(s.substr(i));
        return b;
    }
    bool is_zero() const { return d.empty(); }
    void mul_pow2(unsigned e)                          // e <= 60
    {
        if (is_zero() || e == 0) return;
        const unsigned __int128 m = (unsigned __int128)1 << e;
        std::string out;
        unsigned __int128 carry = 0;
        for (size_t k = d.size(); k-- > 0;) {
            unsigned __int128 cur = (unsigned __int128)(d[k] - '0') * m + carry;
            out.push_back((char)('0' + (int)(cur % 10)));
            carry = cur / 10;
        }
        while (carry) { out.push_back((char)('0' + (int)(carry % 10))); carry /= 10; }
        d.assign(out.rbegin(), out.rend());
    }
    void mul_pow10(size_t n) { if (!is_zero()) d.append(n, '0'); }
    // floor division by 10^n; returns whether the remainder was nonzero
    bool div_pow10(size_t n)
    {
        if (n == 0) return false;
        bool rem = false;
        if (n >= d.size()) {
            rem = !is_zero();
            d.clear();
            return rem;
        }
        for (size_t k = d.size() - n; k < d.size(); k++) rem |= d[k] != '0';
        d.resize(d.size() - n);
        return rem;
    }
    void add_one()
    {
        size_t k = d.size();
        while (k > 0) {
            if (d[k - 1] != '9') { d[k - 1]++; return; }
            d[k - 1] = '0';
            k--;
        }
        d.insert(d.begin(), '1');
    }
    int cmp(const BigDec &o) const
    {
        if (d.size() != o.d.size()) return d.size() < o.d.size() ? -1 : 1;
        int c = d.compare(o.d);
        return c < 0 ? -1 : (c > 0 ? 1 : 0);
    }
    uint64_t low64() const                              // value mod 2^64
    {
        uint64_t v = 0;
        for (char c : d) v = v * 10u + (uint64_t)(c - '0');
        return v;
    }
};

const BigDec &max_int64()
{
    static const BigDec m = BigDec::from_digits("9223372036854775807");
    return m;
}

// big.Int.Int64() of (neg ? -mag : mag): low 64 bits of the magnitude, negated.
int64_t big_int64(const BigDec &mag, bool neg)
{
    uint64_t v = mag.low64();
    if (neg) v = 0u - v;
    return (int64_t)v;
}

const int64_t kPow10[19] = {1ll, 10ll, 100ll, 1000ll, 10000ll, 100000ll, 1000000ll, 10000000ll, 100000000ll,
                            1000000000ll, 10000000000ll, 100000000000ll, 1000000000000ll, 10000000000000ll,
                            100000000000000ll, 1000000000000000ll, 10000000000000000ll, 100000000000000000ll,
                            1000000000000000000ll};

// scale_int.go:43-95 with newScale == 0.  `mag`/`neg` = unscaled big.Int.
int64_t scaled_value(const BigDec &mag, bool neg, int64_t dscale)
{
    const int64_t dif = dscale;
    if (dif == 0) return big_int64(mag, neg);
    if (dif < 0) {
        // unscaled.Int64() * int64(math.Pow10(-dif)): only reachable with a zero amount
        if (mag.is_zero()) return 0;
        if (-dif > 18) return 0;   // float->int64 conversion out of range in the reference; unreachable
        return (int64_t)((uint64_t)big_int64(mag, neg) * (uint64_t)kPow10[-dif]);
    }
    const bool below_max = neg ? true : (mag.cmp(max_int64()) < 0);      // unscaled.Cmp(maxInt64) < 0
    if (below_max && dif < 19) {
        const int64_t u = big_int64(mag, neg);
        const int64_t divide = kPow10[dif];
        const int64_t result = u / divide;
        const int64_t mod = u % divide;
        if (mod != 0) return (int64_t)((uint64_t)result + 1u);
        return result;
    }
    // Euclidean DivMod on the big value; only non-negative values get here
    BigDec q = mag;
    const bool rem = q.div_pow10((size_t)dif);
    int64_t r = big_int64(q, false);
    if (neg) {
        // unreachable (negative values always satisfy below_max; dif >= 19 only for zero)
        r = (int64_t)(0u - (uint64_t)r);
        if (rem) r = (int64_t)((uint64_t)r - 1u);       // floor toward -inf, then +1 below
    }
    if (rem) r = (int64_t)((uint64_t)r + 1u);
    return r;
}

struct Scan {
    bool positive = true;
    std::string_view value, num, denom, suffix;
};

// quantity.go:159-273
bool parse_quantity_string(std::string_view s, Scan &o)
{
    size_t pos = 0;
    const size_t end = s.size();
    if (pos < end) {
        if (s[0] == '-') { o.positive = false; pos++; }
        else if (s[0] == '+') pos++;
    }
    for (size_t i = pos;; i++) {                         // strip leading zeros
        if (i >= end) { o.num = "0"; o.value = o.num; return true; }
        if (s[i] == '0') pos++;
        else break;
    }
    bool done = false;
    for (size_t i = pos;; i++) {                         // numerator
        if (i >= end) { o.num = s.substr(pos, end - pos); o.value = s.substr(0, end); done = true; break; }
        if (!is_digit(s[i])) { o.num = s.substr(pos, i - pos); pos = i; break; }
    }
    if (done) return true;
    if (o.num.empty()) o.num = "0";
    if (pos < end && s[pos] == '.') {                    // denominator
        pos++;
        for (size_t i = pos;; i++) {
            if (i >= end) { o.denom = s.substr(pos, end - pos); o.value = s.substr(0, end); return true; }
            if (!is_digit(s[i])) { o.denom = s.substr(pos, i - pos); pos = i; break; }
        }
    }
    o.value = s.substr(0, pos);
    const size_t suffix_start = pos;
    for (size_t i = pos;; i++) {
        if (i >= end) { o.suffix = s.substr(suffix_start, end - suffix_start); return true; }
        if (std::strchr("eEinumkKMGTP", s[i]) == nullptr) { pos = i; break; }
    }
    if (pos < end && (s[pos] == '-' || s[pos] == '+')) pos++;
    for (size_t i = pos;; i++) {
        if (i >= end) { o.suffix = s.substr(suffix_start, end - suffix_start); return true; }
        if (!is_digit(s[i])) break;
    }
    return false;                                        // ErrFormatWrong
}

enum class Fmt { DecimalSI, BinarySI, DecimalExponent };

// strconv.ParseInt(s, 10, 64)
bool go_parse_int64(std::string_view s, int64_t &out)
{
    if (s.empty()) return false;
    bool neg = false;
    size_t i = 0;
    if (s[0] == '+' || s[0] == '-') { neg = s[0] == '-'; i = 1; }
    if (i >= s.size()) return false;
    unsigned __int128 v = 0;
    for (; i < s.size(); i++) {
        if (!is_digit(s[i])) return false;
        v = v * 10 + (unsigned)(s[i] - '0');
        if (v > ((unsigned __int128)1 << 64)) return false;
    }
    if (neg) {
        if (v > ((unsigned __int128)1 << 63)) return false;
        out = (int64_t)(0u - (uint64_t)v);
    } else {
        if (v > (unsigned __int128)kI64Max) return false;
        out = (int64_t)v;
    }
    return true;
}

// suffix.go:88-106, 180-198
bool interpret_suffix(std::string_view suf, int &base, int32_t &exponent, Fmt &fmt)
{
    struct E { const char *s; int base; int32_t e; Fmt f; };
    static const E table[] = {
        {"", 10, 0, Fmt::DecimalSI},  {"n", 10, -9, Fmt::DecimalSI}, {"u", 10, -6, Fmt::DecimalSI},
        {"m", 10, -3, Fmt::DecimalSI}, {"k", 10, 3, Fmt::DecimalSI},  {"M", 10, 6, Fmt::DecimalSI},
        {"G", 10, 9, Fmt::DecimalSI},  {"T", 10, 12, Fmt::DecimalSI}, {"P", 10, 15, Fmt::DecimalSI},
        {"E", 10, 18, Fmt::DecimalSI}, {"Ki", 2, 10, Fmt::BinarySI},  {"Mi", 2, 20, Fmt::BinarySI},
        {"Gi", 2, 30, Fmt::BinarySI},  {"Ti", 2, 40, Fmt::BinarySI},  {"Pi", 2, 50, Fmt::BinarySI},
        {"Ei", 2, 60, Fmt::BinarySI}};
    for (const E &e : table)
        if (suf == e.s) { base = e.base; exponent = e.e; fmt = e.f; return true; }
    if (suf.size() > 1 && (suf[0] == 'E' || suf[0] == 'e')) {
        int64_t parsed;
        if (!go_parse_int64(suf.substr(1), parsed)) return false;
        base = 10;
        exponent = (int32_t)(uint32_t)(uint64_t)parsed;   // int32(parsed): truncation
        fmt = Fmt::DecimalExponent;
        return true;
    }
    return false;
}

// math.go:78-88
bool int64_multiply(int64_t a, int64_t b, int64_t &out)
{
    if (a == 0 || b == 0 || a == 1 || b == 1) { out = (int64_t)((uint64_t)a * (uint64_t)b); return true; }
    if (a == kI64Min || b == kI64Min) return false;
    __int128 c = (__int128)a * (__int128)b;
    if (c > (__int128)kI64Max || c < (__int128)kI64Min) return false;
    out = (int64_t)c;
    return true;
}

}  // namespace

// amount.go:95-101 with scale 0; the dropped `ok` (quantity.go:713-714) turns overflow into 0.
int64_t amount_value(int64_t value, int32_t scale)
{
    if (scale == 0) return value;
    if (scale < 0) {
        const int64_t k = -(int64_t)scale;
        if (k > 18) return value > 0 ? 1 : (value < 0 ? -1 : 0);
        const int64_t d = kPow10[k];
        int64_t q = value / d;
        if (value % d != 0) q += value > 0 ? 1 : -1;      // round away from zero (math.go:173-203)
        return q;
    }
    if (scale > 18) return 0;                              // 0 stays 0, anything else overflows (math.go:158-167)
    if (value == kI64Min) return 0;                        // mostNegative: (0,false) for every scale >= 1
    __int128 p = (__int128)value * (__int128)kPow10[scale];
    // scales 1,2,3,6,9 are one checked multiply whose WRAPPED product is returned
    // with ok=false (math.go:118-140); the looped scales return 0 on overflow
    if (scale == 1 || scale == 2 || scale == 3 || scale == 6 || scale == 9) return (int64_t)(uint64_t)p;
    if (p > (__int128)kI64Max || p < (__int128)kI64Min) return 0;
    return (int64_t)p;
}

bool parse_quantity(std::string_view s, ParsedQuantity &out)
{
    out = ParsedQuantity();
    if (s.empty()) return false;
    if (s == "0") return true;                             // quantity.go:280-282

    Scan sc;
    if (!parse_quantity_string(s, sc)) return false;
    int base;
    int32_t exponent;
    Fmt fmt;
    if (!interpret_suffix(sc.suffix, base, exponent, fmt)) return false;

    int32_t precision = 0;
    int64_t scale = 0;                                     // int32 in the reference; widened, same values
    int64_t mantissa = 1;
    const int32_t digits = (int32_t)(sc.num.size() + sc.denom.size());
    if (fmt == Fmt::BinarySI) {
        scale = 0;
        if (exponent >= 0 && sc.denom.empty()) {
            mantissa = (int64_t)1 << exponent;
            // int32(float32(exponent)*3/10): exponent in {10,...,60} => exact
            precision = 15 - (int32_t)sc.num.size() - (exponent * 3) / 10 - 1;
        } else {
            precision = -1;
        }
    } else {
        scale = exponent;
        precision = 18 - digits;
    }

    if (precision >= 0) {
        scale -= (int64_t)sc.denom.size();
        scale = (int64_t)(int32_t)scale;                   // int32 arithmetic wraps in Go
        if (scale >= -9) {
            // shifted = num + denom: at most 18 digits, ParseInt cannot fail
            int64_t v = 0;
            for (char c : sc.num) v = v * 10 + (c - '0');
            for (char c : sc.denom) v = v * 10 + (c - '0');
            int64_t result;
            if (int64_multiply(v, mantissa, result)) {
                if (!sc.positive) result = -result;
                out.slow = false;
                out.amount = result;
                out.scale = (int32_t)scale;
                out.value = amount_value(result, (int32_t)scale);
                return true;
            }
        }
    }

    // ---- slow path: inf.Dec (quantity.go:346-389) ---------------------------------
    // amount.SetString(value): [sign] digits [. digits]
    std::string_view body = sc.value;
    bool neg = false;
    if (!body.empty() && (body[0] == '+' || body[0] == '-')) { neg = body[0] == '-'; body.remove_prefix(1); }
    std::string digits_all;
    int64_t dscale = 0;
    bool seen_dot = false;
    for (char c : body) {
        if (c == '.') {
            if (seen_dot) return false;
            seen_dot = true;
            continue;
        }
        if (!is_digit(c)) return false;
        digits_all.push_back(c);
        if (seen_dot) dscale++;
    }
    if (digits_all.empty()) return false;                  // "no digits read" => ErrNumeric
    BigDec mag = BigDec::from_digits(digits_all);
    if (mag.is_zero()) neg = false;
    if (base == 10) dscale -= exponent;                    // SetScale(scale + infScale(exponent))
    else mag.mul_pow2((unsigned)exponent);

    // sign handling: work on |amount| (quantity.go:362-365, 386-388)
    if (!mag.is_zero()) {                                  // Round(amount, 9, RoundUp)
        if (dscale < 9) mag.mul_pow10((size_t)(9 - dscale));
        else if (dscale > 9) { if (mag.div_pow10((size_t)(dscale - 9))) mag.add_one(); }
        dscale = 9;
    }
    if (fmt == Fmt::BinarySI) {                            // cap at maxAllowed (math.go:48)
        // amount.Cmp(maxAllowed) > 0  <=>  mag > maxInt64 * 10^dscale
        BigDec lim = max_int64();
        if (dscale > 0) lim.mul_pow10((size_t)dscale);
        bool greater;
        if (dscale >= 0) greater = mag.cmp(lim) > 0;
        else { BigDec m2 = mag; m2.mul_pow10((size_t)(-dscale)); greater = m2.cmp(max_int64()) > 0; }
        if (greater) { mag = max_int64(); dscale = 0; }
    }
    out.slow = true;
    out.scale = 0;
    out.value = scaled_value(mag, neg, dscale);
    out.amount = out.value;
    return true;
}

namespace {
// unicode.IsSpace over UTF-8 input: returns the byte length of a leading /
// trailing space rune at position i (0 if none).
size_t space_len_at(std::string_view s, size_t i)
{
    unsigned char c = (unsigned char)s[i];
    if (c == '\t' || c == '\n' || c == '\v' || c == '\f' || c == '\r' || c == ' ') return 1;
    if (c == 0xC2 && i + 1 < s.size()) {
        unsigned char d = (unsigned char)s[i + 1];
        if (d == 0x85 || d == 0xA0) return 2;
    }
    if (i + 2 < s.size()) {
        unsigned char d = (unsigned char)s[i + 1], e = (unsigned char)s[i + 2];
        if (c == 0xE1 && d == 0x9A && e == 0x80) return 3;                       // U+1680
        if (c == 0xE2 && d == 0x80 && (e >= 0x80 && e <= 0x8A)) return 3;        // U+2000-200A
        if (c == 0xE2 && d == 0x80 && (e == 0xA8 || e == 0xA9 || e == 0xAF)) return 3;   // U+2028/2029/202F
        if (c == 0xE2 && d == 0x81 && e == 0x9F) return 3;                       // U+205F
        if (c == 0xE3 && d == 0x80 && e == 0x80) return 3;                       // U+3000
    }
    return 0;
}

std::string_view go_trim_space(std::string_view s)
{
    size_t a = 0;
    while (a < s.size()) {
        size_t n = space_len_at(s, a);
        if (!n) break;
        a += n;
    }
    size_t b = s.size();
    while (b > a) {
        // find the start of the last rune
        size_t k = b - 1;
        while (k > a && ((unsigned char)s[k] & 0xC0) == 0x80) k--;
        size_t n = space_len_at(s, k);
        if (!n || k + n != b) break;
        b = k;
    }
    return s.substr(a, b - a);
}
}  // namespace

// quantity.go:648-667
bool unmarshal_quantity_json(std::string_view raw, ParsedQuantity &out)
{
    if (raw == "null") { out = ParsedQuantity(); return true; }
    if (raw.size() >= 2 && raw.front() == '"' && raw.back() == '"') raw = raw.substr(1, raw.size() - 2);
    return parse_quantity(go_trim_space(raw), out);
}

void column_encoding(const ParsedQuantity &q, int64_t &val, int8_t &scale)
{
    if (q.slow) { val = q.value; scale = 0; return; }
    if (q.scale > 18 || q.scale < -18) { val = q.value; scale = 0; return; }   // value is fully determined
    val = q.amount;
    scale = (int8_t)q.scale;
}

}  // namespace arena_host
