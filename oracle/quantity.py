"""ORACLE (test infrastructure, never shipped or timed as the product).

CPU restatement, in plain Python with arbitrary-precision ints, of the
reference's `resource.Quantity` string -> int64 arithmetic.  Only `tests/`,
`__graft_entry__.smoke()` and `bench.py`'s cpu_baseline / `--impl reference`
legs may import this module.

Everything below follows the vendored k8s.io/apimachinery `release-1.9` sources
(paths relative to /root/reference/vendor/k8s.io/apimachinery/pkg/api/resource/
unless stated otherwise):

  parse_quantity_string   quantity.go:159-273   (parseQuantityString)
  interpret_suffix        suffix.go:88-106, 108-135, 180-198
  parse_quantity          quantity.go:276-390   (ParseQuantity)
  unmarshal_json          quantity.go:648-667   (Quantity.UnmarshalJSON)
  value                   quantity.go:699-718   (Value / ScaledValue)
  as_scaled_int64         amount.go:95-101
  positive_scale_int64    math.go:144-168 (+ int64MultiplyScale* math.go:92-140)
  negative_scale_int64    math.go:173-203
  scaled_value            scale_int.go:43-95
  inf.Dec SetString/Round vendor/gopkg.in/inf.v0/dec.go:243-335, 455-537;
                          RoundUp vendor/gopkg.in/inf.v0/rounder.go:109-116

Parity status: the reference holds NO test or golden vector for this code
(vendored tests are pruned, Gopkg.toml:56-58) and no Go toolchain exists in the
build container, so the known-answer table in tests/golden/quantity_kat.json
was derived by hand from the lines cited above ("parity unpinned" by the
reference's own tests; pinned by hand-derived vectors).  The only answers the
reference itself states are the worked examples in its doc comments
(quantity.go:52, 55-58, 69-74, 524-526; amount.go:90-94; math.go:170-172;
suffix.go:88-106): tests/test_reference_examples.py holds this module, the C
oracle and the product's parser to them.
"""

from __future__ import annotations

from dataclasses import dataclass
from typing import Optional, Tuple

INT64_MAX = (1 << 63) - 1
INT64_MIN = -(1 << 63)
_MASK64 = (1 << 64) - 1

NANO = -9                 # amount.go:37
MAX_INT64_FACTORS = 18    # math.go:28


class QuantityError(ValueError):
    """ErrFormatWrong / ErrNumeric / ErrSuffix (quantity.go:149-153)."""


def wrap_i64(x: int) -> int:
    """Go int64 arithmetic wraps modulo 2**64."""
    x &= _MASK64
    return x - (1 << 64) if x >> 63 else x


def go_quo(a: int, b: int) -> int:
    """Go `/` on integers truncates toward zero."""
    q = abs(a) // abs(b)
    return -q if (a < 0) != (b < 0) else q


def go_rem(a: int, b: int) -> int:
    """Go `%`: sign follows the dividend."""
    return a - b * go_quo(a, b)


def big_int64(x: int) -> int:
    """math/big (*Int).Int64(): low 64 bits of |x|, negated when x < 0."""
    v = wrap_i64(abs(x) & _MASK64)
    return wrap_i64(-v) if x < 0 else v


# --------------------------------------------------------------------------
# Quantity value: either int64Amount{value, scale} or an inf.Dec (unscaled big
# int, inf scale).  quantity.go:98-109
# --------------------------------------------------------------------------
@dataclass(frozen=True)
class Quantity:
    i_value: int = 0              # int64Amount.value
    i_scale: int = 0              # int64Amount.scale (power of ten)
    d_unscaled: Optional[int] = None   # inf.Dec unscaled (None <=> d.Dec == nil)
    d_scale: int = 0              # inf.Dec scale: value = unscaled * 10**(-d_scale)

    @property
    def is_fast(self) -> bool:
        return self.d_unscaled is None


_DIGITS = "0123456789"
_SUFFIX_CHARS = "eEinumkKMGTP"


def parse_quantity_string(s: str) -> Tuple[bool, str, str, str, str]:
    """quantity.go:159-273.  Returns (positive, value, num, denom, suffix)."""
    positive = True
    pos = 0
    end = len(s)
    num = denom = suffix = ""

    if pos < end:                       # :165-174
        if s[0] == "-":
            positive = False
            pos += 1
        elif s[0] == "+":
            pos += 1

    i = pos                             # :176-190  strip leading zeros
    while True:
        if i >= end:
            return positive, "0", "0", "", ""
        if s[i] == "0":
            pos += 1
        else:
            break
        i += 1

    i = pos                             # :192-208  numerator
    while True:
        if i >= end:
            return positive, s[0:end], s[pos:end], "", ""
        if s[i] not in _DIGITS:
            num = s[pos:i]
            pos = i
            break
        i += 1

    if len(num) == 0:                   # :210-213
        num = "0"

    if pos < end and s[pos] == ".":     # :215-238  denominator
        pos += 1
        i = pos
        while True:
            if i >= end:
                return positive, s[0:end], num, s[pos:end], ""
            if s[i] not in _DIGITS:
                denom = s[pos:i]
                pos = i
                break
            i += 1
    value = s[0:pos]                    # :239

    suffix_start = pos                  # :241-252
    i = pos
    while True:
        if i >= end:
            return positive, value, num, denom, s[suffix_start:end]
        if s[i] not in _SUFFIX_CHARS:
            pos = i
            break
        i += 1
    if pos < end and s[pos] in "-+":    # :253-258
        pos += 1
    i = pos                             # :259-269
    while True:
        if i >= end:
            return positive, value, num, denom, s[suffix_start:end]
        if s[i] not in _DIGITS:
            break
        i += 1
    raise QuantityError("ErrFormatWrong")   # :270-272


_DEC_SUFFIX = {"n": -9, "u": -6, "m": -3, "": 0, "k": 3, "M": 6, "G": 9,
               "T": 12, "P": 15, "E": 18}                    # suffix.go:122-132
_BIN_SUFFIX = {"Ki": 10, "Mi": 20, "Gi": 30, "Ti": 40, "Pi": 50, "Ei": 60}  # :113-118


def _go_parse_int64(s: str) -> int:
    """strconv.ParseInt(s, 10, 64): optional sign, digits only, range-checked."""
    body = s[1:] if s[:1] in ("+", "-") else s
    if not body or any(c not in _DIGITS for c in body):
        raise QuantityError("ErrSuffix")
    v = int(body)
    if s[:1] == "-":
        v = -v
    if v > INT64_MAX or v < INT64_MIN:
        raise QuantityError("ErrSuffix")
    return v


def _to_int32(x: int) -> int:
    x &= 0xFFFFFFFF
    return x - (1 << 32) if x >> 31 else x


def interpret_suffix(suf: str) -> Tuple[int, int, str]:
    """suffix.go:88-106 then :180-198.  Returns (base, exponent, format)."""
    if suf in _DEC_SUFFIX:
        return 10, _DEC_SUFFIX[suf], "DecimalSI"
    if suf in _BIN_SUFFIX:
        return 2, _BIN_SUFFIX[suf], "BinarySI"
    if len(suf) > 1 and suf[0] in "Ee":
        parsed = _go_parse_int64(suf[1:])
        return 10, _to_int32(parsed), "DecimalExponent"     # int32(parsed) truncates
    raise QuantityError("ErrSuffix")


def int64_multiply(a: int, b: int) -> Tuple[int, bool]:
    """math.go:78-88."""
    if a == 0 or b == 0 or a == 1 or b == 1:
        return wrap_i64(a * b), True
    if a == INT64_MIN or b == INT64_MIN:
        return 0, False
    c = wrap_i64(a * b)
    return c, go_quo(c, b) == a


def _round_up_to_nano(unscaled: int, scale: int) -> Tuple[int, int]:
    """inf.Dec.Round(x, 9, RoundUp) for x = unscaled*10**-scale.
    dec.go:243-245, 289-326; rounder.go:109-116."""
    s = 9
    shift = s - scale
    if shift > 0:
        ix, iy = unscaled * 10 ** shift, 1
    elif shift < 0:
        ix, iy = unscaled, 10 ** (-shift)
    else:
        ix, iy = unscaled, 1
    q = go_quo(ix, iy)                      # big.Int.QuoRem truncates
    r = go_rem(ix, iy)
    if r != 0:
        sgn_r = 1 if r > 0 else -1
        sgn_den = 1 if iy > 0 else -1
        q += sgn_r * sgn_den                # intSign[rA.Sign()*rB.Sign()+1]
    return q, s


def parse_quantity(s: str) -> Quantity:
    """quantity.go:276-390."""
    if len(s) == 0:
        raise QuantityError("ErrFormatWrong")
    if s == "0":
        return Quantity()

    positive, value, num, denom, suf = parse_quantity_string(s)
    base, exponent, fmt = interpret_suffix(suf)

    precision = 0
    scale = 0
    mantissa = 1
    if fmt in ("DecimalExponent", "DecimalSI"):             # :298-300
        scale = exponent
        precision = MAX_INT64_FACTORS - (len(num) + len(denom))
    else:                                                   # BinarySI :301-312
        scale = 0
        if exponent >= 0 and len(denom) == 0:
            mantissa = 1 << exponent
            # int32(float32(exponent)*3/10): exact for the table's exponents
            precision = 15 - len(num) - int(exponent * 3 / 10) - 1
        else:
            precision = -1

    if precision >= 0:                                      # :314-343
        scale -= len(denom)
        if scale >= NANO:
            shifted = num + denom
            v = int(shifted)        # <= 18 digits: ParseInt cannot fail
            result, ok = int64_multiply(v, mantissa)
            if ok:
                if not positive:
                    result = -result
                return Quantity(i_value=result, i_scale=scale)

    # slow path :346-389  (inf.Dec)
    unscaled, dscale = _dec_set_string(value)
    if base == 10:
        dscale = dscale + (-exponent)                       # :352-353
    elif base == 2:
        unscaled = unscaled * (1 << exponent)               # :354-359
    sign = (unscaled > 0) - (unscaled < 0)
    if sign == -1:
        unscaled = -unscaled
    if unscaled != 0:                                       # :371-373 (v != 0 || !ok)
        unscaled, dscale = _round_up_to_nano(unscaled, dscale)
    if fmt == "BinarySI" and _dec_cmp(unscaled, dscale, INT64_MAX, 0) > 0:   # :377-379
        unscaled, dscale = INT64_MAX, 0
    if sign == -1:
        unscaled = -unscaled
    return Quantity(d_unscaled=unscaled, d_scale=dscale)


def _dec_cmp(ua: int, sa: int, ub: int, sb: int) -> int:
    m = max(sa, sb)
    a = ua * 10 ** (m - sa)
    b = ub * 10 ** (m - sb)
    return (a > b) - (a < b)


def _dec_set_string(value: str) -> Tuple[int, int]:
    """inf.Dec.SetString: dec.go:455-537.  `value` is sign, digits and at most
    one '.', as produced by parse_quantity_string."""
    body = value
    neg = False
    if body[:1] in ("+", "-"):
        neg = body[0] == "-"
        body = body[1:]
    if body.count(".") > 1:
        raise QuantityError("ErrNumeric")
    if "." in body:
        ip, fp = body.split(".")
        digits = ip + fp
        scale = len(fp)
    else:
        digits = body
        scale = 0
    if not digits or any(c not in _DIGITS for c in digits):
        raise QuantityError("ErrNumeric")       # "no digits read"
    u = int(digits)
    return (-u if neg else u), scale


def unmarshal_json(raw: str) -> Quantity:
    """quantity.go:648-667.  `raw` is the raw JSON token text of the value
    (quotes included for strings; escapes NOT decoded, as in the reference)."""
    if raw == "null":
        return Quantity()
    if len(raw) >= 2 and raw[0] == '"' and raw[-1] == '"':
        raw = raw[1:-1]
    return parse_quantity(_go_trim_space(raw))


# unicode.IsSpace (Go): White_Space property
_GO_SPACE = frozenset(
    [0x09, 0x0A, 0x0B, 0x0C, 0x0D, 0x20, 0x85, 0xA0, 0x1680, 0x2028, 0x2029,
     0x202F, 0x205F, 0x3000] + list(range(0x2000, 0x200B)))


def _go_trim_space(s: str) -> str:
    """strings.TrimSpace; Python's str.strip() set differs (it also strips
    \x1c-\x1f), so be explicit."""
    a, b = 0, len(s)
    while a < b and ord(s[a]) in _GO_SPACE:
        a += 1
    while b > a and ord(s[b - 1]) in _GO_SPACE:
        b -= 1
    return s[a:b]


# --------------------------------------------------------------------------
# Value()
# --------------------------------------------------------------------------
def _mul_scale(a: int, b: int) -> Tuple[int, bool]:
    """int64MultiplyScale / Scale10 / Scale100 / Scale1000: math.go:92-140."""
    if a == 0 or a == 1:
        return wrap_i64(a * b), True
    if a == INT64_MIN and b != 1:
        return 0, False
    c = wrap_i64(a * b)
    return c, go_quo(c, b) == a


def positive_scale_int64(base: int, scale: int) -> Tuple[int, bool]:
    """math.go:144-168."""
    if scale == 0:
        return base, True
    if scale == 1:
        return _mul_scale(base, 10)
    if scale == 2:
        return _mul_scale(base, 100)
    if scale == 3:
        return _mul_scale(base, 1000)
    if scale == 6:
        return _mul_scale(base, 1000000)
    if scale == 9:
        return _mul_scale(base, 1000000000)
    value = base
    for _ in range(scale):
        value, ok = _mul_scale(value, 10)
        if not ok:
            return 0, False
    return value, True


def negative_scale_int64(base: int, scale: int) -> Tuple[int, bool]:
    """math.go:173-203."""
    if scale == 0:
        return base, True
    value = base
    fraction = False
    for _ in range(scale):
        if not fraction and go_rem(value, 10) != 0:
            fraction = True
        value = go_quo(value, 10)
        if value == 0:
            if fraction:
                return (1, False) if base > 0 else (-1, False)
            return 0, True
    if fraction:
        value += 1 if base > 0 else -1
    return value, not fraction


def as_scaled_int64(i_value: int, i_scale: int, scale: int) -> int:
    """amount.go:95-101; the `ok` result is dropped by the caller
    (quantity.go:713-714), so overflow yields 0."""
    if i_scale < scale:
        r, _ = negative_scale_int64(i_value, scale - i_scale)
        return r
    r, _ = positive_scale_int64(i_value, i_scale - scale)
    return r


def scaled_value(unscaled: int, scale: int, new_scale: int) -> int:
    """scale_int.go:43-95."""
    dif = scale - new_scale
    if dif == 0:
        return big_int64(unscaled)
    if dif < 0:
        # unscaled.Int64() * int64(math.Pow10(-dif)); on this path the
        # quantity is zero (see parse_quantity), but keep the arithmetic.
        if -dif > 18:
            if unscaled == 0:
                return 0
            raise QuantityError("scaledValue: float->int64 conversion out of range")
        return wrap_i64(big_int64(unscaled) * 10 ** (-dif))
    if unscaled < INT64_MAX and dif < 19:               # :60-69
        divide = 10 ** dif
        u = big_int64(unscaled)
        result = go_quo(u, divide)
        mod = go_rem(u, divide)
        if mod != 0:
            return wrap_i64(result + 1)
        return result
    divisor = 10 ** dif                                 # :71-94, Euclidean DivMod
    q, m = divmod(unscaled, divisor)                    # Python divmod == Euclidean for divisor > 0
    if m != 0:
        return wrap_i64(big_int64(q) + 1)
    return big_int64(q)


def value(q: Quantity) -> int:
    """Quantity.Value() == ScaledValue(0): quantity.go:699-718."""
    if q.is_fast:
        return as_scaled_int64(q.i_value, q.i_scale, 0)
    return scaled_value(q.d_unscaled, q.d_scale, 0)


def value_of_json(raw: str) -> int:
    return value(unmarshal_json(raw))


def value_of_fixed(i_value: int, i_scale: int) -> int:
    """Value() of an int64Amount{value, scale}: what the columnar snapshot's
    (val, scale) pair decodes to."""
    return as_scaled_int64(i_value, i_scale, 0)
