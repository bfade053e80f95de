"""The worked examples the REFERENCE itself states, as known answers.

/root/reference ships no test files (the vendored k8s.io/apimachinery tree is pruned of *_test.go) and the
Go toolchain is absent, so nothing executable pins the oracle.  What the reference does hold are the examples
its own doc comments spell out for the functions on this path; they are few, but they are the reference's
words, not ours.  Each is checked against the Python oracle (oracle/quantity.py), the C oracle
(oracle/acct_ref.c through oracle.cref) and the product's host parser (arena_quantity_parse); the kernels'
decode is checked against the same C oracle by tests/test_gpu_top_node.py.

Paths are relative to /root/reference/vendor/k8s.io/apimachinery/pkg/api/resource/."""

import ctypes as C

import pytest

from oracle import cref
from oracle import quantity as Q


def product_value(built, raw: str):
    from arena_b200 import lib
    b = raw.encode()
    v, a, s, slow = C.c_int64(), C.c_int64(), C.c_int32(), C.c_int()
    rc = lib().arena_quantity_parse(b, len(b), C.byref(v), C.byref(a), C.byref(s), C.byref(slow))
    assert rc == 0, raw
    return v.value, a.value, s.value


def test_more_than_three_decimal_places_round_up(built):
    """quantity.go:55-58: "nor may it have more than 3 decimal places.  Numbers larger or more precise will be
    capped or rounded up.  (E.g.: 0.1m will rounded up to 1m.)"  -- since this version keeps nano precision
    (quantity.go:360-373 rounds to scale -9), 0.1m = 1e-4 is held exactly and only Value() rounds it up; one
    digit past nano is what gets rounded up on the way in."""
    q = Q.parse_quantity("0.1m")
    assert (q.i_value, q.i_scale) == (1, -4) and Q.value(q) == 1
    assert product_value(built, "0.1m")[0] == 1
    q = Q.parse_quantity("0.0000000001")                 # 1e-10: below nano, rounded UP to 1n on the way in
    assert not q.is_fast and (q.d_unscaled, q.d_scale) == (1, 9) and Q.value(q) == 1
    assert product_value(built, "0.0000000001")[0] == 1


def test_canonical_form_examples(built):
    """quantity.go:69-74: "1.5 will be serialized as "1500m"; 1.5Gi will be serialized as "1536Mi"": the parsed
    amounts are 1500 * 10^-3 and 1536 * 2^20; Value() (quantity.go:699-718) rounds the first up."""
    q = Q.parse_quantity("1.5")
    assert q.i_value * 10 ** (q.i_scale + 3) == 1500 and Q.value(q) == 2
    assert product_value(built, "1.5")[0] == 2
    q = Q.parse_quantity("1.5Gi")
    assert Q.value(q) == 1536 * 2 ** 20
    assert product_value(built, "1.5Gi")[0] == 1536 * 2 ** 20
    assert Q.value(Q.parse_quantity("1500m")) == 2 and Q.value(Q.parse_quantity("1536Mi")) == 1536 * 2 ** 20


def test_rounding_is_away_from_zero(built):
    """quantity.go:524-526 (RoundUp): "Negative numbers are rounded away from zero (-9 scale 1 rounds to -10)";
    math.go:170-172 (negativeScaleInt64): "The value returned, if not exact, is rounded away from zero."
    -9 taken one decimal scale up is -0.9 tens -> -1 ten."""
    assert Q.negative_scale_int64(-9, 1) == (-1, False)
    assert cref.quantity_value(-9, -1) == -1 and Q.value_of_fixed(-9, -1) == -1
    assert product_value(built, "-0.9")[0] == -1
    assert cref.quantity_value(9, -1) == 1 and product_value(built, "0.9")[0] == 1


def test_as_scaled_int64_examples(built):
    """amount.go:90-94: "(1e20).AsScaledInt64(1) would result in overflow because 1e19 is not representable as an
    int64 ... (1e-6).AsScaledInt64(0) would return 1, because 0.000001 is rounded up to 1."  Value() drops the
    ok flag (quantity.go:713-714), so the overflow reads as 0."""
    assert Q.as_scaled_int64(1, -6, 0) == 1 and cref.quantity_value(1, -6) == 1
    assert product_value(built, "0.000001")[0] == 1 and product_value(built, "1u")[0] == 1
    r, ok = Q.positive_scale_int64(1, 19)                   # int64Amount{1, 20} at scale 1: times 10^19
    assert not ok
    assert Q.as_scaled_int64(1, 20, 0) == 0 and cref.quantity_value(1, 20) == 0


@pytest.mark.parametrize("suffix,exp", [("n", -9), ("u", -6), ("m", -3), ("", 0), ("k", 3), ("M", 6), ("G", 9), ("T", 12),
                                        ("P", 15), ("E", 18)])
def test_decimal_si_table(built, suffix, exp):
    """suffix.go:88-106, the decimalSI table, and quantity.go:50-52 "<decimalSI> ::= m | "" | k | M | G | T | P | E
    (Note that 1024 = 1Ki but 1000 = 1k ...)"."""
    q = Q.parse_quantity("3" + suffix)
    assert q.is_fast and q.i_value * 10 ** q.i_scale == 3 * 10 ** exp if exp >= 0 else (q.i_value, q.i_scale) == (3, exp)
    want = 3 * 10 ** exp if exp >= 0 else 1                 # 3n, 3u, 3m round up to 1
    assert Q.value(q) == want and product_value(built, "3" + suffix)[0] == want


def test_1024_is_1ki_but_1000_is_1k(built):
    """quantity.go:52."""
    assert Q.value(Q.parse_quantity("1Ki")) == 1024 and Q.value(Q.parse_quantity("1k")) == 1000
    assert product_value(built, "1Ki")[0] == 1024 and product_value(built, "1k")[0] == 1000
