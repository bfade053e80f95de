"""Quantity string -> int64: the Python oracle against the hand-derived
known-answer table, the C oracle against the Python oracle, and the PRODUCT's
C++ parser (arena_quantity_parse, no GPU needed) against both."""

import ctypes as C
import json
import os
import random

import pytest

from oracle import cref
from oracle import quantity as Q

HERE = os.path.dirname(os.path.abspath(__file__))
KAT = json.load(open(os.path.join(HERE, "golden", "quantity_kat.json")))["vectors"]


def product_parse(lib, raw: str):
    b = raw.encode("utf-8")
    v, a, s, slow = C.c_int64(), C.c_int64(), C.c_int32(), C.c_int()
    rc = lib.arena_quantity_parse(b, len(b), C.byref(v), C.byref(a), C.byref(s), C.byref(slow))
    if rc != 0:
        assert rc == -6, f"unexpected status {rc} for {raw!r}"
        return None
    return v.value, a.value, s.value, bool(slow.value)


@pytest.mark.parametrize("vec", KAT, ids=[v["raw"] for v in KAT])
def test_oracle_matches_hand_vectors(vec):
    if vec.get("error"):
        with pytest.raises(Q.QuantityError):
            Q.value_of_json(vec["raw"])
    else:
        assert Q.value_of_json(vec["raw"]) == vec["value"]


@pytest.mark.parametrize("vec", KAT, ids=[v["raw"] for v in KAT])
def test_product_parser_matches_hand_vectors(built, vec):
    from arena_b200 import _lib
    got = product_parse(_lib.lib(), vec["raw"])
    if vec.get("error"):
        assert got is None
    else:
        assert got is not None and got[0] == vec["value"]


def test_c_oracle_value_matches_python_oracle():
    rnd = random.Random(7)
    cases = [(0, 0), (1, 18), (10, 18), (-(2**63), 0), (-(2**63), 1), (2**63 - 1, 0), (2**63 - 1, -18), (500, -3),
             (-500, -3), (1, -30), (-1, -30), (0, -30), (5, 19), (0, 40), (922337203685477580, 1), (922337203685477581, 1)]
    for _ in range(20000):
        cases.append((rnd.choice([rnd.randrange(-2**63, 2**63), rnd.randrange(-5000, 5000)]), rnd.randrange(-21, 22)))
    for v, s in cases:
        assert cref.quantity_value(v, s) == Q.value_of_fixed(v, s), (v, s)


def _random_quantity(rnd):
    sign = rnd.choice(["", "", "", "-", "+"])
    ndig = rnd.choice([0, 1, 1, 2, 3, 6, 12, 18, 19, 20, 25])
    num = "".join(rnd.choice("0123456789") for _ in range(ndig))
    if rnd.random() < 0.3:
        num = "0" * rnd.randrange(0, 3) + num
    frac = ""
    if rnd.random() < 0.4:
        frac = "." + "".join(rnd.choice("0123456789") for _ in range(rnd.choice([0, 1, 2, 3, 9, 10, 19])))
    suf = rnd.choice(["", "", "", "m", "k", "M", "G", "T", "P", "E", "n", "u", "Ki", "Mi", "Gi", "Ti", "Pi", "Ei",
                      "e3", "E-3", "e+2", "e-12", "e18", "e19", "e0", "e", "i", "K", "mm", "e3e", "Kix", "e-", "E+"])
    s = sign + num + frac + suf
    if rnd.random() < 0.05:
        s = " " + s + "\t"
    return s


def test_product_parser_matches_oracle_on_random_strings(built):
    from arena_b200 import _lib
    lib = _lib.lib()
    rnd = random.Random(20180723)
    n_err = n_slow = 0
    for _ in range(30000):
        s = _random_quantity(rnd)
        raw = '"' + s + '"'
        try:
            q = Q.unmarshal_json(raw)
            want = Q.value(q)
        except Q.QuantityError:
            want = None
        got = product_parse(lib, raw)
        if want is None:
            n_err += 1
            assert got is None, f"{raw}: product accepted, oracle rejected"
            continue
        assert got is not None, f"{raw}: product rejected, oracle accepted"
        assert got[0] == want, f"{raw}: product {got[0]} oracle {want}"
        assert got[3] == (not q.is_fast), raw
        if q.is_fast:
            assert (got[1], got[2]) == (q.i_value, q.i_scale), raw
        else:
            n_slow += 1
    assert n_err > 1000 and n_slow > 1000       # the generator reaches both paths
