"""Exhaustive check of the reciprocal tables the kernels use to round scaled
quantities (kCeilDiv10 in arena_b200/csrc/kernels.cuh): the table in the source
is parsed and verified for every x < 2^23."""

import os
import re

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _table(name):
    src = open(os.path.join(ROOT, "arena_b200", "csrc", "kernels.cuh")).read()
    body = src[src.index(name + "[10] = {"):]
    body = body[:body.index("};")]
    rows = re.findall(r"\{\s*(\d+)u,\s*(\d+)u,\s*(\d+)u,\s*0u\s*\}", body)
    assert len(rows) == 10
    return [tuple(int(x) for x in r) for r in rows]


def test_ceil_div10_table_is_exact_on_its_whole_domain():
    x = np.arange(1 << 23, dtype=np.uint64)
    for k, (m, sh, dm1) in enumerate(_table("kCeilDiv10")):
        d = 10 ** k
        assert dm1 == d - 1 and m < 2 ** 32
        t = x + np.uint64(dm1)
        assert int(t.max()) * m < 2 ** 64                       # the 64-bit product cannot wrap
        got = (t * np.uint64(m)) >> np.uint64(sh)
        assert np.array_equal(got, (x + np.uint64(d - 1)) // np.uint64(d)), k


def test_packed_table_matches_and_shift_formula():
    src = open(os.path.join(ROOT, "arena_b200", "csrc", "kernels.cuh")).read()
    body = src[src.index("kCeilDiv10Packed[16] = {"):]
    pairs = re.findall(r"\{(\d+)u, (\d+)u\}", body[:body.index("};")])
    assert len(pairs) == 16
    full = _table("kCeilDiv10")
    for k in range(1, 10):
        m, sh, dm1 = full[k]
        assert (int(pairs[k][0]), int(pairs[k][1])) == (m, dm1)
        assert sh >= 35 and sh - 32 == (k * 3402) >> 10          # the kernel derives the shift from k
    # (x + dm1) * m >> sh == umulhi(x + dm1, m) >> (sh - 32) needs x + dm1 < 2^32
    assert (1 << 23) + 999_999_999 < 2 ** 32


def test_recip10_table():
    src = open(os.path.join(ROOT, "arena_b200", "csrc", "kernels.cuh")).read()
    body = src[src.index("kRecip10[10] = {"):]
    vals = [int(v) for v in re.findall(r"(\d+)ull", body[:body.index("};")])]
    assert vals[0] == 0
    for k in range(1, 10):
        assert vals[k] == (1 << 64) // 10 ** k + 1
