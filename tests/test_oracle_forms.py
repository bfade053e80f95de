"""The oracle's three forms of `top node` agree with one another: the faithful nested loop over
array-of-structs (top_node.go:77-97), the string-hash join over the same structs, and the columnar
pass (single thread and one thread per node range).  The structs spread every limit over the
spellings gpu.go:61-79 has to see through -- deprecated key only, both keys with a decoy under the
deprecated one, an unrelated key in front -- so the CPU baseline bench.py times exercises the
key-selection rule, not just the JSON fixture."""

import numpy as np
import pytest

from oracle import cref


@pytest.mark.parametrize("P,N,seed,skewed", [(20_000, 200, 1, False), (3_000, 7, 2, False), (50_000, 400, 5, True)])
def test_aos_forms_equal_the_columnar_pass(built, P, N, seed, skewed):
    from arena_b200 import HostSnapshot
    h = HostSnapshot.synth(P, N, seed=seed, skewed=skewed)
    tot, alloc, cl, _ = cref.top_node_columnar(h.cols)
    t2, a2, c2 = cref.top_node_columnar_threads(h.cols, 5)
    assert np.array_equal(tot, t2) and np.array_equal(alloc, a2) and np.array_equal(cl, c2)
    aos = cref.AoS(h.cols)
    for nested, threads in ((True, 1), (False, 1), (False, 4)):
        t3, a3, c3 = aos.run(nested, threads)
        assert np.array_equal(tot, t3) and np.array_equal(alloc, a3) and np.array_equal(cl, c3), (nested, threads)
    aos.close()
    assert alloc.sum() > 0 and tot.sum() > 0
