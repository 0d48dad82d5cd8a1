"""E6: the device port of libstdc++'s std::nth_element (csrc/orbx_kernels_select.inl) against the REAL std::nth_element
(oracle/orb_oracle.cc::retain_best calls it).  The reference keeps the first n entries in the order nth_element leaves
them (src/ORBextractor.cc:683-685, 697-701), so the ORDER is compared, not just the set.  Both retainBest variants
(OpenCV 2.4: nth = n, >= 3.4: nth = n - 1) and both scan forms of the port."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _cases(oracle):
    rng = np.random.default_rng(42)
    lists, keep = [], []

    def add(r, k):
        lists.append(np.asarray(r, np.float32)); keep.append(int(k))

    for n in list(range(1, 40)) + [63, 64, 65, 100, 217, 500, 1000, 1999, 2000]:
        for _ in range(3):
            k = int(rng.integers(1, n + 1))
            add(rng.integers(7, 120, n), k)                       # FAST scores: small integers, many ties
            add(rng.random(n), k)                                  # all distinct
        add(np.full(n, 31.0), max(1, n // 2))                      # all ties
        add(np.arange(n), max(1, n // 3))                          # ascending
        add(np.arange(n)[::-1], max(1, n // 3))                    # descending
        add(rng.integers(0, 2, n), max(1, n - 1))                  # two values
        add(np.r_[np.arange(n // 2), np.arange(n - n // 2)[::-1]], max(1, n // 2))   # organ pipe
    # adversarial: libstdc++'s own pivot choices turned against it (depth limit -> __heap_select)
    for n, nth in ((200, 100), (1000, 217), (2000, 1000), (2000, 60)):
        r, ncmp = oracle.nth_element_killer(n, nth)
        assert ncmp > 6 * n, (n, nth, ncmp)                        # a healthy quickselect needs ~3 n comparisons
        add(r, nth)                                                # retain_mode 0: nth_element(begin + keep)
        add(r, nth + 1)                                            # retain_mode 1: nth_element(begin + keep - 1)
    return lists, keep


@pytest.mark.parametrize("retain_mode", [0, 1])
@pytest.mark.parametrize("batch", [0, 1])
def test_nth_element_port_equals_libstdcxx(oracle, retain_mode, batch):
    from orb_slam_b200 import ORBextractor
    ex = ORBextractor(max_batch=1)
    lists, keep = _cases(oracle)
    got = ex.debug_retain_best(lists, keep, retain_mode, batch)
    nchecked = 0
    for r, k, order in zip(lists, keep, got):
        assert sorted(order.tolist()) == list(range(len(r)))          # a permutation of the list
        ref_r, ref_tag = oracle.retain_best(r, k, retain_mode)
        m = min(k, len(r))
        assert len(ref_tag) == m
        assert np.array_equal(order[:m], ref_tag), (len(r), k, retain_mode, batch)
        nchecked += 1
    assert nchecked > 450
