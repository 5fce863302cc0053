"""sdrx_agc_adjust_noalloc (include/sdrx_b200.h): the AGC entry point that allocates nothing -- its scratch is one fixed device
array inside the library.  No counterpart in the reference.  Checked through the C
ABI against the existing entry point (sdrx_agc_adjust, agc.hpp:48-69) on two calls in a row: the second call starts from the
state the first one left and finds the first call's data in the shared array.  The file is named to be collected last."""
import numpy as np
import pytest

from sdrx_b200.objects import AGC, AGC_NOALLOC_MAX

pytestmark = pytest.mark.gpu


def _signal(n, seed):
    rng = np.random.default_rng(seed)
    env = np.repeat([0.01, 2.0, 0.0005, 0.3], (n + 3) // 4)[:n]
    return ((rng.standard_normal(n) + 1j * rng.standard_normal(n)) * env).astype(np.complex64)


def test_agc_adjust_noalloc_matches_agc_adjust_on_two_calls_in_a_row():
    a, b = AGC(), AGC()                                    # same parameters, same initial state
    first, second = _signal(AGC_NOALLOC_MAX, 21), _signal(3001, 22)    # a full array, then a shorter call over its leftovers
    for x in (first, second):
        want = a.adjust(x)                                 # sdrx_agc_adjust: grow-only cudaMalloc'ed scratch
        got = b.adjust_noalloc(x)                          # sdrx_agc_adjust_noalloc: the static device array
        # the bar of the AGC tests (test_gpu_objects.py): bit for bit
        assert np.array_equal(got, want)
        assert b.gain() == a.gain() and float(b.s.power) == float(a.s.power)
    assert np.any(want != 0)


def test_agc_adjust_noalloc_rejects_more_than_the_array_holds():
    b = AGC()
    g0 = b.gain()
    with pytest.raises(Exception, match="SDRX_AGC_NOALLOC_MAX"):
        b.adjust_noalloc(_signal(AGC_NOALLOC_MAX + 1, 23))
    assert b.gain() == g0                                  # a refused call leaves the state alone
    assert b.adjust_noalloc(np.zeros(0, np.complex64)).size == 0
