"""Host logic of the specialised channel bank (bank_fast.cuh): the static work schedule must tile
(slot group x time) exactly once for any channel count / block length / SM count, and stay balanced."""
import ctypes as C

import pytest

from sdrx_b200 import _lib


def _plan(groups, U, sms):
    lib = _lib.load()
    a, b, c = C.c_int(), C.c_int(), C.c_int()
    bad = lib.sdrx_debug_bank_plan_check(groups, U, sms, C.byref(a), C.byref(b), C.byref(c))
    return bad, a.value, b.value, c.value


@pytest.mark.parametrize("groups", [1, 2, 3, 8, 30, 31, 60, 61, 125, 256])
@pytest.mark.parametrize("U", [1, 5, 8, 16, 40, 320, 640, 2560, 2561, 5000])
def test_schedule_tiles_every_output_exactly_once(groups, U):
    for sms in (1, 7, 132, 148):
        bad, n_ctas, n_pieces, _ = _plan(groups, U, sms)
        assert bad == 0, (groups, U, sms, bad)
        assert 1 <= n_ctas <= sms and n_pieces >= n_ctas


def test_full_band_schedule_is_balanced():
    """960 channels as 481 conjugate-pair slots = 31 groups, one 32 ms block = 2560 units, 148 SMs: the busiest CTA
    must stay within 15 % of the ideal share (all SMs finish together, block-start warm-up included)."""
    bad, n_ctas, _, worst = _plan(31, 2560, 148)
    assert bad == 0 and 130 <= n_ctas <= 148
    ideal = 31 * 2560 / (148 * 32)          # units per segment if the work divided evenly over 148 x 32 segments
    assert worst <= 1.15 * (ideal + 1), (worst, ideal)
    bad, n_ctas, _, worst = _plan(60, 2560, 148)
    ideal = 60 * 2560 / (148 * 32)
    assert bad == 0 and worst <= 1.15 * (ideal + 1), (worst, ideal)
