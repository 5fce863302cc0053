"""The operand split of the tensor-core table GEMM (sdrx_b200/csrc/tgemm_tc.cuh), restated in numpy: an FP32 value is exactly
the sum of three bf16 values, a plane product is exact in an FP32 accumulator, and the three plane products the default drops
are below FP32 resolution of the term.  (The device and host splits themselves are exercised by the GPU parity tests.)"""
import numpy as np


def bf16_rn(x: np.ndarray) -> np.ndarray:
    """float32 -> the nearest bfloat16 (ties to even), returned as float32; finite inputs."""
    u = x.astype(np.float32).view(np.uint32).astype(np.uint64)
    u = (u + 0x7FFF + ((u >> 16) & 1)) & 0xFFFF0000
    return u.astype(np.uint32).view(np.float32)


def split3(x: np.ndarray):
    p0 = bf16_rn(x)
    r1 = (x - p0).astype(np.float32)
    p1 = bf16_rn(r1)
    r2 = (r1 - p1).astype(np.float32)
    return p0, p1, bf16_rn(r2)


def _values(n, seed):
    rng = np.random.default_rng(seed)
    return (rng.standard_normal(n) * 10.0 ** rng.uniform(-6, 2, n)).astype(np.float32)


def test_three_bf16_planes_hold_an_fp32_value_exactly():
    x = np.concatenate([_values(200000, 1), np.float32([0.0, 1.0, -1.0, 1.0 + 2.0 ** -23, 1.0e30, 1.2e-30])])     # (within 2^-8 of FLT_MAX the first plane would round to infinity: tables and folds are far from it)
    p0, p1, p2 = split3(x)
    assert np.array_equal(p0.astype(np.float64) + p1.astype(np.float64) + p2.astype(np.float64), x.astype(np.float64))
    for p in (p0, p1, p2):                                   # each plane really is a bf16 value
        assert not np.any(p.view(np.uint32) & 0xFFFF)


def test_plane_products_are_exact_and_the_dropped_ones_are_below_fp32_resolution():
    a, b = _values(100000, 2), _values(100000, 3)
    pa, pb = split3(a), split3(b)
    exact = a.astype(np.float64) * b.astype(np.float64)
    total = np.zeros_like(exact)
    for i in range(3):
        for j in range(3):
            prod = pa[i].astype(np.float64) * pb[j].astype(np.float64)
            assert np.array_equal(prod.astype(np.float32).astype(np.float64), prod)      # 8 x 8 significant bits: exact in FP32
            total += prod
    assert np.array_equal(total, exact)                                                  # nine products: the exact product
    dropped = sum(pa[i].astype(np.float64) * pb[j].astype(np.float64) for i, j in ((1, 2), (2, 1), (2, 2)))
    ok = exact != 0
    assert np.max(np.abs(dropped[ok] / exact[ok])) < 2.0 ** -23                           # six products: within ~1 ulp of the term
