"""The hot loop replaces `/`, roundf and floorf by cheaper forms that must give the SAME bits
(csrc/exactmath.cuh).  This brute-forces them on the device against the IEEE operators:
~3e9 random + adversarial operand pairs, and the integer-weight colour path exhaustively."""
import ctypes as C

import pytest

pytestmark = pytest.mark.gpu


def test_exact_forms_match_ieee_operators(pkg):
    fails = (C.c_int64 * 5)()
    for seed in (1, 0xB200):
        pkg.capi.check(pkg.load(dev=True).tsdf_selftest_exactmath(0, 5000, seed, fails))   # dev library only
        assert list(fails) == [0, 0, 0, 0, 0], dict(zip(["rcp", "div", "round", "blend", "blend_intw"], fails))
