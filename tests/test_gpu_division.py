"""The guided filter divides without the IEEE routines' range check (k_arith.cuh: div_rn_normal,
rcp_rn_normal), and the percentile stretches divide by a loop-invariant divisor through its correctly rounded
reciprocal (k_arith.cuh: InvDiv, float32 and float64).  Their results must have the bits of the compiler's own
__ddiv_rn / __drcp_rn / __fdiv_rn on the operand ranges the kernels guarantee; the library compares them on the device."""
import ctypes as C

import pytest

from underwater_image_enhancement_b200 import _lib

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("seed,count", [(1, 1 << 26), (20261104, 1 << 26), (7, 1 << 29)])
def test_branch_free_division_matches_ieee(seed, count):
    bad = C.c_ulonglong(123)
    _lib.check(_lib.load().uwie_selftest_division(count, seed, C.byref(bad)))
    assert bad.value == 0
