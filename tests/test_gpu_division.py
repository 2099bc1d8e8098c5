"""The guided filter divides without the IEEE routines' range check (k_arith.cuh: div_rn_normal,
rcp_rn_normal).  Their results must have the bits of the compiler's own __ddiv_rn / __drcp_rn on the
operand ranges the kernels guarantee; the library compares them on the device."""
import ctypes as C

import pytest

from underwater_image_enhancement_b200 import _lib

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("seed", [1, 20261104])
def test_branch_free_division_matches_ieee(seed):
    bad = C.c_ulonglong(123)
    _lib.check(_lib.load().uwie_selftest_division(1 << 26, seed, C.byref(bad)))
    assert bad.value == 0
