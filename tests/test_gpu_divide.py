"""The exact-math force pass divides through shared refined reciprocals (csrc/common.cuh div_by); this checks on the
B200 that the quotient is the correctly rounded one (== the oracle's `/`) over every divisor significand."""
import ctypes as C

import pytest

pytestmark = pytest.mark.gpu


def test_shared_reciprocal_division_is_correctly_rounded(fluid):
    lib = fluid.load()
    tested, bad, inexact = C.c_uint64(0), C.c_uint64(0), C.c_uint64(0)
    first = (C.c_float * 4)()
    rc = lib.fluid_selftest_division(0, 0xD1F1D, 24, C.byref(tested), C.byref(bad), C.byref(inexact), first)
    assert rc == 0, lib.fluid_last_error()
    assert tested.value == 13 * (1 << 23) * 24
    assert bad.value == 0, f"{bad.value} of {tested.value} quotients differ from __fdiv_rn; first: a={first[0]!r} b={first[1]!r} want={first[2]!r} got={first[3]!r}"
    print(f"division self-test: {tested.value} quotients exact; refined reciprocal != RN(1/b) for {inexact.value} of {13 << 23} divisors")
