"""The reference's own known answers: the 20 Vec3 unit tests of raytracer/src/vec3.rs:425-564, asserted with exact
equality like the originals (`assert_eq!` on f64).  These pin the oracle's arithmetic layer."""
import ctypes as C

import numpy as np
import pytest

from oracle_ffi import lib

ADD, ADDS, SUB, SUBS, DOT, MULS, DIVS, ELEMUL, CROSS, NEG, SQLEN, LEN, UNIT, REFLECT, REFRACT = range(15)


def op(code, a, b=None, out0=0.0):
    a = np.array(a, np.float64)
    bb = None if b is None else np.array(np.atleast_1d(b), np.float64)
    out = np.array([out0, 0.0, 0.0], np.float64)
    rc = lib().oracle_vec3_op(code, a.ctypes.data_as(C.c_void_p), None if bb is None else bb.ctypes.data_as(C.c_void_p), out.ctypes.data_as(C.c_void_p))
    return rc, out


def vec(code, a, b=None):
    rc, out = op(code, a, b)
    assert rc == 0
    return tuple(out)


def test_new():  # vec3.rs:429-432
    assert vec(ADDS, (1.0, 2.0, 3.0), 0.0) == (1.0, 2.0, 3.0)


def test_add_and_add_assign():  # vec3.rs:434-447
    assert vec(ADD, (1.0, 0.0, -1.0), (2.0, 4.0, 6.0)) == (3.0, 4.0, 5.0)


def test_add_f64_and_assign():  # vec3.rs:449-462
    assert vec(ADDS, (1.0, 0.0, -1.0), 233.0) == (234.0, 233.0, 232.0)


def test_sub_and_sub_assign():  # vec3.rs:464-477
    assert vec(SUB, (1.0, 0.0, -1.0), (2.0, 4.0, 6.0)) == (-1.0, -4.0, -7.0)


def test_sub_f64_and_assign():  # vec3.rs:479-489
    assert vec(SUBS, (1.0, 0.0, -1.0), 1.0) == (0.0, -1.0, -2.0)


def test_mul_is_dot():  # vec3.rs:491-494
    rc, out = op(DOT, (1.0, 0.0, -1.0), (1.0, 1.0, 1.0))
    assert rc == 0 and out[0] == 0.0


def test_mul_assign_and_mul_f64():  # vec3.rs:496-506
    assert vec(MULS, (1.0, 0.0, -1.0), 2.0) == (2.0, 0.0, -2.0)
    assert vec(MULS, (1.0, 0.0, -1.0), 1.0) == (1.0, 0.0, -1.0)


def test_div():  # vec3.rs:508-514
    assert vec(DIVS, (1.0, -2.0, 0.0), 2.0) == (0.5, -1.0, 0.0)


def test_elemul():  # vec3.rs:516-522
    assert vec(ELEMUL, (1.0, 2.0, 3.0), (1.0, 2.0, 3.0)) == (1.0, 4.0, 9.0)


def test_cross():  # vec3.rs:524-530
    assert vec(CROSS, (1.0, 2.0, 3.0), (2.0, 3.0, 4.0)) == (8.0 - 9.0, 6.0 - 4.0, 3.0 - 4.0)


def test_neg():  # vec3.rs:532-535
    assert vec(NEG, (1.0, -2.0, 3.0)) == (-1.0, 2.0, -3.0)


def test_squared_length():  # vec3.rs:537-540
    rc, out = op(SQLEN, (1.0, 2.0, 3.0))
    assert rc == 0 and out[0] == 14.0


def test_length():  # vec3.rs:542-548
    rc, out = op(LEN, (3.0, 4.0, 5.0))
    assert rc == 0 and out[0] == np.sqrt(3.0 * 3.0 + 4.0 * 4.0 + 5.0 * 5.0)


def test_unit():  # vec3.rs:550-557
    assert vec(UNIT, (233.0, 0.0, 0.0)) == (1.0, 0.0, 0.0)
    assert vec(UNIT, (-233.0, 0.0, 0.0)) == (-1.0, 0.0, 0.0)


def test_unit_panics_on_zero():  # vec3.rs:559-563 (#[should_panic])
    rc, _ = op(UNIT, (0.0, 0.0, 0.0))
    assert rc == 1


def test_reflect_refract_closed_form():  # vec3.rs:163-172 (no upstream test; closed-form values)
    assert vec(REFLECT, (1.0, -1.0, 0.0), (0.0, 1.0, 0.0)) == (1.0, 1.0, 0.0)
    rc, out = op(REFRACT, (0.0, -1.0, 0.0), (0.0, 1.0, 0.0), out0=1.0 / 1.5)  # normal incidence passes straight through
    assert rc == 0 and tuple(out) == (0.0, -1.0, 0.0)
