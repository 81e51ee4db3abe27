"""fb2_vec_dot_no_alloc (no counterpart in the reference): the allocation-free
dot product whose block partials live in a file-scope __device__ array, called through the C-ABI and compared with the existing
fb2_vec_dot on the same device vectors, two calls in a row (the scratch array keeps the previous call's contents).
Collected last on purpose (a failure here must not hide other tests)."""
import ctypes as C

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

RTOL = 1e-12      # the bar of the other dot / norm tests: relative to |x|_2 |y|_2 (summation order differs between the two kernels)


def _dots(fb2, s, x, y):
    L = fb2._lib.lib()
    vx, vy = s.create_dof_values(), s.create_dof_values()
    vx.set_entries(x)
    vy.set_entries(y)
    ref = C.c_double(0)
    fb2._lib.check(L.fb2_vec_dot(vx.h, vy.h, C.byref(ref)))
    L.fb2_vec_device_ptr.restype = C.c_void_p
    px, py = L.fb2_vec_device_ptr(vx.h), L.fb2_vec_device_ptr(vy.h)
    assert px and py
    out = C.c_double(-1.0)
    fb2._lib.check(L.fb2_vec_dot_no_alloc(C.c_void_p(px), C.c_void_p(py), C.c_int64(x.size), C.byref(out)))
    return ref.value, out.value, (vx, vy)


def test_no_alloc_dot_matches_vec_dot_two_calls_in_a_row(fb2):
    s_big = fb2.FeSpace((40, 36, 33), order=2)        # 455 k DOFs: 112 block partials
    s_small = fb2.FeSpace((5, 4, 3), order=1)         # fewer DOFs than one chunk: stale partials of the first call stay behind it
    rng = np.random.default_rng(2026)
    keep = []
    for s in (s_big, s_small, s_big):
        n = s.num_free_dofs
        x, y = rng.standard_normal(n), rng.standard_normal(n)
        ref, out, vecs = _dots(fb2, s, x, y)
        keep.append(vecs)
        scale = np.linalg.norm(x) * np.linalg.norm(y)
        assert abs(out - ref) <= RTOL * scale, (n, out, ref)
        assert abs(out - float(np.dot(x, y))) <= RTOL * scale
        # <x, x>: no cancellation, plain relative bar
        ref2, out2, vecs2 = _dots(fb2, s, x, x)
        keep.append(vecs2)
        assert abs(out2 - ref2) <= RTOL * abs(ref2)
    # the same inputs again: bitwise repeatable (fixed summation order, nothing left over from the calls in between)
    n = s_big.num_free_dofs
    x, y = rng.standard_normal(n), rng.standard_normal(n)
    _, a, v1 = _dots(fb2, s_big, x, y)
    _, b, v2 = _dots(fb2, s_big, x, y)
    assert a == b


def test_no_alloc_dot_rejects_sizes_beyond_the_static_scratch(fb2):
    L = fb2._lib.lib()
    s = fb2.FeSpace((4, 4, 4), order=1)
    v = s.create_dof_values()
    L.fb2_vec_device_ptr.restype = C.c_void_p
    p = L.fb2_vec_device_ptr(v.h)
    out = C.c_double(123.0)
    too_long = 268435456 + 1                          # FB2_NO_ALLOC_MAX_N + 1 (include/fempar_b200.h); rejected before any launch
    rc = L.fb2_vec_dot_no_alloc(C.c_void_p(p), C.c_void_p(p), C.c_int64(too_long), C.byref(out))
    assert rc != 0 and out.value == 123.0
    assert b"static scratch" in L.fb2_last_error_string()
    # empty vectors are fine and the entry point still works after the rejected call
    fb2._lib.check(L.fb2_vec_dot_no_alloc(None, None, C.c_int64(0), C.byref(out)))
    assert out.value == 0.0
    v.set_entries(np.full(s.num_free_dofs, 2.0))
    fb2._lib.check(L.fb2_vec_dot_no_alloc(C.c_void_p(p), C.c_void_p(p), C.c_int64(s.num_free_dofs), C.byref(out)))
    assert out.value == 4.0 * s.num_free_dofs
