"""GPU: ``ba_triangulate_static`` - the no-allocation variant of ``ba_triangulate`` (scratch in a file-scope ``__device__``
array, see include/ba_b200.h) - against ``ba_triangulate`` through the C ABI, on two calls in a row.

Named test_zz_* so that pytest collects it last: a failure here must not hide the other
tests.  Runs alone in a fresh process (it creates its own handle) and after the other tests (it shares nothing with them).
Tolerance: both entry points launch the same kernel on the same inputs, so the outputs must agree bit for bit; the assertion
uses the 1e-12 absolute bound of the float64 triangulation tests plus exact equality of the cheirality flags.
"""
import ctypes as C

import numpy as np
import pytest

from bundle_adjustment_b200 import FlatSolver, _capi as api, synthetic as S

pytestmark = pytest.mark.gpu


def _two_view(n, seed):
    rng = np.random.default_rng(seed)
    K = S.CAMERA_MATRIX
    rvec = np.array([0.02, -0.05, 0.01])
    R = S.rodrigues(rvec)
    t = np.array([0.4, 0.02, -0.03])
    X = np.stack([rng.uniform(-3, 3, n), rng.uniform(-2, 2, n), rng.uniform(6, 20, n)], -1)
    X[::7, 2] *= -1.0                                      # some points behind the cameras: the cheirality flag must differ
    P1 = K @ np.hstack((np.eye(3), np.zeros((3, 1))))
    P2 = K @ np.hstack((R, t.reshape(3, 1)))

    def proj(P):
        h = (P @ np.hstack((X, np.ones((n, 1)))).T).T
        return h[:, :2] / h[:, 2:3] + rng.normal(0, 0.3, (n, 2))
    return P1, P2, R, t, proj(P1), proj(P2)


def _call(fn, solver, P1, P2, R, t, p1, p2):
    n = p1.shape[0]
    hom, xyz, valid = np.full((n, 4), np.nan), np.full((n, 3), np.nan), np.full(n, 255, dtype=np.uint8)
    P1, P2, R, t = api.f64(P1).reshape(12), api.f64(P2).reshape(12), api.f64(R).reshape(9), api.f64(t).reshape(3)
    p1, p2 = api.f64(p1), api.f64(p2)
    api.check(fn(solver._h, api.dptr(P1), api.dptr(P2), api.dptr(R), api.dptr(t), n, api.dptr(p1), api.dptr(p2),
                 api.dptr(hom), api.dptr(xyz), valid.ctypes.data_as(C.POINTER(C.c_uint8))))
    return hom, xyz, valid


def test_triangulate_static_matches_triangulate_on_two_calls_in_a_row():
    L = api.lib()
    with FlatSolver(0) as solver:
        # second call: FEWER correspondences and another geometry, so leftovers of the first call in the shared array would show
        for n, seed in ((5000, 1), (1234, 2)):
            args = _two_view(n, seed)
            ref = _call(L.ba_triangulate, solver, *args)
            got = _call(L.ba_triangulate_static, solver, *args)
            assert np.isfinite(got[0]).all() and np.isfinite(got[1]).all()
            assert np.abs(got[0] - ref[0]).max() <= 1e-12
            assert np.abs(got[1] - ref[1]).max() <= 1e-12 * max(1.0, np.abs(ref[1]).max())
            assert np.array_equal(got[2], ref[2]) and 0 < int(ref[2].sum()) < n
        # outputs are optional, as in ba_triangulate
        P1, P2, R, t, p1, p2 = _two_view(100, 3)
        hom = np.zeros((100, 4))
        null_d, null_b = C.POINTER(C.c_double)(), C.POINTER(C.c_uint8)()
        a = [api.f64(x).ravel() for x in (P1, P2, p1, p2)]
        api.check(L.ba_triangulate_static(solver._h, api.dptr(a[0]), api.dptr(a[1]), null_d, null_d, 100, api.dptr(a[2]),
                                          api.dptr(a[3]), api.dptr(hom), null_d, null_b))
        assert np.allclose(np.linalg.norm(hom, axis=1), 1.0, atol=1e-12)
        # at the limit it works, beyond it the call is refused with a message that names the limit
        M = api.TRIANGULATE_STATIC_MAX
        args = _two_view(M, 4)
        ref, got = _call(L.ba_triangulate, solver, *args), _call(L.ba_triangulate_static, solver, *args)
        assert np.abs(got[1] - ref[1]).max() <= 1e-12 * max(1.0, np.abs(ref[1]).max()) and np.array_equal(got[2], ref[2])
        with pytest.raises(api.BaError, match="BA_TRIANGULATE_STATIC_MAX"):
            _call(L.ba_triangulate_static, solver, *_two_view(M + 1, 5))
        # and the refused call left the entry point usable
        args = _two_view(64, 6)
        assert np.array_equal(_call(L.ba_triangulate_static, solver, *args)[2], _call(L.ba_triangulate, solver, *args)[2])
