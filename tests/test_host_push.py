"""Delivery of the CSC values into a pinned caller buffer without the structural zeros crossing PCIe (mocob200.h:
mocob200_host_push_stats): the caller's buffer must be bit-identical to the plain copy at every iterate, including the
iterates at which an entry outside the kept set turns non-zero (the call then falls back and widens the set)."""
import numpy as np
import pytest

from conftest import golden


def _pinned(n):
    import torch

    t = torch.empty(n, dtype=torch.float64).pin_memory()
    return t, t.numpy()


@pytest.mark.gpu
@pytest.mark.parametrize("name,opts", [("gait2d_tracking", "num_mesh_intervals=20"), ("gait2d_tracking", None)])
def test_pinned_delivery_is_bit_identical_to_the_plain_copy(pkg, name, opts):
    p = pkg.Problem.load(golden(name + ".mb20"), opts)
    d = p.dims()
    plain, pushed = pkg.Evaluator(p), pkg.Evaluator(p)
    pushed.set_host_delivery(True)
    keep, vals = _pinned(d.nnz_jac)
    xs = [p.make_iterate(seed=s, spread=sp) for s, sp in ((11, 0.1), (12, 0.5), (13, 1.0), (14, 0.02), (15, 0.8), (11, 0.1))]
    xs.append(xs[-1].copy())  # same x twice in a row: the memoised path re-assembles only
    for x in xs:
        ref = np.array(plain.eval_jac_g(x))  # pageable buffer: plain device-to-host copy
        vals[:] = np.nan  # whatever the buffer held before must be gone
        x = np.ascontiguousarray(x)
        pushed.eval_jac_g_ptr(x.ctypes.data, vals.ctypes.data)
        assert np.array_equal(vals.view(np.int64), ref.view(np.int64))  # bitwise, signs of zeros and NaN payloads included
    n_pushed, n_fallback, kept = pushed.host_push_stats()
    assert n_pushed + n_fallback == len(xs)
    assert n_pushed >= 1 and 0 < kept < 0.6 * d.nnz_jac  # the masked path was the one that ran
    assert plain.host_push_stats()[0] == 0  # pageable buffers never take it
    other = pkg.Evaluator(p)  # nor does a pinned buffer unless the caller asked for it
    other.eval_jac_g_ptr(np.ascontiguousarray(xs[0]).ctypes.data, vals.ctypes.data)
    assert other.host_push_stats() == (0, 0, 0)
    del keep


@pytest.mark.gpu
def test_small_and_constrained_problems_take_the_plain_copy(pkg):
    for name in ("sliding_mass", "double_pendulum_coupler", "pendulum_cb_both"):
        p = pkg.Problem.load(golden(name + ".mb20"))
        d = p.dims()
        ev = pkg.Evaluator(p)
        ev.set_host_delivery(True)
        keep, vals = _pinned(d.nnz_jac)
        x = np.ascontiguousarray(p.make_iterate(seed=2, spread=0.3))
        ref = np.array(pkg.Evaluator(p).eval_jac_g(x))
        ev.eval_jac_g_ptr(x.ctypes.data, vals.ctypes.data)
        assert np.array_equal(vals.view(np.int64), ref.view(np.int64))
        assert ev.host_push_stats() == (0, 0, 0)
        del keep
