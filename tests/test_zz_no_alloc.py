"""mocob200_eval_g_static: the allocation-free variant of eval_g (scratch in a static __device__ array of the library;
include/mocob200.h). No counterpart in the reference, so the check is against
the existing entry points: same g as mocob200_eval_g / mocob200_eval_g_device within the value tolerance of the parity tests
(1e-10 relative), on two calls in a row (the array keeps the previous call's contents) and across problems that share it.
Collected last on purpose (pytest -x at the end of a round must not lose the other tests to a failure here)."""
import numpy as np
import pytest

from conftest import golden

RTOL = 1e-10  # same as tests/test_gpu_parity.py: values


def _close(a, b):
    scale = max(1.0, float(np.max(np.abs(b))))
    assert np.all(np.isfinite(a))
    assert float(np.max(np.abs(a - b))) <= RTOL * scale, float(np.max(np.abs(a - b))) / scale


def _static_g(ev, torch, x, stream=None):
    d_x = torch.from_numpy(x).cuda()
    d_g = torch.full((ev.dims.n_con,), float("nan"), dtype=torch.float64, device="cuda")
    torch.cuda.synchronize()
    ev.eval_g_static(d_x.data_ptr(), d_g.data_ptr(), stream.cuda_stream if stream is not None else None)
    if stream is not None:
        stream.synchronize()
    else:
        ev.sync()
    return d_g.cpu().numpy()


@pytest.mark.gpu
@pytest.mark.parametrize("name", ["gait2d_tracking", "pendulum3_implicit", "hanging_muscle"])
def test_static_scratch_eval_g_matches_eval_g_twice_in_a_row(pkg, name):
    import torch

    p = pkg.Problem.load(golden(name + ".mb20"))
    ev = pkg.Evaluator(p)
    xs = [p.make_iterate(seed=s, spread=0.1) for s in (1, 2)]
    want = [ev.eval_g(x).copy() for x in xs]
    mem0 = torch.cuda.memory_allocated()
    # two calls in a row at different iterates, then the first again: nothing of an earlier call may leak into a later one
    for k in (0, 1, 0):
        _close(_static_g(ev, torch, xs[k]), want[k])
    # on a caller's stream (everything is queued there, in order)
    st = torch.cuda.Stream()
    _close(_static_g(ev, torch, xs[1], st), want[1])
    # the device-pointer entry point agrees too
    d_x = torch.from_numpy(xs[0]).cuda()
    d_g = torch.empty(ev.dims.n_con, dtype=torch.float64, device="cuda")
    torch.cuda.synchronize()
    ev.eval_g_device(d_x.data_ptr(), d_g.data_ptr())
    ev.sync()
    _close(_static_g(ev, torch, xs[0]), d_g.cpu().numpy())
    assert torch.cuda.memory_allocated() >= mem0  # (torch's own tensors only; the library call allocates nothing)


@pytest.mark.gpu
def test_static_scratch_is_shared_between_handles_without_leaking(pkg):
    import torch

    pa = pkg.Problem.load(golden("gait2d_tracking.mb20"))
    pb = pkg.Problem.load(golden("pendulum3_implicit.mb20"))
    ea, eb = pkg.Evaluator(pa), pkg.Evaluator(pb)
    xa, xb = pa.make_iterate(seed=3, spread=0.1), pb.make_iterate(seed=3, spread=0.1)
    wa, wb = ea.eval_g(xa).copy(), eb.eval_g(xb).copy()
    for _ in range(2):
        _close(_static_g(ea, torch, xa), wa)
        _close(_static_g(eb, torch, xb), wb)


@pytest.mark.gpu
def test_static_scratch_limit_is_enforced_with_a_clear_error(pkg):
    import torch

    # 2200 Hermite-Simpson intervals = 4401 grid points x 31 point outputs x 8 bytes > the 1 MiB static array
    p = pkg.Problem.load(golden("gait2d_tracking.mb20"), "num_mesh_intervals=2200")
    ev = pkg.Evaluator(p)
    d_x = torch.from_numpy(p.make_iterate(seed=1, spread=0.1)).cuda()
    d_g = torch.zeros(ev.dims.n_con, dtype=torch.float64, device="cuda")
    with pytest.raises(RuntimeError, match="static limit"):
        ev.eval_g_static(d_x.data_ptr(), d_g.data_ptr())
    assert float(d_g.abs().max()) == 0.0  # nothing was launched
