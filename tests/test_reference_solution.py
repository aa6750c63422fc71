"""A reference-produced NLP SOLUTION as a pin of the whole MocoInverse path.

Moco/Tests/std_testMocoInverse_subject_18musc_solution.sto is what the reference's own regression test of MocoInverse
(testMocoInverse.cpp:118-148: Rajagopal-type 3-D model, 18 DeGrooteFregly2016 muscles of which 10 with compliant tendons in
implicit mode, external loads, coordinates low-pass filtered at 6 Hz and prescribed, Hermite-Simpson, 11 intervals) compares
its result with: IPOPT converged there (tolerances 1e-3) with objective 1.087741. The problem is rebuilt from the reference's
input files by this repo's host code (examples.cpp: inverse18Muscles; fixtures by scripts/make_inverse_pins.py) and evaluated
AT that solution: the objective must come out the same and every constraint must hold to the solver's tolerance. That pins,
end to end and against the real OpenSim / Simbody / CasADi / IPOPT stack: the flattener (Version 30000 model, CustomJoints,
welded joints, DGF replacement of Millard muscles), the table processing (degrees, padding, Butterworth filter, GCV splines),
prescribed kinematics, external loads, the DGF muscle inside a multibody model (activation dynamics, implicit tendon
equilibrium), the multibody residuals, the Hermite-Simpson defects, the quadrature and the auxiliary-derivative penalty."""
import ctypes as C

import numpy as np
import pytest

from conftest import golden


@pytest.fixture(scope="module")
def sol():
    return np.load(golden("moco_inverse_18musc_solution.npz"))


def blocks(p, g):
    """g split into (multibody residuals [grid][nq], auxiliary residuals [grid][naux], defects [interval][2][ns])"""
    d, t = p.dims(), p.table_sizes()
    nq, na, ns = t.nq, t.nderiv_aux, d.ns
    row, res, dfc = 18, [], []  # 18 initial-activation rows first
    for _ in range(d.n_mesh - 1):
        res.append(g[row:row + 2 * (nq + na)]); row += 2 * (nq + na)
        dfc.append(g[row:row + 2 * ns]); row += 2 * ns
    res.append(g[row:row + nq + na]); row += nq + na
    assert row == len(g)
    R = np.concatenate(res).reshape(d.n_grid, nq + na)
    return R[:, :nq], R[:, nq:], np.array(dfc).reshape(d.n_mesh - 1, 2, ns)


def check(p, f, terms, g, sol):
    # the objective: 7 printed digits of the reference's value
    assert abs(f - float(sol["objective"])) < 1e-6, (f, float(sol["objective"]))
    assert len(terms) == 2 and abs(terms[0] + terms[1] - f) < 1e-12  # effort + 0.01 x implicit tendon derivative penalty
    xl, xh, gl, gh = p.bounds()
    x = sol["x"]
    assert np.all(x >= xl - 1e-9) and np.all(x <= xh + 1e-9)
    resid, aux, defects = blocks(p, g)
    assert np.abs(g[:18]).max() < 1e-6            # initial activation = initial excitation
    # IPOPT stopped at tol = 1e-3 on its scaled problem; the loads in these rows are ~1e3 N / N m
    assert np.abs(defects).max() < 5e-4           # activation and tendon-force dynamics, Hermite and Simpson rows
    assert np.abs(aux).max() < 3e-2               # tendon equilibrium residuals (normalised forces, O(1) terms)
    assert np.abs(resid).max() < 2.0              # multibody residuals: < 0.2 % of the ~900 N ground reactions
    assert np.sqrt((resid ** 2).mean()) < 0.2


def test_prescribed_kinematics_match_what_the_reference_wrote(pkg, sol):
    """The solution file carries the coordinate values and speeds the reference prescribed (its PositionMotion, exported):
    ours agree to the file's 6-7 digits - degrees conversion, padding, Butterworth filter and spline are the reference's."""
    p = pkg.Problem.load(golden("moco_inverse_18musc.mb20"))
    t = p.table_sizes()
    dbl = np.ctypeslib.as_array(C.cast(t.dbl, C.POINTER(C.c_double)), shape=(t.n_dbl,))
    idx = np.ctypeslib.as_array(C.cast(t.idx, C.POINTER(C.c_int)), shape=(t.n_int,))

    def ppoly(pid, tt, deriv):
        off, npc, order = (int(v) for v in idx[t.o_ppoly_i + 3 * pid:t.o_ppoly_i + 3 * pid + 3])
        br = dbl[t.o_ppoly_data + off:t.o_ppoly_data + off + npc + 1]
        co = dbl[t.o_ppoly_data + off + npc + 1:t.o_ppoly_data + off + npc + 1 + npc * order].reshape(npc, order)
        k = int(np.clip(np.searchsorted(br, tt, side="right") - 1, 0, npc - 1))
        pol = np.polynomial.Polynomial(co[k])
        return (pol.deriv(deriv) if deriv else pol)(tt - br[k])

    Q = np.array([[ppoly(int(idx[t.o_presc_i + k]), tt, 0) for k in range(t.nq)] for tt in sol["time"]])
    U = np.array([[ppoly(int(idx[t.o_presc_i + k]), tt, 1) for k in range(t.nq)] for tt in sol["time"]])
    assert sol["q"].shape[1] == t.nq == 21
    for c in range(t.nq):  # the file's columns are in the model's coordinate order too, up to the tree ordering: match by value
        e = np.abs(Q - sol["q"][:, c:c + 1]).max(axis=0)
        k = int(e.argmin())
        assert e[k] < 1e-6, (sol["coordinates"][c], e[k])
        assert np.abs(U[:, k] - sol["u"][:, c]).max() < 1e-4, sol["coordinates"][c]


def test_oracle_accepts_the_reference_solution(pkg, orc, sol):
    p = pkg.Problem.load(golden("moco_inverse_18musc.mb20"))
    o = orc.Oracle(p)
    f, terms = o.eval_f(sol["x"], terms=True)
    check(p, f, terms, o.eval_g(sol["x"]), sol)


@pytest.mark.gpu
def test_gpu_accepts_the_reference_solution(pkg, orc, sol):
    p = pkg.Problem.load(golden("moco_inverse_18musc.mb20"))
    ev = pkg.Evaluator(p)
    f, terms = ev.eval_f(sol["x"], terms=True)
    g = ev.eval_g(sol["x"])
    check(p, f, terms, g, sol)
    go = orc.Oracle(p).eval_g(sol["x"])
    assert np.max(np.abs(g - go)) < 1e-10 * max(1.0, np.abs(go).max())
