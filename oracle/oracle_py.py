"""TEST INFRASTRUCTURE ONLY — ctypes binding of the CPU oracle (oracle/liborc.so).

Imported by tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs, never by
the product package. The oracle consumes the ABI's own POD tables (include/mocob200.h), obtained through
the product's HOST-side loader (no device code involved).
"""
import ctypes as C
import importlib.util
import os
import subprocess
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
LIB = os.path.join(HERE, "liborc.so")


def _pkg():
    name = "opensim_moco_b200"
    if name in sys.modules:
        return sys.modules[name]
    spec = importlib.util.spec_from_file_location(name, os.path.join(ROOT, "opensim-moco_b200", "__init__.py"))
    mod = importlib.util.module_from_spec(spec)
    sys.modules[name] = mod
    spec.loader.exec_module(mod)
    return mod


_lib = None


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(LIB):
            subprocess.check_call(["make", "-s", "-C", HERE])
        L = C.CDLL(LIB)
        vp, dp, ip = C.c_void_p, C.POINTER(C.c_double), C.POINTER(C.c_int)
        L.orc_create.argtypes = [vp, vp]
        L.orc_create.restype = vp
        L.orc_destroy.argtypes = [vp]
        L.orc_set_threads.argtypes = [C.c_int]
        L.orc_max_threads.restype = C.c_int
        L.orc_dims.argtypes = [vp, ip]
        L.orc_set_fd.argtypes = [vp, C.c_int, C.c_double, C.c_int, C.c_double]
        L.orc_eval_point.argtypes = [vp, dp, dp, C.c_uint]
        L.orc_eval_points.argtypes = [vp, dp, dp]
        L.orc_eval_point_jacobians.argtypes = [vp, dp, dp, dp]
        L.orc_eval_g.argtypes = [vp, dp, dp]
        L.orc_eval_f.argtypes = [vp, dp, dp]
        L.orc_eval_f.restype = C.c_double
        L.orc_eval_grad_f.argtypes = [vp, dp, dp]
        L.orc_nnz.argtypes = [vp]
        L.orc_nnz.restype = C.c_int
        L.orc_structure.argtypes = [vp, ip, ip]
        L.orc_eval_jac_g.argtypes = [vp, dp, dp]
        L.orc_eval_endpoint.argtypes = [vp, C.c_int, dp, dp]
        L.orc_count_ops.argtypes = [vp, dp, C.c_uint, C.POINTER(C.c_longlong)]
        L.orc_dgf_curve.argtypes = [C.c_int, C.c_double, C.c_double, C.c_double]
        L.orc_dgf_curve.restype = C.c_double
        L.orc_activation_derivative.argtypes = [C.c_double] * 4
        L.orc_activation_derivative.restype = C.c_double
        L.orc_muscle.argtypes = [dp, C.c_int, C.c_double, C.c_double, C.c_double, C.c_double, C.c_double, dp]
        L.orc_mass_matrix.argtypes = [vp, dp, dp]
        L.orc_inverse_dynamics.argtypes = [vp, dp, dp, dp, C.c_int, dp]
        L.orc_energy.argtypes = [vp, dp, dp, dp, dp]
        L.orc_path.argtypes = [vp, C.c_int, dp, dp, dp]
        L.orc_path_generalized_force.argtypes = [vp, C.c_int, dp, C.c_double, dp]
        L.orc_contact.argtypes = [vp, C.c_int, dp, dp, dp, dp]
        _lib = L
    return _lib


def _d(a):
    return a.ctypes.data_as(C.POINTER(C.c_double))


def _a(x):
    return np.ascontiguousarray(x, np.float64)


MUSCLE_FIELDS = ("tendonForce equilibriumResidual normTendonForceRate linearizedResidualRate normFiberLength cosPennation "
                 "normFiberVelocity fiberForce activeFiberForce passiveFiberForce fiberStiffness tendonStiffness "
                 "fiberStiffnessAlongTendon muscleStiffness fiberVelocity tendonVelocity fal fpe fv pennationAngle "
                 "fiberLength tendonLength").split()


def dgf_curve(which, x, p0=1.0, p1=0.0):
    names = {"active": 0, "passive": 1, "fv": 2, "fv_inv": 3, "tendon": 4, "tendon_inv": 5, "active_d": 6, "passive_d": 7,
             "tendon_d": 8}
    return lib().orc_dgf_curve(names[which], float(x), float(p0), float(p1))


def activation_derivative(e, a, tau_a=0.015, tau_d=0.060):
    return lib().orc_activation_derivative(e, a, tau_a, tau_d)


def muscle(params14, flags, mt_length, mt_velocity, activation, ntf=float("nan"), ntf_rate=float("nan")):
    p = _a(params14)
    out = np.empty(22)
    lib().orc_muscle(_d(p), flags, mt_length, mt_velocity, activation, ntf, ntf_rate, _d(out))
    return dict(zip(MUSCLE_FIELDS, out))


class Oracle:
    """CPU oracle of one problem: same tables, same transcription, independent arithmetic."""

    def __init__(self, problem_or_path, options=None):
        m = _pkg()
        self.problem = problem_or_path if isinstance(problem_or_path, m.Problem) else m.Problem.load(problem_or_path, options)
        self._h = C.c_void_p(lib().orc_create(self.problem.tables, self.problem.transcription))
        if not self._h:
            raise RuntimeError("orc_create failed (model exceeds the oracle's limits?)")
        d = (C.c_int * 17)()
        lib().orc_dims(self._h, d)
        (self.n_grid, self.n_mesh, self.n_var, self.n_con, self.ns, self.nc, self.nd, self.n_in, self.n_out, self.n_dyn_out,
         self.n_integrand, self.n_cost_terms, self.n_ein, self.nq, self.nz, self.nmus, self.ncontact) = list(d)

    def set_fd(self, scheme, h=-1.0, h_iter=-1, u_aim=-1.0):
        lib().orc_set_fd(self._h, scheme, h, h_iter, u_aim)

    def eval_point(self, inputs, mask=0xFFFFFFFF):
        i = _a(inputs)
        out = np.full(self.n_out, np.nan)
        lib().orc_eval_point(self._h, _d(i), _d(out), mask)
        return out

    def eval_points(self, x):
        x = _a(x)
        Y = np.empty((self.n_grid, self.n_out))
        lib().orc_eval_points(self._h, _d(x), _d(Y))
        return Y

    def eval_point_jacobians(self, x):
        x = _a(x)
        Y = np.empty((self.n_grid, self.n_out))
        J = np.empty((self.n_grid, self.n_out, self.n_in))
        lib().orc_eval_point_jacobians(self._h, _d(x), _d(Y), _d(J))
        return Y, J

    def eval_g(self, x):
        x = _a(x)
        g = np.empty(self.n_con)
        lib().orc_eval_g(self._h, _d(x), _d(g))
        return g

    def eval_f(self, x, terms=False):
        x = _a(x)
        t = np.empty(max(self.n_cost_terms, 1))
        f = lib().orc_eval_f(self._h, _d(x), _d(t))
        return (f, t[:self.n_cost_terms]) if terms else f

    def eval_grad_f(self, x):
        x = _a(x)
        g = np.empty(self.n_var)
        lib().orc_eval_grad_f(self._h, _d(x), _d(g))
        return g

    def structure(self):
        nnz = lib().orc_nnz(self._h)
        colptr, rowidx = np.empty(self.n_var + 1, np.int32), np.empty(nnz, np.int32)
        lib().orc_structure(self._h, colptr.ctypes.data_as(C.POINTER(C.c_int)), rowidx.ctypes.data_as(C.POINTER(C.c_int)))
        return colptr, rowidx

    def eval_jac_g(self, x):
        x = _a(x)
        v = np.zeros(lib().orc_nnz(self._h))
        lib().orc_eval_jac_g(self._h, _d(x), _d(v))
        return v

    def eval_endpoint(self, goal, ein, nout):
        e = _a(ein)
        v = np.zeros(max(nout, 1))
        lib().orc_eval_endpoint(self._h, goal, _d(e), _d(v))
        return v[:nout]

    def count_ops(self, inputs, mask=0xFFFFFFFF):
        i = _a(inputs)
        c = (C.c_longlong * 2)()
        lib().orc_count_ops(self._h, _d(i), mask, c)
        return int(c[0]), int(c[1])

    def mass_matrix(self, q):
        q = _a(q)
        M = np.empty((q.size, q.size))
        lib().orc_mass_matrix(self._h, _d(q), _d(M))
        return M

    def inverse_dynamics(self, q, u, udot, gravity=True):
        q, u, udot = _a(q), _a(u), _a(udot)
        tau = np.empty(q.size)
        lib().orc_inverse_dynamics(self._h, _d(q), _d(u), _d(udot), 1 if gravity else 0, _d(tau))
        return tau

    def energy(self, q, u):
        q, u = _a(q), _a(u)
        e, c = np.empty(2), np.empty(3)
        lib().orc_energy(self._h, _d(q), _d(u), _d(e), _d(c))
        return e[0], e[1], c

    def path(self, m, q, u):
        q, u = _a(q), _a(u)
        o = np.empty(2)
        lib().orc_path(self._h, m, _d(q), _d(u), _d(o))
        return o[0], o[1]

    def path_generalized_force(self, m, q, tension=1.0):
        q = _a(q)
        tau = np.empty(q.size)
        lib().orc_path_generalized_force(self._h, m, _d(q), tension, _d(tau))
        return tau

    def contact(self, c, q, u):
        q, u = _a(q), _a(u)
        f, p = np.empty(3), np.empty(3)
        lib().orc_contact(self._h, c, _d(q), _d(u), _d(f), _d(p))
        return f, p

    def gather_inputs(self, x, j):
        """inputs of grid point j from the NLP vector (same layout as the ABI)."""
        x = np.asarray(x)
        grid = self.grid()
        off_s, off_c = 2, 2 + self.ns * self.n_grid
        off_d = off_c + self.nc * self.n_grid
        return np.concatenate([[x[0] + (x[1] - x[0]) * grid[j]], x[off_s + j * self.ns: off_s + (j + 1) * self.ns],
                               x[off_c + j * self.nc: off_c + (j + 1) * self.nc], x[off_d + j * self.nd: off_d + (j + 1) * self.nd]])

    def grid(self):
        n_int = self.n_mesh - 1
        mesh = np.arange(self.n_mesh) / n_int
        if self.n_grid == self.n_mesh:
            return mesh
        g = np.empty(self.n_grid)
        g[0::2] = mesh
        g[1::2] = 0.5 * (mesh[:-1] + mesh[1:])
        return g

    def close(self):
        if self._h:
            lib().orc_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


def set_threads(n):
    lib().orc_set_threads(n)


def max_threads():
    return lib().orc_max_threads()
