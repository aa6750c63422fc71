"""ctypes harness over libmocob200.so (the C ABI in include/mocob200.h).

The product is the shared library: host C++ that flattens a fixed-topology model + MocoProblem
into device tables, and FP64 sm_100a kernels behind `mocob200_*`.  This module only binds those
entry points for tests / bench.py; it holds no arithmetic and has no CPU path: creating an
`Evaluator` without the library or without a CUDA device raises.
"""
import ctypes as C
import os

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(HERE, "libmocob200.so")

FD_FORWARD, FD_CENTRAL, FD_BACKWARD = 0, 1, 2
POINTS_OUTPUTS, POINTS_JACOBIAN, POINTS_PATHS = 0, 1, 2

# every symbol include/mocob200.h declares (checked by tests/test_abi.py)
ABI_SYMBOLS = [
    "mocob200_problem_create_example", "mocob200_problem_load", "mocob200_problem_load_opts", "mocob200_problem_save",
    "mocob200_problem_names", "mocob200_problem_tables", "mocob200_problem_transcription", "mocob200_problem_destroy",
    "mocob200_layout", "mocob200_layout_bounds", "mocob200_layout_structure", "mocob200_make_iterate",
    "mocob200_create", "mocob200_nccl_unique_id", "mocob200_destroy", "mocob200_last_error", "mocob200_get_dims",
    "mocob200_set_fd", "mocob200_jac_g_structure", "mocob200_eval_points", "mocob200_eval_f", "mocob200_eval_g",
    "mocob200_eval_grad_f", "mocob200_eval_jac_g", "mocob200_set_host_delivery", "mocob200_host_push_stats", "mocob200_eval_jac_g_device", "mocob200_eval_g_device",
    "mocob200_eval_g_static",
    "mocob200_eval_jac_g_batch",
    "mocob200_sync", "mocob200_mark", "mocob200_elapsed_ms", "mocob200_launch_count", "mocob200_last_kernel_ms",
    "mocob200_point_evals_per_jacobian", "mocob200_executed_evals", "mocob200_working_ranks", "mocob200_peer_gather", "mocob200_uses_warp_kernels",
    "mocob200_measure_fp64_peak", "mocob200_shard_range", "mocob200_effective_world", "mocob200_set_min_points_per_rank", "mocob200_set_sparsity_detection",
    "mocob200_iterate_write_sto", "mocob200_iterate_read_sto",
]


class Dims(C.Structure):
    _fields_ = [(n, C.c_int) for n in (
        "n_grid n_mesh n_var n_con nnz_jac ns nc nd n_in n_out n_dyn_out n_integrand n_cost_terms "
        "first_interval n_intervals_local rank world nm nsl n_kc").split()]

    def asdict(self):
        return {n: getattr(self, n) for n, _ in self._fields_}


_lib = None


def lib():
    """Load libmocob200.so; fails loudly when it has not been built (python opensim-moco_b200/build.py)."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise RuntimeError("libmocob200.so is missing: run __graft_entry__.build() (there is no CPU fallback)")
        L = C.CDLL(LIB_PATH)
        vp, ip, dp, cp = C.c_void_p, C.POINTER(C.c_int), C.POINTER(C.c_double), C.c_char_p
        L.mocob200_problem_create_example.argtypes = [cp, cp, cp, C.POINTER(vp), cp, C.c_int]
        L.mocob200_problem_load.argtypes = [cp, C.POINTER(vp), cp, C.c_int]
        L.mocob200_problem_load_opts.argtypes = [cp, cp, C.POINTER(vp), cp, C.c_int]
        L.mocob200_problem_save.argtypes = [vp, cp]
        L.mocob200_problem_names.argtypes = [vp, C.c_int, cp, C.c_int]
        L.mocob200_problem_tables.argtypes = [vp]
        L.mocob200_problem_tables.restype = vp
        L.mocob200_problem_transcription.argtypes = [vp]
        L.mocob200_problem_transcription.restype = vp
        L.mocob200_problem_destroy.argtypes = [vp]
        L.mocob200_problem_destroy.restype = None
        L.mocob200_layout.argtypes = [vp, vp, C.POINTER(Dims)]
        L.mocob200_layout_bounds.argtypes = [vp, vp, dp, dp, dp, dp]
        L.mocob200_layout_structure.argtypes = [vp, vp, ip, ip]
        L.mocob200_make_iterate.argtypes = [vp, vp, C.c_long, C.c_double, dp]
        L.mocob200_effective_world.argtypes = [C.c_int, C.c_int]
        L.mocob200_set_min_points_per_rank.argtypes = [C.c_int]
        L.mocob200_set_min_points_per_rank.restype = None
        L.mocob200_set_sparsity_detection.argtypes = [vp, C.c_int, C.c_int]
        L.mocob200_shard_range.argtypes = [C.c_int, C.c_int, C.c_int, C.POINTER(C.c_int), C.POINTER(C.c_int)]
        L.mocob200_create.argtypes = [vp, vp, C.c_int, C.c_int, C.c_int, vp, C.POINTER(vp)]
        L.mocob200_nccl_unique_id.argtypes = [vp]
        L.mocob200_destroy.argtypes = [vp]
        L.mocob200_destroy.restype = None
        L.mocob200_last_error.argtypes = [vp]
        L.mocob200_last_error.restype = cp
        L.mocob200_get_dims.argtypes = [vp, C.POINTER(Dims)]
        L.mocob200_set_fd.argtypes = [vp, C.c_int, C.c_double, C.c_int, C.c_double]
        L.mocob200_jac_g_structure.argtypes = [vp, ip, ip]
        L.mocob200_eval_points.argtypes = [vp, C.c_int, vp, vp]
        L.mocob200_eval_f.argtypes = [vp, vp, dp, vp]
        L.mocob200_eval_g.argtypes = [vp, vp, vp]
        L.mocob200_eval_grad_f.argtypes = [vp, vp, vp]
        L.mocob200_eval_jac_g.argtypes = [vp, vp, vp]
        L.mocob200_set_host_delivery.argtypes = [vp, C.c_int]
        L.mocob200_host_push_stats.argtypes = [vp, C.POINTER(C.c_long), C.POINTER(C.c_long), C.POINTER(C.c_long)]
        L.mocob200_eval_jac_g_device.argtypes = [vp, vp, vp, vp]
        L.mocob200_eval_g_device.argtypes = [vp, vp, vp, vp]
        L.mocob200_eval_g_static.argtypes = [vp, vp, vp, vp]
        L.mocob200_eval_jac_g_batch.argtypes = [C.POINTER(vp), C.c_int, C.POINTER(vp), C.POINTER(vp)]
        L.mocob200_sync.argtypes = [vp]
        L.mocob200_mark.argtypes = [vp, C.c_int]
        L.mocob200_elapsed_ms.argtypes = [vp, C.POINTER(C.c_float)]
        L.mocob200_launch_count.argtypes = [vp]
        L.mocob200_launch_count.restype = C.c_long
        L.mocob200_last_kernel_ms.argtypes = [vp, C.POINTER(C.c_float), C.POINTER(C.c_float)]
        L.mocob200_point_evals_per_jacobian.argtypes = [vp]
        L.mocob200_point_evals_per_jacobian.restype = C.c_long
        L.mocob200_measure_fp64_peak.argtypes = [C.c_int, dp]
        L.mocob200_executed_evals.argtypes = [vp, C.POINTER(C.c_long)]
        L.mocob200_working_ranks.argtypes = [vp]
        L.mocob200_peer_gather.argtypes = [vp]
        L.mocob200_uses_warp_kernels.argtypes = [vp]
        L.mocob200_iterate_write_sto.argtypes = [vp, dp, C.c_char_p]
        L.mocob200_iterate_read_sto.argtypes = [vp, C.c_char_p, dp, C.c_char_p, C.c_int]
        _lib = L
    return _lib


def _dptr(a):
    return a.ctypes.data_as(C.POINTER(C.c_double))


class Tables(C.Structure):
    """mocob200_tables (include/mocob200.h), for reading sizes from Python."""
    _fields_ = ([("n_dbl", C.c_int), ("n_int", C.c_int), ("dbl", C.c_void_p), ("idx", C.c_void_p)] +
                [(n, C.c_int) for n in ("nb nq nz nc nderiv_aux nmus npp nspline ncontact ncact ngoal nppoly o_body_d o_gravity o_pp_d "
                                        "o_mus_d o_contact_d o_cact_d o_spline_data o_ppoly_data o_bounds o_time_bounds o_body_i o_pp_i "
                                        "o_mus_i o_contact_i o_cact_i o_spline_i o_goal_i o_ppoly_i ncustom o_cj_d o_cj_i o_body_cj nxforce "
                                        "o_xforce_i ncoupler o_cc_i o_cc_d prescribed o_presc_i nwrapobj o_wo_d o_wo_i npathwrap o_pw_i").split()])


def _vptr(a):
    return C.c_void_p(a.ctypes.data) if a is not None else None


class Problem:
    """A flattened model + optimal control problem + transcription settings (host only; no GPU needed)."""

    def __init__(self, handle):
        self._h = handle
        self.tables = C.c_void_p(lib().mocob200_problem_tables(handle))
        self.transcription = C.c_void_p(lib().mocob200_problem_transcription(handle))

    @classmethod
    def example(cls, name, data_dir=".", options=None):
        p, err = C.c_void_p(), C.create_string_buffer(1024)
        rc = lib().mocob200_problem_create_example(name.encode(), data_dir.encode(), options.encode() if options else None,
                                                   C.byref(p), err, 1024)
        if rc:
            raise RuntimeError("mocob200_problem_create_example(%s): %s" % (name, err.value.decode()))
        return cls(p)

    @classmethod
    def load(cls, path, options=None):
        p, err = C.c_void_p(), C.create_string_buffer(1024)
        rc = lib().mocob200_problem_load_opts(path.encode(), options.encode() if options else None, C.byref(p), err, 1024)
        if rc:
            raise RuntimeError("mocob200_problem_load(%s): %s" % (path, err.value.decode()))
        return cls(p)

    def save(self, path):
        rc = lib().mocob200_problem_save(self._h, path.encode())
        if rc:
            raise RuntimeError("mocob200_problem_save failed (%d)" % rc)

    def table_sizes(self):
        return C.cast(self.tables, C.POINTER(Tables)).contents

    def names(self, which):
        buf = C.create_string_buffer(1 << 16)
        lib().mocob200_problem_names(self._h, which, buf, 1 << 16)
        return [s for s in buf.value.decode().split("\n") if s]

    def dims(self):
        d = Dims()
        rc = lib().mocob200_layout(self.tables, self.transcription, C.byref(d))
        if rc:
            raise RuntimeError("mocob200_layout failed (%d)" % rc)
        return d

    def bounds(self):
        d = self.dims()
        xl, xh = np.empty(d.n_var), np.empty(d.n_var)
        gl, gh = np.empty(d.n_con), np.empty(d.n_con)
        rc = lib().mocob200_layout_bounds(self.tables, self.transcription, _dptr(xl), _dptr(xh), _dptr(gl), _dptr(gh))
        if rc:
            raise RuntimeError("mocob200_layout_bounds failed (%d)" % rc)
        return xl, xh, gl, gh

    def structure(self):
        d = self.dims()
        colptr, rowidx = np.empty(d.n_var + 1, np.int32), np.empty(d.nnz_jac, np.int32)
        rc = lib().mocob200_layout_structure(self.tables, self.transcription, colptr.ctypes.data_as(C.POINTER(C.c_int)),
                                             rowidx.ctypes.data_as(C.POINTER(C.c_int)))
        if rc:
            raise RuntimeError("mocob200_layout_structure failed (%d)" % rc)
        return colptr, rowidx

    def make_iterate(self, seed=-1, spread=1.0):
        x = np.empty(self.dims().n_var)
        rc = lib().mocob200_make_iterate(self.tables, self.transcription, seed, spread, _dptr(x))
        if rc:
            raise RuntimeError("mocob200_make_iterate failed (%d)" % rc)
        return x

    def write_sto(self, x, path):
        """MocoTrajectory::write of the iterate x (one row per grid point)."""
        x = np.ascontiguousarray(x, dtype=np.float64)
        rc = lib().mocob200_iterate_write_sto(self._h, _dptr(x), str(path).encode())
        if rc:
            raise RuntimeError("mocob200_iterate_write_sto failed (%d)" % rc)

    def read_sto(self, path):
        """setGuessFile: a MocoTrajectory table resampled onto this problem's grid, as the NLP vector x."""
        x = np.zeros(self.dims().n_var)
        err = C.create_string_buffer(512)
        rc = lib().mocob200_iterate_read_sto(self._h, str(path).encode(), _dptr(x), err, 512)
        if rc:
            raise RuntimeError("mocob200_iterate_read_sto: %s" % err.value.decode())
        return x

    def close(self):
        if self._h:
            lib().mocob200_problem_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


class Evaluator:
    """Device evaluator of one problem (mocob200_create ... mocob200_destroy). Host numpy in, host numpy out."""

    def __init__(self, problem, device=0, rank=0, world=1, nccl_id=None):
        self.problem = problem
        h = C.c_void_p()
        idbuf = C.c_char_p(nccl_id) if nccl_id is not None else None
        rc = lib().mocob200_create(problem.tables, problem.transcription, device, rank, world, idbuf, C.byref(h))
        if rc:
            raise RuntimeError("mocob200_create: %s" % (lib().mocob200_last_error(None) or b"").decode())
        self._h = h
        d = Dims()
        lib().mocob200_get_dims(h, C.byref(d))
        self.dims = d

    def _ck(self, rc, what):
        if rc:
            raise RuntimeError("%s: %s" % (what, (lib().mocob200_last_error(self._h) or b"").decode()))

    def set_fd(self, scheme, h=-1.0, h_iter=-1, u_aim=-1.0):
        self._ck(lib().mocob200_set_fd(self._h, scheme, h, h_iter, u_aim), "set_fd")

    def set_sparsity_detection(self, mode, n_iterates=0):
        """mode "none" (dense callback blocks) or "random" (optim_sparsity_detection); changes nnz_jac and the structure."""
        code = {"none": 0, "random": 1}[mode] if isinstance(mode, str) else int(mode)
        self._ck(lib().mocob200_set_sparsity_detection(self._h, code, n_iterates), "set_sparsity_detection")
        d = Dims()
        lib().mocob200_get_dims(self._h, C.byref(d))
        self.dims = d

    def structure(self):
        colptr, rowidx = np.empty(self.dims.n_var + 1, np.int32), np.empty(self.dims.nnz_jac, np.int32)
        self._ck(lib().mocob200_jac_g_structure(self._h, colptr.ctypes.data_as(C.POINTER(C.c_int)),
                                                rowidx.ctypes.data_as(C.POINTER(C.c_int))), "jac_g_structure")
        return colptr, rowidx

    def eval_points(self, x, jacobian=False):
        d = self.dims
        x = np.ascontiguousarray(x, np.float64)
        out = np.empty((d.n_grid, d.n_out, d.n_in) if jacobian else (d.n_grid, d.n_out))
        self._ck(lib().mocob200_eval_points(self._h, POINTS_JACOBIAN if jacobian else POINTS_OUTPUTS, _vptr(x), _vptr(out)),
                 "eval_points")
        return out

    def eval_paths(self, x):
        """[n_grid][nmus][2]: muscle-tendon path length and lengthening speed at every grid point."""
        x = np.ascontiguousarray(x, np.float64)
        nmus = self.problem.table_sizes().nmus
        out = np.empty((self.dims.n_grid, nmus, 2))
        self._ck(lib().mocob200_eval_points(self._h, POINTS_PATHS, _vptr(x), _vptr(out)), "eval_points(paths)")
        return out

    def eval_f(self, x, terms=False):
        x = np.ascontiguousarray(x, np.float64)
        f = C.c_double()
        t = np.empty(max(self.dims.n_cost_terms, 1))
        self._ck(lib().mocob200_eval_f(self._h, _vptr(x), C.byref(f), _vptr(t)), "eval_f")
        return (f.value, t[:self.dims.n_cost_terms]) if terms else f.value

    def eval_g(self, x, out=None):
        x = np.ascontiguousarray(x, np.float64)
        g = out if out is not None else np.empty(self.dims.n_con)
        self._ck(lib().mocob200_eval_g(self._h, _vptr(x), _vptr(g)), "eval_g")
        return g

    def eval_grad_f(self, x, out=None):
        x = np.ascontiguousarray(x, np.float64)
        g = out if out is not None else np.empty(self.dims.n_var)
        self._ck(lib().mocob200_eval_grad_f(self._h, _vptr(x), _vptr(g)), "eval_grad_f")
        return g

    def eval_jac_g(self, x, out=None):
        x = np.ascontiguousarray(x, np.float64)
        v = out if out is not None else np.empty(self.dims.nnz_jac)
        self._ck(lib().mocob200_eval_jac_g(self._h, _vptr(x), _vptr(v)), "eval_jac_g")
        return v

    # raw-pointer variants (pinned host buffers / device pointers from torch)
    def set_host_delivery(self, masked):
        self._ck(lib().mocob200_set_host_delivery(self._h, 1 if masked else 0), "set_host_delivery")

    def host_push_stats(self):
        """(calls delivered by the masked push, calls that fell back to the plain copy, entries crossing PCIe)"""
        a, b, c = C.c_long(), C.c_long(), C.c_long()
        lib().mocob200_host_push_stats(self._h, C.byref(a), C.byref(b), C.byref(c))
        return a.value, b.value, c.value

    def eval_jac_g_ptr(self, x_ptr, vals_ptr):
        self._ck(lib().mocob200_eval_jac_g(self._h, C.c_void_p(x_ptr), C.c_void_p(vals_ptr)), "eval_jac_g")

    def eval_g_ptr(self, x_ptr, g_ptr):
        self._ck(lib().mocob200_eval_g(self._h, C.c_void_p(x_ptr), C.c_void_p(g_ptr)), "eval_g")

    def eval_jac_g_device(self, d_x_ptr, d_vals_ptr=None):
        self._ck(lib().mocob200_eval_jac_g_device(self._h, C.c_void_p(d_x_ptr), C.c_void_p(d_vals_ptr) if d_vals_ptr else None, None),
                 "eval_jac_g_device")

    def eval_g_device(self, d_x_ptr, d_g_ptr=None):
        self._ck(lib().mocob200_eval_g_device(self._h, C.c_void_p(d_x_ptr), C.c_void_p(d_g_ptr) if d_g_ptr else None, None),
                 "eval_g_device")

    def eval_g_static(self, d_x_ptr, d_g_ptr, stream_ptr=None):
        """eval_g on device pointers with the library's static scratch array: allocates nothing (include/mocob200.h)."""
        self._ck(lib().mocob200_eval_g_static(self._h, C.c_void_p(d_x_ptr), C.c_void_p(d_g_ptr),
                                              C.c_void_p(stream_ptr) if stream_ptr else None), "eval_g_static")

    def sync(self):
        self._ck(lib().mocob200_sync(self._h), "sync")

    def mark(self, which):
        self._ck(lib().mocob200_mark(self._h, which), "mark")

    def elapsed_ms(self):
        ms = C.c_float()
        self._ck(lib().mocob200_elapsed_ms(self._h, C.byref(ms)), "elapsed_ms")
        return ms.value

    def launch_count(self):
        return lib().mocob200_launch_count(self._h)

    def last_kernel_ms(self):
        a, b = C.c_float(), C.c_float()
        lib().mocob200_last_kernel_ms(self._h, C.byref(a), C.byref(b))
        return a.value, b.value

    def point_evals_per_jacobian(self):
        return lib().mocob200_point_evals_per_jacobian(self._h)

    def executed_evals(self):
        out = (C.c_long * 4)()
        self._ck(lib().mocob200_executed_evals(self._h, out), "executed_evals")
        return list(out)

    def working_ranks(self):
        return lib().mocob200_working_ranks(self._h)

    def peer_gather(self):
        return bool(lib().mocob200_peer_gather(self._h))

    def uses_warp_kernels(self):
        return bool(lib().mocob200_uses_warp_kernels(self._h))

    def close(self):
        if self._h:
            lib().mocob200_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


def measure_fp64_peak(device=0):
    t = C.c_double()
    rc = lib().mocob200_measure_fp64_peak(device, C.byref(t))
    if rc:
        raise RuntimeError("mocob200_measure_fp64_peak failed (%d)" % rc)
    return t.value


def eval_jac_g_batch(evaluators, d_x_ptrs, d_vals_ptrs=None):
    """mocob200_eval_jac_g_batch: independent problems side by side; only enqueues (sync / mark on evaluators[0])."""
    n = len(evaluators)
    hs = (C.c_void_p * n)(*[e._h for e in evaluators])
    xs = (C.c_void_p * n)(*[C.c_void_p(p) for p in d_x_ptrs])
    vs = (C.c_void_p * n)(*[C.c_void_p(p) if p else None for p in d_vals_ptrs]) if d_vals_ptrs is not None else None
    rc = lib().mocob200_eval_jac_g_batch(hs, n, xs, vs)
    if rc != 0:
        raise RuntimeError("eval_jac_g_batch failed (%d): %s" % (rc, "; ".join(
            (lib().mocob200_last_error(e._h) or b"").decode() for e in evaluators)))
