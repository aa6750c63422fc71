"""bench.py — NLP Jacobian evaluations per second on the 2D-gait problems (BASELINE.json metric).

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl reference] [--workload config3|config5]

One "step" = one complete eval_jac_g (all CSC values of jac_g, finite differences through the point functions) at one of a
set of seeded iterates larger than L2.

Workloads
  config3  example2DWalking gaitTracking: Hermite-Simpson, 50 mesh intervals = 101 grid points, forward FD — the
           configuration BASELINE.json's metric is quoted on; the default of a plain `python bench.py` (N = 1).
  config5  example2DWalking gaitPrediction at 800 mesh intervals = 1601 grid points, central FD — the configuration
           BASELINE.json names for 1/2/4/8 GPUs. It is the default whenever the run is launched by torch.distributed.run
           (WORLD_SIZE in the environment, any N): grid points are sharded over the ranks by the library, one process per GPU.
           The N = 1 line of such a run carries a `config3` sub-record, and every N > 1 line an `n1` sub-record (the same
           workload on rank 0's GPU alone, measured in the same process), so that each line is self-contained.

`value`   : device-resident throughput (x already in HBM, values left in HBM), CUDA events on the library's stream, max over ranks.
`e2e`     : the same metric through the host-pointer C-ABI call mocob200_eval_jac_g with pinned host buffers
            (H2D of x and D2H of the CSC values inside the timed region).
`roofline`: the point-function FD stage against the FP64 FMA peak measured on this GPU in this process (MEASURED_PEAKS.json has
            no FP64 figure). `frac` counts the evaluations the kernels EXECUTE (partial evaluations counted as whole ones: an
            upper bound); `frac_credited` counts what differencing the whole function would cost ((n_in + 1) evaluations per
            grid point) — the work the stage delivers. Algorithmic FLOPs from profiles/flop_counts.json (oracle op counts).
`cpu_baseline` / `--impl reference`: the CPU oracle (oracle/, a from-scratch restatement pinned to reference-produced data —
            the reference itself cannot be built without Simbody/CasADi/IPOPT) on all host threads.
`solve`   : (config 3, one GPU, after the timed region; `--solve-iters 0` skips it) a stand-in for BASELINE's second metric,
            IPOPT time-to-solve, which cannot be measured here (no IPOPT): a fixed number of iterations of an interior-point
            loop (scripts/ip_solve.py) on 10-interval gait tracking with the GPU callbacks and with the oracle's callbacks —
            callback seconds of both arms and the distance between the two optimisation paths. Not a converged solve.
"""
import argparse
import ctypes
import json
import os
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
DATA = os.path.join(ROOT, "opensim-moco_b200", "data")  # flattened problems (numbers only) shipped with the package

WORKLOADS = {
    # name: (fixture, options, fd scheme, description)
    "config3": ("gait2d_tracking.mb20", None, 0,
                "example2DWalking gaitTracking: 2D_gait.osim, Hermite-Simpson, 50 intervals (101 grid points), forward FD"),
    "config5": ("gait2d_prediction.mb20", "num_mesh_intervals=800", 1,
                "example2DWalking gaitPrediction at 800 intervals (1601 grid points), central FD"),
    # BASELINE config 4 (--workload config4): exampleMocoInverse, the armless full-body Rajagopal model, 80 muscles, kinematics prescribed
    "config4": ("rajagopal_inverse.mb20", None, 0,
                "exampleMocoInverse: Rajagopal 3D model, 80 DeGrooteFregly2016 muscles, prescribed kinematics, Hermite-Simpson, "
                "49 intervals (99 grid points), implicit mode, forward FD"),
}
METRIC = "NLP Jacobian evals/s (2D gait, 101 mesh pts)"
UNIT = "jacobian evals/s"
FD_NAMES = ["forward", "central", "backward"]
L2_BYTES = 126 << 20


def host_threads():
    try:
        return max(1, len(os.sched_getaffinity(0)))
    except Exception:
        return max(1, os.cpu_count() or 1)


# ------------------------------------------------------------------------------------------------ iterates
def _pools(m, prob):
    t = prob.table_sizes()
    dbl = np.ctypeslib.as_array(ctypes.cast(t.dbl, ctypes.POINTER(ctypes.c_double)), shape=(t.n_dbl,))
    idx = np.ctypeslib.as_array(ctypes.cast(t.idx, ctypes.POINTER(ctypes.c_int)), shape=(t.n_int,))
    return t, dbl, idx


def _ppoly(t, dbl, idx, pid, times, deriv=0):
    rec = idx[t.o_ppoly_i + 3 * pid:t.o_ppoly_i + 3 * pid + 3]
    off, npc, order = int(rec[0]), int(rec[1]), int(rec[2])
    br = dbl[t.o_ppoly_data + off:t.o_ppoly_data + off + npc + 1]
    co = dbl[t.o_ppoly_data + off + npc + 1:t.o_ppoly_data + off + npc + 1 + npc * order].reshape(npc, order)
    k = np.clip(np.searchsorted(br, times, side="right") - 1, 0, npc - 1)
    s = times - br[k]
    v = np.zeros_like(times)
    for j in range(order - 1, -1, -1):
        v = v * s + co[k, j]
    return v


def grid_of(prob):
    """normalised grid times of the problem's transcription (mesh points, plus the interval midpoints for Hermite-Simpson)"""
    tr = ctypes.cast(prob.transcription, ctypes.POINTER(ctypes.c_int))
    scheme, n_mesh = tr[0], tr[2]
    mesh_ptr = ctypes.cast(ctypes.c_void_p(prob.transcription.value + 16), ctypes.POINTER(ctypes.c_void_p))[0]
    if mesh_ptr:
        mesh = np.ctypeslib.as_array(ctypes.cast(mesh_ptr, ctypes.POINTER(ctypes.c_double)), shape=(n_mesh,)).copy()
    else:
        mesh = np.linspace(0.0, 1.0, n_mesh)
    if scheme == 0:
        return mesh
    g = np.empty(2 * n_mesh - 1)
    g[0::2] = mesh
    g[1::2] = 0.5 * (mesh[:-1] + mesh[1:])
    return g


def make_iterates(m, prob, n, kind):
    """kind "trajectory" (SURVEY 8(d)): the reference walking trajectory — example2DWalking's referenceCoordinates.sto, held as the
    tracking-goal splines of the gaitTracking problem — resampled onto the grid, plus U(-0.02, 0.02) (upper - lower) on every
    coordinate and speed; activations and excitations U(0.01, 1); the lumbar actuator U(-10, 10); t0 = 0, tf = 0.47008941; the feet
    are in contact with the ground as in a solve. kind "bounds": bounds midpoint + 0.1 U(-1, 1) half-range (round 1's iterates)."""
    d = prob.dims()
    if kind == "bounds" or prob.table_sizes().prescribed:  # (prescribed kinematics: no coordinate states to put on a trajectory)
        return np.stack([prob.make_iterate(seed=k, spread=0.1) for k in range(n)])
    track = m.Problem.load(os.path.join(DATA, "gait2d_tracking.mb20"))
    t, dbl, idx = _pools(m, track)
    # goal 0 of gaitTracking is the MocoStateTrackingGoal: payload = n refs x (state index, ppoly id)
    g0 = idx[t.o_goal_i:t.o_goal_i + 8]
    assert g0[0] == 1, "expected the state tracking goal first"
    ip = idx[g0[4]:g0[4] + g0[6]]
    refs = {int(ip[1 + 2 * k]): int(ip[2 + 2 * k]) for k in range(int(ip[0]))}
    tf = 0.47008941
    grid = grid_of(prob)
    N, ns, nc, nd = d.n_grid, d.ns, d.nc, d.nd
    assert len(grid) == N
    base = np.zeros((N, ns))
    for s, pid in refs.items():
        base[:, s] = _ppoly(t, dbl, idx, pid, tf * grid)
    xl, xh, _, _ = prob.bounds()
    sl = xl[2:2 + N * ns].reshape(N, ns)
    sh = xh[2:2 + N * ns].reshape(N, ns)
    cl = xl[2 + N * ns:2 + N * ns + N * nc].reshape(N, nc)
    ch = xh[2 + N * ns:2 + N * ns + N * nc].reshape(N, nc)
    nq2 = 2 * track.table_sizes().nq
    xs = np.zeros((n, d.n_var))
    for k in range(n):
        rng = np.random.default_rng(k)
        st = base.copy()
        width = np.where(np.isfinite(sh - sl), sh - sl, 1.0)
        st[:, :nq2] += rng.uniform(-0.02, 0.02, (N, nq2)) * width[:, :nq2]
        st[:, nq2:] = rng.uniform(0.01, 1.0, (N, ns - nq2))
        ct = rng.uniform(0.01, 1.0, (N, nc))
        wide = (ch - cl) > 2.0  # the lumbar coordinate actuator ([-150, 150]); muscles are [0, 1]
        ct = np.where(wide, rng.uniform(-10.0, 10.0, (N, nc)), ct)
        st = np.minimum(np.maximum(st, sl), sh)
        ct = np.minimum(np.maximum(ct, cl), ch)
        x = xs[k]
        x[0], x[1] = 0.0, tf
        x[2:2 + N * ns] = st.reshape(-1)
        x[2 + N * ns:2 + N * ns + N * nc] = ct.reshape(-1)
        if nd:
            x[2 + N * (ns + nc):] = 0.0
        x[:] = np.minimum(np.maximum(x, xl), xh)
    return xs


def config_dict(workload, desc, dims, fd, sparsity, iterates_kind):
    """the workload description: identical in both arms (the driver compares the two dicts); run details go under "run" """
    return {"workload": workload + ": " + desc, "n_grid": int(dims["n_grid"]), "n_var": int(dims["n_var"]), "n_con": int(dims["n_con"]),
            "nnz_jac": int(dims["nnz_jac"]), "sparsity_detection": sparsity, "fd": FD_NAMES[fd],
            "iterates": ("seeded iterates: reference walking trajectory (referenceCoordinates.sto splines) + U(-0.02,0.02)*(upper-lower) "
                         "on coordinates and speeds, activations/excitations U(0.01,1), lumbar U(-10,10), tf = 0.47008941"
                         if iterates_kind == "trajectory" and not workload.startswith("config4") else
                         "seeded iterates: bounds midpoint + 0.1*U(-1,1)*half-range")}


# ------------------------------------------------------------------------------------------------ clocks
class ClockSampler:
    """SM clock / throttle reasons sampled through NVML inside this process (a polling thread), from before the warm-up to
    the end of the timed regions; `under_load` keeps the samples taken while a timed region was open."""

    def __init__(self, index):
        self.index, self.rows, self.stop_flag, self.thread, self.h = index, [], False, None, None
        self.in_region = False

    def start(self):
        try:
            import pynvml

            pynvml.nvmlInit()
            idx = self.index
            vis = os.environ.get("CUDA_VISIBLE_DEVICES")
            if vis:
                try:
                    idx = int(vis.split(",")[self.index])
                except Exception:
                    pass
            self.h = pynvml.nvmlDeviceGetHandleByIndex(idx)
            self.nv = pynvml
            self.smmax = float(pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM))
        except Exception:
            self.h = None
            return
        self.thread = threading.Thread(target=self._run, daemon=True)
        self.thread.start()

    def _run(self):
        nv = self.nv
        while not self.stop_flag:
            try:
                sm = float(nv.nvmlDeviceGetClockInfo(self.h, nv.NVML_CLOCK_SM))
                reasons = int(nv.nvmlDeviceGetCurrentClocksThrottleReasons(self.h))
                self.rows.append((sm, reasons, self.in_region))
            except Exception:
                pass
            time.sleep(0.0005)

    def stop(self):
        self.stop_flag = True
        if self.thread:
            self.thread.join(timeout=1.0)
        if not self.h or not self.rows:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": [], "samples": 0}
        nv = self.nv
        names = {"hw_slowdown": nv.nvmlClocksThrottleReasonHwSlowdown, "hw_thermal_slowdown": nv.nvmlClocksThrottleReasonHwThermalSlowdown,
                 "sw_thermal_slowdown": nv.nvmlClocksThrottleReasonSwThermalSlowdown, "sw_power_cap": nv.nvmlClocksThrottleReasonSwPowerCap}
        load = [r for r in self.rows if r[2]] or self.rows
        seen = set()
        for _, reasons, _ in load:
            for n, bit in names.items():
                if reasons & bit:
                    seen.add(n)
        return {"sm_mhz": float(np.median([r[0] for r in load])), "sm_max_mhz": self.smmax, "reasons": sorted(seen),
                "samples": len(self.rows), "samples_under_load": len([r for r in self.rows if r[2]]), "source": "NVML polled in-process"}


# ------------------------------------------------------------------------------------------------ CPU arm
def cpu_oracle_rate(m, prob, fd, xs, seconds, steps=None, warmup=1):
    """the oracle's full eval_jac_g on all host threads; returns (evals/s, n calls, wall seconds, threads)"""
    sys.path.insert(0, os.path.join(ROOT, "oracle"))
    import oracle_py as O

    threads = host_threads()
    O.lib().orc_set_threads(threads)  # torch.distributed.run sets OMP_NUM_THREADS=1: ask for the cores explicitly
    o = O.Oracle(prob)
    o.set_fd(fd)
    o.structure()
    for w in range(max(warmup, 1)):
        o.eval_jac_g(xs[w % len(xs)])
    n, t0 = 0, time.perf_counter()
    while (n < steps) if steps is not None else (time.perf_counter() - t0 < seconds or n < 2):
        o.eval_jac_g(xs[n % len(xs)])
        n += 1
    dt = time.perf_counter() - t0
    return n / dt, n, dt, threads, int(o.structure()[1].size)


def run_reference(args, rank, workload):
    """--impl reference: the CPU arm on rank 0 (all host threads), same workload / config keys / metric as the GPU arm."""
    if rank != 0:
        return
    import __graft_entry__ as ge

    m = ge.package()
    fixture, opts, fd, desc = WORKLOADS[workload]
    prob = m.Problem.load(os.path.join(DATA, fixture), opts)
    d = prob.dims().asdict()
    xs = make_iterates(m, prob, 10, args.iterates)
    # each step is one full eval_jac_g; for config 5 (0.3 s per call on 16 threads) the steps are capped so that the run ends in minutes
    steps = args.steps if workload == "config3" else min(args.steps, 12)
    val, n, dt, threads, nnz = cpu_oracle_rate(m, prob, fd, xs, 0.0, steps=steps, warmup=min(args.warmup, 2))
    d["nnz_jac"] = nnz
    line = {"metric": METRIC, "value": val, "unit": UNIT, "impl": "reference", "n_gpus": args.gpus, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": 1e3 * dt / n, "higher_is_better": True, "scaling": "strong" if workload == "config5" else "weak",
            "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "config": config_dict(workload, desc, d, fd, "none", args.iterates),
            "run": {"n_iterates": 10},
            "cpu_baseline": {"value": val, "unit": UNIT, "cores": threads, "kind": "port",
                             "sample": "%d full eval_jac_g calls (all %d grid points) in %.1f s wall; CPU restatement (oracle/) pinned to "
                                       "reference-produced CMC results, not the reference binary: Simbody/CasADi/IPOPT are absent so the "
                                       "reference cannot be built" % (n, d["n_grid"], dt)},
            "e2e": {"value": val, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    print(json.dumps(line))


# ------------------------------------------------------------------------------------------------ GPU arm
def flop_counts(problem_key):
    return json.load(open(os.path.join(ROOT, "profiles", "flop_counts.json")))["problems"][problem_key]


def credited_flops(fc, dims, fd, h_iter):
    n_int = dims["n_integrand"]
    if fd != 1:  # forward/backward: one fused evaluation per direction + the nominal one
        per_point = (dims["n_in"] + 1) * fc["all"]["flop"]
    else:  # central: every callback group refines its own step: 2 evaluations x (1 + h_iter) per direction per group
        groups = fc["dynamics"]["flop"] + sum(fc["integrand%d" % k]["flop"] for k in range(n_int))
        per_point = fc["all"]["flop"] + dims["n_in"] * 2 * (1 + h_iter) * groups
    return per_point * dims["n_grid"]


def measure(m, torch, dist, args, workload, rank, world, local_rank, nccl_id, sampler, single_rank=False):
    """times `workload`; returns the dict of measured numbers (rank 0) — device-resident and end-to-end"""
    fixture, opts, fd, desc = WORKLOADS[workload]
    prob = m.Problem.load(os.path.join(DATA, fixture), opts)
    if single_rank:
        ev = m.Evaluator(prob, device=local_rank)
        w = 1
    else:
        ev = m.Evaluator(prob, device=local_rank, rank=rank, world=world, nccl_id=nccl_id)
        w = world
    if args.sparsity != "none":
        ev.set_sparsity_detection(args.sparsity)
    if args.host_delivery == "masked":
        ev.set_host_delivery(True)
    d = ev.dims.asdict()
    ev.set_fd(fd)
    h_iter = 1 if fd == 1 else 0
    if args.l2 == "rotate":  # inputs (and outputs) larger than L2: no step finds its x or its CSC buffer in cache
        n_x = min(4096, -(-int(1.05 * L2_BYTES) // (8 * d["n_var"])))
        n_v = max(2, -(-int(1.05 * L2_BYTES) // (8 * d["nnz_jac"])))
    else:
        n_x, n_v = 10, 1
    n_distinct = min(n_x, 64)  # distinct seeded iterates; the rotation set repeats them with a per-copy perturbation of 1e-9
    base = make_iterates(m, prob, n_distinct, args.iterates)
    xl, xh, _, _ = prob.bounds()
    xs_all = np.empty((n_x, d["n_var"]))
    for k in range(n_x):
        xs_all[k] = base[k % n_distinct]
        if k >= n_distinct:
            xs_all[k, 2:] = np.minimum(np.maximum(xs_all[k, 2:] * (1.0 + 1e-9 * (k // n_distinct)), xl[2:]), xh[2:])
    xs_dev_all = torch.from_numpy(xs_all).cuda()
    vals_dev_all = torch.empty((n_v, d["nnz_jac"]), dtype=torch.float64, device="cuda")
    x_pin_all = torch.from_numpy(xs_all).pin_memory()
    vals_pin = torch.empty(d["nnz_jac"], dtype=torch.float64).pin_memory()
    flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda") if args.l2 == "flush" else None  # > 126 MB L2

    def barrier():
        if w > 1:
            dist.barrier()
        torch.cuda.synchronize()
        ev.sync()

    def dev_step(k):
        ev.eval_jac_g_device(xs_dev_all[k % n_x].data_ptr(), vals_dev_all[k % n_v].data_ptr())

    def host_step(k):
        ev.eval_jac_g_ptr(x_pin_all[k % n_x].data_ptr(), vals_pin.data_ptr())

    for k in range(max(args.warmup, 3)):
        dev_step(k)
        host_step(k)
    barrier()
    launches0 = ev.launch_count()
    # device-resident: per-step CUDA-event time on the library's stream
    sampler.in_region = True
    dev_ms, kernel_ms, asm_ms = 0.0, [], []
    for k in range(args.steps):
        if flush is not None:
            flush.fill_(k & 0xFF)
        torch.cuda.synchronize()
        if w > 1:
            dist.barrier()
        ev.mark(0)
        dev_step(k)
        ev.mark(1)
        ev.sync()
        dev_ms += ev.elapsed_ms()
        a, b = ev.last_kernel_ms()
        kernel_ms.append(a)
        asm_ms.append(b)
    sampler.in_region = False
    barrier()
    launches = ev.launch_count() - launches0
    # end-to-end: host pointers in and out (wall clock brackets the host call, which synchronises internally)
    sampler.in_region = True
    e2e_s = 0.0
    for k in range(args.steps):
        if flush is not None:
            flush.fill_(k & 0xFF)
        torch.cuda.synchronize()
        if w > 1:
            dist.barrier()
        t0 = time.perf_counter()
        host_step(k)
        e2e_s += time.perf_counter() - t0
    sampler.in_region = False
    barrier()
    t = torch.tensor([dev_ms, e2e_s * 1e3, float(np.mean(kernel_ms)), float(np.mean(asm_ms))], dtype=torch.float64, device="cuda")
    if w > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    dev_ms, e2e_ms, kern_ms, asm_ms = [float(v) for v in t.cpu()]
    res = {"prob": prob, "ev": ev, "dims": d, "fd": fd, "h_iter": h_iter, "desc": desc, "fixture": fixture, "n_x": n_x, "n_v": n_v,
           "dev_ms": dev_ms, "e2e_ms": e2e_ms, "kern_ms": kern_ms, "asm_ms": asm_ms, "launches": launches, "xs_host": base, "world": w,
           "push": ev.host_push_stats()}
    return res


def roofline(m, r, args, local_rank, workload, peak):
    d, ev, w = r["dims"], r["ev"], r["world"]
    key = r["fixture"][:-5]
    fc = flop_counts(key)
    credited = credited_flops(fc, d, r["fd"], r["h_iter"])
    ex = ev.executed_evals()  # evaluations one Jacobian launches on ALL ranks: full, pose-reusing, velocity-reusing, local-column warps
    n_exec = int(sum(ex))
    executed = n_exec * fc["all"]["flop"]  # partial evaluations counted as whole ones: an upper bound of the executed FLOPs
    workers = ev.working_ranks()
    kern_s = r["kern_ms"] * 1e-3
    algo_bytes = 8.0 * (d["n_grid"] * (d["n_in"] + d["n_out"] + d["n_out"] * d["n_in"])) / workers
    peaks = {}
    try:
        peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except Exception:
        pass
    hbm_peak = peaks.get("hbm_gbs", 6650.0)
    traffic, traffic_src = None, None
    try:  # DRAM bytes of the stage per launch from the committed `ncu --set full` capture of this workload (not measured live)
        nc = json.load(open(os.path.join(ROOT, "profiles", "ncu_r02_summary.json")))
        pick = nc.get(workload, {}).get("stage_kernels", {})
        if pick:
            traffic = float(sum(v["dram__bytes_read.sum"] + v["dram__bytes_write.sum"] for v in pick.values())) / max(workers, 1)
            traffic_src = "profiles/ncu_r02_summary.json (%s): sum over the stage's kernels, one launch each (single-GPU capture%s)" % (
                workload, "" if workers == 1 else ", divided by the %d working ranks" % workers)
    except Exception:
        pass
    warp = ev.uses_warp_kernels()
    return {"bound": "fp64",
            "kernel": "point-function FD stage: " + ("k_wpoints_fd1<0> then <1>,<2>,<3> side by side (warp per evaluation)" if warp else
                                                     "k_wpoints_nominal, then the thread-per-evaluation FD kernels + k_wlocal_central side by side"),
            "achieved": executed / workers / kern_s / 1e12, "peak": peak, "unit": "TFLOP/s",
            "frac": executed / workers / kern_s / 1e12 / peak,
            "frac_note": "EXECUTED evaluations x FLOP per whole evaluation (partial evaluations counted whole: upper bound) / stage time / peak",
            "achieved_credited": credited / workers / kern_s / 1e12, "frac_credited": credited / workers / kern_s / 1e12 / peak,
            "peak_source": "FP64 FMA saturation kernel measured in this process (MEASURED_PEAKS.json has no FP64 entry)",
            "kernel_ms": r["kern_ms"], "assemble_kernel_ms": r["asm_ms"],
            "flop_per_launch": executed / workers, "flop_per_launch_credited": credited / workers, "flop_per_point_eval": fc["all"]["flop"],
            "evals_executed_per_step": {"full": int(ex[0]), "pose_reusing": int(ex[1]), "velocity_reusing": int(ex[2]), "local_column_warps": int(ex[3])},
            "evals_credited_per_step": int(ev.point_evals_per_jacobian()),
            "hbm": {"achieved_gbs": algo_bytes / kern_s / 1e9, "peak_gbs": hbm_peak, "frac": algo_bytes / kern_s / 1e9 / hbm_peak,
                    "algorithmic_bytes": algo_bytes},
            "traffic": traffic, "traffic_source": traffic_src}, n_exec


def solve_leg(m, iters):
    """BASELINE metric (ii) stand-in, outside the timed region: `iters` iterations of an interior-point loop (scripts/ip_solve.py,
    NOT IPOPT - absent from this image) on 10-interval gait tracking, once with the GPU library's four TNLP callbacks and once
    with the CPU oracle's, from the same starting guess. A fixed iteration budget, not a converged solve: it reports callback
    seconds of both arms and how far apart the two optimisation paths are. Failures are reported, never raised."""
    try:
        sys.path.insert(0, os.path.join(ROOT, "scripts"))
        from ip_solve import solve
        from solve_standin import quiet_guess

        sys.path.insert(0, os.path.join(ROOT, "oracle"))
        import oracle_py as o

        prob = m.Problem.load(os.path.join(DATA, "gait2d_tracking.mb20"), "num_mesh_intervals=10")
        x0 = quiet_guess(sys.modules[__name__], m, prob)
        out = {"what": "interior-point stand-in (scripts/ip_solve.py), not IPOPT; fixed iteration budget, not converged; "
                       "example2DWalking gaitTracking at 10 intervals, forward FD", "iterations": iters}
        paths = {}
        for arm in ("gpu", "cpu"):
            if arm == "gpu":
                ev = m.Evaluator(prob, device=0)
            else:
                o.lib().orc_set_threads(host_threads())
                ev = o.Oracle(prob)
            t0 = time.perf_counter()
            res, st = solve(prob, ev, x0=x0, maxiter=iters)
            out[arm] = {"objective": float(res.fun), "constraint_violation": float(res.constraint_violation), "iterations": st["iterations"],
                        "callbacks": st["f"] + st["grad"] + st["g"] + st["jac"], "callback_seconds": st["t_eval"],
                        "total_seconds": time.perf_counter() - t0}
            paths[arm] = [h[1] for h in st["history"]]
        n = min(len(paths["gpu"]), len(paths["cpu"]))
        out["objective_path_max_rel_diff"] = float(max(abs(a - b) / max(1e-300, abs(b)) for a, b in zip(paths["gpu"][:n], paths["cpu"][:n])))
        out["callback_speedup"] = out["cpu"]["callback_seconds"] / max(1e-12, out["gpu"]["callback_seconds"])
        return out
    except Exception as e:  # the stand-in is a side record: it must not take the bench line with it
        return {"error": "%s: %s" % (type(e).__name__, e)}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=200)
    ap.add_argument("--warmup", type=int, default=10)
    ap.add_argument("--impl", default="b200")
    ap.add_argument("--workload", default=None, choices=sorted(WORKLOADS),
                    help="default: config3 for a plain run, config5 (the sharded configuration) under torch.distributed.run")
    ap.add_argument("--cpu-seconds", type=float, default=10.0, help="wall budget of the cpu_baseline leg")
    ap.add_argument("--sparsity", default="none", choices=["none", "random"],
                    help="optim_sparsity_detection: none = dense callback blocks (the reference solver's default, the headline "
                         "configuration); random = detected block sparsity (fewer CSC entries to assemble and copy back)")
    ap.add_argument("--l2", default="rotate", choices=["rotate", "flush"],
                    help="rotate: every step reads a different iterate out of a set larger than L2 and writes a different "
                         "output buffer (also > L2 in total); flush: 10 iterates, 256 MiB fill between steps (cold code too)")
    ap.add_argument("--iterates", default="trajectory", choices=["trajectory", "bounds"])
    ap.add_argument("--host-delivery", default="copy", choices=["copy", "masked"],
                    help="masked: mocob200_set_host_delivery(1) for the end-to-end leg (structural zeros filled by the host)")
    ap.add_argument("--solve-iters", type=int, default=40,
                    help="iterations of the interior-point stand-in run with each callback arm after the timed region (config 3, one GPU; 0 = skip)")
    ap.add_argument("--no-subrecords", action="store_true", help="skip the config3 / n1 sub-records of a distributed run")
    args = ap.parse_args()
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    distributed = "WORLD_SIZE" in os.environ
    workload = args.workload or ("config5" if distributed else "config3")
    if args.impl == "reference":
        run_reference(args, rank, workload)
        return

    import torch
    import torch.distributed as dist

    import __graft_entry__ as ge

    m = ge.package()
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device: the evaluator has no CPU path")
    torch.cuda.set_device(local_rank)
    nccl_id = None
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", rank=rank, world_size=world, device_id=torch.device("cuda", local_rank))
        idbuf = torch.zeros(128, dtype=torch.uint8)
        if rank == 0:
            raw = ctypes.create_string_buffer(128)
            assert m.lib().mocob200_nccl_unique_id(raw) == 0
            idbuf = torch.frombuffer(bytearray(raw.raw), dtype=torch.uint8).clone()
        idbuf = idbuf.cuda()
        dist.broadcast(idbuf, 0)
        nccl_id = bytes(idbuf.cpu().numpy().tobytes())

    sampler = ClockSampler(local_rank)
    sampler.start()
    r = measure(m, torch, dist, args, workload, rank, world, local_rank, nccl_id, sampler)
    sub = {}
    if distributed and not args.no_subrecords and rank == 0:
        # self-contained lines: the headline configuration next to the sharded one at N = 1, the single-GPU rate next to every N > 1
        if world == 1 and workload != "config3":
            s = measure(m, torch, dist, args, "config3", 0, 1, local_rank, None, sampler, single_rank=True)
            sub["config3"] = {"value": args.steps / (s["dev_ms"] * 1e-3), "e2e": args.steps / (s["e2e_ms"] * 1e-3), "unit": UNIT,
                              "ms_per_step": s["dev_ms"] / args.steps, "workload": "config3: " + s["desc"]}
            s["ev"].close()
        if world > 1:
            s = measure(m, torch, dist, args, workload, 0, 1, local_rank, None, sampler, single_rank=True)
            sub["n1"] = {"value": args.steps / (s["dev_ms"] * 1e-3), "e2e": args.steps / (s["e2e_ms"] * 1e-3), "unit": UNIT,
                         "ms_per_step": s["dev_ms"] / args.steps, "workload": workload + ": " + s["desc"],
                         "note": "the same workload on rank 0's GPU alone, timed in this process after the sharded run"}
            s["ev"].close()
    if world > 1:
        dist.barrier()
    clocks = sampler.stop()

    if rank == 0:
        d, ev = r["dims"], r["ev"]
        value = args.steps / (r["dev_ms"] * 1e-3)
        e2e_val = args.steps / (r["e2e_ms"] * 1e-3)
        peak = m.measure_fp64_peak(local_rank)
        roof, n_exec = roofline(m, r, args, local_rank, workload, peak)
        workers = ev.working_ranks()
        cfg = config_dict(workload, r["desc"], d, r["fd"], args.sparsity, args.iterates)
        run = ({"n_iterates": r["n_x"],
            "l2": ("inputs larger than L2: step k reads iterate k %% %d out of %.0f MB of iterates and writes CSC buffer k %% %d of %.0f MB "
                   "(L2 = 126 MB); no flush, code and model tables stay warm as in an IPOPT loop"
                   % (r["n_x"], r["n_x"] * 8e-6 * d["n_var"], r["n_v"], r["n_v"] * 8e-6 * d["nnz_jac"])) if args.l2 == "rotate" else
                  "256 MiB device fill between steps (outside the per-step CUDA events): cold code and tables",
            "point_evals_per_step": n_exec,
            "sharding": "grid points over %d of %d rank(s); per-rank FD blocks and CSC column runs, %s" % (
                workers, world, "n/a (one rank works)" if workers <= 1 else
                "stored by each rank's assembly kernel straight into rank 0's value array over NVLink (CUDA IPC peer memory, flag words)"
                if ev.peer_gather() else "NCCL send/recv to rank 0")})
        line = {
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": max(args.warmup, 3),
            "ms_per_step": r["dev_ms"] / args.steps, "higher_is_better": True, "scaling": "strong" if workload == "config5" else "weak",
            "vs_baseline": None, "dtype": "f64", "data": "synthetic", "config": cfg, "run": run, "clocks": clocks,
            "e2e": {"value": e2e_val, "unit": UNIT, "h2d_bytes_per_step": 8 * d["n_var"],
                    "d2h_bytes_per_step": 8 * (r["push"][2] if r["push"][2] > 0 else d["nnz_jac"]),
                    "ms_per_step": r["e2e_ms"] / args.steps, "api": "mocob200_eval_jac_g(handle, x_host_pinned, values_host_pinned)",
                    "delivery": ("%d of the %d CSC values stored by a kernel straight into the caller's pinned buffer, the entries that are +0.0 at "
                                 "every iterate zero-filled by the host meanwhile and checked on the device (%d calls, %d fell back to the plain copy)"
                                 % (r["push"][2], d["nnz_jac"], r["push"][0], r["push"][1])) if r["push"][2] > 0 else
                                "device-to-host copy of the whole value array"},
            "gpu_launches": int(r["launches"]), "roofline": roof}
        line.update(sub)
        if world > 1:
            line["scaling_note"] = ("this line is config 5 (the configuration that shards, BASELINE.json configs[4]); the plain N = 1 bench line is "
                                    "config 3 (101 grid points, does not shard): strong-scaling efficiency = value / (n_gpus * n1.value)")
        # CPU baseline leg (rank 0): the oracle on all host threads, bounded by --cpu-seconds
        val, n, dt, threads, _ = cpu_oracle_rate(m, r["prob"], r["fd"], r["xs_host"], args.cpu_seconds)
        line["cpu_baseline"] = {"value": val, "unit": UNIT, "cores": threads, "kind": "port",
                                "sample": "%d full eval_jac_g calls of the same workload in %.1f s wall" % (n, dt)}
        if world == 1 and workload == "config3" and args.solve_iters > 0:
            line["solve"] = solve_leg(m, args.solve_iters)
        print(json.dumps(line))
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
