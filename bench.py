"""bench.py — NLP Jacobian evaluations per second on the 2D-gait tracking problem (BASELINE.json metric).

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl reference] [--workload config3|config5]

One "step" = one complete eval_jac_g (all CSC values of jac_g, coloured-FD through the point functions) at one of ten
seeded iterates. N=1 workload = BASELINE config 3 (example2DWalking gaitTracking: Hermite–Simpson, 50 mesh intervals =
101 grid points, forward FD), the configuration the metric is quoted on. N>1: grid points are sharded over ranks by the
library (no data-path collective except the gather of the per-point blocks to rank 0, NCCL), one process per GPU.

`value`  : device-resident throughput (x already in HBM, values left in HBM), CUDA events on the library's stream.
`e2e`    : the same metric through the host-pointer C-ABI call mocob200_eval_jac_g with pinned host buffers
           (H2D of x and D2H of the CSC values inside the timed region).
`roofline`: the point-function FD kernel (k_points_fd) against the FP64 FMA peak measured on this GPU in this process
           (MEASURED_PEAKS.json has no FP64 figure), algorithmic FLOPs from profiles/flop_counts.json.
`cpu_baseline` / `--impl reference`: the CPU oracle (oracle/, a from-scratch restatement — the reference itself
           cannot be built without Simbody/CasADi/IPOPT) on all host threads.
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
GOLD = os.path.join(ROOT, "tests", "golden")

WORKLOADS = {
    # name: (fixture, options, description)
    "config3": ("gait2d_tracking.mb20", None,
                "example2DWalking gaitTracking: 2D_gait.osim, Hermite-Simpson, 50 intervals (101 grid points), forward FD"),
    "config5": ("gait2d_prediction.mb20", "num_mesh_intervals=800",
                "example2DWalking gaitPrediction at 800 intervals (1601 grid points), central FD"),
}
METRIC = "NLP Jacobian evals/s (2D gait, 101 mesh pts)"
UNIT = "jacobian evals/s"


def flop_per_jacobian(problem_key, dims, fd_scheme, h_iter):
    fc = json.load(open(os.path.join(ROOT, "profiles", "flop_counts.json")))["problems"][problem_key]
    n_int = dims["n_integrand"]
    if fd_scheme != 1:  # forward/backward: one fused evaluation per direction + the nominal one
        per_point = (dims["n_in"] + 1) * fc["all"]["flop"]
    else:  # central: every callback group refines its own step: 2 evaluations x (1 + h_iter) per direction per group
        groups = fc["dynamics"]["flop"] + sum(fc["integrand%d" % k]["flop"] for k in range(n_int))
        per_point = fc["all"]["flop"] + dims["n_in"] * 2 * (1 + h_iter) * groups
    return per_point * dims["n_grid"], fc


class ClockSampler:
    """nvidia-smi clocks / throttle reasons during the timed region (B200_PROFILING.md recipe)."""

    def __init__(self, index):
        self.rows, self.proc, self.index = [], None, index

    def start(self):
        q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
             "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), "--query-gpu=" + q, "--format=csv,noheader,nounits",
                                          "-lms", "20"], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            threading.Thread(target=self._read, daemon=True).start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([c.strip() for c in line.split(",")])

    def stop(self):
        if self.proc:
            self.proc.terminate()
        sm, reasons, smmax = [], set(), None
        for r in self.rows:
            try:
                sm.append(float(r[0]))
                smmax = float(r[1])
                for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), r[3:7]):
                    if v.lower().startswith("active"):
                        reasons.add(name)
            except Exception:
                pass
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": smmax, "reasons": sorted(reasons),
                "samples": len(sm)}


def run_reference(args, rank):
    """The CPU arm: the oracle's full eval_jac_g on all host threads (rank 0 only)."""
    if rank != 0:
        return
    sys.path.insert(0, os.path.join(ROOT, "oracle"))
    import oracle_py as O

    fixture, opts, desc = WORKLOADS[args.workload]
    o = O.Oracle(os.path.join(GOLD, fixture), opts)
    threads = O.max_threads()
    xs = [o.problem.make_iterate(seed=k, spread=0.1) for k in range(10)]
    # bounded sample: for config 5 one step covers a slice of the grid points' FD blocks (the dominant cost)
    sample = "full eval_jac_g (all %d grid points)" % o.n_grid
    o.structure()
    for w in range(max(args.warmup, 1)):
        o.eval_jac_g(xs[w % 10])
    t0 = time.perf_counter()
    for k in range(args.steps):
        o.eval_jac_g(xs[k % 10])
    dt = time.perf_counter() - t0
    val = args.steps / dt
    line = {"metric": METRIC, "value": val, "unit": UNIT, "impl": "reference", "n_gpus": args.gpus, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": 1e3 * dt / args.steps, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "config": {"workload": args.workload + ": " + desc, "n_grid": o.n_grid, "n_var": o.n_var, "nnz_jac": int(o.structure()[1].size)},
            "cpu_baseline": {"value": val, "unit": UNIT, "cores": threads, "kind": "port",
                             "sample": sample + "; CPU restatement (oracle/), not the reference binary: Simbody/CasADi/IPOPT "
                                       "are absent so the reference cannot be built"},
            "e2e": {"value": val, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    print(json.dumps(line))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=200)
    ap.add_argument("--warmup", type=int, default=10)
    ap.add_argument("--impl", default="b200")
    ap.add_argument("--workload", default="config3", choices=sorted(WORKLOADS))
    ap.add_argument("--cpu-seconds", type=float, default=10.0, help="CPU work budget of the cpu_baseline leg")
    ap.add_argument("--sparsity", default="none", choices=["none", "random"],
                    help="optim_sparsity_detection: none = dense callback blocks (the reference solver's default, the headline "
                         "configuration); random = detected block sparsity (fewer CSC entries to assemble and copy back)")
    ap.add_argument("--l2", default="rotate", choices=["rotate", "flush"],
                    help="rotate: every step reads a different iterate out of a set larger than L2 and writes a different "
                         "output buffer (also > L2 in total); flush: 10 iterates, 256 MiB fill between steps (cold code too)")
    args = ap.parse_args()
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if args.impl == "reference":
        run_reference(args, rank)
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
            import ctypes

            raw = ctypes.create_string_buffer(128)
            assert m.lib().mocob200_nccl_unique_id(raw) == 0
            idbuf = torch.frombuffer(bytearray(raw.raw), dtype=torch.uint8).clone()
        idbuf = idbuf.cuda()
        dist.broadcast(idbuf, 0)
        nccl_id = bytes(idbuf.cpu().numpy().tobytes())

    fixture, opts, desc = WORKLOADS[args.workload]
    prob = m.Problem.load(os.path.join(GOLD, fixture), opts)
    ev = m.Evaluator(prob, device=local_rank, rank=rank, world=world, nccl_id=nccl_id)
    if args.sparsity != "none":
        ev.set_sparsity_detection(args.sparsity)
    d = ev.dims.asdict()
    tr_fd = {"config3": 0, "config5": 1}[args.workload]
    ev.set_fd(tr_fd)
    h_iter = 1 if tr_fd == 1 else 0
    L2_BYTES = 126 << 20
    if args.l2 == "rotate":  # inputs (and outputs) larger than L2: no step finds its x or its CSC buffer in cache
        n_x = min(4096, -(-int(1.05 * L2_BYTES) // (8 * d["n_var"])))
        n_v = max(2, -(-int(1.05 * L2_BYTES) // (8 * d["nnz_jac"])))
    else:
        n_x, n_v = 10, 1
    xs_all = np.stack([prob.make_iterate(seed=k, spread=0.1) for k in range(n_x)])
    xs_host = [xs_all[k] for k in range(min(n_x, 10))]
    xs_dev_all = torch.from_numpy(xs_all).cuda()
    vals_dev_all = torch.empty((n_v, d["nnz_jac"]), dtype=torch.float64, device="cuda")
    x_pin_all = torch.from_numpy(xs_all).pin_memory()
    vals_pin = torch.empty(d["nnz_jac"], dtype=torch.float64).pin_memory()
    flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda") if args.l2 == "flush" else None  # > 126 MB L2

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()
        ev.sync()

    def dev_step(k):
        ev.eval_jac_g_device(xs_dev_all[k % n_x].data_ptr(), vals_dev_all[k % n_v].data_ptr())

    def host_step(k):
        ev.eval_jac_g_ptr(x_pin_all[k % n_x].data_ptr(), vals_pin.data_ptr())

    def timed(step, steps):
        """per-step CUDA-event time on the library's stream; L2 flushed between steps (outside the events)."""
        total_ms, kernel_ms, asm_ms = 0.0, [], []
        for k in range(steps):
            if flush is not None:
                flush.fill_(k & 0xFF)
            torch.cuda.synchronize()
            if world > 1:
                dist.barrier()
            ev.mark(0)
            step(k)
            ev.mark(1)
            ev.sync()
            total_ms += ev.elapsed_ms()
            a, b = ev.last_kernel_ms()
            kernel_ms.append(a)
            asm_ms.append(b)
        return total_ms, float(np.mean(kernel_ms)), float(np.mean(asm_ms))

    for w in range(max(args.warmup, 3)):
        dev_step(w)
        host_step(w)
    barrier()
    sampler = ClockSampler(local_rank)
    sampler.start()
    launches0 = ev.launch_count()
    barrier()
    dev_ms, kern_ms, asm_ms = timed(dev_step, args.steps)
    barrier()
    launches = ev.launch_count() - launches0
    # end-to-end: host pointers in and out (wall clock brackets the host call, which synchronises internally)
    e2e_s = 0.0
    for k in range(args.steps):
        if flush is not None:
            flush.fill_(k & 0xFF)
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        t0 = time.perf_counter()
        host_step(k)
        e2e_s += time.perf_counter() - t0
    barrier()
    clocks = sampler.stop()
    t = torch.tensor([dev_ms, e2e_s * 1e3, kern_ms, asm_ms], dtype=torch.float64, device="cuda")
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    dev_ms, e2e_ms, kern_ms, asm_ms = [float(v) for v in t.cpu()]

    if rank == 0:
        value = args.steps / (dev_ms * 1e-3)
        e2e_val = args.steps / (e2e_ms * 1e-3)
        # roofline of the dominant kernel: this rank's share of the point-function FD work
        key = fixture[:-5]
        flops, fc = flop_per_jacobian(key, d, tr_fd, h_iter)
        eff = m.lib().mocob200_effective_world(d["n_grid"], world)  # ranks that own grid points
        flops_rank = flops / eff
        peak = m.measure_fp64_peak(local_rank)
        achieved = flops_rank / (kern_ms * 1e-3) / 1e12
        algo_bytes = 8.0 * (d["n_grid"] * (d["n_in"] + d["n_out"] + d["n_out"] * d["n_in"])) / eff
        peaks = {}
        try:
            peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
        except Exception:
            pass
        hbm_peak = peaks.get("hbm_gbs", 6650.0)
        # DRAM bytes of the stage per launch from the committed `ncu --set full` capture of this workload (not measured live)
        traffic, traffic_src = None, None
        try:
            nc = json.load(open(os.path.join(ROOT, "profiles", "ncu_r01_summary.json")))
            pick = [v for k, v in nc.items() if ("k_wpoints_fd1" in k if args.workload == "config3" else ("k_points_fd_flat" in k or "k_wlocal_central" in k or "k_wpoints_nominal" in k))]
            if pick and world == 1:
                traffic = float(sum(v["dram__bytes_read.sum"] + v["dram__bytes_write.sum"] for v in pick))
                traffic_src = "profiles/ncu_r01_summary.json: sum over the stage's kernels, one launch each"
        except Exception:
            pass
        line = {
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": max(args.warmup, 3),
            "ms_per_step": dev_ms / args.steps, "higher_is_better": True, "scaling": "strong" if world > 1 else "weak",
            "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "config": {"workload": args.workload + ": " + desc, "n_grid": d["n_grid"], "n_var": d["n_var"], "n_con": d["n_con"],
                       "nnz_jac": d["nnz_jac"], "sparsity_detection": args.sparsity, "fd": ["forward", "central", "backward"][tr_fd],
                       "iterates": "%d seeded iterates, bounds midpoint + 0.1*U(-1,1)*half-range (mt19937_64 seeds 0-%d)" % (n_x, n_x - 1),
                       "l2": ("inputs larger than L2: step k reads iterate k %% %d out of %.0f MB of distinct iterates and writes CSC "
                              "buffer k %% %d of %.0f MB (L2 = 126 MB); no flush, code and model tables stay warm as in an IPOPT loop"
                              % (n_x, n_x * 8e-6 * d["n_var"], n_v, n_v * 8e-6 * d["nnz_jac"])) if args.l2 == "rotate" else
                             "256 MiB device fill between steps (outside the per-step CUDA events): cold code and tables",
                       "point_evals_per_step": int(ev.point_evals_per_jacobian()),
                       "sharding": "grid points over %d of %d rank(s) (a rank needs >= 384 grid points to be worth its launch + "
                                   "gather latency, mocob200_effective_world); NCCL send/recv of x and of the per-point blocks" % (eff, world)},
            "clocks": clocks,
            "e2e": {"value": e2e_val, "unit": UNIT, "h2d_bytes_per_step": 8 * d["n_var"], "d2h_bytes_per_step": 8 * d["nnz_jac"],
                    "ms_per_step": e2e_ms / args.steps, "api": "mocob200_eval_jac_g(handle, x_host_pinned, values_host_pinned)"},
            "gpu_launches": int(launches),
            "roofline": {"bound": "fp64", "kernel": "point-function FD stage: " + (
                             "k_wpoints_fd1<0> then <1>,<2>,<3> side by side (warp per evaluation)" if ev.point_evals_per_jacobian() / eff <= 30000
                             else "k_wpoints_nominal then k_points_fd_flat_pair<8> (thread per (point, direction, side)) + k_wlocal_central side by side"), "achieved": achieved, "peak": peak, "unit": "TFLOP/s",
                         "frac": achieved / peak, "peak_source": "FP64 FMA saturation kernel measured in this process "
                         "(MEASURED_PEAKS.json has no FP64 entry)", "kernel_ms": kern_ms, "assemble_kernel_ms": asm_ms,
                         "flop_per_launch": flops_rank, "flop_per_point_eval": fc["all"]["flop"],
                         "hbm": {"achieved_gbs": algo_bytes / (kern_ms * 1e-3) / 1e9, "peak_gbs": hbm_peak,
                                 "frac": algo_bytes / (kern_ms * 1e-3) / 1e9 / hbm_peak, "algorithmic_bytes": algo_bytes},
                         "traffic": traffic, "traffic_source": traffic_src},
        }
        # CPU baseline leg (rank 0, N=1 only): the oracle on all host threads, bounded by --cpu-seconds
        if world == 1:
            sys.path.insert(0, os.path.join(ROOT, "oracle"))
            import oracle_py as O

            o = O.Oracle(prob)
            o.set_fd(tr_fd)
            o.structure()
            o.eval_jac_g(xs_host[0])
            n, t0 = 0, time.perf_counter()
            while time.perf_counter() - t0 < args.cpu_seconds / max(O.max_threads(), 1) * 4 or n < 2:
                o.eval_jac_g(xs_host[n % 10])
                n += 1
            dt = time.perf_counter() - t0
            line["cpu_baseline"] = {"value": n / dt, "unit": UNIT, "cores": O.max_threads(), "kind": "port",
                                    "sample": "%d full eval_jac_g calls of the same workload in %.1f s wall" % (n, dt)}
        print(json.dumps(line))
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
