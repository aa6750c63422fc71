"""K independent config-3 problems evaluated side by side on one B200 (SURVEY 8(d) sizing caveat: one 101-point Jacobian
leaves most of the GPU idle; the batched figure shows the throughput regime of the same kernels).

  python scripts/gpu_concurrent.py [--steps 100] [--ks 1,2,4,8,16,32] [--workload config3]  -> one JSON line per K

A step = one mocob200_eval_jac_g_batch over K handles (K complete Jacobians, each at its own iterate, each into its own
CSC buffer), timed with CUDA events on handles[0]'s stream (the batch forks from and joins to that stream). Inputs are
larger than L2: iterates rotate through > 126 MB, CSC buffers through > 126 MB.
"""
import argparse
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench as B  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--steps", type=int, default=100)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--ks", default="1,2,4,8,16,32")
    ap.add_argument("--workload", default="config3", choices=sorted(B.WORKLOADS))
    args = ap.parse_args()
    import torch

    import __graft_entry__ as ge

    m = ge.package()
    fixture, opts, desc = B.WORKLOADS[args.workload]
    prob = m.Problem.load(os.path.join(B.GOLD, fixture), opts)
    tr_fd = {"config3": 0, "config5": 1}[args.workload]
    ks = [int(k) for k in args.ks.split(",")]
    evs = []
    for k in range(max(ks)):
        ev = m.Evaluator(prob, device=0)
        ev.set_fd(tr_fd)
        evs.append(ev)
    d = evs[0].dims.asdict()
    L2 = 126 << 20
    n_x = min(4096, -(-int(1.05 * L2) // (8 * d["n_var"])))
    xs = torch.from_numpy(np.stack([prob.make_iterate(seed=k, spread=0.1) for k in range(n_x)])).cuda()
    flops, _ = B.flop_per_jacobian(fixture[:-5], d, tr_fd, 1 if tr_fd == 1 else 0)
    peak = m.measure_fp64_peak(0)
    for K in ks:
        n_v = max(2, -(-int(1.05 * L2) // (8 * d["nnz_jac"] * K)))
        vals = torch.empty((n_v, K, d["nnz_jac"]), dtype=torch.float64, device="cuda")
        hs = evs[:K]

        def step(s):
            m.eval_jac_g_batch(hs, [xs[(s * K + k) % n_x].data_ptr() for k in range(K)],
                               [vals[s % n_v, k].data_ptr() for k in range(K)])

        for w in range(max(args.warmup, 3)):
            step(w)
        hs[0].sync()
        torch.cuda.synchronize()
        launches0 = sum(e.launch_count() for e in hs)
        total_ms = 0.0
        for s in range(args.steps):
            torch.cuda.synchronize()
            hs[0].mark(0)
            step(s)
            hs[0].mark(1)
            hs[0].sync()
            total_ms += hs[0].elapsed_ms()
        launches = sum(e.launch_count() for e in hs) - launches0
        jac_s = K * args.steps / (total_ms * 1e-3)
        # a pipelined variant: no sync between batches (what a multi-study driver does while the host factorises)
        torch.cuda.synchronize()
        hs[0].mark(0)
        for s in range(args.steps):
            step(s)
        hs[0].mark(1)
        hs[0].sync()
        pipe_ms = hs[0].elapsed_ms()
        tf = flops * K * args.steps / (total_ms * 1e-3) / 1e12
        print(json.dumps({
            "metric": B.METRIC, "unit": B.UNIT, "concurrent_problems": K, "value": jac_s,
            "ms_per_batch": total_ms / args.steps, "back_to_back_value": K * args.steps / (pipe_ms * 1e-3),
            "point_evals_per_s": jac_s * evs[0].point_evals_per_jacobian(), "steps": args.steps,
            "whole_step_fp64": {"achieved_tflops": tf, "peak_tflops": peak, "frac": tf / peak,
                                "note": "algorithmic FLOPs of K Jacobians / whole batch time (copies, endpoint and assembly kernels included)"},
            "gpu_launches": int(launches),
            "config": {"workload": args.workload + ": " + desc, "l2": "iterates rotate through %.0f MB, CSC buffers through %.0f MB"
                       % (n_x * 8e-6 * d["n_var"], n_v * K * 8e-6 * d["nnz_jac"])}}), flush=True)
        del vals


if __name__ == "__main__":
    main()
