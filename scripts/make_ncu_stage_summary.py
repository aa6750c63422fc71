"""profiles/ncu_r02_summary.json: the DRAM bytes of the point-function FD stage per bench workload, one launch of each of the
stage's kernels, picked from the per-kernel summaries scripts/ncu_summary.py wrote (bench.py reads `traffic` from here).
usage: make_ncu_stage_summary.py <tag>   (reads profiles/ncu_<tag>_config{3,4,5}.json)"""
import json, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
tag = sys.argv[1] if len(sys.argv) > 1 else "r02"
STAGE = {"config3": ["k_wpoints_fd1<0", "k_wpoints_fd1<1", "k_wpoints_fd1<2", "k_wpoints_fd1<3"],  # <KIND, WRAP>
         "config4": ["k_wpoints_fd1<0", "k_wpoints_fd1<1", "k_wpoints_fd1<2", "k_wpoints_fd1<3"],
         "config5": ["k_wpoints_nominal", "k_points_fd_flat_pair", "k_points_fd_vel_pair", "k_wlocal_central"]}
out = {"source": "ncu --set full --clock-control none, scripts/gpu_ncu_r02.sh, summaries profiles/ncu_%s_config{3,4,5}.json" % tag}
for wl, names in STAGE.items():
    if not os.path.exists(os.path.join(ROOT, "profiles", "ncu_%s_%s.json" % (tag, wl))): continue
    d = json.load(open(os.path.join(ROOT, "profiles", "ncu_%s_%s.json" % (tag, wl))))
    pick, other = {}, {}
    for k, v in d.items():
        if "#" in k: continue  # first launch of each kernel only
        short = k.split(" grid")[0]
        rec = {m: v[m] for m in ("dram__bytes_read.sum", "dram__bytes_write.sum", "gpu__time_duration.sum")}
        hit = [n for n in names if short.startswith(n)]
        (pick if hit else other)[short] = rec
    out[wl] = {"stage_kernels": pick, "other_kernels": other,
               "stage_dram_bytes": sum(v["dram__bytes_read.sum"] + v["dram__bytes_write.sum"] for v in pick.values())}
json.dump(out, open(os.path.join(ROOT, "profiles", "ncu_r02_summary.json"), "w"), indent=1)
print(json.dumps({k: v["stage_dram_bytes"] for k, v in out.items() if isinstance(v, dict)}))
