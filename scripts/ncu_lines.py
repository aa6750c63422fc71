"""Attribute ncu SASS-level samples to CUDA source lines: ncu -i X.ncu-rep --page source --csv > src.csv, plus the cubin.
usage: ncu_lines.py src.csv cubin mangled_kernel_substring"""
import re, csv, collections, subprocess, sys
src, cubin, kern = sys.argv[1:4]
sass = subprocess.run(["nvdisasm", "-c", "-g", cubin], capture_output=True, text=True).stdout.split("\n")
start = [i for i, l in enumerate(sass) if (".text." in l and kern in l and l.strip().startswith(".section"))][0]
stack_re = re.compile(r'//## File "([^"]+)", line (\d+)(?: inlined at "([^"]+)", line (\d+))?')
cur, seq = None, []
HELPER = range(40, 70)  # small vector helpers in point_eval.cuh: attribute to the caller
for l in sass[start + 1:]:
    if l.strip().startswith(".section"): break
    m = stack_re.search(l)
    if m:
        f, ln = m.group(1).split("/")[-1], int(m.group(2))
        if m.group(3) and (f == "point_eval.cuh" and ln in HELPER or f.endswith(".hpp") or f.endswith(".h")):
            f, ln = m.group(3).split("/")[-1], int(m.group(4))
        cur = (f, ln)
        continue
    if re.match(r'\s+/\*[0-9a-f]{4,}\*/', l): seq.append(cur)
rows = list(csv.reader(open(src)))
hdr = rows[1]; ie = hdr.index("Instructions Executed"); isamp = hdr.index("# Samples")
data = rows[2:]
print("sass instr", len(seq), "csv rows", len(data))
agg, samp = collections.Counter(), collections.Counter()
for k, r in enumerate(data):
    if k >= len(seq): break
    agg[seq[k]] += int(r[ie]); samp[seq[k]] += int(r[isamp])
tot, ts = sum(agg.values()), sum(samp.values())
print("total inst", tot, "samples", ts)
for k, v in samp.most_common(int(sys.argv[4]) if len(sys.argv) > 4 else 40):
    print(k, "samples %.1f%%" % (100 * v / ts), "inst %.1f%%" % (100 * agg[k] / tot))
# phase grouping for the warp kernels
groups = [("kinematics", "point_warp.cuh", 51, 126), ("contact", "point_warp.cuh", 127, 171), ("muscle path", "point_warp.cuh", 172, 260),
          ("simm_spline", "point_eval.cuh", 74, 101), ("ppoly", "point_eval.cuh", 103, 124), ("dgf curves", "point_eval.cuh", 126, 250),
          ("activation", "point_eval.cuh", 438, 450), ("symmul", "point_eval.cuh", 451, 456),
          ("gather", "point_warp_dyn.cuh", 10, 53), ("rnea fwd", "point_warp_dyn.cuh", 54, 92), ("inward", "point_warp_dyn.cuh", 93, 144),
          ("mass", "point_warp_dyn.cuh", 145, 165), ("ldlt", "point_warp_dyn.cuh", 166, 199), ("driver+integrands", "point_warp_dyn.cuh", 200, 400),
          ("kernel", "kernels_warp.cu", 1, 400)]
g_s, g_i = collections.Counter(), collections.Counter()
for k, v in samp.items():
    name = "other"
    if k:
        for n, f, a, b in groups:
            if k[0] == f and a <= k[1] <= b: name = n
    g_s[name] += v; g_i[name] += agg[k]
print("---- phases")
for n, v in g_s.most_common(): print("%-20s samples %5.1f%%  inst %5.1f%%" % (n, 100 * v / ts, 100 * g_i[n] / tot))
