"""Reference-produced NLP solution as a pin (run in the build container, where /root/reference exists).

Source: Moco/Tests/std_testMocoInverse_subject_18musc_solution.sto - the solution the reference's own MocoInverse regression test
(testMocoInverse.cpp:118-148) is compared with: objective 1.087741 after 52 IPOPT iterations, 23 Hermite-Simpson grid points,
28 auxiliary states, 33 controls, 2 multipliers, 10 implicit tendon derivatives, plus the prescribed coordinate values / speeds
the reference inserted into the file. The problem itself is rebuilt here from the reference's inputs
(subject_walk_armless_18musc.osim, ..._external_loads.xml, ..._grfs.mot, ..._coordinates.mot) by the example
"moco_inverse_18musc" (opensim-moco_b200/csrc/examples.cpp: inverse18Muscles).
Writes
  tests/golden/moco_inverse_18musc.mb20          the flattened problem (numbers only)
  tests/golden/moco_inverse_18musc_solution.npz  x (the NLP vector at the reference's solution, in this problem's variable order),
                                                 objective, and the file's coordinate values / speeds per grid point
"""
import importlib.util
import os

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
spec = importlib.util.spec_from_file_location("opensim_moco_b200", os.path.join(ROOT, "opensim-moco_b200", "__init__.py"))
m = importlib.util.module_from_spec(spec)
spec.loader.exec_module(m)

REF = os.environ.get("MOCO_REFERENCE", "/root/reference")
T = os.path.join(REF, "Moco/Tests")
OUT = os.path.join(ROOT, "tests", "golden")
p = m.Problem.example("moco_inverse_18musc", T, None)
p.save(os.path.join(OUT, "moco_inverse_18musc.mb20"))
d = p.dims()
lines = open(os.path.join(T, "std_testMocoInverse_subject_18musc_solution.sto")).read().splitlines()
hdr = {l.split("=")[0]: l.split("=")[1] for l in lines if "=" in l and not l.startswith("\t")}
i = [k for k, l in enumerate(lines) if l.startswith("endheader")][0]
lab = lines[i + 1].split("\t")
D = np.array([[float(v) for v in l.split()] for l in lines[i + 2:] if l.strip()])
assert D.shape[0] == d.n_grid
col = {n: k for k, n in enumerate(lab)}
sn, cn, dn = p.names(0), p.names(1), p.names(2)
assert (len(sn), len(cn), len(dn), d.nm) == (28, 33, 10, 2)
x = np.zeros(d.n_var)
x[0], x[1] = D[0, 0], D[-1, 0]
st = np.array([D[:, col[n]] for n in sn]).T
ct = np.array([D[:, col[n]] for n in cn]).T
mu = np.array([D[:, col[n]] for n in lab if n.startswith("lambda")]).T
dv = np.array([D[:, col[n]] for n in dn]).T
x[2:] = np.concatenate([st.ravel(), ct.ravel(), mu.ravel(), dv.ravel()])
vals = [n for n in lab if n.endswith("/value")]
np.savez_compressed(os.path.join(OUT, "moco_inverse_18musc_solution.npz"), x=x, objective=float(hdr["objective"]),
                    iterations=int(hdr["num_iterations"]), time=D[:, 0], coordinates=np.array([n.split("/")[-2] for n in vals]),
                    q=np.array([D[:, col[n]] for n in vals]).T, u=np.array([D[:, col[n.replace("/value", "/speed")]] for n in vals]).T)
print(d.asdict(), hdr["objective"], os.path.getsize(os.path.join(OUT, "moco_inverse_18musc.mb20")))
