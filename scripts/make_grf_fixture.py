"""tests/golden/gait2d_reference_grf.json: the coordinates and ground-reaction forces of example2DWalking's reference
trajectory (referenceCoordinates.sto / referenceGRF.sto, produced by the reference's own SmoothSphereHalfSpaceForce in
a Moco simulation of 2D_gait.osim), re-ordered to the NLP state order. Run in the build container (/root/reference);
the JSON holds numbers only. Used by tests/test_oracle_known_answers.py to pin the oracle's contact model."""
import importlib.util
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
spec = importlib.util.spec_from_file_location("opensim_moco_b200", os.path.join(ROOT, "opensim-moco_b200", "__init__.py"))
m = importlib.util.module_from_spec(spec)
spec.loader.exec_module(m)
REF = os.path.join(os.environ.get("MOCO_REFERENCE", "/root/reference"), "Moco/Examples/C++/example2DWalking")


def sto(path):
    L = open(path).read().split("\n")
    i = [k for k, l in enumerate(L) if l.startswith("endheader")][0]
    lab = L[i + 1].strip().split("\t")
    rows = [[float(v) for v in l.split("\t") if v.strip()] for l in L[i + 2:] if l.strip()]
    return lab, rows


names = m.Problem.example("gait2d_tracking", REF, None).names(0)
if isinstance(names, str):
    names = names.split("\n")
lab, C = sto(os.path.join(REF, "referenceCoordinates.sto"))
glab, G = sto(os.path.join(REF, "referenceGRF.sto"))
cols = [lab.index(n) for n in names[:10]]
gcols = [glab.index(n) for n in ("ground_force_r_vx", "ground_force_r_vy", "ground_force_l_vx", "ground_force_l_vy")]
out = {"source": "Moco/Examples/C++/example2DWalking/referenceCoordinates.sto + referenceGRF.sto",
       "coordinates": names[:10], "time": [r[0] for r in C], "q": [[r[c] for c in cols] for r in C],
       "grf_r_xy_l_xy": [[r[c] for c in gcols] for r in G]}
assert [r[0] for r in C] == [r[0] for r in G]
path = os.path.join(ROOT, "tests", "golden", "gait2d_reference_grf.json")
json.dump(out, open(path, "w"))
print(path, os.path.getsize(path))
