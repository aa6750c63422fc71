"""Reference-produced golden data for cylinder wrapping (run in the build container, where /root/reference exists).

Source: opensim-models/Pipelines/Rajagopal/outputReference — the reference's own CMC run on the full-body Rajagopal model
(80 muscles, 40 wrap cylinders): Results_CMC/cmc_states.sto (q, u in radians / metres), cmc_MuscleAnalysis_Length.sto
(musculotendon path lengths) and cmc_Actuation_speed.sto (path lengthening speeds), 457 rows, 8 decimals.
Writes
  tests/golden/rajagopal_replay.mb20     the replay problem (host-flattened tables, numbers only)
  tests/golden/rajagopal_cmc_rows.npz    every 4th row: time, q, u (the problem's coordinate order), lengths, speeds
                                         (the problem's muscle order)
"""
import importlib.util
import os

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
spec = importlib.util.spec_from_file_location("opensim_moco_b200", os.path.join(ROOT, "opensim-moco_b200", "__init__.py"))
m = importlib.util.module_from_spec(spec)
spec.loader.exec_module(m)

REF = os.environ.get("MOCO_REFERENCE", "/root/reference")
D = os.path.join(REF, "opensim-models/Pipelines/Rajagopal")
OUT = os.path.join(ROOT, "tests", "golden")
STRIDE = 4


def sto(path):
    lines = open(path).read().splitlines()
    i = [k for k, l in enumerate(lines) if l.strip().startswith("endheader")][0]
    labels = lines[i + 1].split()
    data = np.array([[float(v) for v in l.split()] for l in lines[i + 2:] if l.strip()])
    return labels, data


p = m.Problem.example("rajagopal_replay", D, None)
p.save(os.path.join(OUT, "rajagopal_replay.mb20"))
names_s, names_c = p.names(0), p.names(1)
t = p.table_sizes()
nq = t.nq
coords = [n.split("/")[-2] for n in names_s[:nq]]
R = os.path.join(D, "outputReference/Results_CMC")
ls, S = sto(os.path.join(R, "cmc_states.sto"))
ll, L = sto(os.path.join(R, "cmc_MuscleAnalysis_Length.sto"))
lv, V = sto(os.path.join(R, "cmc_Actuation_speed.sto"))
assert len(S) == len(L) and np.allclose(S[:, 0], L[:, 0], atol=1e-9)
mus = [n for n in ll[1:]]
order = [i for i, n in enumerate(names_c) if n.split("/")[-1] in mus]
mnames = [names_c[i].split("/")[-1] for i in order]
assert len(mnames) == t.nmus == 80


def states_at(rows):
    return (S[rows, 0], np.array([[S[r][ls.index(c)] for c in coords] for r in rows]),
            np.array([[S[r][ls.index(c + "_u")] for c in coords] for r in rows]))


# lengths: MuscleAnalysis reports on every integrator step, like the states; every 4th row is kept
rows = list(range(0, len(S), STRIDE))
time, q, u = states_at(rows)
lengths = np.array([[L[r][ll.index(n)] for n in mnames] for r in rows])
# speeds: the Actuation analysis reports every 10th step; find the states row of each of its times
srow = []
for tt in V[:, 0]:
    k = int(np.argmin(np.abs(S[:, 0] - tt)))
    assert abs(S[k, 0] - tt) < 2e-8, (tt, S[k, 0])
    srow.append(k)
time_v, q_v, u_v = states_at(srow)
speeds = np.array([[V[k][lv.index(n)] for n in mnames] for k in range(len(V))])
np.savez_compressed(os.path.join(OUT, "rajagopal_cmc_rows.npz"), time=time, q=q, u=u, lengths=lengths,
                    time_v=time_v, q_v=q_v, u_v=u_v, speeds=speeds, muscles=np.array(mnames), coords=np.array(coords),
                    control_index=np.array(order))
print(len(rows), "length rows,", len(srow), "speed rows", os.path.getsize(os.path.join(OUT, "rajagopal_cmc_rows.npz")))

# BASELINE config 4 (exampleMocoInverse on the armless Rajagopal model): the flattened problem for the tests and the bench
p4 = m.Problem.example("rajagopal_inverse", os.path.join(REF, "Moco/Examples/C++/example3DWalking"), None)
p4.save(os.path.join(OUT, "rajagopal_inverse.mb20"))
print("rajagopal_inverse", p4.dims().asdict(), os.path.getsize(os.path.join(OUT, "rajagopal_inverse.mb20")))
