"""Reference-produced golden data for the oracle / kernel pins (run in the build container, where /root/reference exists).

Source: opensim-models/Pipelines/Gait10dof18musc/OutputReference — results the reference's own CMC tool wrote with real
OpenSim/Simbody (states, accelerations, actuator forces and path lengthening speeds along a walking trial):
  CMC/ResultsCMC/walk_subject_Kinematics_{q,u,dudt}.sto, walk_subject_Actuation_{force,speed}.sto   (4416 rows)
model CMC/subject01_metabolics.osim + gait10dof_Reserve_Actuators.xml + ExperimentalData/subject01_walk_grf.{xml,mot}.

Writes
  tests/golden/gait10dof_replay.mb20     the replay problem (host-flattened tables, numbers only)
  tests/golden/gait10dof_cmc_rows.npz    every 8th row: time, q, u, udot (radians / metres, in the problem's coordinate
                                         order), controls (recorded actuator forces in the problem's control order; the
                                         FX/FY/MZ residuals mapped onto the pelvis generalized forces), path speeds
"""
import importlib.util
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
spec = importlib.util.spec_from_file_location("opensim_moco_b200", os.path.join(ROOT, "opensim-moco_b200", "__init__.py"))
m = importlib.util.module_from_spec(spec)
spec.loader.exec_module(m)

REF = os.environ.get("MOCO_REFERENCE", "/root/reference")
D = os.path.join(REF, "opensim-models/Pipelines/Gait10dof18musc/OutputReference")
OUT = os.path.join(ROOT, "tests", "golden")
STRIDE = 8


def sto(path):
    lines = open(path).read().splitlines()
    i = [k for k, l in enumerate(lines) if l.strip().startswith("endheader")][0]
    labels = lines[i + 1].split()
    data = np.array([[float(v) for v in l.split()] for l in lines[i + 2:] if l.strip()])
    return labels, data


p = m.Problem.example("gait10dof_replay", D, None)
p.save(os.path.join(OUT, "gait10dof_replay.mb20"))
names_s, names_c = p.names(0), p.names(1)
nq = p.table_sizes().nq
coords = [n.split("/")[-2] for n in names_s[:nq]]
rot = np.array([0 if c in ("pelvis_tx", "pelvis_ty") else 1 for c in coords])  # the files hold degrees for rotations

R = os.path.join(D, "CMC/ResultsCMC")
lq, Q = sto(os.path.join(R, "walk_subject_Kinematics_q.sto"))
lu, U = sto(os.path.join(R, "walk_subject_Kinematics_u.sto"))
la, A = sto(os.path.join(R, "walk_subject_Kinematics_dudt.sto"))
lf, F = sto(os.path.join(R, "walk_subject_Actuation_force.sto"))
ls, S = sto(os.path.join(R, "walk_subject_Actuation_speed.sto"))
assert np.array_equal(Q[:, 0], U[:, 0]) and np.array_equal(Q[:, 0], A[:, 0]) and np.array_equal(Q[:, 0], F[:, 0]) and np.array_equal(Q[:, 0], S[:, 0])


def conv(labels, row):
    v = np.array([row[labels.index(c)] for c in coords])
    return np.where(rot == 1, np.deg2rad(v), v)


rows = list(range(0, len(Q), STRIDE))
time = Q[rows, 0]
q = np.array([conv(lq, Q[r]) for r in rows])
u = np.array([conv(lu, U[r]) for r in rows])
udot = np.array([conv(la, A[r]) for r in rows])
muscles = [n for n in names_c if n.split("/")[-1] in ls and not n.endswith("_reserve")]
speeds = np.array([[S[r][ls.index(n.split("/")[-1])] for n in muscles] for r in rows])
controls = np.zeros((len(rows), len(names_c)))
for k, r in enumerate(rows):
    for i, n in enumerate(names_c):
        leaf = n.split("/")[-1]
        if leaf in lf:
            controls[k, i] = F[r][lf.index(leaf)]
    # residual actuators of gait10dof_Reserve_Actuators.xml: PointActuators FX, FY at (-0.034, 0, 0) of the pelvis, global
    # directions x and y; TorqueActuator MZ about global z. The pelvis joint is (rotation about z, translation x, y) at the pelvis
    # origin, so their generalized forces are (MZ + (r x F)_z, FX, FY) with r = R(tilt) (-0.034, 0, 0).
    FX, FY, MZ = (F[r][lf.index(n)] for n in ("FX", "FY", "MZ"))
    th = q[k][coords.index("pelvis_tilt")]
    rx, ry = np.cos(th) * (-0.034), np.sin(th) * (-0.034)
    controls[k, names_c.index("/forceset/residual_pelvis_tx")] = FX
    controls[k, names_c.index("/forceset/residual_pelvis_ty")] = FY
    controls[k, names_c.index("/forceset/residual_pelvis_tilt")] = MZ + rx * FY - ry * FX
np.savez_compressed(os.path.join(OUT, "gait10dof_cmc_rows.npz"), time=time, q=q, u=u, udot=udot, controls=controls, speeds=speeds,
                    coords=np.array(coords), control_names=np.array(names_c), muscle_names=np.array(muscles))
print("rows", len(rows), "coords", coords, "muscles", len(muscles))
for f in ("gait10dof_replay.mb20", "gait10dof_cmc_rows.npz"):
    print(f, os.path.getsize(os.path.join(OUT, f)))
