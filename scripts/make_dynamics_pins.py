"""Fixtures of tests/test_reference_dynamics_pins.py: the reference's own RRA and CMC records of the full-body running trial
(opensim-models/Pipelines/Rajagopal/outputReference), to be replayed through the implicit multibody residuals.

Run in the build container only (reads /root/reference); commits
  tests/golden/rajagopal_replay_loads.mb20   the RRA-adjusted model flattened with muscles as path actuators (tension = control),
                                             locked coordinates left free, the CMC reserve actuators (Actuators_CMC.xml) and the
                                             measured ground reactions (externalLoads_run.xml, data rows 0.75 s - 1.25 s), implicit mode
  tests/golden/rajagopal_dynamics_rows.npz   per recorded instant: time, q, u, udot (radians; the files hold degrees for
                                             Rotational coordinates, radians for the Coupled patella coordinates), and
      cmc_*: the 47 rows of Results_CMC (cmc_states.sto, cmc_Kinematics_dudt.sto, cmc_Actuation_force.sto): controls = recorded
             actuator forces / optimal force, in the problem's control order
      rra_*: every 9th row of Results_RRA (rra_run_1_states.sto, ..._Kinematics_dudt.sto, ..._Actuation_force.sto): the recorded
             CoordinateActuator force of every coordinate that has one (rra_tau, NaN elsewhere). RRA replaces the force set, so
             the replay runs with all controls at zero and the residual rows must equal minus these forces.
What is NOT replayed: the six pelvis rows (the records' residual Point/TorqueActuators FX..MZ are not in the flattened model) and
the rows of the coordinates the reference locks (subtalar, mtp, wrist: their lock reactions are not recorded).
"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import __graft_entry__ as ge  # noqa: E402

D = "/root/reference/opensim-models/Pipelines/Rajagopal"
OUT = os.path.join(ROOT, "tests", "golden")
RADIANS_IN_FILES = {"pelvis_tx", "pelvis_ty", "pelvis_tz", "knee_angle_r_beta", "knee_angle_l_beta"}  # Translational / Coupled


def sto(path):
    lines = open(path).read().splitlines()
    i = [k for k, l in enumerate(lines) if l.strip().startswith("endheader")][0]
    labels = lines[i + 1].split()
    return labels, np.array([[float(v) for v in l.split()] for l in lines[i + 2:] if l.strip()])


def main():
    m = ge.package()
    p = m.Problem.example("rajagopal_replay_loads", D, "multibody_dynamics_mode=implicit")
    p.save(os.path.join(OUT, "rajagopal_replay_loads.mb20"))
    t = p.table_sizes()
    sn, cn = p.names(0), p.names(1)
    coords = [n.split("/")[-2] for n in sn[:t.nq]]
    # optimal forces of the CoordinateActuators, from the XML files (control = force / optimal force)
    import re
    opt = {}
    for f in (D + "/outputReference/subject01_RRA_Adjusted.osim", D + "/Actuators_CMC.xml"):
        txt = open(f).read()
        for blk in re.findall(r'<CoordinateActuator name="([^"]+)">(.*?)</CoordinateActuator>', txt, re.S):
            v = re.search(r"<optimal_force>([^<]+)</optimal_force>", blk[1])
            if v and blk[0] != "default":
                opt[blk[0]] = float(v.group(1))

    def rows_of(states, dudt, force, pick):
        ls, S = sto(states)
        la, A = sto(dudt)
        lf, F = sto(force)
        out = {"time": [], "q": [], "u": [], "udot": [], "F": []}
        for k in pick(A):
            tt = A[k, 0]
            r = int(np.argmin(np.abs(S[:, 0] - tt)))
            f = int(np.argmin(np.abs(F[:, 0] - tt)))
            assert abs(S[r, 0] - tt) < 2e-8 and abs(F[f, 0] - tt) < 2e-8
            out["time"].append(tt)
            out["q"].append([S[r][ls.index(c)] for c in coords])  # the states files are in radians throughout
            out["u"].append([S[r][ls.index(c + "_u")] for c in coords])
            out["udot"].append([A[k][la.index(c)] * (1.0 if c in RADIANS_IN_FILES else np.pi / 180) for c in coords])
            out["F"].append(F[f])
        return lf, {k: np.array(v) for k, v in out.items()}

    R = D + "/outputReference/Results_CMC"
    lf, cmc = rows_of(R + "/cmc_states.sto", R + "/cmc_Kinematics_dudt.sto", R + "/cmc_Actuation_force.sto", lambda A: range(len(A)))
    controls = np.zeros((len(cmc["time"]), len(cn)))
    for i, n in enumerate(cn):
        leaf = n.split("/")[-1]
        controls[:, i] = cmc["F"][:, lf.index(leaf)] / opt.get(leaf, 1.0)  # every control of the problem is in the record
    R = D + "/outputReference/Results_RRA"
    lf, rra = rows_of(R + "/rra_run_1_states.sto", R + "/rra_run_1_Kinematics_dudt.sto", R + "/rra_run_1_Actuation_force.sto",
                      lambda A: range(0, len(A), 9))
    tau = np.full((len(rra["time"]), len(coords)), np.nan)
    for i, c in enumerate(coords):
        if c in lf:
            tau[:, i] = rra["F"][:, lf.index(c)]
    np.savez_compressed(os.path.join(OUT, "rajagopal_dynamics_rows.npz"), coords=np.array(coords),
                        cmc_time=cmc["time"], cmc_q=cmc["q"], cmc_u=cmc["u"], cmc_udot=cmc["udot"], cmc_controls=controls,
                        rra_time=rra["time"], rra_q=rra["q"], rra_u=rra["u"], rra_udot=rra["udot"], rra_tau=tau)
    print("rows: cmc", len(cmc["time"]), "rra", len(rra["time"]), "| problem file",
          os.path.getsize(os.path.join(OUT, "rajagopal_replay_loads.mb20")), "bytes")


if __name__ == "__main__":
    main()
