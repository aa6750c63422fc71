"""Flatten the reference's example problems into tests/golden/*.mb20 (run in the build container,
where /root/reference holds 2D_gait.osim and the reference .sto data; the GPU box has neither).
The .mb20 files hold only numbers (device tables + transcription), no reference source."""
import importlib.util
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
spec = importlib.util.spec_from_file_location("opensim_moco_b200", os.path.join(ROOT, "opensim-moco_b200", "__init__.py"))
m = importlib.util.module_from_spec(spec)
spec.loader.exec_module(m)

REF = os.environ.get("MOCO_REFERENCE", "/root/reference")
GAIT = os.path.join(REF, "Moco/Examples/C++/example2DWalking")
OUT = os.path.join(ROOT, "tests", "golden")
CASES = {
    "sliding_mass": ("sliding_mass", None),
    "sliding_mass_trap": ("sliding_mass", "num_mesh_intervals=19;transcription_scheme=trapezoidal"),
    "hanging_muscle": ("hanging_muscle", None),
    "hanging_muscle_rigidtendon": ("hanging_muscle_rigidtendon", None),
    "pendulum2": ("pendulum:2", None),
    "pendulum3_implicit": ("pendulum:3", "multibody_dynamics_mode=implicit"),
    # MocoControlBoundConstraint sections of testConstraints.cpp:1458-1541 (+ one with both bounds)
    "pendulum_cb_lower": ("pendulum_control_bound:lower", None),
    "pendulum_cb_upper_trap": ("pendulum_control_bound:upper", "transcription_scheme=trapezoidal"),
    "pendulum_cb_equality": ("pendulum_control_bound:equality", None),
    "pendulum_cb_both": ("pendulum_control_bound:both", "num_mesh_intervals=10;optim_finite_difference_scheme=forward"),
    # kinematic constraints (testConstraints.cpp:1371-1416, 620-693): multipliers, slacks, velocity correction
    "point_mass_coupler": ("planar_point_mass_coupler", None),
    "point_mass_coupler_implicit": ("planar_point_mass_coupler", "multibody_dynamics_mode=implicit"),
    "double_pendulum_coupler": ("double_pendulum_coupler", None),
    "double_pendulum_coupler_noderiv_trap": ("double_pendulum_coupler", "enforce_constraint_derivatives=0;transcription_scheme=trapezoidal;num_mesh_intervals=8"),
    # MocoInverse on the 2D gait model: prescribed kinematics (PositionMotion), and the same problem with free coordinates
    "gait2d_inverse": ("gait2d_inverse", None),
    "gait2d_inverse_free": ("gait2d_inverse_free", None),
    "gait2d_inverse_trap_central": ("gait2d_inverse", "transcription_scheme=trapezoidal;num_mesh_intervals=12;optim_finite_difference_scheme=central"),
    # three-link leg with a knee-like function-based joint (one coordinate, spline secondary rotations and translations)
    "knee_chain": ("knee_chain", None),
    "knee_chain_implicit_trap": ("knee_chain", "multibody_dynamics_mode=implicit;transcription_scheme=trapezoidal;optim_finite_difference_scheme=central"),
    "gait2d_tracking": ("gait2d_tracking", None),
    "gait2d_prediction": ("gait2d_prediction", None),
}
for out, (name, opts) in CASES.items():
    p = m.Problem.example(name, GAIT, opts)
    path = os.path.join(OUT, out + ".mb20")
    p.save(path)
    print(out, p.dims().asdict(), os.path.getsize(path))
