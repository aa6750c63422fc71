// The reference's example / test problems, set up through the host mirror of the
// MocoProblem + MocoCasADiSolver surface.  Each builder cites the reference program it follows.
#include <cmath>
#include <stdexcept>

#include "host_model.h"

namespace mocob200 {
namespace {

const double kPi = 3.14159265358979323846;

bool endsWith(const std::string& s, const std::string& e) {
    return s.size() >= e.size() && s.compare(s.size() - e.size(), e.size(), e) == 0;
}
std::string swapSuffix(const std::string& path, const std::string& from, const std::string& to) {
    // std::regex_replace(name, regex("_r"), "_l") of the reference replaces every occurrence
    std::string o;
    for (size_t i = 0; i < path.size();) {
        if (path.compare(i, from.size(), from) == 0) { o += to; i += from.size(); }
        else o += path[i++];
    }
    return o;
}

// exampleSlidingMass.cpp:42-66 (createSlidingMassModel)
ModelSpec slidingMassModel() {
    ModelSpec M;
    M.name = "sliding_mass";
    M.gravity = Vec3();
    BodySpec b;
    b.name = "body";
    b.mass = 2.0;
    M.bodies.push_back(b);
    JointSpec j;
    j.name = "slider";
    j.container = "/";
    j.kind = MOCOB200_JOINT_SLIDER;
    j.parent.body = -1;
    j.child.body = 0;
    j.coords.push_back({"position", -10, 10});
    M.joints.push_back(j);
    CoordActSpec a;
    a.path = "/actuator";
    a.coord = "position";
    a.optimalForce = 1;
    M.coordActs.push_back(a);
    M.actuatorOrder.emplace_back(0, 0);
    return M;
}

// Examples/C++/exampleSlidingMass/exampleSlidingMass.cpp:68-116 ; trapezoidal variant testMocoInterface.cpp:66-83
std::unique_ptr<FlatProblem> slidingMass(const std::string& options) {
    Problem P;
    P.model = slidingMassModel();
    P.setTimeBounds(Bounds(0), Bounds(0, 5));
    P.setStateInfo("/slider/position/value", Bounds(-5, 5), Bounds(0), Bounds(1));
    P.setStateInfo("/slider/position/speed", Bounds(-50, 50), Bounds(0), Bounds(0));
    P.setControlInfo("/actuator", Bounds(-50, 50));
    P.addGoal(MOCOB200_GOAL_FINAL_TIME, "cost");
    SolverSettings S;
    S.num_mesh_intervals = 50;
    S.applyOptions(options);
    return flatten(P, S);
}

// Moco/Tests/testMocoActuators.cpp:1050-1085 (createHangingMuscleModel)
ModelSpec hangingMuscleModel(bool ignoreActivationDynamics, bool ignoreTendonCompliance, bool tendonExplicit) {
    ModelSpec M;
    M.name = "isometric_muscle";
    M.gravity = Vec3();
    M.gravity.v[0] = 9.81;
    BodySpec b;
    b.name = "body";
    b.mass = 0.5;
    M.bodies.push_back(b);
    JointSpec j;
    j.name = "joint";
    j.container = "/";
    j.kind = MOCOB200_JOINT_SLIDER;
    j.parent.body = -1;
    j.child.body = 0;
    j.coords.push_back({"height", -10, 10});
    M.joints.push_back(j);
    MuscleSpec mu;
    mu.path = "/forceset/actuator";
    mu.maxIsometricForce = 30.0;
    mu.optimalFiberLength = 0.10;
    mu.tendonSlackLength = 0.05;
    mu.tendonStrainAtOneNormForce = 0.10;
    mu.ignoreActivationDynamics = ignoreActivationDynamics;
    mu.ignoreTendonCompliance = ignoreTendonCompliance;
    mu.fiberDamping = 0.01;
    mu.tendonImplicit = !tendonExplicit;
    mu.maxContractionVelocity = 10;
    mu.pennationAngleAtOptimal = 0.10;
    PathPointSpec o, i;
    o.body = -1;
    i.body = 0;
    mu.points = {o, i};
    M.muscles.push_back(mu);
    M.actuatorOrder.emplace_back(1, 0);
    return M;
}

// "Hanging muscle minimum time", testMocoActuators.cpp:1087-1150
std::unique_ptr<FlatProblem> hangingMuscle(const std::string& options, bool ignoreAct, bool ignoreTendon) {
    Problem P;
    P.model = hangingMuscleModel(ignoreAct, ignoreTendon, false);
    P.setTimeBounds(Bounds(0), Bounds(0.05, 1.0));
    P.setStateInfo("/joint/height/value", Bounds(0.14, 0.16), Bounds(0.15), Bounds(0.14));
    P.setStateInfo("/joint/height/speed", Bounds(-1, 1), Bounds(0), Bounds(0));
    P.setControlInfo("/forceset/actuator", Bounds(0.01, 1));
    if (!ignoreAct) P.addGoal(MOCOB200_GOAL_INITIAL_ACTIVATION, "initial_activation");
    if (!ignoreTendon) {
        GoalSpec& g = P.addGoal(MOCOB200_GOAL_INITIAL_VELOCITY_EQUILIBRIUM_DGF, "initial_velocity_equilibrium", 0.001);
        g.mode = MOCOB200_MODE_COST;
    }
    P.addGoal(MOCOB200_GOAL_FINAL_TIME, "cost");
    SolverSettings S;
    S.num_mesh_intervals = 25;
    S.multibody_dynamics_mode = "implicit";
    S.optim_convergence_tolerance = 1e-4;
    S.optim_constraint_tolerance = 1e-4;
    S.applyOptions(options);
    return flatten(P, S);
}

// Moco/Moco/Components/ModelFactory.cpp:33-89 (createNLinkPendulum) with a swing-up problem
// in the style of testImplicit.cpp:29-112.
std::unique_ptr<FlatProblem> pendulum(int n, const std::string& options) {
    if (n < 1) throw std::runtime_error("pendulum:<N> needs N >= 1");
    Problem P;
    ModelSpec& M = P.model;
    M.name = std::to_string(n) + "_link_pendulum";
    for (int i = 0; i < n; ++i) {
        const std::string is = std::to_string(i);
        BodySpec b;
        b.name = "b" + is;
        b.mass = 1;
        b.inertiaCom[0] = b.inertiaCom[1] = b.inertiaCom[2] = 1;
        M.bodies.push_back(b);
        JointSpec j;
        j.name = "j" + is;
        j.kind = MOCOB200_JOINT_PIN;
        j.parent.body = i - 1;
        j.child.body = i;
        j.child.X.p.v[0] = -1;
        j.coords.push_back({"q" + is, -10, 10});
        M.joints.push_back(j);
        CoordActSpec a;
        a.path = "/tau" + is;
        a.coord = "q" + is;
        M.coordActs.push_back(a);
        M.actuatorOrder.emplace_back(0, i);
    }
    P.setTimeBounds(Bounds(0), Bounds(0.5, 5));
    for (int i = 0; i < n; ++i) {
        const std::string is = std::to_string(i);
        P.setStateInfo("/jointset/j" + is + "/q" + is + "/value", Bounds(-10, 10), Bounds(0), Bounds(i == 0 ? kPi : 0.0));
        P.setStateInfo("/jointset/j" + is + "/q" + is + "/speed", Bounds(-50, 50), Bounds(0), Bounds(0));
        P.setControlInfo("/tau" + is, Bounds(-100, 100));
    }
    P.addGoal(MOCOB200_GOAL_CONTROL, "effort");
    P.addGoal(MOCOB200_GOAL_FINAL_TIME, "time", 0.1);
    SolverSettings S;
    S.num_mesh_intervals = 20;
    S.applyOptions(options);
    return flatten(P, S);
}

// The three solvable sections of TEMPLATE_TEST_CASE("MocoControlBoundConstraint") (Moco/Tests/testConstraints.cpp:1458-1541)
// on ModelFactory::createPendulum(): which = "lower" | "upper" | "equality" | "both" (lower and upper, this repo's own case).
std::unique_ptr<FlatProblem> pendulumControlBound(const std::string& which, const std::string& options) {
    Problem P;
    ModelSpec& M = P.model;
    M.name = "pendulum";
    BodySpec b;
    b.name = "b0";
    b.mass = 1;
    b.inertiaCom[0] = b.inertiaCom[1] = b.inertiaCom[2] = 1;
    M.bodies.push_back(b);
    JointSpec j;
    j.name = "j0";
    j.kind = MOCOB200_JOINT_PIN;
    j.parent.body = -1;
    j.child.body = 0;
    j.child.X.p.v[0] = -1;
    j.coords.push_back({"q0", -10, 10});
    M.joints.push_back(j);
    CoordActSpec a;
    a.path = "/tau0";
    a.coord = "q0";
    M.coordActs.push_back(a);
    M.actuatorOrder.emplace_back(0, 0);
    if (which == "upper") {
        P.setTimeBounds(Bounds(0), Bounds(0.1, 10));
        P.setStateInfo("/jointset/j0/q0/value", Bounds(0, 1), Bounds(0), Bounds(0.53));
        P.setStateInfo("/jointset/j0/q0/speed", Bounds(-10, 10), Bounds(0), Bounds(0));
        P.setControlInfo("/tau0", Bounds(-20, 20));
        P.addGoal(MOCOB200_GOAL_FINAL_TIME, "time");
    } else {
        P.setTimeBounds(Bounds(0), Bounds(1));
        P.setStateInfo("/jointset/j0/q0/value", Bounds(-10, 10), Bounds(0));
        P.setStateInfo("/jointset/j0/q0/speed", Bounds(-10, 10), Bounds(0));
        P.setControlInfo("/tau0", Bounds(-5, 5));
        P.addGoal(MOCOB200_GOAL_CONTROL, "effort");
    }
    GoalSpec& constr = P.addPathConstraint(MOCOB200_PATHCON_CONTROL_BOUND, "control_bound");
    constr.addControlPath("/tau0");
    const PiecewisePoly violateLower = piecewiseLinearFunction({0, 0.2, 0.7, 1}, {0, 0.5316, -0.3137, .0319});
    if (which == "lower") constr.lowerBound = constantFunction(0.1318);
    else if (which == "upper") constr.upperBound = constantFunction(11.236);
    else if (which == "equality") { constr.lowerBound = violateLower; constr.equalityWithLower = true; }
    else if (which == "both") { constr.lowerBound = violateLower; constr.upperBound = constantFunction(0.75); }
    else if (which != "none") throw std::runtime_error("pendulum_control_bound:<lower|upper|equality|both|none>");
    SolverSettings S;
    S.num_mesh_intervals = 25;
    S.applyOptions(options);
    return flatten(P, S);
}

// symmetry goal shared by gaitTracking / gaitPrediction (example2DWalking.cpp:80-126, 225-271)
void addSymmetryGoal(Problem& P) {
    GoalSpec& g = P.addGoal(MOCOB200_GOAL_PERIODICITY, "symmetryGoal");
    for (const JointSpec& J : P.model.joints)
        for (const CoordSpec& c : J.coords) {
            const std::string base = J.container + J.name + "/" + c.name;
            const std::string sv[2] = {base + "/value", base + "/speed"};
            if (endsWith(c.name, "_r"))
                for (auto& s : sv) g.statePairs.push_back({s, swapSuffix(s, "_r", "_l"), 1});
            if (endsWith(c.name, "_l"))
                for (auto& s : sv) g.statePairs.push_back({s, swapSuffix(s, "_l", "_r"), 1});
            if (!endsWith(c.name, "_l") && !endsWith(c.name, "_r") && !endsWith(c.name, "_tx"))
                for (auto& s : sv) g.statePairs.push_back({s, s, 1});
        }
    g.statePairs.push_back({"/jointset/groundPelvis/pelvis_tx/speed", "", 1});
    g.controlPairs.push_back({"/lumbarAct", "", 1});
    for (const MuscleSpec& mu : P.model.muscles) {
        const std::string s = mu.path + "/activation";
        if (endsWith(mu.path, "_r")) g.statePairs.push_back({s, swapSuffix(s, "_r", "_l"), 1});
        if (endsWith(mu.path, "_l")) g.statePairs.push_back({s, swapSuffix(s, "_l", "_r"), 1});
    }
}
void setGaitBounds(Problem& P) { // example2DWalking.cpp:151-169
    const double d = kPi / 180;
    P.setStateInfo("/jointset/groundPelvis/pelvis_tilt/value", Bounds(-20 * d, -10 * d));
    P.setStateInfo("/jointset/groundPelvis/pelvis_tx/value", Bounds(0, 1));
    P.setStateInfo("/jointset/groundPelvis/pelvis_ty/value", Bounds(0.75, 1.25));
    P.setStateInfo("/jointset/hip_l/hip_flexion_l/value", Bounds(-10 * d, 60 * d));
    P.setStateInfo("/jointset/hip_r/hip_flexion_r/value", Bounds(-10 * d, 60 * d));
    P.setStateInfo("/jointset/knee_l/knee_angle_l/value", Bounds(-50 * d, 0));
    P.setStateInfo("/jointset/knee_r/knee_angle_r/value", Bounds(-50 * d, 0));
    P.setStateInfo("/jointset/ankle_l/ankle_angle_l/value", Bounds(-15 * d, 25 * d));
    P.setStateInfo("/jointset/ankle_r/ankle_angle_r/value", Bounds(-15 * d, 25 * d));
    P.setStateInfo("/jointset/lumbar/lumbar/value", Bounds(0, 20 * d));
}

// example2DWalking.cpp:58-204 (gaitTracking through MocoTrack, MocoTrack.cpp:54-233).
// The reference low-pass filters the coordinate table at 6 Hz before splining
// (TabOpLowPassFilter); table pre-processing is outside the evaluator and the
// raw table is splined here.
std::unique_ptr<FlatProblem> gaitTracking(const std::string& dir, const std::string& options) {
    Problem P;
    P.model = loadOsim(dir + "/2D_gait.osim");
    StoTable ref = readSto(dir + "/referenceCoordinates.sto");
    // state tracking goal: values from the table, speeds = derivative of the value spline at the
    // table times (track_reference_position_derivatives, MocoTrack.cpp:180-191), all re-splined (degree 5)
    GoalSpec& st = P.addGoal(MOCOB200_GOAL_STATE_TRACKING, "state_tracking", 1.0);
    std::vector<GoalSpec::Ref> speeds;
    for (size_t c = 0; c < ref.labels.size(); ++c) {
        GoalSpec::Ref r;
        r.name = ref.labels[c];
        r.spline = fitNaturalSpline(ref.time, ref.cols[c], 5);
        st.refs.push_back(r);
    }
    for (const JointSpec& J : P.model.joints)
        for (const CoordSpec& c : J.coords) {
            const std::string base = J.container + J.name + "/" + c.name;
            const int ci = ref.col(base + "/value");
            if (ci < 0 || ref.col(base + "/speed") >= 0) continue;
            std::vector<double> sp(ref.time.size());
            for (size_t k = 0; k < ref.time.size(); ++k) sp[k] = evalPiecewisePoly(st.refs[ci].spline, ref.time[k], 1);
            GoalSpec::Ref r;
            r.name = base + "/speed";
            r.spline = fitNaturalSpline(ref.time, sp, 5);
            speeds.push_back(r);
        }
    st.refs.insert(st.refs.end(), speeds.begin(), speeds.end());
    P.addGoal(MOCOB200_GOAL_CONTROL, "control_effort", 10.0);
    P.setTimeBounds(Bounds(0.0), Bounds(0.47008941));
    addSymmetryGoal(P);
    {
        StoTable grf = readSto(dir + "/referenceGRF.sto");
        GoalSpec& ct = P.addGoal(MOCOB200_GOAL_CONTACT_TRACKING, "contact", 1.0);
        auto group = [&](std::vector<std::string> cs, const std::string& body, const std::string& id) {
            GoalSpec::ContactGroup g;
            g.contacts = cs;
            g.appliedToBody = body;
            const char* ax[3] = {"x", "y", "z"};
            for (int a = 0; a < 3; ++a) {
                const int ci = grf.col(id + ax[a]);
                if (ci < 0) throw std::runtime_error("referenceGRF.sto lacks column " + id + ax[a]);
                g.ref[a] = fitNaturalSpline(grf.time, grf.cols[ci], 5);
            }
            ct.groups.push_back(g);
        };
        // referenceGRF.xml: Right_GRF applied_to_body calcn_r, force_identifier ground_force_r_v
        group({"contactHeel_r", "contactFront_r"}, "calcn_r", "ground_force_r_v");
        group({"contactHeel_l", "contactFront_l"}, "calcn_l", "ground_force_l_v");
        ct.projection = 2; // "plane"
        ct.projectionVector.v[2] = 1;
    }
    setGaitBounds(P);
    SolverSettings S;
    S.num_mesh_intervals = 50;
    S.multibody_dynamics_mode = "explicit";
    S.optim_finite_difference_scheme = "forward"; // MocoTrack.cpp:112
    S.optim_convergence_tolerance = 1e-4;
    S.optim_constraint_tolerance = 1e-4;
    S.optim_max_iterations = 1000;
    S.applyOptions(options);
    return flatten(P, S);
}

// example2DWalking.cpp:211-340 (gaitPrediction). BASELINE config 5 runs it at 800 mesh intervals.
std::unique_ptr<FlatProblem> gaitPrediction(const std::string& dir, const std::string& options) {
    Problem P;
    P.model = loadOsim(dir + "/2D_gait.osim");
    addSymmetryGoal(P);
    GoalSpec& sp = P.addGoal(MOCOB200_GOAL_AVERAGE_SPEED, "speed");
    sp.desiredAverageSpeed = 1.2;
    GoalSpec& ef = P.addGoal(MOCOB200_GOAL_CONTROL, "effort", 10);
    ef.exponent = 3;
    ef.divideByDisplacement = true;
    P.setTimeBounds(Bounds(0), Bounds(0.4, 0.6));
    setGaitBounds(P);
    SolverSettings S;
    S.num_mesh_intervals = 50;
    S.optim_convergence_tolerance = 1e-4;
    S.optim_constraint_tolerance = 1e-4;
    S.optim_max_iterations = 1000;
    S.applyOptions(options);
    return flatten(P, S);
}

// "Planar point mass with CoordinateCouplerConstraint" (Moco/Tests/testConstraints.cpp:1371-1416; model
// ModelFactory::createPlanarPointMass, ModelFactory.cpp:91-132): the coupler ty = tx makes the multiplier 0.5 (Fx - Fy).
std::unique_ptr<FlatProblem> planarPointMassCoupler(const std::string& options) {
    Problem P;
    ModelSpec& M = P.model;
    M.name = "planar_point_mass";
    M.gravity = Vec3();
    BodySpec intermed, body;
    intermed.name = "intermed";
    intermed.mass = 0;
    body.name = "body";
    body.mass = 1;
    M.bodies = {intermed, body};
    JointSpec jx, jy;
    jx.name = "tx";
    jx.kind = MOCOB200_JOINT_SLIDER;
    jx.parent.body = -1;
    jx.child.body = 0;
    jx.coords.push_back({"tx", -10, 10});
    jy.name = "ty"; // the joint's x axis points in the global +y direction
    jy.kind = MOCOB200_JOINT_SLIDER;
    jy.parent.body = 0;
    jy.child.body = 1;
    jy.parent.X.R = rotXYZ(0, 0, 0.5 * kPi);
    jy.child.X.R = rotXYZ(0, 0, 0.5 * kPi);
    jy.coords.push_back({"ty", -10, 10});
    M.joints = {jx, jy};
    const char* names[2] = {"force_x", "force_y"};
    const char* coords[2] = {"tx", "ty"};
    for (int k = 0; k < 2; ++k) {
        CoordActSpec a;
        a.path = std::string("/forceset/") + names[k];
        a.coord = coords[k];
        M.coordActs.push_back(a);
        M.actuatorOrder.emplace_back(0, k);
    }
    CouplerSpec c;
    c.name = "coupler";
    c.independent = "tx";
    c.dependent = "ty";
    c.fn.fnKind = MOCOB200_FN_LINEAR;
    c.fn.c0 = 1.0;
    c.fn.c1 = 0.0;
    M.couplers.push_back(c);
    P.setTimeBounds(Bounds(0), Bounds(1));
    P.setStateInfo("/jointset/tx/tx/value", Bounds(-5, 5), Bounds(0), Bounds(3));
    P.setStateInfo("/jointset/tx/tx/speed", Bounds(-5, 5), Bounds(0), Bounds(0));
    P.setControlInfo("/forceset/force_x", Bounds(0.5));
    P.addGoal(MOCOB200_GOAL_CONTROL, "effort");
    SolverSettings S;
    S.num_mesh_intervals = 10;
    S.enforce_constraint_derivatives = true;
    S.applyOptions(options);
    return flatten(P, S);
}

// testDoublePendulumCoordinateCoupler (Moco/Tests/testConstraints.cpp:620-693; model createDoublePendulumModel :391-440):
// q1 = -2 q0 + pi, boundary conditions on q0 only, multipliers minimised with weight 10.
std::unique_ptr<FlatProblem> doublePendulumCoupler(const std::string& options) {
    Problem P;
    ModelSpec& M = P.model;
    M.name = "double_pendulum";
    for (int i = 0; i < 2; ++i) {
        const std::string is = std::to_string(i);
        BodySpec b;
        b.name = "b" + is;
        b.mass = 1;
        b.inertiaCom[0] = b.inertiaCom[1] = b.inertiaCom[2] = 1;
        M.bodies.push_back(b);
        JointSpec j;
        j.name = "j" + is;
        j.kind = MOCOB200_JOINT_PIN;
        j.parent.body = i - 1;
        j.child.body = i;
        j.child.X.p.v[0] = -1;
        j.coords.push_back({"q" + is, -10, 10});
        M.joints.push_back(j);
        CoordActSpec a;
        a.path = "/tau" + is;
        a.coord = "q" + is;
        M.coordActs.push_back(a);
        M.actuatorOrder.emplace_back(0, i);
    }
    CouplerSpec c;
    c.name = "coupler";
    c.independent = "q0";
    c.dependent = "q1";
    c.fn.fnKind = MOCOB200_FN_LINEAR;
    c.fn.c0 = -2.0;
    c.fn.c1 = kPi;
    M.couplers.push_back(c);
    P.setTimeBounds(Bounds(0), Bounds(1));
    P.setStateInfo("/jointset/j0/q0/value", Bounds(-5, 5), Bounds(0), Bounds(kPi / 2));
    P.setStateInfo("/jointset/j0/q0/speed", Bounds(-10, 10), Bounds(0), Bounds(0));
    P.setStateInfo("/jointset/j1/q1/value", Bounds(-10, 10));
    P.setStateInfo("/jointset/j1/q1/speed", Bounds(-5, 5), Bounds(0), Bounds(0));
    P.setControlInfo("/tau0", Bounds(-50, 50));
    P.setControlInfo("/tau1", Bounds(-50, 50));
    P.addGoal(MOCOB200_GOAL_CONTROL, "effort");
    SolverSettings S;
    S.num_mesh_intervals = 20;
    S.enforce_constraint_derivatives = true;
    S.minimize_lagrange_multipliers = true;
    S.lagrange_multiplier_weight = 10;
    S.applyOptions(options);
    return flatten(P, S);
}

// A three-link leg whose middle joint is a function-based joint driven by ONE coordinate, shaped like the knee of the
// full-body models (primary rotation linear in the coordinate, two secondary rotations and two translations SimmSplines of
// it): small enough for the thread-per-evaluation kernels, so that the one-coordinate case of the joint's velocity-product
// term (point_eval.cuh: custom_joint_bias) is exercised on both kernel families. Torque-driven, effort goal.
std::unique_ptr<FlatProblem> kneeChain(const std::string& options) {
    Problem P;
    ModelSpec& M = P.model;
    M.name = "knee_chain";
    const double mass[3] = {8.0, 3.5, 1.2}, len[3] = {0.42, 0.40, 0.15};
    for (int i = 0; i < 3; ++i) {
        const std::string is = std::to_string(i);
        BodySpec b;
        b.name = "seg" + is;
        b.mass = mass[i];
        b.com = Vec3{0.01 * (i + 1), -0.45 * len[i], 0.005 * (2 - i)};
        b.inertiaCom[0] = 0.08 * mass[i] * len[i] * len[i] + 0.001;
        b.inertiaCom[1] = 0.02 * mass[i] * len[i] * len[i] + 0.0005;
        b.inertiaCom[2] = 0.085 * mass[i] * len[i] * len[i] + 0.001;
        b.inertiaCom[3] = 0.0004 * (i + 1); // products of inertia: a general tensor
        b.inertiaCom[4] = -0.0003;
        b.inertiaCom[5] = 0.0002 * (i + 1);
        M.bodies.push_back(b);
        JointSpec j;
        j.name = "j" + is;
        j.parent.body = i - 1;
        j.child.body = i;
        if (i > 0) j.parent.X.p.v[1] = -len[i - 1];
        j.coords.push_back({"q" + is, -3, 3});
        if (i == 1) {
            j.kind = MOCOB200_JOINT_CUSTOM;
            const double xs[7] = {0, 0.35, 0.70, 1.05, 1.40, 1.75, 2.10};
            const double r2[7] = {0, 0.0227, 0.0332, 0.0309, 0.0189, 0.0044, -0.0017};
            const double r3[7] = {0, -0.1094, -0.1839, -0.2300, -0.2540, -0.2624, -0.2617};
            const double t1[7] = {0, 0.000835, 0.001251, 0.001391, 0.0014, 0.001421, 0.001599};
            const double t2[7] = {0, -0.001899, -0.003492, -0.004777, -0.005756, -0.006427, -0.006792};
            auto spline = [&](TransformAxisSpec& a, const double* y, double scale) {
                a.fnKind = MOCOB200_FN_SIMMSPLINE;
                a.spline.x.assign(xs, xs + 7);
                a.spline.y.assign(y, y + 7);
                a.spline.scale = scale;
                a.coord = "q1";
            };
            j.axes[0].axis = Vec3{1, 0, 0}; j.axes[0].fnKind = MOCOB200_FN_LINEAR; j.axes[0].c0 = 1; j.axes[0].coord = "q1";
            j.axes[1].axis = Vec3{0, 0, 1}; spline(j.axes[1], r2, 1.0);
            j.axes[2].axis = Vec3{0, 1, 0}; spline(j.axes[2], r3, 1.0);
            j.axes[3].axis = Vec3{0, 1, 0}; spline(j.axes[3], t1, 1.1);
            j.axes[4].axis = Vec3{0, 0, 1}; spline(j.axes[4], t2, 1.1);
            j.axes[5].axis = Vec3{1, 0, 0}; // constant 0
        } else j.kind = MOCOB200_JOINT_PIN;
        M.joints.push_back(j);
        CoordActSpec a;
        a.path = "/tau" + is;
        a.coord = "q" + is;
        a.optimalForce = 20;
        M.coordActs.push_back(a);
        M.actuatorOrder.emplace_back(0, i);
    }
    P.setTimeBounds(Bounds(0), Bounds(0.6));
    const double q0[3] = {0.3, 0.2, -0.1}, q1[3] = {-0.4, 1.3, 0.3};
    for (int i = 0; i < 3; ++i) {
        const std::string is = std::to_string(i), base = "/jointset/j" + is + "/q" + is;
        P.setStateInfo(base + "/value", Bounds(-2.5, 2.5), Bounds(q0[i]), Bounds(q1[i]));
        P.setStateInfo(base + "/speed", Bounds(-20, 20), Bounds(0), Bounds());
        P.setControlInfo("/tau" + is, Bounds(-5, 5));
    }
    P.addGoal(MOCOB200_GOAL_CONTROL, "effort");
    SolverSettings S;
    S.num_mesh_intervals = 12;
    S.applyOptions(options);
    return flatten(P, S);
}

// Replay model for the reference-held CMC / ID results of opensim-models/Pipelines/Gait10dof18musc/OutputReference
// (CMC/walk_Setup_CMC.xml): CMC/subject01_metabolics.osim - the variant of subject01.osim (document Version 30000, CustomJoint
// pelvis and knees; its gluteus maximus path differs) whose path speeds reproduce ResultsCMC/walk_subject_Actuation_speed.sto
// to the file's 8 decimals, i.e. the model those results were written with - with every muscle read as a
// PathActuator of optimal force 1 (control = recorded tension), the reserve CoordinateActuators of
// gait10dof_Reserve_Actuators.xml, the measured ground reactions of ExperimentalData/subject01_walk_grf.xml as
// ExternalForces, and three CoordinateActuators on the pelvis coordinates standing for the FX / FY / MZ residual
// Point/TorqueActuators (the caller maps their recorded forces onto pelvis generalized forces). dir = .../OutputReference
std::unique_ptr<FlatProblem> gait10dofReplay(const std::string& dir, const std::string& options) {
    Problem P;
    P.model = loadOsim(dir + "/CMC/subject01_metabolics.osim", /*musclesAsPathActuators=*/true);
    appendForceSetFile(P.model, dir + "/gait10dof_Reserve_Actuators.xml");
    const char* pelvis[3] = {"pelvis_tilt", "pelvis_tx", "pelvis_ty"};
    for (const char* c : pelvis) {
        CoordActSpec a;
        a.path = std::string("/forceset/residual_") + c;
        a.coord = c;
        P.model.actuatorOrder.emplace_back(0, (int)P.model.coordActs.size());
        P.model.coordActs.push_back(a);
    }
    appendExternalLoads(P.model, dir + "/ExperimentalData/subject01_walk_grf.xml");
    P.setTimeBounds(Bounds(0.03), Bounds(2.494));
    P.defaultSpeedBounds = Bounds(-50, 50);
    for (auto& a : P.model.coordActs) P.setControlInfo(a.path, Bounds(-1e4, 1e4));
    for (auto& m : P.model.muscles) P.setControlInfo(m.path, Bounds(0, 1e4));
    P.addGoal(MOCOB200_GOAL_CONTROL, "effort");
    SolverSettings S;
    S.num_mesh_intervals = 10;
    S.transcription_scheme = "trapezoidal";
    S.applyOptions(options);
    return flatten(P, S);
}

// Replay of the reference's CMC run on the full-body Rajagopal model (opensim-models/Pipelines/Rajagopal, model
// outputReference/subject01_RRA_Adjusted.osim: 22 bodies, 39 coordinates, 80 muscles, 40 wrap cylinders, document Version
// 30000): the muscles become path actuators driven by the recorded tensions; used by the tests to pin cylinder wrapping
// (path lengths and lengthening speeds) against what the reference printed (Results_CMC/cmc_MuscleAnalysis_Length.sto,
// cmc_Actuation_speed.sto). Locked coordinates are read as free ones (they stay constant in the recorded states).
// withLoads: also the CMC actuator set (Actuators_CMC.xml: reserve CoordinateActuators; its pelvis Point / Torque actuators are
// left out - they load the six pelvis rows only) and the measured ground reactions (externalLoads_run.xml), for the replay of
// the recorded accelerations through the implicit multibody residuals.
std::unique_ptr<FlatProblem> rajagopalReplay(const std::string& dir, const std::string& options, bool withLoads = false) {
    Problem P;
    P.model = loadOsim(dir + "/outputReference/subject01_RRA_Adjusted.osim", /*musclesAsPathActuators=*/true, /*lockedAsFree=*/true);
    if (withLoads) {
        appendForceSetFile(P.model, dir + "/Actuators_CMC.xml");
        appendExternalLoads(P.model, dir + "/externalLoads_run.xml", 0.75, 1.25);
    }
    P.setTimeBounds(Bounds(0.87), Bounds(1.13));
    P.defaultSpeedBounds = Bounds(-50, 50);
    for (auto& a : P.model.coordActs) P.setControlInfo(a.path, Bounds(-1e4, 1e4));
    for (auto& m : P.model.muscles) P.setControlInfo(m.path, Bounds(0, 1e4));
    P.addGoal(MOCOB200_GOAL_CONTROL, "effort");
    SolverSettings S;
    S.num_mesh_intervals = 10;
    S.transcription_scheme = "trapezoidal";
    S.applyOptions(options);
    return flatten(P, S);
}

// ---- pieces of MocoInverse / TableProcessor shared by the full-body inverse problems
// The value columns of a coordinates table as OpenSim reads them for a model: labels are state paths (".../value") or bare
// coordinate names (older files; TableProcessor::process -> updateStateLabels40); with inDegrees = yes only the columns of
// Rotational coordinates are converted (SimbodyEngine.cpp:905-945; a CustomJoint coordinate is Rotational when a rotation axis
// follows it with slope +-1, CustomJoint.cpp:150-200).
std::map<std::string, std::vector<double>> coordinateColumns(const ModelSpec& M, const StoTable& tab) {
    const double deg = tab.inDegrees ? kPi / 180.0 : 1.0;
    std::map<std::string, std::vector<double>> values;
    for (const JointSpec& J : M.joints)
        for (const CoordSpec& c : J.coords) {
            const std::string path = J.container + J.name + "/" + c.name;
            int ci = tab.col(path + "/value");
            if (ci < 0) ci = tab.col(c.name);
            if (ci < 0) throw std::runtime_error("the coordinates table lacks '" + c.name + "'");
            bool rotational = J.kind == MOCOB200_JOINT_PIN;
            if (J.kind == MOCOB200_JOINT_CUSTOM)
                for (int a = 0; a < 3; ++a)
                    rotational = rotational || (J.axes[a].coord == c.name && J.axes[a].fnKind == MOCOB200_FN_LINEAR && std::fabs(J.axes[a].c0) == 1.0);
            std::vector<double> v = tab.cols[ci];
            if (rotational) for (double& x : v) x *= deg;
            values[c.name] = v;
        }
    return values;
}
// TabOpLowPassFilter (Common/TableProcessor.h:160-175 -> MocoUtilities.cpp:199-208): Storage::pad(rows / 2) - the signal
// continued at both ends by its reflection through the end point (Signal.cpp:418-450) - then a third-order Butterworth
// filter run forwards and backwards (Signal::LowpassIIR, Signal.cpp:160-245). The padded rows stay in the table.
void lowpassFilter(std::vector<double>& time, std::map<std::string, std::vector<double>>& values, double cutoff) {
    const int n = (int)time.size(), pad = n / 2;
    auto padded = [&](const std::vector<double>& s) {
        std::vector<double> r(n + 2 * pad);
        for (int i = 0, j = pad; i < pad; ++i, --j) r[i] = 2.0 * s[0] - s[j];
        for (int i = 0; i < n; ++i) r[pad + i] = s[i];
        for (int i = pad + n, j = n - 2; i < 2 * pad + n; ++i, --j) r[i] = 2.0 * s[n - 1] - s[j];
        return r;
    };
    time = padded(time);
    const int N = (int)time.size();
    double dt = time[1] - time[0];
    for (int i = 1; i + 1 < N; ++i) dt = std::min(dt, time[i + 1] - time[i]); // Storage::getMinTimeStep (uniform tables: no resampling)
    const double wa = std::tan(2.0 * kPi * cutoff * dt / 2.0), wa2 = wa * wa, wa3 = wa2 * wa;
    const double denom = (wa + 1) * (wa * wa + wa + 1.0);
    const double a[4] = {wa3 / denom, 3 * wa3 / denom, 3 * wa3 / denom, wa3 / denom};
    const double b[4] = {1, (3 * wa3 + 2 * wa2 - 2 * wa - 3) / denom, (3 * wa3 - 2 * wa2 - 2 * wa + 3) / denom, (wa - 1) * (wa2 - wa + 1) / denom};
    auto pass = [&](const std::vector<double>& in) {
        std::vector<double> out(N);
        for (int i = 0; i <= 3 && i < N; ++i) out[i] = in[i];
        for (int i = 3; i < N; ++i)
            out[i] = a[0] * in[i] + a[1] * in[i - 1] + a[2] * in[i - 2] + a[3] * in[i - 3] - b[1] * out[i - 1] - b[2] * out[i - 2] - b[3] * out[i - 3];
        return std::vector<double>(out.rbegin(), out.rend());
    };
    for (auto& kv : values) kv.second = pass(pass(padded(kv.second)));
}
// ModOpReplaceJointsWithWelds (ModelFactory.cpp: replaceJointWithWeldJoint): same frames, no coordinates
void replaceJointsWithWelds(ModelSpec& M, const std::vector<std::string>& names) {
    for (const std::string& n : names) {
        bool found = false;
        for (JointSpec& J : M.joints)
            if (J.name == n) {
                found = true;
                J.kind = MOCOB200_JOINT_WELD;
                J.coords.clear();
                for (auto& A : J.axes) A = TransformAxisSpec();
            }
        if (!found) throw std::runtime_error("no joint named '" + n + "'");
    }
}
// MocoInverse::initializeInternal (MocoInverse.cpp:44-140) from here on: dependent coordinates follow their coupler functions
// (StatesTrajectory::createFromStatesStorage with assemble = true moves the coordinates that are free to satisfy constraints),
// PositionMotion from GCV splines (degree 5) of the value columns, MocoControlGoal + MocoInitialActivationGoal, solver settings.
std::unique_ptr<FlatProblem> finishInverse(Problem& P, const std::vector<double>& time, std::map<std::string, std::vector<double>>& values,
                                           bool addReserves, double t0, double tf, int intervals, const std::string& options) {
    for (const CouplerSpec& C : P.model.couplers) {
        if (C.fn.fnKind == MOCOB200_FN_SIMMSPLINE) throw std::runtime_error("coupler '" + C.name + "': spline-coupled kinematics are unsupported here");
        const std::vector<double>& qi = values.at(C.independent);
        std::vector<double>& qd = values.at(C.dependent);
        for (size_t r = 0; r < qd.size(); ++r)
            qd[r] = C.scaleFactor * (C.fn.fnKind == MOCOB200_FN_LINEAR ? C.fn.c0 * qi[r] + C.fn.c1 : C.fn.c1);
    }
    for (const JointSpec& J : P.model.joints)
        for (const CoordSpec& c : J.coords) {
            const std::string path = J.container + J.name + "/" + c.name;
            P.model.positionMotion.emplace_back(path, fitNaturalSpline(time, values.at(c.name), 5));
            if (!addReserves) continue;
            bool has = false; // ModOpAddReserves(1.0): ModelFactory.cpp:265-320
            for (const CoordActSpec& a : P.model.coordActs) has = has || a.coord == c.name;
            for (const CouplerSpec& C : P.model.couplers) has = has || C.dependent == c.name; // constrained: no reserve (:294)
            if (has) continue;
            CoordActSpec a;
            std::string n = path;
            for (char& ch : n) if (ch == '/') ch = '_';
            a.path = "/forceset/reserve" + n;
            a.coord = c.name;
            a.optimalForce = 1.0;
            P.model.actuatorOrder.emplace_back(0, (int)P.model.coordActs.size());
            P.model.coordActs.push_back(a);
        }
    P.setTimeBounds(Bounds(t0), Bounds(tf));
    P.addGoal(MOCOB200_GOAL_CONTROL, "excitation_effort", 1.0);
    P.addGoal(MOCOB200_GOAL_INITIAL_ACTIVATION, "initial_activation");
    SolverSettings S;
    S.num_mesh_intervals = intervals;
    S.multibody_dynamics_mode = "implicit";
    S.interpolate_control_midpoints = false;
    S.minimize_implicit_auxiliary_derivatives = true;
    S.implicit_auxiliary_derivatives_weight = 0.01;
    S.optim_finite_difference_scheme = "forward";
    // (MocoInverse also sets optim_sparsity_detection = "random": a property of the evaluator handle here, mocob200_set_sparsity_detection)
    S.optim_convergence_tolerance = 1e-3;
    S.optim_constraint_tolerance = 1e-3;
    S.applyOptions(options);
    return flatten(P, S);
}

// BASELINE config 4: exampleMocoInverse (Moco/Examples/C++/example3DWalking/exampleMocoInverse.cpp:34-77) - MocoInverse
// on the armless full-body Rajagopal model, 80 muscles, kinematics of coordinates.sto prescribed.
// Model operators, in the example's order: ModOpAddExternalLoads(grf_walk.xml), ModOpIgnoreTendonCompliance,
// ModOpReplaceMusclesWithDeGrooteFregly2016 (copies the path POINTS only: the DeGrooteFregly2016 muscles of this problem have
// no wrap surfaces, in the reference too - DeGrooteFregly2016Muscle.cpp:1009-1021), ModOpIgnorePassiveFiberForcesDGF,
// ModOpScaleActiveFiberForceCurveWidthDGF(1.5), ModOpAddReserves(1.0). Time range [0.81, 1.79], mesh interval 0.02 s.
std::unique_ptr<FlatProblem> rajagopalInverse(const std::string& dir, const std::string& options) {
    Problem P;
    P.model = loadOsim(dir + "/subject_walk_armless.osim", false, false, /*musclesAsDGF=*/true);
    appendExternalLoads(P.model, dir + "/grf_walk.xml");
    for (MuscleSpec& mu : P.model.muscles) {
        mu.ignoreTendonCompliance = true;
        mu.ignorePassiveFiberForce = true;
        mu.activeForceWidthScale = 1.5;
    }
    const StoTable tab = readSto(dir + "/coordinates.sto");
    auto values = coordinateColumns(P.model, tab);
    return finishInverse(P, tab.time, values, /*addReserves=*/true, 0.81, 1.79, 49, options);
}

// The reference's regression test of MocoInverse (Moco/Tests/testMocoInverse.cpp:118-148, "MocoInverse Rajagopal2016, 18
// muscles"), whose SOLUTION the reference ships (std_testMocoInverse_subject_18musc_solution.sto: objective 1.087741 after 52
// IPOPT iterations): subject_walk_armless_18musc.osim with the subtalar and toe joints welded, DeGrooteFregly2016 muscles
// without passive fibre force and with implicit tendon dynamics where the tendon is compliant, external loads; coordinates
// low-pass filtered at 6 Hz; time range [0.45, 1.0], mesh interval 0.05 s. The tests evaluate f and g at the shipped solution.
std::unique_ptr<FlatProblem> inverse18Muscles(const std::string& dir, const std::string& options) {
    Problem P;
    P.model = loadOsim(dir + "/subject_walk_armless_18musc.osim", false, /*lockedAsFree=*/true, /*musclesAsDGF=*/true);
    replaceJointsWithWelds(P.model, {"subtalar_r", "subtalar_l", "mtp_r", "mtp_l"});
    for (MuscleSpec& mu : P.model.muscles) {
        mu.ignorePassiveFiberForce = true;
        mu.tendonImplicit = true;
    }
    appendExternalLoads(P.model, dir + "/subject_walk_armless_external_loads.xml");
    const StoTable tab = readSto(dir + "/subject_walk_armless_coordinates.mot");
    auto values = coordinateColumns(P.model, tab);
    std::vector<double> time = tab.time;
    lowpassFilter(time, values, 6.0);
    return finishInverse(P, time, values, /*addReserves=*/false, 0.45, 1.0, 11, options);
}

// MocoInverse (Moco/Moco/MocoInverse.cpp:44-140) on the example2DWalking model: the kinematics of referenceCoordinates.sto are
// prescribed (PositionMotion::createFromTable, Components/PositionMotion.cpp:121-150: GCVSplines of the value columns), reserve
// actuators of optimal force 1 on every coordinate without an actuator (ModelFactory.cpp:265-320, ModOpAddReserves(1.0)),
// MocoControlGoal "excitation_effort" + MocoInitialActivationGoal, implicit dynamics, no control interpolation at the
// midpoints, implicit auxiliary derivatives minimised with weight 0.01, forward differences. The calf and shin muscles get
// compliant tendons in implicit mode (ModOpTendonComplianceDynamicsModeDGF("implicit"), as testMocoInverse.cpp:122-129 does)
// so that the problem carries auxiliary states, derivatives and residuals; the feet's contact spheres supply the ground reactions.
// prescribe = false: the same problem with the coordinates left as states (used by the tests to check the prescribed point
// function against the implicit one evaluated at the prescribed values).
std::unique_ptr<FlatProblem> gaitInverse(const std::string& dir, const std::string& options, bool prescribe = true) {
    Problem P;
    P.model = loadOsim(dir + "/2D_gait.osim");
    for (MuscleSpec& mu : P.model.muscles)
        if (mu.path.find("soleus") != std::string::npos || mu.path.find("gastroc") != std::string::npos ||
            mu.path.find("tib_ant") != std::string::npos) {
            mu.ignoreTendonCompliance = false;
            mu.tendonImplicit = true;
        }
    const StoTable ref = readSto(dir + "/referenceCoordinates.sto");
    for (const JointSpec& J : P.model.joints)
        for (const CoordSpec& c : J.coords) {
            const std::string path = J.container + J.name + "/" + c.name;
            const int ci = ref.col(path + "/value");
            if (ci < 0) throw std::runtime_error("referenceCoordinates.sto lacks " + path + "/value");
            if (prescribe) P.model.positionMotion.emplace_back(path, fitNaturalSpline(ref.time, ref.cols[ci], 5));
            bool has = false;
            for (const CoordActSpec& a : P.model.coordActs) has = has || a.coord == c.name;
            if (has) continue;
            CoordActSpec a;
            std::string n = path;
            for (char& ch : n) if (ch == '/') ch = '_';
            a.path = "/forceset/reserve" + n;
            a.coord = c.name;
            a.optimalForce = 1.0;
            P.model.actuatorOrder.emplace_back(0, (int)P.model.coordActs.size());
            P.model.coordActs.push_back(a);
        }
    P.setTimeBounds(Bounds(ref.time.front()), Bounds(ref.time.back()));
    P.addGoal(MOCOB200_GOAL_CONTROL, "excitation_effort", 1.0);
    P.addGoal(MOCOB200_GOAL_INITIAL_ACTIVATION, "initial_activation");
    SolverSettings S;
    S.num_mesh_intervals = 24; // mesh_interval 0.02 s over 0.47 s
    S.multibody_dynamics_mode = "implicit";
    S.interpolate_control_midpoints = false;
    S.minimize_implicit_auxiliary_derivatives = true;
    S.implicit_auxiliary_derivatives_weight = 0.01;
    S.optim_finite_difference_scheme = "forward";
    S.optim_convergence_tolerance = 1e-3;
    S.optim_constraint_tolerance = 1e-3;
    S.applyOptions(options);
    return flatten(P, S);
}

} // namespace

std::unique_ptr<FlatProblem> buildExample(const std::string& name, const std::string& dataDir, const std::string& options) {
    if (name == "sliding_mass") return slidingMass(options);
    if (name == "hanging_muscle") return hangingMuscle(options, false, false);
    if (name == "hanging_muscle_rigidtendon") return hangingMuscle(options, false, true);
    if (name == "hanging_muscle_noactdyn") return hangingMuscle(options, true, false);
    if (name.rfind("pendulum_control_bound:", 0) == 0) return pendulumControlBound(name.substr(23), options);
    if (name.rfind("pendulum:", 0) == 0) return pendulum(std::stoi(name.substr(9)), options);
    if (name == "gait2d_tracking") return gaitTracking(dataDir, options);
    if (name == "gait2d_prediction") return gaitPrediction(dataDir, options);
    if (name == "gait10dof_replay") return gait10dofReplay(dataDir, options);
    if (name == "rajagopal_replay") return rajagopalReplay(dataDir, options);
    if (name == "rajagopal_replay_loads") return rajagopalReplay(dataDir, options, true);
    if (name == "rajagopal_inverse") return rajagopalInverse(dataDir, options);
    if (name == "moco_inverse_18musc") return inverse18Muscles(dataDir, options);
    if (name == "gait2d_inverse") return gaitInverse(dataDir, options);
    if (name == "gait2d_inverse_free") return gaitInverse(dataDir, options, false);
    if (name == "planar_point_mass_coupler") return planarPointMassCoupler(options);
    if (name == "double_pendulum_coupler") return doublePendulumCoupler(options);
    if (name == "knee_chain") return kneeChain(options);
    throw std::runtime_error("unknown example '" + name + "'");
}

} // namespace mocob200
