// Host-side description of a fixed-topology model + optimal control problem, and
// its flattening into the POD pools of include/mocob200.h.
//
// The classes mirror, name for name, the slice of the reference's surface that the
// BASELINE configs touch (MocoProblem::setTimeBounds/setStateInfo/setControlInfo/addGoal,
// MocoCasADiSolver's properties) so that problem set-up code reads like the reference's
// examples; they hold plain data and never evaluate dynamics.
#pragma once
#include <array>
#include <cmath>
#include <limits>
#include <map>
#include <memory>
#include <string>
#include <vector>

#include "../../include/mocob200.h"

namespace mocob200 {

struct Vec3 {
    double v[3] = {0, 0, 0};
    double& operator[](int i) { return v[i]; }
    double operator[](int i) const { return v[i]; }
};
struct Mat3 { // row-major
    double m[9] = {1, 0, 0, 0, 1, 0, 0, 0, 1};
};
struct Xform { // x_outer = R * x_inner + p
    Mat3 R;
    Vec3 p;
};
Mat3 mul(const Mat3& a, const Mat3& b);
Vec3 mul(const Mat3& a, const Vec3& x);
Xform compose(const Xform& a, const Xform& b); // a∘b
Xform inverse(const Xform& a);
Mat3 rotXYZ(double a, double b, double c); // body-fixed X-Y-Z (PhysicalOffsetFrame orientation)

struct FrameRef { // a frame fixed on a body: body index (-1 = ground) + transform in the body's base frame
    int body = -1;
    Xform X;
};

struct BodySpec {
    std::string name;
    double mass = 0;
    Vec3 com;
    double inertiaCom[6] = {0, 0, 0, 0, 0, 0}; // xx yy zz xy xz yz about the mass centre
};
struct CoordSpec {
    std::string name;
    double lo = -10, hi = 10;
};
struct SplineSpec { // SimmSpline wrapped in a MultiplierFunction
    std::vector<double> x, y;
    double scale = 1;
};
// One TransformAxis of a CustomJoint's SpatialTransform (SimbodyEngine/TransformAxis.cpp, SpatialTransform.cpp:100-176):
// an axis and a scalar function of (at most) one of the joint's coordinates.
struct TransformAxisSpec {
    Vec3 axis;
    int fnKind = MOCOB200_FN_CONSTANT;
    double c0 = 0, c1 = 0;   // linear: c0 x + c1 (LinearFunction coefficients) ; constant: c1
    SplineSpec spline;       // MOCOB200_FN_SIMMSPLINE (scale = product of the enclosing MultiplierFunctions)
    std::string coord;       // empty: constant
};
struct JointSpec {
    std::string name;
    std::string container = "/jointset/"; // "/" when added with Model::addComponent
    int kind = MOCOB200_JOINT_WELD;
    FrameRef parent, child;
    std::vector<CoordSpec> coords;
    TransformAxisSpec axes[6]; // MOCOB200_JOINT_CUSTOM: rotation1..3, translation1..3
};
struct PathPointSpec {
    int kind = MOCOB200_PP_FIXED;
    int body = -1;
    Vec3 location; // already expressed in the body's base frame
    std::string condCoord;
    double range[2] = {0, 0};
    SplineSpec fn[3];
    std::string coord[3];
};
struct MuscleSpec { // DeGrooteFregly2016Muscle properties (Components/DeGrooteFregly2016Muscle.cpp:52-63)
    std::string path; // absolute path, e.g. "/hamstrings_r" or "/forceset/muscle"
    double maxIsometricForce = 1000, optimalFiberLength = 0.1, tendonSlackLength = 0.2;
    double pennationAngleAtOptimal = 0, maxContractionVelocity = 10;
    double activationTimeConstant = 0.015, deactivationTimeConstant = 0.060;
    double defaultActivation = 0.5, defaultNormalizedTendonForce = 0.5;
    double activeForceWidthScale = 1.0, fiberDamping = 0.0;
    double passiveFiberStrainAtOneNormForce = 0.6, tendonStrainAtOneNormForce = 0.049;
    double minControl = 0, maxControl = 1;
    bool ignoreTendonCompliance = false, ignoreActivationDynamics = false, ignorePassiveFiberForce = false;
    bool tendonImplicit = false; // tendon_compliance_dynamics_mode == "implicit"
    // PathActuator (Actuators/PathActuator.cpp:156-197): tension = optimal_force (maxIsometricForce here) x control, no states
    bool pathActuator = false;
    std::vector<PathPointSpec> points;
    struct PathWrapSpec { // Simulation/Wrap/PathWrap.cpp: one (path, wrap object) pairing, in PathWrapSet order
        std::string object;
        int rangeStart = -1, rangeEnd = -1; // 1-based path-point numbers, < 1 = first / last
    };
    std::vector<PathWrapSpec> wraps;
};
struct ContactSpec { // SmoothSphereHalfSpaceForce + its ContactSphere / ContactHalfSpace
    std::string name;
    FrameRef sphere; // X.p = sphere centre
    double radius = 0.01;
    FrameRef halfSpace;
    double stiffness = 1, dissipation = 0, muS = 0, muD = 0, muV = 0, vT = 0.01, cF = 1e-5, bD = 300, bV = 50;
};
struct CoordActSpec {
    std::string path;
    std::string coord;
    double optimalForce = 1, minControl = -std::numeric_limits<double>::infinity(),
           maxControl = std::numeric_limits<double>::infinity();
};

struct PiecewisePoly { // value(t) on [breaks[k], breaks[k+1]) = sum_j coef[k*order+j] (t-breaks[k])^j
    std::vector<double> breaks, coef;
    int order = 0;
    bool isSpline = false; // came from a GCVSpline (domain-checked where the reference checks, e.g. MocoControlBoundConstraint.cpp:74-92)
};
// ExternalForce (Model/ExternalForce.cpp:276-355): force / point / torque as functions of time (GCVSpline degree 3 for
// tables of >= 4 rows, piecewise linear for 2-3 rows, constant for 1), applied to a body.
struct ExternalForceSpec {
    std::string name;
    int appliedTo = -1, forceExpressedIn = -1, pointExpressedIn = -1; // body indices (-1 = ground)
    bool hasForce = false, hasPoint = false, hasTorque = false;
    PiecewisePoly force[3], point[3], torque[3];
};
// CoordinateCouplerConstraint (SimbodyEngine/CoordinateCouplerConstraint.cpp): q_dependent = scale_factor * f(q_independent)
struct CouplerSpec {
    std::string name;
    std::string independent, dependent;
    TransformAxisSpec fn; // kind / c0 / c1 / spline (axis unused)
    double scaleFactor = 1;
};
struct WrapCylinderSpec { // Simulation/Wrap/WrapCylinder.cpp + WrapObject.cpp:97-141 (pose, quadrant)
    std::string name;
    int body = -1;
    Xform X;                 // wrap frame in the body: xyz_body_rotation (body-fixed X-Y-Z), translation
    double radius = 0, length = 0;
    int wrapAxis = 0, wrapSign = 0; // quadrant "x" / "-x" / "y" / "-y" -> axis 0/1, sign +-1; "all" -> sign 0
    bool active = true;
    bool supported = true;   // false: a wrap object of another kind (an error only when a path refers to it)
};
struct ModelSpec {
    std::string name;
    std::vector<WrapCylinderSpec> wrapObjects;
    Vec3 gravity;
    std::vector<BodySpec> bodies;
    std::vector<JointSpec> joints;
    std::vector<MuscleSpec> muscles;
    std::vector<ContactSpec> contacts;
    std::vector<CoordActSpec> coordActs;
    std::vector<ExternalForceSpec> externalForces;
    std::vector<CouplerSpec> couplers;
    // PositionMotion (Components/PositionMotion.cpp): when non-empty, one function of time for EVERY coordinate, keyed by the
    // coordinate's absolute path (createFromTable: GCVSplines of the kinematics table) - all kinematics are prescribed
    std::vector<std::pair<std::string, PiecewisePoly>> positionMotion;
    std::vector<std::pair<int, int>> actuatorOrder; // (0 = coordinate actuator, 1 = muscle ; index) in component order
    ModelSpec() { gravity.v[1] = -9.80665; }
    int bodyIndex(const std::string& n) const;
};

// .osim: document Version 40000 subset and the Version 30000 layout (joints nested in bodies) of the
// opensim-models pipelines. musclesAsPathActuators: every muscle of any class is read as a PathActuator with
// optimal_force 1 on its GeometryPath (its tension becomes the control) - used to replay recorded actuator forces.
// lockedAsFree: read locked coordinates as ordinary ones (enough for replaying recorded states, where they stay constant)
// musclesAsDGF: Millard2012EquilibriumMuscle / Thelen2003Muscle are replaced by DeGrooteFregly2016Muscles the way
// ModOpReplaceMusclesWithDeGrooteFregly2016 does (DeGrooteFregly2016Muscle.cpp:933-1050) - which copies the path POINTS only:
// the replaced muscles lose their PathWraps, in the reference too
ModelSpec loadOsim(const std::string& path, bool musclesAsPathActuators = false, bool lockedAsFree = false, bool musclesAsDGF = false);
// ForceSet file (e.g. reserve actuators) appended to a model: CoordinateActuators only; other classes are returned by name
std::vector<std::string> appendForceSetFile(ModelSpec& M, const std::string& path);
// ExternalLoads .xml (+ its data file, resolved relative to the .xml) -> M.externalForces (ExternalLoads.cpp, ExternalForce.cpp:200-317)
// [tmin, tmax]: only the data rows inside the window are splined (a window that overhangs the times of interest by 100 rows
// changes the interior of an interpolating cubic spline by less than round-off; it keeps the flattened problem small)
void appendExternalLoads(ModelSpec& M, const std::string& xmlPath, double tmin = -1e300, double tmax = 1e300);

// ---- the optimal control problem (MocoProblem / MocoPhase surface) ----
struct Bounds {
    double lo = std::numeric_limits<double>::quiet_NaN(), hi = std::numeric_limits<double>::quiet_NaN();
    Bounds() {}
    Bounds(double v) : lo(v), hi(v) {}
    Bounds(double a, double b) : lo(a), hi(b) {}
    bool isSet() const { return !std::isnan(lo); }
};
struct VariableInfo {
    Bounds bounds, initial, final;
};
// natural interpolating spline of odd degree (what GCVSPL returns for zero smoothing), as a piecewise polynomial
PiecewisePoly fitNaturalSpline(const std::vector<double>& t, const std::vector<double>& y, int degree);
std::vector<PiecewisePoly> fitNaturalSplines(const std::vector<double>& t, const std::vector<const std::vector<double>*>& ys, int degree);
double evalPiecewisePoly(const PiecewisePoly& pp, double t, int deriv = 0);

struct GoalSpec {
    int kind = MOCOB200_GOAL_CONTROL;
    std::string name;
    double weight = 1;
    int mode = -1; // -1: the goal's default mode
    // control goal
    std::map<std::string, double> controlWeights;
    int exponent = 2;
    bool divideByDisplacement = false;
    // tracking goals: (state/control name, weight, reference)
    struct Ref {
        std::string name;
        double weight = 1;
        PiecewisePoly spline;
    };
    std::vector<Ref> refs;
    // contact tracking
    struct ContactGroup {
        std::vector<std::string> contacts;
        std::string appliedToBody;
        PiecewisePoly ref[3];
    };
    std::vector<ContactGroup> groups;
    int projection = 0; // 0 none, 1 vector, 2 plane
    Vec3 projectionVector;
    // periodicity: (initial name, final name, sign)
    struct Pair {
        std::string a, b;
        double sign = 1;
    };
    std::vector<Pair> statePairs, controlPairs;
    // average speed
    double desiredAverageSpeed = 0;
    // endpoint-constraint bounds (default 0,0)
    Bounds ecBounds{0, 0};
    // MocoControlBoundConstraint (kind MOCOB200_PATHCON_CONTROL_BOUND): control_paths, lower_bound / upper_bound
    // functions of time (empty breaks = not set), equality_with_lower (MocoControlBoundConstraint.h)
    std::vector<std::string> controlPaths;
    PiecewisePoly lowerBound, upperBound;
    bool equalityWithLower = false;
    void addControlPath(const std::string& p) { controlPaths.push_back(p); }
};
PiecewisePoly constantFunction(double v);                                                    // OpenSim::Constant
PiecewisePoly piecewiseLinearFunction(const std::vector<double>& x, const std::vector<double>& y); // PiecewiseLinearFunction.cpp:429-463

struct Problem {
    ModelSpec model;
    Bounds timeInitial{0}, timeFinal{1};
    std::map<std::string, VariableInfo> stateInfos, controlInfos;
    Bounds defaultSpeedBounds{-50, 50};
    Bounds multiplierBounds{-1000, 1000}; // MocoProblem::setMultiplierBounds (MocoProblem.h:197; MocoPhase default)
    Bounds kinematicConstraintBounds{0, 0}; // MocoProblem::setKinematicConstraintBounds
    bool boundActivationFromExcitation = true;
    std::vector<GoalSpec> goals;

    void setTimeBounds(Bounds i, Bounds f) { timeInitial = i; timeFinal = f; }
    void setStateInfo(const std::string& n, Bounds b, Bounds i = {}, Bounds f = {}) { stateInfos[n] = {b, i, f}; }
    void setControlInfo(const std::string& n, Bounds b, Bounds i = {}, Bounds f = {}) { controlInfos[n] = {b, i, f}; }
    GoalSpec& addGoal(int kind, const std::string& name, double weight = 1) {
        goals.emplace_back();
        goals.back().kind = kind;
        goals.back().name = name;
        goals.back().weight = weight;
        return goals.back();
    }
    // MocoProblem::addPathConstraint<T>() ; evaluated after the goals' records (their own index space in the reference)
    GoalSpec& addPathConstraint(int kind, const std::string& name) {
        GoalSpec& g = addGoal(kind, name);
        g.mode = MOCOB200_MODE_PATH_CONSTRAINT;
        return g;
    }
};

// MocoCasADiSolver / MocoDirectCollocationSolver properties that shape the NLP.
struct SolverSettings {
    int num_mesh_intervals = 100;
    std::vector<double> mesh;
    std::string transcription_scheme = "hermite-simpson";
    bool interpolate_control_midpoints = true;
    std::string multibody_dynamics_mode = "explicit";
    std::string optim_finite_difference_scheme = "central";
    std::string optim_sparsity_detection = "none";
    Bounds implicit_multibody_acceleration_bounds{-1000, 1000};
    Bounds implicit_auxiliary_derivative_bounds{-1000, 1000};
    bool minimize_implicit_multibody_accelerations = false;
    double implicit_multibody_accelerations_weight = 1.0;
    bool minimize_implicit_auxiliary_derivatives = false;
    double implicit_auxiliary_derivatives_weight = 1.0;
    // IPOPT-side properties are carried but unused by the evaluator
    double optim_convergence_tolerance = -1, optim_constraint_tolerance = -1;
    int optim_max_iterations = -1;
    // kinematic constraints (MocoCasADiSolver.h:115-159, MocoDirectCollocationSolver.h:90-168)
    bool enforce_constraint_derivatives = false;
    bool minimize_lagrange_multipliers = false;
    double lagrange_multiplier_weight = 1.0;
    Bounds velocity_correction_bounds{-0.1, 0.1};
    void applyOptions(const std::string& opts); // "key=value;key=value"
    std::vector<std::string> keysGiven;         // canonical names of the keys applyOptions has seen
    bool given(const std::string& k) const;
};

// Owner of one flattened problem (pools + transcription).
struct FlatProblem {
    std::vector<double> dbl;
    std::vector<int> idx;
    std::vector<double> mesh;
    mocob200_tables tables{};
    mocob200_transcription tr{};
    std::vector<std::string> stateNames, controlNames, derivativeNames, multiplierNames, slackNames;
    void rebind(); // re-point tables.dbl/idx and tr.mesh after a copy/load
};

std::unique_ptr<FlatProblem> flatten(const Problem& p, const SolverSettings& s);
// solver settings -> fp.tr / fp.mesh (all of them, or only the keys applyOptions was given); throws on invalid values
void applySettingsToTranscription(const SolverSettings& s, FlatProblem& fp, bool onlyKeysGiven);
// throws unless every offset / count of the tables lies inside its pool
void validateTables(const mocob200_tables& T);
std::unique_ptr<FlatProblem> buildExample(const std::string& name, const std::string& dataDir, const std::string& options);

// simple .sto/.mot table
struct StoTable {
    std::vector<std::string> labels; // without "time"
    std::vector<double> time;
    std::vector<std::vector<double>> cols; // [label][row]
    bool inDegrees = false;                // header "inDegrees=yes": rotational coordinates are in degrees
    int col(const std::string& l) const;
};
StoTable readSto(const std::string& path);

} // namespace mocob200
