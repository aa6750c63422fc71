// .osim -> ModelSpec, for the component subset of SURVEY.md Appendix E:
// Body; Pin/Slider/Planar/Weld/Custom joints (PhysicalOffsetFrames in document Version 40000, joint-in-body with
// location/orientation pairs in Version 30000); DeGrooteFregly2016Muscle (or any path-carrying force read as a
// PathActuator) with PathPoint / ConditionalPathPoint / MovingPathPoint(SimmSpline in MultiplierFunction);
// ContactSphere / ContactHalfSpace / SmoothSphereHalfSpaceForce; CoordinateActuator; ExternalLoads files.
// Anything else that would change the dynamics is refused loudly (no silent approximation).
#include <algorithm>
#include <cmath>
#include <cstdlib>
#include <limits>
#include <stdexcept>

#include "host_model.h"
#include "xml_reader.h"

namespace mocob200 {
namespace {

Vec3 vec3Of(const std::string& s, const char* what) {
    auto v = parseDoubles(s);
    if (v.size() != 3) throw std::runtime_error(std::string("expected 3 numbers for ") + what);
    Vec3 r;
    for (int i = 0; i < 3; ++i) r.v[i] = v[i];
    return r;
}
bool boolOf(const std::string& s, bool def) {
    if (s.empty()) return def;
    return s == "true" || s == "1" || s == "True";
}
double numOf(const std::string& s) {
    if (s == "Inf" || s == "inf" || s == "+Inf" || s == "infinity") return std::numeric_limits<double>::infinity();
    if (s == "-Inf" || s == "-inf" || s == "-infinity") return -std::numeric_limits<double>::infinity();
    return std::strtod(s.c_str(), nullptr);
}
double dblOf(const XmlNode& n, const char* tag, double def) {
    std::string s = n.childText(tag);
    return s.empty() ? def : numOf(s);
}
std::string leaf(const std::string& path) {
    size_t sl = path.rfind('/');
    return sl == std::string::npos ? path : path.substr(sl + 1);
}

struct Ctx {
    ModelSpec& M;
    bool v3 = false; // document Version < 40000: frames are bodies referenced by bare name
    bool asPathActuators = false;
    bool lockedAsFree = false;
    bool musclesAsDGF = false;
    // resolve "/bodyset/x", "/ground", "ground", and (Version 30000) a bare body name
    bool baseFrame(const std::string& path, int& body) const {
        if (path == "/ground" || path == "ground") { body = -1; return true; }
        const std::string pre = "/bodyset/";
        if (path.compare(0, pre.size(), pre) == 0 && path.find('/', pre.size()) == std::string::npos) {
            body = M.bodyIndex(path.substr(pre.size()));
            return true;
        }
        if (v3 && !path.empty() && path.find('/') == std::string::npos) {
            body = M.bodyIndex(path);
            return true;
        }
        return false;
    }
    FrameRef offsetFrame(const XmlNode& f) const {
        FrameRef r;
        std::string parent = f.childText("socket_parent");
        if (!baseFrame(parent, r.body))
            throw std::runtime_error("PhysicalOffsetFrame '" + f.name() + "': parent '" + parent + "' is not a body or ground");
        Vec3 t = vec3Of(f.childText("translation", "0 0 0"), "translation");
        Vec3 o = vec3Of(f.childText("orientation", "0 0 0"), "orientation");
        r.X.R = rotXYZ(o.v[0], o.v[1], o.v[2]);
        r.X.p = t;
        return r;
    }
    FrameRef jointFrame(const XmlNode& joint, const std::string& socket) const {
        FrameRef r;
        if (baseFrame(socket, r.body)) return r;
        const XmlNode* frames = joint.child("frames");
        const std::string want = leaf(socket);
        if (frames)
            for (auto& f : frames->children)
                if (f->tag == "PhysicalOffsetFrame" && f->name() == want) return offsetFrame(*f);
        throw std::runtime_error("joint '" + joint.name() + "': cannot resolve frame '" + socket + "'");
    }
};

// <holder> [<MultiplierFunction> <function>] <SimmSpline | LinearFunction | Constant> ... ; the scales of nested
// MultiplierFunctions multiply (Common/MultiplierFunction.cpp)
const XmlNode* unwrapFunction(const XmlNode& holder, double& scale) {
    const XmlNode* cur = &holder;
    scale = 1.0;
    for (;;) {
        if (const XmlNode* mf = cur->child("MultiplierFunction")) {
            scale *= dblOf(*mf, "scale", 1.0);
            const XmlNode* inner = mf->child("function");
            if (!inner) throw std::runtime_error("MultiplierFunction without a function");
            cur = inner;
            continue;
        }
        for (auto& c : cur->children)
            if (c->tag == "SimmSpline" || c->tag == "LinearFunction" || c->tag == "Constant") return c.get();
        return nullptr;
    }
}

SplineSpec simmSplineOf(const XmlNode& fn, double scale, const std::string& what) {
    SplineSpec s;
    s.scale = scale;
    s.x = parseDoubles(fn.childText("x"));
    s.y = parseDoubles(fn.childText("y"));
    if (s.x.size() != s.y.size() || s.x.size() < 2) throw std::runtime_error(what + ": bad SimmSpline knots");
    return s;
}

SplineSpec splineOf(const XmlNode& holder, const std::string& what) {
    double scale = 1;
    const XmlNode* fn = unwrapFunction(holder, scale);
    if (!fn || fn->tag != "SimmSpline")
        throw std::runtime_error(what + ": only SimmSpline (optionally inside a MultiplierFunction) is supported");
    return simmSplineOf(*fn, scale, what);
}

// SpatialTransform of a CustomJoint: six TransformAxis children in order rotation1..3, translation1..3
// (SpatialTransform.cpp:100-176). An axis without a function but with one coordinate gets LinearFunction(1, 0), without
// coordinates Constant(0) (SpatialTransform::connectToJoint).
void readSpatialTransform(const XmlNode& joint, JointSpec& J) {
    const XmlNode* st = joint.child("SpatialTransform");
    if (!st) throw std::runtime_error("CustomJoint '" + J.name + "' has no SpatialTransform");
    int k = 0;
    for (auto& ta : st->children) {
        if (ta->tag != "TransformAxis") continue;
        if (k >= 6) throw std::runtime_error("CustomJoint '" + J.name + "': more than 6 TransformAxis entries");
        TransformAxisSpec& A = J.axes[k++];
        A.axis = vec3Of(ta->childText("axis"), "TransformAxis axis");
        std::vector<std::string> coords;
        {
            std::string cs = ta->childText("coordinates");
            size_t i = 0;
            while (i < cs.size()) {
                while (i < cs.size() && isspace((unsigned char)cs[i])) ++i;
                size_t j = i;
                while (j < cs.size() && !isspace((unsigned char)cs[j])) ++j;
                if (j > i) coords.push_back(cs.substr(i, j - i));
                i = j;
            }
        }
        if (coords.size() > 1) throw std::runtime_error("CustomJoint '" + J.name + "': multi-coordinate TransformAxis functions are unsupported");
        if (!coords.empty()) A.coord = coords[0];
        double scale = 1;
        const XmlNode* holder = ta->child("function");
        const XmlNode* fn = holder ? unwrapFunction(*holder, scale) : nullptr;
        if (!holder || holder->children.empty()) {
            if (A.coord.empty()) { A.fnKind = MOCOB200_FN_CONSTANT; A.c1 = 0; }
            else { A.fnKind = MOCOB200_FN_LINEAR; A.c0 = 1; A.c1 = 0; }
        } else if (!fn) {
            throw std::runtime_error("CustomJoint '" + J.name + "': unsupported TransformAxis function (SimmSpline, LinearFunction, Constant)");
        } else if (fn->tag == "Constant") {
            A.fnKind = MOCOB200_FN_CONSTANT;
            A.c1 = scale * dblOf(*fn, "value", 0.0);
            A.coord.clear();
        } else if (fn->tag == "LinearFunction") {
            auto c = parseDoubles(fn->childText("coefficients", "1 0"));
            if (c.size() != 2) throw std::runtime_error("LinearFunction needs 2 coefficients");
            if (A.coord.empty()) { A.fnKind = MOCOB200_FN_CONSTANT; A.c1 = scale * c[1]; } // SimTK::Function::Linear of no argument
            else { A.fnKind = MOCOB200_FN_LINEAR; A.c0 = scale * c[0]; A.c1 = scale * c[1]; }
        } else {
            if (A.coord.empty()) throw std::runtime_error("CustomJoint '" + J.name + "': SimmSpline axis without a coordinate");
            A.fnKind = MOCOB200_FN_SIMMSPLINE;
            A.spline = simmSplineOf(*fn, scale, "CustomJoint '" + J.name + "'");
        }
        const double n2 = A.axis.v[0] * A.axis.v[0] + A.axis.v[1] * A.axis.v[1] + A.axis.v[2] * A.axis.v[2];
        if (!(n2 > 0)) throw std::runtime_error("CustomJoint '" + J.name + "': zero TransformAxis axis");
        const double inv = 1.0 / std::sqrt(n2); // SimTK::UnitVec3 normalises
        for (int a = 0; a < 3; ++a) A.axis.v[a] *= inv;
    }
    if (k != 6) throw std::runtime_error("CustomJoint '" + J.name + "': SpatialTransform needs 6 TransformAxis entries");
    // the mobilizer's coordinate i is the i-th distinct coordinate name met in the axes (SpatialTransform::getCoordinateNames,
    // CustomJoint.cpp:150-160); the joint's own Coordinate list must be in that order for the names to mean what they say
    std::vector<std::string> order;
    for (auto& A : J.axes)
        if (!A.coord.empty() && std::find(order.begin(), order.end(), A.coord) == order.end()) order.push_back(A.coord);
    if (order.size() != J.coords.size()) throw std::runtime_error("CustomJoint '" + J.name + "': coordinates and SpatialTransform disagree");
    for (size_t i = 0; i < order.size(); ++i)
        if (order[i] != J.coords[i].name)
            throw std::runtime_error("CustomJoint '" + J.name + "': coordinate order differs from the SpatialTransform's (unsupported)");
}

int jointKindOf(const std::string& tag, const std::string& name) {
    if (tag == "PinJoint") return MOCOB200_JOINT_PIN;
    if (tag == "SliderJoint") return MOCOB200_JOINT_SLIDER;
    if (tag == "PlanarJoint") return MOCOB200_JOINT_PLANAR;
    if (tag == "WeldJoint") return MOCOB200_JOINT_WELD;
    if (tag == "CustomJoint") return MOCOB200_JOINT_CUSTOM;
    if (tag == "UniversalJoint") return MOCOB200_JOINT_CUSTOM; // two rotations, about X then about the rotated Y (universalAxes below)
    throw std::runtime_error("unsupported joint type '" + tag + "' (" + name + ")");
}
// SimTK::MobilizedBody::Universal (UniversalJoint.cpp:60-80): rotation about the parent frame's X by the first coordinate, then
// about the rotated Y by the second - the function-based mobilizer with axes X, Y (linear) and everything else constant
void universalAxes(JointSpec& J) {
    if (J.coords.size() != 2) throw std::runtime_error("UniversalJoint '" + J.name + "' needs two coordinates");
    const double ax[6][3] = {{1, 0, 0}, {0, 1, 0}, {0, 0, 1}, {1, 0, 0}, {0, 1, 0}, {0, 0, 1}};
    for (int k = 0; k < 6; ++k) {
        TransformAxisSpec& A = J.axes[k];
        for (int a = 0; a < 3; ++a) A.axis.v[a] = ax[k][a];
        if (k < 2) { A.fnKind = MOCOB200_FN_LINEAR; A.c0 = 1; A.c1 = 0; A.coord = J.coords[k].name; }
        else { A.fnKind = MOCOB200_FN_CONSTANT; A.c1 = 0; A.coord.clear(); }
    }
}
bool g_lockedAsFree = false; // set for the duration of one loadOsim call
void readCoordinates(const XmlNode* holder, JointSpec& J) {
    if (!holder) return;
    for (auto& c : holder->children) {
        if (c->tag != "Coordinate") continue;
        CoordSpec C;
        C.name = c->name();
        auto r = parseDoubles(c->childText("range", "-10 10"));
        if (r.size() == 2) { C.lo = r[0]; C.hi = r[1]; }
        if ((boolOf(c->childText("locked"), false) && !g_lockedAsFree) || boolOf(c->childText("prescribed"), false))
            throw std::runtime_error("coordinate '" + C.name + "': locked/prescribed coordinates are unsupported");
        J.coords.push_back(C);
    }
}

void readPath(const Ctx& cx, const XmlNode& n, MuscleSpec& mu, bool dropWraps = false) {
    const XmlNode* gp = n.child("GeometryPath");
    if (!gp) throw std::runtime_error("'" + mu.path + "' has no GeometryPath");
    if (const XmlNode* ws = dropWraps ? nullptr : gp->child("PathWrapSet"))
        if (const XmlNode* wo = ws->child("objects"))
            for (auto& w : wo->children) {
                if (w->tag != "PathWrap") continue;
                MuscleSpec::PathWrapSpec pw;
                pw.object = trim(w->childText("wrap_object"));
                auto r = parseDoubles(w->childText("range", "-1 -1"));
                if (r.size() == 2) { pw.rangeStart = (int)r[0]; pw.rangeEnd = (int)r[1]; }
                const std::string method = trim(w->childText("method", "hybrid"));
                // the method only matters for ellipsoids (WrapEllipsoid.cpp); cylinders ignore it
                bool found = false;
                for (const WrapCylinderSpec& o : cx.M.wrapObjects)
                    if (o.name == pw.object) {
                        found = true;
                        if (!o.supported)
                            throw std::runtime_error("muscle '" + mu.path + "' wraps over '" + pw.object + "': only WrapCylinder is supported (SURVEY §8 a18)");
                    }
                if (!found) throw std::runtime_error("muscle '" + mu.path + "': no wrap object named '" + pw.object + "'");
                mu.wraps.push_back(pw);
            }
    const XmlNode* pps = gp->child("PathPointSet");
    const XmlNode* objs = pps ? pps->child("objects") : nullptr;
    if (!objs) throw std::runtime_error("muscle '" + mu.path + "' has no path points");
    for (auto& pn : objs->children) {
        PathPointSpec pt;
        // (files between Version 30000 and 40000 mix the two layouts: take whichever the point carries)
        std::string frame = pn->childText("socket_parent_frame");
        if (frame.empty()) frame = pn->childText("body");
        if (!cx.baseFrame(frame, pt.body))
            throw std::runtime_error("path point '" + pn->name() + "': parent frame '" + frame + "' must be a body");
        if (pn->tag == "PathPoint") {
            pt.kind = MOCOB200_PP_FIXED;
            pt.location = vec3Of(pn->childText("location"), "location");
        } else if (pn->tag == "ConditionalPathPoint") {
            pt.kind = MOCOB200_PP_CONDITIONAL;
            pt.location = vec3Of(pn->childText("location"), "location");
            auto r = parseDoubles(pn->childText("range"));
            if (r.size() != 2) throw std::runtime_error("ConditionalPathPoint '" + pn->name() + "': bad range");
            pt.range[0] = r[0];
            pt.range[1] = r[1];
            pt.condCoord = leaf(cx.v3 ? pn->childText("coordinate") : pn->childText("socket_coordinate"));
        } else if (pn->tag == "MovingPathPoint") {
            pt.kind = MOCOB200_PP_MOVING;
            const char* ax[3] = {"x", "y", "z"};
            for (int a = 0; a < 3; ++a) {
                const XmlNode* loc = pn->child(std::string(ax[a]) + "_location");
                if (!loc) throw std::runtime_error("MovingPathPoint '" + pn->name() + "': missing location function");
                pt.fn[a] = splineOf(*loc, "MovingPathPoint '" + pn->name() + "'");
                pt.coord[a] = leaf(cx.v3 ? pn->childText(std::string(ax[a]) + "_coordinate")
                                         : pn->childText(std::string("socket_") + ax[a] + "_coordinate"));
            }
        } else
            throw std::runtime_error("unsupported path point type '" + pn->tag + "'");
        mu.points.push_back(pt);
    }
}

void readMuscle(const Ctx& cx, const XmlNode& n, const std::string& prefix) {
    MuscleSpec mu;
    mu.path = prefix + n.name();
    mu.minControl = dblOf(n, "min_control", 0.0);
    mu.maxControl = dblOf(n, "max_control", 1.0);
    mu.maxIsometricForce = dblOf(n, "max_isometric_force", 1000.0);
    mu.optimalFiberLength = dblOf(n, "optimal_fiber_length", 0.1);
    mu.tendonSlackLength = dblOf(n, "tendon_slack_length", 0.2);
    mu.pennationAngleAtOptimal = dblOf(n, "pennation_angle_at_optimal", 0.0);
    mu.maxContractionVelocity = dblOf(n, "max_contraction_velocity", 10.0);
    mu.ignoreTendonCompliance = boolOf(n.childText("ignore_tendon_compliance"), false);
    mu.ignoreActivationDynamics = boolOf(n.childText("ignore_activation_dynamics"), false);
    mu.activationTimeConstant = dblOf(n, "activation_time_constant", 0.015);
    mu.deactivationTimeConstant = dblOf(n, "deactivation_time_constant", 0.060);
    mu.defaultActivation = dblOf(n, "default_activation", 0.5);
    mu.defaultNormalizedTendonForce = dblOf(n, "default_normalized_tendon_force", 0.5);
    mu.activeForceWidthScale = dblOf(n, "active_force_width_scale", 1.0);
    mu.fiberDamping = dblOf(n, "fiber_damping", 0.0);
    mu.passiveFiberStrainAtOneNormForce = dblOf(n, "passive_fiber_strain_at_one_norm_force", 0.6);
    mu.tendonStrainAtOneNormForce = dblOf(n, "tendon_strain_at_one_norm_force", 0.049);
    mu.ignorePassiveFiberForce = boolOf(n.childText("ignore_passive_fiber_force"), false);
    mu.tendonImplicit = n.childText("tendon_compliance_dynamics_mode", "explicit") == "implicit";
    readPath(cx, n, mu);
    cx.M.actuatorOrder.emplace_back(1, (int)cx.M.muscles.size());
    cx.M.muscles.push_back(std::move(mu));
}

// DeGrooteFregly2016Muscle::replaceMuscles (DeGrooteFregly2016Muscle.cpp:933-1050): what a Millard2012EquilibriumMuscle or a
// Thelen2003Muscle becomes under ModOpReplaceMusclesWithDeGrooteFregly2016. Properties absent from the file take the source
// class's defaults (Millard2012EquilibriumMuscle.cpp:58-75, Thelen2003Muscle.cpp:62-80). Only the path points are copied.
void readMuscleAsDGF(const Ctx& cx, const XmlNode& n, const std::string& prefix) {
    const bool millard = n.tag == "Millard2012EquilibriumMuscle";
    MuscleSpec mu;
    mu.path = prefix + n.name();
    mu.minControl = dblOf(n, "min_control", 0.0); // Muscle::constructProperties
    mu.maxControl = dblOf(n, "max_control", 1.0);
    mu.maxIsometricForce = dblOf(n, "max_isometric_force", 1000.0);
    mu.optimalFiberLength = dblOf(n, "optimal_fiber_length", 0.1);
    mu.tendonSlackLength = dblOf(n, "tendon_slack_length", 0.2);
    mu.pennationAngleAtOptimal = dblOf(n, "pennation_angle_at_optimal", 0.0);
    mu.maxContractionVelocity = dblOf(n, "max_contraction_velocity", 10.0);
    mu.ignoreTendonCompliance = boolOf(n.childText("ignore_tendon_compliance"), false);
    mu.ignoreActivationDynamics = boolOf(n.childText("ignore_activation_dynamics"), false);
    mu.defaultActivation = dblOf(n, "default_activation", 0.05);
    mu.activationTimeConstant = dblOf(n, "activation_time_constant", millard ? 0.010 : 0.015);
    mu.deactivationTimeConstant = dblOf(n, "deactivation_time_constant", millard ? 0.040 : 0.050);
    if (millard) {
        mu.fiberDamping = dblOf(n, "fiber_damping", 0.1);
        auto curveStrain = [&](const char* curve, double dflt) {
            if (const XmlNode* c = n.child(curve)) {
                const XmlNode* inner = c->children.empty() ? c : c->children[0].get();
                if (inner->has("strain_at_one_norm_force")) return dblOf(*inner, "strain_at_one_norm_force", dflt);
                if (c->has("strain_at_one_norm_force")) return dblOf(*c, "strain_at_one_norm_force", dflt);
            }
            return dflt;
        };
        mu.passiveFiberStrainAtOneNormForce = curveStrain("FiberForceLengthCurve", 0.7);
        mu.tendonStrainAtOneNormForce = curveStrain("TendonForceLengthCurve", 0.049);
    } else {
        mu.fiberDamping = 0.0;
        mu.passiveFiberStrainAtOneNormForce = dblOf(n, "FmaxMuscleStrain", 0.6);
        mu.tendonStrainAtOneNormForce = dblOf(n, "FmaxTendonStrain", 0.033);
    }
    readPath(cx, n, mu, /*dropWraps=*/true);
    cx.M.actuatorOrder.emplace_back(1, (int)cx.M.muscles.size());
    cx.M.muscles.push_back(std::move(mu));
}

// PathActuator (Actuators/PathActuator.cpp:156-197), or — with musclesAsPathActuators — any force that carries a GeometryPath
void readPathActuator(const Ctx& cx, const XmlNode& n, const std::string& prefix, bool isPathActuator) {
    MuscleSpec mu;
    mu.path = prefix + n.name();
    mu.pathActuator = true;
    mu.ignoreActivationDynamics = true;
    mu.ignoreTendonCompliance = true;
    mu.maxIsometricForce = isPathActuator ? dblOf(n, "optimal_force", 1.0) : 1.0;
    mu.minControl = isPathActuator ? dblOf(n, "min_control", -std::numeric_limits<double>::infinity()) : -std::numeric_limits<double>::infinity();
    mu.maxControl = isPathActuator ? dblOf(n, "max_control", std::numeric_limits<double>::infinity()) : std::numeric_limits<double>::infinity();
    readPath(cx, n, mu);
    cx.M.actuatorOrder.emplace_back(1, (int)cx.M.muscles.size());
    cx.M.muscles.push_back(std::move(mu));
}

void readCoordinateActuator(const Ctx& cx, const XmlNode& n, const std::string& prefix) {
    CoordActSpec a;
    a.path = prefix + n.name();
    a.coord = n.childText("coordinate");
    a.optimalForce = dblOf(n, "optimal_force", 1.0);
    a.minControl = dblOf(n, "min_control", a.minControl);
    a.maxControl = dblOf(n, "max_control", a.maxControl);
    cx.M.actuatorOrder.emplace_back(0, (int)cx.M.coordActs.size());
    cx.M.coordActs.push_back(a);
}

void readForceLike(const Ctx& cx, const XmlNode& n, const std::string& prefix, const XmlNode* contactGeom) {
    if (boolOf(n.childText("isDisabled"), false) || !boolOf(n.childText("appliesForce"), true)) return;
    if (cx.asPathActuators && n.child("GeometryPath")) {
        readPathActuator(cx, n, prefix, n.tag == "PathActuator");
    } else if (cx.musclesAsDGF && (n.tag == "Millard2012EquilibriumMuscle" || n.tag == "Thelen2003Muscle")) {
        readMuscleAsDGF(cx, n, prefix);
    } else if (n.tag == "DeGrooteFregly2016Muscle") {
        readMuscle(cx, n, prefix);
    } else if (n.tag == "PathActuator") {
        readPathActuator(cx, n, prefix, true);
    } else if (n.tag == "CoordinateActuator") {
        readCoordinateActuator(cx, n, prefix);
    } else if (n.tag == "SmoothSphereHalfSpaceForce") {
        if (!contactGeom) throw std::runtime_error("SmoothSphereHalfSpaceForce without a ContactGeometrySet");
        ContactSpec c;
        c.name = n.name();
        auto geom = [&](const std::string& socket, const char* tag) -> const XmlNode& {
            const std::string want = leaf(socket);
            for (auto& g : contactGeom->children)
                if (g->tag == tag && g->name() == want) return *g;
            throw std::runtime_error("contact '" + c.name + "': no " + tag + " named '" + want + "'");
        };
        const XmlNode& sp = geom(n.childText("socket_sphere"), "ContactSphere");
        const XmlNode& hs = geom(n.childText("socket_half_space"), "ContactHalfSpace");
        if (!cx.baseFrame(sp.childText("socket_frame"), c.sphere.body) || !cx.baseFrame(hs.childText("socket_frame"), c.halfSpace.body))
            throw std::runtime_error("contact '" + c.name + "': geometry frames must be bodies or ground");
        c.sphere.X.p = vec3Of(sp.childText("location", "0 0 0"), "location");
        c.radius = dblOf(sp, "radius", 0.0);
        Vec3 o = vec3Of(hs.childText("orientation", "0 0 0"), "orientation");
        c.halfSpace.X.R = rotXYZ(o.v[0], o.v[1], o.v[2]);
        c.halfSpace.X.p = vec3Of(hs.childText("location", "0 0 0"), "location");
        c.stiffness = dblOf(n, "stiffness", 1.0);
        c.dissipation = dblOf(n, "dissipation", 0.0);
        c.muS = dblOf(n, "static_friction", 0.0);
        c.muD = dblOf(n, "dynamic_friction", 0.0);
        c.muV = dblOf(n, "viscous_friction", 0.0);
        c.vT = dblOf(n, "transition_velocity", 0.01);
        c.cF = dblOf(n, "constant_contact_force", 1e-5); // older files call it derivative_smoothing: ignored, default applies
        c.bD = dblOf(n, "hertz_smoothing", 300.0);
        c.bV = dblOf(n, "hunt_crossley_smoothing", 50.0);
        cx.M.contacts.push_back(c);
    } else {
        throw std::runtime_error("unsupported force/component type '" + n.tag + "' (" + n.name() + ")");
    }
}

} // namespace

ModelSpec loadOsim(const std::string& path, bool musclesAsPathActuators, bool lockedAsFree, bool musclesAsDGF) {
    struct Restore { bool old; ~Restore() { g_lockedAsFree = old; } } restore{g_lockedAsFree};
    g_lockedAsFree = lockedAsFree;
    auto doc = parseXmlFile(path);
    if (doc->tag != "OpenSimDocument") throw std::runtime_error(path + ": not an OpenSimDocument");
    const std::string* ver = doc->attr("Version");
    const int version = ver ? std::atoi(ver->c_str()) : 0;
    if (version < 30000) throw std::runtime_error(path + ": only document Version >= 30000 is supported");
    const XmlNode* model = doc->child("Model");
    if (!model) throw std::runtime_error(path + ": no <Model>");
    ModelSpec M;
    M.name = model->name();
    if (model->has("gravity")) M.gravity = vec3Of(model->childText("gravity"), "gravity");
    Ctx cx{M};
    cx.v3 = version < 40000;
    cx.asPathActuators = musclesAsPathActuators;
    cx.musclesAsDGF = musclesAsDGF;

    const XmlNode* bs = model->child("BodySet");
    const XmlNode* bobjs = bs ? bs->child("objects") : nullptr;
    if (!bobjs) throw std::runtime_error(path + ": no BodySet");
    std::vector<const XmlNode*> bodyNodes;
    for (auto& b : bobjs->children) {
        if (b->tag != "Body") throw std::runtime_error("unsupported BodySet entry '" + b->tag + "'");
        if (cx.v3 && b->name() == "ground") continue; // Version 30000 lists ground as a body
        BodySpec B;
        B.name = b->name();
        B.mass = dblOf(*b, "mass", 0.0);
        B.com = vec3Of(b->childText("mass_center", "0 0 0"), "mass_center");
        if (cx.v3) {
            const char* tags[6] = {"inertia_xx", "inertia_yy", "inertia_zz", "inertia_xy", "inertia_xz", "inertia_yz"};
            for (int i = 0; i < 6; ++i) B.inertiaCom[i] = dblOf(*b, tags[i], 0.0);
        } else {
            auto in = parseDoubles(b->childText("inertia", "0 0 0 0 0 0"));
            if (in.size() != 6) throw std::runtime_error("body '" + B.name + "': inertia needs 6 numbers");
            for (int i = 0; i < 6; ++i) B.inertiaCom[i] = in[i];
        }
        if (const XmlNode* ws = b->child("WrapObjectSet"))
            if (const XmlNode* wo = ws->child("objects"))
                for (auto& w : wo->children) {
                    WrapCylinderSpec W;
                    W.name = w->name();
                    W.body = (int)M.bodies.size();
                    W.supported = w->tag == "WrapCylinder";
                    const Vec3 rot = vec3Of(w->childText("xyz_body_rotation", "0 0 0"), "xyz_body_rotation");
                    W.X.R = rotXYZ(rot.v[0], rot.v[1], rot.v[2]);
                    W.X.p = vec3Of(w->childText("translation", "0 0 0"), "translation");
                    W.active = boolOf(w->childText("active"), true);
                    W.radius = dblOf(*w, "radius", 0.0);
                    W.length = dblOf(*w, "length", 0.0);
                    const std::string qd = trim(w->childText("quadrant", "all"));
                    if (qd == "x" || qd == "+x" || qd == "X" || qd == "+X") { W.wrapAxis = 0; W.wrapSign = 1; }
                    else if (qd == "-x" || qd == "-X") { W.wrapAxis = 0; W.wrapSign = -1; }
                    else if (qd == "y" || qd == "+y" || qd == "Y" || qd == "+Y") { W.wrapAxis = 1; W.wrapSign = 1; }
                    else if (qd == "-y" || qd == "-Y") { W.wrapAxis = 1; W.wrapSign = -1; }
                    else if (qd == "z" || qd == "+z" || qd == "Z" || qd == "+Z") { W.wrapAxis = 2; W.wrapSign = 1; }
                    else if (qd == "-z" || qd == "-Z") { W.wrapAxis = 2; W.wrapSign = -1; }
                    else if (qd == "all" || qd == "ALL" || qd == "All" || qd == "Unassigned") { W.wrapAxis = 0; W.wrapSign = 0; }
                    else throw std::runtime_error("wrap object '" + W.name + "': bad quadrant '" + qd + "'");
                    M.wrapObjects.push_back(W);
                }
        M.bodies.push_back(B);
        bodyNodes.push_back(b.get());
    }

    if (cx.v3) {
        // Version 30000: every body owns the joint that connects it to its parent (Joint.cpp updateFromXMLNode, 30500 conversion):
        // parent frame = (location_in_parent, orientation_in_parent) on parent_body, child frame = (location, orientation)
        for (size_t bi = 0; bi < bodyNodes.size(); ++bi) {
            const XmlNode* jh = bodyNodes[bi]->child("Joint");
            if (!jh || jh->children.empty()) throw std::runtime_error("body '" + M.bodies[bi].name + "' has no joint");
            const XmlNode* j = jh->children[0].get();
            JointSpec J;
            J.name = j->name();
            J.kind = jointKindOf(j->tag, J.name);
            if (boolOf(j->childText("reverse"), false)) throw std::runtime_error("joint '" + J.name + "': reversed joints are unsupported");
            if (!cx.baseFrame(j->childText("parent_body"), J.parent.body))
                throw std::runtime_error("joint '" + J.name + "': unknown parent_body");
            J.child.body = (int)bi;
            Vec3 o = vec3Of(j->childText("orientation_in_parent", "0 0 0"), "orientation_in_parent");
            J.parent.X.R = rotXYZ(o.v[0], o.v[1], o.v[2]);
            J.parent.X.p = vec3Of(j->childText("location_in_parent", "0 0 0"), "location_in_parent");
            o = vec3Of(j->childText("orientation", "0 0 0"), "orientation");
            J.child.X.R = rotXYZ(o.v[0], o.v[1], o.v[2]);
            J.child.X.p = vec3Of(j->childText("location", "0 0 0"), "location");
            if (const XmlNode* cs = j->child("CoordinateSet")) readCoordinates(cs->child("objects"), J);
            if (j->tag == "UniversalJoint") universalAxes(J);
            else if (J.kind == MOCOB200_JOINT_CUSTOM) readSpatialTransform(*j, J);
            M.joints.push_back(J);
        }
    } else {
        const XmlNode* js = model->child("JointSet");
        const XmlNode* jobjs = js ? js->child("objects") : nullptr;
        if (!jobjs) throw std::runtime_error(path + ": no JointSet");
        for (auto& j : jobjs->children) {
            JointSpec J;
            J.name = j->name();
            J.kind = jointKindOf(j->tag, J.name);
            J.parent = cx.jointFrame(*j, j->childText("socket_parent_frame"));
            J.child = cx.jointFrame(*j, j->childText("socket_child_frame"));
            readCoordinates(j->child("coordinates"), J);
            if (j->tag == "UniversalJoint") universalAxes(J);
            else if (J.kind == MOCOB200_JOINT_CUSTOM) readSpatialTransform(*j, J);
            M.joints.push_back(J);
        }
    }
    if (const XmlNode* cs = model->child("ConstraintSet"))
        if (const XmlNode* o = cs->child("objects"))
            for (auto& c : o->children) {
                if (boolOf(c->childText("isDisabled"), false) || !boolOf(c->childText("isEnforced"), true)) continue;
                if (c->tag != "CoordinateCouplerConstraint")
                    throw std::runtime_error("unsupported kinematic constraint '" + c->tag + "' (only CoordinateCouplerConstraint, SURVEY §8 a9)");
                CouplerSpec C;
                C.name = c->name();
                C.dependent = trim(c->childText("dependent_coordinate_name"));
                {
                    std::string names = c->childText("independent_coordinate_names");
                    std::vector<std::string> v;
                    size_t i = 0;
                    while (i < names.size()) {
                        while (i < names.size() && isspace((unsigned char)names[i])) ++i;
                        size_t j = i;
                        while (j < names.size() && !isspace((unsigned char)names[j])) ++j;
                        if (j > i) v.push_back(names.substr(i, j - i));
                        i = j;
                    }
                    if (v.size() != 1) throw std::runtime_error("CoordinateCouplerConstraint '" + C.name + "': exactly one independent coordinate is supported");
                    C.independent = v[0];
                }
                C.scaleFactor = dblOf(*c, "scale_factor", 1.0);
                const XmlNode* holder = c->child("coupled_coordinates_function");
                double scale = 1;
                const XmlNode* fn = holder ? unwrapFunction(*holder, scale) : nullptr;
                if (!fn) throw std::runtime_error("CoordinateCouplerConstraint '" + C.name + "': unsupported coupled_coordinates_function");
                if (fn->tag == "Constant") { C.fn.fnKind = MOCOB200_FN_CONSTANT; C.fn.c1 = scale * dblOf(*fn, "value", 0.0); }
                else if (fn->tag == "LinearFunction") {
                    auto co = parseDoubles(fn->childText("coefficients", "1 0"));
                    if (co.size() != 2) throw std::runtime_error("LinearFunction needs 2 coefficients");
                    C.fn.fnKind = MOCOB200_FN_LINEAR; C.fn.c0 = scale * co[0]; C.fn.c1 = scale * co[1];
                } else { C.fn.fnKind = MOCOB200_FN_SIMMSPLINE; C.fn.spline = simmSplineOf(*fn, scale, "CoordinateCouplerConstraint '" + C.name + "'"); }
                M.couplers.push_back(C);
            }

    const XmlNode* cg = model->child("ContactGeometrySet");
    const XmlNode* cgobjs = cg ? cg->child("objects") : nullptr;
    if (const XmlNode* comps = model->child("components"))
        for (auto& c : comps->children) readForceLike(cx, *c, "/", cgobjs);
    if (const XmlNode* fs = model->child("ForceSet"))
        if (const XmlNode* o = fs->child("objects"))
            for (auto& c : o->children) readForceLike(cx, *c, "/forceset/", cgobjs);
    return M;
}

// A ForceSet file appended to the model (AbstractTool force_set_files, e.g. reserve actuators). CoordinateActuators are
// added; the names of entries of other classes (PointActuator, TorqueActuator residuals ...) are returned to the caller.
std::vector<std::string> appendForceSetFile(ModelSpec& M, const std::string& path) {
    auto doc = parseXmlFile(path);
    const XmlNode* fs = doc->child("ForceSet");
    const XmlNode* objs = fs ? fs->child("objects") : nullptr;
    if (!objs) throw std::runtime_error(path + ": no ForceSet");
    Ctx cx{M};
    cx.v3 = true;
    std::vector<std::string> skipped;
    for (auto& c : objs->children) {
        if (boolOf(c->childText("isDisabled"), false)) continue;
        if (c->tag == "CoordinateActuator") readCoordinateActuator(cx, *c, "/forceset/");
        else skipped.push_back(c->tag + ":" + c->name());
    }
    return skipped;
}

// ExternalLoads (Simulation/Model/ExternalLoads.cpp) + ExternalForce::setDataSource (ExternalForce.cpp:200-317): the three
// columns whose labels start with the force / point / torque identifier become functions of time.
void appendExternalLoads(ModelSpec& M, const std::string& xmlPath, double tmin, double tmax) {
    auto doc = parseXmlFile(xmlPath);
    const XmlNode* el = doc->child("ExternalLoads");
    const XmlNode* objs = el ? el->child("objects") : nullptr;
    if (!objs) throw std::runtime_error(xmlPath + ": no ExternalLoads");
    std::string datafile = el->childText("datafile");
    if (datafile.empty()) throw std::runtime_error(xmlPath + ": no datafile");
    if (datafile[0] != '/') {
        size_t sl = xmlPath.rfind('/');
        datafile = (sl == std::string::npos ? std::string() : xmlPath.substr(0, sl + 1)) + datafile;
    }
    StoTable tab = readSto(datafile);
    if (tmin > -1e299 || tmax < 1e299) {
        StoTable w;
        w.labels = tab.labels;
        w.inDegrees = tab.inDegrees;
        w.cols.resize(tab.cols.size());
        for (size_t r = 0; r < tab.time.size(); ++r)
            if (tab.time[r] >= tmin && tab.time[r] <= tmax) {
                w.time.push_back(tab.time[r]);
                for (size_t c = 0; c < tab.cols.size(); ++c) w.cols[c].push_back(tab.cols[c][r]);
            }
        if (w.time.size() < 4) throw std::runtime_error(xmlPath + ": fewer than 4 data rows inside the requested time window");
        tab = std::move(w);
    }
    Ctx cx{M};
    cx.v3 = true;
    auto functionsOf = [&](const std::string& id, PiecewisePoly* out) {
        // getDataForIdentifier: the columns whose label starts with the identifier, in file order
        std::vector<int> cols;
        for (size_t c = 0; c < tab.labels.size(); ++c)
            if (tab.labels[c].compare(0, id.size(), id) == 0) cols.push_back((int)c);
        if (cols.size() != 3) throw std::runtime_error("ExternalForce: 3 unique components could not be identified for identifier: " + id);
        const size_t nt = tab.time.size();
        for (int a = 0; a < 3; ++a) {
            const std::vector<double>& y = tab.cols[cols[a]];
            if (nt == 1) out[a] = constantFunction(y[0]);
            else if (nt <= 3) out[a] = piecewiseLinearFunction(tab.time, y);
            else out[a] = fitNaturalSpline(tab.time, y, 3);
        }
    };
    for (auto& f : objs->children) {
        if (f->tag != "ExternalForce") continue;
        if (boolOf(f->childText("isDisabled"), false) || !boolOf(f->childText("appliesForce"), true)) continue;
        ExternalForceSpec X;
        X.name = f->name();
        if (!cx.baseFrame(f->childText("applied_to_body"), X.appliedTo) || X.appliedTo < 0)
            throw std::runtime_error("ExternalForce '" + X.name + "': applied_to_body must be a body");
        if (!cx.baseFrame(f->childText("force_expressed_in_body", "ground"), X.forceExpressedIn) ||
            !cx.baseFrame(f->childText("point_expressed_in_body", "ground"), X.pointExpressedIn))
            throw std::runtime_error("ExternalForce '" + X.name + "': unknown expressed-in body");
        const std::string fid = f->childText("force_identifier"), pid = f->childText("point_identifier"), tid = f->childText("torque_identifier");
        X.hasForce = !fid.empty();
        X.hasPoint = X.hasForce && !pid.empty();
        X.hasTorque = !tid.empty();
        if (X.hasForce) functionsOf(fid, X.force);
        if (X.hasPoint) functionsOf(pid, X.point);
        if (X.hasTorque) functionsOf(tid, X.torque);
        M.externalForces.push_back(std::move(X));
    }
}

} // namespace mocob200
