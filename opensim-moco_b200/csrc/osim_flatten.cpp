// .osim (OpenSimDocument Version 40000) -> ModelSpec, for the component subset of SURVEY.md Appendix E:
// Body, Pin/Slider/Planar/Weld joints with PhysicalOffsetFrames, DeGrooteFregly2016Muscle with
// PathPoint / ConditionalPathPoint / MovingPathPoint(SimmSpline in MultiplierFunction),
// ContactSphere / ContactHalfSpace / SmoothSphereHalfSpaceForce, CoordinateActuator.
// Anything else that would change the dynamics is refused loudly (no silent approximation).
#include <cstdlib>
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
double dblOf(const XmlNode& n, const char* tag, double def) {
    std::string s = n.childText(tag);
    return s.empty() ? def : std::strtod(s.c_str(), nullptr);
}
std::string leaf(const std::string& path) {
    size_t sl = path.rfind('/');
    return sl == std::string::npos ? path : path.substr(sl + 1);
}

struct Ctx {
    ModelSpec& M;
    // resolve "/bodyset/x", "/ground", "ground"
    bool baseFrame(const std::string& path, int& body) const {
        if (path == "/ground" || path == "ground") { body = -1; return true; }
        const std::string pre = "/bodyset/";
        if (path.compare(0, pre.size(), pre) == 0 && path.find('/', pre.size()) == std::string::npos) {
            body = M.bodyIndex(path.substr(pre.size()));
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

SplineSpec splineOf(const XmlNode& holder, const std::string& what) {
    // <x_location><MultiplierFunction><function><SimmSpline><x/><y/></SimmSpline></function><scale/></MultiplierFunction>
    SplineSpec s;
    const XmlNode* fn = nullptr;
    if (const XmlNode* mf = holder.child("MultiplierFunction")) {
        s.scale = dblOf(*mf, "scale", 1.0);
        const XmlNode* inner = mf->child("function");
        if (inner) fn = inner->child("SimmSpline");
    } else
        fn = holder.child("SimmSpline");
    if (!fn) throw std::runtime_error(what + ": only SimmSpline (optionally inside a MultiplierFunction) is supported");
    s.x = parseDoubles(fn->childText("x"));
    s.y = parseDoubles(fn->childText("y"));
    if (s.x.size() != s.y.size() || s.x.size() < 2) throw std::runtime_error(what + ": bad SimmSpline knots");
    return s;
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
    const XmlNode* gp = n.child("GeometryPath");
    if (!gp) throw std::runtime_error("muscle '" + mu.path + "' has no GeometryPath");
    if (const XmlNode* ws = gp->child("PathWrapSet"))
        if (const XmlNode* wo = ws->child("objects"))
            if (!wo->children.empty())
                throw std::runtime_error("muscle '" + mu.path + "': PathWrap objects are unsupported (SURVEY §8 a18)");
    const XmlNode* pps = gp->child("PathPointSet");
    const XmlNode* objs = pps ? pps->child("objects") : nullptr;
    if (!objs) throw std::runtime_error("muscle '" + mu.path + "' has no path points");
    for (auto& pn : objs->children) {
        PathPointSpec pt;
        std::string frame = pn->childText("socket_parent_frame");
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
            pt.condCoord = leaf(pn->childText("socket_coordinate"));
        } else if (pn->tag == "MovingPathPoint") {
            pt.kind = MOCOB200_PP_MOVING;
            const char* ax[3] = {"x", "y", "z"};
            for (int a = 0; a < 3; ++a) {
                const XmlNode* loc = pn->child(std::string(ax[a]) + "_location");
                if (!loc) throw std::runtime_error("MovingPathPoint '" + pn->name() + "': missing location function");
                pt.fn[a] = splineOf(*loc, "MovingPathPoint '" + pn->name() + "'");
                pt.coord[a] = leaf(pn->childText(std::string("socket_") + ax[a] + "_coordinate"));
            }
        } else
            throw std::runtime_error("unsupported path point type '" + pn->tag + "'");
        mu.points.push_back(pt);
    }
    cx.M.actuatorOrder.emplace_back(1, (int)cx.M.muscles.size());
    cx.M.muscles.push_back(std::move(mu));
}

void readForceLike(const Ctx& cx, const XmlNode& n, const std::string& prefix, const XmlNode* contactGeom) {
    if (n.tag == "DeGrooteFregly2016Muscle") {
        readMuscle(cx, n, prefix);
    } else if (n.tag == "CoordinateActuator") {
        CoordActSpec a;
        a.path = prefix + n.name();
        a.coord = n.childText("coordinate");
        a.optimalForce = dblOf(n, "optimal_force", 1.0);
        a.minControl = dblOf(n, "min_control", a.minControl);
        a.maxControl = dblOf(n, "max_control", a.maxControl);
        cx.M.actuatorOrder.emplace_back(0, (int)cx.M.coordActs.size());
        cx.M.coordActs.push_back(a);
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

ModelSpec loadOsim(const std::string& path) {
    auto doc = parseXmlFile(path);
    if (doc->tag != "OpenSimDocument") throw std::runtime_error(path + ": not an OpenSimDocument");
    const std::string* ver = doc->attr("Version");
    if (!ver || std::atoi(ver->c_str()) < 40000)
        throw std::runtime_error(path + ": only document Version >= 40000 is supported");
    const XmlNode* model = doc->child("Model");
    if (!model) throw std::runtime_error(path + ": no <Model>");
    ModelSpec M;
    M.name = model->name();
    if (model->has("gravity")) M.gravity = vec3Of(model->childText("gravity"), "gravity");
    Ctx cx{M};

    const XmlNode* bs = model->child("BodySet");
    const XmlNode* bobjs = bs ? bs->child("objects") : nullptr;
    if (!bobjs) throw std::runtime_error(path + ": no BodySet");
    for (auto& b : bobjs->children) {
        if (b->tag != "Body") throw std::runtime_error("unsupported BodySet entry '" + b->tag + "'");
        BodySpec B;
        B.name = b->name();
        B.mass = dblOf(*b, "mass", 0.0);
        B.com = vec3Of(b->childText("mass_center", "0 0 0"), "mass_center");
        auto in = parseDoubles(b->childText("inertia", "0 0 0 0 0 0"));
        if (in.size() != 6) throw std::runtime_error("body '" + B.name + "': inertia needs 6 numbers");
        for (int i = 0; i < 6; ++i) B.inertiaCom[i] = in[i];
        M.bodies.push_back(B);
    }

    const XmlNode* js = model->child("JointSet");
    const XmlNode* jobjs = js ? js->child("objects") : nullptr;
    if (!jobjs) throw std::runtime_error(path + ": no JointSet");
    for (auto& j : jobjs->children) {
        JointSpec J;
        J.name = j->name();
        if (j->tag == "PinJoint") J.kind = MOCOB200_JOINT_PIN;
        else if (j->tag == "SliderJoint") J.kind = MOCOB200_JOINT_SLIDER;
        else if (j->tag == "PlanarJoint") J.kind = MOCOB200_JOINT_PLANAR;
        else if (j->tag == "WeldJoint") J.kind = MOCOB200_JOINT_WELD;
        else throw std::runtime_error("unsupported joint type '" + j->tag + "' (" + J.name + ")");
        J.parent = cx.jointFrame(*j, j->childText("socket_parent_frame"));
        J.child = cx.jointFrame(*j, j->childText("socket_child_frame"));
        if (const XmlNode* cs = j->child("coordinates"))
            for (auto& c : cs->children) {
                if (c->tag != "Coordinate") continue;
                CoordSpec C;
                C.name = c->name();
                auto r = parseDoubles(c->childText("range", "-10 10"));
                if (r.size() == 2) { C.lo = r[0]; C.hi = r[1]; }
                if (boolOf(c->childText("locked"), false) || boolOf(c->childText("prescribed"), false))
                    throw std::runtime_error("coordinate '" + C.name + "': locked/prescribed coordinates are unsupported");
                J.coords.push_back(C);
            }
        M.joints.push_back(J);
    }
    if (const XmlNode* cs = model->child("ConstraintSet"))
        if (const XmlNode* o = cs->child("objects"))
            if (!o->children.empty()) throw std::runtime_error("kinematic constraints are unsupported (SURVEY §8 a9)");

    const XmlNode* cg = model->child("ContactGeometrySet");
    const XmlNode* cgobjs = cg ? cg->child("objects") : nullptr;
    if (const XmlNode* comps = model->child("components"))
        for (auto& c : comps->children) readForceLike(cx, *c, "/", cgobjs);
    if (const XmlNode* fs = model->child("ForceSet"))
        if (const XmlNode* o = fs->child("objects"))
            for (auto& c : o->children) readForceLike(cx, *c, "/forceset/", cgobjs);
    return M;
}

} // namespace mocob200
