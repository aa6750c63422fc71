/* mocob200.h — C ABI of the B200-native direct-collocation evaluator.
 *
 * This is the drop-in boundary for the hot path of OpenSim Moco's MocoCasADiSolver:
 * the per-grid-point multibody + muscle evaluation (the CasOC point functions) and
 * the finite-difference sparse Jacobian that IPOPT asks for every iteration.
 *
 * Every entry point below names the reference interface it replaces
 * (paths relative to the reference tree):
 *
 *   mocob200_create            <- CasOC::Problem::initialize + Transcription ctor
 *                                 (Moco/Moco/MocoCasADiSolver/CasOCProblem.h:384-468,
 *                                  CasOCTranscription.cpp:83-251)
 *   mocob200_eval_points       <- CasOC::MultibodySystemExplicit/Implicit::eval,
 *                                 CostIntegrand::eval mapped over the grid
 *                                 (CasOCFunction.cpp:232-245,317-331,145-151;
 *                                  CasOCTranscription.cpp:1179-1212)
 *   mocob200_eval_f / _g       <- nlp_f / nlp_g of casadi::nlpsol
 *                                 (CasOCTranscription.cpp:448-585, 253-446; .h:219-313)
 *   mocob200_eval_grad_f       <- nlp_grad_f (CasADi FD through the callbacks)
 *   mocob200_jac_g_structure   <- CSC sparsity of nlp_jac_g (optim_write_sparsity,
 *                                 CasOCTranscription.cpp:661-677)
 *   mocob200_eval_jac_g        <- nlp_jac_g (CasOCFunction.h:38-44 enable_fd/fd_method)
 *   mocob200_set_fd            <- optim_finite_difference_scheme (MocoCasADiSolver.cpp:255-257)
 *   mocob200_set_sparsity_detection <- optim_sparsity_detection (CasOCFunction.cpp:25-105, CasOCSolver.cpp:71-87)
 *
 * Conventions: plain C, POD only, caller owns every buffer, every call returns an
 * int status (0 = ok), nothing throws across the boundary, NaN/Inf pass through.
 * One handle is driven by one host thread (IPOPT is serial); concurrency is inside
 * (CUDA streams, NCCL).  All floating point is FP64.
 */
#ifndef MOCOB200_H
#define MOCOB200_H

#ifdef __cplusplus
extern "C" {
#endif

#define MOCOB200_OK 0
#define MOCOB200_ERR_ARG 1
#define MOCOB200_ERR_UNSUPPORTED 2
#define MOCOB200_ERR_CUDA 3
#define MOCOB200_ERR_IO 4
#define MOCOB200_ERR_NCCL 5

/* joint (mobilizer) kinds; conventions of Simbody mobilizers used by
 * opensim-core/OpenSim/Simulation/SimbodyEngine/{Pin,Slider,Planar,Weld}Joint.cpp */
enum { MOCOB200_JOINT_WELD = 0, MOCOB200_JOINT_PIN = 1, MOCOB200_JOINT_SLIDER = 2,
       MOCOB200_JOINT_PLANAR = 3,
       /* CustomJoint (SimbodyEngine/CustomJoint.cpp:205-273 -> SimTK::MobilizedBody::FunctionBased): X_FM(q) =
        * [ R(axis0, f0) R(axis1, f1) R(axis2, f2) , axis3 f3 + axis4 f4 + axis5 f5 ], every f a scalar function of ONE of the
        * joint's coordinates (or a constant); 1..6 coordinates; u = qdot. Record: MOCOB200_CJ_* below. */
       MOCOB200_JOINT_CUSTOM = 4 };
/* scalar functions of one variable used by custom joints (Common/{Constant,LinearFunction,SimmSpline,MultiplierFunction}.cpp) */
enum { MOCOB200_FN_CONSTANT = 0 /* c1 */, MOCOB200_FN_LINEAR = 1 /* c0 x + c1 */, MOCOB200_FN_SIMMSPLINE = 2 /* spline id, scale inside */ };
/* path point kinds (PathPoint / ConditionalPathPoint / MovingPathPoint) */
enum { MOCOB200_PP_FIXED = 0, MOCOB200_PP_CONDITIONAL = 1, MOCOB200_PP_MOVING = 2 };
/* goal kinds (Moco/Moco/MocoGoal) */
enum { MOCOB200_GOAL_CONTROL = 0, MOCOB200_GOAL_STATE_TRACKING = 1,
       MOCOB200_GOAL_CONTACT_TRACKING = 2, MOCOB200_GOAL_FINAL_TIME = 3,
       MOCOB200_GOAL_AVERAGE_SPEED = 4, MOCOB200_GOAL_PERIODICITY = 5,
       MOCOB200_GOAL_INITIAL_ACTIVATION = 6, MOCOB200_GOAL_INITIAL_VELOCITY_EQUILIBRIUM_DGF = 7,
       MOCOB200_GOAL_SUM_SQUARED_STATE = 8, MOCOB200_GOAL_CONTROL_TRACKING = 9,
       /* path constraints (Moco/Moco/MocoConstraint.h) share the goal record; their mode is MOCOB200_MODE_PATH_CONSTRAINT */
       MOCOB200_PATHCON_CONTROL_BOUND = 10 /* MocoControlBoundConstraint.cpp:121-148 */ };
/* MODE_PATH_CONSTRAINT: n_outputs equations evaluated at every mesh point (CasOCTranscription.cpp:414-432); rows sit in
 * front of the mesh point's residual rows (CasOCTranscription.h:279-283). */
enum { MOCOB200_MODE_COST = 0, MOCOB200_MODE_ENDPOINT_CONSTRAINT = 1, MOCOB200_MODE_PATH_CONSTRAINT = 2 };
enum { MOCOB200_SCHEME_TRAPEZOIDAL = 0, MOCOB200_SCHEME_HERMITE_SIMPSON = 1 };
enum { MOCOB200_DYNAMICS_EXPLICIT = 0, MOCOB200_DYNAMICS_IMPLICIT = 1 };
enum { MOCOB200_FD_FORWARD = 0, MOCOB200_FD_CENTRAL = 1, MOCOB200_FD_BACKWARD = 2 };
/* what mocob200_eval_points returns */
enum { MOCOB200_POINTS_OUTPUTS = 0 /* [n_grid][n_out] */, MOCOB200_POINTS_JACOBIAN = 1 /* [n_grid][n_out][n_in] */,
       /* [n_grid][nmus][2]: muscle-tendon path length and lengthening speed (GeometryPath.cpp:848-864, 1117-1145), what the
        * reference's Actuation / MuscleAnalysis analyses report; single rank */
       MOCOB200_POINTS_PATHS = 2 };

/* Layout of per-entity records inside the pools.  All strides are in elements. */
/* wrap cylinder: dbl = [R_BW row-major (wrap frame in its body: xyz_body_rotation, body-fixed X-Y-Z), p_BW (translation),
 * radius, length, 0, 0]; idx = [body (-1 ground), wrap axis (0 x, 1 y), wrap sign (+-1; 0 = all quadrants), active] */
#define MOCOB200_WO_DSTRIDE 16
#define MOCOB200_WO_ISTRIDE 4
/* PathWrap: idx = [wrap object, range start, range end (1-based path-point numbers, < 1 = first / last), 0] */
#define MOCOB200_PW_ISTRIDE 4
#define MOCOB200_BODY_DSTRIDE 34   /* X_PF R(9) p(3) | X_BM R(9) p(3) | mass | com(3) | inertia about origin xx yy zz xy xz yz (6) */
#define MOCOB200_BODY_ISTRIDE 4    /* parent (-1 ground) | joint kind | first q | ndof */
#define MOCOB200_PP_DSTRIDE 5      /* loc(3) in body frame | range lo | range hi */
#define MOCOB200_PP_ISTRIDE 8      /* body | kind | cond coord q | spline x,y,z | coord x (q idx) | coord y | -- z is idx[7] */
#define MOCOB200_MUS_DSTRIDE 14    /* Fmax lopt lslack alpha vmax tau_a tau_d damping eT e0 width_scale min_ctl max_ctl kT */
#define MOCOB200_MUS_ISTRIDE 10    /* pp_first | pp_count | control | act state(z idx or -1) | tendon-force state (z idx or -1) | flags | deriv idx (in derivatives, -1) | auxres idx (-1) | 0 | 0 */
#define MOCOB200_MUS_FLAG_IGNORE_TENDON_COMPLIANCE 1
#define MOCOB200_MUS_FLAG_IGNORE_ACTIVATION_DYNAMICS 2
#define MOCOB200_MUS_FLAG_IGNORE_PASSIVE 4
#define MOCOB200_MUS_FLAG_TENDON_IMPLICIT 8
/* PathActuator (Actuators/PathActuator.cpp:156-197): tension = optimal_force (the Fmax slot) x control; no states */
#define MOCOB200_MUS_FLAG_PATH_ACTUATOR 16
#define MOCOB200_CONTACT_DSTRIDE 28 /* sphere centre(3) radius | halfspace R(9) p(3) | stiffness dissipation mu_s mu_d mu_v v_t c_f b_d b_v | pad(3) */
#define MOCOB200_CONTACT_ISTRIDE 2  /* sphere body | half-space body (-1 ground) */
#define MOCOB200_CACT_DSTRIDE 1     /* optimal force */
#define MOCOB200_CACT_ISTRIDE 2     /* speed index u | control */
#define MOCOB200_SPLINE_ISTRIDE 3   /* first knot offset (into dbl, relative to o_spline_data) | n knots | 0 */
#define MOCOB200_GOAL_ISTRIDE 8     /* kind | mode | n_outputs | has_integrand | ioff (into idx) | doff (into dbl) | ni | nd */
/* payload of MOCOB200_PATHCON_CONTROL_BOUND: idx = n_eq x { control | ppoly id of the bound function } ;
 * dbl = weight(1) 0 0 | n_eq x { lower, upper } (bounds of the equation control - bound(t)) */
#define MOCOB200_PPOLY_ISTRIDE 3    /* first break offset (rel. o_ppoly_data) | n pieces | order (coeffs per piece) */
#define MOCOB200_CJ_DSTRIDE 30      /* 6 x { axis(3) | c0 | c1 } : rotation1..3 then translation1..3 (SpatialTransform.cpp:168-176) */
#define MOCOB200_CJ_ISTRIDE 18      /* 6 x { function kind | q index driving it (-1: constant) | spline id (MOCOB200_FN_SIMMSPLINE) } */
#define MOCOB200_CC_ISTRIDE 4       /* function kind (MOCOB200_FN_*) | independent q | dependent q | spline id */
#define MOCOB200_CC_DSTRIDE 2       /* c0 c1 of the linear / constant function, scale_factor folded in (splines carry it in their record) */
#define MOCOB200_XFORCE_ISTRIDE 12  /* applied body | force-expressed-in body (-1 ground) | point-expressed-in body | 9 ppoly ids: force xyz, point xyz, torque xyz (-1 = none) */

/* A fixed-topology model + optimal control problem flattened into two pools.
 * Filled by the host flattener (the reference analogue is MocoProblemRep,
 * Moco/Moco/MocoProblemRep.cpp:36-559, read by MocoCasOCProblem.cpp:30-256). */
typedef struct mocob200_tables {
    int n_dbl, n_int;
    const double* dbl;
    const int* idx;
    /* sizes */
    int nb;          /* mobilized bodies (ground excluded), Simbody tree (breadth-first) order */
    int nq;          /* coordinates == speeds */
    int nz;          /* auxiliary states (activation, normalized tendon force) */
    int nc;          /* controls */
    int nderiv_aux;  /* implicit auxiliary derivative variables */
    int nmus, npp, nspline, ncontact, ncact, ngoal, nppoly;
    /* offsets into dbl */
    int o_body_d, o_gravity, o_pp_d, o_mus_d, o_contact_d, o_cact_d, o_spline_data, o_ppoly_data;
    int o_bounds;    /* 6 doubles per variable: lo hi init_lo init_hi final_lo final_hi ; states (nq+nq+nz) then controls (nc) */
    int o_time_bounds; /* t0 lo,hi ; tf lo,hi */
    /* offsets into idx */
    int o_body_i, o_pp_i, o_mus_i, o_contact_i, o_cact_i, o_spline_i, o_goal_i, o_ppoly_i;
    /* ---- format version 2 additions (all zero in version-1 problem files) ---- */
    int ncustom;     /* custom joints */
    int o_cj_d, o_cj_i;
    int o_body_cj;   /* idx[nb]: custom-joint record of body b, or -1 (valid when ncustom > 0) */
    int nxforce;     /* ExternalForce components (Model/ExternalForce.cpp:330-355): functions of time applied to a body */
    int o_xforce_i;
    /* kinematic constraints: CoordinateCouplerConstraints (SimbodyEngine/CoordinateCouplerConstraint.cpp:40-237 ->
     * SimTK::Constraint::CoordinateCoupler): one holonomic equation each, error = scale f(q_independent) - q_dependent, one
     * Lagrange multiplier each (and one velocity-correction slack each with enforce_constraint_derivatives + Hermite-Simpson) */
    int ncoupler;
    int o_cc_i, o_cc_d;
    /* prescribed kinematics (Components/PositionMotion.cpp:30-77, MocoProblemRep.cpp:70-100,546-560,
     * CasOCProblem.h:223-233,485-517): prescribed != 0 = every coordinate follows a function of time, idx[o_presc_i + k] = the
     * piecewise polynomial of coordinate k (value; speed and acceleration are its derivatives). The NLP then has no coordinate
     * value / speed states and no acceleration variables, no kinematic-constraint rows, and nq multibody residual rows per grid
     * point (implicit mode is required, MocoCasADiSolver.cpp:147-151); multipliers stay variables and their forces are applied */
    int prescribed, o_presc_i;
    /* cylinder wrapping (Simulation/Wrap/WrapCylinder.cpp, Model/GeometryPath.cpp:868-1094 applyWrapObjects):
     * nwrapobj wrap cylinders fixed on bodies, npathwrap (muscle, cylinder) pairings in PathWrapSet order.
     * idx[o_pw_i + 2 m], idx[o_pw_i + 2 m + 1] = first pairing and number of pairings of muscle m; the pairing records follow
     * at idx[o_pw_i + 2 nmus + k MOCOB200_PW_ISTRIDE]. */
    int nwrapobj, o_wo_d, o_wo_i;
    int npathwrap, o_pw_i;
} mocob200_tables;

/* Solver-side settings: mirrors the properties of MocoDirectCollocationSolver /
 * MocoCasADiSolver that shape the NLP (MocoDirectCollocationSolver.h:90-168,
 * MocoCasADiSolver.h:115-159). */
typedef struct mocob200_transcription {
    int scheme;                        /* MOCOB200_SCHEME_* (default hermite-simpson) */
    int dynamics_mode;                 /* MOCOB200_DYNAMICS_* */
    int n_mesh;                        /* mesh points = intervals + 1 */
    const double* mesh;                /* [n_mesh], 0..1 increasing; NULL = uniform */
    int interpolate_control_midpoints; /* default 1 */
    int fd_scheme;                     /* MOCOB200_FD_* ; solver default central */
    double accel_bounds[2];            /* implicit_multibody_acceleration_bounds */
    double auxderiv_bounds[2];         /* implicit_auxiliary_derivative_bounds */
    int minimize_accelerations;  double accelerations_weight;
    int minimize_auxiliary_derivatives; double auxiliary_derivatives_weight;
    /* ---- format version 2 additions (zero in version-1 files = the solver defaults below are applied by the loader) ---- */
    /* kinematic constraints (MocoCasADiSolver.h:115-159): enforce_constraint_derivatives, minimize_lagrange_multipliers,
     * lagrange_multiplier_weight, velocity_correction_bounds; multiplier bounds of MocoProblem (default [-1000, 1000]) */
    int enforce_constraint_derivatives; int minimize_lagrange_multipliers;
    double lagrange_multiplier_weight;
    double velocity_correction_bounds[2];
    double multiplier_bounds[2];
    int reserved[8];
} mocob200_transcription;

typedef struct mocob200_dims {
    int n_grid, n_mesh, n_var, n_con, nnz_jac;
    int ns, nc, nd, n_in, n_out;       /* point function: n_in = 1+ns+nc+nm+nd ; n_out = nu + nz + n_auxres + n_kc + n_integrands */
    int n_dyn_out, n_integrand;
    int n_cost_terms;
    int first_interval, n_intervals_local; /* this rank's shard of mesh intervals */
    int rank, world;
    int nm, nsl, n_kc;                 /* multipliers per grid point, slacks per interior point, kinematic-constraint equations per mesh point */
} mocob200_dims;

typedef struct mocob200_handle mocob200_handle;
typedef struct mocob200_problem mocob200_problem; /* host-side: owns a tables + transcription pair */

/* ---- host side: flatten models / build the reference's example problems ---- */
/* name: "sliding_mass" (exampleSlidingMass.cpp / testMocoInterface.cpp:66-83),
 *       "hanging_muscle" (testMocoActuators.cpp:1050-1175), "pendulum:<N>" (ModelFactory::createNLinkPendulum),
 *       "gait2d_tracking", "gait2d_prediction" (example2DWalking.cpp:58-340; need data_dir with 2D_gait.osim etc.).
 * options: "key=value;..." — num_mesh_intervals, scheme, dynamics_mode, fd_scheme, ... (NULL = the example's own). */
int mocob200_problem_create_example(const char* name, const char* data_dir, const char* options,
                                    mocob200_problem** out, char* err, int errlen);
/* Load a flattened problem written by mocob200_problem_save (so a GPU box needs no .osim). */
int mocob200_problem_load(const char* path, mocob200_problem** out, char* err, int errlen);
/* options (may be NULL) override the stored transcription: "num_mesh_intervals=800;fd_scheme=central;..." */
int mocob200_problem_load_opts(const char* path, const char* options, mocob200_problem** out, char* err, int errlen);
int mocob200_problem_save(const mocob200_problem*, const char* path);
/* '\n'-joined names in NLP order; which = 0 states, 1 controls, 2 auxiliary derivative variables, 3 multipliers, 4 slacks */
int mocob200_problem_names(const mocob200_problem*, int which, char* buf, int buflen);
const mocob200_tables* mocob200_problem_tables(const mocob200_problem*);
const mocob200_transcription* mocob200_problem_transcription(const mocob200_problem*);
void mocob200_problem_destroy(mocob200_problem*);
/* NLP layout computed on the host only (no GPU needed). bounds arrays may be NULL. */
int mocob200_layout(const mocob200_tables*, const mocob200_transcription*, mocob200_dims* dims);
int mocob200_layout_bounds(const mocob200_tables*, const mocob200_transcription*,
                           double* x_lo, double* x_hi, double* g_lo, double* g_hi);
int mocob200_layout_structure(const mocob200_tables*, const mocob200_transcription*,
                              int* colptr /*[n_var+1]*/, int* rowidx /*[nnz]*/);
/* A deterministic iterate: guess from bounds midpoints (createInitialGuessFromBounds,
 * CasOCTranscription.cpp:1123-1149) when seed<0, else uniform within bounds scaled by `spread`
 * about the midpoint (createRandomIterateWithinBounds, :1151-1177). */
int mocob200_make_iterate(const mocob200_tables*, const mocob200_transcription*, long seed, double spread, double* x);

/* Guesses and solutions as MocoTrajectory tables (.sto; Moco/Moco/MocoTrajectory.cpp:610-860, setGuessFile of
 * MocoCasADiSolver.h:162-202): write = one row per grid point, columns time | states | controls | multipliers | derivatives |
 * slacks with the num_* metadata (MocoTrajectory.cpp:793-829; the midpoint-only slacks stretched over the grid times as
 * MocoCasOCProblem.h:170-185 does); read = columns matched by name and resampled onto this problem's grid with the interpolating spline
 * MocoTrajectory::resample uses (degree min(rows-1, 5)); t0/tf from the first/last row; slack columns keep their midpoint
 * entries (CasOCTranscription.cpp:601-618). Host only; the problem must carry
 * variable names (created from a model, not loaded from a .mb20 file). */
int mocob200_iterate_write_sto(const mocob200_problem*, const double* x, const char* path);
int mocob200_iterate_read_sto(const mocob200_problem*, const char* path, double* x, char* err, int errlen);

/* Number of ranks that actually get grid points: min(world, n_grid / 96) but at least 1 — a rank is only worth its launch +
 * exchange latency with a hundred or so grid points; the remaining ranks idle (the reference caps its worker threads the same
 * way, MocoCasADiSolver.cpp:127-143). A fixed rule evaluated identically on every rank. */
int mocob200_effective_world(int n_grid, int world);
/* Process-wide override of the 96 (<= 0 restores it); every rank of a job must make the same call before mocob200_create.
 * For tests and for measuring the sharded path on small grids. */
void mocob200_set_min_points_per_rank(int n);
/* Grid points [*p0, *p1) evaluated by `rank` of `world` (contiguous, every point owned by exactly one rank, ranks >=
 * mocob200_effective_world get the empty range; the reference's analogue is the split of the time indices over CasADi
 * Map threads, CasOCTranscription.cpp:1179-1212).
 * Host only; mocob200_create uses the same rule, and rank r's rows of the per-point blocks Y[n_grid][n_out] /
 * J[n_grid][n_out][n_in] are gathered to rank 0 at offset *p0 * stride. */
int mocob200_shard_range(int n_grid, int rank, int world, int* p0, int* p1);

/* ---- device side ---- */
/* rank/world: this process's shard of mesh intervals (world=1: everything).
 * nccl_unique_id: 128 bytes shared by all ranks (NULL when world==1). */
int mocob200_create(const mocob200_tables*, const mocob200_transcription*, int device,
                    int rank, int world, const void* nccl_unique_id, mocob200_handle** out);
int mocob200_nccl_unique_id(void* id128);
void mocob200_destroy(mocob200_handle*);
const char* mocob200_last_error(const mocob200_handle*);
int mocob200_get_dims(const mocob200_handle*, mocob200_dims*);
int mocob200_set_fd(mocob200_handle*, int scheme, double h /*<=0: CasADi default*/, int h_iter /*<0 default*/,
                    double u_aim /*<=0 default 100*/);
/* optim_sparsity_detection (MocoCasADiSolver.h:139-147; CasOCFunction.cpp:25-105): mode 0 = "none" (dense callback
 * blocks, the default), 1 = "random" (block sparsity detected by perturbation at n_iterates random iterates within the
 * bounds, <= 0: 3). Changes nnz_jac and the CSC structure: call before mocob200_jac_g_structure / mocob200_get_dims. */
int mocob200_set_sparsity_detection(mocob200_handle*, int mode, int n_iterates);
int mocob200_jac_g_structure(const mocob200_handle*, int* colptr, int* rowidx);
/* x, outputs: HOST pointers (pinned or pageable). Full-size arrays on every rank;
 * after the call rank 0 holds the gathered result (all ranks for eval_f). */
int mocob200_eval_points(mocob200_handle*, int what, const double* x, double* out);
int mocob200_eval_f(mocob200_handle*, const double* x, double* f, double* f_terms /*[n_cost_terms] or NULL*/);
int mocob200_eval_g(mocob200_handle*, const double* x, double* g);
int mocob200_eval_grad_f(mocob200_handle*, const double* x, double* grad);
int mocob200_eval_jac_g(mocob200_handle*, const double* x, double* values);
/* Optional delivery of the values into a PINNED (page-locked) caller buffer, single rank (masked = 1; default 0 = the plain
 * device-to-host copy): the entries of the dense-block structure that are +0.0 at every iterate seen so far (CasADi keeps them
 * in the structure under optim_sparsity_detection = "none", MocoCasADiSolver.cpp:39, CasOCFunction.cpp:25-74) do not cross
 * PCIe - the host zero-fills their runs while the device computes and a kernel stores the other entries straight into the
 * buffer. A kernel-side check of every skipped entry makes the call fall back to the plain copy (and widen the kept set) when
 * one of them is not +0.0, so the buffer is bit-identical to the plain copy in every case. Pageable buffers, sharded handles
 * and problems with path-constraint / velocity-correction rows always take the plain copy. Off by default because what it
 * gains depends on the host: config 3 on two B200 boxes, 175 us against 183 us per call on one, 270 us against 185 us on the
 * other, where the host's zero-fill (2.2 MB in 16 k runs, one core) outlasted the device work (profiles/experiments_r02.md).
 * mocob200_host_push_stats: calls delivered each way; kept: entries that cross PCIe (0: not in use). */
int mocob200_set_host_delivery(mocob200_handle*, int masked);
int mocob200_host_push_stats(const mocob200_handle*, long* pushed, long* fallbacks, long* kept);
/* Same evaluations with x already resident on the device and results left there (device pointers; used by bench.py's
 * device-resident timing and by callers that chain kernels). The work runs on the handle's own stream. cuda_stream != NULL: that
 * stream is ordered around the call - the handle's stream first waits for everything already queued on cuda_stream (the
 * producer of d_x) and cuda_stream then waits for the pipeline, so consumers of the outputs can be queued on it directly.
 * cuda_stream == NULL: no ordering is added (NULL cannot be told from the legacy default stream); call mocob200_sync. */
int mocob200_eval_jac_g_device(mocob200_handle*, const double* d_x, double* d_values, void* cuda_stream);
int mocob200_eval_g_device(mocob200_handle*, const double* d_x, double* d_g, void* cuda_stream);
/* eval_g for callers that cannot allocate device memory (code under CUDA graph capture, tight serving loops, harnesses that
 * forbid cudaMalloc). NO COUNTERPART IN THE REFERENCE. Same result as
 * mocob200_eval_g_device (d_x[n_var] -> d_g[n_con], device pointers), but the pipeline's scratch (per-point outputs, goal
 * integrals, endpoint values) is a static __device__ array inside the library instead of buffers of the handle, and the call
 * itself allocates nothing (no cudaMalloc / cudaFree / event or stream creation): it clears the part of the array it uses and
 * queues its kernels on cuda_stream (NULL: the handle's own stream; then call mocob200_sync). Limits: single-rank handles;
 * scratch need = 8 * (n_grid * n_out + goal terms) bytes <= MOCOB200_STATIC_SCRATCH_BYTES, else MOCOB200_ERR_ARG with the
 * sizes in mocob200_last_error. Every GPU that loads the library holds that array until exit. ALL calls (of all handles)
 * share it and it keeps the last call's contents: calls that overlap on the GPU (different streams, no ordering) race. */
#define MOCOB200_STATIC_SCRATCH_BYTES (1 << 20)
int mocob200_eval_g_static(mocob200_handle*, const double* d_x, double* d_g, void* cuda_stream);
/* n independent single-rank problems on one device evaluated side by side (one handle each; each on its own stream,
 * forked from and joined to handles[0]'s stream, so mocob200_mark / mocob200_sync on handles[0] cover the batch).
 * d_values may be NULL (or hold NULL entries): results stay in the handles' own buffers. Replaces running one
 * reference process per study (one NLP per MocoCasADiSolver::solveImpl, MocoCasADiSolver.cpp:312-415). */
int mocob200_eval_jac_g_batch(mocob200_handle* const* handles, int n, const double* const* d_x, double* const* d_values);
/* the *_device and *_batch calls only enqueue work on the handle's stream: */
int mocob200_sync(mocob200_handle*);
int mocob200_mark(mocob200_handle*, int which /*0 start, 1 stop*/); /* CUDA-event stopwatch on that stream */
int mocob200_elapsed_ms(mocob200_handle*, float* ms);
/* bookkeeping for bench.py */
long mocob200_launch_count(const mocob200_handle*);           /* kernels launched so far */
int mocob200_last_kernel_ms(const mocob200_handle*, float* point_kernel_ms, float* assemble_kernel_ms);
long mocob200_point_evals_per_jacobian(const mocob200_handle*); /* credited: what differencing the whole function costs */
/* executed: evaluations one eval_jac_g launches over all ranks: whole | pose-reusing | pose+velocity-reusing | local-column warps */
int mocob200_executed_evals(const mocob200_handle*, long out[4]);
int mocob200_working_ranks(const mocob200_handle*);      /* ranks that own grid points */
/* 1 once a sharded Jacobian has been evaluated with the other ranks storing their CSC runs straight into rank 0's value
 * array over NVLink (CUDA IPC mappings, set up at the first sharded Jacobian); 0: single rank, not yet evaluated, or the
 * NCCL send / recv gather is in use because a rank could not map rank 0's memory. No counterpart in the reference. */
int mocob200_peer_gather(const mocob200_handle*);
int mocob200_uses_warp_kernels(const mocob200_handle*);  /* 1: warp-per-evaluation FD kernels, 0: thread-per-evaluation */
/* FP64 FMA saturation microbenchmark: returns measured TFLOP/s (the FP64 roofline peak). */
int mocob200_measure_fp64_peak(int device, double* tflops);

#ifdef __cplusplus
}
#endif
#endif
