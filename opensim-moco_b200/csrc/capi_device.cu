// Device half of the C ABI: handle life cycle, evaluation pipelines, NCCL gather of the point blocks.
#include <cuda_runtime.h>
#include <emmintrin.h> // SSE2 stores of the host zero-fill (x86-64 baseline)
#include <dlfcn.h>

#include <algorithm>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <mutex>
#include <string>
#include <vector>

#include "kernels.cuh"
#include "layout.h"

using namespace mocob200;

namespace {

// ---- minimal NCCL binding, resolved at run time so single-GPU use needs no NCCL at all
typedef struct ncclComm* ncclComm_t;
struct NcclUid { char internal[128]; };
struct NcclApi {
    void* lib = nullptr;
    int (*GetUniqueId)(NcclUid*) = nullptr;
    int (*CommInitRank)(ncclComm_t*, int, NcclUid, int) = nullptr;
    int (*CommDestroy)(ncclComm_t) = nullptr;
    int (*Broadcast)(const void*, void*, size_t, int, int, ncclComm_t, cudaStream_t) = nullptr;
    int (*Send)(const void*, size_t, int, int, ncclComm_t, cudaStream_t) = nullptr;
    int (*Recv)(void*, size_t, int, int, ncclComm_t, cudaStream_t) = nullptr;
    int (*GroupStart)() = nullptr;
    int (*GroupEnd)() = nullptr;
    const char* (*GetErrorString)(int) = nullptr;
    bool load(std::string& err) {
        if (lib) return true;
        const char* names[] = {"libnccl.so.2", "libnccl.so"};
        for (const char* n : names) { lib = dlopen(n, RTLD_NOW | RTLD_GLOBAL); if (lib) break; }
        if (!lib) { err = std::string("cannot load NCCL: ") + dlerror(); return false; }
#define NCCL_SYM(field, sym) field = (decltype(field))dlsym(lib, sym); if (!field) { err = std::string("NCCL symbol missing: ") + sym; return false; }
        NCCL_SYM(GetUniqueId, "ncclGetUniqueId");
        NCCL_SYM(CommInitRank, "ncclCommInitRank");
        NCCL_SYM(CommDestroy, "ncclCommDestroy");
        NCCL_SYM(Broadcast, "ncclBroadcast");
        NCCL_SYM(Send, "ncclSend");
        NCCL_SYM(Recv, "ncclRecv");
        NCCL_SYM(GroupStart, "ncclGroupStart");
        NCCL_SYM(GroupEnd, "ncclGroupEnd");
        NCCL_SYM(GetErrorString, "ncclGetErrorString");
#undef NCCL_SYM
        return true;
    }
};
NcclApi g_nccl;
const int kNcclDouble = 8; // ncclFloat64

template <class T>
T* devAlloc(size_t n) {
    T* p = nullptr;
    if (n == 0) n = 1;
    if (cudaMalloc(&p, n * sizeof(T)) != cudaSuccess) return nullptr;
    return p;
}
template <class T>
T* devUpload(const std::vector<T>& v) {
    T* p = devAlloc<T>(v.size());
    if (p && !v.empty()) cudaMemcpy(p, v.data(), v.size() * sizeof(T), cudaMemcpyHostToDevice);
    return p;
}

} // namespace

// Zero-fill of runs [i0, i1) of a (start, length) run list: two doubles per store
static void fillZeroRuns(double* values, const unsigned* r, size_t i0, size_t i1) {
    const __m128d z = _mm_setzero_pd();
    for (size_t i = i0; i < i1; i += 2) {
        double* p = values + r[i];
        const unsigned n = r[i + 1];
        unsigned k = 0;
        for (; k + 2 <= n; k += 2) _mm_storeu_pd(p + k, z);
        if (k < n) p[k] = 0.0;
    }
}
struct mocob200_handle {
    std::string err;
    int device = 0, rank = 0, world = 1;
    int eff = 1; // ranks that own grid points (mocob200_effective_world)
    bool velKernel = true; // thread path: pose-reusing directions through eval_point_vel (MOCOB200_VEL_KERNEL=0: whole evaluations, A-B)
    int mergedBodies = 0; // welded bodies folded into their parents in the device tables
    double* d_Jt = nullptr; // sharded evaluation: time column of every grid point's FD block, gathered to rank 0
    DevLayout DLt{};        // DL with Jt set (rank 0's assembly of the t0 / tf columns)
    // CSC value runs of the columns each rank assembles itself: [rank][block: states, controls, derivatives] = [k0, k1)
    std::vector<int> runs;
    Layout L;
    DevTables DT{};
    DevLayout DL{};
    FdSettings fd{};
    WarpAux WA{};
    // point-function kernels: warp-per-evaluation (latency: few evaluations) or thread-per-evaluation articulated-body
    // (throughput: enough evaluations to fill the machine with threads). MOCOB200_KERNEL=warp|thread forces one.
    bool useWarp = true;
    int kernelMode = 0; // 0 auto, 1 warp, 2 thread
    long warpMaxEvals = 30000;
    bool beyondThreadKernels = false;
    bool wrapInWarpKernels = true; // the warp kernels carry the wrap tangent points (point_warp.cuh: wk_muscle_wraps)
    // Central differences: every direction costs 2 x (1 + h_iter) evaluations and the warp kernels run them one after the
    // other, while the paired thread kernel runs the two sides on adjacent lanes - measured on gait prediction (B200, ms per
    // Jacobian, thread / warp): 101 grid points 0.47 / 0.70, 201: 0.49 / 1.28, 401: 0.71 / 2.40, 1601: 1.24 / 8.81
    // (scripts/gpu_kernel_crossover.py). So the warp kernels are chosen for one-sided schemes only.
    bool warpFor(long evals) const {
        return kernelMode == 1 || (kernelMode == 0 && evals <= warpMaxEvals && fd.scheme != MOCOB200_FD_CENTRAL);
    }
    // the nominal evaluations alone (eval_f / eval_g / eval_points): one evaluation per grid point, whatever the FD scheme
    bool warpForValues(long evals) const { return kernelMode == 1 || (kernelMode == 0 && evals <= warpMaxEvals); }
    std::vector<void*> owned;
    double* d_cfg = nullptr; // per-point blobs of the nominal evaluation (warp kernels)
    double *d_x = nullptr, *d_Y = nullptr, *d_J = nullptr, *d_I = nullptr, *d_G = nullptr, *d_E = nullptr;
    double *d_g = nullptr, *d_vals = nullptr, *d_grad = nullptr, *d_f = nullptr;
    cudaStream_t stream = nullptr;
    // side stream: the endpoint functions (which need only x and the integrals of the nominal point outputs) run
    // beside the FD evaluations of the point functions
    cudaStream_t stream2 = nullptr;
    SideStreams side{};
    // host copies of the problem (the layout is rebuilt when the sparsity detection mode changes)
    std::vector<double> hostDbl, hostMesh;
    std::vector<int> hostIdx;
    mocob200_tables hostT{};
    mocob200_transcription hostR{};
    cudaGraphExec_t jacGraph = nullptr; // captured Jacobian pipeline (d_x -> d_vals), rebuilt after mocob200_set_fd
    long graphLaunches = 0;
    bool graphFailed = false;
    bool allowGraph = true; // read at create: MOCOB200_GRAPH=0 or MOCOB200_TIMING=1 turn the graph off
    cudaEvent_t evY = nullptr, evE = nullptr, evJ = nullptr, evI = nullptr, evX = nullptr;
    cudaEvent_t evBatch = nullptr; // fork / join event of mocob200_eval_jac_g_batch
    cudaEvent_t evCaller = nullptr; // ordering against a caller-supplied stream (*_device entry points)
    cudaEvent_t marks[2] = {nullptr, nullptr}; // mocob200_mark stopwatch (per handle: handles may sit on different devices)
    bool ecHasIntegrand = false; // some endpoint-constraint goal has an integrand: jac_g needs dI/dt0, dI/dtf
    // Peer-memory gather of the sharded Jacobian (one node, one process per GPU): the working ranks other than 0 map rank 0's
    // value array, point outputs, time columns and flag words (CUDA IPC) and their assembly kernels store straight into
    // them over NVLink - no staging copy, no collective on the data path. See peerGatherSetup / jacPipelineSharded.
    struct PeerGather {
        bool wanted = false; // mocob200_set_host_delivery: off unless the caller asks for it
        bool tried = false, on = false;
        double *vals = nullptr, *Y = nullptr, *Jt = nullptr; // rank > 0: rank 0's buffers in this process's address space
        unsigned long long* flags = nullptr; // rank 0: its flag words ([0] consumed, [r] done of rank r); rank > 0: their mapping
        unsigned long long seq = 0;          // Jacobians evaluated through the peer path since the setup
        double* d_msg = nullptr;             // setup exchange (NCCL send / recv)
        cudaEvent_t trace[8] = {};           // MOCOB200_PEER_TRACE=1 (development): segment stopwatch of the last Jacobian
        bool tracing = false;
    } peer;
    double* d_I2 = nullptr;
    cudaEvent_t ev[4] = {nullptr, nullptr, nullptr, nullptr};
    long launches = 0;
    int p0 = 0, p1 = 0; // this rank's grid points
    std::vector<int> shard; // world+1 boundaries
    ncclComm_t comm = nullptr;
    // host-side memo: IPOPT asks for grad_f and jac_g at the same x
    std::vector<double> x_cached;
    bool haveJ = false;
    // Host delivery of the CSC values into a PINNED caller buffer without the structural zeros crossing PCIe (pushSetup /
    // pushValues below): byte mask over the value array (1 = the entry was not +0.0 at one of the detection iterates), a flag
    // word in mapped host memory the kernel raises when an entry outside the mask is not +0.0 (then the call falls back to
    // the plain copy, so the caller's buffer is bit-identical either way).
    struct HostPush {
        bool wanted = false; // mocob200_set_host_delivery: off unless the caller asks for it
        bool tried = false, on = false;
        unsigned char* d_mask = nullptr;
        int* flag_h = nullptr;
        int* flag_d = nullptr;
        const void* lastHost = nullptr; // last caller buffer that was looked up ...
        double* lastDev = nullptr;      // ... and its device alias (nullptr: not pinned)
        long pushed = 0, fallbacks = 0;
        size_t kept = 0;
        std::vector<unsigned char> mask; // host copy (extended when a call finds a new non-zero entry)
        std::vector<unsigned> zeroRuns;  // (start, length) of the runs of entries outside the mask: what the host zero-fills
        std::vector<unsigned char> devMask; // mask + the entries of zero runs too short to be worth a host loop iteration
        size_t sent = 0;                    // entries of devMask: what crosses PCIe
        int shortRun = 8;                   // zero runs shorter than this go with the kernel's stores (MOCOB200_PUSH_SHORT_RUN: A-B)
        void buildRuns() {
            zeroRuns.clear();
            devMask = mask;
            const size_t n = mask.size();
            for (size_t e = 0; e < n;) {
                if (mask[e]) { ++e; continue; }
                size_t b = e;
                while (e < n && !mask[e]) ++e;
                if (e - b < (size_t)shortRun) { for (size_t k = b; k < e; ++k) devMask[k] = 1; }
                else { zeroRuns.push_back((unsigned)b); zeroRuns.push_back((unsigned)(e - b)); }
            }
            sent = 0;
            for (unsigned char m : devMask) sent += m;
        }
    } push;

    bool ck(cudaError_t e, const char* what) {
        if (e == cudaSuccess) return true;
        err = std::string(what) + ": " + cudaGetErrorString(e);
        return false;
    }
};

extern "C" long mocob200_point_evals_per_jacobian(const mocob200_handle* h);

extern "C" { static void peerGatherClose(mocob200_handle* h); } // defined with the sharded pipeline

namespace {

bool ncclOk(mocob200_handle* h, int rc, const char* what) {
    if (rc == 0) return true;
    h->err = std::string(what) + ": " + (g_nccl.GetErrorString ? g_nccl.GetErrorString(rc) : "nccl error");
    return false;
}

// Topology tables + shared-memory layout of the warp-per-evaluation kernels.
bool buildWarpAux(mocob200_handle* h, const mocob200_tables& T, std::string& err) {
    const Layout& L = h->L;
    const int nb = T.nb, nq = T.nq, npp = T.npp;
    auto bodyI = [&](int b) { return T.idx + T.o_body_i + b * MOCOB200_BODY_ISTRIDE; };
    std::vector<int> level(nb, 0), levelStart{0}, childStart(nb + 1, 0), childList, dofBody(nq, -1);
    for (int b = 0; b < nb; ++b) {
        const int par = bodyI(b)[0];
        if (par >= b) { err = "bodies are not stored parent-before-child"; return false; }
        level[b] = par < 0 ? 0 : level[par] + 1;
        if (b > 0 && level[b] < level[b - 1]) { err = "bodies are not stored level by level"; return false; }
        if (b > 0 && level[b] != level[b - 1]) levelStart.push_back(b);
        for (int k = 0; k < bodyI(b)[3]; ++k) dofBody[bodyI(b)[2] + k] = b;
    }
    levelStart.push_back(nb);
    for (int b = 0; b < nb; ++b) {
        childStart[b] = (int)childList.size();
        for (int c = b + 1; c < nb; ++c) if (bodyI(c)[0] == b) childList.push_back(c);
    }
    childStart[nb] = (int)childList.size();
    std::vector<int> bodyPpStart(nb + 1, 0), bodyPpList, ppMpp(npp > 0 ? npp : 1, -1), mppPp, mppCoord;
    for (int b = 0; b < nb; ++b) {
        bodyPpStart[b] = (int)bodyPpList.size();
        for (int p = 0; p < npp; ++p) if (T.idx[T.o_pp_i + p * MOCOB200_PP_ISTRIDE] == b) bodyPpList.push_back(p);
        // the two tangent points of every (muscle, wrap cylinder) pairing are stations of the cylinder's body: slots npp + 2 k, + 1
        for (int k = 0; k < T.npathwrap; ++k) {
            const int wo = T.idx[T.o_pw_i + 2 * T.nmus + k * MOCOB200_PW_ISTRIDE];
            if (T.idx[T.o_wo_i + wo * MOCOB200_WO_ISTRIDE] == b) { bodyPpList.push_back(npp + 2 * k); bodyPpList.push_back(npp + 2 * k + 1); }
        }
    }
    bodyPpStart[nb] = (int)bodyPpList.size();
    for (int p = 0; p < npp; ++p) {
        const int* pi = T.idx + T.o_pp_i + p * MOCOB200_PP_ISTRIDE;
        if (pi[1] == MOCOB200_PP_MOVING) { ppMpp[p] = (int)mppPp.size(); mppPp.push_back(p); mppCoord.push_back(pi[6]); }
    }
    std::vector<int> bodyFlags(nb, 0);
    for (int b = 0; b < nb; ++b) {
        const double* bd = T.dbl + T.o_body_d + b * MOCOB200_BODY_DSTRIDE;
        auto isIdentity = [](const double* R) {
            for (int k = 0; k < 9; ++k) if (R[k] != ((k % 4 == 0) ? 1.0 : 0.0)) return false;
            return true;
        };
        if (isIdentity(bd)) bodyFlags[b] |= 1;
        if (isIdentity(bd + 12)) bodyFlags[b] |= 2;
    }
    std::vector<int> pairI, pairJ;
    for (int i = 0; i < nq; ++i) {
        int b = dofBody[i];
        while (b >= 0) {
            for (int k = 0; k < bodyI(b)[3]; ++k) {
                const int j = bodyI(b)[2] + k;
                if (j <= i) { pairI.push_back(i); pairJ.push_back(j); }
            }
            b = bodyI(b)[0];
        }
    }
    // finite-difference work lists
    std::vector<int> cwDirs;
    std::vector<int> p0Dirs, p1Dirs, p2Dirs, cdDirs, locDir, locKind, locOwner, locRole, dirClass(L.n_in, 4);
    {
        const char* e = devEnv("MOCOB200_FD_LOCAL");
        bool local = !(e && e[0] == '0');
        for (int g = 0; g < T.ngoal; ++g) {
            const int* gr = T.idx + T.o_goal_i + g * MOCOB200_GOAL_ISTRIDE;
            const int k = gr[0];
            if (gr[3] && k != MOCOB200_GOAL_CONTROL && k != MOCOB200_GOAL_STATE_TRACKING && k != MOCOB200_GOAL_CONTACT_TRACKING &&
                k != MOCOB200_GOAL_SUM_SQUARED_STATE && k != MOCOB200_GOAL_CONTROL_TRACKING)
                local = false; // an integrand this analysis does not know
        }
        const int nqs = L.nqs; // coordinate states among the inputs (0 with prescribed kinematics)
        const int nacc = (L.implicit && !L.prescribed) ? nq : 0;
        const int ns = 2 * nqs + T.nz, iz = 1 + 2 * nqs, ic = 1 + ns, ider = 1 + ns + T.nc + L.nm, iaux = ider + nacc;
        for (int d = 0; d < L.n_in; ++d) {
            if (d >= 1 && d <= nqs) { dirClass[d] = 0; continue; }
            if ((d > nqs && d <= 2 * nqs) || (d >= ider && d < ider + nacc)) { dirClass[d] = 1; continue; }
            if (d == 0) { dirClass[d] = L.prescribed ? 0 : 2; continue; } // prescribed kinematics: time moves every coordinate
            if (!local) continue;
            int owners = 0, kind = -1, owner = -1, role = -1;
            for (int m = 0; m < T.nmus; ++m) {
                const int* mi = T.idx + T.o_mus_i + m * MOCOB200_MUS_ISTRIDE;
                if (d == ic + mi[2]) { ++owners; kind = 0; owner = m; role = 0; }
                if (mi[3] >= 0 && d == iz + mi[3]) { ++owners; kind = 0; owner = m; role = 1; }
                if (mi[4] >= 0 && d == iz + mi[4]) { ++owners; kind = 0; owner = m; role = 2; }
                if (mi[6] >= 0 && d == iaux + mi[6]) { ++owners; kind = 0; owner = m; role = 3; }
            }
            for (int a = 0; a < T.ncact; ++a)
                if (d == ic + T.idx[T.o_cact_i + a * MOCOB200_CACT_ISTRIDE + 1]) { ++owners; kind = 1; owner = a; role = 0; }
            if (owners != 1) continue; // shared or unread input: difference the whole function
            dirClass[d] = 3;
            locDir.push_back(d); locKind.push_back(kind); locOwner.push_back(owner); locRole.push_back(role);
        }
        for (int d = 0; d < L.n_in; ++d) {
            if (dirClass[d] != 3) cdDirs.push_back(d);
            if (dirClass[d] != 3 && dirClass[d] != 1) cwDirs.push_back(d);
            if (dirClass[d] == 0) p0Dirs.push_back(d);
            else if (dirClass[d] == 1) p1Dirs.push_back(d);
            else if (dirClass[d] != 3) p2Dirs.push_back(d);
        }
        h->WA.time_integrands_only = (local && T.nxforce == 0 && !L.prescribed) ? 1 : 0; // external forces are functions of time
    }
    // all topology arrays live in one int pool so that a block can stage them with one copy loop
    std::vector<int> pool;
    std::vector<size_t> at;
    auto put = [&](const std::vector<int>& v) { at.push_back(pool.size()); pool.insert(pool.end(), v.begin(), v.end()); pool.push_back(0); };
    put(levelStart); put(childStart); put(childList); put(bodyPpStart); put(bodyPpList); put(dofBody); put(bodyFlags);
    put(ppMpp); put(mppPp); put(mppCoord); put(pairI); put(pairJ);
    put(p0Dirs); put(p1Dirs); put(p2Dirs); put(cdDirs); put(locDir); put(locKind); put(locOwner); put(locRole); put(dirClass);
    put(cwDirs);
    const int* base = devUpload(pool);
    h->owned.push_back((void*)base);
    WarpAux& A = h->WA;
    A.aux_base = base;
    A.st_naux = (int)pool.size();
    A.nlevels = (int)levelStart.size() - 1;
    A.level_start = base + at[0];
    A.child_start = base + at[1];
    A.child_list = base + at[2];
    A.body_pp_start = base + at[3];
    A.body_pp_list = base + at[4];
    A.dof_body = base + at[5];
    A.body_flags = base + at[6];
    A.pp_mpp = base + at[7];
    A.mpp_pp = base + at[8];
    A.mpp_coord = base + at[9];
    A.nmpp = (int)mppPp.size();
    A.npairs = (int)pairI.size();
    A.pair_i = base + at[10];
    A.pair_j = base + at[11];
    A.n_p0 = (int)p0Dirs.size(); A.n_p1 = (int)p1Dirs.size(); A.n_p2 = (int)p2Dirs.size(); A.n_local = (int)locDir.size();
    A.p0_dirs = base + at[12]; A.p1_dirs = base + at[13]; A.p2_dirs = base + at[14]; A.cd_dirs = base + at[15];
    A.n_cd = (int)cdDirs.size();
    A.loc_dir = base + at[16]; A.loc_kind = base + at[17]; A.loc_owner = base + at[18]; A.loc_role = base + at[19];
    A.dir_class = base + at[20];
    A.cw_dirs = base + at[21]; A.n_cw = (int)cwDirs.size();
    // staged part of the pools: everything below the bounds block (host-only), plus the goal payload
    {
        int goal0 = T.n_dbl;
        for (int g = 0; g < T.ngoal; ++g) {
            const int* gr = T.idx + T.o_goal_i + g * MOCOB200_GOAL_ISTRIDE;
            if (gr[7] > 0 && gr[5] < goal0) goal0 = gr[5];
        }
        const bool tail = goal0 >= T.o_ppoly_data && T.o_bounds <= T.o_ppoly_data && T.o_spline_data <= T.o_bounds &&
                          T.o_body_d <= T.o_bounds && T.o_pp_d <= T.o_bounds && T.o_mus_d <= T.o_bounds && T.o_contact_d <= T.o_bounds &&
                          T.o_cact_d <= T.o_bounds && T.o_gravity <= T.o_bounds;
        if (!tail) { err = "unexpected pool order (model | bounds | references | goal payload)"; return false; }
        A.st_ndbl = T.o_bounds;
        A.st_goal0 = goal0;
        A.st_ngoal = T.n_dbl - goal0;
        A.st_nint = T.n_int;
        const int ints = A.st_nint + A.st_naux;
        A.st_doubles = ((A.st_ndbl + A.st_ngoal + 1) & ~1) + ((ints + 3) / 4) * 2;
    }
    int o = 0;
    auto take = [&](int n) { const int at = o; o += (n + 1) & ~1; return at; };
    A.o_in = take(L.n_in); A.o_out = take(L.n_out); A.o_out2 = take(L.n_out);
    // position-dependent block (saved/reloaded as one range)
    A.blob_begin = o;
    A.o_R = take(9 * nb); A.o_p = take(3 * nb); A.o_oM = take(3 * nb); A.o_a = take(3 * nq); A.o_l = take(3 * nq);
    const int nppx = npp + 2 * T.npathwrap; // path points + wrap tangent points
    A.npp_ext = nppx;
    A.o_cg = take(3 * nb); A.o_Io = take(6 * nb); A.o_ppos = take(3 * nppx); A.o_prel = take(3 * nppx);
    A.o_warc = take(T.npathwrap);                       // helix length of each pairing
    A.o_mpath = take(T.npathwrap > 0 ? 4 * T.nmus : 0); // current path of each muscle: 32 bytes = count + up to 31 entries
    A.o_dpl = take(3 * A.nmpp); A.o_mloc = take(3 * A.nmpp); A.o_ref = take(T.nppoly); A.o_mact = take(T.nmus); A.o_M = take(nq * nq);
    A.o_cjs = take(30 * T.ncustom);
    A.o_pk = take(L.prescribed ? 3 * nq : 0); // q, u, udot of prescribed kinematics
    A.blob_pos_end = o;
    // velocity-dependent block
    A.o_w = take(3 * nb); A.o_v = take(3 * nb); A.o_al = take(3 * nb); A.o_ac = take(3 * nb);
    A.o_cF = take(3 * T.ncontact); A.o_cP = take(3 * T.ncontact);
    A.o_mlen = take(T.nmus); A.o_mspd = take(T.nmus); A.o_mout = take(4 * T.nmus); // nominal muscle lengths, speeds, outputs
    A.blob_end = o;
    // per-evaluation scratch
    A.o_F = take(3 * nb); A.o_N = take(3 * nb); A.o_cm = take(nb); A.o_ch = take(3 * nb); A.o_cJ = take(6 * nb);
    A.o_tau = take(nq); A.o_taub = take(nq); A.o_lin = take(3 * nq); A.o_ang = take(3 * nq);
    A.o_pf = take(3 * nppx); A.o_dp = take(3 * A.nmpp);
    // right-hand sides / solutions of the local columns: reuse the wrench + composite-inertia scratch, dead by then
    A.o_Rm = ((T.nmus + T.ncact) * nq <= A.o_tau - A.o_F) ? A.o_F : take((T.nmus + T.ncact) * nq);
    A.n_lparts = std::max(1, (T.nmus + T.ncact + 31) / 32);
    { const char* e = devEnv("MOCOB200_LOCAL_PARTS"); if (e && atoi(e) > 0) A.n_lparts = std::min(A.n_lparts, atoi(e)); }
    A.ws_doubles = o;
    { const char* e = devEnv("MOCOB200_POISON_SCRATCH"); A.poison = (e && e[0] == '1') ? 1 : 0; }
    return warp_kernels_configure(A, err);
}

void defaultFd(FdSettings& fd, int scheme) {
    const double eps = 2.220446049250313e-16;
    fd.scheme = scheme;
    fd.reltol = eps;
    fd.abstol = eps;
    fd.u_aim = 100.0;
    if (scheme == MOCOB200_FD_CENTRAL) { fd.h = cbrt(eps); fd.h_iter = 1; }
    else { fd.h = sqrt(eps); fd.h_iter = 0; }
}

// x on the device (rank 0 holds the caller's x; other ranks receive it)
bool stageX(mocob200_handle* h, const double* x_host) {
    if (h->rank == 0) {
        if (!x_host) { h->err = "x is NULL"; return false; }
        if (!h->ck(cudaMemcpyAsync(h->d_x, x_host, sizeof(double) * h->L.n_var, cudaMemcpyHostToDevice, h->stream), "H2D x")) return false;
    }
    return true;
}
bool bcastX(mocob200_handle* h, const double* d_src) {
    if (h->rank == 0 && d_src != h->d_x) {
        if (!h->ck(cudaMemcpyAsync(h->d_x, d_src, sizeof(double) * h->L.n_var, cudaMemcpyDeviceToDevice, h->stream), "D2D x")) return false;
    }
    if (h->eff > 1 && h->rank < h->eff) { // only the ranks that own grid points need x
        if (!ncclOk(h, g_nccl.GroupStart(), "ncclGroupStart")) return false;
        int rc = 0;
        if (h->rank == 0) { for (int r = 1; r < h->eff && rc == 0; ++r) rc = g_nccl.Send(h->d_x, (size_t)h->L.n_var, kNcclDouble, r, h->comm, h->stream); }
        else rc = g_nccl.Recv(h->d_x, (size_t)h->L.n_var, kNcclDouble, 0, h->comm, h->stream);
        const int rcEnd = g_nccl.GroupEnd(); // always close the group, then report the first failure
        if (!ncclOk(h, rc, "ncclSend/ncclRecv(x)") || !ncclOk(h, rcEnd, "ncclGroupEnd")) return false;
    }
    return true;
}
// gather per-point blocks (stride doubles per grid point) to rank 0
bool gatherBlocks(mocob200_handle* h, double* d_buf, size_t stride) {
    if (h->eff == 1 || h->rank >= h->eff) return true;
    if (!ncclOk(h, g_nccl.GroupStart(), "ncclGroupStart")) return false;
    int rc = 0;
    if (h->rank == 0) {
        for (int r = 1; r < h->eff && rc == 0; ++r) {
            const size_t n = (size_t)(h->shard[r + 1] - h->shard[r]) * stride;
            if (n) rc = g_nccl.Recv(d_buf + (size_t)h->shard[r] * stride, n, kNcclDouble, r, h->comm, h->stream);
        }
    } else {
        const size_t n = (size_t)(h->p1 - h->p0) * stride;
        if (n) rc = g_nccl.Send(d_buf + (size_t)h->p0 * stride, n, kNcclDouble, 0, h->comm, h->stream);
    }
    const int rcEnd = g_nccl.GroupEnd();
    return ncclOk(h, rc, "ncclSend/ncclRecv(blocks)") && ncclOk(h, rcEnd, "ncclGroupEnd");
}

// stopwatch events also work inside a captured graph (as external event-record nodes)
void recTiming(cudaEvent_t e, cudaStream_t s) {
    cudaStreamCaptureStatus st = cudaStreamCaptureStatusNone;
    cudaStreamIsCapturing(s, &st);
    if (st == cudaStreamCaptureStatusActive) cudaEventRecordWithFlags(e, s, cudaEventRecordExternal);
    else cudaEventRecord(e, s);
}

// stages shared by the pipelines; x must already be in d_x
// Wrap cylinders are evaluated by the path kernel (mocob200_eval_points(..., MOCOB200_POINTS_PATHS)) and by the point functions'
// thread-level muscle code; the warp kernels' staged path geometry does not carry them yet: refuse rather than ignore them
bool dynamicsSupported(mocob200_handle* h) {
    if (h->DT.npathwrap > 0 && !h->wrapInWarpKernels) {
        h->err = "models with wrap cylinders: only path lengths / speeds (MOCOB200_POINTS_PATHS) are available on the device so far (SURVEY 8 a18)";
        return false;
    }
    return true;
}
bool runNominal(mocob200_handle* h) {
    if (!dynamicsSupported(h)) return false;
    if (h->rank >= h->eff) return true; // idle rank
    if (h->warpForValues(h->p1 - h->p0)) launch_wpoints_nominal(h->DT, h->WA, h->DL, h->d_x, h->d_Y, nullptr, h->p0, h->p1 - h->p0, h->stream);
    else launch_points_nominal(h->DT, h->DL, h->d_x, h->d_Y, h->p0, h->p1 - h->p0, h->stream);
    h->launches++;
    if (!gatherBlocks(h, h->d_Y, (size_t)h->L.n_out)) return false;
    if (h->rank == 0) {
        launch_integrals(h->DL, h->d_x, h->d_Y, h->d_J, h->d_I, false, h->stream);
        launch_endpoint(h->DT, h->DL, h->fd, h->d_x, h->d_I, h->d_G, h->d_E, false, h->stream);
        h->launches += (h->L.n_integrand > 0) + (h->DL.n_eout_total > 0);
    }
    return true;
}
// point-function FD kernels of this rank's grid points (Y and J rows [p0, p1)); after_nominal as in launch_wpoints_fd
void launchFdKernels(mocob200_handle* h, cudaEvent_t after_nominal) {
    if (h->warpFor(mocob200_point_evals_per_jacobian(h) / h->eff)) {
        h->launches += launch_wpoints_fd(h->DT, h->WA, h->DL, h->fd, h->d_x, h->d_Y, h->d_J, h->d_cfg, h->p0, h->p1 - h->p0, h->stream,
                                         after_nominal, &h->side);
        return;
    }
    // thread-per-evaluation kernel for the inputs that need whole evaluations; the local columns beside it
    const bool local = h->WA.n_local > 0 && h->side.s[0];
    if (local) {
        // warp nominal evaluation first (Y and the blobs), then the local-column kernel next to the flat FD kernel
        launch_wpoints_nominal(h->DT, h->WA, h->DL, h->d_x, h->d_Y, h->d_cfg, h->p0, h->p1 - h->p0, h->stream);
        if (after_nominal) cudaEventRecord(after_nominal, h->stream);
        cudaEventRecord(h->side.fork, h->stream);
        cudaStreamWaitEvent(h->side.s[0], h->side.fork, 0);
        h->launches += 1 + launch_wlocal(h->DT, h->WA, h->DL, h->fd, h->d_x, h->d_Y, h->d_J, h->d_cfg, h->p0, h->p1 - h->p0, h->side.s[0]);
        cudaEventRecord(h->side.join[0], h->side.s[0]);
        if (h->velKernel && h->WA.n_p1 > 0 && h->side.s[1]) {
            // pose-reusing directions on their own (lighter) kernel beside the whole-evaluation ones
            cudaStreamWaitEvent(h->side.s[1], h->side.fork, 0);
            launch_points_fd_vel(h->DT, h->WA, h->DL, h->fd, h->d_x, h->d_Y, h->d_J, h->d_cfg, h->p0, h->p1 - h->p0, h->side.s[1]);
            cudaEventRecord(h->side.join[1], h->side.s[1]);
            launch_points_fd_flat(h->DT, h->DL, h->fd, h->d_x, h->d_Y, h->d_J, h->p0, h->p1 - h->p0, h->stream, h->WA.cw_dirs, h->WA.n_cw);
            h->launches += 2;
            cudaStreamWaitEvent(h->stream, h->side.join[1], 0);
        } else {
            launch_points_fd_flat(h->DT, h->DL, h->fd, h->d_x, h->d_Y, h->d_J, h->p0, h->p1 - h->p0, h->stream, h->WA.cd_dirs, h->WA.n_cd);
            h->launches++;
        }
        cudaStreamWaitEvent(h->stream, h->side.join[0], 0);
    } else {
        launch_points_fd(h->DT, h->DL, h->fd, h->d_x, h->d_Y, h->d_J, h->p0, h->p1 - h->p0, h->stream);
        h->launches++;
        if (after_nominal) cudaEventRecord(after_nominal, h->stream);
    }
}

// deferIntegralDerivs: the caller (the jac_g pipeline) does not read dI/dt0, dI/dtf — they are computed on the side stream
// beside its assembly kernel and the caller joins on h->evI afterwards. Returns through *deferred whether that happened.
bool runFd(mocob200_handle* h, bool deferIntegralDerivs = false, bool* deferred = nullptr) {
    if (deferred) *deferred = false;
    if (!dynamicsSupported(h)) return false;
    recTiming(h->ev[0], h->stream);
    if (h->rank >= h->eff) { recTiming(h->ev[1], h->stream); return true; } // idle rank
    const bool overlap = h->eff == 1; // single rank: Y is complete as soon as the nominal evaluations are
    // jac_g pipeline without integrand-carrying endpoint constraints: the endpoint functions that read only x (periodicity,
    // initial activation ...) start now, beside the point kernels; the integrand goals' blocks feed grad_f only and run
    // beside the assembly (measured: waiting for one k_endpoint launch after the FD stage cost 14 us of a 131 us step)
    const bool split = overlap && deferIntegralDerivs && deferred && h->DL.n_eout_total > 0;
    if (split) {
        cudaEventRecord(h->evX, h->stream);
        cudaStreamWaitEvent(h->stream2, h->evX, 0);
        launch_endpoint(h->DT, h->DL, h->fd, h->d_x, h->d_I2, h->d_G, h->d_E, true, h->stream2, 1);
        cudaEventRecord(h->evE, h->stream2);
        h->launches++;
    }
    launchFdKernels(h, overlap ? h->evY : nullptr);
    recTiming(h->ev[1], h->stream);
    if (overlap) {
        cudaStreamWaitEvent(h->stream2, h->evY, 0);
        launch_integrals(h->DL, h->d_x, h->d_Y, h->d_J, h->d_I2, false, h->stream2);
        launch_endpoint(h->DT, h->DL, h->fd, h->d_x, h->d_I2, h->d_G, h->d_E, true, h->stream2, split ? 2 : 0);
        if (!split) cudaEventRecord(h->evE, h->stream2);
        if (deferIntegralDerivs && deferred) {
            cudaEventRecord(h->evJ, h->stream);
            cudaStreamWaitEvent(h->stream2, h->evJ, 0);
            launch_integrals(h->DL, h->d_x, h->d_Y, h->d_J, h->d_I, true, h->stream2);
            cudaEventRecord(h->evI, h->stream2); // everything queued on stream2 is done: the caller joins on it
            *deferred = true;
        } else launch_integrals(h->DL, h->d_x, h->d_Y, h->d_J, h->d_I, true, h->stream);
        cudaStreamWaitEvent(h->stream, h->evE, 0);
        h->launches += 2 * (h->L.n_integrand > 0) + (h->DL.n_eout_total > 0);
        return true;
    }
    if (!gatherBlocks(h, h->d_Y, (size_t)h->L.n_out)) return false;
    if (!gatherBlocks(h, h->d_J, (size_t)h->L.n_out * h->L.n_in)) return false;
    if (h->rank == 0) {
        launch_integrals(h->DL, h->d_x, h->d_Y, h->d_J, h->d_I, true, h->stream);
        launch_endpoint(h->DT, h->DL, h->fd, h->d_x, h->d_I, h->d_G, h->d_E, true, h->stream);
        h->launches += (h->L.n_integrand > 0) + (h->DL.n_eout_total > 0);
    }
    return true;
}
// values[k] -> host[k] for the entries the mask keeps (stores into the caller's pinned buffer over PCIe); any other entry that is
// not +0.0 raises the flag in host memory
__global__ void k_push_masked(const double* __restrict__ vals, const unsigned char* __restrict__ mask, double* __restrict__ host,
                              int* __restrict__ flag, int nnz) {
    const int k = blockIdx.x * blockDim.x + threadIdx.x;
    if (k >= nnz) return;
    const double v = vals[k];
    if (mask[k]) host[k] = v;
    else if (__double_as_longlong(v) != 0) *flag = 1;
}
bool finish(mocob200_handle* h, const char* what) {
    if (!h->ck(cudaStreamSynchronize(h->stream), what)) return false;
    return h->ck(cudaGetLastError(), what);
}
bool sameX(mocob200_handle* h, const double* x) {
    return h->rank == 0 && x && h->x_cached.size() == (size_t)h->L.n_var &&
           std::memcmp(h->x_cached.data(), x, sizeof(double) * h->L.n_var) == 0;
}
void rememberX(mocob200_handle* h, const double* x, bool withJ) {
    if (h->rank != 0 || !x) return;
    h->x_cached.assign(x, x + h->L.n_var);
    h->haveJ = withJ;
}

} // namespace

extern "C" {

int mocob200_nccl_unique_id(void* id128) {
    std::string err;
    if (!id128 || !g_nccl.load(err)) return MOCOB200_ERR_NCCL;
    NcclUid uid;
    if (g_nccl.GetUniqueId(&uid) != 0) return MOCOB200_ERR_NCCL;
    std::memcpy(id128, &uid, 128);
    return MOCOB200_OK;
}

void mocob200_destroy(mocob200_handle* h) {
    if (!h) return;
    cudaSetDevice(h->device);
    peerGatherClose(h);
    if (h->push.flag_h) cudaFreeHost(h->push.flag_h);
    if (h->comm && g_nccl.CommDestroy) g_nccl.CommDestroy(h->comm);
    for (void* p : h->owned) cudaFree(p);
    for (auto& e : h->ev) if (e) cudaEventDestroy(e);
    if (h->jacGraph) cudaGraphExecDestroy(h->jacGraph);
    for (auto& st : h->side.s) if (st) cudaStreamDestroy(st);
    if (h->side.fork) cudaEventDestroy(h->side.fork);
    for (auto& e : h->side.join) if (e) cudaEventDestroy(e);
    if (h->evY) cudaEventDestroy(h->evY);
    if (h->evE) cudaEventDestroy(h->evE);
    if (h->evJ) cudaEventDestroy(h->evJ);
    if (h->evX) cudaEventDestroy(h->evX);
    if (h->evBatch) cudaEventDestroy(h->evBatch);
    if (h->evCaller) cudaEventDestroy(h->evCaller);
    for (auto& e : h->marks) if (e) cudaEventDestroy(e);
    if (h->evI) cudaEventDestroy(h->evI);
    if (h->stream2) cudaStreamDestroy(h->stream2);
    if (h->stream) cudaStreamDestroy(h->stream);
    delete h;
}

static thread_local std::string g_create_err;
const char* mocob200_last_error(const mocob200_handle* h) { return h ? h->err.c_str() : g_create_err.c_str(); }

int mocob200_create(const mocob200_tables* T, const mocob200_transcription* R, int device, int rank, int world,
                    const void* nccl_unique_id, mocob200_handle** out) {
    if (!T || !R || !out || world < 1 || rank < 0 || rank >= world) { g_create_err = "bad argument"; return MOCOB200_ERR_ARG; }
    auto h = new mocob200_handle;
    h->device = device; h->rank = rank; h->world = world;
    auto fail = [&](int code, const std::string& m) { g_create_err = m.empty() ? h->err : m; mocob200_destroy(h); return code; };
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev <= 0)
        return fail(MOCOB200_ERR_CUDA, "no CUDA device: the evaluator has no CPU path");
    if (!h->ck(cudaSetDevice(device), "cudaSetDevice")) return fail(MOCOB200_ERR_CUDA, "");
    try {
        h->L = buildLayout(*T, *R, true);
    } catch (const std::exception& e) { return fail(MOCOB200_ERR_UNSUPPORTED, e.what()); }
    Layout& L = h->L;
    if (L.n_in > MOCOB200_MAX_IN || L.n_out > MOCOB200_MAX_OUT || L.n_in + 1 > 256 || L.n_ein + 1 > 512)
        return fail(MOCOB200_ERR_UNSUPPORTED, "point function exceeds the compiled input/output limits");
    for (int m = 0; m < T->nmus; ++m)
        if (T->idx[T->o_mus_i + m * MOCOB200_MUS_ISTRIDE + 1] > MOCOB200_MAXPPM)
            return fail(MOCOB200_ERR_UNSUPPORTED, "a muscle has more path points than the compiled limit");
    if (T->npathwrap > 0)
        for (int m = 0; m < T->nmus; ++m)
            if (T->idx[T->o_pw_i + 2 * m + 1] > MOCOB200_MAXWRAP)
                return fail(MOCOB200_ERR_UNSUPPORTED, "a muscle wraps over more cylinders than the compiled limit (4)");
    for (int g = 0; g < T->ngoal; ++g)
        if (T->idx[T->o_goal_i + g * MOCOB200_GOAL_ISTRIDE + 2] > MOCOB200_MAX_EOUT)
            return fail(MOCOB200_ERR_UNSUPPORTED, "a goal has more outputs than the compiled limit");
    if (L.n_integrand > 30) return fail(MOCOB200_ERR_UNSUPPORTED, "too many integrand goals");

    auto own = [&](void* p) { h->owned.push_back(p); return p; };
    h->hostDbl.assign(T->dbl, T->dbl + T->n_dbl);
    h->hostIdx.assign(T->idx, T->idx + T->n_int);
    h->hostT = *T; h->hostT.dbl = h->hostDbl.data(); h->hostT.idx = h->hostIdx.data();
    h->hostR = *R;
    if (R->mesh) { h->hostMesh.assign(R->mesh, R->mesh + R->n_mesh); h->hostR.mesh = h->hostMesh.data(); }
    // tables: the device works on the model with welded bodies merged into their parents (table_opt.cpp); the NLP layout,
    // the host copies and the ABI keep the caller's tables
    const mocob200_tables* const callerT = T;
    std::vector<double> dbl;
    std::vector<int> idx;
    mocob200_tables reduced;
    {
        const char* e = devEnv("MOCOB200_MERGE_WELDS"); // development A-B switch, read once at create
        if (e && e[0] == '0') { dbl.assign(callerT->dbl, callerT->dbl + callerT->n_dbl); idx.assign(callerT->idx, callerT->idx + callerT->n_int); reduced = *callerT; reduced.dbl = dbl.data(); reduced.idx = idx.data(); }
        else h->mergedBodies = mergeWeldedBodies(*callerT, dbl, idx, reduced);
    }
    T = &reduced;
    // constants of the muscle and contact models that depend on the parameters only, into the device copy (point_eval.cuh DEV_* slots)
    for (int m = 0; m < T->nmus; ++m) {
        double* md = dbl.data() + T->o_mus_d + m * MOCOB200_MUS_DSTRIDE;
        const double kPE = 4.0, offset = std::exp(kPE * (0.2 - 1.0) / md[9]);
        const double fiberWidth = md[1] * std::sin(md[3]);
        md[DEV_MUS_PE_DENOM] = std::exp(kPE) - offset;
        md[DEV_MUS_FIBER_WIDTH] = fiberWidth;
        md[DEV_MUS_PE_OFFSET] = offset;
    }
    for (int c = 0; c < T->ncontact; ++c) {
        double* cd = dbl.data() + T->o_contact_d + c * MOCOB200_CONTACT_DSTRIDE;
        const double k = 0.5 * std::pow(cd[16], 2.0 / 3.0);
        cd[DEV_CONTACT_HERTZ] = (4.0 / 3.0) * k * std::sqrt(cd[3] * k);
    }
    if (T->nb > SizeHuge::NB || T->nq > SizeHuge::NQ || T->ncontact > MOCOB200_MAXCONTACT)
        return fail(MOCOB200_ERR_UNSUPPORTED, "model exceeds the compiled limits (bodies/coordinates/contacts)");
    // beyond the thread-per-evaluation kernels' arrays (16 bodies / 16 coordinates / 160 inputs) only the warp kernels apply
    h->beyondThreadKernels = T->nb > MOCOB200_MAXB || T->nq > MOCOB200_MAXQ || L.n_in > SizeLarge::NIN || L.n_out > SizeLarge::NOUT || T->npathwrap > 0;
    if (h->beyondThreadKernels)
        for (int g = 0; g < T->ngoal; ++g) {
            const int* gr = T->idx + T->o_goal_i + g * MOCOB200_GOAL_ISTRIDE;
            const bool disp = gr[0] == MOCOB200_GOAL_AVERAGE_SPEED || gr[0] == MOCOB200_GOAL_INITIAL_VELOCITY_EQUILIBRIUM_DGF ||
                              (gr[0] == MOCOB200_GOAL_CONTROL && T->idx[gr[4] + 1]);
            if (disp) return fail(MOCOB200_ERR_UNSUPPORTED, "goals that evaluate the model's kinematics at the end points are limited to 16 bodies / 16 coordinates");
        }
    DevTables& D = h->DT;
    D.dbl = (const double*)own(devUpload(dbl));
    D.idx = (const int*)own(devUpload(idx));
    D.dblP = D.dbl;
    D.dblG = D.dbl;
    D.nb = T->nb; D.nq = T->nq; D.nz = T->nz; D.nc = T->nc; D.nderiv_aux = T->nderiv_aux; D.nmus = T->nmus; D.npp = T->npp;
    D.nspline = T->nspline; D.ncontact = T->ncontact; D.ncact = T->ncact; D.ngoal = T->ngoal; D.nppoly = T->nppoly;
    D.o_body_d = T->o_body_d; D.o_gravity = T->o_gravity; D.o_pp_d = T->o_pp_d; D.o_mus_d = T->o_mus_d;
    D.o_contact_d = T->o_contact_d; D.o_cact_d = T->o_cact_d; D.o_spline_data = T->o_spline_data; D.o_ppoly_data = T->o_ppoly_data;
    D.o_body_i = T->o_body_i; D.o_pp_i = T->o_pp_i; D.o_mus_i = T->o_mus_i; D.o_contact_i = T->o_contact_i; D.o_cact_i = T->o_cact_i;
    D.o_spline_i = T->o_spline_i; D.o_goal_i = T->o_goal_i; D.o_ppoly_i = T->o_ppoly_i;
    D.ncustom = T->ncustom; D.o_cj_d = T->o_cj_d; D.o_cj_i = T->o_cj_i; D.o_body_cj = T->o_body_cj;
    D.nxforce = T->nxforce; D.o_xforce_i = T->o_xforce_i;
    D.ncoupler = T->ncoupler; D.o_cc_i = T->o_cc_i; D.o_cc_d = T->o_cc_d;
    D.nm = L.nm; D.n_kc = L.n_kc; D.enforce_derivs = L.enforce_derivs;
    D.prescribed = L.prescribed; D.o_presc_i = T->o_presc_i;
    D.nwrapobj = T->nwrapobj; D.o_wo_d = T->o_wo_d; D.o_wo_i = T->o_wo_i; D.npathwrap = T->npathwrap; D.o_pw_i = T->o_pw_i;
    D.implicit = L.implicit; D.n_auxres = L.n_auxres; D.n_integrand = L.n_integrand;
    // layout
    DevLayout& G = h->DL;
    G.prescribed = L.prescribed; G.nqs = L.nqs;
    G.scheme = L.scheme; G.implicit = L.implicit; G.nq = L.nq; G.nz = L.nz; G.ns = L.ns; G.nc = L.nc; G.nd = L.nd; G.n_in = L.n_in;
    G.n_auxres = L.n_auxres; G.n_integrand = L.n_integrand; G.n_dyn_out = L.n_dyn_out; G.n_out = L.n_out;
    G.n_mesh = L.n_mesh; G.n_grid = L.n_grid; G.n_intervals = L.n_intervals; G.n_var = L.n_var; G.n_con = L.n_con; G.nnz = L.nnz;
    G.off_states = L.off_states; G.off_controls = L.off_controls; G.off_derivs = L.off_derivs;
    G.nm = L.nm; G.nsl = L.nsl; G.n_kc = L.n_kc; G.off_mult = L.off_mult; G.off_slacks = L.off_slacks;
    G.minimize_mult = (L.nm > 0 && R->minimize_lagrange_multipliers) ? 1 : 0; G.mult_weight = R->lagrange_multiplier_weight;
    G.n_endpoint_rows = L.n_endpoint_rows; G.n_resid_rows_per_point = L.n_resid_rows_per_point; G.n_defect_rows = L.n_defect_rows;
    G.n_interp_rows = L.n_interp_rows; G.n_ein = L.n_ein; G.n_cost_terms = L.n_cost_terms;
    G.minimize_accel = R->minimize_accelerations; G.minimize_auxderiv = R->minimize_auxiliary_derivatives;
    G.accel_weight = R->accelerations_weight; G.auxderiv_weight = R->auxiliary_derivatives_weight;
    std::vector<int> eoff(T->ngoal + 1, 0);
    for (int g = 0; g < T->ngoal; ++g) eoff[g + 1] = eoff[g] + T->idx[T->o_goal_i + g * MOCOB200_GOAL_ISTRIDE + 2];
    G.n_eout_total = eoff[T->ngoal];
    G.n_path_rows = L.n_path_rows;
    {
        int maxPpm = 0;
        for (int m = 0; m < T->nmus; ++m) maxPpm = std::max(maxPpm, T->idx[T->o_mus_i + m * MOCOB200_MUS_ISTRIDE + 1]);
        G.small_size = (T->nb <= SizeSmall::NB && T->nq <= SizeSmall::NQ && maxPpm <= SizeSmall::NPPM && L.n_in <= SizeSmall::NIN &&
                        L.n_out <= SizeSmall::NOUT) ? 1 : 0;
    }
    G.mesh = (const double*)own(devUpload(L.mesh));
    G.grid = (const double*)own(devUpload(L.grid));
    G.quad = (const double*)own(devUpload(L.quad));
    G.row_resid = (const int*)own(devUpload(L.row_resid));
    G.row_defect = (const int*)own(devUpload(L.row_defect));
    G.row_interp = (const int*)own(devUpload(L.row_interp));
    G.goal_row = (const int*)own(devUpload(L.goal_row));
    G.goal_integrand = (const int*)own(devUpload(L.goal_integrand));
    G.goal_eoff = (const int*)own(devUpload(eoff));
    static_assert(sizeof(RowMeta) == sizeof(RowMetaD) && sizeof(ColMeta) == sizeof(ColMetaD), "meta layouts");
    {
        RowMetaD* p = devAlloc<RowMetaD>(L.rows.size());
        if (p) cudaMemcpy(p, L.rows.data(), L.rows.size() * sizeof(RowMeta), cudaMemcpyHostToDevice);
        G.rows = (const RowMetaD*)own(p);
        ColMetaD* c = devAlloc<ColMetaD>(L.cols.size());
        if (c) cudaMemcpy(c, L.cols.data(), L.cols.size() * sizeof(ColMeta), cudaMemcpyHostToDevice);
        G.cols = (const ColMetaD*)own(c);
    }
    G.colptr = (const int*)own(devUpload(L.colptr));
    G.rowidx = (const int*)own(devUpload(L.rowidx));
    G.nnzcol = (const int*)own(devUpload(L.nnzcol));
    {
        const char* k = devEnv("MOCOB200_KERNEL");
        h->kernelMode = (k && std::string(k) == "thread") ? 2 : (k && std::string(k) == "warp") ? 1 : 0;
        // the thread-per-evaluation dynamics handle the constant-axis joints only: function-based joints run on the warp kernels
        if (T->ncustom > 0 || h->beyondThreadKernels) {
            if (h->kernelMode == 2)
                return fail(MOCOB200_ERR_UNSUPPORTED, "the thread-per-evaluation kernels do not support custom joints or models beyond 16 bodies / 16 coordinates / 160 inputs");
            h->kernelMode = 1;
        }
        std::string e;
        if (!buildWarpAux(h, *T, e)) return fail(MOCOB200_ERR_UNSUPPORTED, "warp kernels: " + e);
    }
    // work buffers
    h->d_x = (double*)own(devAlloc<double>(L.n_var));
    h->d_Y = (double*)own(devAlloc<double>((size_t)L.n_grid * L.n_out));
    h->d_J = (double*)own(devAlloc<double>((size_t)L.n_grid * L.n_out * L.n_in));
    h->d_cfg = (double*)own(devAlloc<double>((size_t)L.n_grid * (h->WA.blob_end - h->WA.blob_begin) + 1));
    h->d_I = (double*)own(devAlloc<double>(3 * (size_t)L.n_integrand + 1));
    h->d_I2 = (double*)own(devAlloc<double>(3 * (size_t)L.n_integrand + 1));
    h->d_G = (double*)own(devAlloc<double>(G.n_eout_total + 1));
    h->d_E = (double*)own(devAlloc<double>((size_t)(G.n_eout_total + 1) * L.n_ein));
    h->d_g = (double*)own(devAlloc<double>(L.n_con));
    h->d_vals = (double*)own(devAlloc<double>(L.nnz));
    h->d_grad = (double*)own(devAlloc<double>(L.n_var));
    h->d_f = (double*)own(devAlloc<double>(2 + L.n_cost_terms));
    for (void* p : h->owned)
        if (!p) return fail(MOCOB200_ERR_CUDA, "device allocation failed");
    if (!h->ck(cudaGetLastError(), "upload")) return fail(MOCOB200_ERR_CUDA, "");
    if (!h->ck(cudaStreamCreateWithFlags(&h->stream, cudaStreamNonBlocking), "stream")) return fail(MOCOB200_ERR_CUDA, "");
    if (!h->ck(cudaStreamCreateWithFlags(&h->stream2, cudaStreamNonBlocking), "stream")) return fail(MOCOB200_ERR_CUDA, "");
    for (auto& e : h->ev) cudaEventCreate(&e);
    for (auto& st : h->side.s) cudaStreamCreateWithFlags(&st, cudaStreamNonBlocking);
    cudaEventCreateWithFlags(&h->side.fork, cudaEventDisableTiming);
    for (auto& e : h->side.join) cudaEventCreateWithFlags(&e, cudaEventDisableTiming);
    cudaEventCreateWithFlags(&h->evY, cudaEventDisableTiming);
    cudaEventCreateWithFlags(&h->evE, cudaEventDisableTiming);
    cudaEventCreateWithFlags(&h->evJ, cudaEventDisableTiming);
    cudaEventCreateWithFlags(&h->evX, cudaEventDisableTiming);
    cudaEventCreateWithFlags(&h->evBatch, cudaEventDisableTiming);
    cudaEventCreateWithFlags(&h->evI, cudaEventDisableTiming);
    for (int g = 0; g < T->ngoal; ++g) {
        const int* gr = T->idx + T->o_goal_i + g * MOCOB200_GOAL_ISTRIDE;
        if (gr[1] == MOCOB200_MODE_ENDPOINT_CONSTRAINT && gr[3]) h->ecHasIntegrand = true;
    }
    defaultFd(h->fd, R->fd_scheme);
    // shard grid points evenly over ranks
    h->shard.resize(world + 1);
    h->shard[0] = 0;
    for (int r = 0; r < world; ++r) { int a = 0, b = 0; mocob200_shard_range(L.n_grid, r, world, &a, &b); h->shard[r + 1] = b; }
    h->p0 = h->shard[rank];
    h->p1 = h->shard[rank + 1];
    h->eff = mocob200_effective_world(L.n_grid, world);
    if (h->eff > 1) {
        h->d_Jt = (double*)own(devAlloc<double>((size_t)L.n_grid * L.n_out));
        if (!h->d_Jt) return fail(MOCOB200_ERR_CUDA, "device allocation failed");
        h->DLt = h->DL;
        h->DLt.Jt = h->d_Jt;
        h->runs.assign((size_t)world * 6, 0);
        for (int r = 0; r < h->eff; ++r) {
            const int a = h->shard[r], b = h->shard[r + 1];
            const int c0[3] = {L.off_states + a * L.ns, L.off_controls + a * L.nc, L.off_derivs + a * L.nd};
            const int c1[3] = {L.off_states + b * L.ns, L.off_controls + b * L.nc, L.off_derivs + b * L.nd};
            for (int k = 0; k < 3; ++k) { h->runs[6 * r + 2 * k] = L.colptr[c0[k]]; h->runs[6 * r + 2 * k + 1] = L.colptr[c1[k]]; }
        }
    }
    { const char* e = devEnv("MOCOB200_VEL_KERNEL"); h->velKernel = !(e && e[0] == '0'); }
    h->allowGraph = !(devEnv("MOCOB200_GRAPH") && devEnv("MOCOB200_GRAPH")[0] == '0') && !devEnv("MOCOB200_TIMING");
    if (h->eff > 1) { // a communicator only when more than one rank owns grid points
        if (!nccl_unique_id) return fail(MOCOB200_ERR_ARG, "world > 1 needs a NCCL unique id");
        std::string e;
        if (!g_nccl.load(e)) return fail(MOCOB200_ERR_NCCL, e);
        NcclUid uid;
        std::memcpy(&uid, nccl_unique_id, 128);
        if (!ncclOk(h, g_nccl.CommInitRank(&h->comm, world, uid, rank), "ncclCommInitRank")) return fail(MOCOB200_ERR_NCCL, "");
    }
    *out = h;
    return MOCOB200_OK;
}

int mocob200_get_dims(const mocob200_handle* h, mocob200_dims* d) {
    if (!h || !d) return MOCOB200_ERR_ARG;
    const Layout& L = h->L;
    std::memset(d, 0, sizeof *d);
    d->n_grid = L.n_grid; d->n_mesh = L.n_mesh; d->n_var = L.n_var; d->n_con = L.n_con; d->nnz_jac = L.nnz;
    d->ns = L.ns; d->nc = L.nc; d->nd = L.nd; d->n_in = L.n_in; d->n_out = L.n_out;
    d->n_dyn_out = L.n_dyn_out; d->n_integrand = L.n_integrand; d->n_cost_terms = L.n_cost_terms;
    d->first_interval = h->p0; d->n_intervals_local = h->p1 - h->p0; d->rank = h->rank; d->world = h->world;
    return MOCOB200_OK;
}

int mocob200_set_fd(mocob200_handle* h, int scheme, double step, int h_iter, double u_aim) {
    if (!h || scheme < 0 || scheme > 2) return MOCOB200_ERR_ARG;
    defaultFd(h->fd, scheme);
    if (step > 0) h->fd.h = step;
    if (h_iter >= 0) h->fd.h_iter = h_iter;
    if (u_aim > 0) h->fd.u_aim = u_aim;
    if (h->jacGraph) { cudaSetDevice(h->device); cudaStreamSynchronize(h->stream); cudaGraphExecDestroy(h->jacGraph); h->jacGraph = nullptr; }
    h->x_cached.clear();
    h->push.tried = h->push.on = false; // the sign of a zero quotient depends on the scheme: find the mask again
    return MOCOB200_OK;
}

// optim_sparsity_detection (MocoCasADiSolver.h:139-147): "none" = dense callback blocks; "random" = the per-callback block
// sparsity found by perturbing every input (eps = 1e-5, entry kept when the output changes or is NaN) at random iterates
// within the bounds, united over the iterates (CasOCFunction.cpp:25-105, CasOCSolver.cpp:71-87,
// CasOCTranscription.cpp:1151-1177). The reference perturbs at the first time point of each iterate; here every grid
// point of each iterate votes (a superset, so never fewer entries than the reference would keep).
// host copy of ppoly_eval (point_eval.cuh) for the set-up time sparsity probe of the path constraints
static double hostPPoly(const mocob200_tables& T, int id, double t) {
    const int* rec = T.idx + T.o_ppoly_i + id * MOCOB200_PPOLY_ISTRIDE;
    const int np = rec[1], order = rec[2];
    const double* br = T.dbl + T.o_ppoly_data + rec[0];
    const double* co = br + np + 1;
    int k = 0;
    while (k + 1 < np && t >= br[k + 1]) ++k;
    const double s = t - br[k];
    double v = co[(size_t)k * order + order - 1];
    for (int j = order - 2; j >= 0; --j) v = v * s + co[(size_t)k * order + j];
    return v;
}

int mocob200_set_sparsity_detection(mocob200_handle* h, int mode, int n_iterates) {
    if (!h || mode < 0 || mode > 1) return MOCOB200_ERR_ARG;
    if (h->world != 1) { h->err = "sparsity detection is single-rank"; return MOCOB200_ERR_UNSUPPORTED; }
    cudaSetDevice(h->device);
    if (!finish(h, "set_sparsity_detection")) return MOCOB200_ERR_CUDA;
    const Layout& L0 = h->L;
    std::vector<unsigned char> mask;
    if (mode == 1) {
        if (n_iterates <= 0) n_iterates = 3;
        mask.assign((size_t)(L0.n_out + L0.n_path_rows) * L0.n_in, 0);
        const FdSettings keep = h->fd;
        defaultFd(h->fd, MOCOB200_FD_FORWARD);
        h->fd.h = 1e-5;
        std::vector<double> x(L0.n_var), J((size_t)L0.n_grid * L0.n_out * L0.n_in);
        bool ok = true;
        for (int k = 0; k < n_iterates && ok; ++k) {
            mocob200_make_iterate(&h->hostT, &h->hostR, k, 1.0, x.data());
            ok = stageX(h, x.data()) && bcastX(h, h->d_x) && runFd(h) &&
                 h->ck(cudaMemcpyAsync(J.data(), h->d_J, sizeof(double) * J.size(), cudaMemcpyDeviceToHost, h->stream), "D2H J") &&
                 finish(h, "sparsity detection");
            if (!ok) break;
            const size_t blk = (size_t)L0.n_out * L0.n_in;
            for (int j = 0; j < L0.n_grid; ++j)
                for (size_t e = 0; e < blk; ++e) {
                    const double v = J[(size_t)j * blk + e];
                    if (v != 0.0 || v != v) mask[e] = 1;
                }
            // path-constraint callbacks: same perturbation test at the iterate's first time point (CasOCFunction.cpp:25-74)
            const mocob200_tables& T = h->hostT;
            for (int g = 0; g < T.ngoal; ++g) {
                if (L0.goal_path_off[g] < 0) continue;
                const int* gr = T.idx + T.o_goal_i + g * MOCOB200_GOAL_ISTRIDE;
                const int* ip = T.idx + gr[4];
                for (int e = 0; e < gr[2]; ++e) {
                    unsigned char* row = mask.data() + (size_t)(L0.n_out + L0.goal_path_off[g] + e) * L0.n_in;
                    const double u = x[L0.off_controls + ip[2 * e]], b0 = hostPPoly(T, ip[2 * e + 1], x[0]);
                    const double dt = (u - hostPPoly(T, ip[2 * e + 1], x[0] + 1e-5)) - (u - b0);
                    const double du = ((u + 1e-5) - b0) - (u - b0);
                    if (dt != 0.0 || dt != dt) row[0] = 1;
                    if (du != 0.0 || du != du) row[1 + L0.ns + ip[2 * e]] = 1;
                }
            }
        }
        h->fd = keep;
        if (!ok) return MOCOB200_ERR_CUDA;
    }
    Layout L;
    try {
        L = buildLayout(h->hostT, h->hostR, true, mode == 1 ? mask.data() : nullptr);
    } catch (const std::exception& e) { h->err = e.what(); return MOCOB200_ERR_UNSUPPORTED; }
    h->L.colptr = L.colptr; h->L.rowidx = L.rowidx; h->L.nnzcol = L.nnzcol; h->L.nnz = L.nnz;
    int* c = devUpload(L.colptr); int* r = devUpload(L.rowidx); int* n = devUpload(L.nnzcol);
    double* v = devAlloc<double>(L.nnz);
    if (!c || !r || !n || !v) { h->err = "device allocation failed"; return MOCOB200_ERR_CUDA; }
    h->owned.push_back(c); h->owned.push_back(r); h->owned.push_back(n); h->owned.push_back(v);
    h->DL.colptr = c; h->DL.rowidx = r; h->DL.nnzcol = n; h->DL.nnz = L.nnz;
    h->d_vals = v;
    peerGatherClose(h); // the value array the other ranks had mapped is no longer the one in use: set up again at the next Jacobian
    if (h->jacGraph) { cudaGraphExecDestroy(h->jacGraph); h->jacGraph = nullptr; }
    h->x_cached.clear();
    h->push.tried = h->push.on = false; // another value array: the mask of the host delivery is found again at the next call
    return MOCOB200_OK;
}

int mocob200_jac_g_structure(const mocob200_handle* h, int* colptr, int* rowidx) {
    if (!h || !colptr || !rowidx) return MOCOB200_ERR_ARG;
    std::memcpy(colptr, h->L.colptr.data(), sizeof(int) * (h->L.n_var + 1));
    std::memcpy(rowidx, h->L.rowidx.data(), sizeof(int) * h->L.nnz);
    return MOCOB200_OK;
}

int mocob200_eval_points(mocob200_handle* h, int what, const double* x, double* out) {
    if (!h) return MOCOB200_ERR_ARG;
    cudaSetDevice(h->device);
    if (!stageX(h, x) || !bcastX(h, h->d_x)) return MOCOB200_ERR_CUDA;
    const Layout& L = h->L;
    if (what == MOCOB200_POINTS_OUTPUTS) {
        if (!runNominal(h)) return MOCOB200_ERR_CUDA;
        if (h->rank == 0 && out) cudaMemcpyAsync(out, h->d_Y, sizeof(double) * L.n_grid * L.n_out, cudaMemcpyDeviceToHost, h->stream);
        rememberX(h, nullptr, false);
        h->x_cached.clear();
    } else if (what == MOCOB200_POINTS_PATHS) {
        if (h->world != 1) { h->err = "MOCOB200_POINTS_PATHS is single-rank"; return MOCOB200_ERR_UNSUPPORTED; }
        const size_t n = (size_t)L.n_grid * h->DT.nmus * 2;
        double* d_out = nullptr; // a diagnostic output: allocated per call
        if (n == 0) return MOCOB200_OK;
        if (!h->ck(cudaMalloc(&d_out, n * sizeof(double)), "cudaMalloc paths")) return MOCOB200_ERR_CUDA;
        launch_paths(h->DT, h->DL, h->d_x, d_out, 0, L.n_grid, h->stream);
        h->launches++;
        if (out) cudaMemcpyAsync(out, d_out, n * sizeof(double), cudaMemcpyDeviceToHost, h->stream);
        const bool ok = finish(h, "eval_points(paths)");
        cudaFree(d_out);
        return ok ? MOCOB200_OK : MOCOB200_ERR_CUDA;
    } else {
        if (!runFd(h)) return MOCOB200_ERR_CUDA;
        if (h->rank == 0 && out) cudaMemcpyAsync(out, h->d_J, sizeof(double) * L.n_grid * L.n_out * L.n_in, cudaMemcpyDeviceToHost, h->stream);
        h->x_cached.clear();
    }
    return finish(h, "eval_points") ? MOCOB200_OK : MOCOB200_ERR_CUDA;
}

int mocob200_eval_g(mocob200_handle* h, const double* x, double* g) {
    if (!h) return MOCOB200_ERR_ARG;
    cudaSetDevice(h->device);
    // IPOPT asks for f, g, grad_f and jac_g at the same x one after the other: the point outputs and endpoint values of
    // the last evaluation are reused when x has not changed (x_cached non-empty <=> d_Y, d_G hold the values at x_cached)
    const bool reuse = h->eff == 1 && sameX(h, x);
    if (!reuse && (!stageX(h, x) || !bcastX(h, h->d_x) || !runNominal(h))) return MOCOB200_ERR_CUDA;
    if (h->rank == 0) {
        launch_assemble_g(h->DT, h->DL, h->d_x, h->d_Y, h->d_G, h->d_g, h->stream);
        h->launches += 1 + (h->DL.n_path_rows > 0);
        if (g) cudaMemcpyAsync(g, h->d_g, sizeof(double) * h->L.n_con, cudaMemcpyDeviceToHost, h->stream);
    }
    if (!reuse) h->x_cached.clear();
    if (!finish(h, "eval_g")) return MOCOB200_ERR_CUDA;
    if (!reuse && h->eff == 1) rememberX(h, x, false);
    return MOCOB200_OK;
}

int mocob200_eval_f(mocob200_handle* h, const double* x, double* f, double* f_terms) {
    if (!h) return MOCOB200_ERR_ARG;
    cudaSetDevice(h->device);
    const bool reuse = h->eff == 1 && sameX(h, x);
    if (!reuse && (!stageX(h, x) || !bcastX(h, h->d_x) || !runNominal(h))) return MOCOB200_ERR_CUDA;
    std::vector<double> host(1 + h->L.n_cost_terms, 0.0);
    if (h->rank == 0) {
        launch_assemble_f(h->DT, h->DL, h->d_x, h->d_G, h->d_f, h->stream);
        h->launches++;
        cudaMemcpyAsync(host.data(), h->d_f, sizeof(double) * host.size(), cudaMemcpyDeviceToHost, h->stream);
    }
    if (!reuse) h->x_cached.clear();
    if (!finish(h, "eval_f")) return MOCOB200_ERR_CUDA;
    if (!reuse && h->eff == 1) rememberX(h, x, false);
    if (h->rank == 0) {
        if (f) *f = host[0];
        if (f_terms) for (int k = 0; k < h->L.n_cost_terms; ++k) f_terms[k] = host[1 + k];
    }
    return MOCOB200_OK;
}

int mocob200_eval_grad_f(mocob200_handle* h, const double* x, double* grad) {
    if (!h) return MOCOB200_ERR_ARG;
    cudaSetDevice(h->device);
    const bool reuse = h->eff == 1 && h->haveJ && sameX(h, x);
    if (!reuse) {
        if (!stageX(h, x) || !bcastX(h, h->d_x) || !runFd(h)) return MOCOB200_ERR_CUDA;
    }
    if (h->rank == 0) {
        launch_assemble_grad(h->DT, h->DL, h->d_x, h->d_Y, h->d_J, h->d_I, h->d_E, h->d_grad, h->stream);
        h->launches++;
        if (grad) cudaMemcpyAsync(grad, h->d_grad, sizeof(double) * h->L.n_var, cudaMemcpyDeviceToHost, h->stream);
    }
    if (!finish(h, "eval_grad_f")) return MOCOB200_ERR_CUDA;
    rememberX(h, x, true);
    return MOCOB200_OK;
}

// ---- peer-memory gather -------------------------------------------------------------------------------------------------
// Flag words live in rank 0's memory. flags[0] = s - 1 is stored by rank 0 when it STARTS Jacobian s (stream-ordered after
// whatever consumed Jacobian s - 1): from then on the other ranks may overwrite its buffers. flags[r] = s is stored by rank r
// after its last store of Jacobian s (a kernel boundary and a system-scope fence precede it); rank 0 waits for all of them
// before it reads what they wrote. Waits are bounded (10 s): a rank that never arrives traps the context - loud, not a hang.
__global__ void k_peer_signal(unsigned long long* flag, unsigned long long v) {
    __threadfence_system();
    *(volatile unsigned long long*)flag = v;
    __threadfence_system();
}
__global__ void k_peer_wait(const unsigned long long* flags, int first, int count, unsigned long long v) {
    const int i = threadIdx.x;
    if (i < count) {
        const volatile unsigned long long* f = flags + first + i;
        unsigned long long t0, t1;
        asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t0));
        while (*f < v) {
            __nanosleep(200);
            asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t1));
            if (t1 - t0 > 10000000000ull) __trap();
        }
    }
    __threadfence_system();
}
static void peerGatherClose(mocob200_handle* h) {
    auto& P = h->peer;
    if (P.tracing) { // segments of the last Jacobian on this rank
        cudaStreamSynchronize(h->stream);
        const char* name[6] = {"FD stage", "wait for rank 0", "pack + endpoint join + own runs", "rank 0: wait for the others / rank r: Y copy + signal", "t0 / tf columns", "D2D of the others' runs"};
        for (int k = 0; k < 6; ++k) {
            float ms = 0;
            if (cudaEventElapsedTime(&ms, P.trace[k], P.trace[k + 1]) == cudaSuccess) fprintf(stderr, "[peer trace] rank %d  %-58s %.4f ms\n", h->rank, name[k], ms);
            else cudaGetLastError();
        }
        for (auto& e : P.trace) cudaEventDestroy(e);
    }
    if (h->rank > 0) {
        if (P.vals) cudaIpcCloseMemHandle(P.vals);
        if (P.Y) cudaIpcCloseMemHandle(P.Y);
        if (P.Jt) cudaIpcCloseMemHandle(P.Jt);
        if (P.flags) cudaIpcCloseMemHandle(P.flags);
    } else if (P.flags) cudaFree(P.flags);
    if (P.d_msg) cudaFree(P.d_msg);
    P = mocob200_handle::PeerGather{};
}
// Collective over the working ranks (rank < eff), once per handle and value array: rank 0 exports its buffers, the others map
// them, and all agree on the outcome (if any rank cannot, every rank keeps the NCCL gather). MOCOB200_P2P=0 (development
// switch) declines on the rank that reads it, hence on all.
static bool peerGatherSetup(mocob200_handle* h) {
    auto& P = h->peer;
    P.tried = true;
    P.on = false;
    P.seq = 0;
    const int MSG = 40; // doubles per message: 4 IPC handles (4 x 64 bytes = 32 doubles), [32] = the sender's verdict
    static_assert(sizeof(cudaIpcMemHandle_t) == 64, "IPC handle size");
    if (!P.d_msg && cudaMalloc(&P.d_msg, sizeof(double) * MSG * (size_t)(h->world + 1)) != cudaSuccess) { cudaGetLastError(); P.d_msg = nullptr; }
    const char* e = devEnv("MOCOB200_P2P");
    bool mine = P.d_msg && !(e && e[0] == '0');
    if (!P.d_msg) { h->err = "peer gather: device allocation failed"; return false; } // the exchange itself needs the buffer
    std::vector<double> host((size_t)MSG * (h->world + 1), 0.0);
    auto exchange = [&](bool fromZero, int count) { // rank 0 <-> every other working rank, `count` doubles of slot r (rank 0: slot 0 out)
        if (!ncclOk(h, g_nccl.GroupStart(), "ncclGroupStart")) return false;
        int rc = 0;
        if (h->rank == 0) {
            for (int r = 1; r < h->eff && rc == 0; ++r)
                rc = fromZero ? g_nccl.Send(P.d_msg, (size_t)count, kNcclDouble, r, h->comm, h->stream)
                              : g_nccl.Recv(P.d_msg + (size_t)MSG * r, (size_t)count, kNcclDouble, r, h->comm, h->stream);
        } else {
            rc = fromZero ? g_nccl.Recv(P.d_msg, (size_t)count, kNcclDouble, 0, h->comm, h->stream)
                          : g_nccl.Send(P.d_msg + (size_t)MSG * h->rank, (size_t)count, kNcclDouble, 0, h->comm, h->stream);
        }
        const int rcEnd = g_nccl.GroupEnd();
        return ncclOk(h, rc, "ncclSend/ncclRecv(peer setup)") && ncclOk(h, rcEnd, "ncclGroupEnd") &&
               h->ck(cudaStreamSynchronize(h->stream), "peer setup sync");
    };
    // 1. rank 0 -> all: handles of d_vals, d_Y, d_Jt, flags
    if (h->rank == 0) {
        if (mine && !P.flags) {
            if (cudaMalloc(&P.flags, 64 * sizeof(unsigned long long)) != cudaSuccess) { cudaGetLastError(); P.flags = nullptr; mine = false; }
        }
        if (P.flags) cudaMemsetAsync(P.flags, 0, 64 * sizeof(unsigned long long), h->stream);
        cudaIpcMemHandle_t hd[4];
        void* ptrs[4] = {h->d_vals, h->d_Y, h->d_Jt, P.flags};
        for (int k = 0; k < 4 && mine; ++k)
            if (!ptrs[k] || cudaIpcGetMemHandle(&hd[k], ptrs[k]) != cudaSuccess) { cudaGetLastError(); mine = false; }
        if (mine) std::memcpy(host.data(), hd, sizeof hd);
        host[32] = mine ? 1.0 : 0.0;
        if (!h->ck(cudaMemcpyAsync(P.d_msg, host.data(), sizeof(double) * MSG, cudaMemcpyHostToDevice, h->stream), "peer setup H2D")) return false;
    }
    if (!exchange(true, MSG)) return false;
    // 2. the others map them and report
    if (h->rank > 0) {
        if (!h->ck(cudaMemcpy(host.data(), P.d_msg, sizeof(double) * MSG, cudaMemcpyDeviceToHost), "peer setup D2H")) return false;
        mine = mine && host[32] == 1.0;
        if (mine) {
            cudaIpcMemHandle_t hd[4];
            std::memcpy(hd, host.data(), sizeof hd);
            void* ptrs[4] = {nullptr, nullptr, nullptr, nullptr};
            for (int k = 0; k < 4 && mine; ++k)
                if (cudaIpcOpenMemHandle(&ptrs[k], hd[k], cudaIpcMemLazyEnablePeerAccess) != cudaSuccess) { cudaGetLastError(); ptrs[k] = nullptr; mine = false; }
            P.vals = (double*)ptrs[0]; P.Y = (double*)ptrs[1]; P.Jt = (double*)ptrs[2]; P.flags = (unsigned long long*)ptrs[3];
        }
        host[(size_t)MSG * h->rank] = mine ? 1.0 : 0.0;
        if (!h->ck(cudaMemcpy(P.d_msg + (size_t)MSG * h->rank, host.data() + (size_t)MSG * h->rank, sizeof(double), cudaMemcpyHostToDevice), "peer setup H2D")) return false;
    }
    if (!exchange(false, 1)) return false;
    // 3. rank 0 -> all: the common verdict
    if (h->rank == 0) {
        if (!h->ck(cudaMemcpy(host.data(), P.d_msg, sizeof(double) * MSG * (size_t)h->world, cudaMemcpyDeviceToHost), "peer setup D2H")) return false;
        for (int r = 1; r < h->eff; ++r) mine = mine && host[(size_t)MSG * r] == 1.0;
        host[0] = mine ? 1.0 : 0.0;
        if (!h->ck(cudaMemcpy(P.d_msg, host.data(), sizeof(double), cudaMemcpyHostToDevice), "peer setup H2D")) return false;
    }
    if (!exchange(true, 1)) return false;
    if (h->rank > 0) {
        if (!h->ck(cudaMemcpy(host.data(), P.d_msg, sizeof(double), cudaMemcpyDeviceToHost), "peer setup D2H")) return false;
        mine = host[0] == 1.0;
    }
    P.on = mine;
    if (!P.on && h->rank > 0) { // leave nothing mapped
        if (P.vals) cudaIpcCloseMemHandle(P.vals);
        if (P.Y) cudaIpcCloseMemHandle(P.Y);
        if (P.Jt) cudaIpcCloseMemHandle(P.Jt);
        if (P.flags) cudaIpcCloseMemHandle(P.flags);
        P.vals = P.Y = P.Jt = nullptr; P.flags = nullptr;
    }
    return true;
}

// Sharded Jacobian (eff > 1). Every working rank differences its own grid points and assembles the CSC entries of
// the columns that belong to them — a column's entries need the FD block of the column's own grid point only, and CasADi's
// column order (CasOCTranscription.h:192-199: states of all points, controls, derivatives) makes them three contiguous runs of
// the value array per rank. Rank 0 receives the runs straight into the destination buffer, plus the point outputs and the time
// column of every block, from which it finishes the two columns every grid point contributes to (t0, tf). Endpoint-constraint
// goals here read x alone, so every rank evaluates them itself (the caller routes problems whose endpoint constraints carry an
// integral through the gather-everything pipeline instead).
static bool jacPipelineSharded(mocob200_handle* h, double* d_vals) {
    if (!dynamicsSupported(h)) return false;
    recTiming(h->ev[0], h->stream);
    if (h->rank >= h->eff) { recTiming(h->ev[1], h->stream); return true; } // idle rank
    const Layout& L = h->L;
    if (h->DL.n_eout_total > 0) {
        cudaEventRecord(h->evX, h->stream);
        cudaStreamWaitEvent(h->stream2, h->evX, 0);
        launch_endpoint(h->DT, h->DL, h->fd, h->d_x, h->d_I2, h->d_G, h->d_E, true, h->stream2, 1);
        cudaEventRecord(h->evE, h->stream2);
        h->launches++;
    }
    if (!h->peer.tried && !peerGatherSetup(h)) return false;
    const bool peer = h->peer.on;
    if (peer && !h->peer.tracing && devEnv("MOCOB200_PEER_TRACE")) {
        for (auto& e : h->peer.trace) cudaEventCreate(&e);
        h->peer.tracing = true;
    }
    auto trace = [&](int k) { if (h->peer.tracing) cudaEventRecord(h->peer.trace[k], h->stream); };
    trace(0);
    const size_t no = (size_t)L.n_out;
    const int np = h->p1 - h->p0;
    if (peer) { // rank 0 releases its buffers to the writers of this Jacobian (see k_peer_signal)
        h->peer.seq++;
        if (h->rank == 0) { k_peer_signal<<<1, 1, 0, h->stream>>>(h->peer.flags, h->peer.seq - 1); h->launches++; }
    }
    launchFdKernels(h, nullptr);
    recTiming(h->ev[1], h->stream);
    trace(1);
    if (peer && h->rank > 0) { // from here on this rank stores into rank 0's memory
        k_peer_wait<<<1, 32, 0, h->stream>>>(h->peer.flags, 0, 1, h->peer.seq - 1);
        h->launches++;
    }
    trace(2);
    launch_pack_time_column(h->DL, h->d_J, peer && h->rank > 0 ? h->peer.Jt : h->d_Jt, h->p0, np, h->stream);
    h->launches++;
    if (h->DL.n_eout_total > 0) cudaStreamWaitEvent(h->stream, h->evE, 0);
    recTiming(h->ev[2], h->stream);
    // with the peer path the other ranks' runs land in rank 0's own value array (the one they have mapped)
    double* mine = h->rank == 0 ? d_vals : (peer ? h->peer.vals : h->d_vals);
    for (int k = 0; k < 3; ++k) {
        const int k0 = h->runs[6 * h->rank + 2 * k], k1 = h->runs[6 * h->rank + 2 * k + 1];
        if (k1 > k0) { launch_assemble_jac(h->DT, h->DL, h->fd, h->d_x, h->d_Y, h->d_J, h->d_I, h->d_E, mine, h->stream, k0, k1); h->launches += 1 + (h->DL.n_path_rows > 0); }
    }
    trace(3);
    if (peer) {
        if (h->rank > 0) {
            if (np > 0 && !h->ck(cudaMemcpyAsync(h->peer.Y + (size_t)h->p0 * no, h->d_Y + (size_t)h->p0 * no, sizeof(double) * np * no,
                                                 cudaMemcpyDeviceToDevice, h->stream), "point outputs to rank 0")) return false;
            k_peer_signal<<<1, 1, 0, h->stream>>>(h->peer.flags + h->rank, h->peer.seq);
            h->launches++;
        } else {
            k_peer_wait<<<1, 32, 0, h->stream>>>(h->peer.flags, 1, h->eff - 1, h->peer.seq);
            trace(4);
            launch_assemble_jac(h->DT, h->DLt, h->fd, h->d_x, h->d_Y, h->d_J, h->d_I, h->d_E, d_vals, h->stream, 0, L.colptr[2]);
            trace(5);
            h->launches += 2 + (h->DL.n_path_rows > 0);
            if (d_vals != h->d_vals) // a caller-supplied destination: the other ranks' runs (adjacent within each block) follow
                for (int k = 0; k < 3; ++k) {
                    const int k0 = h->runs[6 * 1 + 2 * k], k1 = h->runs[6 * (h->eff - 1) + 2 * k + 1];
                    if (k1 > k0 && !h->ck(cudaMemcpyAsync(d_vals + k0, h->d_vals + k0, sizeof(double) * (size_t)(k1 - k0),
                                                          cudaMemcpyDeviceToDevice, h->stream), "D2D values")) return false;
                }
        }
        trace(6);
        recTiming(h->ev[3], h->stream);
        return h->ck(cudaGetLastError(), "peer gather launches");
    }
    // one NCCL group: point outputs + time columns + value runs to rank 0
    if (!ncclOk(h, g_nccl.GroupStart(), "ncclGroupStart")) return false;
    int rc = 0;
    if (h->rank == 0) {
        for (int r = 1; r < h->eff && rc == 0; ++r) {
            const size_t a = h->shard[r], n = (size_t)(h->shard[r + 1] - h->shard[r]) * no;
            if (n) {
                rc = g_nccl.Recv(h->d_Y + a * no, n, kNcclDouble, r, h->comm, h->stream);
                if (rc == 0) rc = g_nccl.Recv(h->d_Jt + a * no, n, kNcclDouble, r, h->comm, h->stream);
            }
            for (int k = 0; k < 3 && rc == 0; ++k) {
                const int k0 = h->runs[6 * r + 2 * k], k1 = h->runs[6 * r + 2 * k + 1];
                if (k1 > k0) rc = g_nccl.Recv(d_vals + k0, (size_t)(k1 - k0), kNcclDouble, r, h->comm, h->stream);
            }
        }
    } else {
        const size_t a = h->p0, n = (size_t)np * no;
        if (n) {
            rc = g_nccl.Send(h->d_Y + a * no, n, kNcclDouble, 0, h->comm, h->stream);
            if (rc == 0) rc = g_nccl.Send(h->d_Jt + a * no, n, kNcclDouble, 0, h->comm, h->stream);
        }
        for (int k = 0; k < 3 && rc == 0; ++k) {
            const int k0 = h->runs[6 * h->rank + 2 * k], k1 = h->runs[6 * h->rank + 2 * k + 1];
            if (k1 > k0) rc = g_nccl.Send(h->d_vals + k0, (size_t)(k1 - k0), kNcclDouble, 0, h->comm, h->stream);
        }
    }
    const int rcEnd = g_nccl.GroupEnd();
    if (!ncclOk(h, rc, "ncclSend/ncclRecv(values)") || !ncclOk(h, rcEnd, "ncclGroupEnd")) return false;
    if (h->rank == 0) { // the t0 and tf columns: the first two columns of the value array
        launch_assemble_jac(h->DT, h->DLt, h->fd, h->d_x, h->d_Y, h->d_J, h->d_I, h->d_E, d_vals, h->stream, 0, L.colptr[2]);
        h->launches += 1 + (h->DL.n_path_rows > 0);
    }
    recTiming(h->ev[3], h->stream);
    return true;
}

static bool jacPipelineLaunches(mocob200_handle* h, double* d_vals) {
    if (h->eff > 1 && !h->ecHasIntegrand && h->L.nm == 0) return jacPipelineSharded(h, d_vals);
    bool deferred = false;
    if (!runFd(h, !h->ecHasIntegrand, &deferred)) return false;
    if (h->rank == 0) {
        recTiming(h->ev[2], h->stream);
        launch_assemble_jac(h->DT, h->DL, h->fd, h->d_x, h->d_Y, h->d_J, h->d_I, h->d_E, d_vals, h->stream);
        recTiming(h->ev[3], h->stream);
        h->launches += 1 + (h->DL.n_path_rows > 0);
    }
    if (deferred) cudaStreamWaitEvent(h->stream, h->evI, 0); // d_I (for a grad_f at the same x) complete; side stream rejoined
    return true;
}
// The Jacobian pipeline (FD kernels on three streams, integrals, endpoint functions, assembly: 8-9 launches and a dozen
// event dependencies) is captured once per handle and FD setting into a CUDA graph that reads h->d_x and writes h->d_vals;
// every later evaluation is one graph launch. Single-rank handles only (the NCCL gather stays outside graphs);
// MOCOB200_GRAPH=0 or MOCOB200_TIMING=1 launches the kernels one by one.
static bool jacPipeline(mocob200_handle* h, double* d_vals) {
    if (!h->allowGraph || h->eff != 1 || h->rank != 0 || h->graphFailed) return jacPipelineLaunches(h, d_vals);
    if (!h->jacGraph) {
        const long before = h->launches;
        cudaGraph_t g = nullptr;
        if (cudaStreamBeginCapture(h->stream, cudaStreamCaptureModeThreadLocal) != cudaSuccess) { cudaGetLastError(); h->graphFailed = true; return jacPipelineLaunches(h, d_vals); }
        const bool ok = jacPipelineLaunches(h, h->d_vals);
        const cudaError_t e = cudaStreamEndCapture(h->stream, &g);
        h->graphLaunches = h->launches - before;
        h->launches = before;
        if (!ok || e != cudaSuccess || !g || cudaGraphInstantiate(&h->jacGraph, g, 0) != cudaSuccess) {
            cudaGetLastError();
            if (g) cudaGraphDestroy(g);
            h->jacGraph = nullptr;
            h->graphFailed = true;
            return jacPipelineLaunches(h, d_vals);
        }
        cudaGraphDestroy(g);
    }
    if (!h->ck(cudaGraphLaunch(h->jacGraph, h->stream), "cudaGraphLaunch")) return false;
    h->launches += h->graphLaunches;
    if (d_vals != h->d_vals &&
        !h->ck(cudaMemcpyAsync(d_vals, h->d_vals, sizeof(double) * h->L.nnz, cudaMemcpyDeviceToDevice, h->stream), "D2D values")) return false;
    return true;
}

// Device alias of a caller buffer in pinned (page-locked, mapped) host memory, or nullptr.
static double* pinnedAlias(mocob200_handle* h, const double* values) {
    auto& P = h->push;
    if (values == P.lastHost) return P.lastDev;
    cudaPointerAttributes at;
    double* dev = nullptr;
    if (cudaPointerGetAttributes(&at, values) == cudaSuccess && at.type == cudaMemoryTypeHost && at.devicePointer) dev = (double*)at.devicePointer;
    else cudaGetLastError();
    P.lastHost = values; P.lastDev = dev;
    return dev;
}
// Once per handle: which entries of the value array can differ from +0.0? The Jacobian at four seeded iterates within the bounds
// (the iterates of optim_sparsity_detection = "random", CasOCFunction.cpp:25-74, here applied to the assembled values and
// bitwise). Worth it only when the dense-block structure carries enough zeros. Single rank, no path-constraint or
// velocity-correction rows (their kernels patch the value array after the main assembly).
static void pushSetup(mocob200_handle* h, const double* x_caller) {
    auto& P = h->push;
    P.tried = true;
    const size_t nnz = (size_t)h->L.nnz;
    if (h->eff != 1 || h->rank != 0 || h->DL.n_path_rows > 0 || h->L.nsl > 0 || nnz < 65536) return;
    std::vector<double> x(h->L.n_var), v(nnz);
    std::vector<unsigned char> mask(nnz, 0);
    for (int k = 0; k < 5; ++k) { // four seeded iterates and the caller's own
        if (k < 4) mocob200_make_iterate(&h->hostT, &h->hostR, k, k == 0 ? 0.3 : 1.0, x.data());
        else if (x_caller) std::memcpy(x.data(), x_caller, sizeof(double) * x.size());
        else break;
        if (!stageX(h, x.data()) || !jacPipeline(h, h->d_vals) ||
            !h->ck(cudaMemcpyAsync(v.data(), h->d_vals, sizeof(double) * nnz, cudaMemcpyDeviceToHost, h->stream), "D2H values") ||
            !finish(h, "host push setup")) { h->err.clear(); cudaGetLastError(); return; }
        for (size_t e = 0; e < nnz; ++e) {
            long long bits;
            std::memcpy(&bits, &v[e], sizeof bits);
            if (bits != 0) mask[e] = 1;
        }
    }
    h->x_cached.clear(); h->haveJ = false;
    size_t kept = 0;
    for (unsigned char m : mask) kept += m;
    if (kept * 10 > nnz * 6) return; // fewer than 40 % zeros: the plain copy is as good
    if (cudaMalloc(&P.d_mask, nnz) != cudaSuccess) { cudaGetLastError(); P.d_mask = nullptr; return; }
    h->owned.push_back(P.d_mask);
    P.mask.swap(mask);
    if (const char* e = devEnv("MOCOB200_PUSH_SHORT_RUN")) { if (atoi(e) > 0) P.shortRun = atoi(e); }
    P.buildRuns();
    if (cudaMemcpy(P.d_mask, P.devMask.data(), nnz, cudaMemcpyHostToDevice) != cudaSuccess ||
        (!P.flag_h && cudaHostAlloc((void**)&P.flag_h, sizeof(int), cudaHostAllocMapped) != cudaSuccess)) { cudaGetLastError(); P.flag_h = nullptr; return; }
    if (cudaHostGetDevicePointer((void**)&P.flag_d, P.flag_h, 0) != cudaSuccess) { cudaGetLastError(); return; }
    P.kept = kept;
    P.on = true;
}

int mocob200_set_host_delivery(mocob200_handle* h, int masked) {
    if (!h || masked < 0 || masked > 1) return MOCOB200_ERR_ARG;
    h->push.wanted = masked != 0;
    return MOCOB200_OK;
}

int mocob200_host_push_stats(const mocob200_handle* h, long* pushed, long* fallbacks, long* kept) {
    if (!h) return MOCOB200_ERR_ARG;
    if (pushed) *pushed = h->push.pushed;
    if (fallbacks) *fallbacks = h->push.fallbacks;
    if (kept) *kept = h->push.on ? (long)h->push.sent : 0;
    return MOCOB200_OK;
}

int mocob200_eval_jac_g(mocob200_handle* h, const double* x, double* values) {
    if (!h) return MOCOB200_ERR_ARG;
    cudaSetDevice(h->device);
    double* hostAlias = (h->push.wanted && h->rank == 0 && values && h->eff == 1) ? pinnedAlias(h, values) : nullptr; // default: untouched path
    if (hostAlias && !h->push.tried) pushSetup(h, x);
    const bool reuse = h->eff == 1 && h->haveJ && sameX(h, x);
    if (!reuse) {
        if (!stageX(h, x) || !bcastX(h, h->d_x)) return MOCOB200_ERR_CUDA;
        if (!jacPipeline(h, h->d_vals)) return MOCOB200_ERR_CUDA;
    } else {
        launch_assemble_jac(h->DT, h->DL, h->fd, h->d_x, h->d_Y, h->d_J, h->d_I, h->d_E, h->d_vals, h->stream);
        h->launches += 1 + (h->DL.n_path_rows > 0);
    }
    const bool push = hostAlias && h->push.on;
    if (push) {
        // a kernel stores the entries of the mask straight into the caller's buffer; the others are written here, by the
        // host, while the device computes
        *h->push.flag_h = 0;
        k_push_masked<<<(h->L.nnz + 255) / 256, 256, 0, h->stream>>>(h->d_vals, h->push.d_mask, hostAlias, h->push.flag_d, h->L.nnz);
        static const int noFill = (devEnv("MOCOB200_HOST_PUSH") && devEnv("MOCOB200_HOST_PUSH")[0] == '2') ? 1 : 0; // timing diagnostic only
        if (!noFill) { // disjoint from what the kernel stores, so it needs no ordering against it: it overlaps the device work
            const unsigned* r = h->push.zeroRuns.data();
            fillZeroRuns(values, r, 0, h->push.zeroRuns.size());
        }
        h->launches++;
    } else if (h->rank == 0 && values)
        cudaMemcpyAsync(values, h->d_vals, sizeof(double) * h->L.nnz, cudaMemcpyDeviceToHost, h->stream);
    rememberX(h, x, true); // the host copy of x (same-x memoisation) overlaps the device work instead of following the sync
    if (!finish(h, "eval_jac_g")) { h->x_cached.clear(); h->haveJ = false; return MOCOB200_ERR_CUDA; }
    if (push) {
        if (*h->push.flag_h) { // an entry outside the mask was not +0.0 at this x: deliver the plain copy and widen the mask
            auto& P = h->push;
            P.fallbacks++;
            if (!h->ck(cudaMemcpyAsync(values, h->d_vals, sizeof(double) * h->L.nnz, cudaMemcpyDeviceToHost, h->stream), "D2H values") ||
                !finish(h, "eval_jac_g")) { h->x_cached.clear(); h->haveJ = false; return MOCOB200_ERR_CUDA; }
            for (size_t e = 0; e < P.mask.size(); ++e) {
                long long bits;
                std::memcpy(&bits, &values[e], sizeof bits);
                if (bits != 0 && !P.mask[e]) { P.mask[e] = 1; P.kept++; }
            }
            P.buildRuns();
            if (P.kept * 10 > P.mask.size() * 6 || P.fallbacks > 64 ||
                cudaMemcpy(P.d_mask, P.devMask.data(), P.devMask.size(), cudaMemcpyHostToDevice) != cudaSuccess) { cudaGetLastError(); P.on = false; }
        } else h->push.pushed++;
    }
    return MOCOB200_OK;
}

// Ordering against the caller's stream (non-NULL cuda_stream): the handle's stream first waits for everything queued on the
// caller's stream so far (the producer of d_x), and the caller's stream afterwards waits for the pipeline (so a consumer of
// d_values may simply be queued on it). NULL: no ordering is added; the caller synchronises through mocob200_sync.
static bool orderAfterCaller(mocob200_handle* h, void* cuda_stream) {
    if (!cuda_stream) return true;
    if (!h->evCaller && !h->ck(cudaEventCreateWithFlags(&h->evCaller, cudaEventDisableTiming), "event")) return false;
    return h->ck(cudaEventRecord(h->evCaller, (cudaStream_t)cuda_stream), "record on the caller's stream") &&
           h->ck(cudaStreamWaitEvent(h->stream, h->evCaller, 0), "wait for the caller's stream");
}
static bool orderCallerAfter(mocob200_handle* h, void* cuda_stream) {
    if (!cuda_stream) return true;
    return h->ck(cudaEventRecord(h->evCaller, h->stream), "record on the handle's stream") &&
           h->ck(cudaStreamWaitEvent((cudaStream_t)cuda_stream, h->evCaller, 0), "caller's stream waits for the pipeline");
}

int mocob200_eval_jac_g_device(mocob200_handle* h, const double* d_x, double* d_values, void* cuda_stream) {
    if (!h || !d_x) return MOCOB200_ERR_ARG;
    cudaSetDevice(h->device);
    h->x_cached.clear();
    if (!orderAfterCaller(h, cuda_stream)) return MOCOB200_ERR_CUDA;
    if (!bcastX(h, d_x)) return MOCOB200_ERR_CUDA;
    if (!jacPipeline(h, d_values ? d_values : h->d_vals)) return MOCOB200_ERR_CUDA;
    if (!orderCallerAfter(h, cuda_stream)) return MOCOB200_ERR_CUDA;
    return h->ck(cudaGetLastError(), "eval_jac_g_device") ? MOCOB200_OK : MOCOB200_ERR_CUDA;
}

// n independent problems (one handle each, single-rank, same device) evaluated side by side: every handle's pipeline
// is queued on its own stream behind a fork point on hs[0]'s stream, and hs[0]'s stream waits for all of them, so a
// mocob200_mark pair / mocob200_sync on hs[0] brackets the whole batch. The reference runs one NLP per process and
// fans its point functions over threads (CasOCTranscription.cpp:1179-1212); one config-3 Jacobian leaves most of a
// B200 idle (SURVEY 8(d) sizing caveat), independent studies (multi-start, parameter sweeps) fill it.
int mocob200_eval_jac_g_batch(mocob200_handle* const* hs, int n, const double* const* d_x, double* const* d_values) {
    if (!hs || n <= 0 || !d_x) return MOCOB200_ERR_ARG;
    for (int k = 0; k < n; ++k) {
        if (!hs[k] || !d_x[k]) return MOCOB200_ERR_ARG;
        if (hs[k]->world != 1 || hs[k]->device != hs[0]->device) {
            hs[k]->err = "eval_jac_g_batch: handles must be single-rank and on one device";
            return MOCOB200_ERR_ARG;
        }
        for (int j = 0; j < k; ++j) if (hs[j] == hs[k]) { hs[k]->err = "eval_jac_g_batch: a handle appears twice"; return MOCOB200_ERR_ARG; }
    }
    mocob200_handle* h0 = hs[0];
    cudaSetDevice(h0->device);
    if (!h0->ck(cudaEventRecord(h0->evBatch, h0->stream), "batch fork")) return MOCOB200_ERR_CUDA;
    for (int k = 0; k < n; ++k) {
        mocob200_handle* h = hs[k];
        if (k && !h->ck(cudaStreamWaitEvent(h->stream, h0->evBatch, 0), "batch fork wait")) return MOCOB200_ERR_CUDA;
        h->x_cached.clear();
        if (!bcastX(h, d_x[k])) return MOCOB200_ERR_CUDA;
        if (!jacPipeline(h, d_values && d_values[k] ? d_values[k] : h->d_vals)) return MOCOB200_ERR_CUDA;
    }
    for (int k = 1; k < n; ++k) {
        mocob200_handle* h = hs[k];
        if (!h->ck(cudaEventRecord(h->evBatch, h->stream), "batch join")) return MOCOB200_ERR_CUDA;
        if (!h0->ck(cudaStreamWaitEvent(h0->stream, h->evBatch, 0), "batch join wait")) return MOCOB200_ERR_CUDA;
    }
    return h0->ck(cudaGetLastError(), "eval_jac_g_batch") ? MOCOB200_OK : MOCOB200_ERR_CUDA;
}

int mocob200_eval_g_device(mocob200_handle* h, const double* d_x, double* d_g, void* cuda_stream) {
    if (!h || !d_x) return MOCOB200_ERR_ARG;
    cudaSetDevice(h->device);
    h->x_cached.clear();
    if (!orderAfterCaller(h, cuda_stream)) return MOCOB200_ERR_CUDA;
    if (!bcastX(h, d_x) || !runNominal(h)) return MOCOB200_ERR_CUDA;
    if (h->rank == 0) { launch_assemble_g(h->DT, h->DL, h->d_x, h->d_Y, h->d_G, d_g ? d_g : h->d_g, h->stream); h->launches += 1 + (h->DL.n_path_rows > 0); }
    if (!orderCallerAfter(h, cuda_stream)) return MOCOB200_ERR_CUDA;
    return h->ck(cudaGetLastError(), "eval_g_device") ? MOCOB200_OK : MOCOB200_ERR_CUDA;
}

// wait for everything queued on the handle's stream; elapsed_ms (optional) = time between two marks
int mocob200_sync(mocob200_handle* h) {
    if (!h) return MOCOB200_ERR_ARG;
    cudaSetDevice(h->device);
    return finish(h, "sync") ? MOCOB200_OK : MOCOB200_ERR_CUDA;
}
// CUDA-event stopwatch on the handle's stream: mark(0) ... mark(1), then elapsed
int mocob200_mark(mocob200_handle* h, int which) {
    if (!h || which < 0 || which > 1) return MOCOB200_ERR_ARG;
    cudaSetDevice(h->device);
    if (!h->marks[which] && !h->ck(cudaEventCreate(&h->marks[which]), "mark")) return MOCOB200_ERR_CUDA;
    return h->ck(cudaEventRecord(h->marks[which], h->stream), "mark") ? MOCOB200_OK : MOCOB200_ERR_CUDA;
}
int mocob200_elapsed_ms(mocob200_handle* h, float* ms) {
    if (!h || !ms || !h->marks[0] || !h->marks[1]) return MOCOB200_ERR_ARG;
    cudaSetDevice(h->device);
    if (!h->ck(cudaEventSynchronize(h->marks[1]), "elapsed")) return MOCOB200_ERR_CUDA;
    return h->ck(cudaEventElapsedTime(ms, h->marks[0], h->marks[1]), "elapsed") ? MOCOB200_OK : MOCOB200_ERR_CUDA;
}

long mocob200_launch_count(const mocob200_handle* h) { return h ? h->launches : 0; }

int mocob200_last_kernel_ms(const mocob200_handle* h, float* point_ms, float* asm_ms) {
    if (!h) return MOCOB200_ERR_ARG;
    if (point_ms) { *point_ms = 0; cudaEventElapsedTime(point_ms, h->ev[0], h->ev[1]); }
    if (asm_ms) { *asm_ms = 0; if (h->rank == 0) cudaEventElapsedTime(asm_ms, h->ev[2], h->ev[3]); }
    return MOCOB200_OK;
}

long mocob200_point_evals_per_jacobian(const mocob200_handle* h) {
    if (!h) return 0;
    const Layout& L = h->L;
    // evaluations of the dynamics callback group only (integrand-only passes are not counted)
    long per_dir = 1;
    if (h->fd.scheme == MOCOB200_FD_CENTRAL) per_dir = 2L * (1 + h->fd.h_iter);
    return (long)L.n_grid * (1 + per_dir * L.n_in);
}

// Evaluations of the point function that one eval_jac_g launches on all ranks together, by kind:
// out[0] whole evaluations, out[1] evaluations that reload the nominal poses (speeds, implicit-mode accelerations),
// out[2] evaluations that reload poses and velocities (time, unclassified inputs), out[3] local-column warps (per grid point: one for central differences, one per 32 owners of local inputs otherwise).
int mocob200_executed_evals(const mocob200_handle* h, long out[4]) {
    if (!h || !out) return MOCOB200_ERR_ARG;
    const Layout& L = h->L;
    const WarpAux& A = h->WA;
    const long N = L.n_grid;
    const bool central = h->fd.scheme == MOCOB200_FD_CENTRAL;
    const long per = central ? 2L * (1 + h->fd.h_iter) : 1;
    const bool warp = h->warpFor(mocob200_point_evals_per_jacobian(h) / h->eff);
    out[0] = out[1] = out[2] = out[3] = 0;
    if (warp) {
        out[0] = N * (1 + per * A.n_p0);
        out[1] = N * per * A.n_p1;
        out[2] = N * per * A.n_p2;
        out[3] = A.n_local > 0 ? N * A.n_lparts : 0;
    } else if (A.n_local > 0 && h->side.s[0]) {
        if (h->velKernel && A.n_p1 > 0 && h->side.s[1]) { out[0] = N * (1 + per * A.n_cw); out[1] = N * per * A.n_p1; }
        else out[0] = N * (1 + per * A.n_cd); // the thread kernels evaluate every non-local direction whole
        out[3] = N;
    } else {
        out[0] = N * (1 + per * L.n_in);
    }
    return MOCOB200_OK;
}
int mocob200_working_ranks(const mocob200_handle* h) { return h ? h->eff : 0; }
int mocob200_peer_gather(const mocob200_handle* h) { return h && h->peer.on ? 1 : 0; }
int mocob200_uses_warp_kernels(const mocob200_handle* h) {
    return h && h->warpFor(mocob200_point_evals_per_jacobian(h) / h->eff) ? 1 : 0;
}

int mocob200_measure_fp64_peak(int device, double* tflops) {
    if (!tflops) return MOCOB200_ERR_ARG;
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev <= 0) return MOCOB200_ERR_CUDA;
    *tflops = measure_fp64_peak_tflops(device);
    return *tflops > 0 ? MOCOB200_OK : MOCOB200_ERR_CUDA;
}

} // extern "C"

// ---------------------------------------------------------------------------------------------
// mocob200_eval_g_static: eval_g for callers that cannot allocate (CUDA graph capture, tight serving loops, harnesses that
// forbid cudaMalloc). No counterpart in the reference. The pipeline's scratch
// (per-point outputs Y, goal integrals, endpoint values) lives in the file-scope array below instead of the handle's buffers:
// every GPU that loads the library holds MOCOB200_STATIC_SCRATCH_BYTES until exit, all calls (all handles) share it, it
// keeps the last call's contents, and calls that overlap on the GPU race. In host code the array's name is a host-side
// stand-in: the kernels get the address cudaGetSymbolAddress returns, looked up once per device with that device current.
static __device__ __align__(256) double g_static_scratch[MOCOB200_STATIC_SCRATCH_BYTES / sizeof(double)];

namespace {
std::mutex g_static_mu;
double* g_static_ptr[64] = {nullptr}; // device address of g_static_scratch, per device ordinal
double* staticScratch(mocob200_handle* h) {
    int dev = -1;
    if (!h->ck(cudaGetDevice(&dev), "cudaGetDevice")) return nullptr;
    if (dev < 0 || dev >= 64) { h->err = "eval_g_static: device ordinal out of range"; return nullptr; }
    std::lock_guard<std::mutex> lock(g_static_mu);
    if (!g_static_ptr[dev]) {
        double* ptr = nullptr;
        if (!h->ck(cudaGetSymbolAddress((void**)&ptr, g_static_scratch), "cudaGetSymbolAddress(static scratch)")) return nullptr;
        g_static_ptr[dev] = ptr;
    }
    return g_static_ptr[dev];
}
inline size_t roundUp32(size_t n) { return (n + 31) / 32 * 32; } // 256-byte sections
} // namespace

extern "C" int mocob200_eval_g_static(mocob200_handle* h, const double* d_x, double* d_g, void* cuda_stream) {
    if (!h || !d_x || !d_g) return MOCOB200_ERR_ARG;
    if (h->world != 1 || h->eff != 1) { h->err = "eval_g_static: single-rank handles only"; return MOCOB200_ERR_UNSUPPORTED; }
    if (!dynamicsSupported(h)) return MOCOB200_ERR_UNSUPPORTED;
    const Layout& L = h->L;
    const size_t nY = roundUp32((size_t)L.n_grid * L.n_out), nI = roundUp32(3 * (size_t)L.n_integrand + 1),
                 nG = roundUp32((size_t)h->DL.n_eout_total + 1), nE = roundUp32((size_t)(h->DL.n_eout_total + 1) * L.n_ein);
    const size_t need = (nY + nI + nG + nE) * sizeof(double);
    if (need > (size_t)MOCOB200_STATIC_SCRATCH_BYTES) {
        h->err = "eval_g_static: the problem needs " + std::to_string(need) + " bytes of scratch (" + std::to_string(L.n_grid) +
                 " grid points x " + std::to_string(L.n_out) + " point outputs), more than the static limit of " +
                 std::to_string((size_t)MOCOB200_STATIC_SCRATCH_BYTES) + " bytes (MOCOB200_STATIC_SCRATCH_BYTES); use mocob200_eval_g_device";
        return MOCOB200_ERR_ARG;
    }
    if (!h->ck(cudaSetDevice(h->device), "cudaSetDevice")) return MOCOB200_ERR_CUDA;
    double* base = staticScratch(h);
    if (!base) return MOCOB200_ERR_CUDA;
    double *sY = base, *sI = sY + nY, *sG = sI + nI, *sE = sG + nG;
    cudaStream_t s = cuda_stream ? (cudaStream_t)cuda_stream : h->stream; // everything is queued on ONE stream, in order
    // the array keeps the previous call's contents: clear the part this call uses
    if (!h->ck(cudaMemsetAsync(base, 0, need, s), "clear static scratch")) return MOCOB200_ERR_CUDA;
    const int np = h->p1 - h->p0;
    if (h->warpForValues(np)) launch_wpoints_nominal(h->DT, h->WA, h->DL, d_x, sY, nullptr, h->p0, np, s);
    else launch_points_nominal(h->DT, h->DL, d_x, sY, h->p0, np, s);
    launch_integrals(h->DL, d_x, sY, nullptr, sI, false, s);
    launch_endpoint(h->DT, h->DL, h->fd, d_x, sI, sG, sE, false, s);
    launch_assemble_g(h->DT, h->DL, d_x, sY, sG, d_g, s);
    h->launches += 2 + (L.n_integrand > 0) + (h->DL.n_eout_total > 0) + (h->DL.n_path_rows > 0);
    return h->ck(cudaGetLastError(), "eval_g_static") ? MOCOB200_OK : MOCOB200_ERR_CUDA;
}
