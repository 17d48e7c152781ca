// libomo.so -- C ABI (include/omo.h) over the sm_100a kernels in omo_kernels.cuh / omo_objective.cuh.
// Host side of the CUDA library only: scene-table flattening for constant memory, buffer management, launches.
// No CPU implementation of the objective lives here: without a CUDA device every entry point fails.
#include <algorithm>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <mutex>
#include <string>
#include <vector>

#include "omo_objective.cuh"
#include "omo_adjoint_kernels.cuh"

#include <chrono>

#include "omo_jit.hpp"
#include "omo_jit_sources.inc"

static_assert(OMO_MAX_JOINTS == 128 && OMO_MAX_IK == 8 && OMO_MAX_IK_CHAIN == 8, "keep the NVRTC branch of omo_scene.cuh in sync with include/omo.h");

namespace {

std::string g_create_error;
std::mutex g_scene_mutex;
const omo_ctx* g_active_scene[64] = {nullptr};  // per device: whose table sits in c_scene
// The table-driven kernels read ONE constant-memory scene table per device (c_scene). Every entry point that launches holds
// its device's mutex from the upload of its table to its last launch, so two contexts on one device driven from two host
// threads (two Cores) cannot read each other's table; a context that finds another one's table active waits for the device
// to drain before it overwrites it (activate_scene). Contexts on different devices never contend.
std::recursive_mutex g_device_mutex[64];

constexpr int kBlockThreads = 128;

}  // namespace

struct omo_ctx {
    int device = 0;
    int sm_count = 148;
    cudaStream_t stream = nullptr;
    cudaStream_t copy_stream = nullptr;      // host -> device copies of omo_eval, pipelined against the kernels on `stream`
    cudaStream_t stream2 = nullptr;          // second compute stream: odd chunks of a pipelined batch, so that the partial last wave of chunk i
                                             // overlaps the first waves of chunk i+1 (a chunk's launches lose most of a wave otherwise)
    cudaEvent_t join_ev = nullptr, fork_ev = nullptr;
    cudaEvent_t chunk_ev[8] = {nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr};
    cudaEvent_t done_ev = nullptr;
    std::string err;

    // scene
    int J = 0, n_ik = 0, n_cp = 0, D = 0, n_channels = 0;
    OmoDevScene hs;               // host copy of the constant table
    bool scene_dirty = true;      // hs changed since the last upload
    std::vector<int> cp_of_pos, first_of_pos, last_of_pos, next_same;
    std::vector<double> joint_const_q;  // dynamics q of constant channels (for residues)
    int *d_cp_of_pos = nullptr, *d_first_of_pos = nullptr, *d_last_of_pos = nullptr, *d_next_same = nullptr;
    int n_slots_nofk[2] = {0, 0}, n_slots_fk[2] = {0, 0};   // [0] full tree, [1] pruned tree
    std::vector<char> pruned;                                // [J] body left out of the pruned tree
    unsigned long long scene_version = 1;                    // bumped whenever hs changes

    // scene-specialised kernels (omo_jit.hpp), one per (fk, ik, count) instance
    struct JitKernel {
        bool tried = false, ok = false;
        omo_jit::CUmodule mod = nullptr;
        omo_jit::CUfunction fn = nullptr;
        omo_jit::CUdeviceptr d_scene = 0;
        size_t smem = 0, item_smem = 0;
        int sub = 1;
        int regs = 0;
        bool from_disk = false, loop_form = false;
        unsigned long long uploaded_version = 0;
        double compile_ms = 0.0;
    };
    JitKernel jit[16];   // index: fk | ik << 1 | count << 2 | self << 3
    int jit_mode = -1;            // -1: auto (large launches), 0: never, 1: always (OMO_JIT)
    bool prefer_jit = false;      // this launch takes the specialised kernel whatever its size (gradient columns: never by shard size)
    int batch_hint = 0;           // omo_set_batch_hint: candidates of the whole batch this context evaluates a slice of
    std::string jit_log;
    unsigned long long jit_launches = 0;

    // window
    bool have_window = false;
    double min_t = 0, max_t = 0, dt = 0.2;
    std::vector<double> tgrid;
    int N_s = 0, n_tiles = 0, T_out = 0;
    double* d_tgrid = nullptr;
    bool residues_valid = false; double residues_dt = 0.0;   // the stencil residues of constants depend on dt only

    // weights
    bool have_weights = false;
    std::vector<double> wparam;
    std::vector<omo_gauss> kernels;
    double w_traj = 0.0;
    int cyclic = 0;
    std::vector<double> wcost;
    double *d_wparam = nullptr, *d_wcost = nullptr;
    bool wcost_dirty = true;

    // original motion
    bool have_original = false;
    std::vector<double> x0;
    double *d_x0 = nullptr, *d_p_orig = nullptr;  // d_p_orig: [3J][N_s]
    bool p_orig_dirty = true;

    // work buffers (capacity in candidates)
    int cap = 0, cap_tiles = 0;   // candidates / tiles per candidate the work buffers were sized for
    int row0 = 0;                 // first scratch row of the launches being queued (a chunk of a pipelined batch works in its own rows, so that
                                  // chunks may run on two streams at once)
    double *d_X = nullptr, *d_prep = nullptr, *d_dparam = nullptr, *d_partial = nullptr, *d_f = nullptr, *d_terms = nullptr;
    unsigned* d_flags = nullptr;
    double* d_seg = nullptr;              // [cap][n_cp][OMO_SEGD] prepared segments
    unsigned short* d_tile_seg = nullptr; // [cap][n_tiles][n_keyed]
    // single-candidate outputs (OptimizationCache)
    double *d_oq = nullptr, *d_oqd = nullptr, *d_oqdd = nullptr, *d_otau = nullptr, *d_olocal = nullptr, *d_op = nullptr;
    int out_cap_samples = 0;
    // gradient scratch
    double* d_xbase = nullptr; int* d_free = nullptr; int free_cap = 0;
    std::vector<int> pos_of_cp;                 // inverse of cp_of_pos
    std::vector<int> h_items, h_lo, h_hi;       // sparse finite-difference columns: listed (row, tile) items, tile range per column
    int *d_items = nullptr, *d_lohi = nullptr;  // their device copies (d_lohi = lo[cols] then hi[cols])
    size_t items_cap = 0, lohi_cap = 0;
    unsigned long long sparse_items = 0, dense_items = 0;   // executed / dense work items of the last sparse launch
    // what the device copies of h_items / h_lo / h_hi / free_idx were built from (an optimisation loop moves theta / t_b / t_f
    // but no key time, so the item list of one gradient is the next one's too)
    std::vector<double> sp_times; std::vector<int> sp_free; int sp_begin = -1, sp_count = -1; bool sp_valid = false, sp_uploaded = false;
    std::vector<int> free_uploaded;
    std::vector<double> h_cols;
    double* d_Xfree = nullptr; size_t xfree_cap = 0;   // omo_eval_free: the candidates' free entries as they arrive from the host
    std::vector<double> xbase_uploaded; bool xbase_is_free_base = false;   // what d_xbase holds, if omo_eval_free put it there (the gradient paths write d_xbase too)
    int* d_inv_free = nullptr; std::vector<int> inv_free_of;   // descriptor index -> position in the free list (-1: frozen), and the free list it was built from

    // single-candidate call (the NLopt callback, core_opt.cpp:42-57): H2D of x -> prepare -> objective -> finish -> D2H of f and
    // the three terms, captured once as a CUDA graph and replayed with one launch while nothing the capture baked in has changed
    int graph_mode = -1;                        // OMO_GRAPH=0 switches it off
    cudaGraphExec_t sg_exec = nullptr;
    std::vector<unsigned long long> sg_key, sg_pending;
    unsigned long long sg_kernels = 0, sg_replays = 0;
    double* h_single = nullptr;                 // pinned: x [D], then f, C, D_param, D_traj
    double* d_single = nullptr;                 // device: f, C, D_param, D_traj of the single candidate (one copy-out)

    // self-preparing launches (Walker SELF): one launch per single candidate / small gradient, x as a kernel parameter
    int self_mode = -1;                         // OMO_SELF=0 switches it off
    double* d_self_dparam = nullptr; int self_rows_cap = 0;
    unsigned* d_self_counter = nullptr;
    double* h_self_out = nullptr;               // mapped pinned host memory the kernel writes f into: [max(4, self_out_cap)]
    double* d_self_out = nullptr;               // its device alias
    int self_out_cap = 0;
    int self_spin_slots = 0;                    // outputs of the self-preparing launch in flight that the host polls in h_self_out (0: wait on the stream)
    int self_host_tiles = 0;                    // > 0: single candidate whose tile partials the HOST adds up (slots 4 .. of h_self_out), see launch_self
    int self_host_rows = 0, self_host_base = -1, self_host_ntiles = 0, self_host_off = 0;   // > 0 rows: a small gradient reduced on the host the same way
    int self_spin_mode = -1;                    // OMO_SPIN=0: always wait on the stream
    unsigned long long self_launches = 0;

    // sorted curve work lists (OmoEvalParams::perm), rebuilt with x_original / the window
    unsigned short* d_perm = nullptr; size_t perm_cap = 0; int perm_stride = 0; bool perm_dirty = true, perm_valid = false; int perm_mode = -1;

    // analytic gradient (omo_adjoint_kernels.cuh)
    int *d_sub_end = nullptr, *d_pos_of_cp = nullptr, *d_ci_of_pos = nullptr, *d_adj_m = nullptr;
    double *d_adj_L = nullptr, *d_adj_G = nullptr, *d_adj_contrib = nullptr, *d_adj_gfull = nullptr, *d_adj_gout = nullptr, *d_oroot = nullptr;
    size_t adj_cap = 0; int adj_F_cap = 0; int adj_max_sub = 0;
    double* d_adj_ws = nullptr; size_t adj_ws_cap = 0, adj_L_cap = 0; int adj_mode = -1;   // 0 forward, 1 reverse, 2 by size (OMO_ADJ_MODE)
    unsigned long long adj_calls = 0;

    // omo_peer_gather_eval in flight: the finish step of the batch launches is the fused finish + peer exchange kernel
    struct omo_peer_gather* peer = nullptr;

    // counters / timing
    unsigned long long* d_counters = nullptr;
    bool counting = false;
    unsigned long long launches = 0;
    bool timing = false;
    std::vector<std::pair<cudaEvent_t, cudaEvent_t>> timed;
    std::vector<cudaEvent_t> event_pool;
};

namespace {

#define OMO_CUDA(call)                                                                                     \
    do {                                                                                                   \
        cudaError_t e_ = (call);                                                                           \
        if (e_ != cudaSuccess) {                                                                           \
            ctx->err = std::string(#call) + ": " + cudaGetErrorString(e_);                                 \
            return (e_ == cudaErrorMemoryAllocation) ? OMO_OOM : OMO_CUDA_ERROR;                           \
        }                                                                                                  \
    } while (0)

struct DeviceLock {
    std::recursive_mutex* m = nullptr;
    explicit DeviceLock(const omo_ctx* ctx) { if (ctx && ctx->device >= 0 && ctx->device < 64) { m = &g_device_mutex[ctx->device]; m->lock(); } }
    ~DeviceLock() { if (m) m->unlock(); }
    DeviceLock(const DeviceLock&) = delete;
    DeviceLock& operator=(const DeviceLock&) = delete;
};

omo_status fail(omo_ctx* ctx, omo_status st, const std::string& msg)
{
    if (ctx) ctx->err = msg; else g_create_error = msg;
    return st;
}

template <typename T> omo_status dev_alloc(omo_ctx* ctx, T** p, size_t n)
{
    if (*p) { cudaFree(*p); *p = nullptr; }
    if (n == 0) n = 1;
    OMO_CUDA(cudaMalloc((void**)p, n * sizeof(T)));
    return OMO_OK;
}

template <typename T> omo_status upload(omo_ctx* ctx, T** p, const std::vector<T>& v)
{
    omo_status st = dev_alloc(ctx, p, v.size());
    if (st != OMO_OK) return st;
    if (!v.empty()) OMO_CUDA(cudaMemcpyAsync(*p, v.data(), v.size() * sizeof(T), cudaMemcpyHostToDevice, ctx->stream));
    return OMO_OK;
}

// SpatialRigidBodyInertia::createFromMassComInertiaC (external/rbdl-v2.5.0/include/rbdl/SpatialAlgebraOperators.h:116-128):
// I = inertia_C + m * (cx * cx^T), h = m * com.
void make_body(OmoBody& B, const double* r, double mass, const double* com, const double* Ic)
{
    B.r[0] = r[0]; B.r[1] = r[1]; B.r[2] = r[2];
    B.m = mass;
    B.h[0] = com[0] * mass; B.h[1] = com[1] * mass; B.h[2] = com[2] * mass;
    double cx[3][3] = {{0.0, -com[2], com[1]}, {com[2], 0.0, -com[0]}, {-com[1], com[0], 0.0}};
    double I[3][3];
    for (int i = 0; i < 3; ++i)
        for (int j = 0; j < 3; ++j) {
            const double cc = cx[i][0] * cx[j][0] + cx[i][1] * cx[j][1] + cx[i][2] * cx[j][2];
            I[i][j] = Ic[3 * i + j] + cc * mass;
        }
    B.Ixx = I[0][0]; B.Iyx = I[1][0]; B.Iyy = I[1][1]; B.Izx = I[2][0]; B.Izy = I[2][1]; B.Izz = I[2][2];
}

// Rounding residue the 5-point stencils leave on a constant channel (core_rbdl.cpp:49,84; SURVEY §7 ix).
void constant_residue(double c, double dx, double* qd, double* qdd)
{
    volatile double v = c;  // keep the reference's left-to-right evaluation
    *qd = (-v + 8.0 * v - 8.0 * v + v) / (12.0 * dx);
    *qdd = (-v + 16.0 * v - 30.0 * v + 16.0 * v - v) / (12.0 * dx * dx);
}

// Host-only part of omo_create: flattens the scene into OmoDevScene and the prepare tables (no CUDA calls).
omo_status build_tables(omo_ctx* ctx, const omo_scene_desc* sc)
{
    const int J = sc->n_joints;
    if (J < 1 || J > OMO_MAXJ) return fail(ctx, OMO_BAD_ARG, "n_joints must be in [1, OMO_MAX_JOINTS]");
    if (sc->n_ik < 0 || sc->n_ik > OMO_MAX_IK) return fail(ctx, OMO_BAD_ARG, "n_ik must be in [0, OMO_MAX_IK]");
    if (sc->n_channels != 3 * J + 3 * sc->n_ik) return fail(ctx, OMO_BAD_ARG, "n_channels must equal 3*J + 3*n_ik");
    if (!sc->parent || !sc->offset || !sc->mass || !sc->com || !sc->inertia || !sc->chan_const || !sc->chan_first_cp || !sc->chan_n_cp)
        return fail(ctx, OMO_BAD_ARG, "null scene table");
    if (sc->n_cp > 0 && !sc->chain) return fail(ctx, OMO_BAD_ARG, "null chain table");
    if (sc->parent[0] != -1) return fail(ctx, OMO_BAD_ARG, "parent[0] must be -1 (root first)");
    OmoDevScene& S = ctx->hs;
    std::memset(&S, 0, sizeof(S));
    S.J = J; S.n_ik = sc->n_ik;
    ctx->J = J; ctx->n_ik = sc->n_ik; ctx->n_cp = sc->n_cp; ctx->D = 5 * sc->n_cp; ctx->n_channels = sc->n_channels;

    std::vector<int> depth(J, 0), nchild(J, 0);
    for (int j = 1; j < J; ++j) {
        const int p = sc->parent[j];
        if (p < 0 || p >= j) return fail(ctx, OMO_BAD_ARG, "joints must be in DFS pre-order (0 <= parent[j] < j)");
        depth[j] = depth[p] + 1;
        nchild[p] += 1;
    }
    // DFS pre-order check: parent[j] must be j-1 or an ancestor of j-1
    for (int j = 2; j < J; ++j) {
        int k = j - 1; bool ok = false;
        while (k >= 0) { if (k == sc->parent[j]) { ok = true; break; } k = sc->parent[k]; }
        if (!ok) return fail(ctx, OMO_BAD_ARG, "joints are not in DFS pre-order");
    }
    for (int j = 0; j < J; ++j) { S.parent[j] = sc->parent[j]; S.n_child[j] = (unsigned char)std::min(nchild[j], 255); }

    // bodies (core_rbdl.cpp:228-252): body j hangs at Xtrans(parent.offset), carries j's own mass properties
    for (int j = 0; j < J; ++j) {
        for (int k = 0; k < 3; ++k) S.offset[j][k] = sc->offset[3 * j + k];
        if (j == 0) { const double z[3] = {0, 0, 0}; const double zi[9] = {0}; make_body(S.body[0], sc->offset, 0.0, z, zi); continue; }
        make_body(S.body[j], sc->offset + 3 * sc->parent[j], sc->mass[j], sc->com + 3 * j, sc->inertia + 9 * j);
    }

    // channels -> chain-ordered positions
    ctx->cp_of_pos.assign(sc->n_cp, 0); ctx->first_of_pos.assign(sc->n_cp, 0); ctx->last_of_pos.assign(sc->n_cp, 0);
    ctx->next_same.assign(sc->n_cp, 0);
    int pos = 0;
    for (int ch = 0; ch < sc->n_channels; ++ch) {
        const int n = sc->chan_n_cp[ch], first = sc->chan_first_cp[ch];
        if (n < 0 || (n > 0 && (first < 0 || first + n > sc->n_cp))) return fail(ctx, OMO_BAD_ARG, "bad channel control-point range");
        if (n > 65535) return fail(ctx, OMO_BAD_ARG, "more than 65535 control points on one channel");
        int cpos = -1;
        if (n > 0) {
            if (first != pos) return fail(ctx, OMO_BAD_ARG, "channels must list their control points contiguously in descriptor order");
            cpos = pos;
            for (int i = 0; i < n; ++i) {
                const int cp = sc->chain[first + i];
                if (cp < first || cp >= first + n) return fail(ctx, OMO_BAD_ARG, "chain entry outside its channel");
                ctx->cp_of_pos[pos] = cp; ctx->first_of_pos[pos] = first; ctx->last_of_pos[pos] = first + n - 1;
                ctx->next_same[pos] = (i + 1 < n) ? 1 : 0;
                ++pos;
            }
        }
        if (ch < 3 * J) {
            S.chan_pos[ch] = cpos; S.chan_n[ch] = n; S.chan_const[ch] = sc->chan_const[ch];
            S.chan_sin[ch] = std::sin(sc->chan_const[ch]); S.chan_cos[ch] = std::cos(sc->chan_const[ch]);
            if (n > 0) S.dyn[ch / 3] |= (unsigned char)(1 << (ch % 3));
        } else {
            const int h = (ch - 3 * J) / 3, a = (ch - 3 * J) % 3;
            S.ik_chan_pos[h][a] = cpos; S.ik_chan_n[h][a] = n; S.ik_chan_const[h][a] = sc->chan_const[ch];
        }
    }
    if (pos != sc->n_cp) return fail(ctx, OMO_BAD_ARG, "n_cp does not match the channels' control points");
    // joints frozen at zero angles take the identity fast path (E = I, qd = qdd = 0 exactly)
    for (int j = 1; j < J; ++j)
        S.ident[j] = (S.dyn[j] == 0 && S.chan_const[3 * j] == 0.0 && S.chan_const[3 * j + 1] == 0.0 && S.chan_const[3 * j + 2] == 0.0) ? 1 : 0;

    // curve groups: bodies are packed greedily, a body's keyed channels never straddle two groups
    {
        int total = 0;
        for (int j = 1; j < J; ++j) for (int a = 0; a < 3; ++a) total += (S.dyn[j] >> a) & 1;
        const int G = std::max(3, std::min(total, 12));
        int g = -1, used = G, n_list = 0;
        for (int j = 0; j < J; ++j) S.grp_first[j] = -1;
        for (int ch = 0; ch < 3 * J; ++ch) S.chan_row[ch] = -1;
        for (int j = 1; j < J; ++j) {
            const int nd = ((S.dyn[j] >> 0) & 1) + ((S.dyn[j] >> 1) & 1) + ((S.dyn[j] >> 2) & 1);
            if (nd == 0) continue;
            if (used + nd > G) { ++g; used = 0; S.grp_first[j] = (short)g; S.grp_begin[g] = (short)n_list; }
            for (int a = 0; a < 3; ++a)
                if ((S.dyn[j] >> a) & 1) {
                    S.grp_list[n_list++] = (short)(3 * j + a);
                    S.chan_row[3 * j + a] = (short)((g & 1) * G + used);
                    ++used;
                }
        }
        S.n_groups = g + 1;
        S.grp_begin[S.n_groups] = (short)n_list;
        S.n_keyed = n_list;
        S.n_ex_rows = (S.n_groups > 1) ? 2 * G : G;
    }

    // IK handlers (ikhandler.cpp:45-67,77-101): chain = path root .. end effector
    for (int j = 0; j < J; ++j) { S.ik_of[j] = -1; S.ik_idx[j] = 0; S.ik_start[j] = -1; }
    for (int h = 0; h < sc->n_ik; ++h) {
        const int root = sc->ik_root[h], end = sc->ik_end[h];
        if (root <= 0 || root >= J || end <= 0 || end >= J) return fail(ctx, OMO_BAD_ARG, "IK root/end must be non-root joints");
        std::vector<int> path;
        for (int j = end;; j = sc->parent[j]) {
            path.insert(path.begin(), j);
            if (j == root) break;
            if (j <= 0) return fail(ctx, OMO_BAD_ARG, "IK end effector is not a descendant of the IK root");
        }
        const int nb = (int)path.size() - 1;
        if (nb < 1 || nb > OMO_MAX_IK_CHAIN) return fail(ctx, OMO_BAD_ARG, "IK chain length must be in [1, OMO_MAX_IK_CHAIN]");
        S.ik_nb[h] = nb;
        for (int i = 0; i <= nb; ++i) S.ik_joint[h][i] = path[i];
        S.ik_start[root] = (short)h;
        for (int i = 0; i < nb; ++i) { S.ik_of[path[i]] = (short)h; S.ik_idx[path[i]] = (short)i; S.ident[path[i]] = 0; }
    }
    // ---- walk plans -------------------------------------------------------------------------------------------
    // Tree 0 = every body (used by the per-sample-output instance). Tree 1 = the bodies whose torque can depend on
    // the candidate: a body is pruned when neither it, nor any ancestor, nor any descendant has a keyed or
    // IK-driven angle -- its v, a, f are the same for every candidate (the stencil residue of constants at
    // interior samples, exactly 0 at boundary samples), so the sum of its tau^2 is a per-scene constant
    // (OmoDevScene::static_tsq, measured once from the full-tree instance in ensure_ready). A keyed root
    // rotation moves every joint position (joint.cpp:64-83), so nothing is pruned then.
    {
        std::vector<char> dynamic(J, 0), anc_dyn(J, 0), sub_dyn(J, 0);
        for (int j = 1; j < J; ++j) dynamic[j] = (S.dyn[j] != 0 || S.ik_of[j] >= 0) ? 1 : 0;
        for (int j = 1; j < J; ++j) anc_dyn[j] = (char)(dynamic[j] | anc_dyn[sc->parent[j]]);
        for (int j = J - 1; j >= 1; --j) { sub_dyn[j] |= dynamic[j]; sub_dyn[sc->parent[j]] |= sub_dyn[j]; }
        ctx->pruned.assign(J, 0);
        const bool allow_prune = (S.dyn[0] == 0) && std::getenv("OMO_NO_PRUNE") == nullptr;
        for (int j = 1; j < J; ++j) ctx->pruned[j] = (allow_prune && !anc_dyn[j] && !sub_dyn[j]) ? 1 : 0;
        for (int tree = 0; tree < 2; ++tree) {
            std::vector<int> nch(J, 0), nb_nonroot(J, 0), nb_all(J, 0), fs(J, -1), va(J, -1), fk(J, -1);
            auto keep = [&](int j) { return tree == 0 || !ctx->pruned[j]; };
            for (int j = 1; j < J; ++j) if (keep(j)) nch[sc->parent[j]] += 1;
            int max_fs = 0, max_va = 0, max_fk = 0, n_act = 0;
            for (int j = 0; j < J; ++j) {
                if (!keep(j)) continue;
                if (j > 0) {
                    const int pj = sc->parent[j];
                    nb_nonroot[j] = nb_nonroot[pj] + ((pj > 0 && nch[pj] >= 2) ? 1 : 0);
                    nb_all[j] = nb_all[pj] + ((nch[pj] >= 2) ? 1 : 0);
                    S.act[tree][n_act++] = (unsigned char)j;
                }
                if (j > 0 && nch[j] > 0) { fs[j] = depth[j] - 1; max_fs = std::max(max_fs, depth[j]); }
                if (j > 0 && nch[j] >= 2) { va[j] = nb_nonroot[j]; max_va = std::max(max_va, nb_nonroot[j] + 1); }
                if (nch[j] >= 2) { fk[j] = nb_all[j]; max_fk = std::max(max_fk, nb_all[j] + 1); }
            }
            S.n_act[tree] = n_act;
            for (int j = 0; j < J; ++j) {
                if (va[j] >= 0) va[j] += max_fs;
                if (fk[j] >= 0) fk[j] += max_fs + max_va;
                auto byte = [](int v) { return v < 0 ? 0xff : (v & 0xff); };
                if (max_fs + max_va + max_fk >= 0xff) return fail(ctx, OMO_BAD_ARG, "joint tree needs too many stack slots");
                int4 w;
                w.x = (j > 0 ? sc->parent[j] : 0) | (byte(fs[j]) << 8) | (byte(va[j]) << 16) | (byte(fk[j]) << 24);
                w.y = (S.dyn[j] & 7) | ((S.ident[j] ? 1 : 0) << 3) | ((nch[j] > 0 ? 1 : 0) << 4) | (((S.grp_first[j] + 1) & 0xff) << 8) |
                      (((S.ik_of[j] + 1) & 0xf) << 16) | ((S.ik_idx[j] & 0xf) << 20) | (((S.ik_start[j] + 1) & 0xf) << 24);
                w.z = byte(S.chan_row[3 * j]) | (byte(S.chan_row[3 * j + 1]) << 8) | (byte(S.chan_row[3 * j + 2]) << 16);
                w.w = 0;
                S.meta[tree][j] = w;
            }
            S.n_slots_tree[tree] = max_fs + max_va + max_fk;
            S.n_slots_nofk_tree[tree] = max_fs + max_va;
            ctx->n_slots_nofk[tree] = max_fs + max_va;
            ctx->n_slots_fk[tree] = max_fs + max_va + max_fk;
        }
        S.n_slots = ctx->n_slots_fk[0];
        S.static_tsq = 0.0;
    }
    ctx->scene_dirty = true;
    ctx->scene_version += 1;
    return OMO_OK;
}

omo_status build_scene(omo_ctx* ctx, const omo_scene_desc* sc)
{
    omo_status st = build_tables(ctx, sc);
    if (st != OMO_OK) return st;
    if ((st = upload(ctx, &ctx->d_cp_of_pos, ctx->cp_of_pos)) != OMO_OK) return st;
    if ((st = upload(ctx, &ctx->d_first_of_pos, ctx->first_of_pos)) != OMO_OK) return st;
    if ((st = upload(ctx, &ctx->d_last_of_pos, ctx->last_of_pos)) != OMO_OK) return st;
    if ((st = upload(ctx, &ctx->d_next_same, ctx->next_same)) != OMO_OK) return st;
    return OMO_OK;
}

void refresh_residues(omo_ctx* ctx)
{
    OmoDevScene& S = ctx->hs;
    for (int ch = 0; ch < 3 * ctx->J; ++ch) constant_residue(S.chan_const[ch], ctx->dt, &S.chan_qd[ch], &S.chan_qdd[ch]);
    for (int k = 0; k < 3; ++k) constant_residue(S.offset[0][k], ctx->dt, &S.root_qd[k], &S.root_qdd[k]);
    ctx->scene_dirty = true;
    ctx->scene_version += 1;
}

omo_status activate_scene(omo_ctx* ctx)
{
    std::lock_guard<std::mutex> g(g_scene_mutex);
    const int dev = ctx->device;
    if (dev < 0 || dev >= 64) return fail(ctx, OMO_BAD_ARG, "device index out of range");
    if (g_active_scene[dev] == ctx && !ctx->scene_dirty) return OMO_OK;
    if (g_active_scene[dev] != nullptr && g_active_scene[dev] != ctx) OMO_CUDA(cudaDeviceSynchronize());  // other ctx may still read it
    OMO_CUDA(cudaMemcpyToSymbolAsync(c_scene, &ctx->hs, sizeof(OmoDevScene), 0, cudaMemcpyHostToDevice, ctx->stream));
    g_active_scene[dev] = ctx;
    ctx->scene_dirty = false;
    return OMO_OK;
}

// prepare_kernel's split of a candidate's chain positions over blocks
constexpr int kPrepChunk = 640;
int prep_parts(const omo_ctx* ctx) { return std::max(1, (ctx->n_cp + 640 - 1) / 640); }

// scratch of the candidate rows [row0, row0 + B) of the launches being queued
double* sc_prep(const omo_ctx* c) { return c->d_prep + (size_t)c->row0 * 6 * c->n_cp; }
unsigned* sc_flags(const omo_ctx* c) { return c->d_flags + c->row0; }
double* sc_seg(const omo_ctx* c) { return c->d_seg + (size_t)c->row0 * c->n_cp * OMO_SEGD; }
unsigned short* sc_tile_seg(const omo_ctx* c) { return c->d_tile_seg + (size_t)c->row0 * std::max(c->n_tiles, 1) * std::max(c->hs.n_keyed, 1); }
double* sc_partial(const omo_ctx* c) { return c->d_partial + (size_t)c->row0 * std::max(c->n_tiles, 1) * 2; }
int prep_parts(const omo_ctx* ctx);
double* sc_dparam(const omo_ctx* c) { return c->d_dparam + (size_t)c->row0 * prep_parts(c); }

omo_status ensure_capacity(omo_ctx* ctx, int B)
{
    if (B <= ctx->cap && ctx->n_tiles <= ctx->cap_tiles) return OMO_OK;
    const int cap = std::max(std::max(B, ctx->cap), 16);
    const int tiles = std::max(std::max(ctx->n_tiles, ctx->cap_tiles), 1);
    omo_status st;
    if ((st = dev_alloc(ctx, &ctx->d_X, (size_t)cap * ctx->D)) != OMO_OK) return st;
    if ((st = dev_alloc(ctx, &ctx->d_prep, (size_t)cap * 6 * ctx->n_cp)) != OMO_OK) return st;
    if ((st = dev_alloc(ctx, &ctx->d_dparam, (size_t)cap * prep_parts(ctx))) != OMO_OK) return st;
    if ((st = dev_alloc(ctx, &ctx->d_flags, (size_t)cap)) != OMO_OK) return st;
    // + OMO_SEGW records of slack: the staging copy of the last channel never reads past the allocation
    if ((st = dev_alloc(ctx, &ctx->d_seg, ((size_t)cap * ctx->n_cp + OMO_SEGW) * OMO_SEGD)) != OMO_OK) return st;
    if ((st = dev_alloc(ctx, &ctx->d_tile_seg, (size_t)cap * tiles * std::max(ctx->hs.n_keyed, 1))) != OMO_OK) return st;
    if ((st = dev_alloc(ctx, &ctx->d_partial, (size_t)cap * tiles * 2)) != OMO_OK) return st;
    if ((st = dev_alloc(ctx, &ctx->d_f, (size_t)cap)) != OMO_OK) return st;
    if ((st = dev_alloc(ctx, &ctx->d_terms, (size_t)cap * 3)) != OMO_OK) return st;
    ctx->cap = cap; ctx->cap_tiles = tiles;
    return OMO_OK;
}

struct LaunchPlan { int ns, n_tiles, t_out; size_t smem; int smem_slots; bool deep; };

size_t smem_bytes(const omo_ctx* ctx, bool fk, int ns, bool out, int slots_override = -1)
{
    const int tree = out ? 0 : 1;
    const int slots = slots_override >= 0 ? slots_override : (fk ? ctx->n_slots_fk[tree] : ctx->n_slots_nofk[tree]);
    const int ex_rows = ctx->hs.n_ex_rows + (ctx->n_ik > 0 ? 6 : 0);
    const size_t seg = (((size_t)(kBlockThreads / 32) * ctx->hs.n_keyed * sizeof(unsigned short)) + 15) & ~(size_t)15;
    return std::max((size_t)(ex_rows + slots * OMO_SLOT) * ns * kBlockThreads * sizeof(double) + seg, (size_t)64 * sizeof(double));
}

// Tiling of the time axis: a block covers 128 consecutive samples of one candidate, 4 of them halo.
LaunchPlan make_plan(const omo_ctx* ctx, int B, bool fk, bool ik, bool out)
{
    LaunchPlan p;
    p.ns = 1;   // two samples per thread (NS = 2) measured slower on B200 (register pressure halves the resident warps)
    const int t_max = p.ns * kBlockThreads - 4;
    p.n_tiles = (ctx->N_s + t_max - 1) / t_max;
    p.t_out = (ctx->N_s + p.n_tiles - 1) / p.n_tiles;
    p.smem = smem_bytes(ctx, fk, p.ns, out);
    p.deep = false;
    p.smem_slots = fk ? ctx->n_slots_fk[out ? 0 : 1] : ctx->n_slots_nofk[out ? 0 : 1];
    if (p.smem > (size_t)227 * 1024) {
        // deep tree: as many slots as fit stay in shared memory, the rest go per thread (objective_kernel<..., DEEP>, which
        // always carries the FK steps: one set of deep instances instead of two)
        const int all = ctx->n_slots_fk[out ? 0 : 1];
        int fit = all;
        while (fit > 0 && smem_bytes(ctx, true, p.ns, out, fit) > (size_t)227 * 1024) --fit;
        p.deep = true; p.smem_slots = fit;
        p.smem = smem_bytes(ctx, true, p.ns, out, fit);
        if (all - fit > OMO_DEEP_SLOTS) p.smem = (size_t)1 << 30;   // refused by the launch below
    }
    return p;
}

template <bool FK, bool IK, bool OUT, bool COUNT, bool DEEP = false>
omo_status launch_objective_t(omo_ctx* ctx, const OmoEvalParams& P, int B, const LaunchPlan& plan)
{
    if (plan.smem > 227 * 1024) return fail(ctx, OMO_BAD_ARG, "joint tree needs more stack slots than shared memory plus OMO_DEEP_SLOTS per-thread slots hold");
    OMO_CUDA(cudaFuncSetAttribute(omo::objective_kernel<FK, IK, OUT, COUNT, DEEP>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)plan.smem));
    cudaEvent_t e0 = nullptr, e1 = nullptr;
    if (ctx->timing && !OUT) {
        OMO_CUDA(cudaEventCreate(&e0)); OMO_CUDA(cudaEventCreate(&e1));
        OMO_CUDA(cudaEventRecord(e0, ctx->stream));
    }
    omo::objective_kernel<FK, IK, OUT, COUNT, DEEP><<<P.n_items, kBlockThreads, plan.smem, ctx->stream>>>(P);
    OMO_CUDA(cudaGetLastError());
    if (e0) { OMO_CUDA(cudaEventRecord(e1, ctx->stream)); ctx->timed.emplace_back(e0, e1); }
    ctx->launches += 1;
    return OMO_OK;
}

// ---- scene-specialised kernels (omo_jit.hpp) ------------------------------------------------------------------------
int jit_mode(omo_ctx* ctx)
{
    if (ctx->jit_mode < 0) {
        const char* e = std::getenv("OMO_JIT");
        ctx->jit_mode = e ? (std::atoi(e) == 0 ? 0 : (std::atoi(e) == 1 ? 1 : 2)) : 2;   // 2 = auto
    }
    return ctx->jit_mode;
}

// Compiles (once) and returns the specialised kernel of an instance, or nullptr when it cannot be had -- the caller
// then launches the table-driven kernel, which is the same CUDA code driven by tables.
omo_ctx::JitKernel* acquire_jit(omo_ctx* ctx, bool fk, bool ik, bool count, bool self = false)
{
    omo_ctx::JitKernel& k = ctx->jit[(fk ? 1 : 0) | (ik ? 2 : 0) | (count ? 4 : 0) | (self ? 8 : 0)];
    if (k.tried) return k.ok ? &k : nullptr;
    k.tried = true;
    omo_jit::Api& a = omo_jit::api();
    if (!a.have_nvrtc() || !a.have_driver()) { ctx->jit_log = "JIT unavailable:" + a.error; return nullptr; }
    const auto t0 = std::chrono::steady_clock::now();
    omo_jit::KernelSpec spec = omo_jit::generate(ctx->hs, fk, ik, count, self ? ctx->n_cp : 0);
    if (const char* dump = std::getenv("OMO_JIT_DUMP")) {
        if (FILE* fh = std::fopen((std::string(dump) + (fk ? "_fk" : "") + (ik ? "_ik" : "") + (count ? "_count" : "") + (self ? "_self" : "") + ".cu").c_str(), "w")) {
            std::fputs(spec.source.c_str(), fh); std::fclose(fh);
        }
    }
    // cubins are cached per process by the hash of their source: a second context on the same scene (another Core, another
    // device) skips NVRTC
    static std::mutex cache_mutex;
    static std::map<uint64_t, std::vector<char>> cubin_cache;
    const uint64_t key = omo_jit::fnv1a(spec.source + "#sub" + std::to_string(spec.sub)) ^ omo_jit::headers_hash();
    std::vector<char> cubin;
    {
        std::lock_guard<std::mutex> g(cache_mutex);
        auto it = cubin_cache.find(key);
        if (it != cubin_cache.end()) cubin = it->second;
    }
    if (cubin.empty() && omo_jit::disk_cache_load(key, cubin)) {
        k.from_disk = true;
        std::lock_guard<std::mutex> g(cache_mutex);
        cubin_cache[key] = cubin;
    }
    if (cubin.empty()) {
        std::string log;
        if (!omo_jit::compile(spec.source, cubin, log, spec.sub)) { ctx->jit_log = "NVRTC: " + log; return nullptr; }
        omo_jit::disk_cache_store(key, cubin);
        std::lock_guard<std::mutex> g(cache_mutex);
        cubin_cache[key] = cubin;
    }
    k.loop_form = spec.loop_form;
    if (a.cuModuleLoadData(&k.mod, cubin.data()) != 0) { ctx->jit_log = "cuModuleLoadData failed"; return nullptr; }
    size_t bytes = 0;
    if (a.cuModuleGetFunction(&k.fn, k.mod, "omo_jit_kernel") != 0 || a.cuModuleGetGlobal(&k.d_scene, &bytes, k.mod, "c_scene") != 0 ||
        bytes != sizeof(OmoDevScene)) {
        ctx->jit_log = "specialised module lacks omo_jit_kernel / c_scene";
        a.cuModuleUnload(k.mod); k.mod = nullptr;
        return nullptr;
    }
    k.smem = spec.smem_bytes; k.item_smem = spec.item_smem; k.sub = spec.sub;
    if (a.cuFuncSetAttribute(k.fn, 8 /* CU_FUNC_ATTRIBUTE_MAX_DYNAMIC_SHARED_SIZE_BYTES */, (int)k.smem) != 0) {
        ctx->jit_log = "cuFuncSetAttribute(max dynamic shared memory) failed";
        a.cuModuleUnload(k.mod); k.mod = nullptr;
        return nullptr;
    }
    a.cuFuncGetAttribute(&k.regs, 4 /* CU_FUNC_ATTRIBUTE_NUM_REGS */, k.fn);
    k.compile_ms = std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0).count();
    k.ok = true;
    return &k;
}

omo_status launch_jit(omo_ctx* ctx, omo_ctx::JitKernel* k, const OmoEvalParams& P, int B, int n_tiles)
{
    omo_jit::Api& a = omo_jit::api();
    if (k->uploaded_version != ctx->scene_version) {
        OMO_CUDA(cudaMemcpyAsync(reinterpret_cast<void*>(k->d_scene), &ctx->hs, sizeof(OmoDevScene), cudaMemcpyHostToDevice, ctx->stream));
        OMO_CUDA(cudaStreamSynchronize(ctx->stream));   // hs is pageable host memory that may change after this call
        k->uploaded_version = ctx->scene_version;
    }
    cudaEvent_t e0 = nullptr, e1 = nullptr;
    if (ctx->timing) {
        OMO_CUDA(cudaEventCreate(&e0)); OMO_CUDA(cudaEventCreate(&e1));
        OMO_CUDA(cudaEventRecord(e0, ctx->stream));
    }
    OmoEvalParams Pc = P;
    Pc.smem_stride = (int)(k->item_smem / sizeof(double));
    void* args[] = {&Pc};
    const int rc = a.cuLaunchKernel(k->fn, (unsigned)((Pc.n_items + k->sub - 1) / k->sub), 1, 1, kBlockThreads * k->sub, 1, 1, (unsigned)k->smem,
                                    reinterpret_cast<omo_jit::CUstream>(ctx->stream), args, nullptr);
    if (rc != 0) return fail(ctx, OMO_CUDA_ERROR, "cuLaunchKernel(omo_jit_kernel) failed with CUresult " + std::to_string(rc));
    if (e0) { OMO_CUDA(cudaEventRecord(e1, ctx->stream)); ctx->timed.emplace_back(e0, e1); }
    ctx->launches += 1;
    ctx->jit_launches += 1;
    return OMO_OK;
}

omo_status build_perm(omo_ctx* ctx);

}  // namespace

// Gather of a multi-process job over NVLink / NVSwitch peer memory (include/omo.h: omo_peer_gather_*).
struct omo_peer_gather {
    omo_ctx* ctx = nullptr;
    int rank = 0, world = 1, capacity = 0;
    unsigned char* d_block = nullptr;          // this rank's allocation: [2][capacity] doubles, then [world] flag words
    void* opened[OMO_MAX_PEERS] = {};          // cudaIpcOpenMemHandle mappings of the other ranks' allocations
    double* buf[OMO_MAX_PEERS] = {};
    unsigned long long* flag[OMO_MAX_PEERS] = {};
    unsigned* d_ticket = nullptr;
    int* h_error = nullptr; int* d_error = nullptr;   // mapped host word: a wait that timed out
    unsigned long long seq = 0;
    bool connected = false;
    long long first = 0;                       // arguments of the call in flight
};

namespace {

size_t peer_flag_offset(const int capacity) { return (((size_t)2 * capacity * sizeof(double)) + 255) & ~(size_t)255; }

// The finish step of a batch launch: finish_kernel, or -- inside omo_peer_gather_eval -- its fusion with the peer exchange.
omo_status launch_finish(omo_ctx* ctx, int B, int n_tiles, double* d_f, double* d_terms)
{
    if (omo_peer_gather* g = ctx->peer) {
        omo::OmoPeerParams G;
        std::memset(&G, 0, sizeof(G));
        for (int p = 0; p < g->world; ++p) { G.buf[p] = g->buf[p]; G.flag[p] = g->flag[p]; }
        G.world = g->world; G.rank = g->rank;
        G.offset = (long long)(g->seq & 1ull) * g->capacity + g->first;
        G.seq = g->seq; G.ticket = g->d_ticket; G.error = g->d_error;
        G.timeout_cycles = 4000000000ll;   // ~2 s at 1.9 GHz
        omo::finish_exchange_kernel<<<std::max((B + 127) / 128, 1), 128, 0, ctx->stream>>>(sc_partial(ctx), sc_dparam(ctx), B, n_tiles, prep_parts(ctx), G);
    } else {
        omo::finish_kernel<<<(B + 127) / 128, 128, 0, ctx->stream>>>(sc_partial(ctx), sc_dparam(ctx), B, n_tiles, prep_parts(ctx), d_f, d_terms);
    }
    OMO_CUDA(cudaGetLastError());
    ctx->launches += 1;
    return OMO_OK;
}

// Launches the objective kernel over B candidates and the finish kernel that adds the tile partials.
// d_item_map (optional): the launch covers only the n_items listed (candidate, tile) pairs and leaves the sums to the caller.
omo_status launch_objective(omo_ctx* ctx, OmoEvalParams& P, int B, bool fk, bool out, double* d_f, double* d_terms,
                            const int* d_item_map = nullptr, int n_items = -1)
{
    const bool ik = ctx->n_ik > 0;
    if (ik || out) fk = true;
    const bool count = P.counters != nullptr && !out;
    const LaunchPlan plan = make_plan(ctx, B, fk, ik, out);
    P.n_tiles = plan.n_tiles; P.T_out = plan.t_out; P.smem_stride = 0; P.n_smem_slots = plan.smem_slots;
    P.item_map = d_item_map;
    P.n_items = d_item_map ? n_items : B * plan.n_tiles;
    omo_status st;
    // large launches run the scene-specialised kernel (auto: >= 64 K sample-evals; OMO_JIT=1 always, OMO_JIT=0 never)
    const int mode = jit_mode(ctx);
    // (sparse column launches always do: a gradient's columns must not switch kernels with the size of a rank's shard)
    const long long sized_as = std::max<long long>(P.n_items, (long long)ctx->batch_hint * plan.n_tiles);
    // sorted curve work lists pay for launches of many work items; built lazily, once per (x_original, window)
    if (!out && mode != 0 && ctx->perm_dirty && ctx->perm_mode != 0 && (long long)P.n_items >= 4096 && !ctx->counting) {
        if ((st = build_perm(ctx)) != OMO_OK) return st;
        P.perm = ctx->perm_valid ? ctx->d_perm : nullptr;
        P.perm_stride = ctx->perm_stride;
    }
    if ((long long)P.n_items < 4096 || count || out) P.perm = nullptr;
    if (!out && mode != 0 && (mode == 1 || d_item_map != nullptr || ctx->prefer_jit || sized_as * plan.t_out >= 65536)) {
        if (omo_ctx::JitKernel* k = acquire_jit(ctx, fk, ik, count)) {
            if ((st = launch_jit(ctx, k, P, B, plan.n_tiles)) != OMO_OK) return st;
            if (d_item_map) return OMO_OK;
            return launch_finish(ctx, B, plan.n_tiles, d_f, d_terms);
        }
        if (mode == 1) return fail(ctx, OMO_CUDA_ERROR, "OMO_JIT=1 but the specialised kernel is unavailable: " + ctx->jit_log);
    }
    if (plan.deep) {
        if (count) return fail(ctx, OMO_BAD_ARG, "work counters are not available for trees that need the deep-stack kernel");
        if (out) st = ik ? launch_objective_t<true, true, true, false, true>(ctx, P, B, plan) : launch_objective_t<true, false, true, false, true>(ctx, P, B, plan);
        else st = ik ? launch_objective_t<true, true, false, false, true>(ctx, P, B, plan) : launch_objective_t<true, false, false, false, true>(ctx, P, B, plan);
    }
    else if (out) st = ik ? launch_objective_t<true, true, true, false>(ctx, P, B, plan) : launch_objective_t<true, false, true, false>(ctx, P, B, plan);
    else if (ik) st = count ? launch_objective_t<true, true, false, true>(ctx, P, B, plan) : launch_objective_t<true, true, false, false>(ctx, P, B, plan);
    else if (fk) st = count ? launch_objective_t<true, false, false, true>(ctx, P, B, plan) : launch_objective_t<true, false, false, false>(ctx, P, B, plan);
    else st = count ? launch_objective_t<false, false, false, true>(ctx, P, B, plan) : launch_objective_t<false, false, false, false>(ctx, P, B, plan);
    if (st != OMO_OK) return st;
    if (d_item_map) return OMO_OK;
    return launch_finish(ctx, B, plan.n_tiles, d_f, d_terms);
}

void fill_params(const omo_ctx* ctx, OmoEvalParams& P)
{
    std::memset(&P, 0, sizeof(P));
    P.prep = sc_prep(ctx); P.flags = sc_flags(ctx); P.seg = sc_seg(ctx); P.tile_seg = sc_tile_seg(ctx); P.tgrid = ctx->d_tgrid; P.wcost = ctx->d_wcost; P.partial = sc_partial(ctx);
    P.N_s = ctx->N_s; P.n_cp = ctx->n_cp;
    const double dt = ctx->dt;
    P.dt = dt; P.inv_dt = 1.0 / dt; P.inv_2dt = 1.0 / (2.0 * dt); P.inv_12dt = 1.0 / (12.0 * dt);
    P.inv_dt2 = 1.0 / (dt * dt); P.inv_2dt2 = 1.0 / (2.0 * dt * dt); P.inv_12dt2 = 1.0 / (12.0 * dt * dt);
    P.w_traj = ctx->w_traj;
    P.counters = ctx->counting ? ctx->d_counters : nullptr;
    P.perm = (ctx->perm_valid && !ctx->perm_dirty) ? ctx->d_perm : nullptr;
    P.perm_stride = ctx->perm_stride;
}

omo_status run_prepare(omo_ctx* ctx, const double* dX, int B)
{
    omo::OmoPrepParams Q;
    Q.X = dX; Q.x0 = ctx->d_x0; Q.wparam = ctx->d_wparam;
    Q.cp_of_pos = ctx->d_cp_of_pos; Q.first_of_pos = ctx->d_first_of_pos; Q.last_of_pos = ctx->d_last_of_pos; Q.next_same = ctx->d_next_same;
    Q.prep = sc_prep(ctx); Q.flags = sc_flags(ctx); Q.dparam = sc_dparam(ctx);
    Q.n_cp = ctx->n_cp; Q.D = ctx->D; Q.cyclic = ctx->cyclic;
    Q.seg = sc_seg(ctx); Q.tile_seg = sc_tile_seg(ctx); Q.tgrid = ctx->d_tgrid;
    Q.N_s = ctx->N_s; Q.T_out = ctx->T_out; Q.n_tiles = ctx->n_tiles; Q.n_keyed = ctx->hs.n_keyed;
    Q.chunk = std::min(kPrepChunk, std::max(ctx->n_cp, 1));
    const size_t sh = (size_t)6 * (Q.chunk + 1) * sizeof(double);
    // bit 0 of every flag word = "sorted": preset here when several blocks share a candidate, written by the one block otherwise
    if (prep_parts(ctx) > 1) OMO_CUDA(cudaMemsetAsync(sc_flags(ctx), 1, (size_t)B * sizeof(unsigned), ctx->stream));
    // long chains (several passes over the block's positions: two dependent global round trips and a tan() each) take 256 threads;
    // up to 256 positions stay with 128, which keeps more candidates resident per SM (measured on B200: 4096 candidates of 160
    // control points 49 -> 41 us with 128 threads, 45 us with 256; 256 candidates of 37 989: 1.6 -> 0.7 ms with 256)
    const unsigned prep_threads = (Q.chunk + 1 > 256) ? 256u : 128u;
    omo::prepare_kernel<<<dim3((unsigned)B, (unsigned)prep_parts(ctx)), prep_threads, sh, ctx->stream>>>(Q);
    OMO_CUDA(cudaGetLastError());
    ctx->launches += 1;
    return OMO_OK;
}

omo_status ensure_outputs(omo_ctx* ctx)
{
    if (ctx->out_cap_samples >= ctx->N_s && ctx->d_oq) return OMO_OK;
    const size_t n = (size_t)ctx->N_s * 3 * ctx->J;
    omo_status st;
    if ((st = dev_alloc(ctx, &ctx->d_oq, n)) != OMO_OK) return st;
    if ((st = dev_alloc(ctx, &ctx->d_oqd, n)) != OMO_OK) return st;
    if ((st = dev_alloc(ctx, &ctx->d_oqdd, n)) != OMO_OK) return st;
    if ((st = dev_alloc(ctx, &ctx->d_otau, n)) != OMO_OK) return st;
    if ((st = dev_alloc(ctx, &ctx->d_op, n)) != OMO_OK) return st;
    if ((st = dev_alloc(ctx, &ctx->d_olocal, (size_t)ctx->N_s)) != OMO_OK) return st;
    ctx->out_cap_samples = ctx->N_s;
    return OMO_OK;
}

// One candidate with per-sample outputs (used for the OptimizationCache, positions and p_original).
omo_status run_single_out(omo_ctx* ctx, const double* d_x, bool with_p_orig, double* d_root = nullptr)
{
    omo_status st;
    if ((st = ensure_capacity(ctx, 1)) != OMO_OK) return st;
    if ((st = ensure_outputs(ctx)) != OMO_OK) return st;
    if ((st = activate_scene(ctx)) != OMO_OK) return st;
    if ((st = run_prepare(ctx, d_x, 1)) != OMO_OK) return st;
    OmoEvalParams P;
    fill_params(ctx, P);
    P.p_orig = with_p_orig ? ctx->d_p_orig : nullptr;
    P.o_q = ctx->d_oq; P.o_qd = ctx->d_oqd; P.o_qdd = ctx->d_oqdd; P.o_tau = ctx->d_otau; P.o_local = ctx->d_olocal; P.o_p = ctx->d_op;
    P.o_root = d_root;
    return launch_objective(ctx, P, 1, true, true, ctx->d_f, ctx->d_terms);
}

// Makes `grid` the sample times of every following launch (time-ordered within each warp: the segment search of the
// table-driven kernel warm-starts from its warp's first sample). Everything sized by N_s / n_tiles is invalidated.
omo_status install_grid(omo_ctx* ctx, std::vector<double>& grid, double dt)
{
    OMO_CUDA(cudaStreamSynchronize(ctx->stream));
    ctx->tgrid.swap(grid);
    ctx->N_s = (int)ctx->tgrid.size();
    ctx->dt = dt;
    const int t_max = kBlockThreads - 4;
    ctx->n_tiles = (ctx->N_s + t_max - 1) / t_max;
    ctx->T_out = (ctx->N_s + ctx->n_tiles - 1) / ctx->n_tiles;
    omo_status st = upload(ctx, &ctx->d_tgrid, ctx->tgrid);
    if (st != OMO_OK) return st;
    if (!ctx->residues_valid || ctx->residues_dt != dt) { refresh_residues(ctx); ctx->residues_valid = true; ctx->residues_dt = dt; }
    ctx->wcost_dirty = true; ctx->p_orig_dirty = true;
    ctx->sp_valid = false;  // the sparse-gradient item lists are per time grid
    // work / output buffers are kept: ensure_capacity and ensure_outputs grow them when the new grid needs more tiles / samples
    return OMO_OK;
}

// Sorted curve work lists (OmoEvalParams::perm). For every tile and curve group, the (channel, sample) items of the tile's window
// ordered by the Newton iterations (curve.cpp:43-54) x_original's motion needs for them -- a host-side count, statistics only: the
// candidates of a batch perturb tangents and handle lengths around x_original, so their counts follow its counts closely
// (profiles/newton_sim.py: 24 % fewer lockstep rounds on stick240, 20 % on the 64-joint tree). Correctness does not depend on it.
omo_status build_perm(omo_ctx* ctx)
{
    if (ctx->perm_mode < 0) ctx->perm_mode = omo_jit::sorted_lists_enabled(ctx->hs) ? 1 : 0;
    ctx->perm_dirty = false; ctx->perm_valid = false;
    const OmoDevScene& S = ctx->hs;
    const int nk = S.n_keyed, N = ctx->N_s, n_tiles = ctx->n_tiles, T_out = ctx->T_out;
    // Measured on B200: stick240 (one group of 6 curves) 0.927 -> 0.885 ms; the 64-joint tree (16 groups of 12: a 4 MB list, four
    // triples per group to amortise the per-lane lookups over) 14.9 -> 15.2 ms. Scenes with many curve groups keep the per-sample order.
    if (ctx->perm_mode != 1 || nk == 0 || ctx->n_cp == 0 || N < 3) return OMO_OK;
    // predicted iteration count of every (keyed channel, sample)
    std::vector<unsigned char> cnt((size_t)nk * N, 0);
    const std::vector<double>& x = ctx->x0;
    for (int i = 0; i < nk; ++i) {
        const int ch = S.grp_list[i], pos0 = S.chan_pos[ch], n = S.chan_n[ch];
        std::vector<double> T(n);
        for (int p = 0; p < n; ++p) T[p] = x[5 * ctx->cp_of_pos[pos0 + p]];
        if (!std::is_sorted(T.begin(), T.end())) continue;   // list-order search: no prediction, counts stay 0
        for (int s = 0; s < N; ++s) {
            const double t = ctx->tgrid[s];
            const int m = (int)(std::upper_bound(T.begin(), T.end(), t) - T.begin());
            if (m == 0 || m == n) continue;
            const int cs = ctx->cp_of_pos[pos0 + m - 1], ce = ctx->cp_of_pos[pos0 + m];
            const double p0 = x[5 * cs], p1 = x[5 * cs] + x[5 * cs + 4], p2 = x[5 * ce] + x[5 * ce + 3], p3 = x[5 * ce];
            const double a = -p0 + 3.0 * p1 - 3.0 * p2 + p3, b = 3.0 * (p0 - 2.0 * p1 + p2), c = -3.0 * p0 + 3.0 * p1, d = p0 - t;
            double u = 0.5;
            int k = 0;
            for (; k < 50; ++k) {
                const double f = u * (u * (a * u + b) + c) + d, fp = u * (3.0 * a * u + 2.0 * b) + c;
                if (std::fabs(f) < 1e-09) break;
                u -= f / fp;
            }
            cnt[(size_t)i * N + s] = (unsigned char)std::min(k + 1, 255);   // +1: an in-range curve costs its Bernstein evaluation even at 0 iterations
        }
    }
    int stride = 0;
    for (int g = 0; g < S.n_groups; ++g) stride += ((S.grp_begin[g + 1] - S.grp_begin[g] + 2) / 3) * 3 * kBlockThreads;
    std::vector<unsigned short> perm((size_t)n_tiles * stride, 0xffff);
    std::vector<std::pair<int, unsigned short>> items;
    for (int tile = 0; tile < n_tiles; ++tile) {
        size_t off = (size_t)tile * stride;
        for (int g = 0; g < S.n_groups; ++g) {
            const int g0 = S.grp_begin[g], ng = S.grp_begin[g + 1] - g0;
            items.clear();
            for (int lc = 0; lc < ng; ++lc)
                for (int si = 0; si < kBlockThreads; ++si) {
                    const int s = std::min(std::max(tile * T_out - 2 + si, 0), N - 1);
                    items.emplace_back(-(int)cnt[(size_t)(g0 + lc) * N + s], (unsigned short)((lc << 7) | si));
                }
            std::stable_sort(items.begin(), items.end(), [](const std::pair<int, unsigned short>& p, const std::pair<int, unsigned short>& q) { return p.first < q.first; });
            // chunk q of 96 sorted items goes to triple q / 4 of warp q % 4 -- in snake order (odd rows reversed), so that the four warps
            // of the block carry equal shares of the expensive curves and meet at the group's barrier together
            for (size_t k = 0; k < items.size(); ++k) {
                const size_t q = k / 96, row = q / 4, w = (row & 1) ? 3 - (q % 4) : (q % 4);
                perm[off + (row * 4 + w) * 96 + (k % 96)] = items[k].second;
            }
            off += (size_t)((ng + 2) / 3) * 3 * kBlockThreads;
        }
    }
    omo_status st;
    if (perm.size() > ctx->perm_cap) {
        if ((st = dev_alloc(ctx, &ctx->d_perm, perm.size())) != OMO_OK) return st;
        ctx->perm_cap = perm.size();
    }
    OMO_CUDA(cudaMemcpyAsync(ctx->d_perm, perm.data(), perm.size() * sizeof(unsigned short), cudaMemcpyHostToDevice, ctx->stream));
    OMO_CUDA(cudaStreamSynchronize(ctx->stream));   // perm is a local
    ctx->perm_stride = stride;
    ctx->perm_valid = true;
    return OMO_OK;
}

omo_status ensure_ready(omo_ctx* ctx)
{
    if (!ctx->have_window) return fail(ctx, OMO_NOT_READY, "omo_set_window has not been called");
    if (!ctx->have_weights) return fail(ctx, OMO_NOT_READY, "omo_set_weights has not been called");
    if (!ctx->have_original) return fail(ctx, OMO_NOT_READY, "omo_set_original has not been called");
    OMO_CUDA(cudaSetDevice(ctx->device));
    omo_status st;
    if (ctx->wcost_dirty) {
        // Core::GetCostWeight (core.cpp:191-199) on the host's libm, exactly as the reference evaluates it
        ctx->wcost.resize(ctx->N_s);
        for (int i = 0; i < ctx->N_s; ++i) {
            double w = 1.0;
            for (const omo_gauss& k : ctx->kernels) w -= k.amp * std::exp(-(ctx->tgrid[i] - k.mu) * (ctx->tgrid[i] - k.mu) / (2.0 * k.sigma * k.sigma));
            ctx->wcost[i] = std::max(w, 0.0);
        }
        if ((st = upload(ctx, &ctx->d_wcost, ctx->wcost)) != OMO_OK) return st;
        ctx->wcost_dirty = false;
    }
    if (ctx->p_orig_dirty) {
        // OriginalMotionCache (core_rbdl.cpp:357-368): positions of x_original, kept transposed [3J][N_s]
        if ((st = dev_alloc(ctx, &ctx->d_p_orig, (size_t)ctx->N_s * 3 * ctx->J)) != OMO_OK) return st;
        ctx->hs.static_tsq = 0.0;
        if ((st = run_single_out(ctx, ctx->d_x0, false)) != OMO_OK) return st;
        // candidate-independent share of sum tau^2: the pruned bodies' torques at an interior sample (full-tree run above)
        if (ctx->N_s >= 5) {
            std::vector<double> tau(3 * ctx->J);
            OMO_CUDA(cudaMemcpyAsync(tau.data(), ctx->d_otau + (size_t)2 * 3 * ctx->J, tau.size() * sizeof(double), cudaMemcpyDeviceToHost, ctx->stream));
            OMO_CUDA(cudaStreamSynchronize(ctx->stream));
            double acc = 0.0;
            for (int j = 1; j < ctx->J; ++j)
                if (ctx->pruned[j]) for (int k = 0; k < 3; ++k) acc += tau[3 * j + k] * tau[3 * j + k];
            ctx->hs.static_tsq = acc;
        }
        ctx->scene_dirty = true;
        ctx->scene_version += 1;
        const int n = ctx->N_s * 3 * ctx->J;
        omo::transpose_kernel<<<(n + 255) / 256, 256, 0, ctx->stream>>>(ctx->d_op, ctx->N_s, 3 * ctx->J, ctx->d_p_orig);
        OMO_CUDA(cudaGetLastError());
        ctx->launches += 1;
        ctx->p_orig_dirty = false;
        ctx->perm_dirty = true;   // x_original or the window changed: the sorted curve lists follow
    }
    return activate_scene(ctx);
}

// Sparse finite-difference columns. Fills h_lo / h_hi (tile range each of the columns [col_begin, col_begin + col_count)
// must be evaluated on; lo > hi = none) and h_items (row * n_tiles + tile for those tiles, then every tile of the extra base
// row `col_count`). A control point at chain position pos shapes the segments pos-1 -> pos and pos -> pos+1 of its channel
// (curve.cpp:57-93: each segment reads its two end points only; before the first / after the last key the end value holds),
// so a sample is affected iff T[pos-1] <= t < T[pos+1] (open ends at a channel's first / last point), and through the
// 5-point stencils (core_rbdl.cpp:28-95) so are its +-2 neighbours. A column that moves a key time itself, or the averaged
// end tangent of a cyclic curve, keeps all its tiles. Returns false -- evaluate densely -- only when a chain of x is not
// time-sorted (list-order segment search) or the window has a single tile.
bool affected_tiles(omo_ctx* ctx, const double* x, const int* free_idx, int F, int col_begin, int col_count)
{
    const int n_cp = ctx->n_cp, N = ctx->N_s, T_out = ctx->T_out, n_tiles = ctx->n_tiles;
    if (n_cp == 0 || N < 5 || n_tiles < 2) return false;
    if (ctx->pos_of_cp.size() != (size_t)n_cp) {
        ctx->pos_of_cp.assign(n_cp, 0);
        for (int pos = 0; pos < n_cp; ++pos) ctx->pos_of_cp[ctx->cp_of_pos[pos]] = pos;
    }
    {
        bool same = ctx->sp_valid && ctx->sp_begin == col_begin && ctx->sp_count == col_count && ctx->sp_free.size() == (size_t)F &&
                    std::equal(ctx->sp_free.begin(), ctx->sp_free.end(), free_idx);
        for (int pos = 0; same && pos < n_cp; ++pos) same = ctx->sp_times[pos] == x[5 * ctx->cp_of_pos[pos]];
        if (same) return true;   // device copies are current (sp_uploaded)
    }
    ctx->sp_valid = false; ctx->sp_uploaded = false;
    for (int pos = 0; pos + 1 < n_cp; ++pos)
        if (ctx->next_same[pos] && !(x[5 * ctx->cp_of_pos[pos]] <= x[5 * ctx->cp_of_pos[pos + 1]])) return false;
    ctx->h_lo.assign(col_count, 0); ctx->h_hi.assign(col_count, n_tiles - 1);
    ctx->h_items.clear();
    const std::vector<double>& tg = ctx->tgrid;
    for (int lc = 0; lc < col_count; ++lc) {
        const int col = col_begin + lc;
        int lo = 0, hi = n_tiles - 1;
        if (col > 0) {
            const int idx = free_idx[(col - 1) >> 1], cp = idx / 5, field = idx % 5, pos = ctx->pos_of_cp[cp];
            const bool cyc_end = ctx->cyclic && field == 2 && (cp == ctx->first_of_pos[pos] || cp == ctx->last_of_pos[pos]);
            // A moved key time re-shapes the segment search, a cyclic curve's end tangent is shared by both ends: such a column
            // keeps every tile (lo = 0, hi = n_tiles - 1) but stays in the item-map launch, so that f(x + h e_k) and f(x - h e_k)
            // come from the same kernel whichever rank's shard they fall into.
            if (!(field == 0 || cyc_end)) {
                const double t_lo = (pos > 0 && ctx->next_same[pos - 1]) ? x[5 * ctx->cp_of_pos[pos - 1]] : -INFINITY;
                const double t_hi = ctx->next_same[pos] ? x[5 * ctx->cp_of_pos[pos + 1]] : INFINITY;
                // samples with t_lo <= t <= t_hi, one sample of slack each side, then the stencil's two
                const int s_lo = (int)(std::lower_bound(tg.begin(), tg.end(), t_lo) - tg.begin()) - 3;
                const int s_hi = (int)(std::upper_bound(tg.begin(), tg.end(), t_hi) - tg.begin()) - 1 + 3;
                if (s_hi < 0 || s_lo > N - 1) { lo = 0; hi = 0; }   // out of reach of the window: one tile anyway (it equals the base point's), so that every row has a block that books its D_param
                else { lo = std::max(s_lo, 0) / T_out; hi = std::min(std::min(s_hi, N - 1) / T_out, n_tiles - 1); }
            }
        }
        ctx->h_lo[lc] = lo; ctx->h_hi[lc] = hi;
        for (int t = lo; t <= hi; ++t) ctx->h_items.push_back(lc * n_tiles + t);
    }
    for (int t = 0; t < n_tiles; ++t) ctx->h_items.push_back(col_count * n_tiles + t);
    ctx->sparse_items = ctx->h_items.size();
    ctx->dense_items = (unsigned long long)col_count * n_tiles;
    ctx->sp_times.resize(n_cp);
    for (int pos = 0; pos < n_cp; ++pos) ctx->sp_times[pos] = x[5 * ctx->cp_of_pos[pos]];
    ctx->sp_free.assign(free_idx, free_idx + F);
    ctx->sp_begin = col_begin; ctx->sp_count = col_count; ctx->sp_valid = true;
    return true;
}

// Everything a captured single-candidate graph bakes in: buffer addresses, launch geometry, by-value kernel parameters.
// (What the kernels READ through those addresses -- weights, x_original, p_original, the constant scene table -- may change freely.)
std::vector<unsigned long long> single_key(const omo_ctx* ctx)
{
    auto bits = [](double v) { unsigned long long u; std::memcpy(&u, &v, sizeof(u)); return u; };
    const void* ptrs[] = {ctx->d_X, ctx->d_prep, ctx->d_dparam, ctx->d_partial, ctx->d_f, ctx->d_terms, ctx->d_flags, ctx->d_seg, ctx->d_tile_seg, ctx->d_tgrid,
                          ctx->d_wcost, ctx->d_wparam, ctx->d_x0, ctx->d_p_orig, ctx->d_cp_of_pos, ctx->d_first_of_pos, ctx->d_last_of_pos, ctx->d_next_same,
                          ctx->h_single, ctx->d_single, ctx->d_counters};
    std::vector<unsigned long long> k;
    for (const void* p : ptrs) k.push_back((unsigned long long)reinterpret_cast<uintptr_t>(p));
    const long long ints[] = {ctx->N_s, ctx->n_tiles, ctx->T_out, ctx->n_cp, ctx->D, ctx->cyclic, ctx->counting ? 1 : 0, ctx->jit_mode, ctx->n_ik, ctx->cap};
    for (long long v : ints) k.push_back((unsigned long long)v);
    k.push_back(bits(ctx->dt)); k.push_back(bits(ctx->w_traj)); k.push_back(ctx->scene_version);
    return k;
}

// ---- self-preparing launches ------------------------------------------------------------------------------------------------
constexpr int kSelfMaxCp = 192;     // control points whose prepared form + segment records fit beside the walk's shared memory
constexpr int kSelfMaxRows = 2048;

bool self_usable(omo_ctx* ctx)
{
    if (ctx->self_mode < 0) { const char* e = std::getenv("OMO_SELF"); ctx->self_mode = (e && std::atoi(e) == 0) ? 0 : 1; }
    return ctx->self_mode == 1 && jit_mode(ctx) != 0 && !ctx->counting && ctx->n_cp >= 1 && ctx->n_cp <= kSelfMaxCp;
}

// rows: rows of the launch; slots: doubles the mapped host buffer must hold (0: rows)
omo_status ensure_self_buffers(omo_ctx* ctx, int rows, int slots = 0)
{
    omo_status st;
    if (!ctx->d_self_counter) {
        if ((st = dev_alloc(ctx, &ctx->d_self_counter, 1)) != OMO_OK) return st;
        OMO_CUDA(cudaMemsetAsync(ctx->d_self_counter, 0, sizeof(unsigned), ctx->stream));
    }
    if (rows > ctx->self_rows_cap) {
        const int cap = std::max(rows, 256);
        if ((st = dev_alloc(ctx, &ctx->d_self_dparam, (size_t)cap)) != OMO_OK) return st;
        ctx->self_rows_cap = cap;
    }
    if (std::max(std::max(rows, slots), 4) > ctx->self_out_cap) {
        const int cap = std::max(std::max(rows, slots), 256);
        OMO_CUDA(cudaStreamSynchronize(ctx->stream));
        if (ctx->h_self_out) { cudaFreeHost(ctx->h_self_out); ctx->h_self_out = nullptr; }
        OMO_CUDA(cudaHostAlloc((void**)&ctx->h_self_out, (size_t)cap * sizeof(double), cudaHostAllocMapped));
        OMO_CUDA(cudaHostGetDevicePointer((void**)&ctx->d_self_out, ctx->h_self_out, 0));
        ctx->self_out_cap = cap;
    }
    return OMO_OK;
}

// Waits for the self-preparing launch that was queued last. Where it writes into the mapped host buffer, the host has filled the
// output slots with a sentinel (a NaN payload no computation produces) before the launch and reads them in a loop until every one has
// been overwritten: each value is one 8-byte store, there a PCIe write after the kernel produced it -- no completion word, no
// system-scope fence in the kernel, no trip through the driver's completion path (fox-tail callback 31.0 -> 30.1 us with a completion
// word, -> 28 us with the sentinel). The stream is asked now and then (a faulted kernel never writes), and is what is waited on when
// polling is off (OMO_SPIN=0) or the output lives elsewhere.
constexpr unsigned long long kSelfSentinel = 0x7ff8dead5e1fbeefull;

omo_status wait_self(omo_ctx* ctx)
{
    if (ctx->self_spin_slots > 0 && ctx->self_host_rows > 0) {
        // small gradient reduced on the host: wait for every listed (row, tile) pair and every row's D_param, then finish_sparse_kernel's sums
        const int rows = ctx->self_host_rows, base = ctx->self_host_base, nt = ctx->self_host_ntiles, off = ctx->self_host_off;
        ctx->self_spin_slots = 0; ctx->self_host_rows = 0;
        const volatile unsigned long long* out = reinterpret_cast<const volatile unsigned long long*>(ctx->h_self_out);
        const std::vector<int>& items = ctx->h_items;
        size_t next = 0;
        int next_row = 0;
        auto arrived = [&]() {
            while (next < items.size() && out[off + 2 * items[next]] != kSelfSentinel && out[off + 2 * items[next] + 1] != kSelfSentinel) ++next;
            if (next < items.size()) return false;
            while (next_row < rows - 1 && out[off + 2 * rows * nt + next_row] != kSelfSentinel) ++next_row;   // (the base row's D_param is not used)
            return next_row == rows - 1;
        };
        for (unsigned spin = 1;; ++spin) {
            if (arrived()) break;
            if ((spin & 2047u) == 0u) {
                const cudaError_t q = cudaStreamQuery(ctx->stream);
                if (q == cudaSuccess) {
                    if (arrived()) break;
                    return fail(ctx, OMO_CUDA_ERROR, "self-preparing kernel finished without writing all of its outputs");
                }
                if (q != cudaErrorNotReady) { ctx->err = std::string("self-preparing launch: ") + cudaGetErrorString(q); return OMO_CUDA_ERROR; }
            }
#if defined(__x86_64__) || defined(__i386__)
            __builtin_ia32_pause();
#endif
        }
        volatile double* h = ctx->h_self_out;
        const volatile double* part = h + off;
        const volatile double* dpar = h + off + 2 * rows * nt;
        for (int r = 0; r < rows - 1; ++r) {
            const int l = ctx->h_lo[r], hi = ctx->h_hi[r];
            double C = 0.0, Dt = 0.0;
            for (int tl = 0; tl < nt; ++tl) {
                const int row = (tl >= l && tl <= hi) ? r : base;
                C = C + part[((size_t)row * nt + tl) * 2];
                Dt = Dt + part[((size_t)row * nt + tl) * 2 + 1];
            }
            double Dp = 0.0;
            Dp = Dp + dpar[r];
            h[r] = (C + Dp) + Dt;
        }
        return OMO_OK;
    }
    if (ctx->self_spin_slots > 0) {
        const int n = ctx->self_spin_slots, host_tiles = ctx->self_host_tiles;
        ctx->self_spin_slots = 0; ctx->self_host_tiles = 0;
        const volatile unsigned long long* out = reinterpret_cast<const volatile unsigned long long*>(ctx->h_self_out);
        int next = host_tiles > 0 ? 4 : 0;    // slots below `next` have arrived
        auto reduce_on_host = [&]() {
            // finish_kernel's sums, same order: C and D_traj over the tiles from 0.0 up, then f = C + D_param + D_traj
            volatile double* h = ctx->h_self_out;
            double C = 0.0, Dt = 0.0;
            for (int t = 0; t < host_tiles; ++t) { C = C + h[4 + 2 * t]; Dt = Dt + h[4 + 2 * t + 1]; }
            double Dp = 0.0;
            Dp = Dp + h[4 + 2 * host_tiles];
            const double f = (C + Dp) + Dt;
            h[0] = f; h[1] = C; h[2] = Dp; h[3] = Dt;
        };
        for (unsigned spin = 1;; ++spin) {
            while (next < n && out[next] != kSelfSentinel) ++next;
            if (next == n) { if (host_tiles > 0) reduce_on_host(); return OMO_OK; }
            if ((spin & 2047u) == 0u) {
                const cudaError_t q = cudaStreamQuery(ctx->stream);
                if (q == cudaSuccess) {
                    while (next < n && out[next] != kSelfSentinel) ++next;
                    if (next == n) { if (host_tiles > 0) reduce_on_host(); return OMO_OK; }
                    return fail(ctx, OMO_CUDA_ERROR, "self-preparing kernel finished without writing all of its outputs");
                }
                if (q != cudaErrorNotReady) { ctx->err = std::string("self-preparing launch: ") + cudaGetErrorString(q); return OMO_CUDA_ERROR; }
            }
#if defined(__x86_64__) || defined(__i386__)
            __builtin_ia32_pause();
#endif
        }
    }
    OMO_CUDA(cudaStreamSynchronize(ctx->stream));
    return OMO_OK;
}

// One launch: rows of the finite-difference list (or the one candidate x) x their tiles. `out` must be device-visible.
// item_map / lo / hi as in the sparse column path (null: every tile of every row); base_row = -1 without a base row.
omo_status launch_self(omo_ctx* ctx, const double* x_host, int rows, int base_row, const int* d_item_map, int n_items, const int* d_lo, const int* d_hi,
                       const int* d_free, double h, int col_begin, double* out, bool* launched)
{
    *launched = false;
    const bool ik = ctx->n_ik > 0;
    const bool fk = (ctx->w_traj != 0.0) || ik;
    omo_ctx::JitKernel* k = acquire_jit(ctx, fk, ik, false, true);
    if (!k) return OMO_OK;    // the caller takes the general path
    omo_status st;
    if ((st = ensure_capacity(ctx, rows)) != OMO_OK) return st;
    if ((st = ensure_self_buffers(ctx, rows)) != OMO_OK) return st;
    OmoEvalParams P;
    fill_params(ctx, P);
    P.counters = nullptr;
    P.p_orig = (ctx->w_traj != 0.0) ? ctx->d_p_orig : nullptr;
    const LaunchPlan plan = make_plan(ctx, rows, fk, ik, false);
    P.n_tiles = plan.n_tiles; P.T_out = plan.t_out; P.smem_stride = 0; P.n_smem_slots = plan.smem_slots;
    P.item_map = d_item_map;
    P.n_items = d_item_map ? n_items : rows * plan.n_tiles;
    P.self_free = d_free; P.self_lo = d_lo; P.self_hi = d_hi;
    P.self_cp_of_pos = ctx->d_cp_of_pos; P.self_first_of_pos = ctx->d_first_of_pos; P.self_last_of_pos = ctx->d_last_of_pos; P.self_next_same = ctx->d_next_same;
    P.self_x0 = ctx->d_x0; P.self_wparam = ctx->d_wparam;
    P.self_dparam = ctx->d_self_dparam; P.self_out = out; P.self_counter = ctx->d_self_counter;
    P.self_h = h; P.self_col_begin = col_begin; P.self_rows = rows; P.self_base_row = base_row; P.self_cyclic = ctx->cyclic; P.self_D = ctx->D;
    // results into the context's mapped host buffer: the host polls the output slots themselves (wait_self) instead of paying the stream
    // synchronisation's round trip through the driver
    if (ctx->self_spin_mode < 0) { const char* e = std::getenv("OMO_SPIN"); ctx->self_spin_mode = (e && std::atoi(e) == 0) ? 0 : 1; }
    ctx->self_spin_slots = 0; ctx->self_host_tiles = 0; ctx->self_host_rows = 0;
    if (out == ctx->d_self_out && ctx->self_spin_mode == 1 && !ctx->timing) {
        unsigned long long* slots = reinterpret_cast<unsigned long long*>(ctx->h_self_out);
        if (rows == 1 && d_item_map == nullptr && 5 + 2 * plan.n_tiles <= ctx->self_out_cap) {
            // ONE candidate: every block writes its tile's (C, D_traj) pair -- and the block of tile 0 D_param -- straight into the mapped
            // buffer (slots 4 .., behind the four result slots), and the host adds them in finish_kernel's order (wait_self). The kernel's
            // tail -- fence, ticket, barrier, the last block's loads -- is gone from every BOBYQA callback.
            P.partial = ctx->d_self_out + 4;
            P.self_dparam = ctx->d_self_out + 4 + 2 * plan.n_tiles;
            P.self_counter = nullptr;
            ctx->self_host_tiles = plan.n_tiles;
            for (int i = 4; i < 5 + 2 * plan.n_tiles; ++i) slots[i] = kSelfSentinel;
            ctx->self_spin_slots = 5 + 2 * plan.n_tiles;
        } else if (rows > 1 && d_item_map != nullptr && base_row == rows - 1 && ctx->sp_valid && (int)ctx->h_lo.size() == rows - 1 &&
                   ((rows + 3) & ~3) + 2 * rows * plan.n_tiles + rows <= ctx->self_out_cap) {
            // A small gradient (sparse finite-difference rows + the base row): the listed (row, tile) blocks write their pairs, and each
            // row's first listed block its D_param, into the mapped buffer behind the result slots; the host applies
            // finish_sparse_kernel's rule (tiles outside [lo, hi] of a row come from the base row) -- same numbers, same order.
            const int off = (rows + 3) & ~3;
            P.partial = ctx->d_self_out + off;
            P.self_dparam = ctx->d_self_out + off + 2 * rows * plan.n_tiles;
            P.self_counter = nullptr;
            for (const int item : ctx->h_items) { slots[off + 2 * item] = kSelfSentinel; slots[off + 2 * item + 1] = kSelfSentinel; }
            for (int r = 0; r < rows; ++r) slots[off + 2 * rows * plan.n_tiles + r] = kSelfSentinel;
            ctx->self_host_rows = rows; ctx->self_host_base = base_row; ctx->self_host_ntiles = plan.n_tiles; ctx->self_host_off = off;
            ctx->self_spin_slots = 1;   // wait_self polls the listed slots itself
        } else {
            const int n_out = rows == 1 ? 4 : (base_row >= 0 ? rows - 1 : rows);   // what self_finish writes: (f, C, D_param, D_traj) or one f per row
            for (int i = 0; i < n_out; ++i) slots[i] = kSelfSentinel;
            ctx->self_spin_slots = n_out;
        }
    }
    omo_jit::Api& a = omo_jit::api();
    if (k->uploaded_version != ctx->scene_version) {
        OMO_CUDA(cudaMemcpyAsync(reinterpret_cast<void*>(k->d_scene), &ctx->hs, sizeof(OmoDevScene), cudaMemcpyHostToDevice, ctx->stream));
        OMO_CUDA(cudaStreamSynchronize(ctx->stream));
        k->uploaded_version = ctx->scene_version;
    }
    cudaEvent_t e0 = nullptr, e1 = nullptr;
    if (ctx->timing) {
        OMO_CUDA(cudaEventCreate(&e0)); OMO_CUDA(cudaEventCreate(&e1));
        OMO_CUDA(cudaEventRecord(e0, ctx->stream));
    }
    void* args[] = {&P, const_cast<double*>(x_host)};   // second parameter: the descriptor itself, D doubles by value
    const int rc = a.cuLaunchKernel(k->fn, (unsigned)P.n_items, 1, 1, kBlockThreads, 1, 1, (unsigned)k->smem, reinterpret_cast<omo_jit::CUstream>(ctx->stream), args, nullptr);
    if (rc != 0) return fail(ctx, OMO_CUDA_ERROR, "cuLaunchKernel(self-preparing omo_jit_kernel) failed with CUresult " + std::to_string(rc));
    if (e0) { OMO_CUDA(cudaEventRecord(e1, ctx->stream)); ctx->timed.emplace_back(e0, e1); }
    ctx->launches += 1; ctx->jit_launches += 1; ctx->self_launches += 1;
    *launched = true;
    return OMO_OK;
}

omo_status eval_device_impl(omo_ctx* ctx, const double* dX, int B, double* d_f, double* d_terms)
{
    omo_status st;
    if ((st = ensure_capacity(ctx, B)) != OMO_OK) return st;
    if ((st = run_prepare(ctx, dX, B)) != OMO_OK) return st;
    OmoEvalParams P;
    fill_params(ctx, P);
    const bool fk = (ctx->w_traj != 0.0) || ctx->n_ik > 0;
    P.p_orig = (ctx->w_traj != 0.0) ? ctx->d_p_orig : nullptr;
    return launch_objective(ctx, P, B, fk, false, d_f, d_terms);
}

// Host rows of width W (W = D: whole candidates into d_X; W = F with `expand`: free entries into d_Xfree, expanded on the device)
// copied and evaluated chunk by chunk; f (and the terms) copied back; synchronised on return.
// d_f_dst != nullptr: the values stay on the device (d_f_dst[B], optionally d_terms_dst[3B]) and nothing is copied back or waited for.
omo_status eval_host_pipelined(omo_ctx* ctx, const double* X, int B, int W, double* d_stage, bool expand, int F, double* f_out, omo_terms* terms_out,
                               double* d_f_dst = nullptr, double* d_terms_dst = nullptr)
{
    double* const d_f = d_f_dst ? d_f_dst : ctx->d_f;
    double* const d_terms = d_f_dst ? d_terms_dst : (terms_out ? ctx->d_terms : nullptr);
    omo_status st;
    // Large batches travel in chunks on a second stream so that the PCIe copy of chunk i+1 overlaps the kernels of chunk i
    // (the copy of 4096 stick-figure candidates takes about half as long as evaluating them).
    const size_t bytes = (size_t)B * W * sizeof(double);
    // Chunk sizes double (1 : 2 : 4 : ...): the kernels cannot start before the first chunk has landed, so it is small, and since a
    // chunk's copy takes about half as long as its evaluation on this box (PCIe Gen5 against the FP64 pipe), the copy of chunk i+1 (twice
    // the size) ends when the evaluation of chunk i does -- both engines stay busy. As many chunks (at most 5) as leave the first one
    // ~1.75 waves of objective blocks and 256 KB: measured on stick240 (4096 candidates, 25.6 MB), e2e per step: ramp 1:2:2:2 1.066 ms,
    // double x4 1.073, double x5 1.027, double x6 1.075 (its first chunk is less than a wave); free parameters 1.091 -> 1.035.
    // OMO_EVAL_CHUNKS / OMO_EVAL_CHUNKS_SHAPE=equal | ramp | double override.
    const long long work_items = (long long)B * std::max(ctx->n_tiles, 1);
    int n_chunks = 1;
    while (n_chunks < 5 && work_items / ((1LL << (n_chunks + 1)) - 1) >= 1300 && bytes / (size_t)((1LL << (n_chunks + 1)) - 1) >= ((size_t)256 << 10)) ++n_chunks;
    if (const char* e = std::getenv("OMO_EVAL_CHUNKS")) n_chunks = std::max(1, std::min(8, std::min(B, std::atoi(e))));
    // every chunk takes the kernel the whole batch would (scene-specialised from 64 K sample-evals up): a small first chunk must not switch
    // to the table-driven kernel, whose results differ in the last ulps
    struct HintScope { omo_ctx* c; int saved; HintScope(omo_ctx* c_, int B_) : c(c_), saved(c_->batch_hint) { c->batch_hint = std::max(saved, B_); } ~HintScope() { c->batch_hint = saved; } } hint_scope(ctx, B);
    auto evaluate = [&](int b0, int nb) -> omo_status {
        if (expand) {
            const long long n = (long long)nb * ctx->D;
            omo::expand_kernel<<<(unsigned)((n + 255) / 256), 256, 0, ctx->stream>>>(ctx->d_xbase, ctx->d_inv_free, ctx->D, F, d_stage + (size_t)b0 * W,
                                                                                   ctx->d_X + (size_t)b0 * ctx->D, n);
            OMO_CUDA(cudaGetLastError());
            ctx->launches += 1;
        }
        return eval_device_impl(ctx, ctx->d_X + (size_t)b0 * ctx->D, nb, d_f + b0, d_terms ? d_terms + (size_t)3 * b0 : nullptr);
    };
    if (n_chunks == 1) {
        OMO_CUDA(cudaMemcpyAsync(d_stage, X, bytes, cudaMemcpyHostToDevice, ctx->stream));
        if ((st = evaluate(0, B)) != OMO_OK) return st;
    } else {
        if (!ctx->copy_stream) {
            OMO_CUDA(cudaStreamCreateWithFlags(&ctx->copy_stream, cudaStreamNonBlocking));
            for (auto& e : ctx->chunk_ev) OMO_CUDA(cudaEventCreateWithFlags(&e, cudaEventDisableTiming));
            OMO_CUDA(cudaEventCreateWithFlags(&ctx->done_ev, cudaEventDisableTiming));
        }
        // d_X may still be read by kernels queued earlier on `stream`
        OMO_CUDA(cudaEventRecord(ctx->done_ev, ctx->stream));
        OMO_CUDA(cudaStreamWaitEvent(ctx->copy_stream, ctx->done_ev, 0));
        // OMO_EVAL_CHUNKS_SHAPE=equal | ramp (1 : 2 : 2 : ...) | double (default, 1 : 2 : 4 : ...)
        int bound[9];
        const char* shape = std::getenv("OMO_EVAL_CHUNKS_SHAPE");
        const int mode_c = !shape ? 2 : (std::strcmp(shape, "equal") == 0 ? 0 : (std::strcmp(shape, "ramp") == 0 ? 1 : 2));
        for (int c = 0; c <= n_chunks; ++c) {
            long long num = c, den = n_chunks;
            if (mode_c == 1) { num = c == 0 ? 0 : 2 * c - 1; den = 2 * n_chunks - 1; }
            if (mode_c == 2) { num = (1LL << c) - 1; den = (1LL << n_chunks) - 1; }
            bound[c] = (int)((long long)B * num / den);
        }
        for (int c = 0; c < n_chunks; ++c) {
            const int b0 = bound[c], nb = bound[c + 1] - b0;
            if (nb <= 0) continue;
            OMO_CUDA(cudaMemcpyAsync(d_stage + (size_t)b0 * W, X + (size_t)b0 * W, (size_t)nb * W * sizeof(double), cudaMemcpyHostToDevice, ctx->copy_stream));
            OMO_CUDA(cudaEventRecord(ctx->chunk_ev[c], ctx->copy_stream));
        }
        // chunks alternate between two compute streams (their scratch rows are disjoint); everything joins on `stream` again
        const bool two = std::getenv("OMO_EVAL_ONE_STREAM") == nullptr;
        if (two && !ctx->stream2) {
            OMO_CUDA(cudaStreamCreateWithFlags(&ctx->stream2, cudaStreamNonBlocking));
            OMO_CUDA(cudaEventCreateWithFlags(&ctx->join_ev, cudaEventDisableTiming));
            OMO_CUDA(cudaEventCreateWithFlags(&ctx->fork_ev, cudaEventDisableTiming));
        }
        cudaStream_t const main_stream = ctx->stream;
        if (two) {   // stream2 starts after whatever `stream` has queued so far (uploads of tables, earlier calls)
            OMO_CUDA(cudaEventRecord(ctx->fork_ev, main_stream));
            OMO_CUDA(cudaStreamWaitEvent(ctx->stream2, ctx->fork_ev, 0));
        }
        bool used2 = false;
        for (int c = 0; c < n_chunks; ++c) {
            const int b0 = bound[c], nb = bound[c + 1] - b0;
            if (nb <= 0) continue;
            const bool odd = two && (c & 1);
            if (odd) { ctx->stream = ctx->stream2; used2 = true; }
            ctx->row0 = b0;     // this chunk's own scratch rows
            cudaError_t we = cudaStreamWaitEvent(ctx->stream, ctx->chunk_ev[c], 0);
            st = we == cudaSuccess ? evaluate(b0, nb) : OMO_CUDA_ERROR;
            ctx->stream = main_stream;
            ctx->row0 = 0;
            if (we != cudaSuccess) return fail(ctx, OMO_CUDA_ERROR, cudaGetErrorString(we));
            if (st != OMO_OK) return st;
        }
        if (used2) {
            OMO_CUDA(cudaEventRecord(ctx->join_ev, ctx->stream2));
            OMO_CUDA(cudaStreamWaitEvent(main_stream, ctx->join_ev, 0));
        }
    }
    if (d_f_dst) return OMO_OK;
    OMO_CUDA(cudaMemcpyAsync(f_out, ctx->d_f, (size_t)B * sizeof(double), cudaMemcpyDeviceToHost, ctx->stream));
    if (terms_out) OMO_CUDA(cudaMemcpyAsync(terms_out, ctx->d_terms, (size_t)B * 3 * sizeof(double), cudaMemcpyDeviceToHost, ctx->stream));
    OMO_CUDA(cudaStreamSynchronize(ctx->stream));
    return OMO_OK;
}

}  // namespace

// =====================================================================================================
extern "C" {

const char* omo_version(void) { return "optimo-b200 0.1 (sm_100a)"; }

const char* omo_last_error(const omo_ctx* ctx) { return ctx ? ctx->err.c_str() : g_create_error.c_str(); }

omo_status omo_create(const omo_scene_desc* scene, int device, omo_ctx** out)
{
    if (!scene || !out) return fail(nullptr, OMO_BAD_ARG, "null argument");
    *out = nullptr;
    int n_dev = 0;
    cudaError_t e = cudaGetDeviceCount(&n_dev);
    if (e != cudaSuccess || n_dev == 0) return fail(nullptr, OMO_CUDA_ERROR, std::string("no usable CUDA device (there is no CPU fallback): ") + cudaGetErrorString(e));
    if (device < 0 || device >= n_dev) return fail(nullptr, OMO_BAD_ARG, "device index out of range");
    if ((e = cudaSetDevice(device)) != cudaSuccess) return fail(nullptr, OMO_CUDA_ERROR, cudaGetErrorString(e));
    omo_ctx* ctx = new omo_ctx();
    ctx->device = device;
    if ((e = cudaStreamCreateWithFlags(&ctx->stream, cudaStreamNonBlocking)) != cudaSuccess) {
        g_create_error = cudaGetErrorString(e); delete ctx; return OMO_CUDA_ERROR;
    }
    cudaDeviceGetAttribute(&ctx->sm_count, cudaDevAttrMultiProcessorCount, device);
    omo_status st = build_scene(ctx, scene);
    if (st == OMO_OK) {
        cudaError_t e2 = cudaMalloc((void**)&ctx->d_counters, 4 * sizeof(unsigned long long));
        if (e2 != cudaSuccess) { ctx->err = cudaGetErrorString(e2); st = OMO_CUDA_ERROR; }
        else cudaMemsetAsync(ctx->d_counters, 0, 4 * sizeof(unsigned long long), ctx->stream);
    }
    if (st != OMO_OK) { g_create_error = ctx->err; omo_destroy(ctx); return st; }
    *out = ctx;
    return OMO_OK;
}

void omo_destroy(omo_ctx* ctx)
{
    if (!ctx) return;
    cudaSetDevice(ctx->device);
    if (ctx->stream) cudaStreamSynchronize(ctx->stream);
    {
        std::lock_guard<std::mutex> g(g_scene_mutex);
        if (ctx->device >= 0 && ctx->device < 64 && g_active_scene[ctx->device] == ctx) g_active_scene[ctx->device] = nullptr;
    }
    void* ptrs[] = {ctx->d_cp_of_pos, ctx->d_first_of_pos, ctx->d_last_of_pos, ctx->d_next_same, ctx->d_tgrid, ctx->d_wparam, ctx->d_wcost,
                    ctx->d_x0, ctx->d_p_orig, ctx->d_X, ctx->d_prep, ctx->d_dparam, ctx->d_partial, ctx->d_f, ctx->d_terms, ctx->d_flags,
                    ctx->d_oq, ctx->d_oqd, ctx->d_oqdd, ctx->d_otau, ctx->d_olocal, ctx->d_op, ctx->d_xbase, ctx->d_free, ctx->d_counters, ctx->d_seg, ctx->d_tile_seg, ctx->d_items, ctx->d_lohi};
    for (void* p : ptrs) if (p) cudaFree(p);
    if (ctx->sg_exec) cudaGraphExecDestroy(ctx->sg_exec);
    if (ctx->h_single) cudaFreeHost(ctx->h_single);
    if (ctx->d_Xfree) cudaFree(ctx->d_Xfree);
    if (ctx->d_inv_free) cudaFree(ctx->d_inv_free);
    if (ctx->d_single) cudaFree(ctx->d_single);
    { void* ap[] = {ctx->d_sub_end, ctx->d_pos_of_cp, ctx->d_ci_of_pos, ctx->d_adj_m, ctx->d_adj_L, ctx->d_adj_G, ctx->d_adj_contrib, ctx->d_adj_gfull,
                    ctx->d_adj_gout, ctx->d_oroot, ctx->d_adj_ws};
      for (void* q : ap) if (q) cudaFree(q); }
    if (ctx->d_perm) cudaFree(ctx->d_perm);
    if (ctx->d_self_dparam) cudaFree(ctx->d_self_dparam);
    if (ctx->d_self_counter) cudaFree(ctx->d_self_counter);
    if (ctx->h_self_out) cudaFreeHost(ctx->h_self_out);
    for (auto& k : ctx->jit) if (k.mod) omo_jit::api().cuModuleUnload(k.mod);
    for (auto& pr : ctx->timed) { cudaEventDestroy(pr.first); cudaEventDestroy(pr.second); }
    for (auto& e : ctx->chunk_ev) if (e) cudaEventDestroy(e);
    if (ctx->done_ev) cudaEventDestroy(ctx->done_ev);
    if (ctx->copy_stream) cudaStreamDestroy(ctx->copy_stream);
    if (ctx->join_ev) cudaEventDestroy(ctx->join_ev);
    if (ctx->fork_ev) cudaEventDestroy(ctx->fork_ev);
    if (ctx->stream2) cudaStreamDestroy(ctx->stream2);
    if (ctx->stream) cudaStreamDestroy(ctx->stream);
    delete ctx;
}

omo_status omo_set_window(omo_ctx* ctx, double min_t, double max_t, double dt)
{
    if (!ctx) return OMO_BAD_ARG;
    if (!(dt > 0.0) || !(max_t >= min_t)) return fail(ctx, OMO_BAD_ARG, "need dt > 0 and max_t >= min_t");
    OMO_CUDA(cudaSetDevice(ctx->device));
    if (ctx->have_window && ctx->min_t == min_t && ctx->max_t == max_t && ctx->dt == dt) return OMO_OK;
    std::vector<double> grid;
    for (double t = min_t; t <= max_t; t += dt) {  // core_rbdl.cpp:120-123 -- accumulated, never closed-form
        grid.push_back(t);
        if (grid.size() > (size_t)(1 << 24)) return fail(ctx, OMO_BAD_ARG, "time grid too long");
    }
    if (grid.size() < 3) return fail(ctx, OMO_BAD_ARG, "the window must hold at least 3 samples");
    omo_status st = install_grid(ctx, grid, dt);
    if (st != OMO_OK) return st;
    ctx->min_t = min_t; ctx->max_t = max_t;
    ctx->have_window = true;
    return OMO_OK;
}

omo_status omo_set_weights(omo_ctx* ctx, const double* w_param, int D, double w_traj, const omo_gauss* kernels, int n_kernels, int cyclic)
{
    if (!ctx) return OMO_BAD_ARG;
    if (D != ctx->D || (D > 0 && !w_param) || n_kernels < 0 || (n_kernels > 0 && !kernels)) return fail(ctx, OMO_BAD_ARG, "bad weights");
    OMO_CUDA(cudaSetDevice(ctx->device));
    const bool same_param = ctx->have_weights && std::equal(w_param, w_param + D, ctx->wparam.begin());
    if (!same_param) {
        OMO_CUDA(cudaStreamSynchronize(ctx->stream));
        ctx->wparam.assign(w_param, w_param + D);
        omo_status st = upload(ctx, &ctx->d_wparam, ctx->wparam);
        if (st != OMO_OK) return st;
    }
    bool same_k = ctx->have_weights && (int)ctx->kernels.size() == n_kernels;
    for (int i = 0; same_k && i < n_kernels; ++i)
        same_k = kernels[i].amp == ctx->kernels[i].amp && kernels[i].mu == ctx->kernels[i].mu && kernels[i].sigma == ctx->kernels[i].sigma;
    if (!same_k) { ctx->kernels.assign(kernels, kernels + n_kernels); ctx->wcost_dirty = true; }
    ctx->w_traj = w_traj;
    if (ctx->cyclic != (cyclic ? 1 : 0)) ctx->sp_valid = false;   // cyclic curves average their end tangents: different reach
    ctx->cyclic = cyclic ? 1 : 0;
    ctx->have_weights = true;
    return OMO_OK;
}

omo_status omo_set_original(omo_ctx* ctx, const double* x_original, int D)
{
    if (!ctx) return OMO_BAD_ARG;
    if (D != ctx->D || (D > 0 && !x_original)) return fail(ctx, OMO_BAD_ARG, "bad x_original");
    OMO_CUDA(cudaSetDevice(ctx->device));
    if (ctx->have_original && std::equal(x_original, x_original + D, ctx->x0.begin())) return OMO_OK;
    OMO_CUDA(cudaStreamSynchronize(ctx->stream));
    ctx->x0.assign(x_original, x_original + D);
    omo_status st = upload(ctx, &ctx->d_x0, ctx->x0);
    if (st != OMO_OK) return st;
    ctx->have_original = true;
    ctx->p_orig_dirty = true;
    return OMO_OK;
}

omo_status omo_set_batch_hint(omo_ctx* ctx, int total_candidates)
{
    if (!ctx || total_candidates < 0) return OMO_BAD_ARG;
    ctx->batch_hint = total_candidates;
    return OMO_OK;
}

int omo_num_samples(const omo_ctx* ctx) { return ctx ? ctx->N_s : 0; }
int omo_descriptor_size(const omo_ctx* ctx) { return ctx ? ctx->D : 0; }
int omo_dof(const omo_ctx* ctx) { return ctx ? 3 * ctx->J : 0; }
int omo_num_walked_bodies(const omo_ctx* ctx, int* general_out)
{
    if (!ctx) return 0;
    int gen = 0;
    for (int i = 0; i < ctx->hs.n_act[1]; ++i) gen += ctx->hs.ident[ctx->hs.act[1][i]] ? 0 : 1;
    if (general_out) *general_out = gen;
    return ctx->hs.n_act[1];
}

omo_status omo_get_time_grid(const omo_ctx* ctx, double* t_out)
{
    if (!ctx || !t_out || !ctx->have_window) return OMO_BAD_ARG;
    std::copy(ctx->tgrid.begin(), ctx->tgrid.end(), t_out);
    return OMO_OK;
}

omo_status omo_get_cost_weights(const omo_ctx* cctx, double* w_out)
{
    omo_ctx* ctx = const_cast<omo_ctx*>(cctx);
    DeviceLock device_lock(ctx);
    if (!ctx || !w_out) return OMO_BAD_ARG;
    omo_status st = ensure_ready(ctx);
    if (st != OMO_OK) return st;
    std::copy(ctx->wcost.begin(), ctx->wcost.end(), w_out);
    return OMO_OK;
}

omo_status omo_get_p_original(omo_ctx* ctx, double* p_out)
{
    DeviceLock device_lock(ctx);
    if (!ctx || !p_out) return OMO_BAD_ARG;
    omo_status st = ensure_ready(ctx);
    if (st != OMO_OK) return st;
    const int rows = 3 * ctx->J, cols = ctx->N_s;
    std::vector<double> tmp((size_t)rows * cols);
    OMO_CUDA(cudaMemcpyAsync(tmp.data(), ctx->d_p_orig, tmp.size() * sizeof(double), cudaMemcpyDeviceToHost, ctx->stream));
    OMO_CUDA(cudaStreamSynchronize(ctx->stream));
    for (int r = 0; r < rows; ++r) for (int c = 0; c < cols; ++c) p_out[(size_t)c * rows + r] = tmp[(size_t)r * cols + c];
    return OMO_OK;
}

omo_status omo_eval_device(omo_ctx* ctx, const double* dX, int B, double* d_f_out, double* d_terms_out, int sync)
{
    DeviceLock device_lock(ctx);
    if (!ctx || !dX || !d_f_out || B < 0) return fail(ctx, OMO_BAD_ARG, "bad argument");
    if (B == 0) return OMO_OK;
    omo_status st = ensure_ready(ctx);
    if (st != OMO_OK) return st;
    if ((st = eval_device_impl(ctx, dX, B, d_f_out, d_terms_out)) != OMO_OK) return st;
    if (sync) OMO_CUDA(cudaStreamSynchronize(ctx->stream));
    return OMO_OK;
}

omo_status omo_eval(omo_ctx* ctx, const double* X, int B, double* f_out, omo_terms* terms_out)
{
    DeviceLock device_lock(ctx);
    if (!ctx || (!X && ctx->D > 0) || !f_out || B < 0) return fail(ctx, OMO_BAD_ARG, "bad argument");   // D = 0: a scene without keyed curves
    if (B == 0) return OMO_OK;
    omo_status st = ensure_ready(ctx);
    if (st != OMO_OK) return st;
    if ((st = ensure_capacity(ctx, B)) != OMO_OK) return st;
    const size_t bytes = (size_t)B * ctx->D * sizeof(double);
    if (ctx->graph_mode < 0) { const char* e = std::getenv("OMO_GRAPH"); ctx->graph_mode = (e && std::atoi(e) == 0) ? 0 : 1; }
    // (a context that evaluates ONE row of a larger batch -- a device of omo_multi -- stays on the batch kernels: same bits as the whole batch)
    const bool lone = B == 1 && ctx->batch_hint <= 1;
    if (lone && self_usable(ctx)) {
        // one launch: x travels as a kernel parameter, the kernel prepares, evaluates, reduces and writes (f, C, D_param, D_traj)
        // into mapped host memory
        if ((st = ensure_self_buffers(ctx, 1)) != OMO_OK) return st;
        bool launched = false;
        if ((st = launch_self(ctx, X, 1, -1, nullptr, 0, nullptr, nullptr, nullptr, 0.0, 0, ctx->d_self_out, &launched)) != OMO_OK) return st;
        if (launched) {
            if ((st = wait_self(ctx)) != OMO_OK) return st;
            *f_out = ctx->h_self_out[0];
            if (terms_out) { terms_out->C = ctx->h_self_out[1]; terms_out->D_param = ctx->h_self_out[2]; terms_out->D_traj = ctx->h_self_out[3]; }
            return OMO_OK;
        }
    }
    if (lone && ctx->graph_mode == 1 && !ctx->timing) {
        // The solver's callback: one candidate per call, thousands of calls. The first call with a given state runs the plain
        // path below (it finishes every lazy allocation / module load), the second captures, the rest replay.
        if (!ctx->h_single) OMO_CUDA(cudaHostAlloc((void**)&ctx->h_single, (size_t)(ctx->D + 4) * sizeof(double), cudaHostAllocDefault));
        if (!ctx->d_single) OMO_CUDA(cudaMalloc((void**)&ctx->d_single, 4 * sizeof(double)));
        jit_mode(ctx);
        const std::vector<unsigned long long> key = single_key(ctx);
        bool replay = ctx->sg_exec != nullptr && key == ctx->sg_key;
        if (!replay && key == ctx->sg_pending) {
            if (ctx->sg_exec) { cudaGraphExecDestroy(ctx->sg_exec); ctx->sg_exec = nullptr; }
            const unsigned long long l0 = ctx->launches;
            cudaGraph_t graph = nullptr;
            OMO_CUDA(cudaStreamBeginCapture(ctx->stream, cudaStreamCaptureModeRelaxed));
            cudaError_t ce = cudaMemcpyAsync(ctx->d_X, ctx->h_single, bytes, cudaMemcpyHostToDevice, ctx->stream);
            st = ce == cudaSuccess ? eval_device_impl(ctx, ctx->d_X, 1, ctx->d_single, ctx->d_single + 1) : OMO_CUDA_ERROR;
            if (st == OMO_OK && ce == cudaSuccess) ce = cudaMemcpyAsync(ctx->h_single + ctx->D, ctx->d_single, 4 * sizeof(double), cudaMemcpyDeviceToHost, ctx->stream);
            const cudaError_t ee = cudaStreamEndCapture(ctx->stream, &graph);   // always ends the capture, whatever happened inside
            ctx->sg_kernels = ctx->launches - l0;
            ctx->launches = l0;
            if (st == OMO_OK && ce == cudaSuccess && ee == cudaSuccess && graph != nullptr &&
                cudaGraphInstantiate(&ctx->sg_exec, graph, 0) == cudaSuccess) {
                ctx->sg_key = key;
                replay = true;
            } else {
                cudaGetLastError();
                ctx->sg_exec = nullptr;
                ctx->graph_mode = 0;     // this driver / state cannot capture the chain: plain launches from now on
            }
            if (graph) cudaGraphDestroy(graph);
        }
        if (replay) {
            std::memcpy(ctx->h_single, X, bytes);
            OMO_CUDA(cudaGraphLaunch(ctx->sg_exec, ctx->stream));
            OMO_CUDA(cudaStreamSynchronize(ctx->stream));
            ctx->launches += ctx->sg_kernels;
            ctx->sg_replays += 1;
            *f_out = ctx->h_single[ctx->D];
            if (terms_out) std::memcpy(terms_out, ctx->h_single + ctx->D + 1, 3 * sizeof(double));
            return OMO_OK;
        }
        ctx->sg_pending = key;
    }
    return eval_host_pipelined(ctx, X, B, ctx->D, ctx->d_X, false, 0, f_out, terms_out);
}

omo_status omo_eval_host_in(omo_ctx* ctx, const double* X, int B, double* d_f_out, double* d_terms_out, int sync)
{
    DeviceLock device_lock(ctx);
    if (!ctx || (!X && ctx->D > 0) || !d_f_out || B < 0) return fail(ctx, OMO_BAD_ARG, "bad argument");
    if (B == 0) return OMO_OK;
    omo_status st = ensure_ready(ctx);
    if (st != OMO_OK) return st;
    if ((st = ensure_capacity(ctx, B)) != OMO_OK) return st;
    if ((st = eval_host_pipelined(ctx, X, B, ctx->D, ctx->d_X, false, 0, nullptr, nullptr, d_f_out, d_terms_out)) != OMO_OK) return st;
    if (sync) OMO_CUDA(cudaStreamSynchronize(ctx->stream));
    return OMO_OK;
}

omo_status omo_eval_free(omo_ctx* ctx, const double* x_base, const int* free_idx, int F, const double* Xfree, int B, double* f_out, omo_terms* terms_out)
{
    DeviceLock device_lock(ctx);
    if (!ctx || (!x_base && ctx->D > 0) || F < 0 || (F > 0 && (!free_idx || !Xfree)) || !f_out || B < 0) return fail(ctx, OMO_BAD_ARG, "bad argument");
    for (int k = 0; k < F; ++k) if (free_idx[k] < 0 || free_idx[k] >= ctx->D) return fail(ctx, OMO_BAD_ARG, "free index outside the descriptor");
    if (B == 0) return OMO_OK;
    omo_status st = ensure_ready(ctx);
    if (st != OMO_OK) return st;
    if ((st = ensure_capacity(ctx, B)) != OMO_OK) return st;
    if (!ctx->d_xbase) { if ((st = dev_alloc(ctx, &ctx->d_xbase, (size_t)ctx->D)) != OMO_OK) return st; }
    if ((size_t)B * F > ctx->xfree_cap) { if ((st = dev_alloc(ctx, &ctx->d_Xfree, (size_t)B * F)) != OMO_OK) return st; ctx->xfree_cap = (size_t)B * F; }
    // x_base travels only when it changed (a solve expands every batch over the same x_initial, core_opt.cpp:99-121): the upload from
    // pageable memory and the synchronisation behind it cost ~20 us of a 1.1 ms call
    bool synced = true;
    if (ctx->xbase_uploaded.size() != (size_t)ctx->D || ctx->xbase_is_free_base == false ||
        std::memcmp(ctx->xbase_uploaded.data(), x_base, (size_t)ctx->D * sizeof(double)) != 0) {
        OMO_CUDA(cudaMemcpyAsync(ctx->d_xbase, x_base, (size_t)ctx->D * sizeof(double), cudaMemcpyHostToDevice, ctx->stream));
        ctx->xbase_uploaded.assign(x_base, x_base + ctx->D);
        ctx->xbase_is_free_base = true;
        synced = false;
    }
    // descriptor index -> position in the free list: the expansion is then one gather per element, parallel over the whole batch
    if (!ctx->d_inv_free || ctx->inv_free_of.size() != (size_t)F || !std::equal(ctx->inv_free_of.begin(), ctx->inv_free_of.end(), free_idx)) {
        synced = false;
        std::vector<int> inv((size_t)std::max(ctx->D, 1), -1);
        for (int k = 0; k < F; ++k) inv[free_idx[k]] = k;          // a repeated index: the last entry wins, as a serial scatter would
        OMO_CUDA(cudaStreamSynchronize(ctx->stream));
        if ((st = upload(ctx, &ctx->d_inv_free, inv)) != OMO_OK) return st;
        ctx->inv_free_of.assign(free_idx, free_idx + F);
    }
    if (!synced) OMO_CUDA(cudaStreamSynchronize(ctx->stream));   // x_base and the map came from pageable memory
    return eval_host_pipelined(ctx, Xfree, B, F, ctx->d_Xfree, true, F, f_out, terms_out);
}

omo_status omo_eval_cached(omo_ctx* ctx, const double* x, double* f_out, omo_terms* terms_out, double* q, double* qd, double* qdd,
                           double* tau, double* local_cost)
{
    DeviceLock device_lock(ctx);
    if (!ctx || (!x && ctx->D > 0)) return fail(ctx, OMO_BAD_ARG, "bad argument");
    omo_status st = ensure_ready(ctx);
    if (st != OMO_OK) return st;
    if ((st = ensure_capacity(ctx, 1)) != OMO_OK) return st;
    OMO_CUDA(cudaMemcpyAsync(ctx->d_X, x, (size_t)ctx->D * sizeof(double), cudaMemcpyHostToDevice, ctx->stream));
    if ((st = run_single_out(ctx, ctx->d_X, ctx->w_traj != 0.0)) != OMO_OK) return st;
    const size_t n = (size_t)ctx->N_s * 3 * ctx->J * sizeof(double);
    if (q) OMO_CUDA(cudaMemcpyAsync(q, ctx->d_oq, n, cudaMemcpyDeviceToHost, ctx->stream));
    if (qd) OMO_CUDA(cudaMemcpyAsync(qd, ctx->d_oqd, n, cudaMemcpyDeviceToHost, ctx->stream));
    if (qdd) OMO_CUDA(cudaMemcpyAsync(qdd, ctx->d_oqdd, n, cudaMemcpyDeviceToHost, ctx->stream));
    if (tau) OMO_CUDA(cudaMemcpyAsync(tau, ctx->d_otau, n, cudaMemcpyDeviceToHost, ctx->stream));
    if (local_cost) OMO_CUDA(cudaMemcpyAsync(local_cost, ctx->d_olocal, (size_t)ctx->N_s * sizeof(double), cudaMemcpyDeviceToHost, ctx->stream));
    if (f_out) OMO_CUDA(cudaMemcpyAsync(f_out, ctx->d_f, sizeof(double), cudaMemcpyDeviceToHost, ctx->stream));
    if (terms_out) OMO_CUDA(cudaMemcpyAsync(terms_out, ctx->d_terms, 3 * sizeof(double), cudaMemcpyDeviceToHost, ctx->stream));
    OMO_CUDA(cudaStreamSynchronize(ctx->stream));
    return OMO_OK;
}

omo_status omo_positions(omo_ctx* ctx, const double* x, double* p_out)
{
    DeviceLock device_lock(ctx);
    if (!ctx || (!x && ctx->D > 0) || !p_out) return fail(ctx, OMO_BAD_ARG, "bad argument");
    omo_status st = ensure_ready(ctx);
    if (st != OMO_OK) return st;
    if ((st = ensure_capacity(ctx, 1)) != OMO_OK) return st;
    OMO_CUDA(cudaMemcpyAsync(ctx->d_X, x, (size_t)ctx->D * sizeof(double), cudaMemcpyHostToDevice, ctx->stream));
    if ((st = run_single_out(ctx, ctx->d_X, false)) != OMO_OK) return st;
    OMO_CUDA(cudaMemcpyAsync(p_out, ctx->d_op, (size_t)ctx->N_s * 3 * ctx->J * sizeof(double), cudaMemcpyDeviceToHost, ctx->stream));
    OMO_CUDA(cudaStreamSynchronize(ctx->stream));
    return OMO_OK;
}

omo_status omo_eval_states_at(omo_ctx* ctx, const double* x, const double* times, int M, double dt, double* q, double* qd,
                              double* qdd, double* tau, double* root_euler, double* p)
{
    DeviceLock device_lock(ctx);
    if (!ctx || (!x && ctx->D > 0) || M < 0 || (M > 0 && !times) || !(dt > 0.0)) return fail(ctx, OMO_BAD_ARG, "bad argument");
    if (!ctx->have_weights || !ctx->have_original) return fail(ctx, OMO_NOT_READY, "omo_set_weights / omo_set_original have not been called");
    if (M == 0) return OMO_OK;
    for (int i = 0; i < M; ++i) if (!std::isfinite(times[i])) return fail(ctx, OMO_BAD_ARG, "non-finite time");
    OMO_CUDA(cudaSetDevice(ctx->device));
    // passes of time-ordered, non-overlapping triples (t_k + dt <= t_{k+1} - dt): the frames of a motion are one pass
    std::vector<int> order(M);
    for (int i = 0; i < M; ++i) order[i] = i;
    std::stable_sort(order.begin(), order.end(), [&](int a, int b) { return times[a] < times[b]; });
    const size_t dof = (size_t)3 * ctx->J;
    constexpr int kMaxTriples = 1 << 18;
    std::vector<char> done(M, 0);
    std::vector<int> take;
    std::vector<double> grid, h[6];
    int left = M;
    omo_status st = OMO_OK;
    double* d_w = nullptr; double* d_root = nullptr; double* d_grid = nullptr;
    // The triples' grid is swapped in for the launches below and the objective window is put back untouched afterwards: its
    // cost weights, p_original, sparse-gradient item lists, work buffers and the captured single-candidate graph all stay
    // valid (a torque query from the UI thread in the middle of a solve costs its own launches and nothing else).
    OMO_CUDA(cudaStreamSynchronize(ctx->stream));
    std::vector<double> saved_grid;
    saved_grid.swap(ctx->tgrid);
    double* const saved_d_tgrid = ctx->d_tgrid;
    const int saved_N = ctx->N_s, saved_tiles = ctx->n_tiles, saved_T = ctx->T_out;
    const double saved_dt = ctx->dt;
    const bool swap_residues = ctx->residues_valid && ctx->residues_dt != dt;
    if (!ctx->residues_valid || swap_residues) { ctx->dt = dt; refresh_residues(ctx); ctx->residues_valid = true; ctx->residues_dt = dt; }
    while (left > 0 && st == OMO_OK) {
        take.clear(); grid.clear();
        double last_end = -INFINITY;
        for (int k = 0; k < M && (int)take.size() < kMaxTriples; ++k) {
            const int i = order[k];
            if (done[i]) continue;
            const double t = times[i];
            if (!(t - dt >= last_end)) continue;
            grid.push_back(t - dt); grid.push_back(t); grid.push_back(t + dt);   // core_rbdl.cpp:186-188
            last_end = t + dt;
            take.push_back(i); done[i] = 1;
        }
        left -= (int)take.size();
        const int n = (int)take.size(), N = 3 * n;
        ctx->tgrid = grid;
        ctx->N_s = N; ctx->dt = dt;
        ctx->n_tiles = (N + kBlockThreads - 5) / (kBlockThreads - 4);
        ctx->T_out = (N + ctx->n_tiles - 1) / ctx->n_tiles;
        if ((st = dev_alloc(ctx, &d_grid, (size_t)N)) != OMO_OK) break;
        ctx->d_tgrid = d_grid;
        if ((st = ensure_capacity(ctx, 1)) != OMO_OK) break;
        if ((st = ensure_outputs(ctx)) != OMO_OK) break;
        if ((st = dev_alloc(ctx, &d_w, (size_t)N)) != OMO_OK) break;
        if ((st = dev_alloc(ctx, &d_root, (size_t)N * 3)) != OMO_OK) break;
        cudaError_t ce = cudaMemcpyAsync(d_grid, grid.data(), (size_t)N * sizeof(double), cudaMemcpyHostToDevice, ctx->stream);
        if (ce == cudaSuccess) ce = cudaMemsetAsync(d_w, 0, (size_t)N * sizeof(double), ctx->stream);
        if (ce == cudaSuccess) ce = cudaMemcpyAsync(ctx->d_X, x, (size_t)ctx->D * sizeof(double), cudaMemcpyHostToDevice, ctx->stream);
        if (ce != cudaSuccess) { st = fail(ctx, OMO_CUDA_ERROR, cudaGetErrorString(ce)); break; }
        if ((st = activate_scene(ctx)) != OMO_OK) break;
        if ((st = run_prepare(ctx, ctx->d_X, 1)) != OMO_OK) break;
        OmoEvalParams P;
        fill_params(ctx, P);
        P.wcost = d_w; P.p_orig = nullptr; P.triplet = 1; P.o_root = d_root;
        P.o_q = ctx->d_oq; P.o_qd = ctx->d_oqd; P.o_qdd = ctx->d_oqdd; P.o_tau = ctx->d_otau; P.o_local = ctx->d_olocal; P.o_p = ctx->d_op;
        if ((st = launch_objective(ctx, P, 1, true, true, ctx->d_f, ctx->d_terms)) != OMO_OK) break;
        // only the middle sample of each triple is read back (strided copies, row = one sample)
        double* outs[6] = {q, qd, qdd, tau, p, root_euler};
        const double* srcs[6] = {ctx->d_oq, ctx->d_oqd, ctx->d_oqdd, ctx->d_otau, ctx->d_op, d_root};
        for (int a = 0; a < 6 && ce == cudaSuccess; ++a) {
            if (!outs[a]) continue;
            const size_t w = (a == 5 ? 3 : dof) * sizeof(double);
            h[a].resize((size_t)n * (a == 5 ? 3 : dof));
            ce = cudaMemcpy2DAsync(h[a].data(), w, srcs[a] + (a == 5 ? 3 : dof), 3 * w, w, (size_t)n, cudaMemcpyDeviceToHost, ctx->stream);
        }
        if (ce == cudaSuccess) ce = cudaStreamSynchronize(ctx->stream);
        if (ce != cudaSuccess) { st = fail(ctx, OMO_CUDA_ERROR, cudaGetErrorString(ce)); break; }
        for (int a = 0; a < 6; ++a) {
            if (!outs[a]) continue;
            const size_t w = (a == 5 ? 3 : dof);
            for (int k = 0; k < n; ++k) std::memcpy(outs[a] + (size_t)take[k] * w, h[a].data() + (size_t)k * w, w * sizeof(double));
        }
    }
    cudaStreamSynchronize(ctx->stream);
    if (d_w) cudaFree(d_w);
    if (d_root) cudaFree(d_root);
    if (d_grid) cudaFree(d_grid);
    // put the objective window back exactly as it was
    ctx->tgrid.swap(saved_grid);
    ctx->d_tgrid = saved_d_tgrid; ctx->N_s = saved_N; ctx->n_tiles = saved_tiles; ctx->T_out = saved_T; ctx->dt = saved_dt;
    if (swap_residues) { refresh_residues(ctx); ctx->residues_dt = saved_dt; }
    return st;
}

omo_status omo_eval_grad_columns(omo_ctx* ctx, const double* x, const int* free_idx, int F, double h, int col_begin, int col_count,
                                 double* d_f_cols, int sync)
{
    DeviceLock device_lock(ctx);
    if (!ctx || (!x && ctx->D > 0) || (F > 0 && !free_idx) || F < 0 || !d_f_cols) return fail(ctx, OMO_BAD_ARG, "bad argument");
    if (col_begin < 0 || col_count < 0 || col_begin + col_count > 2 * F + 1) return fail(ctx, OMO_BAD_ARG, "column range outside [0, 2F+1)");
    for (int k = 0; k < F; ++k) if (free_idx[k] < 0 || free_idx[k] >= ctx->D) return fail(ctx, OMO_BAD_ARG, "free index outside the descriptor");
    if (col_count == 0) return OMO_OK;
    omo_status st = ensure_ready(ctx);
    if (st != OMO_OK) return st;
    if ((st = ensure_capacity(ctx, col_count)) != OMO_OK) return st;
    if (!ctx->d_xbase) { if ((st = dev_alloc(ctx, &ctx->d_xbase, (size_t)ctx->D)) != OMO_OK) return st; }
    if (F > ctx->free_cap) { if ((st = dev_alloc(ctx, &ctx->d_free, (size_t)F)) != OMO_OK) return st; ctx->free_cap = F; ctx->free_uploaded.clear(); }
    if (F > 0 && (ctx->free_uploaded.size() != (size_t)F || !std::equal(ctx->free_uploaded.begin(), ctx->free_uploaded.end(), free_idx))) {
        ctx->free_uploaded.assign(free_idx, free_idx + F);
        OMO_CUDA(cudaMemcpyAsync(ctx->d_free, ctx->free_uploaded.data(), (size_t)F * sizeof(int), cudaMemcpyHostToDevice, ctx->stream));
        OMO_CUDA(cudaStreamSynchronize(ctx->stream));
    }
    // Sparse columns: column 2k+1 / 2k+2 moves ONE control point, which only the samples inside its two adjacent segments
    // (plus the stencil's +-2) can see. Only the tiles holding those samples are evaluated; the sums take every other
    // tile from the base point, evaluated once as an extra row. Bit-identical to the dense evaluation of every column.
    const bool sparse = std::getenv("OMO_GRAD_DENSE") == nullptr && affected_tiles(ctx, x, free_idx, F, col_begin, col_count);
    const int rows = sparse ? col_count + 1 : col_count;
    if (sparse && (st = ensure_capacity(ctx, rows)) != OMO_OK) return st;
    auto upload_items = [&]() -> omo_status {
        if (ctx->h_items.size() > ctx->items_cap) {
            if ((st = dev_alloc(ctx, &ctx->d_items, ctx->h_items.size())) != OMO_OK) return st;
            ctx->items_cap = ctx->h_items.size(); ctx->sp_uploaded = false;
        }
        if ((size_t)2 * col_count > ctx->lohi_cap) {
            if ((st = dev_alloc(ctx, &ctx->d_lohi, (size_t)2 * col_count)) != OMO_OK) return st;
            ctx->lohi_cap = (size_t)2 * col_count; ctx->sp_uploaded = false;
        }
        if (!ctx->sp_uploaded) {
            OMO_CUDA(cudaMemcpyAsync(ctx->d_items, ctx->h_items.data(), ctx->h_items.size() * sizeof(int), cudaMemcpyHostToDevice, ctx->stream));
            OMO_CUDA(cudaMemcpyAsync(ctx->d_lohi, ctx->h_lo.data(), (size_t)col_count * sizeof(int), cudaMemcpyHostToDevice, ctx->stream));
            OMO_CUDA(cudaMemcpyAsync(ctx->d_lohi + col_count, ctx->h_hi.data(), (size_t)col_count * sizeof(int), cudaMemcpyHostToDevice, ctx->stream));
            OMO_CUDA(cudaStreamSynchronize(ctx->stream));   // the host vectors are pageable
            ctx->sp_uploaded = true;
        }
        return OMO_OK;
    };
    // (every block prepares its row's n_cp control points itself: worth it while that stays small beside the blocks' own work)
    if (sparse && rows <= kSelfMaxRows && (long long)ctx->h_items.size() * ctx->n_cp <= 60000 && self_usable(ctx)) {
        // small gradients: ONE launch -- x as a kernel parameter, every block builds its column's control points itself, the last
        // block adds the tile partials (same numbers in the same order as the launches below: bit-identical)
        if ((st = upload_items()) != OMO_OK) return st;
        bool launched = false;
        if ((st = launch_self(ctx, x, rows, col_count, ctx->d_items, (int)ctx->h_items.size(), ctx->d_lohi, ctx->d_lohi + col_count, ctx->d_free, h, col_begin,
                              d_f_cols, &launched)) != OMO_OK) return st;
        if (launched) {
            if (sync) return wait_self(ctx);
            ctx->self_spin_slots = 0;   // the caller synchronises the stream itself
            return OMO_OK;
        }
    }
    ctx->xbase_is_free_base = false;   // d_xbase no longer holds omo_eval_free's base
    OMO_CUDA(cudaMemcpyAsync(ctx->d_xbase, x, (size_t)ctx->D * sizeof(double), cudaMemcpyHostToDevice, ctx->stream));
    omo::perturb_kernel<<<rows, 128, 0, ctx->stream>>>(ctx->d_xbase, ctx->d_free, ctx->D, h, col_begin, col_count, ctx->d_X);
    OMO_CUDA(cudaGetLastError());
    ctx->launches += 1;
    if (!sparse) {
        ctx->prefer_jit = true;    // same kernel as the sparse columns of the other shards
        st = eval_device_impl(ctx, ctx->d_X, col_count, d_f_cols, nullptr);
        ctx->prefer_jit = false;
        if (st != OMO_OK) return st;
    } else {
        if ((st = upload_items()) != OMO_OK) return st;
        if ((st = run_prepare(ctx, ctx->d_X, rows)) != OMO_OK) return st;
        OmoEvalParams P;
        fill_params(ctx, P);
        const bool fk = (ctx->w_traj != 0.0) || ctx->n_ik > 0;
        P.p_orig = (ctx->w_traj != 0.0) ? ctx->d_p_orig : nullptr;
        if ((st = launch_objective(ctx, P, rows, fk, false, nullptr, nullptr, ctx->d_items, (int)ctx->h_items.size())) != OMO_OK) return st;
        omo::finish_sparse_kernel<<<(col_count + 127) / 128, 128, 0, ctx->stream>>>(ctx->d_partial, ctx->d_dparam, col_count, ctx->n_tiles, prep_parts(ctx),
                                                                                    ctx->d_lohi, ctx->d_lohi + col_count, col_count, d_f_cols);
        OMO_CUDA(cudaGetLastError());
        ctx->launches += 1;
    }
    if (sync) OMO_CUDA(cudaStreamSynchronize(ctx->stream));
    return OMO_OK;
}

omo_status omo_grad_from_columns(const double* f_cols, int F, double h, double* f0_out, double* grad_out)
{
    if (!f_cols || F < 0 || (F > 0 && !grad_out)) return OMO_BAD_ARG;
    if (f0_out) *f0_out = f_cols[0];
    for (int k = 0; k < F; ++k) grad_out[k] = (f_cols[2 * k + 1] - f_cols[2 * k + 2]) / (2.0 * h);
    return OMO_OK;
}

omo_status omo_eval_grad(omo_ctx* ctx, const double* x, const int* free_idx, int F, double h, double* f0_out, double* grad_out, double* fpm_out)
{
    DeviceLock device_lock(ctx);
    if (!ctx) return OMO_BAD_ARG;
    const int P = 2 * F + 1;
    omo_status st = ensure_ready(ctx);
    if (st != OMO_OK) return st;
    if ((st = ensure_capacity(ctx, P + 1)) != OMO_OK) return st;   // + the base row of the sparse evaluation: d_f must not move below
    ctx->h_cols.resize(P);
    if (P + 1 <= kSelfMaxRows && self_usable(ctx)) {
        // the column values are written straight into mapped host memory (by the self-preparing kernel, or by finish_sparse_kernel
        // if that launch is not to be had): no copy-out
        // + the base row, + room for the host-side reduction (launch_self): the buffer `out` points into must not move under the launch
        if ((st = ensure_self_buffers(ctx, P + 1, ((P + 4) & ~3) + (2 * std::max(ctx->n_tiles, 1) + 1) * (P + 1))) != OMO_OK) return st;
        if ((st = omo_eval_grad_columns(ctx, x, free_idx, F, h, 0, P, ctx->d_self_out, 1)) != OMO_OK) return st;
        std::copy(ctx->h_self_out, ctx->h_self_out + P, ctx->h_cols.begin());
    } else {
    if ((st = omo_eval_grad_columns(ctx, x, free_idx, F, h, 0, P, ctx->d_f, 0)) != OMO_OK) return st;
    OMO_CUDA(cudaMemcpyAsync(ctx->h_cols.data(), ctx->d_f, (size_t)P * sizeof(double), cudaMemcpyDeviceToHost, ctx->stream));
    OMO_CUDA(cudaStreamSynchronize(ctx->stream));
    }
    if (fpm_out) std::copy(ctx->h_cols.begin() + 1, ctx->h_cols.end(), fpm_out);
    return omo_grad_from_columns(ctx->h_cols.data(), F, h, f0_out, grad_out);
}

omo_status omo_eval_grad_analytic(omo_ctx* ctx, const double* x, const int* free_idx, int F, double* f0_out, double* grad_out)
{
    DeviceLock device_lock(ctx);
    if (!ctx || (!x && ctx->D > 0) || F < 0 || (F > 0 && (!free_idx || !grad_out))) return fail(ctx, OMO_BAD_ARG, "bad argument");
    for (int k = 0; k < F; ++k) if (free_idx[k] < 0 || free_idx[k] >= ctx->D) return fail(ctx, OMO_BAD_ARG, "free index outside the descriptor");
    if (ctx->n_ik > 0) return fail(ctx, OMO_UNSUPPORTED, "the analytic gradient does not cover IK handles (the angles of an IK chain come out of an iterative solve); use omo_eval_grad");
    omo_status st = ensure_ready(ctx);
    if (st != OMO_OK) return st;
    if ((st = ensure_capacity(ctx, 1)) != OMO_OK) return st;
    const int N = ctx->N_s, nk = ctx->hs.n_keyed, J = ctx->J;
    // tables, once per context
    if (!ctx->d_sub_end) {
        std::vector<int> sub_end(J), ci_of_pos(std::max(ctx->n_cp, 1), -1), pos_of_cp(std::max(ctx->n_cp, 1), 0);
        for (int j = J - 1; j >= 0; --j) {
            sub_end[j] = j + 1;
            for (int k = j + 1; k < J; ++k) if (ctx->hs.parent[k] == j) sub_end[j] = std::max(sub_end[j], sub_end[k]);
        }
        ctx->adj_max_sub = 1;
        for (int i = 0; i < nk; ++i) {
            const int ch = ctx->hs.grp_list[i], j = ch / 3;
            ctx->adj_max_sub = std::max(ctx->adj_max_sub, sub_end[j] - j);
            for (int p = 0; p < ctx->hs.chan_n[ch]; ++p) ci_of_pos[ctx->hs.chan_pos[ch] + p] = i;
        }
        for (int pos = 0; pos < ctx->n_cp; ++pos) pos_of_cp[ctx->cp_of_pos[pos]] = pos;
        if ((st = upload(ctx, &ctx->d_sub_end, sub_end)) != OMO_OK) return st;
        if ((st = upload(ctx, &ctx->d_ci_of_pos, ci_of_pos)) != OMO_OK) return st;
        if ((st = upload(ctx, &ctx->d_pos_of_cp, pos_of_cp)) != OMO_OK) return st;
        OMO_CUDA(cudaStreamSynchronize(ctx->stream));   // the vectors above are locals
    }
    const size_t cells = (size_t)std::max(nk, 1) * N;
    // Reverse mode (one thread per sample, O(1) passes over the tree) or forward mode (one thread per (sample, channel, kind), each a
    // restricted dual-number pass)? Forward mode's many short threads win while they do not fill the GPU -- measured on B200:
    // stick-figure 216 us against 267, fox-tail 191 against 275 -- reverse mode as soon as they would: the 64-joint x 2000-frame tree
    // (5.7 M tangent threads) 35.3 -> 3.96 ms. OMO_ADJ_MODE=forward / reverse forces either.
    if (ctx->adj_mode < 0) { const char* e = std::getenv("OMO_ADJ_MODE"); ctx->adj_mode = !e ? 2 : (std::strcmp(e, "forward") == 0 ? 0 : (std::strcmp(e, "reverse") == 0 ? 1 : 2)); }
    const bool reverse = ctx->adj_mode == 1 || (ctx->adj_mode == 2 && 3LL * nk * N >= 200000LL);
    const bool traj_on = ctx->w_traj != 0.0;
    const size_t L_cells = 3 * (reverse ? (size_t)3 * J * N : cells);
    if (L_cells > ctx->adj_L_cap) {
        if ((st = dev_alloc(ctx, &ctx->d_adj_L, L_cells)) != OMO_OK) return st;
        ctx->adj_L_cap = L_cells;
    }
    const size_t ws_cells = reverse ? (size_t)J * (traj_on ? omo_ad::kRevFieldsTraj : omo_ad::kRevFields) * N : 0;
    if (ws_cells > ctx->adj_ws_cap) {
        if ((st = dev_alloc(ctx, &ctx->d_adj_ws, ws_cells)) != OMO_OK) return st;
        ctx->adj_ws_cap = ws_cells;
    }
    if (cells > ctx->adj_cap) {
        if ((st = dev_alloc(ctx, &ctx->d_adj_G, cells)) != OMO_OK) return st;
        if ((st = dev_alloc(ctx, &ctx->d_adj_contrib, 10 * cells)) != OMO_OK) return st;
        if ((st = dev_alloc(ctx, &ctx->d_adj_m, cells)) != OMO_OK) return st;
        if ((st = dev_alloc(ctx, &ctx->d_adj_gfull, (size_t)std::max(ctx->D, 1))) != OMO_OK) return st;
        if ((st = dev_alloc(ctx, &ctx->d_oroot, (size_t)3 * N)) != OMO_OK) return st;
        ctx->adj_cap = cells;
    }
    if (F > ctx->adj_F_cap) { if ((st = dev_alloc(ctx, &ctx->d_adj_gout, (size_t)F)) != OMO_OK) return st; ctx->adj_F_cap = F; }
    if (F > ctx->free_cap) { if ((st = dev_alloc(ctx, &ctx->d_free, (size_t)F)) != OMO_OK) return st; ctx->free_cap = F; ctx->free_uploaded.clear(); }
    if (F > 0 && (ctx->free_uploaded.size() != (size_t)F || !std::equal(ctx->free_uploaded.begin(), ctx->free_uploaded.end(), free_idx))) {
        ctx->free_uploaded.assign(free_idx, free_idx + F);
        OMO_CUDA(cudaMemcpyAsync(ctx->d_free, ctx->free_uploaded.data(), (size_t)F * sizeof(int), cudaMemcpyHostToDevice, ctx->stream));
        OMO_CUDA(cudaStreamSynchronize(ctx->stream));
    }
    // the evaluated point: f, and the per-sample q, qd, qdd, tau of the OptimizationCache (per-sample-output kernel, full tree)
    OMO_CUDA(cudaMemcpyAsync(ctx->d_X, x, (size_t)ctx->D * sizeof(double), cudaMemcpyHostToDevice, ctx->stream));
    const bool traj = ctx->w_traj != 0.0;
    if ((st = run_single_out(ctx, ctx->d_X, traj, ctx->d_oroot)) != OMO_OK) return st;
    unsigned flag0 = 1u;
    double f0 = 0.0;
    OMO_CUDA(cudaMemcpyAsync(&flag0, ctx->d_flags, sizeof(unsigned), cudaMemcpyDeviceToHost, ctx->stream));
    OMO_CUDA(cudaMemcpyAsync(&f0, ctx->d_f, sizeof(double), cudaMemcpyDeviceToHost, ctx->stream));
    omo::OmoAdjParams A;
    std::memset(&A, 0, sizeof(A));
    A.sub_end = ctx->d_sub_end; A.q = ctx->d_oq; A.qd = ctx->d_oqd; A.qdd = ctx->d_oqdd; A.tau = ctx->d_otau;
    A.root_e = ctx->hs.dyn[0] ? ctx->d_oroot : nullptr;
    A.wcost = ctx->d_wcost; A.p_orig = traj ? ctx->d_p_orig : nullptr; A.tgrid = ctx->d_tgrid; A.w_traj = ctx->w_traj; A.dt = ctx->dt;
    A.N_s = N; A.n_keyed = nk; A.n_cp = ctx->n_cp; A.D = ctx->D; A.cyclic = ctx->cyclic;
    A.x = ctx->d_X; A.cp_of_pos = ctx->d_cp_of_pos; A.first_of_pos = ctx->d_first_of_pos; A.last_of_pos = ctx->d_last_of_pos;
    A.pos_of_cp = ctx->d_pos_of_cp; A.ci_of_pos = ctx->d_ci_of_pos;
    A.L_rows = reverse ? 3 * J : nk; A.L_by_channel = reverse ? 1 : 0; A.ws = ctx->d_adj_ws;
    A.L = ctx->d_adj_L; A.G = ctx->d_adj_G; A.contrib = ctx->d_adj_contrib; A.seg_m = ctx->d_adj_m; A.gfull = ctx->d_adj_gfull;
    A.free_idx = ctx->d_free; A.F = F; A.x0 = ctx->d_x0; A.wparam = ctx->d_wparam; A.gout = ctx->d_adj_gout;
    if (nk > 0) {
        if (reverse) {
            // one thread per sample: all 3 x 3J partials from O(1) passes over the tree
            const unsigned gr = (unsigned)((N + 127) / 128);
            if (traj_on) omo::adj_reverse_kernel<true><<<gr, 128, 0, ctx->stream>>>(A);
            else omo::adj_reverse_kernel<false><<<gr, 128, 0, ctx->stream>>>(A);
        } else {
            const long long nt = 3LL * nk * N;
            const unsigned gb = (unsigned)((nt + 127) / 128);
            if (ctx->adj_max_sub <= 16) omo::adj_tangent_kernel<16><<<gb, 128, 0, ctx->stream>>>(A);
            else if (ctx->adj_max_sub <= 64) omo::adj_tangent_kernel<64><<<gb, 128, 0, ctx->stream>>>(A);
            else omo::adj_tangent_kernel<OMO_MAX_JOINTS><<<gb, 128, 0, ctx->stream>>>(A);
        }
        OMO_CUDA(cudaGetLastError());
        const unsigned gc = (unsigned)((cells + 127) / 128);
        omo::adj_stencil_kernel<<<gc, 128, 0, ctx->stream>>>(A);
        omo::adj_bezier_kernel<<<gc, 128, 0, ctx->stream>>>(A);
        OMO_CUDA(cudaGetLastError());
        ctx->launches += 3;
    }
    if (ctx->n_cp > 0) {
        omo::adj_reduce_kernel<<<(5 * ctx->n_cp + 127) / 128, 128, 0, ctx->stream>>>(A);
        OMO_CUDA(cudaGetLastError());
        ctx->launches += 1;
    }
    if (F > 0) {
        omo::adj_output_kernel<<<(F + 127) / 128, 128, 0, ctx->stream>>>(A);
        OMO_CUDA(cudaGetLastError());
        OMO_CUDA(cudaMemcpyAsync(grad_out, ctx->d_adj_gout, (size_t)F * sizeof(double), cudaMemcpyDeviceToHost, ctx->stream));
        ctx->launches += 1;
    }
    OMO_CUDA(cudaStreamSynchronize(ctx->stream));
    if ((flag0 & 1u) == 0u)
        return fail(ctx, OMO_UNSUPPORTED, "a curve's key times are not in ascending order at this x: its segments are found in list order (curve.cpp:100-128), which the analytic gradient does not cover; use omo_eval_grad");
    if (f0_out) *f0_out = f0;
    ctx->adj_calls += 1;
    return OMO_OK;
}

omo_status omo_debug_scene_table(const omo_scene_desc* scene, void* out, unsigned long long cap, int* sub_end_out)
{
    if (!scene || !out || cap < sizeof(OmoDevScene)) return OMO_BAD_ARG;
    omo_ctx* ctx = new omo_ctx();
    omo_status st = build_tables(ctx, scene);
    if (st == OMO_OK) {
        ctx->dt = 0.2;
        refresh_residues(ctx);
        std::memcpy(out, &ctx->hs, sizeof(OmoDevScene));
        if (sub_end_out) {
            const int J = ctx->J;
            for (int j = J - 1; j >= 0; --j) {
                sub_end_out[j] = j + 1;
                for (int k = j + 1; k < J; ++k) if (ctx->hs.parent[k] == j) sub_end_out[j] = std::max(sub_end_out[j], sub_end_out[k]);
            }
        }
    }
    delete ctx;
    return st;
}

omo_status omo_grad_sparsity(const omo_ctx* ctx, unsigned long long* out2)
{
    if (!ctx || !out2) return OMO_BAD_ARG;
    out2[0] = ctx->sparse_items; out2[1] = ctx->dense_items;
    return OMO_OK;
}

void omo_shard_range(int n, int rank, int world, int* begin, int* count)
{
    if (world < 1) world = 1;
    const int base = n / world, rem = n % world;
    const int b = rank * base + std::min(rank, rem);
    if (begin) *begin = b;
    if (count) *count = base + (rank < rem ? 1 : 0);
}

omo_status omo_device_alloc(omo_ctx* ctx, unsigned long long bytes, void** out)
{
    if (!ctx || !out) return OMO_BAD_ARG;
    OMO_CUDA(cudaSetDevice(ctx->device));
    OMO_CUDA(cudaMalloc(out, bytes ? bytes : 1));
    return OMO_OK;
}
omo_status omo_device_free(omo_ctx* ctx, void* p)
{
    if (!ctx) return OMO_BAD_ARG;
    OMO_CUDA(cudaSetDevice(ctx->device));
    OMO_CUDA(cudaFree(p));
    return OMO_OK;
}
omo_status omo_host_alloc_pinned(omo_ctx* ctx, unsigned long long bytes, void** out)
{
    if (!ctx || !out) return OMO_BAD_ARG;
    OMO_CUDA(cudaSetDevice(ctx->device));
    OMO_CUDA(cudaHostAlloc(out, bytes ? bytes : 1, cudaHostAllocDefault));
    return OMO_OK;
}
omo_status omo_host_free_pinned(omo_ctx* ctx, void* p)
{
    if (!ctx) return OMO_BAD_ARG;
    OMO_CUDA(cudaFreeHost(p));
    return OMO_OK;
}
omo_status omo_memcpy_h2d(omo_ctx* ctx, void* dst, const void* src, unsigned long long bytes)
{
    if (!ctx) return OMO_BAD_ARG;
    OMO_CUDA(cudaMemcpyAsync(dst, src, bytes, cudaMemcpyHostToDevice, ctx->stream));
    OMO_CUDA(cudaStreamSynchronize(ctx->stream));
    return OMO_OK;
}
omo_status omo_memcpy_d2h(omo_ctx* ctx, void* dst, const void* src, unsigned long long bytes)
{
    if (!ctx) return OMO_BAD_ARG;
    OMO_CUDA(cudaMemcpyAsync(dst, src, bytes, cudaMemcpyDeviceToHost, ctx->stream));
    OMO_CUDA(cudaStreamSynchronize(ctx->stream));
    return OMO_OK;
}
omo_status omo_synchronize(omo_ctx* ctx)
{
    if (!ctx) return OMO_BAD_ARG;
    OMO_CUDA(cudaStreamSynchronize(ctx->stream));
    return OMO_OK;
}
void* omo_stream(omo_ctx* ctx) { return ctx ? (void*)ctx->stream : nullptr; }

omo_status omo_set_counting(omo_ctx* ctx, int enable)
{
    if (!ctx) return OMO_BAD_ARG;
    ctx->counting = enable != 0;
    OMO_CUDA(cudaMemsetAsync(ctx->d_counters, 0, 4 * sizeof(unsigned long long), ctx->stream));
    return OMO_OK;
}

omo_status omo_get_counters(omo_ctx* ctx, unsigned long long* out5)
{
    if (!ctx || !out5) return OMO_BAD_ARG;
    OMO_CUDA(cudaMemcpyAsync(out5, ctx->d_counters, 4 * sizeof(unsigned long long), cudaMemcpyDeviceToHost, ctx->stream));
    OMO_CUDA(cudaStreamSynchronize(ctx->stream));
    out5[4] = ctx->launches;
    return OMO_OK;
}

unsigned long long omo_graph_replays(const omo_ctx* ctx) { return ctx ? ctx->sg_replays : 0; }

omo_status omo_kernel_timing(omo_ctx* ctx, int reset, double* total_ms, int* launches)
{
    if (!ctx) return OMO_BAD_ARG;
    OMO_CUDA(cudaStreamSynchronize(ctx->stream));
    double tot = 0.0;
    for (auto& pr : ctx->timed) {
        float ms = 0.f;
        OMO_CUDA(cudaEventElapsedTime(&ms, pr.first, pr.second));
        tot += ms;
    }
    if (total_ms) *total_ms = tot;
    if (launches) *launches = (int)ctx->timed.size();
    if (reset) {
        for (auto& pr : ctx->timed) { cudaEventDestroy(pr.first); cudaEventDestroy(pr.second); }
        ctx->timed.clear();
        ctx->timing = reset > 0;  // reset = 1: (re)start timing; reset = -1: stop
    }
    return OMO_OK;
}

omo_status omo_jit_probe(const omo_scene_desc* scene, int fk, int count, const char* dump_prefix, char* log, int log_cap)
{
    if (!scene) return OMO_BAD_ARG;
    omo_ctx* ctx = new omo_ctx();
    omo_status st = build_tables(ctx, scene);
    std::string msg;
    if (st == OMO_OK) {
        const bool ik = ctx->n_ik > 0;
        const bool self = (count & 2) != 0;   // `count` bit 1: the self-preparing instance
        omo_jit::KernelSpec spec = omo_jit::generate(ctx->hs, fk != 0 || ik, ik, (count & 1) != 0 && !self, self ? ctx->n_cp : 0);
        std::vector<char> cubin;
        std::string clog;
        // the same key acquire_jit uses: a probe on the build machine pre-populates the on-disk cubin cache of the run machine
        const uint64_t key = omo_jit::fnv1a(spec.source + "#sub" + std::to_string(spec.sub)) ^ omo_jit::headers_hash();
        bool ok = omo_jit::disk_cache_load(key, cubin);
        if (ok) clog = "(cubin taken from the on-disk cache)";
        else {
            ok = omo_jit::compile(spec.source, cubin, clog, spec.sub);
            if (ok) omo_jit::disk_cache_store(key, cubin);
        }
        if (dump_prefix) {
            if (FILE* fh = std::fopen((std::string(dump_prefix) + ".cu").c_str(), "w")) { std::fputs(spec.source.c_str(), fh); std::fclose(fh); }
            if (ok) if (FILE* fh = std::fopen((std::string(dump_prefix) + ".cubin").c_str(), "wb")) { std::fwrite(cubin.data(), 1, cubin.size(), fh); std::fclose(fh); }
        }
        msg = "smem_slots=" + std::to_string(spec.n_smem_slots) + " thread_slots=" + std::to_string(spec.n_reg_slots) + " min_blocks=" +
              std::to_string(spec.min_blocks) + " smem_bytes=" + std::to_string(spec.smem_bytes) + " cubin_bytes=" + std::to_string(cubin.size()) + "\n" + clog;
        if (!ok) st = OMO_CUDA_ERROR;
    } else {
        msg = ctx->err;
    }
    if (log && log_cap > 0) { std::strncpy(log, msg.c_str(), (size_t)log_cap - 1); log[log_cap - 1] = 0; }
    delete ctx;
    return st;
}

omo_status omo_jit_info(omo_ctx* ctx, unsigned long long* out4, char* log, int log_cap)
{
    if (!ctx) return OMO_BAD_ARG;
    if (out4) {
        out4[0] = ctx->jit_launches;
        out4[1] = 0; out4[2] = 0; out4[3] = 0;
        for (const auto& k : ctx->jit)
            if (k.ok) { out4[1] += 1; out4[2] = std::max<unsigned long long>(out4[2], (unsigned long long)k.regs); out4[3] += (unsigned long long)k.compile_ms; }
    }
    if (log && log_cap > 0) { std::strncpy(log, ctx->jit_log.c_str(), (size_t)log_cap - 1); log[log_cap - 1] = 0; }
    return OMO_OK;
}

omo_status omo_jit_describe(omo_ctx* ctx, char* out, int cap)
{
    if (!ctx || !out || cap <= 0) return OMO_BAD_ARG;
    std::string t = "[";
    bool first = true;
    for (int i = 0; i < 16; ++i) {
        const omo_ctx::JitKernel& k = ctx->jit[i];
        if (!k.ok) continue;
        char buf[360];
        std::snprintf(buf, sizeof(buf), "%s{\"fk\": %s, \"ik\": %s, \"count\": %s, \"self_preparing\": %s, \"form\": \"%s\", \"registers\": %d, \"smem_bytes\": %zu, "
                      "\"items_per_block\": %d, \"build_ms\": %.1f, \"from_disk_cache\": %s}",
                      first ? "" : ", ", (i & 1) ? "true" : "false", (i & 2) ? "true" : "false", (i & 4) ? "true" : "false", (i & 8) ? "true" : "false",
                      k.loop_form ? "loop" : "unrolled", k.regs, k.smem, k.sub, k.compile_ms, k.from_disk ? "true" : "false");
        t += buf;
        first = false;
    }
    t += "]";
    std::strncpy(out, t.c_str(), (size_t)cap - 1); out[cap - 1] = 0;
    return OMO_OK;
}

omo_status omo_measure_dfma_peak(omo_ctx* ctx, double* flops_per_s)
{
    if (!ctx || !flops_per_s) return OMO_BAD_ARG;
    OMO_CUDA(cudaSetDevice(ctx->device));
    cudaDeviceProp prop;
    OMO_CUDA(cudaGetDeviceProperties(&prop, ctx->device));
    const int blocks = prop.multiProcessorCount * 8, threads = 256, iters = 4096;
    double* d_out = nullptr;
    OMO_CUDA(cudaMalloc((void**)&d_out, (size_t)blocks * threads * sizeof(double)));
    cudaEvent_t e0, e1;
    OMO_CUDA(cudaEventCreate(&e0)); OMO_CUDA(cudaEventCreate(&e1));
    double best = 0.0;
    for (int rep = 0; rep < 6; ++rep) {
        OMO_CUDA(cudaEventRecord(e0, ctx->stream));
        omo::dfma_peak_kernel<<<blocks, threads, 0, ctx->stream>>>(d_out, iters, 1.0 + rep);
        OMO_CUDA(cudaEventRecord(e1, ctx->stream));
        OMO_CUDA(cudaStreamSynchronize(ctx->stream));
        float ms = 0.f;
        OMO_CUDA(cudaEventElapsedTime(&ms, e0, e1));
        const double flops = 2.0 * 64.0 * (double)iters * (double)blocks * threads;
        if (rep > 0) best = std::max(best, flops / (ms * 1e-3));
    }
    cudaEventDestroy(e0); cudaEventDestroy(e1); cudaFree(d_out);
    *flops_per_s = best;
    return OMO_OK;
}

omo_status omo_selftest_division(omo_ctx* ctx, const double* num, const double* den, int n, double* q_fast, double* q_ref, int* ok)
{
    if (!ctx || !num || !den || !q_fast || !q_ref || !ok || n <= 0) return OMO_BAD_ARG;
    OMO_CUDA(cudaSetDevice(ctx->device));
    double *d_n = nullptr, *d_d = nullptr, *d_q = nullptr, *d_r = nullptr;
    int* d_ok = nullptr;
    const size_t bytes = (size_t)n * sizeof(double);
    OMO_CUDA(cudaMalloc((void**)&d_n, bytes)); OMO_CUDA(cudaMalloc((void**)&d_d, bytes));
    OMO_CUDA(cudaMalloc((void**)&d_q, bytes)); OMO_CUDA(cudaMalloc((void**)&d_r, bytes));
    OMO_CUDA(cudaMalloc((void**)&d_ok, (size_t)n * sizeof(int)));
    OMO_CUDA(cudaMemcpyAsync(d_n, num, bytes, cudaMemcpyHostToDevice, ctx->stream));
    OMO_CUDA(cudaMemcpyAsync(d_d, den, bytes, cudaMemcpyHostToDevice, ctx->stream));
    omo::division_selftest_kernel<<<(n + 255) / 256, 256, 0, ctx->stream>>>(d_n, d_d, n, d_q, d_r, d_ok);
    OMO_CUDA(cudaGetLastError());
    OMO_CUDA(cudaMemcpyAsync(q_fast, d_q, bytes, cudaMemcpyDeviceToHost, ctx->stream));
    OMO_CUDA(cudaMemcpyAsync(q_ref, d_r, bytes, cudaMemcpyDeviceToHost, ctx->stream));
    OMO_CUDA(cudaMemcpyAsync(ok, d_ok, (size_t)n * sizeof(int), cudaMemcpyDeviceToHost, ctx->stream));
    OMO_CUDA(cudaStreamSynchronize(ctx->stream));
    cudaFree(d_n); cudaFree(d_d); cudaFree(d_q); cudaFree(d_r); cudaFree(d_ok);
    ctx->launches += 1;
    return OMO_OK;
}

}  // extern "C"

// ---- multi-process gather over NVLink / NVSwitch peer memory (include/omo.h: omo_peer_gather_*) -----------------------------------
extern "C" omo_status omo_peer_gather_create(omo_ctx* ctx, int rank, int world, int capacity, omo_peer_gather** out, unsigned char* handle_out)
{
    DeviceLock device_lock(ctx);
    if (!ctx || !out || !handle_out || world < 1 || world > OMO_MAX_PEERS || rank < 0 || rank >= world || capacity < 1)
        return fail(ctx, OMO_BAD_ARG, "omo_peer_gather_create: bad argument (1 <= world <= 16, 0 <= rank < world, capacity >= 1)");
    static_assert(sizeof(cudaIpcMemHandle_t) <= OMO_IPC_HANDLE_BYTES, "OMO_IPC_HANDLE_BYTES is too small for cudaIpcMemHandle_t");
    *out = nullptr;
    OMO_CUDA(cudaSetDevice(ctx->device));
    omo_peer_gather* g = new omo_peer_gather();
    g->ctx = ctx; g->rank = rank; g->world = world; g->capacity = capacity;
    const size_t bytes = peer_flag_offset(capacity) + (size_t)OMO_MAX_PEERS * sizeof(unsigned long long);
    cudaError_t e = cudaMalloc((void**)&g->d_block, bytes);
    if (e == cudaSuccess) e = cudaMemset(g->d_block, 0, bytes);
    if (e == cudaSuccess) e = cudaMalloc((void**)&g->d_ticket, sizeof(unsigned));
    if (e == cudaSuccess) e = cudaMemset(g->d_ticket, 0, sizeof(unsigned));
    if (e == cudaSuccess) e = cudaHostAlloc((void**)&g->h_error, sizeof(int), cudaHostAllocMapped);
    if (e == cudaSuccess) { *g->h_error = 0; e = cudaHostGetDevicePointer((void**)&g->d_error, g->h_error, 0); }
    cudaIpcMemHandle_t h;
    if (e == cudaSuccess) e = cudaIpcGetMemHandle(&h, g->d_block);
    if (e == cudaSuccess) e = cudaDeviceSynchronize();   // the zeroed flags are in place before any peer can learn the handle
    if (e != cudaSuccess) {
        ctx->err = std::string("omo_peer_gather_create: ") + cudaGetErrorString(e);
        omo_peer_gather_destroy(g);
        return e == cudaErrorMemoryAllocation ? OMO_OOM : OMO_CUDA_ERROR;
    }
    std::memset(handle_out, 0, OMO_IPC_HANDLE_BYTES);
    std::memcpy(handle_out, &h, sizeof(h));
    g->buf[rank] = reinterpret_cast<double*>(g->d_block);
    g->flag[rank] = reinterpret_cast<unsigned long long*>(g->d_block + peer_flag_offset(capacity));
    *out = g;
    return OMO_OK;
}

extern "C" omo_status omo_peer_gather_connect(omo_peer_gather* g, const unsigned char* handles)
{
    if (!g || !handles) return OMO_BAD_ARG;
    omo_ctx* ctx = g->ctx;
    DeviceLock device_lock(ctx);
    OMO_CUDA(cudaSetDevice(ctx->device));
    for (int p = 0; p < g->world; ++p) {
        if (p == g->rank || g->opened[p]) continue;
        cudaIpcMemHandle_t h;
        std::memcpy(&h, handles + (size_t)p * OMO_IPC_HANDLE_BYTES, sizeof(h));
        void* ptr = nullptr;
        const cudaError_t e = cudaIpcOpenMemHandle(&ptr, h, cudaIpcMemLazyEnablePeerAccess);
        if (e != cudaSuccess) {
            cudaGetLastError();
            ctx->err = "omo_peer_gather_connect: cudaIpcOpenMemHandle(rank " + std::to_string(p) + "): " + cudaGetErrorString(e) +
                       " (no peer access between the devices, or the ranks do not share an IPC namespace): gather with NCCL instead";
            return OMO_NCCL_ERROR;
        }
        g->opened[p] = ptr;
        g->buf[p] = reinterpret_cast<double*>(ptr);
        g->flag[p] = reinterpret_cast<unsigned long long*>(static_cast<unsigned char*>(ptr) + peer_flag_offset(g->capacity));
    }
    g->connected = true;
    return OMO_OK;
}

extern "C" omo_status omo_peer_gather_eval(omo_peer_gather* g, const double* dX, int B_local, int first, int sync)
{
    if (!g) return OMO_BAD_ARG;
    omo_ctx* ctx = g->ctx;
    DeviceLock device_lock(ctx);
    if (!g->connected) return fail(ctx, OMO_NOT_READY, "omo_peer_gather_connect has not been called");
    if ((!dX && B_local > 0) || B_local < 0 || first < 0 || (long long)first + B_local > g->capacity)
        return fail(ctx, OMO_BAD_ARG, "omo_peer_gather_eval: the slice [first, first + B_local) does not fit the gathered array");
    omo_status st = ensure_ready(ctx);
    if (st != OMO_OK) return st;
    g->seq += 1;
    g->first = first;
    if (B_local > 0) {
        ctx->peer = g;
        st = eval_device_impl(ctx, dX, B_local, ctx->d_f, nullptr);
        ctx->peer = nullptr;
        if (st != OMO_OK) return st;
    } else {
        // a rank without candidates still takes part in the exchange: flags only
        if ((st = ensure_capacity(ctx, 1)) != OMO_OK) return st;
        ctx->peer = g;
        st = launch_finish(ctx, 0, std::max(ctx->n_tiles, 1), nullptr, nullptr);
        ctx->peer = nullptr;
        if (st != OMO_OK) return st;
    }
    if (sync) {
        OMO_CUDA(cudaStreamSynchronize(ctx->stream));
        if (*static_cast<volatile int*>(g->h_error) != 0) {
            *g->h_error = 0;
            return fail(ctx, OMO_NCCL_ERROR, "omo_peer_gather_eval: a rank's slice did not arrive within the time-out");
        }
    }
    return OMO_OK;
}

extern "C" const double* omo_peer_gather_values(const omo_peer_gather* g)
{
    return g ? g->buf[g->rank] + (size_t)(g->seq & 1ull) * g->capacity : nullptr;
}

extern "C" int omo_peer_gather_error(omo_peer_gather* g)
{
    if (!g || !g->h_error) return 0;
    const int e = *static_cast<volatile int*>(g->h_error);
    *g->h_error = 0;
    return e;
}

extern "C" void omo_peer_gather_destroy(omo_peer_gather* g)
{
    if (!g) return;
    if (g->ctx) { cudaSetDevice(g->ctx->device); if (g->ctx->stream) cudaStreamSynchronize(g->ctx->stream); }
    for (int p = 0; p < OMO_MAX_PEERS; ++p) if (g->opened[p]) cudaIpcCloseMemHandle(g->opened[p]);
    if (g->d_block) cudaFree(g->d_block);
    if (g->d_ticket) cudaFree(g->d_ticket);
    if (g->h_error) cudaFreeHost(g->h_error);
    delete g;
}

// ---- the batch objective without a single allocation (include/omo.h: omo_eval_device_static) -------------------------------------
// Scratch of the prepare -> objective -> finish chain in a file-scope array instead of the context's cudaMalloc'ed work buffers.
// Every device that loads libomo.so holds OMO_STATIC_SCRATCH_BYTES of it until the process exits; every call on a device shares it.
static __device__ __align__(256) unsigned char g_static_scratch[OMO_STATIC_SCRATCH_BYTES];

namespace {

// In host code `g_static_scratch` is only the host-side stand-in of the symbol: the device address comes from
// cudaGetSymbolAddress, asked once per device with that device current.
std::mutex g_static_scratch_mutex;
unsigned char* g_static_scratch_addr[64] = {};

omo_status static_scratch_address(omo_ctx* ctx, unsigned char** out)
{
    std::lock_guard<std::mutex> g(g_static_scratch_mutex);
    const int dev = ctx->device;
    if (dev < 0 || dev >= 64) return fail(ctx, OMO_BAD_ARG, "device index out of range");
    if (!g_static_scratch_addr[dev]) {
        OMO_CUDA(cudaSetDevice(dev));
        void* p = nullptr;
        OMO_CUDA(cudaGetSymbolAddress(&p, g_static_scratch));
        if (!p || (reinterpret_cast<uintptr_t>(p) & 255u) != 0) return fail(ctx, OMO_CUDA_ERROR, "cudaGetSymbolAddress(g_static_scratch) returned an unusable address");
        g_static_scratch_addr[dev] = static_cast<unsigned char*>(p);
    }
    *out = g_static_scratch_addr[dev];
    return OMO_OK;
}

// The work buffers of ensure_capacity for B candidates, carved out of one block (each 256-byte aligned).
struct StaticLayout {
    size_t prep = 0, dparam = 0, seg = 0, partial = 0, flags = 0, tile_seg = 0, total = 0;
    StaticLayout(const omo_ctx* ctx, size_t B)
    {
        const size_t tiles = (size_t)std::max(ctx->n_tiles, 1), n_cp = (size_t)ctx->n_cp;
        size_t off = 0;
        auto take = [&off](size_t bytes) { const size_t at = off; off = (off + std::max<size_t>(bytes, 1) + 255) & ~(size_t)255; return at; };
        prep = take(B * 6 * n_cp * sizeof(double));
        dparam = take(B * (size_t)prep_parts(ctx) * sizeof(double));
        seg = take((B * n_cp + OMO_SEGW) * OMO_SEGD * sizeof(double));
        partial = take(B * tiles * 2 * sizeof(double));
        flags = take(B * sizeof(unsigned));
        tile_seg = take(B * tiles * (size_t)std::max(ctx->hs.n_keyed, 1) * sizeof(unsigned short));
        total = off;
    }
};

// Points the context's work-buffer fields at the static block for the duration of one call and takes the choices that would
// allocate (specialised-kernel module load, sorted work lists, timing events) off the table; everything is put back on exit.
struct StaticScratchScope {
    omo_ctx* c;
    double *prep, *dparam, *partial, *seg; unsigned* flags; unsigned short* tile_seg;
    int row0, jit_mode, batch_hint; bool prefer_jit, timing;
    StaticScratchScope(omo_ctx* ctx, unsigned char* base, const StaticLayout& L)
        : c(ctx), prep(ctx->d_prep), dparam(ctx->d_dparam), partial(ctx->d_partial), seg(ctx->d_seg), flags(ctx->d_flags), tile_seg(ctx->d_tile_seg),
          row0(ctx->row0), jit_mode(ctx->jit_mode), batch_hint(ctx->batch_hint), prefer_jit(ctx->prefer_jit), timing(ctx->timing)
    {
        c->d_prep = reinterpret_cast<double*>(base + L.prep); c->d_dparam = reinterpret_cast<double*>(base + L.dparam);
        c->d_seg = reinterpret_cast<double*>(base + L.seg); c->d_partial = reinterpret_cast<double*>(base + L.partial);
        c->d_flags = reinterpret_cast<unsigned*>(base + L.flags); c->d_tile_seg = reinterpret_cast<unsigned short*>(base + L.tile_seg);
        c->row0 = 0; c->jit_mode = 0; c->batch_hint = 0; c->prefer_jit = false; c->timing = false;
    }
    ~StaticScratchScope()
    {
        c->d_prep = prep; c->d_dparam = dparam; c->d_partial = partial; c->d_seg = seg; c->d_flags = flags; c->d_tile_seg = tile_seg;
        c->row0 = row0; c->jit_mode = jit_mode; c->batch_hint = batch_hint; c->prefer_jit = prefer_jit; c->timing = timing;
    }
};

}  // namespace

extern "C" omo_status omo_eval_device_static(omo_ctx* ctx, const double* dX, int B, double* d_f_out, double* d_terms_out, int sync)
{
    DeviceLock device_lock(ctx);
    if (!ctx || !dX || !d_f_out || B < 0) return fail(ctx, OMO_BAD_ARG, "bad argument");
    if (B == 0) return OMO_OK;
    if (!ctx->have_window || !ctx->have_weights || !ctx->have_original) return fail(ctx, OMO_NOT_READY, "window / weights / x_original not set");
    // the state derived from them (cost weights, p_original, the pruned bodies' share of the cost) lives in buffers that the allocating
    // entry points build: this one refuses to
    if (ctx->wcost_dirty || ctx->p_orig_dirty)
        return fail(ctx, OMO_NOT_READY, "omo_eval_device_static allocates nothing: after omo_set_window / omo_set_weights / omo_set_original call "
                                         "omo_get_p_original (or any omo_eval*) once, outside the no-allocation region, to build the derived state");
    const StaticLayout L(ctx, (size_t)B);
    if (L.total > (size_t)OMO_STATIC_SCRATCH_BYTES) {
        int fit = 0, over = B;   // largest batch of this scene and window that fits (the layout grows with B)
        while (over - fit > 1) {
            const int mid = fit + (over - fit) / 2;
            if (StaticLayout(ctx, (size_t)mid).total <= (size_t)OMO_STATIC_SCRATCH_BYTES) fit = mid; else over = mid;
        }
        return fail(ctx, OMO_BAD_ARG, "omo_eval_device_static: " + std::to_string(B) + " candidates need " + std::to_string(L.total) + " bytes of scratch, the static array holds " +
                                          std::to_string((size_t)OMO_STATIC_SCRATCH_BYTES) + " (at most " + std::to_string(fit) + " candidates of this scene and window per call)");
    }
    OMO_CUDA(cudaSetDevice(ctx->device));
    unsigned char* base = nullptr;
    omo_status st = static_scratch_address(ctx, &base);
    if (st != OMO_OK) return st;
    if ((st = activate_scene(ctx)) != OMO_OK) return st;   // a copy into constant memory when another context used the device in between; no allocation
    {
        // Nothing in the scratch has to start at zero: prepare_kernel writes every prep / seg / tile_seg / dparam / flag entry of the B rows
        // (flag words shared by several blocks are preset by run_prepare's memset) and the objective kernel every partial sum before
        // finish_kernel reads them, so the previous call's contents are never observed.
        StaticScratchScope scope(ctx, base, L);
        if ((st = run_prepare(ctx, dX, B)) != OMO_OK) return st;
        OmoEvalParams P;
        fill_params(ctx, P);
        P.perm = nullptr;   // the sorted work lists belong to the specialised kernel's launches
        const bool fk = (ctx->w_traj != 0.0) || ctx->n_ik > 0;
        P.p_orig = (ctx->w_traj != 0.0) ? ctx->d_p_orig : nullptr;
        if ((st = launch_objective(ctx, P, B, fk, false, d_f_out, d_terms_out)) != OMO_OK) return st;
    }
    if (sync) OMO_CUDA(cudaStreamSynchronize(ctx->stream));
    return OMO_OK;
}
