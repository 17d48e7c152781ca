/*
 * omo.h -- C ABI of the B200 objective engine (libomo.so, built from optimo_b200/csrc/).
 *
 * This is the drop-in boundary for OptiMo's objective(+gradient) hot path. Every entry point takes plain
 * pointers and sizes; no C++ / torch types cross it. All paths below are relative to the reference checkout.
 *
 *   What it replaces                                             Reference
 *   ------------------------------------------------------------ ------------------------------------------
 *   Core::CalculateFullObjective (f = C + D_param + D_traj)      OptiMo/core_rbdl.cpp:348-401
 *   Core::SetStates (time grid, curve / IK / FK sampling,        OptiMo/core_rbdl.cpp:108-166, :28-95
 *       4th-order stencils)
 *   Core::CalculateCost (RNEA loop, time-varying weight)         OptiMo/core_rbdl.cpp:307-346
 *   RBDL InverseDynamics / jcalc / InverseKinematics             external/rbdl-v2.5.0/src/Dynamics.cc:26-105,
 *                                                                src/Joint.cc:21-210, src/Kinematics.cc:615-709
 *   Curve::GetValue (segment lookup, Newton, Bernstein)          OptiMo/curve.cpp:43-93,100-128
 *   Object::SetDescriptor / MakeMotionCyclic                     OptiMo/object.cpp:147-186, curve.cpp:179-188
 *   Core::InitializeDynamicsModel (body table)                   OptiMo/core_rbdl.cpp:200-262
 *   Core::OptimizationCache (per-sample q, qd, qdd, tau, cost)   OptiMo/core.h:120-152, core_rbdl.cpp:341-343
 *
 * plus one NEW output the reference does not have: a central finite-difference gradient
 * on the free parameters, evaluated as one batch of 2F+1 objective columns.
 *
 * Conventions
 *   - every function returns an omo_status; omo_last_error(ctx) gives the message of the last failure.
 *   - the caller owns every host buffer it passes; an omo_ctx owns all device memory, streams and staging.
 *   - one omo_ctx is bound to one CUDA device and may be used by one host thread at a time.
 *   - there is NO CPU fallback: without a usable CUDA device omo_create fails with OMO_CUDA_ERROR.
 *   - all arithmetic is IEEE binary64.
 */
#ifndef OMO_H
#define OMO_H

#ifdef __cplusplus
extern "C" {
#endif

typedef enum omo_status {
    OMO_OK = 0,
    OMO_BAD_ARG = 1,     /* null pointer, size mismatch, unsupported scene (too many joints, ...) */
    OMO_CUDA_ERROR = 2,  /* any CUDA runtime failure, including "no device" */
    OMO_NCCL_ERROR = 3,
    OMO_OOM = 4,
    OMO_NOT_READY = 5,   /* window / weights / original parameters not set yet */
    OMO_UNSUPPORTED = 6  /* the request is outside what this entry point covers (omo_eval_grad_analytic: IK scenes, unsorted keys) */
} omo_status;

typedef struct omo_ctx omo_ctx;

/* Limits of the sm_100a kernels (scene tables live in constant memory). Any tree shape within OMO_MAX_JOINTS is served:
 * the walk parks one stack slot per open ancestor and per open branching; what does not fit a block's 227 KB of shared
 * memory (a binary tree of depth 7 with positions, a chain of more than ~17 joints) goes to per-thread slots. A tree that
 * needs more than shared memory + 128 per-thread slots (only degenerate "caterpillar" shapes do) is refused with OMO_BAD_ARG. */
#define OMO_MAX_JOINTS 128
#define OMO_MAX_IK 8
#define OMO_MAX_IK_CHAIN 8 /* rotating joints per IK chain (end effector excluded) */

/*
 * Flattened scene == what Core::ReadJson + Core::InitializeDynamicsModel leave behind
 * (OptiMo/core_io.cpp:228-250, core_rbdl.cpp:200-262).
 *
 * Joints are listed in Object::GetJoints() order (DFS pre-order, OptiMo/object.cpp:27-40), which is also
 * RBDL body id - 1 and q_index / 3. parent[0] == -1 is the root.
 *
 * Channels are the keyframable variables in Object::GetItems() order (object.cpp:47-61): channel 3*j+a is
 * rot_{x,y,z}[a] of joint j; channel 3*J + 3*h + a is pos_{x,y,z}[a] of IK handler h. The descriptor
 * x in R^{5*n_cp} lists, channel by channel, the control points of every KEYED channel in insertion order,
 * 5 doubles each: (time, value, theta, time_backward, time_forward) (object.cpp:110-145, curve.h:39-54).
 */
typedef struct omo_scene_desc {
    int n_joints;              /* J, root included */
    const int* parent;         /* [J] */
    const double* offset;      /* [J*3] Joint::offset_ */
    const double* mass;        /* [J]   Joint::mass_ (joint.cpp:90-119); root: 0 */
    const double* com;         /* [J*3] centre of mass in the joint frame */
    const double* inertia;     /* [J*9] row-major inertia about the COM */
    int n_ik;                  /* IK handlers */
    const int* ik_root;        /* [n_ik] joint index of the chain root (ikhandler.cpp:77-101) */
    const int* ik_end;         /* [n_ik] joint index of the end effector */
    int n_channels;            /* 3*J + 3*n_ik */
    const double* chan_const;  /* [n_channels] Variable::var_ (value of an un-keyed channel) */
    const int* chan_first_cp;  /* [n_channels] index of the channel's first control point in the descriptor, -1 if un-keyed */
    const int* chan_n_cp;      /* [n_channels] number of control points (0 if un-keyed) */
    int n_cp;                  /* total control points; descriptor length D = 5*n_cp */
    const int* chain;          /* [n_cp] for each keyed channel, at [first_cp, first_cp+n), the channel's control
                                  points (descriptor-global indices) in segment-list order, i.e. sorted by the
                                  time they had when they were inserted (curve.cpp:130-148) */
} omo_scene_desc;

typedef struct omo_gauss { double amp, mu, sigma; } omo_gauss; /* GaussianKernel, OptiMo/core.h:27-33 */

/* Per-candidate breakdown of f (core_rbdl.cpp:381-400). */
typedef struct omo_terms { double C, D_param, D_traj; } omo_terms;

/* ---- life cycle ------------------------------------------------------------------------------------ */
omo_status omo_create(const omo_scene_desc* scene, int device, omo_ctx** out);
void omo_destroy(omo_ctx* ctx);
const char* omo_last_error(const omo_ctx* ctx); /* ctx may be NULL: error of the last failed omo_create */
const char* omo_version(void);

/* ---- state the objective reads (snapshotted on the device) ----------------------------------------- */
/* Time grid: replays `for (t = min_t; t <= max_t; t += dt)` (core_rbdl.cpp:120-123). Needs >= 3 samples. */
omo_status omo_set_window(omo_ctx* ctx, double min_t, double max_t, double dt);
/* w_param: Core::CalculateParameterWeights() (core.cpp:161-182), length D. kernels: Core::kernels_ (time-varying
 * weight, core.cpp:191-199). cyclic: Object::is_cyclic_. */
omo_status omo_set_weights(omo_ctx* ctx, const double* w_param, int D, double w_traj, const omo_gauss* kernels,
                           int n_kernels, int cyclic);
/* x_original (core.h:237-243): uploads it and rebuilds p_original (N_s x 3J) on the device. */
omo_status omo_set_original(omo_ctx* ctx, const double* x_original, int D);

/* Which kernel evaluates a batch (scene-specialised or table-driven; they agree to ~1e-13, not bit for bit) depends on the size
 * of the launch. A context that evaluates a SLICE of a larger batch (one device of omo_multi, one rank of a multi-process job)
 * names the whole batch here, so that the choice -- and with it every bit of the result -- does not depend on how the batch was
 * split. 0 = no hint. */
omo_status omo_set_batch_hint(omo_ctx* ctx, int total_candidates);

/* ---- queries ---------------------------------------------------------------------------------------- */
int omo_num_samples(const omo_ctx* ctx);  /* N_s */
int omo_descriptor_size(const omo_ctx* ctx); /* D */
int omo_dof(const omo_ctx* ctx);          /* 3*J */
/* Bodies the batch kernels actually run per sample-eval: bodies whose torque cannot depend on the candidate (no
 * keyed / IK-driven angle on the body, its ancestors or its descendants) are evaluated once per window instead.
 * general_out (optional): how many of the walked bodies take the general EulerXYZ step (the rest are joints
 * frozen at zero angles, E = I). Used to report EXECUTED FLOPs. */
int omo_num_walked_bodies(const omo_ctx* ctx, int* general_out);
omo_status omo_get_time_grid(const omo_ctx* ctx, double* t_out /* [N_s] */);
omo_status omo_get_cost_weights(const omo_ctx* ctx, double* w_out /* [N_s] */);
omo_status omo_get_p_original(omo_ctx* ctx, double* p_out /* [N_s * 3J], sample-major */);

/* ---- the hot path ----------------------------------------------------------------------------------- */
/* B candidates, X row-major [B][D] in HOST memory. f_out[B]; terms_out[B] optional (NULL). */
omo_status omo_eval(omo_ctx* ctx, const double* X, int B, double* f_out, omo_terms* terms_out);

/* The same for candidates given in the callback's own parameter space (EvaluateObjective, OptiMo/core_opt.cpp:42-57, receives the
 * F free entries and expands them over x_initial through the select / revert maps of core_opt.cpp:99-121): row b of Xfree[B][F]
 * replaces entries free_idx[0..F) of x_base[D]; the expansion runs on the device, so only 8 B F bytes cross PCIe instead of
 * 8 B D (the reference's default freezes t and v of every control point: F = 0.6 D). Bit-identical to omo_eval on the expanded rows. */
omo_status omo_eval_free(omo_ctx* ctx, const double* x_base, const int* free_idx, int F, const double* Xfree, int B, double* f_out,
                         omo_terms* terms_out);

/* Same with X, f_out (and optional terms_out, 3 doubles per candidate) already in DEVICE memory of ctx's
 * device. Asynchronous on the ctx stream unless `sync` != 0. */
omo_status omo_eval_device(omo_ctx* ctx, const double* dX, int B, double* d_f_out, double* d_terms_out, int sync);

/* Host candidates in (X[B][D], pinned for a truly asynchronous copy; chunked and overlapped with the kernels exactly like omo_eval),
 * values left in DEVICE memory: d_f_out[B], optionally d_terms_out[3B]. What a rank of a multi-process job calls before it
 * gathers its shard with NCCL. Asynchronous on the ctx stream unless `sync` != 0. */
omo_status omo_eval_host_in(omo_ctx* ctx, const double* X, int B, double* d_f_out, double* d_terms_out, int sync);

/* NO COUNTERPART IN THE REFERENCE -- added for callers that cannot call cudaMalloc (code
 * under CUDA graph capture, tight serving loops, harnesses that forbid it). omo_eval_device with the scratch of the
 * prepare -> objective -> finish chain (prepared control points, segment records, per-tile segment indices, tile partial sums) in a
 * file-scope `static __device__` array of OMO_STATIC_SCRATCH_BYTES instead of the context's work buffers: the call makes no
 * cudaMalloc / cudaFree / module load and creates no events; dX, d_f_out[B] and the optional d_terms_out[3B] are the caller's
 * device memory. It always runs the table-driven kernel (loading a scene-specialised module allocates), so it agrees with
 * omo_eval_device bit for bit where that one takes the table-driven kernel too (launches below 64 K sample-evals) and to ~1e-13
 * relative otherwise.
 *   - OMO_BAD_ARG when B candidates of this scene and window need more scratch than the array holds (the message names the largest
 *     B that fits: ~21 KB per candidate for the 160 control points x 10 tiles of BASELINE configs[3], i.e. ~390 candidates).
 *   - OMO_NOT_READY until the state derived from window / weights / x_original has been built by an allocating entry point
 *     (omo_get_p_original or any omo_eval*, once after the omo_set_* calls).
 *   - every device that loads libomo.so holds the array until the process exits, and every call on a device shares it: two calls that
 *     overlap on the GPU (two contexts, or two streams) race. Nothing in it has to start at zero; it keeps the last call's contents.
 * Asynchronous on the ctx stream unless `sync` != 0. */
#define OMO_STATIC_SCRATCH_BYTES (8u << 20)
omo_status omo_eval_device_static(omo_ctx* ctx, const double* dX, int B, double* d_f_out, double* d_terms_out, int sync);

/* One evaluation that also fills the OptimizationCache of the reference (core.h:120-152): each of
 * q, qd, qdd, tau is [N_s * dof] sample-major, local_cost is [N_s]; any of them may be NULL. */
omo_status omo_eval_cached(omo_ctx* ctx, const double* x, double* f_out, omo_terms* terms_out, double* q, double* qd,
                           double* qdd, double* tau, double* local_cost);

/* Joint world positions p (N_s x 3J, sample-major) of descriptor x (SetStates' p output, core_rbdl.cpp:150-153). */
omo_status omo_positions(omo_ctx* ctx, const double* x, double* p_out);

/* Next row N2/N3: the state of descriptor x at M arbitrary times, as Core::CalculateTorques /
 * Core::SetCurrentState compute it one t at a time on the UI thread (core_rbdl.cpp:168-198, 264-289): Euler angles at
 * t - dt, t, t + dt (curves and IK, joint.cpp:36-52), 3-point central differences (core_rbdl.cpp:97-106), one RBDL
 * InverseDynamics per time. All M times are one launch of the per-sample-output kernel over a grid of (t - dt, t,
 * t + dt) triples. q, qd, qdd, tau: [M][dof] (q[0:3] = root offset as in SetCurrentState); root_euler: [M][3], the
 * root joint's own Joint::GetEulerAngle(t) (what Core::ExportMotion writes for it, core_io.cpp:252-294); p: [M][3J]
 * joint world positions. Any output may be NULL. The objective window set by omo_set_window is left as it was. */
omo_status omo_eval_states_at(omo_ctx* ctx, const double* x, const double* times, int M, double dt, double* q, double* qd,
                              double* qdd, double* tau, double* root_euler, double* p);

/* NEW (no reference counterpart): central-difference gradient on the free parameters,
 *   g_k = (f(x + h e_idx[k]) - f(x - h e_idx[k])) / (2h),
 * evaluated as ONE batch of 2F+1 columns built on the device from x. f0_out, fpm_out ([F][2] = f+, f-) optional. */
omo_status omo_eval_grad(omo_ctx* ctx, const double* x, const int* free_idx, int F, double h, double* f0_out,
                         double* grad_out, double* fpm_out);

/* NEW, next row N4: the ANALYTIC gradient of f on the free parameters -- forward-mode derivatives of the RNEA
 * (Dynamics.cc:26-105) restricted to what each keyed channel reaches, the transposed stencils (core_rbdl.cpp:28-95), the Bezier
 * value's partials with the implicit derivative of its Newton root (curve.cpp:43-93), D_param and D_traj -- from ONE evaluation of
 * the per-sample arrays instead of 2F+1 objective columns. Agrees with the central-difference gradient of omo_eval_grad to its
 * truncation error (~1e-7 relative at h = 1e-4). f0_out optional. OMO_UNSUPPORTED for scenes with IK handles and for an x whose
 * key times are not ascending within a curve: call omo_eval_grad there. Large problems take the REVERSE mode of the same chain -- one
 * forward sweep, one force fold-back and their hand-written adjoints per sample, every joint channel's partials at once (64 joints x
 * 2000 frames, 113 967 parameters: 35 -> 4 ms); small ones keep the tangents (OMO_ADJ_MODE=forward / reverse forces either). */
omo_status omo_eval_grad_analytic(omo_ctx* ctx, const double* x, const int* free_idx, int F, double* f0_out, double* grad_out);
/* Host-only (no GPU): the flattened device scene table of `scene` (optimo_b200/csrc/omo_scene.cuh: OmoDevScene, cap >= its size) and
 * each joint's subtree end. For the CPU check of the analytic gradient's arithmetic (tests/native); not a compute path. */
omo_status omo_debug_scene_table(const omo_scene_desc* scene, void* out, unsigned long long cap, int* sub_end_out);

/* Column-sharded form for multi-GPU: evaluates columns [col_begin, col_begin+col_count) of the (2F+1)-column
 * perturbation list (column 0 = x, column 2k+1 = x + h e_k, column 2k+2 = x - h e_k) and leaves the values in
 * DEVICE memory d_f_cols[col_count] (gather them with NCCL, then call omo_grad_from_columns). */
omo_status omo_eval_grad_columns(omo_ctx* ctx, const double* x, const int* free_idx, int F, double h, int col_begin,
                                 int col_count, double* d_f_cols, int sync);
/* grad_k = (cols[2k+1] - cols[2k+2]) / (2h) on the host, from the gathered 2F+1 column values. */
/* Work actually executed by the last omo_eval_grad / omo_eval_grad_columns call: a finite-difference column moves one control
 * point, so only the (column, tile) work items whose samples can see it are evaluated and the rest of each sum is taken from
 * the base point (bit-identical to evaluating every column on every tile; OMO_GRAD_DENSE=1 forces that). out2[0] = work items
 * executed (tiles of 124 samples, base row included), out2[1] = columns x tiles of the dense evaluation. */
omo_status omo_grad_sparsity(const omo_ctx* ctx, unsigned long long* out2);
omo_status omo_grad_from_columns(const double* f_cols, int F, double h, double* f0_out, double* grad_out);

/* Contiguous shard [begin, begin+count) of n items for `rank` of `world` (first ranks take the remainder). */
void omo_shard_range(int n, int rank, int world, int* begin, int* count);

/* ---- multi-device engine: ONE host thread, several GPUs of one box ---------
 * The reference's callback runs on one thread of one process (NLopt's serial loop, OptiMo/core_opt.cpp:141-152; its only
 * parallel construct is the parallel_for of Core::SetStates, OptiMo/core_rbdl.cpp:157-161). An omo_multi lets that one
 * thread use every GPU: it owns one omo_ctx per device, a worker thread per device, the pinned gather buffer and the NCCL
 * communicators. Batches are sharded over candidates, gradients over their 2F+1 finite-difference columns
 * (omo_shard_range); scene, weights and x_original are replicated; the only exchange is the gather of one double per
 * candidate / column. Results are bit-identical to the single-device entry points of the same names.
 * device_ids == NULL: devices 0 .. n_dev-1; n_dev <= 0: every visible device. Not re-entrant (one caller at a time). */
typedef struct omo_multi omo_multi;
#define OMO_GATHER_HOST 0 /* every device copies its slice into the host array itself (default) */
#define OMO_GATHER_NCCL 1 /* ncclAllGather over NVLink / NVSwitch, then one device-to-host copy */
omo_status omo_create_multi(const omo_scene_desc* scene, const int* device_ids, int n_dev, omo_multi** out);
void omo_destroy_multi(omo_multi* m);
const char* omo_multi_last_error(const omo_multi* m); /* m may be NULL: error of the last failed omo_create_multi */
int omo_multi_num_devices(const omo_multi* m);
int omo_multi_device(const omo_multi* m, int i);      /* CUDA device index of slot i */
omo_ctx* omo_multi_ctx(omo_multi* m, int i);          /* the engine of slot i (queries only while no multi call runs) */
/* OMO_NCCL_ERROR when libnccl cannot be loaded or the communicators cannot be created. Env OMO_GATHER=nccl selects it at creation. */
omo_status omo_multi_set_gather(omo_multi* m, int mode);
int omo_multi_get_gather(const omo_multi* m);
omo_status omo_multi_set_window(omo_multi* m, double min_t, double max_t, double dt);
omo_status omo_multi_set_weights(omo_multi* m, const double* w_param, int D, double w_traj, const omo_gauss* kernels, int n_kernels, int cyclic);
omo_status omo_multi_set_original(omo_multi* m, const double* x_original, int D);
/* omo_eval / omo_eval_free with the B candidates split over the devices; every device writes its slice of f_out / terms_out. */
omo_status omo_multi_eval(omo_multi* m, const double* X, int B, double* f_out, omo_terms* terms_out);
omo_status omo_multi_eval_free(omo_multi* m, const double* x_base, const int* free_idx, int F, const double* Xfree, int B, double* f_out,
                               omo_terms* terms_out);
/* omo_eval_grad with the 2F+1 columns split over the devices (omo_eval_grad_columns per device + gather + omo_grad_from_columns). */
omo_status omo_multi_eval_grad(omo_multi* m, const double* x, const int* free_idx, int F, double h, double* f0_out, double* grad_out,
                               double* fpm_out);

/* ---- multi-PROCESS gather over NVLink / NVSwitch peer memory (no reference counterpart) ----------
 * One process per GPU (torchrun, MPI): every rank evaluates a slice of ONE batch -- the candidates of a population, the 2F+1
 * columns of a gradient (OptiMo/core_opt.cpp:42-57 consumes all values on one thread) -- and every rank needs all values. The
 * kernel that forms f (finish_kernel's sums, same order, same bits) stores each value straight into EVERY rank's gathered array
 * through cudaIpc-mapped peer memory, its last block publishes "rank r's slice of call k is complete" in every rank's flag
 * array and waits for the other ranks' flags: compute step and exchange are one launch, no NCCL in the data path (8 B200s,
 * one 4096-candidate batch: 0.163 -> 0.156 ms per step against finish_kernel + ncclAllGather, same bits). Calls alternate
 * between two halves of the array, so a rank may still read call k while a faster one delivers k + 1.
 *   1. every rank: omo_peer_gather_create(ctx, rank, world, capacity, &g, handle)      capacity = values of the whole batch
 *   2. exchange the OMO_IPC_HANDLE_BYTES-byte handles between the processes (any transport), in rank order
 *   3. every rank: omo_peer_gather_connect(g, handles)        OMO_NCCL_ERROR: no peer access / IPC between the devices -> use NCCL
 *   4. per batch, every rank (a rank with B_local == 0 too): omo_peer_gather_eval(g, dX, B_local, first, sync)
 *      then omo_peer_gather_values(g)[i] = f of candidate i of the whole batch, in this rank's device memory.
 * A rank whose peers do not arrive within ~2 s (a crashed process) gets OMO_NCCL_ERROR from the synchronous form, or a non-zero
 * omo_peer_gather_error(g) after its own stream synchronisation; the wait never hangs the GPU. */
#define OMO_IPC_HANDLE_BYTES 64
typedef struct omo_peer_gather omo_peer_gather;
omo_status omo_peer_gather_create(omo_ctx* ctx, int rank, int world, int capacity, omo_peer_gather** out, unsigned char* handle_out);
omo_status omo_peer_gather_connect(omo_peer_gather* g, const unsigned char* handles /* [world][OMO_IPC_HANDLE_BYTES] */);
omo_status omo_peer_gather_eval(omo_peer_gather* g, const double* dX, int B_local, int first, int sync);
const double* omo_peer_gather_values(const omo_peer_gather* g);   /* device pointer, [capacity], of the last call */
int omo_peer_gather_error(omo_peer_gather* g);                     /* 1 once after a wait that timed out */
void omo_peer_gather_destroy(omo_peer_gather* g);

/* ---- device helpers for harnesses (no reference counterpart) ---------------------------------------- */
omo_status omo_device_alloc(omo_ctx* ctx, unsigned long long bytes, void** out);
omo_status omo_device_free(omo_ctx* ctx, void* p);
omo_status omo_host_alloc_pinned(omo_ctx* ctx, unsigned long long bytes, void** out);
omo_status omo_host_free_pinned(omo_ctx* ctx, void* p);
omo_status omo_memcpy_h2d(omo_ctx* ctx, void* dst, const void* src, unsigned long long bytes);
omo_status omo_memcpy_d2h(omo_ctx* ctx, void* dst, const void* src, unsigned long long bytes);
omo_status omo_synchronize(omo_ctx* ctx);
void* omo_stream(omo_ctx* ctx); /* cudaStream_t the hot path launches on */

/* Counters of the last batch (summed over its sample-evals): out[0] Bezier evaluations, out[1] Newton
 * iterations, out[2] IK solves, out[3] IK iterations, out[4] kernels launched since omo_create. Counting is
 * done only when `enable` was set by omo_set_counting(ctx, 1) (it costs a few atomics per block). */
omo_status omo_set_counting(omo_ctx* ctx, int enable);
omo_status omo_get_counters(omo_ctx* ctx, unsigned long long* out5);

/* Single-candidate calls (omo_eval with B = 1 -- what the NLopt callback EvaluateObjective, OptiMo/core_opt.cpp:42-57, makes
 * thousands of times per solve): the copy-in / prepare / objective / finish / copy-out chain is captured once as a CUDA
 * graph and replayed with one launch for as long as window, buffers and launch geometry stay what they were (weights,
 * x_original and kernels may change between calls). OMO_GRAPH=0 switches it off. Returns the number of replays so far. */
unsigned long long omo_graph_replays(const omo_ctx* ctx);

/* Timing of the main objective kernel: accumulates CUDA-event time of every launch between reset and read. */
omo_status omo_kernel_timing(omo_ctx* ctx, int reset, double* total_ms, int* launches);

/* Scene-specialised kernels (optimo_b200/csrc/omo_jit.hpp): for launches of >= 64 K sample-evals the engine unrolls
 * the tree walk of its scene into straight-line CUDA (the same device code as the table-driven kernel), compiles it
 * for sm_100a with NVRTC and launches that. OMO_JIT=0 disables it, OMO_JIT=1 makes it mandatory for every launch.
 * omo_jit_info: out4[0] launches that went through a specialised kernel, [1] instances compiled, [2] their largest
 * register count, [3] total compile time in ms; log = why specialisation was not possible, if it was not.
 * omo_jit_probe (host only, needs no GPU): generates and compiles the specialised kernel of `scene`; writes
 * <dump_prefix>.cu / .cubin when dump_prefix is not NULL. count: bit 0 = the work-counting instance, bit 1 = the
 * self-preparing instance (single candidates / small gradients in one launch). The cubin also lands in the on-disk cache. */
omo_status omo_jit_info(omo_ctx* ctx, unsigned long long* out4, char* log, int log_cap);
omo_status omo_jit_probe(const omo_scene_desc* scene, int fk, int count, const char* dump_prefix, char* log, int log_cap);
/* JSON list of the specialised instances this context has built: form (unrolled / loop), registers, shared memory, build time
 * (NVRTC, or the load from the on-disk cubin cache: $OMO_JIT_CACHE_DIR, else <dir of libomo.so>/jit_cache; OMO_JIT_CACHE=0 disables it). */
omo_status omo_jit_describe(omo_ctx* ctx, char* out, int cap);

/* Device self-test of the hand-scheduled FP64 division used by the curve solver (omo_kernels.cuh: div_rn_fast;
 * it replaces the `f / f'` of SolveCubic, OptiMo/curve.cpp:50): q_fast[i] = the branch-free sequence, ok[i] = its
 * validity test, q_ref[i] = the compiler's IEEE division. Wherever ok[i] != 0, q_fast[i] must equal q_ref[i] bit for bit. */
omo_status omo_selftest_division(omo_ctx* ctx, const double* num, const double* den, int n, double* q_fast, double* q_ref, int* ok);

/* Sustained dependent-free DFMA throughput of the device in FLOP/s (roofline denominator; runs ~ms). */
omo_status omo_measure_dfma_peak(omo_ctx* ctx, double* flops_per_s);

#ifdef __cplusplus
}
#endif
#endif /* OMO_H */
