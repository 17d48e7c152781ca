// Device-side scene tables and kernel parameter blocks of libomo (sm_100a).
//
// The joint tree is stored topology-sorted (DFS pre-order == RBDL body order, OptiMo/object.cpp:27-40,
// OptiMo/core_rbdl.cpp:214-256) in CONSTANT memory so that every per-body quantity is a uniform operand of
// the FP64 pipe. One scene is active per device at a time (omo.cu re-uploads the table when the active
// context changes).
#ifndef OMO_SCENE_CUH
#define OMO_SCENE_CUH

#ifdef __CUDACC_RTC__
// NVRTC build (omo_jit.hpp): only the limits of include/omo.h are needed; omo.cu static_asserts that they agree.
#define OMO_MAX_JOINTS 128
#define OMO_MAX_IK 8
#define OMO_MAX_IK_CHAIN 8
#else
#include "../../include/omo.h"
#endif

#define OMO_MAXJ OMO_MAX_JOINTS
#define OMO_MAX_PEERS 16   // ranks of one omo_peer_gather (GPUs of one NVLink / NVSwitch domain)
#define OMO_NT 128   // threads per block (compile-time: every stack / exchange access gets an immediate offset)
// OMO_SUB (scene-specialised kernels): (candidate, tile) work items fused into one thread block, each on its own
// OMO_NT threads and its own slice of dynamic shared memory; they only share the block barriers -- and therefore
// march through the straight-line code together, so that one instruction-cache line serves all of them.
#ifndef OMO_SUB
#define OMO_SUB 1
#endif
#define OMO_DEEP_SLOTS 128  // per-thread (local-memory) stack slots of the deep-tree instance of the table-driven kernel
#define OMO_SLOT 12  // doubles per stack slot: (f, sincos) or (v, a) or (L, t)
// Prepared Bezier segments (prepare_kernel -> objective kernels): one record of OMO_SEGD doubles per chain position
// `pos` that closes a segment (pos-1 -> pos of the same channel):
//   0: p0 = t_start   1: te = t_end   2: a   3: b   4: c   5: 3a   (time cubic, curve.cpp:85-88)   6..9: v0 v1 v2 v3
// The scene-specialised kernel stages, per keyed channel, the OMO_SEGW records starting at the segment that holds the
// block's first sample into shared memory (cp.async) and solves from there.
#define OMO_SEGD 10
#define OMO_SEGW 4

// Per-body constants: the RBDL quantities InverseDynamics reads (Model::X_T[i].r, Model::I[i];
// external/rbdl-v2.5.0/src/Model.cc:258-464, include/rbdl/SpatialAlgebraOperators.h:116-128).
struct OmoBody {
    double r[3];                         // X_T.r == parent joint's offset (core_rbdl.cpp:228-252)
    double m;                            // mass
    double h[3];                         // m * com
    double Ixx, Iyx, Iyy, Izx, Izy, Izz; // rotational inertia about the joint frame origin
};

struct OmoDevScene {
    int J, n_ik;
    int n_slots;  // stack slots per thread (each OMO_SLOT doubles)
    int n_groups;   // curve groups: keyed joint channels are sampled group by group, one block barrier per group
    int n_ex_rows;  // rows of the stencil exchange buffer (2 * group size when there are several groups)
    int pad0;
    int parent[OMO_MAXJ];
    short fs_slot[OMO_MAXJ];  // slot of (f, sincos) while the body's subtree is open; -1 for leaves and the root
    short va_slot[OMO_MAXJ];  // slot of (v, a) for non-root bodies with >= 2 children; -1 otherwise
    short fk_slot[OMO_MAXJ];  // slot of the world affine (L, t) for bodies with >= 2 children; -1 otherwise
    short ik_of[OMO_MAXJ];    // IK handler driving this joint's angles (joint.cpp:38-44), -1 if none
    short ik_idx[OMO_MAXJ];   // position of the joint inside that handler's chain
    short ik_start[OMO_MAXJ]; // IK handler whose chain root is this joint, -1 if none
    unsigned char n_child[OMO_MAXJ];
    unsigned char dyn[OMO_MAXJ];  // bit a: channel 3j+a is keyframed
    unsigned char ident[OMO_MAXJ];  // 1: all three angles are the constant 0 and the joint is not IK-driven (E = I, qd = qdd = 0)
    // Packed per-body walk metadata, one 16-byte uniform load per body (index 0: full tree, 1: pruned tree):
    //   x = parent | fs_slot << 8 | va_slot << 16 | fk_slot << 24                  (0xFF = none)
    //   y = dyn (bits 0-2) | ident (3) | has_child (4) | (grp_first + 1) << 8 | (ik_of + 1) << 16 | ik_idx << 20 | (ik_start + 1) << 24
    //   z = chan_row[3b] | chan_row[3b+1] << 8 | chan_row[3b+2] << 16                (0xFF = none)
    // The pruned tree leaves out every body whose whole subtree AND whole ancestor chain are constant (no keyed
    // channel, no IK): their torques do not depend on the candidate (see omo.cu, static_tsq).
    int4 meta[2][OMO_MAXJ];
    unsigned char act[2][OMO_MAXJ];   // bodies >= 1 visited by the walk, DFS order
    int n_act[2];
    int n_slots_tree[2];        // stack slots with the FK slots
    int n_slots_nofk_tree[2];   // and without
    int n_keyed;                      // entries of grp_list
    int pad1;
    double static_tsq;                // sum of tau^2 of the pruned bodies at an interior sample (0 at boundary samples)
    short grp_first[OMO_MAXJ];        // group that starts at this body, -1 if none
    short grp_begin[OMO_MAXJ + 1];    // group g owns grp_list[grp_begin[g] .. grp_begin[g+1])
    short grp_list[3 * OMO_MAXJ];     // keyed joint channels (3*j + axis, j >= 1) in body order
    short chan_row[3 * OMO_MAXJ];     // exchange row of a keyed joint channel (parity offset included), -1 otherwise
    // joint channels (3 per joint). pos = first position of the channel in the chain-ordered prepared arrays.
    int chan_pos[3 * OMO_MAXJ];
    int chan_n[3 * OMO_MAXJ];
    double chan_const[3 * OMO_MAXJ];                      // Variable::var_
    double chan_sin[3 * OMO_MAXJ], chan_cos[3 * OMO_MAXJ]; // of chan_const (host libm)
    double chan_qd[3 * OMO_MAXJ], chan_qdd[3 * OMO_MAXJ];  // interior-stencil rounding residue of a constant (SURVEY §7 ix)
    double root_qd[3], root_qdd[3];                        // same for q[0:3] = root offset (core_rbdl.cpp:137-141)
    OmoBody body[OMO_MAXJ];
    double offset[OMO_MAXJ][3];
    // IK handlers (ikhandler.cpp:77-188)
    int ik_nb[OMO_MAX_IK];                           // rotating joints in the chain
    int ik_joint[OMO_MAX_IK][OMO_MAX_IK_CHAIN + 1];  // chain joints root..end effector
    int ik_chan_pos[OMO_MAX_IK][3];
    int ik_chan_n[OMO_MAX_IK][3];
    double ik_chan_const[OMO_MAX_IK][3];
};

struct OmoEvalParams {
    const double* prep;     // [B][6][n_cp] prepared control points, chain order (see prepare_kernel)
    const unsigned* flags;  // [B] bit 0: every channel's chain is sorted by time
    const double* tgrid;    // [N_s]
    const double* wcost;    // [N_s] time-varying cost weight (core.cpp:191-199)
    const double* p_orig;   // [3J][N_s] joint positions of the original motion
    double* partial;        // [B][n_tiles][2]: (sum w*local_cost, sum w_traj*|p-p0|^2*dt)
    const double* seg;      // [B][n_cp][OMO_SEGD] prepared segments, chain order
    const unsigned short* tile_seg;  // [B][n_tiles][n_keyed] number of keys with time <= the tile's first sample time
    int N_s, n_tiles, T_out, n_cp;
    const int* item_map;  // null: work item i is (candidate, tile) number i; else item_map[i] = candidate * n_tiles + tile
    int n_items;      // work items of this launch (one per OMO_NT threads); B * n_tiles without an item_map
    int smem_stride;  // doubles of dynamic shared memory per work item (used when OMO_SUB > 1)
    double dt, inv_dt, inv_2dt, inv_12dt, inv_dt2, inv_2dt2, inv_12dt2;
    double w_traj;
    // optional per-sample outputs of candidate 0 (OptimizationCache, core.h:120-152); sample-major
    double *o_q, *o_qd, *o_qdd, *o_tau, *o_local, *o_p;
    unsigned long long* counters;  // [4] or null
    // OUT instance only (next row N2, Core::CalculateTorques / SetCurrentState, core_rbdl.cpp:168-198,264-289):
    // triplet != 0: the grid is a list of (t - dt, t, t + dt) triples and every sample takes the 3-point central forms
    // (core_rbdl.cpp:97-106); only the middle sample of each triple is meaningful. o_root: [N_s][3] Euler angles of the
    // root joint's own rot channels (Joint::GetEulerAngle of the root, joint.cpp:36-52; q[0:3] holds its offset instead).
    int triplet;
    int n_smem_slots;  // deep-tree instance of the table-driven kernel: stack slots held in shared memory, the rest per thread
    double* o_root;
    // ---- sorted curve work lists (scene-specialised batch kernels) ------------------------------------------------------------
    // perm[tile][perm_stride]: for every curve group, its (local channel << 7 | sample slot) items of the tile's 128-sample window
    // sorted by the Newton iteration count the ORIGINAL motion needs there (descending; 0xFFFF pads a group to whole triples).
    // A warp's lockstep solve then runs curves of similar cost together instead of waiting for its slowest lane
    // (omo_walk.cuh: curve_group). null: every lane solves the curves of its own sample.
    const unsigned short* perm;
    int perm_stride;
    // ---- self-preparing instances (Walker SELF; single candidates and small finite-difference gradients) --------------------
    // One launch does what the batch path spreads over perturb_kernel -> prepare_kernel -> objective -> finish_kernel and two
    // copies: the base descriptor x arrives as a KERNEL PARAMETER (constant bank), every block prepares its row's control
    // points in shared memory itself (row r = x, or x with entry self_free[(col-1)>>1] moved by +-h for column col =
    // self_col_begin + r > 0), the block that finishes last adds the tile partials in order (finish_sparse_kernel's rule:
    // tiles outside [lo, hi] of a row come from self_base_row) and writes f straight into mapped host memory.
    const int* self_free;       // [F] free descriptor indices (gradient rows), null for plain candidates
    const int* self_lo;         // [rows] first / last tile evaluated for the row (null: every tile)
    const int* self_hi;
    const int *self_cp_of_pos, *self_first_of_pos, *self_last_of_pos, *self_next_same;   // prepare tables (see OmoPrepParams)
    const double *self_x0, *self_wparam;   // [D]
    double* self_dparam;        // [rows] scratch
    double* self_out;           // [rows] f (rows == 1: f, C, D_param, D_traj), device-visible (mapped host or device memory)
    unsigned* self_counter;     // blocks finished; the last one resets it
    double self_h;
    int self_col_begin, self_rows, self_base_row, self_cyclic, self_D;
};

#endif
