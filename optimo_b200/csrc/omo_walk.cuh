// The per-thread tree walk of the objective kernels, as a set of force-inlined steps.
//
// One THREAD = one (candidate, time sample); one BLOCK = (candidate, tile of consecutive samples + 2-sample halo
// each side). Two kernels drive these steps:
//   * objective_kernel (omo_objective.cuh): a loop over the scene tables in constant memory -- works for any scene,
//     compiled ahead of time by nvcc;
//   * the scene-specialised kernel (omo_jit.cpp): the SAME steps emitted as straight-line code with every body index,
//     stack slot, channel row and flag a literal, compiled for sm_100a at run time with NVRTC. Every table lookup,
//     branch and address computation of the loop disappears and the body constants become constant-bank operands of
//     the DFMAs.
//
// Restates, per (candidate, time sample):
//   Core::SetStates           OptiMo/core_rbdl.cpp:108-166   (q from curves / IK, p from FK)
//   CalculateFirst/SecondCentral  core_rbdl.cpp:28-95        (SECOND_ORDER_ACCURATE off)
//   Joint::GetAffineRelativeTo{Parent,World}  OptiMo/joint.cpp:64-83   (T(offset) Rz Ry Rx)
//   Core::CalculateCost       core_rbdl.cpp:307-346          (local_cost = |tau[3:]|^2 dt, weight w(t))
//   D_traj                    core_rbdl.cpp:387-398
// Template flags: FK = also run forward kinematics (positions; for D_traj and IK), IK = scene has IK chains,
// OUT = write the per-sample OptimizationCache arrays (single-candidate launches; walks the FULL tree, every other
// instance walks the pruned tree, see OmoDevScene::meta), COUNT = book executed curve / IK work.
#ifndef OMO_WALK_CUH
#define OMO_WALK_CUH

#include "omo_kernels.cuh"
#include "omo_rnea_z.cuh"

// OMO_PLAN = 1: staged curves whose lanes work on their own samples (no sorted work lists) find their segment through per-channel
// thresholds in sample-index space (eval_channels_planned); 0: through five FP64 compares per curve and sample (eval_channels_staged).
// Same results. The generator sets it per scene (omo_jit.hpp: plan_enabled) and reserves the plan area accordingly.
#ifndef OMO_PLAN
#define OMO_PLAN 0
#endif
#ifndef OMO_SINCOS_CALL
#define OMO_SINCOS_CALL 0   // 1: a body with three keyed angles calls sincos_angle3 (out of line); 2: every keyed angle calls sincos_angle1
#endif
#define OMO_PLAN_PAD 16   // plan area rounded up to 16 words = 128 bytes (omo_jit.hpp: place_slots reserves the same)

namespace omo {

// NREG > 0 (scene-specialised kernels only): stack slots numbered 256 + r live in the per-thread array rs[r][] instead of
// shared memory; with literal slot numbers ptxas keeps them in registers or spills them to (L1-cached) local memory.
// That is how deep trees keep several blocks per SM: only the most-used slots take shared memory.
// SEGG > 0 (scene-specialised kernels only): keyed curves are solved from segment records staged in shared memory
// (stage_issue / stage_wait / curve_staged); SEGG = channels per curve group = records rows of one staging buffer.
// CPT (scene-specialised kernels only): compact stack. A slot argument is then the slot's first ROW of the shared-memory
// stack (slots are as wide as what is parked in them, 6 to 12 rows), or 4096 + r for per-thread slot r.
// DEEP (table-driven kernel only): trees whose stack does not fit shared memory (a 128-joint chain parks 127 bodies). Slots
// below `nsm` live in shared memory as usual, slot nsm + r is the per-thread rs[r] (run-time indexed = local memory, which
// the driver provisions per RESIDENT thread -- the pooling a global scratch buffer would have to re-invent).
// SELF (scene-specialised kernels only; needs SEGG > 0, OMO_SUB == 1): the block prepares its row's control points itself, in
// shared memory, from the base descriptor handed over as a kernel parameter; the last block to finish reduces the tile
// partials and writes f (OmoEvalParams::self_*). One launch per call: what a single candidate (every BOBYQA callback,
// core_opt.cpp:145) and the 109-157-column gradients of the shipped scenes need, where launches and copies, not FLOPs, set the time.
template <bool FK, bool IK, bool OUT, bool COUNT, int NREG = 0, int SEGG = 0, bool CPT = false, bool DEEP = false, bool SELF = false>
struct Walker {
    static constexpr int NT = OMO_NT;
    static constexpr int TREE = OUT ? 0 : 1;   // 0: every body, 1: candidate-independent subtrees pruned

    const OmoEvalParams& P;
    double* __restrict__ ex;    // [ex_rows][NT] stencil exchange: q of the dynamic channels
    double* __restrict__ stk;   // [n_slots][OMO_SLOT][NT], already offset by tid
    unsigned short* __restrict__ seg0;  // this warp's segment warm starts, [n_keyed]
    const double* __restrict__ pc;      // this candidate's prepared control points
    int tid, N, tile_end;
    int item;                   // (candidate, tile) work item of this thread's OMO_NT-thread group
    bool item_valid;            // false for the padding groups of the last block (OMO_SUB > 1)
    double* sm0;                // this work item's slice of dynamic shared memory
    size_t dof;
    bool sorted;
    unsigned n_bez, n_newton, n_iksolve, n_ikiter;
    int xbuf;
    int nsm;                    // DEEP: number of stack slots in shared memory

    // state of the body visited last (v, a: also the parent state for the next one)
    SV6 v, a, f;
    double s0, c0, s1, c1, s2, c2;        // sin / cos of its Euler angles
    double L[FK ? 9 : 1], W[FK ? 3 : 1];  // world affine handed to the next body's FK step
    double tsq, dtr;                      // sum tau^2 (non-root DoFs), sum |p - p_original|^2
    double t;                             // time of this sample
    int s;                                // sample index (clamped into [0, N_s))
    int idx;                              // position inside the block's window (exchange-buffer column)
    bool owned, interior;
    SV6 g;                                // force being folded back towards the root
    SV6 groot;                            // OUT only: force folded into the root
    double ikq[IK ? OMO_MAX_IK * 3 * OMO_MAX_IK_CHAIN : 1];
    double rs[NREG > 0 ? NREG : 1][OMO_SLOT];
    const double* __restrict__ sgc;   // SEGG > 0: this candidate's segment records
    double* wsb;                      //           staging buffers [buf][SEGG][OMO_SEGW][OMO_SEGD]
    int* wbase;                       //           [n_keyed] chain position of each keyed channel's first staged record
    int *wcn, *wrow;                  //           [n_keyed] control points / exchange row of each keyed channel (sorted work lists: per-lane lookups)
    double* wts;                      //           [NT] the sample times of the block's window
    unsigned long long* wplan;        //           [warp][SEGG] segment-search plan of the group being solved (eval_channels_planned)
    int perm_cursor;                  //           entries of the tile's sorted list consumed by the groups solved so far
    int perm_staged;                  //           ... and staged so far (the list of a group travels with its segment records)
    unsigned short* wpm;              //           staged slices of the sorted list, [buf][ceil(SEGG / 3) * 3 * NT]
    const unsigned short* __restrict__ tsg;   //   this (candidate, tile)'s row of the tile table
    const double* xs;                 // SELF: the base descriptor (kernel parameter, constant bank)
    double* seg_sm;                   //       this row's segment records [n_cp + OMO_SEGW][OMO_SEGD] in shared memory
    int cand_, tile_;
    double* self_red;                 //       D_param's per-warp partial sums (static shared memory), added up in finish()
    bool self_books_dparam;

    __device__ __forceinline__ Walker(const OmoEvalParams& P_) : P(P_) {}

    // stack element c of slot sl
    __device__ __forceinline__ double ld(const int sl, const int c) const
    {
        if (DEEP) { if (sl >= nsm) return rs[sl - nsm][c]; }
        else if (NREG > 0 && sl >= (CPT ? 4096 : 256)) return rs[sl - (CPT ? 4096 : 256)][c];
        return stk[(size_t)((CPT ? sl : sl * OMO_SLOT) + c) * NT];
    }
    __device__ __forceinline__ void st(const int sl, const int c, const double x)
    {
        if (DEEP && sl >= nsm) rs[sl - nsm][c] = x;
        else if (!DEEP && NREG > 0 && sl >= (CPT ? 4096 : 256)) rs[sl - (CPT ? 4096 : 256)][c] = x;
        else stk[(size_t)((CPT ? sl : sl * OMO_SLOT) + c) * NT] = x;
    }

    // root state (RBDL body 1, TranslationXYZ, q = root offset: core_rbdl.cpp:137-141,212): v = (0, qd), a = (0, qdd)
    __device__ __forceinline__ void set_root()
    {
        const OmoDevScene& S = c_scene;
        v.w0 = v.w1 = v.w2 = 0.0; a.w0 = a.w1 = a.w2 = 0.0;
        v.l0 = interior ? S.root_qd[0] : 0.0; v.l1 = interior ? S.root_qd[1] : 0.0; v.l2 = interior ? S.root_qd[2] : 0.0;
        a.l0 = interior ? S.root_qdd[0] : 0.0; a.l1 = interior ? S.root_qdd[1] : 0.0; a.l2 = interior ? S.root_qdd[2] : 0.0;
    }

    // ---- block / thread set-up -------------------------------------------------------------------------------------
    // n_smem_slots: stack slots that live in shared memory (the scene's plan, or the specialised kernel's own); with CPT
    // it is the number of ROWS of the shared-memory stack
    __device__ __forceinline__ void begin(double* smem, const int n_smem_slots, const bool stage_in_stack = false)
    {
        const OmoDevScene& S = c_scene;
        if (OMO_SUB > 1) {
            const int sub = threadIdx.x / NT;
            tid = threadIdx.x - sub * NT;
            item = blockIdx.x * OMO_SUB + sub;
            item_valid = item < P.n_items;
            item = min(item, P.n_items - 1);
            smem += (size_t)sub * P.smem_stride;
        } else {
            tid = threadIdx.x; item = blockIdx.x; item_valid = true;
        }
        sm0 = smem;
        nsm = n_smem_slots;
        if (P.item_map) item = __ldg(P.item_map + item);   // sparse launches (finite-difference columns): listed (candidate, tile) pairs only
        const int cand = item / P.n_tiles;
        const int tile = item - cand * P.n_tiles;
        cand_ = cand; tile_ = tile;
        const int ex_rows = S.n_ex_rows + (IK ? 6 : 0);
        ex = smem;
        double* stk0 = smem + ex_rows * NT;
        stk = stk0 + tid;
        N = P.N_s;
        tile_end = min(N, (tile + 1) * P.T_out);
        dof = (size_t)3 * S.J;
        n_bez = n_newton = n_iksolve = n_ikiter = 0u;
        xbuf = 0;
        if (SELF) {
            seg0 = nullptr; sgc = nullptr; tsg = nullptr; wsb = nullptr;
            double* tail = stk0 + (size_t)n_smem_slots * ((CPT ? 1 : OMO_SLOT) * NT);
            const int wb_doubles = ((S.n_keyed * (int)sizeof(int) + 15) & ~15) / (int)sizeof(double);
            wbase = reinterpret_cast<int*>(tail);
            double* prep_sm = tail + wb_doubles;
            seg_sm = prep_sm + 6 * P.n_cp;
            pc = prep_sm;
            __shared__ double self_red_store[NT / 32];
            self_red = self_red_store;
            self_prepare(prep_sm, tile);
        } else {
        pc = P.prep + (size_t)cand * 6 * P.n_cp;
        sorted = (__ldg(P.flags + cand) & 1u) != 0u;
        }
        if (SELF) {
        } else if (SEGG > 0) {
            // staged curves: wbase[] and the staging buffers sit behind the stack -- or, when the only curve group opens
            // the walk (stage_in_stack), in the stack area itself, which no body has touched yet
            seg0 = nullptr;
            sgc = P.seg + (size_t)cand * P.n_cp * OMO_SEGD;
            double* tail = stk0 + (size_t)n_smem_slots * ((CPT ? 1 : OMO_SLOT) * NT);
            // [wbase | wcn | wrow : n_keyed ints each][wts : NT doubles][wplan : SEGG words per warp][staging buffers]
            constexpr int kPlanWords = OMO_PLAN != 0 ? (((NT / 32) * SEGG + OMO_PLAN_PAD - 1) / OMO_PLAN_PAD) * OMO_PLAN_PAD : 0;   // whole 128-byte lines
            const int wb_doubles = ((3 * S.n_keyed * (int)sizeof(int) + 15) & ~15) / (int)sizeof(double) + NT + kPlanWords;
            double* const area = stage_in_stack ? stk0 : tail;
            wbase = reinterpret_cast<int*>(area); wcn = wbase + S.n_keyed; wrow = wcn + S.n_keyed;
            wts = area + (wb_doubles - NT - kPlanWords);
            wplan = reinterpret_cast<unsigned long long*>(wts + NT);
            wsb = area + wb_doubles;
            wpm = reinterpret_cast<unsigned short*>(wsb + (size_t)(S.n_groups > 1 ? 2 : 1) * SEGG * OMO_SEGW * OMO_SEGD);
            perm_cursor = 0; perm_staged = 0;
            // wbase[] is published by the barrier that follows the first stage_issue / stage_wait; that first stage reads
            // the tile table itself, so the block start is two dependent global round trips deep, not three
            tsg = P.tile_seg + ((size_t)cand * P.n_tiles + tile) * S.n_keyed;
            if (sorted) {
#pragma unroll 1
                for (int i = tid; i < S.n_keyed; i += NT) {
                    wbase[i] = stage_base(i);
                    const int ch = S.grp_list[i];
                    wcn[i] = S.chan_n[ch]; wrow[i] = S.chan_row[ch];
                }
            }
        } else {
        // seg0[warp][n_keyed] (16-bit) sits behind the stack in dynamic shared memory
        seg0 = reinterpret_cast<unsigned short*>(stk0 + (size_t)n_smem_slots * (OMO_SLOT * NT)) + (tid >> 5) * S.n_keyed;

        // segment warm start: one binary search per keyed channel per WARP (for the time of its first lane; the lanes'
        // times ascend), so each thread later only walks the few keys its own time lies beyond
        if (sorted) {
            const double t_first = __ldg(P.tgrid + min(max(tile * P.T_out - 2 + (tid & ~31), 0), N - 1));
            for (int i = tid & 31; i < S.n_keyed; i += 32) {
                const int ch = S.grp_list[i];
                const double* __restrict__ T = pc + S.chan_pos[ch];
                int lo = 0, hi = S.chan_n[ch];
                while (lo < hi) {
                    const int mid = (lo + hi) >> 1;
                    if (__ldg(T + mid) <= t_first) lo = mid + 1; else hi = mid;
                }
                seg0[i] = (unsigned short)lo;
            }
        }
        __syncwarp();
        }

        idx = tid;
        const int s_raw = tile * P.T_out - 2 + idx;
        s = min(max(s_raw, 0), N - 1);
        owned = (idx >= 2) && (s_raw < tile_end) && (s_raw >= 0) && item_valid;
        interior = (s >= 2) && (s <= N - 3) && !(OUT && P.triplet != 0);
        t = __ldg(P.tgrid + s);
        if (SEGG > 0 && !SELF) wts[idx] = t;           // published by the barrier that follows the first stage
        tsq = (!OUT && interior) ? S.static_tsq : 0.0;   // the pruned bodies' share (candidate-independent)
        dtr = 0.0;
        set_root();
        f.w0 = f.w1 = f.w2 = f.l0 = f.l1 = f.l2 = 0.0;
        g = f;
        s0 = s1 = s2 = 0.0; c0 = c1 = c2 = 1.0;
        if (OUT) groot = f;
        if (OUT && owned) {
            for (int k = 0; k < 3; ++k) {
                P.o_q[s * dof + k] = S.offset[0][k];
                P.o_qd[s * dof + k] = interior ? S.root_qd[k] : 0.0;
                P.o_qdd[s * dof + k] = interior ? S.root_qdd[k] : 0.0;
            }
        }
        if (FK) {
            // root world affine = T(offset) Rz Ry Rx with the root's own rot channels (joint.cpp:64-74)
            double e[3];
#pragma unroll
            for (int k = 0; k < 3; ++k)
                e[k] = (S.dyn[0] >> k & 1) ? eval_channel<!SELF>(pc, P.n_cp, S.chan_pos[k], S.chan_n[k], t, sorted, n_bez, n_newton) : S.chan_const[k];
            if (OUT && owned && P.o_root) { P.o_root[s * 3] = e[0]; P.o_root[s * 3 + 1] = e[1]; P.o_root[s * 3 + 2] = e[2]; }
            double sx, cx, sy, cy, sz, cz;
            sincos_angle(e[0], &sx, &cx); sincos_angle(e[1], &sy, &cy); sincos_angle(e[2], &sz, &cz);
            L[0] = cz * cy; L[1] = cz * sy * sx - sz * cx; L[2] = cz * sy * cx + sz * sx;
            L[3] = sz * cy; L[4] = sz * sy * sx + cz * cx; L[5] = sz * sy * cx - cz * sx;
            L[6] = -sy;     L[7] = cy * sx;                L[8] = cy * cx;
            W[0] = S.offset[0][0]; W[1] = S.offset[0][1]; W[2] = S.offset[0][2];
            if (P.p_orig) {
                const double d0 = W[0] - __ldg(P.p_orig + (size_t)0 * N + s), d1 = W[1] - __ldg(P.p_orig + (size_t)1 * N + s),
                             d2 = W[2] - __ldg(P.p_orig + (size_t)2 * N + s);
                dtr += d0 * d0 + d1 * d1 + d2 * d2;
            }
            if (OUT && owned && P.o_p) { P.o_p[s * dof] = W[0]; P.o_p[s * dof + 1] = W[1]; P.o_p[s * dof + 2] = W[2]; }
        }
    }

    // ---- SELF: Object::SetDescriptor + MakeMotionCyclic + handle values + segment records + D_param of this block's row, in
    // shared memory (the arithmetic of perturb_kernel and prepare_kernel, operation for operation) ----------------------------
    __device__ __forceinline__ void self_prepare(double* __restrict__ prep_sm, const int tile)
    {
        // Latency is all that matters here (one wave of blocks, one warp per scheduler): every thread works from the descriptor in
        // the constant bank alone -- its own control point AND the one before it, so that the segment record needs no neighbour's
        // result -- and the block meets at ONE barrier. D_param's partial sums wait in shared memory until finish().
        const OmoDevScene& S = c_scene;
        const int n_cp = P.n_cp, D = P.self_D;
        __shared__ int unsorted;
        int kpert = -1;
        double step = 0.0;
        if (P.self_free && cand_ != P.self_base_row) {   // column col of the finite-difference list: col 0 = x, 2k+1 / 2k+2 = x +- h e_k
            const int col = P.self_col_begin + cand_;
            if (col > 0) { kpert = __ldg(P.self_free + ((col - 1) >> 1)); step = ((col - 1) & 1) ? -P.self_h : P.self_h; }
        }
        auto X = [&](const int i) { double v = xs[i]; if (i == kpert) v = v + step; return v; };
        // prepared point (t, v, t + t_f, t + t_b, v + t_f tan(theta), v + t_b tan(theta)) of chain position `pos`
        auto point = [&](const int pos, double (&o)[6]) {
            const int cp = __ldg(P.self_cp_of_pos + pos);
            const double t0 = X(5 * cp), v = X(5 * cp + 1), tb = X(5 * cp + 3), tf = X(5 * cp + 4);
            double th = X(5 * cp + 2);
            if (P.self_cyclic) {
                const int fi = __ldg(P.self_first_of_pos + pos), la = __ldg(P.self_last_of_pos + pos);
                if (cp == fi || cp == la) th = __dmul_rn(0.5, __dadd_rn(X(5 * fi + 2), X(5 * la + 2)));
            }
            const double tn = tan(th);
            o[0] = t0; o[1] = v; o[2] = __dadd_rn(t0, tf); o[3] = __dadd_rn(t0, tb);
            o[4] = __dadd_rn(v, __dmul_rn(tf, tn)); o[5] = __dadd_rn(v, __dmul_rn(tb, tn));
        };
        if (tid == 0) unsorted = 0;
        // the tile's first sample time and D_param's operands: issued before anything waits on them
        const double t_first = __ldg(P.tgrid + min(max(tile * P.T_out - 2, 0), N - 1));
        const int lo_tile = P.self_lo ? __ldg(P.self_lo + cand_) : 0;
        double acc = 0.0;
        if (tile == lo_tile) {   // D_param (core_rbdl.cpp:385), in prepare_kernel's order: strided partial sums, shuffle tree, the four warps in turn
            for (int k = tid; k < D; k += NT) {
                const double d = X(k) - __ldg(P.self_x0 + k);
                acc = __fma_rn(__dmul_rn(d, __ldg(P.self_wparam + k)), d, acc);
            }
        }
        bool bad = false;
        for (int pos = tid; pos < n_cp; pos += NT) {
            double cur[6];
            point(pos, cur);
#pragma unroll
            for (int k = 0; k < 6; ++k) prep_sm[k * n_cp + pos] = cur[k];
            if (__ldg(P.self_next_same + pos) && !(cur[0] <= X(5 * __ldg(P.self_cp_of_pos + pos + 1)))) bad = true;
            if (pos >= 1 && __ldg(P.self_next_same + pos - 1)) {
                double prv[6];
                point(pos - 1, prv);
                const double p0 = prv[0], p1 = prv[2], p2 = cur[3], p3 = cur[0];
                const double a = __dadd_rn(__dsub_rn(__dadd_rn(-p0, __dmul_rn(3.0, p1)), __dmul_rn(3.0, p2)), p3);
                double* __restrict__ r = seg_sm + (size_t)pos * OMO_SEGD;
                r[0] = p0; r[1] = p3; r[2] = a;
                r[3] = __dmul_rn(3.0, __dadd_rn(__dsub_rn(p0, __dmul_rn(2.0, p1)), p2));
                r[4] = __dadd_rn(__dmul_rn(-3.0, p0), __dmul_rn(3.0, p1));
                r[5] = __dmul_rn(3.0, a);
                r[6] = prv[1]; r[7] = prv[4]; r[8] = cur[5]; r[9] = cur[1];
            }
        }
        if (bad) unsorted = 1;
        // first staged record of every keyed channel for this tile (prepare_kernel's tile table + stage_base), searched on the
        // key times of the descriptor itself; the last warps take it, the first ones are busy with the points
        for (int i = NT - 1 - tid; i < S.n_keyed; i += NT) {
            const int ch = S.grp_list[i];
            const int* __restrict__ cpl = P.self_cp_of_pos + S.chan_pos[ch];
            const int n = S.chan_n[ch];
            int lo = 0, hi = n;
            while (lo < hi) {
                const int mid = (lo + hi) >> 1;
                if (X(5 * __ldg(cpl + mid)) <= t_first) lo = mid + 1; else hi = mid;
            }
            wbase[i] = (n >= 2) ? min(max(lo, 1), n - 1) : 1;
        }
        if (tile == lo_tile) {
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) acc += __shfl_down_sync(0xffffffffu, acc, o);
            if ((tid & 31) == 0) self_red[tid >> 5] = acc;
        }
        self_books_dparam = tile == lo_tile;
        __syncthreads();
        sorted = unsorted == 0;
    }

    // SELF: the block that finishes last adds the tile partials of every row in tile order (finish_kernel / finish_sparse_kernel)
    __device__ __forceinline__ void self_finish()
    {
        __shared__ int last_block;
        if (P.self_counter == nullptr) {
            // single candidate with the host as the reducer (omo.cu: launch_self): `partial` and `self_dparam` ARE slots of the mapped host
            // buffer, the host adds them in finish_kernel's order once they have arrived -- no ticket, no fence, no last block
            if (tid == 0 && self_books_dparam) {
                double sum = 0.0;
                for (int w = 0; w < NT / 32; ++w) sum += self_red[w];
                P.self_dparam[cand_] = sum;
            }
            return;
        }
        if (tid == 0) {
            if (self_books_dparam) {
                double sum = 0.0;
                for (int w = 0; w < NT / 32; ++w) sum += self_red[w];
                P.self_dparam[cand_] = sum;
            }
            __threadfence();
            const unsigned ticket = atomicAdd(P.self_counter, 1u);
            last_block = (ticket == (unsigned)P.n_items - 1u) ? 1 : 0;
        }
        __syncthreads();
        if (!last_block) return;
        __threadfence();
        const int rows_out = P.self_base_row >= 0 ? P.self_rows - 1 : P.self_rows;
        for (int r = tid; r < rows_out; r += NT) {
            const int l = P.self_lo ? __ldg(P.self_lo + r) : 0, h = P.self_hi ? __ldg(P.self_hi + r) : P.n_tiles - 1;
            double C = 0.0, Dt = 0.0;
            for (int tl = 0; tl < P.n_tiles; ++tl) {
                const int row = (tl >= l && tl <= h) ? r : P.self_base_row;
                C += __ldcg(P.partial + ((size_t)row * P.n_tiles + tl) * 2);
                Dt += __ldcg(P.partial + ((size_t)row * P.n_tiles + tl) * 2 + 1);
            }
            const double Dp = __ldcg(P.self_dparam + r);
            P.self_out[r] = C + Dp + Dt;
            if (P.self_rows == 1) { P.self_out[1] = C; P.self_out[2] = Dp; P.self_out[3] = Dt; }
        }
        if (tid == 0) *P.self_counter = 0u;
        // (no completion word and no system-scope fence behind the outputs: where they go to mapped host memory the host has filled the
        // slots with a sentinel and polls until every one of them has been overwritten -- omo.cu: wait_self)
    }

    // ---- stack traffic ---------------------------------------------------------------------------------------------
    __device__ __forceinline__ void park_fs(const int sl, const bool ident)   // (f, sincos) of a body whose subtree opens
    {
        st(sl, 0, f.w0); st(sl, 1, f.w1); st(sl, 2, f.w2); st(sl, 3, f.l0); st(sl, 4, f.l1); st(sl, 5, f.l2);
        if (!ident) { st(sl, 6, s0); st(sl, 7, c0); st(sl, 8, s1); st(sl, 9, c1); st(sl, 10, s2); st(sl, 11, c2); }
    }
    __device__ __forceinline__ void park_va(const int sl)   // (v, a) of a body with >= 2 children
    {
        st(sl, 0, v.w0); st(sl, 1, v.w1); st(sl, 2, v.w2); st(sl, 3, v.l0); st(sl, 4, v.l1); st(sl, 5, v.l2);
        st(sl, 6, a.w0); st(sl, 7, a.w1); st(sl, 8, a.w2); st(sl, 9, a.l0); st(sl, 10, a.l1); st(sl, 11, a.l2);
    }
    __device__ __forceinline__ void load_va(const int sl)
    {
        v.w0 = ld(sl, 0); v.w1 = ld(sl, 1); v.w2 = ld(sl, 2); v.l0 = ld(sl, 3); v.l1 = ld(sl, 4); v.l2 = ld(sl, 5);
        a.w0 = ld(sl, 6); a.w1 = ld(sl, 7); a.w2 = ld(sl, 8); a.l0 = ld(sl, 9); a.l1 = ld(sl, 10); a.l2 = ld(sl, 11);
    }
    __device__ __forceinline__ void park_fk(const int sl)   // world affine (L, W) of a body with >= 2 children
    {
        if (FK) {
#pragma unroll
            for (int c = 0; c < 9; ++c) st(sl, c, L[c]);
#pragma unroll
            for (int c = 0; c < 3; ++c) st(sl, 9 + c, W[c]);
        }
    }
    __device__ __forceinline__ void load_fk(const int sl)
    {
        if (FK) {
#pragma unroll
            for (int c = 0; c < 9; ++c) L[c] = ld(sl, c);
#pragma unroll
            for (int c = 0; c < 3; ++c) W[c] = ld(sl, 9 + c);
        }
    }

    // ---- backward pass ---------------------------------------------------------------------------------------------
    // tau = S^T fk of body k (books |tau|^2), g = X_lambda^T fk (Dynamics.cc:87-104)
    __device__ __forceinline__ void finalize(const int k, const bool ident, const SV6& fk, const double a0, const double b0, const double a1,
                                             const double b1, const double a2, const double b2)
    {
        const OmoDevScene& S = c_scene;
        double t0, t1, t2;
        const double r0 = S.body[k].r[0], r1 = S.body[k].r[1], r2 = S.body[k].r[2];
        SV6 gn;
        if (ident) rnea_backward_identity(r0, r1, r2, fk, gn, t0, t1, t2);
        else rnea_backward(r0, r1, r2, a0, b0, a1, b1, a2, b2, fk, gn, t0, t1, t2);
        g = gn;
        tsq = fma(t0, t0, fma(t1, t1, fma(t2, t2, tsq)));
        if (OUT && owned) { P.o_tau[s * dof + 3 * k] = t0; P.o_tau[s * dof + 3 * k + 1] = t1; P.o_tau[s * dof + 3 * k + 2] = t2; }
    }
    // the body just visited is a leaf of the walk: fold it
    __device__ __forceinline__ void fold_leaf(const int b, const bool ident) { finalize(b, ident, f, s0, c0, s1, c1, s2, c2); }
    // ancestor k (parked in slot sl) is finished: add the force coming up and fold it
    __device__ __forceinline__ void fold_up(const int k, const int sl, const bool ident)
    {
        SV6 fk;
        fk.w0 = ld(sl, 0) + g.w0; fk.w1 = ld(sl, 1) + g.w1; fk.w2 = ld(sl, 2) + g.w2;
        fk.l0 = ld(sl, 3) + g.l0; fk.l1 = ld(sl, 4) + g.l1; fk.l2 = ld(sl, 5) + g.l2;
        if (ident) finalize(k, true, fk, 0.0, 1.0, 0.0, 1.0, 0.0, 1.0);
        else finalize(k, false, fk, ld(sl, 6), ld(sl, 7), ld(sl, 8), ld(sl, 9), ld(sl, 10), ld(sl, 11));
    }
    // the walk continues below ancestor `stop` (slot sl): leave the force with it
    __device__ __forceinline__ void fold_into(const int sl)
    {
        st(sl, 0, ld(sl, 0) + g.w0); st(sl, 1, ld(sl, 1) + g.w1); st(sl, 2, ld(sl, 2) + g.w2);
        st(sl, 3, ld(sl, 3) + g.l0); st(sl, 4, ld(sl, 4) + g.l1); st(sl, 5, ld(sl, 5) + g.l2);
    }
    __device__ __forceinline__ void fold_root()
    {
        if (OUT) { groot.w0 += g.w0; groot.w1 += g.w1; groot.w2 += g.w2; groot.l0 += g.l0; groot.l1 += g.l1; groot.l2 += g.l2; }
    }

    // ---- sampling --------------------------------------------------------------------------------------------------
    // IK chain h is rooted at the body about to be visited: solve it for this sample (joint.cpp:38-44, ikhandler.cpp:103-164)
    __device__ __forceinline__ void ik_chain(const int h)
    {
        if (!IK) return;
        const OmoDevScene& S = c_scene;
        double tw[3];
#pragma unroll
        for (int k = 0; k < 3; ++k)
            tw[k] = (S.ik_chan_n[h][k] > 0) ? eval_channel<!SELF>(pc, P.n_cp, S.ik_chan_pos[h][k], S.ik_chan_n[h][k], t, sorted, n_bez, n_newton)
                                            : S.ik_chan_const[h][k];
        double Lp[9], tp[3];
#pragma unroll
        for (int c = 0; c < 9; ++c) Lp[c] = FK ? L[c] : 0.0;
#pragma unroll
        for (int c = 0; c < 3; ++c) tp[c] = FK ? W[c] : 0.0;
        unsigned it = 0;
        ik_solve(h, Lp, tp, tw, ikq + h * 3 * OMO_MAX_IK_CHAIN, it);
        if (owned) { n_ikiter += it; n_iksolve += 1u; }
    }

    // up to three keyed joint channels (entries i0.. of grp_list; ch < 0 = unused) sampled in lockstep into the exchange rows
    __device__ __forceinline__ void curve_triple(const int i0, const int ch0, const int ch1, const int ch2)
    {
        const OmoDevScene& S = c_scene;
        const int ch[3] = {ch0, ch1, ch2};
        int pos[3], nn[3], row[3], sg[3];
#pragma unroll
        for (int k = 0; k < 3; ++k) {
            const bool on = ch[k] >= 0;
            const int c = on ? ch[k] : 0;
            pos[k] = on ? S.chan_pos[c] : 0; nn[k] = on ? S.chan_n[c] : 0; row[k] = on ? S.chan_row[c] : 0;
            sg[k] = on ? seg0[i0 + k] : 0;
        }
        double val[3];
        unsigned eb = 0, en = 0;
        eval_channels<3, COUNT>(pc, P.n_cp, pos, nn, sg, t, sorted, val, eb, en);
        if (COUNT && owned) { n_bez += eb; n_newton += en; }   // halo samples repeat a neighbour tile's work
#pragma unroll
        for (int k = 0; k < 3; ++k)
            if (nn[k] > 0) ex[row[k] * NT + idx] = val[k];
    }
    __device__ __forceinline__ void group_sync() { __syncthreads(); }

    // ---- staged sampling (SEGG > 0) ----------------------------------------------------------------------------------
    // Asynchronous copy (cp.async, 16 B per request) of the OMO_SEGW segment records of each of the nch keyed channels
    // grp_list[g0 ..] into staging buffer `buf`. Issued one group ahead; stage_wait + the group barrier publish it.
    // first staged record of keyed channel i: the segment that holds the tile's first sample, clamped into [1, n - 1]
    __device__ __forceinline__ int stage_base(const int i) const
    {
        const int n = c_scene.chan_n[c_scene.grp_list[i]];
        return (n >= 2) ? min(max((int)__ldg(tsg + i), 1), n - 1) : 1;
    }
    // FIRST: the block's first stage -- wbase[] is not visible yet, read the tile table instead
    template <bool FIRST = false>
    __device__ __forceinline__ void stage_issue(const int buf, const int g0, const int nch)
    {
        if (SEGG == 0 || SELF || !sorted) return;
        const OmoDevScene& S = c_scene;
        constexpr int U = OMO_SEGD / 2;   // 16-byte units per record
        const int items = nch * OMO_SEGW * U;
        for (int it = tid; it < items; it += NT) {
            const int lc = it / (OMO_SEGW * U), r = it - lc * (OMO_SEGW * U), j = r / U, u = r - j * U;
            const int ch = S.grp_list[g0 + lc], base = FIRST ? stage_base(g0 + lc) : wbase[g0 + lc];
            if (base + j <= S.chan_n[ch] - 1) {
                const unsigned dst = (unsigned)__cvta_generic_to_shared(wsb + (size_t)((buf * SEGG + lc) * OMO_SEGW + j) * OMO_SEGD + 2 * u);
                const double* src = sgc + (size_t)(S.chan_pos[ch] + base + j) * OMO_SEGD + 2 * u;
                asm volatile("cp.async.ca.shared.global [%0], [%1], 16;" ::"r"(dst), "l"(src) : "memory");
            }
        }
        if (P.perm != nullptr) {   // this group's slice of the tile's sorted work list
            constexpr int kPermBuf = ((SEGG + 2) / 3) * 3 * NT;
            const int entries = ((nch + 2) / 3) * 3 * NT;
            const unsigned short* src = P.perm + (size_t)tile_ * P.perm_stride + perm_staged;
            for (int u = tid; u < entries / 8; u += NT) {
                const unsigned dst = (unsigned)__cvta_generic_to_shared(wpm + (size_t)buf * kPermBuf + 8 * u);
                asm volatile("cp.async.ca.shared.global [%0], [%1], 16;" ::"r"(dst), "l"(src + 8 * u) : "memory");
            }
            perm_staged += entries;
        }
        asm volatile("cp.async.commit_group;" ::: "memory");
    }
    __device__ __forceinline__ void stage_wait()
    {
        if (SEGG > 0 && !SELF) asm volatile("cp.async.wait_all;" ::: "memory");
    }

    // The segment-search plan of the nch channels grp_list[g0 ..] staged in buffer buf (eval_channels_planned), built by every warp
    // for itself -- 5 thresholds per channel, one 7-step search over the window's sample times each, spread over the lanes -- so that
    // no block-wide barrier is needed between the staging of a group and its solve.
    __device__ __forceinline__ void plan_build(const int buf, const int g0, const int nch)
    {
        unsigned char* __restrict__ pl = reinterpret_cast<unsigned char*>(wplan + (tid >> 5) * SEGG);
#pragma unroll 1
        for (int it = tid & 31; it < nch * 5; it += 32) {
            const int lc = it / 5, j = it - lc * 5;
            const int n = wcn[g0 + lc], nvalid = min(OMO_SEGW, n - wbase[g0 + lc]);
            const double* __restrict__ rec = wsb + (size_t)((buf * SEGG + lc) * OMO_SEGW) * OMO_SEGD;
            int th = 128;                                   // never: no record j - 1 in the window
            if (j - 1 < nvalid && nvalid >= 1) {
                const double T = (j == 0) ? rec[0] : rec[(j - 1) * OMO_SEGD + 1];   // the key opening record 0 / closing record j - 1
                int lo = 0, hi = NT;                        // th = number of window samples with !(T <= t): the times ascend
                while (lo < hi) {
                    const int mid = (lo + hi) >> 1;
                    if (!(T <= wts[mid])) lo = mid + 1; else hi = mid;
                }
                th = lo;
            }
            // byte 4 = th_in (j == 0), bytes 0-3 = th_0..3 (j = 1..4)
            pl[lc * 8 + (j == 0 ? 4 : j - 1)] = (unsigned char)th;
            if (j == 0) {
                pl[lc * 8 + 5] = (unsigned char)max(nvalid, 0);
                pl[lc * 8 + 6] = (unsigned char)((n < 2 ? 1 : 0) | (n - wbase[g0 + lc] > OMO_SEGW ? 2 : 0));
                pl[lc * 8 + 7] = 0;
            }
        }
        __syncwarp();
    }

    // K (<= 3) keyed joint channels (entries i0.. of grp_list, rows lc0.. of staging buffer buf) sampled in lockstep from
    // the staged records into the exchange rows
    template <int K>
    __device__ __forceinline__ void curve_staged(const int buf, const int i0, const int lc0, const int ch0, const int ch1, const int ch2)
    {
        const OmoDevScene& S = c_scene;
        const int ch[3] = {ch0, ch1, ch2};
        double val[K];
        if (!sorted) {   // a key moved past its neighbour: list-order search through global memory, one curve at a time
#pragma unroll
            for (int k = 0; k < K; ++k) {
                unsigned eb = 0, en = 0;
                val[k] = COUNT ? eval_channel<!SELF>(pc, P.n_cp, S.chan_pos[ch[k]], S.chan_n[ch[k]], t, false, eb, en)
                               : eval_channel_slow<!SELF>(pc, P.n_cp, S.chan_pos[ch[k]], S.chan_n[ch[k]], t, false);
                if (COUNT && owned) { n_bez += eb; n_newton += en; }
            }
        } else {
            const double* sk[K];
            int nn[K], base[K];
#pragma unroll
            for (int k = 0; k < K; ++k) {
                sk[k] = SELF ? seg_sm + (size_t)(S.chan_pos[ch[k]] + wbase[i0 + k]) * OMO_SEGD
                             : wsb + (size_t)((buf * SEGG + lc0 + k) * OMO_SEGW) * OMO_SEGD;
                nn[k] = S.chan_n[ch[k]];
                base[k] = wbase[i0 + k];
            }
            unsigned fb = 0, eb = 0, en = 0;
            double tk[K];
#pragma unroll
            for (int k = 0; k < K; ++k) tk[k] = t;
            if constexpr (OMO_PLAN != 0 && !SELF) {
                unsigned long long plan[K];
                int sik[K];
                const unsigned long long* __restrict__ wp = wplan + (tid >> 5) * SEGG + lc0;
#pragma unroll
                for (int k = 0; k < K; ++k) { plan[k] = wp[k]; sik[k] = idx; }
                eval_channels_planned<K, COUNT>(sk, plan, sik, tk, val, fb, eb, en);
            } else
            eval_channels_staged<K, COUNT>(sk, nn, base, tk, val, fb, eb, en);
            if (fb) {   // this thread's segment lies beyond the staged window (keys denser than OMO_SEGW per block)
#pragma unroll
                for (int k = 0; k < K; ++k)
                    if (fb >> k & 1u)
                        val[k] = COUNT ? eval_channel<!SELF>(pc, P.n_cp, S.chan_pos[ch[k]], nn[k], t, true, eb, en)
                                       : eval_channel_slow<!SELF>(pc, P.n_cp, S.chan_pos[ch[k]], nn[k], t, true);
            }
            if (COUNT && owned) { n_bez += eb; n_newton += en; }
        }
#pragma unroll
        for (int k = 0; k < K; ++k) ex[S.chan_row[ch[k]] * NT + idx] = val[k];
    }

    // ---- forward step of a joint frozen at zero angles: E = I, qd = qdd = 0 exactly --------------------------------
    __device__ __forceinline__ void body_ident(const int b)
    {
        const OmoDevScene& S = c_scene;
        if (OUT && owned) {
            const size_t o = s * dof + 3 * b;
#pragma unroll
            for (int k = 0; k < 3; ++k) { P.o_q[o + k] = 0.0; P.o_qd[o + k] = 0.0; P.o_qdd[o + k] = 0.0; }
        }
        SV6 vn, an;
        rnea_forward_identity(S.body[b], v, a, vn, an, f);
        v = vn; a = an;
    }

    // ---- forward step of a general EulerXYZ joint: angles (exchange rows / constants / IK), stencils
    // (core_rbdl.cpp:28-95), sin / cos, jcalc + RNEA forward (Joint.cc:100-139, Dynamics.cc:39-73) ------------------
    // mask bit k: channel k comes from exchange row rowK; ikh >= 0: all three come from IK handler ikh, chain position ik_idx.
    __device__ __forceinline__ void body_general(const int b, const int mask_in, const int row0, const int row1, const int row2,
                                                 const int ikh, const int ik_idx)
    {
        const OmoDevScene& S = c_scene;
        const int mask = (IK && ikh >= 0) ? 7 : mask_in;
        double q[3], qd[3], qdd[3], ss[3], cc[3];
        int rowk[3] = {row0, row1, row2};
        if (IK && ikh >= 0) {
            // IK-driven joint: its angles were solved per thread; exchange them through 3 dedicated double-buffered rows
#pragma unroll
            for (int k = 0; k < 3; ++k) {
                rowk[k] = S.n_ex_rows + xbuf * 3 + k;
                ex[rowk[k] * NT + idx] = ikq[ikh * 3 * OMO_MAX_IK_CHAIN + 3 * ik_idx + k];
            }
            __syncthreads();
            xbuf ^= 1;
        }
#pragma unroll
        for (int k = 0; k < 3; ++k) {
            if (mask >> k & 1) {
                const double* __restrict__ row = ex + rowk[k] * NT;
                // stencil neighbours inside the block (clamped for the halo samples, whose results are unused)
                const double m2 = row[max(idx - 2, 0)], m1 = row[max(idx - 1, 0)], x = row[idx];
                const double p1 = row[min(idx + 1, NT - 1)], p2 = row[min(idx + 2, NT - 1)];
                q[k] = x;
                if (interior) {
                    qd[k] = (-p2 + 8.0 * p1 - 8.0 * m1 + m2) * P.inv_12dt;
                    qdd[k] = (-p2 + 16.0 * p1 - 30.0 * x + 16.0 * m1 - m2) * P.inv_12dt2;
                } else if (s == 0) {
                    qd[k] = (p1 - x) * P.inv_dt;
                    qdd[k] = (p2 - 2.0 * p1 + x) * P.inv_2dt2;
                } else if (s == N - 1) {
                    qd[k] = (x - m1) * P.inv_dt;
                    qdd[k] = (x - 2.0 * m1 + m2) * P.inv_2dt2;
                } else {
                    qd[k] = (p1 - m1) * P.inv_2dt;
                    qdd[k] = (p1 - 2.0 * x + m1) * P.inv_dt2;
                }
                sincos_angle(x, &ss[k], &cc[k]);
            } else {
                q[k] = S.chan_const[3 * b + k]; ss[k] = S.chan_sin[3 * b + k]; cc[k] = S.chan_cos[3 * b + k];
                qd[k] = interior ? S.chan_qd[3 * b + k] : 0.0;
                qdd[k] = interior ? S.chan_qdd[3 * b + k] : 0.0;
            }
        }
        s0 = ss[0]; s1 = ss[1]; s2 = ss[2]; c0 = cc[0]; c1 = cc[1]; c2 = cc[2];
        if (OUT && owned) {
            const size_t o = s * dof + 3 * b;
#pragma unroll
            for (int k = 0; k < 3; ++k) { P.o_q[o + k] = q[k]; P.o_qd[o + k] = qd[k]; P.o_qdd[o + k] = qdd[k]; }
        }
        SV6 vn, an;
        rnea_forward(S.body[b], s0, c0, s1, c1, s2, c2, qd[0], qd[1], qd[2], qdd[0], qdd[1], qdd[2], v, a, vn, an, f);
        v = vn; a = an;
    }

    // A whole curve group: n3 lockstep triples as ONE loop body (the channel numbers come from grp_list) plus a tail of
    // `rem` curves. Keeps the code of a 189-curve scene the size of one triple -- the specialised kernel is bound by
    // instruction supply (profiles/micro/icache_mix), straight-line repetition of the solver is what it cannot afford.
    // Sorted form of a curve group (OmoEvalParams::perm): the group's (channel, sample) items of this tile come in the order of
    // the Newton iterations the ORIGINAL motion needs for them; chunk p of 96 items is solved by warp p % 4 as one lockstep triple
    // (32 lanes x 3 curves), so curves of similar cost share a warp instead of every warp waiting for the slowest of its 96. Which
    // lane solves which curve changes nothing in the curve's own arithmetic: the exchange rows hold the same bits.
    __device__ __forceinline__ void curve_group_sorted(const int buf, const int g0, const int n3tot)
    {
        const OmoDevScene& S = c_scene;
        constexpr int kPermBuf = ((SEGG + 2) / 3) * 3 * NT;
        const unsigned short* __restrict__ list = wpm + (size_t)buf * kPermBuf + ((tid >> 5) * 3) * 32 + (tid & 31);
        perm_cursor += n3tot * 3 * NT;
#pragma unroll 1
        for (int c = 0; c < n3tot; ++c) {
            const unsigned short* __restrict__ cur = list + (size_t)c * (3 * NT);
            const unsigned e[3] = {cur[0], cur[32], cur[64]};
            const double* sk[3];
            int nn[3], base[3], si[3], lc[3];
            double tk[3], val[3];
            bool on[3];
#pragma unroll
            for (int k = 0; k < 3; ++k) {
                on[k] = e[k] != 0xffffu;
                lc[k] = on[k] ? (int)(e[k] >> 7) : 0;
                si[k] = (int)(e[k] & 127u);
                sk[k] = wsb + (size_t)((buf * SEGG + lc[k]) * OMO_SEGW) * OMO_SEGD;
                nn[k] = on[k] ? wcn[g0 + lc[k]] : 0;
                base[k] = wbase[g0 + lc[k]];
                tk[k] = wts[si[k]];
            }
            unsigned fb = 0, eb = 0, en = 0;
            // (the thresholds of eval_channels_planned do not pay here: every lane is on another channel, so the plan words are per-lane
            // loads like the records, and the slot's byte pattern has to be made per item -- measured on stick240: 0.8776 -> 0.8848 ms)
            eval_channels_staged<3, COUNT>(sk, nn, base, tk, val, fb, eb, en);
            if (fb) {   // the item's segment lies beyond the staged window (keys denser than OMO_SEGW per block)
#pragma unroll
                for (int k = 0; k < 3; ++k)
                    if (on[k] && (fb >> k & 1u)) {
                        const int ch = S.grp_list[g0 + lc[k]];
                        val[k] = COUNT ? eval_channel(pc, P.n_cp, S.chan_pos[ch], nn[k], tk[k], true, eb, en)
                                       : eval_channel_slow(pc, P.n_cp, S.chan_pos[ch], nn[k], tk[k], true);
                    }
            }
#pragma unroll
            for (int k = 0; k < 3; ++k)
                if (on[k]) ex[wrow[g0 + lc[k]] * NT + si[k]] = val[k];
        }
    }

    template <int REM>
    __device__ __forceinline__ void curve_group(const int buf, const int g0, const int n3)
    {
        const OmoDevScene& S = c_scene;
        if (SEGG > 0 && !SELF && !COUNT && P.perm != nullptr && sorted) { curve_group_sorted(buf, g0, n3 + (REM > 0 ? 1 : 0)); return; }
        if (OMO_PLAN != 0 && SEGG > 0 && !SELF && sorted) plan_build(buf, g0, 3 * n3 + REM);
#pragma unroll 1
        for (int c = 0; c < n3; ++c) {
            const int i0 = g0 + 3 * c;
            curve_staged<3>(buf, i0, 3 * c, S.grp_list[i0], S.grp_list[i0 + 1], S.grp_list[i0 + 2]);
        }
        if (REM > 0) {
            const int i0 = g0 + 3 * n3;
            curve_staged<(REM > 0 ? REM : 1)>(buf, i0, 3 * n3, S.grp_list[i0], REM > 1 ? S.grp_list[i0 + 1] : -1, -1);
        }
    }

    // The per-sample form of a curve group with K curves in lockstep per thread instead of three (the generator's choice per scene,
    // omo_jit.hpp curve_k: more independent Newton chains per warp where only 12 warps fit an SM). Same curves, same arithmetic per
    // curve: which curves share a lockstep round changes no bit.
    template <int K>
    __device__ __forceinline__ void curve_staged_n(const int buf, const int i0, const int lc0)
    {
        const OmoDevScene& S = c_scene;
        int ch[K];
#pragma unroll
        for (int k = 0; k < K; ++k) ch[k] = S.grp_list[i0 + k];
        double val[K];
        if (!sorted) {
#pragma unroll
            for (int k = 0; k < K; ++k) {
                unsigned eb = 0, en = 0;
                val[k] = COUNT ? eval_channel<!SELF>(pc, P.n_cp, S.chan_pos[ch[k]], S.chan_n[ch[k]], t, false, eb, en)
                               : eval_channel_slow<!SELF>(pc, P.n_cp, S.chan_pos[ch[k]], S.chan_n[ch[k]], t, false);
                if (COUNT && owned) { n_bez += eb; n_newton += en; }
            }
        } else {
            const double* sk[K];
            int nn[K], base[K];
#pragma unroll
            for (int k = 0; k < K; ++k) {
                sk[k] = SELF ? seg_sm + (size_t)(S.chan_pos[ch[k]] + wbase[i0 + k]) * OMO_SEGD
                             : wsb + (size_t)((buf * SEGG + lc0 + k) * OMO_SEGW) * OMO_SEGD;
                nn[k] = S.chan_n[ch[k]];
                base[k] = wbase[i0 + k];
            }
            unsigned fb = 0, eb = 0, en = 0;
            double tk[K];
#pragma unroll
            for (int k = 0; k < K; ++k) tk[k] = t;
            if constexpr (OMO_PLAN != 0 && !SELF) {
                unsigned long long plan[K];
                int sik[K];
                const unsigned long long* __restrict__ wp = wplan + (tid >> 5) * SEGG + lc0;
#pragma unroll
                for (int k = 0; k < K; ++k) { plan[k] = wp[k]; sik[k] = idx; }
                eval_channels_planned<K, COUNT>(sk, plan, sik, tk, val, fb, eb, en);
            } else
            eval_channels_staged<K, COUNT>(sk, nn, base, tk, val, fb, eb, en);
            if (fb) {
#pragma unroll
                for (int k = 0; k < K; ++k)
                    if (fb >> k & 1u)
                        val[k] = COUNT ? eval_channel<!SELF>(pc, P.n_cp, S.chan_pos[ch[k]], nn[k], t, true, eb, en)
                                       : eval_channel_slow<!SELF>(pc, P.n_cp, S.chan_pos[ch[k]], nn[k], t, true);
            }
            if (COUNT && owned) { n_bez += eb; n_newton += en; }
        }
#pragma unroll
        for (int k = 0; k < K; ++k) ex[S.chan_row[ch[k]] * NT + idx] = val[k];
    }

    template <int K, int REM>
    __device__ __forceinline__ void curve_group_n(const int buf, const int g0, const int nk)
    {
        if (OMO_PLAN != 0 && SEGG > 0 && !SELF && sorted) plan_build(buf, g0, K * nk + REM);
#pragma unroll 1
        for (int c = 0; c < nk; ++c) curve_staged_n<K>(buf, g0 + K * c, K * c);
        if (REM > 0) curve_staged_n<(REM > 0 ? REM : 1)>(buf, g0 + K * nk, K * nk);
    }

    // ---- the same steps with the scene's structural zeros resolved at compile time (scene-specialised kernels) ---------
    // Channel kind Kk of the body's angle k: 0 = keyed (exchange row, stencil, sincos), 1 = constant (sin / cos / stencil
    // residue from the scene table), 3 = the constant 0 (sin = ZeroT, cos = OneT, qd = qdd = ZeroT). ZB: omo_rnea_z.cuh.
    template <int KIND> static __device__ __forceinline__ auto zs(const double x) { if constexpr (KIND == 3) return ZeroT{}; else return x; }
    template <int KIND> static __device__ __forceinline__ auto zc(const double x) { if constexpr (KIND == 3) return OneT{}; else return x; }
    template <int K0, int K1, int K2>
    static __device__ __forceinline__ auto make_sc(const double a0, const double b0, const double a1, const double b1, const double a2, const double b2)
    {
        return SinCosZ<decltype(zs<K0>(0.0)), decltype(zc<K0>(0.0)), decltype(zs<K1>(0.0)), decltype(zc<K1>(0.0)), decltype(zs<K2>(0.0)),
                       decltype(zc<K2>(0.0))>{zs<K0>(a0), zc<K0>(b0), zs<K1>(a1), zc<K1>(b1), zs<K2>(a2), zc<K2>(b2)};
    }

    template <int K0, int K1, int K2, int ZB>
    __device__ __forceinline__ void body_z(const int b, const int row0, const int row1, const int row2)
    {
        const OmoDevScene& S = c_scene;
        constexpr int kind[3] = {K0, K1, K2};
        const int rowk[3] = {row0, row1, row2};
        double qd[3] = {0.0, 0.0, 0.0}, qdd[3] = {0.0, 0.0, 0.0}, ss[3] = {0.0, 0.0, 0.0}, cc[3] = {1.0, 1.0, 1.0};
        [[maybe_unused]] double xq[3] = {0.0, 0.0, 0.0};
#pragma unroll
        for (int k = 0; k < 3; ++k) {
            if (kind[k] == 0) {
                const double* __restrict__ row = ex + rowk[k] * NT;
                const double m2 = row[max(idx - 2, 0)], m1 = row[max(idx - 1, 0)], x = row[idx];
                const double p1 = row[min(idx + 1, NT - 1)], p2 = row[min(idx + 2, NT - 1)];
                if (interior) {
                    qd[k] = (-p2 + 8.0 * p1 - 8.0 * m1 + m2) * P.inv_12dt;
                    qdd[k] = (-p2 + 16.0 * p1 - 30.0 * x + 16.0 * m1 - m2) * P.inv_12dt2;
                } else if (s == 0) {
                    qd[k] = (p1 - x) * P.inv_dt;
                    qdd[k] = (p2 - 2.0 * p1 + x) * P.inv_2dt2;
                } else if (s == N - 1) {
                    qd[k] = (x - m1) * P.inv_dt;
                    qdd[k] = (x - 2.0 * m1 + m2) * P.inv_2dt2;
                } else {
                    qd[k] = (p1 - m1) * P.inv_2dt;
                    qdd[k] = (p1 - 2.0 * x + m1) * P.inv_dt2;
                }
                if constexpr (OMO_SINCOS_CALL == 1 && K0 == 0 && K1 == 0 && K2 == 0) xq[k] = x;
                else if constexpr (OMO_SINCOS_CALL == 2) { const double2 r = sincos_angle1(x); ss[k] = r.x; cc[k] = r.y; }
                else sincos_angle(x, &ss[k], &cc[k]);
            } else if (kind[k] == 1) {
                ss[k] = S.chan_sin[3 * b + k]; cc[k] = S.chan_cos[3 * b + k];
                qd[k] = interior ? S.chan_qd[3 * b + k] : 0.0;
                qdd[k] = interior ? S.chan_qdd[3 * b + k] : 0.0;
            }
        }
        if constexpr (OMO_SINCOS_CALL == 1 && K0 == 0 && K1 == 0 && K2 == 0) {
            const SinCos3 r = sincos_angle3(xq[0], xq[1], xq[2]);
            ss[0] = r.s0; cc[0] = r.c0; ss[1] = r.s1; cc[1] = r.c1; ss[2] = r.s2; cc[2] = r.c2;
        }
        s0 = ss[0]; s1 = ss[1]; s2 = ss[2]; c0 = cc[0]; c1 = cc[1]; c2 = cc[2];
        const auto sc = make_sc<K0, K1, K2>(ss[0], cc[0], ss[1], cc[1], ss[2], cc[2]);
        const Vec3Z<decltype(zs<K0>(0.0)), decltype(zs<K1>(0.0)), decltype(zs<K2>(0.0))> vqd{zs<K0>(qd[0]), zs<K1>(qd[1]), zs<K2>(qd[2])},
            vqdd{zs<K0>(qdd[0]), zs<K1>(qdd[1]), zs<K2>(qdd[2])};
        SV6 vn, an;
        rnea_forward_z(BodyZ<ZB>(S.body[b]), sc, vqd, vqdd, v, a, vn, an, f);
        v = vn; a = an;
    }
    // (f, sincos of the KEYED angles) of a body whose subtree opens; constants are re-read from the scene table on the way back
    template <int K0, int K1, int K2> __device__ __forceinline__ void park_fs_z(const int sl)
    {
        st(sl, 0, f.w0); st(sl, 1, f.w1); st(sl, 2, f.w2); st(sl, 3, f.l0); st(sl, 4, f.l1); st(sl, 5, f.l2);
        // keyed angles only, packed behind f: rows 6.., two per keyed channel
        constexpr int o1 = 6 + (K0 == 0 ? 2 : 0), o2 = o1 + (K1 == 0 ? 2 : 0);
        if (K0 == 0) { st(sl, 6, s0); st(sl, 7, c0); }
        if (K1 == 0) { st(sl, o1, s1); st(sl, o1 + 1, c1); }
        if (K2 == 0) { st(sl, o2, s2); st(sl, o2 + 1, c2); }
    }
    template <int K0, int K1, int K2, int ZB>
    __device__ __forceinline__ void finalize_z(const int k, const SV6& fk, const double a0, const double b0, const double a1, const double b1,
                                               const double a2, const double b2)
    {
        const OmoDevScene& S = c_scene;
        double t0, t1, t2;
        SV6 gn;
        rnea_backward_z(BodyZ<ZB>(S.body[k]), make_sc<K0, K1, K2>(a0, b0, a1, b1, a2, b2), fk, gn, t0, t1, t2);
        g = gn;
        tsq = fma(t0, t0, fma(t1, t1, fma(t2, t2, tsq)));
    }
    template <int K0, int K1, int K2, int ZB> __device__ __forceinline__ void fold_leaf_z(const int b)
    {
        finalize_z<K0, K1, K2, ZB>(b, f, s0, c0, s1, c1, s2, c2);
    }
    template <int K0, int K1, int K2, int ZB> __device__ __forceinline__ void fold_up_z(const int k, const int sl)
    {
        const OmoDevScene& S = c_scene;
        SV6 fk;
        fk.w0 = ld(sl, 0) + g.w0; fk.w1 = ld(sl, 1) + g.w1; fk.w2 = ld(sl, 2) + g.w2;
        fk.l0 = ld(sl, 3) + g.l0; fk.l1 = ld(sl, 4) + g.l1; fk.l2 = ld(sl, 5) + g.l2;
        constexpr int o1 = 6 + (K0 == 0 ? 2 : 0), o2 = o1 + (K1 == 0 ? 2 : 0);
        const double a0 = K0 == 0 ? ld(sl, 6) : (K0 == 1 ? S.chan_sin[3 * k] : 0.0), b0 = K0 == 0 ? ld(sl, 7) : (K0 == 1 ? S.chan_cos[3 * k] : 1.0);
        const double a1 = K1 == 0 ? ld(sl, o1) : (K1 == 1 ? S.chan_sin[3 * k + 1] : 0.0), b1 = K1 == 0 ? ld(sl, o1 + 1) : (K1 == 1 ? S.chan_cos[3 * k + 1] : 1.0);
        const double a2 = K2 == 0 ? ld(sl, o2) : (K2 == 1 ? S.chan_sin[3 * k + 2] : 0.0), b2 = K2 == 0 ? ld(sl, o2 + 1) : (K2 == 1 ? S.chan_cos[3 * k + 2] : 1.0);
        finalize_z<K0, K1, K2, ZB>(k, fk, a0, b0, a1, b1, a2, b2);
    }

    // ---- forward kinematics: p_b and the world affine handed to b's children (joint.cpp:64-83) --------------------------
    __device__ __forceinline__ void fk_step(const int b, const bool has_child, const bool ident)
    {
        if (!FK) return;
        const OmoDevScene& S = c_scene;
        const double o0 = S.offset[b][0], o1 = S.offset[b][1], o2 = S.offset[b][2];
        const double p0 = fma(L[0], o0, fma(L[1], o1, fma(L[2], o2, W[0])));
        const double p1 = fma(L[3], o0, fma(L[4], o1, fma(L[5], o2, W[1])));
        const double p2 = fma(L[6], o0, fma(L[7], o1, fma(L[8], o2, W[2])));
        if (P.p_orig) {
            const double d0 = p0 - __ldg(P.p_orig + (size_t)(3 * b) * N + s), d1 = p1 - __ldg(P.p_orig + (size_t)(3 * b + 1) * N + s),
                         d2 = p2 - __ldg(P.p_orig + (size_t)(3 * b + 2) * N + s);
            dtr += d0 * d0 + d1 * d1 + d2 * d2;
        }
        if (OUT && owned && P.o_p) { P.o_p[s * dof + 3 * b] = p0; P.o_p[s * dof + 3 * b + 1] = p1; P.o_p[s * dof + 3 * b + 2] = p2; }
        if (has_child) {
            if (!ident) {
                // L <- L * Rz(q2) Ry(q1) Rx(q0)
                const double R00 = c2 * c1, R01 = c2 * s1 * s0 - s2 * c0, R02 = c2 * s1 * c0 + s2 * s0;
                const double R10 = s2 * c1, R11 = s2 * s1 * s0 + c2 * c0, R12 = s2 * s1 * c0 - c2 * s0;
                const double R20 = -s1, R21 = c1 * s0, R22 = c1 * c0;
                double n[9];
#pragma unroll
                for (int r = 0; r < 3; ++r) {
                    n[3 * r] = fma(L[3 * r], R00, fma(L[3 * r + 1], R10, L[3 * r + 2] * R20));
                    n[3 * r + 1] = fma(L[3 * r], R01, fma(L[3 * r + 1], R11, L[3 * r + 2] * R21));
                    n[3 * r + 2] = fma(L[3 * r], R02, fma(L[3 * r + 1], R12, L[3 * r + 2] * R22));
                }
#pragma unroll
                for (int c = 0; c < 9; ++c) L[c] = n[c];
            }
            W[0] = p0; W[1] = p1; W[2] = p2;
        }
    }

    // ---- cost of the sample and block reduction (local_cost = |tau[3:]|^2 dt, core_rbdl.cpp:331-334) ----------------
    __device__ __forceinline__ void finish(double*)
    {
        double* smem = sm0;
        double wc = 0.0, wt = 0.0;
        const double local = tsq * P.dt;
        if (OUT && owned) {
            // tau[0:3] of the TranslationXYZ root = linear part of the force folded into body 1 (Dynamics.cc:91-92)
            P.o_tau[s * dof] = groot.l0; P.o_tau[s * dof + 1] = groot.l1; P.o_tau[s * dof + 2] = groot.l2;
            P.o_local[s] = local;
        }
        if (owned) {
            wc = __ldg(P.wcost + s) * local;
            if (FK) wt = P.w_traj * dtr * P.dt;
        }
        __syncthreads();  // everyone is done with the exchange buffers
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            wc += __shfl_down_sync(0xffffffffu, wc, o);
            wt += __shfl_down_sync(0xffffffffu, wt, o);
        }
        double* red = smem;
        if ((tid & 31) == 0) { red[2 * (tid >> 5)] = wc; red[2 * (tid >> 5) + 1] = wt; }
        __syncthreads();
        if (tid == 0 && item_valid) {
            double C = 0.0, D = 0.0;
            for (int w = 0; w < (NT >> 5); ++w) { C += red[2 * w]; D += red[2 * w + 1]; }
            P.partial[(size_t)item * 2] = C;
            P.partial[(size_t)item * 2 + 1] = D;
        }
        if (SELF) self_finish();
        if (COUNT && P.counters) {
            unsigned cb = n_bez, cn = n_newton, ci = n_iksolve, cj = n_ikiter;
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) {
                cb += __shfl_down_sync(0xffffffffu, cb, o); cn += __shfl_down_sync(0xffffffffu, cn, o);
                ci += __shfl_down_sync(0xffffffffu, ci, o); cj += __shfl_down_sync(0xffffffffu, cj, o);
            }
            if ((tid & 31) == 0) {
                atomicAdd(P.counters + 0, (unsigned long long)cb); atomicAdd(P.counters + 1, (unsigned long long)cn);
                atomicAdd(P.counters + 2, (unsigned long long)ci); atomicAdd(P.counters + 3, (unsigned long long)cj);
            }
        }
    }
};

}  // namespace omo
#endif
