// objective_kernel: the fused sampling -> stencil -> RNEA -> cost kernel (see omo_kernels.cuh for the mapping).
//
// Restates, per (candidate, time sample):
//   Core::SetStates           OptiMo/core_rbdl.cpp:108-166   (q from curves / IK, p from FK)
//   CalculateFirst/SecondCentral  core_rbdl.cpp:28-95        (SECOND_ORDER_ACCURATE off)
//   Joint::GetAffineRelativeTo{Parent,World}  OptiMo/joint.cpp:64-83   (T(offset) Rz Ry Rx)
//   Core::CalculateCost       core_rbdl.cpp:307-346          (local_cost = |tau[3:]|^2 dt, weight w(t))
//   D_traj                    core_rbdl.cpp:387-398
//
// Template parameters: NS = time samples per thread (2 interleaves two independent dependency chains per
// thread for the FP64 pipe and halves the per-body table / address overhead; 1 is the low-latency form for
// small batches), FK = also run forward kinematics (positions; needed for D_traj and for IK), IK = scene has
// IK chains, OUT = write the per-sample OptimizationCache arrays (single-candidate launches only; walks the FULL
// tree, every other instance walks the pruned tree, see OmoDevScene::meta), COUNT = book executed curve / IK work.
#ifndef OMO_OBJECTIVE_CUH
#define OMO_OBJECTIVE_CUH

#include "omo_kernels.cuh"

namespace omo {

template <bool FK> struct Smp {
    SV6 v, a, f;                          // state of the body visited last (v, a: also the parent state for the next one)
    double s0, c0, s1, c1, s2, c2;        // sin / cos of its Euler angles
    double L[FK ? 9 : 1], W[FK ? 3 : 1];  // world affine handed to the next body's FK step
    double tsq, dtr;                      // sum tau^2 (non-root DoFs), sum |p - p_original|^2
    double t;                             // time of this sample
    int s;                                // sample index (clamped into [0, N_s))
    int idx;                              // position inside the block's window (exchange-buffer column)
    bool owned, interior;
};

#define OMO_NT 128  // threads per block (compile-time: every stack / exchange access gets an immediate offset)
#define OMO_STK(slot, comp, j) stk[(((slot) * OMO_SLOT + (comp)) * NS + (j)) * OMO_NT + tid]

struct BodyMeta {
    int parent, fs, va, fk;   // slots: -1 = none
    int dyn, grp, ik_of, ik_idx, ik_start;
    bool ident, has_child;
    int row[3];
};

__device__ __forceinline__ int meta_byte(const int w, const int sh) { const int v = (w >> sh) & 0xff; return v == 0xff ? -1 : v; }

__device__ __forceinline__ BodyMeta load_meta(const int tree, const int b)
{
    const int4 w = c_scene.meta[tree][b];
    BodyMeta M;
    M.parent = w.x & 0xff; M.fs = meta_byte(w.x, 8); M.va = meta_byte(w.x, 16); M.fk = meta_byte(w.x, 24);
    M.dyn = w.y & 7; M.ident = (w.y >> 3) & 1; M.has_child = (w.y >> 4) & 1;
    M.grp = ((w.y >> 8) & 0xff) - 1; M.ik_of = ((w.y >> 16) & 0xf) - 1; M.ik_idx = (w.y >> 20) & 0xf; M.ik_start = ((w.y >> 24) & 0xf) - 1;
    M.row[0] = w.z & 0xff; M.row[1] = (w.z >> 8) & 0xff; M.row[2] = (w.z >> 16) & 0xff;
    return M;
}

template <int NS, bool FK, bool IK, bool OUT, bool COUNT>
__global__ void __launch_bounds__(OMO_NT, (NS == 1 ? 4 : 2)) objective_kernel(const OmoEvalParams P)
{
    static_assert(!(IK && NS != 1), "IK scenes run one sample per thread");
    extern __shared__ double smem[];
    const OmoDevScene& S = c_scene;
    constexpr int NT = OMO_NT;
    constexpr int TREE = OUT ? 0 : 1;           // 0: every body, 1: candidate-independent subtrees pruned
    const int tid = threadIdx.x;
    constexpr int WIN = NS * NT;                // samples in this block's window, halo included
    const int cand = blockIdx.x / P.n_tiles;
    const int tile = blockIdx.x - cand * P.n_tiles;
    const int ex_rows = S.n_ex_rows + (IK ? 6 : 0);
    double* __restrict__ ex = smem;                    // [ex_rows][WIN] stencil exchange: q of the dynamic channels
    double* __restrict__ stk = smem + ex_rows * WIN;   // [n_slots][OMO_SLOT][NS][NT]
    const int N = P.N_s;
    const int J = S.J;
    const int tile_end = min(N, (tile + 1) * P.T_out);
    const double* __restrict__ pc = P.prep + (size_t)cand * 6 * P.n_cp;
    const bool sorted = (__ldg(P.flags + cand) & 1u) != 0u;
    const size_t dof = (size_t)3 * J;

    unsigned n_bez = 0, n_newton = 0, n_iksolve = 0, n_ikiter = 0;
    double ikq[IK ? OMO_MAX_IK * 3 * OMO_MAX_IK_CHAIN : 1];
    SV6 groot[OUT ? NS : 1];
    Smp<FK> sm[NS];

    // ---- segment warm start: one binary search per keyed channel per WARP (for the time of its first lane; the
    // lanes' times ascend), so each thread later only walks the few keys its own time lies beyond ---------------
    // seg0[warp][n_keyed] (16-bit) sits behind the stack in dynamic shared memory
    unsigned short* __restrict__ seg0 =
        reinterpret_cast<unsigned short*>(stk + (size_t)(FK ? S.n_slots_tree[TREE] : S.n_slots_nofk_tree[TREE]) * (OMO_SLOT * NS * NT)) + (tid >> 5) * S.n_keyed;
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

    // root state (RBDL body 1, TranslationXYZ, q = root offset: core_rbdl.cpp:137-141,212): v = (0, qd), a = (0, qdd)
    auto set_root = [&](Smp<FK>& m) {
        m.v.w0 = m.v.w1 = m.v.w2 = 0.0; m.a.w0 = m.a.w1 = m.a.w2 = 0.0;
        m.v.l0 = m.interior ? S.root_qd[0] : 0.0; m.v.l1 = m.interior ? S.root_qd[1] : 0.0; m.v.l2 = m.interior ? S.root_qd[2] : 0.0;
        m.a.l0 = m.interior ? S.root_qdd[0] : 0.0; m.a.l1 = m.interior ? S.root_qdd[1] : 0.0; m.a.l2 = m.interior ? S.root_qdd[2] : 0.0;
    };

#pragma unroll
    for (int j = 0; j < NS; ++j) {
        Smp<FK>& m = sm[j];
        m.idx = tid + j * NT;
        const int s_raw = tile * P.T_out - 2 + m.idx;
        m.s = min(max(s_raw, 0), N - 1);
        m.owned = (m.idx >= 2) && (s_raw < tile_end) && (s_raw >= 0);
        m.interior = (m.s >= 2) && (m.s <= N - 3);
        m.t = __ldg(P.tgrid + m.s);
        m.tsq = (!OUT && m.interior) ? S.static_tsq : 0.0;   // the pruned bodies' share (candidate-independent)
        m.dtr = 0.0;
        set_root(m);
        m.f.w0 = m.f.w1 = m.f.w2 = m.f.l0 = m.f.l1 = m.f.l2 = 0.0;
        m.s0 = m.s1 = m.s2 = 0.0; m.c0 = m.c1 = m.c2 = 1.0;
        if (OUT) { groot[j].w0 = groot[j].w1 = groot[j].w2 = groot[j].l0 = groot[j].l1 = groot[j].l2 = 0.0; }
        if (OUT && m.owned) {
            for (int k = 0; k < 3; ++k) {
                P.o_q[m.s * dof + k] = S.offset[0][k];
                P.o_qd[m.s * dof + k] = m.interior ? S.root_qd[k] : 0.0;
                P.o_qdd[m.s * dof + k] = m.interior ? S.root_qdd[k] : 0.0;
            }
        }
        if (FK) {
            // root world affine = T(offset) Rz Ry Rx with the root's own rot channels (joint.cpp:64-74)
            double e[3];
#pragma unroll
            for (int k = 0; k < 3; ++k)
                e[k] = (S.dyn[0] >> k & 1) ? eval_channel(pc, P.n_cp, S.chan_pos[k], S.chan_n[k], m.t, sorted, n_bez, n_newton)
                                           : S.chan_const[k];
            double sx, cx, sy, cy, sz, cz;
            sincos_angle(e[0], &sx, &cx); sincos_angle(e[1], &sy, &cy); sincos_angle(e[2], &sz, &cz);
            m.L[0] = cz * cy; m.L[1] = cz * sy * sx - sz * cx; m.L[2] = cz * sy * cx + sz * sx;
            m.L[3] = sz * cy; m.L[4] = sz * sy * sx + cz * cx; m.L[5] = sz * sy * cx - cz * sx;
            m.L[6] = -sy;     m.L[7] = cy * sx;                m.L[8] = cy * cx;
            m.W[0] = S.offset[0][0]; m.W[1] = S.offset[0][1]; m.W[2] = S.offset[0][2];
            if (P.p_orig) {
                const double d0 = m.W[0] - __ldg(P.p_orig + (size_t)0 * N + m.s), d1 = m.W[1] - __ldg(P.p_orig + (size_t)1 * N + m.s),
                             d2 = m.W[2] - __ldg(P.p_orig + (size_t)2 * N + m.s);
                m.dtr += d0 * d0 + d1 * d1 + d2 * d2;
            }
            if (OUT && m.owned && P.o_p) { P.o_p[m.s * dof] = m.W[0]; P.o_p[m.s * dof + 1] = m.W[1]; P.o_p[m.s * dof + 2] = m.W[2]; }
        }
    }

    // tau = S^T f of body k from (f, sincos); returns g = X_lambda^T f and books |tau|^2
    auto finalize = [&](const int k, const bool idk, Smp<FK>& m, const SV6& fk, const double s0, const double c0, const double s1,
                        const double c1, const double s2, const double c2, SV6& g) {
        double t0, t1, t2;
        const double r0 = S.body[k].r[0], r1 = S.body[k].r[1], r2 = S.body[k].r[2];
        if (idk) rnea_backward_identity(r0, r1, r2, fk, g, t0, t1, t2);
        else rnea_backward(r0, r1, r2, s0, c0, s1, c1, s2, c2, fk, g, t0, t1, t2);
        m.tsq = fma(t0, t0, fma(t1, t1, fma(t2, t2, m.tsq)));
        if (OUT && m.owned) { P.o_tau[m.s * dof + 3 * k] = t0; P.o_tau[m.s * dof + 3 * k + 1] = t1; P.o_tau[m.s * dof + 3 * k + 2] = t2; }
    };
    // fold the finished leaf b_last and every finished ancestor back, stopping below `stop`
    auto fold_back = [&](const int b_last, const bool id_last, const int par_last, const int stop) {
        SV6 g[NS];
#pragma unroll
        for (int j = 0; j < NS; ++j)
            finalize(b_last, id_last, sm[j], sm[j].f, sm[j].s0, sm[j].c0, sm[j].s1, sm[j].c1, sm[j].s2, sm[j].c2, g[j]);
        int k = par_last;
#pragma unroll 1
        while (k != stop) {
            const int4 w = S.meta[TREE][k];
            const bool idk = (w.y >> 3) & 1;
            double* __restrict__ sp = stk + (size_t)((w.x >> 8) & 0xff) * (OMO_SLOT * NS * NT) + tid;
#pragma unroll
            for (int j = 0; j < NS; ++j) {
                SV6 fk;
                fk.w0 = sp[(0 * NS + j) * NT] + g[j].w0; fk.w1 = sp[(1 * NS + j) * NT] + g[j].w1; fk.w2 = sp[(2 * NS + j) * NT] + g[j].w2;
                fk.l0 = sp[(3 * NS + j) * NT] + g[j].l0; fk.l1 = sp[(4 * NS + j) * NT] + g[j].l1; fk.l2 = sp[(5 * NS + j) * NT] + g[j].l2;
                if (idk) finalize(k, true, sm[j], fk, 0.0, 1.0, 0.0, 1.0, 0.0, 1.0, g[j]);
                else finalize(k, false, sm[j], fk, sp[(6 * NS + j) * NT], sp[(7 * NS + j) * NT], sp[(8 * NS + j) * NT], sp[(9 * NS + j) * NT],
                              sp[(10 * NS + j) * NT], sp[(11 * NS + j) * NT], g[j]);
            }
            k = w.x & 0xff;
        }
        if (stop != 0) {
            double* __restrict__ sp = stk + (size_t)((S.meta[TREE][stop].x >> 8) & 0xff) * (OMO_SLOT * NS * NT) + tid;
#pragma unroll
            for (int j = 0; j < NS; ++j) {
                sp[(0 * NS + j) * NT] += g[j].w0; sp[(1 * NS + j) * NT] += g[j].w1; sp[(2 * NS + j) * NT] += g[j].w2;
                sp[(3 * NS + j) * NT] += g[j].l0; sp[(4 * NS + j) * NT] += g[j].l1; sp[(5 * NS + j) * NT] += g[j].l2;
            }
        } else if (OUT) {
#pragma unroll
            for (int j = 0; j < NS; ++j) {
                groot[j].w0 += g[j].w0; groot[j].w1 += g[j].w1; groot[j].w2 += g[j].w2;
                groot[j].l0 += g[j].l0; groot[j].l1 += g[j].l1; groot[j].l2 += g[j].l2;
            }
        }
    };

    int xbuf = 0;
    const int n_act = S.n_act[TREE];
    BodyMeta Mp = load_meta(TREE, 0);   // the body visited last (starts at the root: fs = va = -1)
    int bp = 0;
#pragma unroll 1
    for (int ai = 0; ai < n_act; ++ai) {
        const int b = S.act[TREE][ai];
        const BodyMeta M = load_meta(TREE, b);
        const int par = M.parent;
        // ---- leave the previous body: park it (descending) or fold the finished subtree(s) back ----------------
        {
            if (par == bp) {
                if (Mp.fs >= 0) {
                    double* __restrict__ sp = stk + (size_t)Mp.fs * (OMO_SLOT * NS * NT) + tid;
#pragma unroll
                    for (int j = 0; j < NS; ++j) {
                        const Smp<FK>& m = sm[j];
                        sp[(0 * NS + j) * NT] = m.f.w0; sp[(1 * NS + j) * NT] = m.f.w1; sp[(2 * NS + j) * NT] = m.f.w2;
                        sp[(3 * NS + j) * NT] = m.f.l0; sp[(4 * NS + j) * NT] = m.f.l1; sp[(5 * NS + j) * NT] = m.f.l2;
                        if (!Mp.ident) {
                            sp[(6 * NS + j) * NT] = m.s0; sp[(7 * NS + j) * NT] = m.c0; sp[(8 * NS + j) * NT] = m.s1;
                            sp[(9 * NS + j) * NT] = m.c1; sp[(10 * NS + j) * NT] = m.s2; sp[(11 * NS + j) * NT] = m.c2;
                        }
                    }
                }
                if (Mp.va >= 0) {
                    double* __restrict__ sp = stk + (size_t)Mp.va * (OMO_SLOT * NS * NT) + tid;
#pragma unroll
                    for (int j = 0; j < NS; ++j) {
                        const Smp<FK>& m = sm[j];
                        sp[(0 * NS + j) * NT] = m.v.w0; sp[(1 * NS + j) * NT] = m.v.w1; sp[(2 * NS + j) * NT] = m.v.w2;
                        sp[(3 * NS + j) * NT] = m.v.l0; sp[(4 * NS + j) * NT] = m.v.l1; sp[(5 * NS + j) * NT] = m.v.l2;
                        sp[(6 * NS + j) * NT] = m.a.w0; sp[(7 * NS + j) * NT] = m.a.w1; sp[(8 * NS + j) * NT] = m.a.w2;
                        sp[(9 * NS + j) * NT] = m.a.l0; sp[(10 * NS + j) * NT] = m.a.l1; sp[(11 * NS + j) * NT] = m.a.l2;
                    }
                }
                if (FK && Mp.fk >= 0) {
                    double* __restrict__ sp = stk + (size_t)Mp.fk * (OMO_SLOT * NS * NT) + tid;
#pragma unroll
                    for (int j = 0; j < NS; ++j) {
#pragma unroll
                        for (int c = 0; c < 9; ++c) sp[(c * NS + j) * NT] = sm[j].L[c];
#pragma unroll
                        for (int c = 0; c < 3; ++c) sp[((9 + c) * NS + j) * NT] = sm[j].W[c];
                    }
                }
            } else {
                fold_back(bp, Mp.ident, Mp.parent, par);
            }
        }
        if (par != bp) {
            // parent state for b (the root's is a function of the sample index only: recomputed, not parked)
            if (par == 0) {
#pragma unroll
                for (int j = 0; j < NS; ++j) set_root(sm[j]);
            } else {
                const double* __restrict__ sp = stk + (size_t)((S.meta[TREE][par].x >> 16) & 0xff) * (OMO_SLOT * NS * NT) + tid;
#pragma unroll
                for (int j = 0; j < NS; ++j) {
                    Smp<FK>& m = sm[j];
                    m.v.w0 = sp[(0 * NS + j) * NT]; m.v.w1 = sp[(1 * NS + j) * NT]; m.v.w2 = sp[(2 * NS + j) * NT];
                    m.v.l0 = sp[(3 * NS + j) * NT]; m.v.l1 = sp[(4 * NS + j) * NT]; m.v.l2 = sp[(5 * NS + j) * NT];
                    m.a.w0 = sp[(6 * NS + j) * NT]; m.a.w1 = sp[(7 * NS + j) * NT]; m.a.w2 = sp[(8 * NS + j) * NT];
                    m.a.l0 = sp[(9 * NS + j) * NT]; m.a.l1 = sp[(10 * NS + j) * NT]; m.a.l2 = sp[(11 * NS + j) * NT];
                }
            }
            if (FK) {
                const double* __restrict__ sp = stk + (size_t)((S.meta[TREE][par].x >> 24) & 0xff) * (OMO_SLOT * NS * NT) + tid;
#pragma unroll
                for (int j = 0; j < NS; ++j) {
#pragma unroll
                    for (int c = 0; c < 9; ++c) sm[j].L[c] = sp[(c * NS + j) * NT];
#pragma unroll
                    for (int c = 0; c < 3; ++c) sm[j].W[c] = sp[((9 + c) * NS + j) * NT];
                }
            }
        }

        // ---- IK chains rooted here: solve once per sample (joint.cpp:38-44, ikhandler.cpp:103-164) ----------
        if (IK) {
            const int h = M.ik_start;
            if (h >= 0) {
                Smp<FK>& m = sm[0];
                double tw[3];
#pragma unroll
                for (int k = 0; k < 3; ++k)
                    tw[k] = (S.ik_chan_n[h][k] > 0)
                                ? eval_channel(pc, P.n_cp, S.ik_chan_pos[h][k], S.ik_chan_n[h][k], m.t, sorted, n_bez, n_newton)
                                : S.ik_chan_const[h][k];
                double Lp[9], tp[3];
#pragma unroll
                for (int c = 0; c < 9; ++c) Lp[c] = FK ? m.L[c] : 0.0;
#pragma unroll
                for (int c = 0; c < 3; ++c) tp[c] = FK ? m.W[c] : 0.0;
                unsigned it = 0;
                ik_solve(h, Lp, tp, tw, ikq + h * 3 * OMO_MAX_IK_CHAIN, it);
                if (m.owned) { n_ikiter += it; n_iksolve += 1u; }
            }
        }

        // ---- curve group starting here: sample its keyed channels for this thread's times (K-way interleaved) ----
        if (M.grp >= 0) {
            const int g0 = S.grp_begin[M.grp], g1 = S.grp_begin[M.grp + 1];
#pragma unroll 1
            for (int c0 = g0; c0 < g1; c0 += 3) {
                int pos[3], nn[3], row[3], sg[3];
#pragma unroll
                for (int k = 0; k < 3; ++k) {
                    const bool on = c0 + k < g1;
                    const int ch = on ? S.grp_list[c0 + k] : 0;
                    pos[k] = on ? S.chan_pos[ch] : 0; nn[k] = on ? S.chan_n[ch] : 0; row[k] = on ? S.chan_row[ch] : 0;
                    sg[k] = on ? seg0[c0 + k] : 0;
                }
#pragma unroll
                for (int j = 0; j < NS; ++j) {
                    double val[3];
                    unsigned eb = 0, en = 0;
                    eval_channels<3, COUNT>(pc, P.n_cp, pos, nn, sg, sm[j].t, sorted, val, eb, en);
                    if (COUNT && sm[j].owned) { n_bez += eb; n_newton += en; }   // halo samples repeat a neighbour tile's work
#pragma unroll
                    for (int k = 0; k < 3; ++k)
                        if (nn[k] > 0) ex[row[k] * WIN + sm[j].idx] = val[k];
                }
            }
            __syncthreads();
        }

        // ---- forward dynamics step (v, a hold the parent's state on entry, the body's on exit) ------------------
        const int ikh = IK ? M.ik_of : -1;
        if (M.ident) {
            // joint frozen at zero angles: E = I, qd = qdd = 0 exactly
#pragma unroll
            for (int j = 0; j < NS; ++j) {
                Smp<FK>& m = sm[j];
                if (OUT && m.owned) {
                    const size_t o = m.s * dof + 3 * b;
#pragma unroll
                    for (int k = 0; k < 3; ++k) { P.o_q[o + k] = 0.0; P.o_qd[o + k] = 0.0; P.o_qdd[o + k] = 0.0; }
                }
                SV6 vn, an;
                rnea_forward_identity(S.body[b], m.v, m.a, vn, an, m.f);
                m.v = vn; m.a = an;
            }
        } else {
            // angles of body b, their time derivatives (core_rbdl.cpp:28-95), sin / cos
            const int mask = (ikh >= 0) ? 7 : M.dyn;
            double q[NS][3], qd[NS][3], qdd[NS][3], ss[NS][3], cc[NS][3];
            int rowk[3];
            if (IK && ikh >= 0) {
                // IK-driven joint: its angles were solved per thread; exchange them through 3 dedicated double-buffered rows
#pragma unroll
                for (int k = 0; k < 3; ++k) {
                    rowk[k] = S.n_ex_rows + xbuf * 3 + k;
                    ex[rowk[k] * WIN + sm[0].idx] = ikq[ikh * 3 * OMO_MAX_IK_CHAIN + 3 * M.ik_idx + k];
                }
                __syncthreads();
                xbuf ^= 1;
            } else {
#pragma unroll
                for (int k = 0; k < 3; ++k) rowk[k] = M.row[k];
            }
#pragma unroll
            for (int k = 0; k < 3; ++k) {
                if (mask >> k & 1) {
                    const double* __restrict__ row = ex + rowk[k] * WIN;
#pragma unroll
                    for (int j = 0; j < NS; ++j) {
                        const Smp<FK>& m = sm[j];
                        // stencil neighbours inside the block (clamped for the halo samples, whose results are unused)
                        const double m2 = row[max(m.idx - 2, 0)], m1 = row[max(m.idx - 1, 0)], x = row[m.idx];
                        const double p1 = row[min(m.idx + 1, WIN - 1)], p2 = row[min(m.idx + 2, WIN - 1)];
                        q[j][k] = x;
                        if (m.interior) {
                            qd[j][k] = (-p2 + 8.0 * p1 - 8.0 * m1 + m2) * P.inv_12dt;
                            qdd[j][k] = (-p2 + 16.0 * p1 - 30.0 * x + 16.0 * m1 - m2) * P.inv_12dt2;
                        } else if (m.s == 0) {
                            qd[j][k] = (p1 - x) * P.inv_dt;
                            qdd[j][k] = (p2 - 2.0 * p1 + x) * P.inv_2dt2;
                        } else if (m.s == N - 1) {
                            qd[j][k] = (x - m1) * P.inv_dt;
                            qdd[j][k] = (x - 2.0 * m1 + m2) * P.inv_2dt2;
                        } else {
                            qd[j][k] = (p1 - m1) * P.inv_2dt;
                            qdd[j][k] = (p1 - 2.0 * x + m1) * P.inv_dt2;
                        }
                        sincos_angle(x, &ss[j][k], &cc[j][k]);
                    }
                } else {
                    const double qc = S.chan_const[3 * b + k], sc_ = S.chan_sin[3 * b + k], cc_ = S.chan_cos[3 * b + k];
                    const double rd = S.chan_qd[3 * b + k], rdd = S.chan_qdd[3 * b + k];
#pragma unroll
                    for (int j = 0; j < NS; ++j) {
                        q[j][k] = qc; ss[j][k] = sc_; cc[j][k] = cc_;
                        qd[j][k] = sm[j].interior ? rd : 0.0;
                        qdd[j][k] = sm[j].interior ? rdd : 0.0;
                    }
                }
            }
#pragma unroll
            for (int j = 0; j < NS; ++j) {
                Smp<FK>& m = sm[j];
                m.s0 = ss[j][0]; m.s1 = ss[j][1]; m.s2 = ss[j][2];
                m.c0 = cc[j][0]; m.c1 = cc[j][1]; m.c2 = cc[j][2];
                if (OUT && m.owned) {
                    const size_t o = m.s * dof + 3 * b;
#pragma unroll
                    for (int k = 0; k < 3; ++k) { P.o_q[o + k] = q[j][k]; P.o_qd[o + k] = qd[j][k]; P.o_qdd[o + k] = qdd[j][k]; }
                }
                SV6 vn, an;
                rnea_forward(S.body[b], m.s0, m.c0, m.s1, m.c1, m.s2, m.c2, qd[j][0], qd[j][1], qd[j][2], qdd[j][0], qdd[j][1], qdd[j][2],
                             m.v, m.a, vn, an, m.f);
                m.v = vn; m.a = an;
            }
        }

        // ---- forward kinematics: p_b and the world affine handed to b's children ----------------------------------
        if (FK) {
            const double o0 = S.offset[b][0], o1 = S.offset[b][1], o2 = S.offset[b][2];
#pragma unroll
            for (int j = 0; j < NS; ++j) {
                Smp<FK>& m = sm[j];
                const double p0 = fma(m.L[0], o0, fma(m.L[1], o1, fma(m.L[2], o2, m.W[0])));
                const double p1 = fma(m.L[3], o0, fma(m.L[4], o1, fma(m.L[5], o2, m.W[1])));
                const double p2 = fma(m.L[6], o0, fma(m.L[7], o1, fma(m.L[8], o2, m.W[2])));
                if (P.p_orig) {
                    const double d0 = p0 - __ldg(P.p_orig + (size_t)(3 * b) * N + m.s), d1 = p1 - __ldg(P.p_orig + (size_t)(3 * b + 1) * N + m.s),
                                 d2 = p2 - __ldg(P.p_orig + (size_t)(3 * b + 2) * N + m.s);
                    m.dtr += d0 * d0 + d1 * d1 + d2 * d2;
                }
                if (OUT && m.owned && P.o_p) { P.o_p[m.s * dof + 3 * b] = p0; P.o_p[m.s * dof + 3 * b + 1] = p1; P.o_p[m.s * dof + 3 * b + 2] = p2; }
                if (M.has_child) {
                    if (!M.ident) {
                        // L <- L * Rz(q2) Ry(q1) Rx(q0)
                        const double s0 = m.s0, c0 = m.c0, s1 = m.s1, c1 = m.c1, s2 = m.s2, c2 = m.c2;
                        const double R00 = c2 * c1, R01 = c2 * s1 * s0 - s2 * c0, R02 = c2 * s1 * c0 + s2 * s0;
                        const double R10 = s2 * c1, R11 = s2 * s1 * s0 + c2 * c0, R12 = s2 * s1 * c0 - c2 * s0;
                        const double R20 = -s1, R21 = c1 * s0, R22 = c1 * c0;
                        double n[9];
#pragma unroll
                        for (int r = 0; r < 3; ++r) {
                            n[3 * r] = fma(m.L[3 * r], R00, fma(m.L[3 * r + 1], R10, m.L[3 * r + 2] * R20));
                            n[3 * r + 1] = fma(m.L[3 * r], R01, fma(m.L[3 * r + 1], R11, m.L[3 * r + 2] * R21));
                            n[3 * r + 2] = fma(m.L[3 * r], R02, fma(m.L[3 * r + 1], R12, m.L[3 * r + 2] * R22));
                        }
#pragma unroll
                        for (int c = 0; c < 9; ++c) m.L[c] = n[c];
                    }
                    m.W[0] = p0; m.W[1] = p1; m.W[2] = p2;
                }
            }
        }
        Mp = M; bp = b;
    }

    // ---- fold everything that is still open back to the root ------------------------------------------------------
    if (n_act > 0) fold_back(bp, Mp.ident, Mp.parent, 0);

    // ---- cost of the samples and block reduction (local_cost = |tau[3:]|^2 dt, core_rbdl.cpp:331-334) ---------------
    double wc = 0.0, wt = 0.0;
#pragma unroll
    for (int j = 0; j < NS; ++j) {
        const Smp<FK>& m = sm[j];
        const double local = m.tsq * P.dt;
        if (OUT && m.owned) {
            // tau[0:3] of the TranslationXYZ root = linear part of the force folded into body 1 (Dynamics.cc:91-92)
            P.o_tau[m.s * dof] = groot[j].l0; P.o_tau[m.s * dof + 1] = groot[j].l1; P.o_tau[m.s * dof + 2] = groot[j].l2;
            P.o_local[m.s] = local;
        }
        if (m.owned) {
            wc += __ldg(P.wcost + m.s) * local;
            if (FK) wt += P.w_traj * m.dtr * P.dt;
        }
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
    if (tid == 0) {
        double C = 0.0, D = 0.0;
        for (int w = 0; w < (NT >> 5); ++w) { C += red[2 * w]; D += red[2 * w + 1]; }
        P.partial[(size_t)blockIdx.x * 2] = C;
        P.partial[(size_t)blockIdx.x * 2 + 1] = D;
    }
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

}  // namespace omo
#endif
