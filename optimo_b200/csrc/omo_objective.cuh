// objective_kernel: the fused sampling -> stencil -> RNEA -> cost kernel (see omo_kernels.cuh for the mapping).
//
// Restates, per (candidate, time sample):
//   Core::SetStates           OptiMo/core_rbdl.cpp:108-166   (q from curves / IK, p from FK)
//   CalculateFirst/SecondCentral  core_rbdl.cpp:28-95        (SECOND_ORDER_ACCURATE off)
//   Joint::GetAffineRelativeTo{Parent,World}  OptiMo/joint.cpp:64-83   (T(offset) Rz Ry Rx)
//   Core::CalculateCost       core_rbdl.cpp:307-346          (local_cost = |tau[3:]|^2 dt, weight w(t))
//   D_traj                    core_rbdl.cpp:387-398
#ifndef OMO_OBJECTIVE_CUH
#define OMO_OBJECTIVE_CUH

#include "omo_kernels.cuh"

namespace omo {

#define OMO_STK(slot, comp) stk[((slot) * OMO_SLOT + (comp)) * NT + tid]

template <bool FK, bool IK, bool OUT>
__global__ void __launch_bounds__(128) objective_kernel(const OmoEvalParams P)
{
    extern __shared__ double smem[];
    const OmoDevScene& S = c_scene;
    const int NT = blockDim.x;
    const int tid = threadIdx.x;
    const int cand = blockIdx.x / P.n_tiles;
    const int tile = blockIdx.x - cand * P.n_tiles;
    double* __restrict__ ex = smem;            // [2][3][NT] stencil exchange
    double* __restrict__ stk = smem + 6 * NT;  // [n_slots][OMO_SLOT][NT]

    const int N = P.N_s;
    const int s_raw = tile * P.T_out - 2 + tid;
    const int s = min(max(s_raw, 0), N - 1);
    const int tile_end = min(N, (tile + 1) * P.T_out);
    const bool owned = (tid >= 2) && (s_raw < tile_end) && (s_raw >= 0);
    const bool interior = (s >= 2) && (s <= N - 3);
    const double t = __ldg(P.tgrid + s);
    const double* __restrict__ pc = P.prep + (size_t)cand * 6 * P.n_cp;
    const bool sorted = (__ldg(P.flags + cand) & 1u) != 0u;
    // stencil neighbours inside the block (clamped for the halo threads, whose results are unused)
    const int im2 = max(tid - 2, 0), im1 = max(tid - 1, 0), ip1 = min(tid + 1, NT - 1), ip2 = min(tid + 2, NT - 1);

    unsigned n_bez = 0, n_newton = 0, n_iksolve = 0, n_ikiter = 0;
    double tsq = 0.0;   // sum of tau^2 over the non-root DoFs
    double dtr = 0.0;   // |p - p_original|^2
    double ikq[IK ? OMO_MAX_IK * 3 * OMO_MAX_IK_CHAIN : 1];

    // ---- root (RBDL body 1, TranslationXYZ, q = root offset: core_rbdl.cpp:137-141,212) -------------
    SV6 v, a, f;  // state of the body visited last
    v.w0 = v.w1 = v.w2 = 0.0; a.w0 = a.w1 = a.w2 = 0.0;
    v.l0 = interior ? S.root_qd[0] : 0.0; v.l1 = interior ? S.root_qd[1] : 0.0; v.l2 = interior ? S.root_qd[2] : 0.0;
    a.l0 = interior ? S.root_qdd[0] : 0.0; a.l1 = interior ? S.root_qdd[1] : 0.0; a.l2 = interior ? S.root_qdd[2] : 0.0;
    f.w0 = f.w1 = f.w2 = f.l0 = f.l1 = f.l2 = 0.0;
    double s0 = 0.0, c0 = 1.0, s1 = 0.0, c1 = 1.0, s2 = 0.0, c2 = 1.0;
    double L00 = 1, L01 = 0, L02 = 0, L10 = 0, L11 = 1, L12 = 0, L20 = 0, L21 = 0, L22 = 1, W0 = 0, W1 = 0, W2 = 0;  // world affine
    SV6 groot; groot.w0 = groot.w1 = groot.w2 = groot.l0 = groot.l1 = groot.l2 = 0.0;

    if (OUT && owned) {
        for (int k = 0; k < 3; ++k) {
            P.o_q[(size_t)s * 3 * S.J + k] = S.offset[0][k];
            P.o_qd[(size_t)s * 3 * S.J + k] = interior ? S.root_qd[k] : 0.0;
            P.o_qdd[(size_t)s * 3 * S.J + k] = interior ? S.root_qdd[k] : 0.0;
        }
    }
    if (FK) {
        // root world affine = T(offset) Rz Ry Rx with the root's own rot channels (joint.cpp:64-74)
        double e[3];
#pragma unroll
        for (int k = 0; k < 3; ++k)
            e[k] = (S.dyn[0] >> k & 1) ? eval_channel(pc, P.n_cp, S.chan_pos[k], S.chan_n[k], t, sorted, n_bez, n_newton)
                                       : S.chan_const[k];
        double sx, cx, sy, cy, sz, cz;
        sincos(e[0], &sx, &cx); sincos(e[1], &sy, &cy); sincos(e[2], &sz, &cz);
        L00 = cz * cy; L01 = cz * sy * sx - sz * cx; L02 = cz * sy * cx + sz * sx;
        L10 = sz * cy; L11 = sz * sy * sx + cz * cx; L12 = sz * sy * cx - cz * sx;
        L20 = -sy;     L21 = cy * sx;                L22 = cy * cx;
        W0 = S.offset[0][0]; W1 = S.offset[0][1]; W2 = S.offset[0][2];
        if (P.p_orig) {
            const double d0 = W0 - __ldg(P.p_orig + (size_t)0 * N + s), d1 = W1 - __ldg(P.p_orig + (size_t)1 * N + s),
                         d2 = W2 - __ldg(P.p_orig + (size_t)2 * N + s);
            dtr += d0 * d0 + d1 * d1 + d2 * d2;
        }
        if (OUT && owned && P.o_p) { P.o_p[(size_t)s * 3 * S.J] = W0; P.o_p[(size_t)s * 3 * S.J + 1] = W1; P.o_p[(size_t)s * 3 * S.J + 2] = W2; }
    }

    int xbuf = 0;
    const int J = S.J;
#pragma unroll 1
    for (int b = 1; b < J; ++b) {
        const int par = S.parent[b];
        // ---- leave body b-1: park it (descending) or fold the finished subtree(s) back ----------------
        if (par == b - 1) {
            const int fs = S.fs_slot[b - 1], va = S.va_slot[b - 1];
            if (fs >= 0) {
                OMO_STK(fs, 0) = f.w0; OMO_STK(fs, 1) = f.w1; OMO_STK(fs, 2) = f.w2;
                OMO_STK(fs, 3) = f.l0; OMO_STK(fs, 4) = f.l1; OMO_STK(fs, 5) = f.l2;
                if (!S.ident[b - 1]) {
                    OMO_STK(fs, 6) = s0; OMO_STK(fs, 7) = c0; OMO_STK(fs, 8) = s1;
                    OMO_STK(fs, 9) = c1; OMO_STK(fs, 10) = s2; OMO_STK(fs, 11) = c2;
                }
            }
            if (va >= 0) {
                OMO_STK(va, 0) = v.w0; OMO_STK(va, 1) = v.w1; OMO_STK(va, 2) = v.w2;
                OMO_STK(va, 3) = v.l0; OMO_STK(va, 4) = v.l1; OMO_STK(va, 5) = v.l2;
                OMO_STK(va, 6) = a.w0; OMO_STK(va, 7) = a.w1; OMO_STK(va, 8) = a.w2;
                OMO_STK(va, 9) = a.l0; OMO_STK(va, 10) = a.l1; OMO_STK(va, 11) = a.l2;
            }
            if (FK) {
                const int fk = S.fk_slot[b - 1];
                if (fk >= 0) {
                    OMO_STK(fk, 0) = L00; OMO_STK(fk, 1) = L01; OMO_STK(fk, 2) = L02;
                    OMO_STK(fk, 3) = L10; OMO_STK(fk, 4) = L11; OMO_STK(fk, 5) = L12;
                    OMO_STK(fk, 6) = L20; OMO_STK(fk, 7) = L21; OMO_STK(fk, 8) = L22;
                    OMO_STK(fk, 9) = W0; OMO_STK(fk, 10) = W1; OMO_STK(fk, 11) = W2;
                }
            }
        } else {
            SV6 g;
            double t0, t1, t2;
            int k = b - 1;
            if (S.ident[k]) rnea_backward_identity(S.body[k].r[0], S.body[k].r[1], S.body[k].r[2], f, g, t0, t1, t2);
            else rnea_backward(S.body[k].r[0], S.body[k].r[1], S.body[k].r[2], s0, c0, s1, c1, s2, c2, f, g, t0, t1, t2);
            tsq += t0 * t0 + t1 * t1 + t2 * t2;
            if (OUT && owned) { P.o_tau[(size_t)s * 3 * J + 3 * k] = t0; P.o_tau[(size_t)s * 3 * J + 3 * k + 1] = t1; P.o_tau[(size_t)s * 3 * J + 3 * k + 2] = t2; }
            k = S.parent[k];
#pragma unroll 1
            while (k != par) {
                const int fs = S.fs_slot[k];
                SV6 fk_;
                fk_.w0 = OMO_STK(fs, 0) + g.w0; fk_.w1 = OMO_STK(fs, 1) + g.w1; fk_.w2 = OMO_STK(fs, 2) + g.w2;
                fk_.l0 = OMO_STK(fs, 3) + g.l0; fk_.l1 = OMO_STK(fs, 4) + g.l1; fk_.l2 = OMO_STK(fs, 5) + g.l2;
                if (S.ident[k]) rnea_backward_identity(S.body[k].r[0], S.body[k].r[1], S.body[k].r[2], fk_, g, t0, t1, t2);
                else rnea_backward(S.body[k].r[0], S.body[k].r[1], S.body[k].r[2], OMO_STK(fs, 6), OMO_STK(fs, 7), OMO_STK(fs, 8),
                                   OMO_STK(fs, 9), OMO_STK(fs, 10), OMO_STK(fs, 11), fk_, g, t0, t1, t2);
                tsq += t0 * t0 + t1 * t1 + t2 * t2;
                if (OUT && owned) { P.o_tau[(size_t)s * 3 * J + 3 * k] = t0; P.o_tau[(size_t)s * 3 * J + 3 * k + 1] = t1; P.o_tau[(size_t)s * 3 * J + 3 * k + 2] = t2; }
                k = S.parent[k];
            }
            if (par != 0) {
                const int fs = S.fs_slot[par];
                OMO_STK(fs, 0) += g.w0; OMO_STK(fs, 1) += g.w1; OMO_STK(fs, 2) += g.w2;
                OMO_STK(fs, 3) += g.l0; OMO_STK(fs, 4) += g.l1; OMO_STK(fs, 5) += g.l2;
            } else if (OUT) {
                groot.w0 += g.w0; groot.w1 += g.w1; groot.w2 += g.w2; groot.l0 += g.l0; groot.l1 += g.l1; groot.l2 += g.l2;
            }
            // parent state for b (the root's is a function of the sample index only: recomputed, not parked)
            if (par == 0) {
                v.w0 = v.w1 = v.w2 = 0.0; a.w0 = a.w1 = a.w2 = 0.0;
                v.l0 = interior ? S.root_qd[0] : 0.0; v.l1 = interior ? S.root_qd[1] : 0.0; v.l2 = interior ? S.root_qd[2] : 0.0;
                a.l0 = interior ? S.root_qdd[0] : 0.0; a.l1 = interior ? S.root_qdd[1] : 0.0; a.l2 = interior ? S.root_qdd[2] : 0.0;
            } else {
                const int va = S.va_slot[par];
                v.w0 = OMO_STK(va, 0); v.w1 = OMO_STK(va, 1); v.w2 = OMO_STK(va, 2);
                v.l0 = OMO_STK(va, 3); v.l1 = OMO_STK(va, 4); v.l2 = OMO_STK(va, 5);
                a.w0 = OMO_STK(va, 6); a.w1 = OMO_STK(va, 7); a.w2 = OMO_STK(va, 8);
                a.l0 = OMO_STK(va, 9); a.l1 = OMO_STK(va, 10); a.l2 = OMO_STK(va, 11);
            }
            if (FK) {
                const int fk = S.fk_slot[par];
                L00 = OMO_STK(fk, 0); L01 = OMO_STK(fk, 1); L02 = OMO_STK(fk, 2);
                L10 = OMO_STK(fk, 3); L11 = OMO_STK(fk, 4); L12 = OMO_STK(fk, 5);
                L20 = OMO_STK(fk, 6); L21 = OMO_STK(fk, 7); L22 = OMO_STK(fk, 8);
                W0 = OMO_STK(fk, 9); W1 = OMO_STK(fk, 10); W2 = OMO_STK(fk, 11);
            }
        }

        // ---- IK chains rooted here: solve once per sample (joint.cpp:38-44, ikhandler.cpp:103-164) ------
        if (IK) {
            const int h = S.ik_start[b];
            if (h >= 0) {
                double tw[3];
#pragma unroll
                for (int k = 0; k < 3; ++k)
                    tw[k] = (S.ik_chan_n[h][k] > 0)
                                ? eval_channel(pc, P.n_cp, S.ik_chan_pos[h][k], S.ik_chan_n[h][k], t, sorted, n_bez, n_newton)
                                : S.ik_chan_const[h][k];
                const double Lp[9] = {L00, L01, L02, L10, L11, L12, L20, L21, L22};
                const double tp[3] = {W0, W1, W2};
                ik_solve(h, Lp, tp, tw, ikq + h * 3 * OMO_MAX_IK_CHAIN, n_ikiter);
                n_iksolve += 1u;
            }
        }

        // ---- angles of body b, their time derivatives, sin/cos ------------------------------------------
        double q0, q1, q2, qd0, qd1, qd2, qdd0, qdd1, qdd2;
        const int dyn = S.dyn[b];
        const int ikh = IK ? S.ik_of[b] : -1;
        const bool ident = S.ident[b] != 0;
        if (ident) {
            q0 = q1 = q2 = 0.0; qd0 = qd1 = qd2 = 0.0; qdd0 = qdd1 = qdd2 = 0.0;
            s0 = s1 = s2 = 0.0; c0 = c1 = c2 = 1.0;
        } else if (dyn == 0 && ikh < 0) {
            q0 = S.chan_const[3 * b]; q1 = S.chan_const[3 * b + 1]; q2 = S.chan_const[3 * b + 2];
            s0 = S.chan_sin[3 * b]; s1 = S.chan_sin[3 * b + 1]; s2 = S.chan_sin[3 * b + 2];
            c0 = S.chan_cos[3 * b]; c1 = S.chan_cos[3 * b + 1]; c2 = S.chan_cos[3 * b + 2];
            qd0 = interior ? S.chan_qd[3 * b] : 0.0; qd1 = interior ? S.chan_qd[3 * b + 1] : 0.0; qd2 = interior ? S.chan_qd[3 * b + 2] : 0.0;
            qdd0 = interior ? S.chan_qdd[3 * b] : 0.0; qdd1 = interior ? S.chan_qdd[3 * b + 1] : 0.0; qdd2 = interior ? S.chan_qdd[3 * b + 2] : 0.0;
        } else {
            double qq[3], qqd[3], qqdd[3], ss[3], cc[3];
            const int mask = (ikh >= 0) ? 7 : dyn;
            double* __restrict__ exb = ex + xbuf * 3 * NT;
#pragma unroll
            for (int k = 0; k < 3; ++k) {
                if (mask >> k & 1) {
                    if (IK && ikh >= 0) qq[k] = ikq[ikh * 3 * OMO_MAX_IK_CHAIN + 3 * S.ik_idx[b] + k];
                    else qq[k] = eval_channel(pc, P.n_cp, S.chan_pos[3 * b + k], S.chan_n[3 * b + k], t, sorted, n_bez, n_newton);
                    exb[k * NT + tid] = qq[k];
                }
            }
            __syncthreads();
#pragma unroll
            for (int k = 0; k < 3; ++k) {
                if (mask >> k & 1) {
                    const double m2 = exb[k * NT + im2], m1 = exb[k * NT + im1], p1 = exb[k * NT + ip1], p2 = exb[k * NT + ip2];
                    const double x = qq[k];
                    if (interior) {
                        qqd[k] = (-p2 + 8.0 * p1 - 8.0 * m1 + m2) * P.inv_12dt;
                        qqdd[k] = (-p2 + 16.0 * p1 - 30.0 * x + 16.0 * m1 - m2) * P.inv_12dt2;
                    } else if (s == 0) {
                        qqd[k] = (p1 - x) * P.inv_dt;
                        qqdd[k] = (p2 - 2.0 * p1 + x) * P.inv_2dt2;
                    } else if (s == N - 1) {
                        qqd[k] = (x - m1) * P.inv_dt;
                        qqdd[k] = (x - 2.0 * m1 + m2) * P.inv_2dt2;
                    } else {
                        qqd[k] = (p1 - m1) * P.inv_2dt;
                        qqdd[k] = (p1 - 2.0 * x + m1) * P.inv_dt2;
                    }
                    sincos(x, &ss[k], &cc[k]);
                } else {
                    qq[k] = S.chan_const[3 * b + k];
                    ss[k] = S.chan_sin[3 * b + k]; cc[k] = S.chan_cos[3 * b + k];
                    qqd[k] = interior ? S.chan_qd[3 * b + k] : 0.0;
                    qqdd[k] = interior ? S.chan_qdd[3 * b + k] : 0.0;
                }
            }
            xbuf ^= 1;
            q0 = qq[0]; q1 = qq[1]; q2 = qq[2];
            qd0 = qqd[0]; qd1 = qqd[1]; qd2 = qqd[2];
            qdd0 = qqdd[0]; qdd1 = qqdd[1]; qdd2 = qqdd[2];
            s0 = ss[0]; s1 = ss[1]; s2 = ss[2]; c0 = cc[0]; c1 = cc[1]; c2 = cc[2];
        }
        if (OUT && owned) {
            const size_t o = (size_t)s * 3 * J + 3 * b;
            P.o_q[o] = q0; P.o_q[o + 1] = q1; P.o_q[o + 2] = q2;
            P.o_qd[o] = qd0; P.o_qd[o + 1] = qd1; P.o_qd[o + 2] = qd2;
            P.o_qdd[o] = qdd0; P.o_qdd[o + 1] = qdd1; P.o_qdd[o + 2] = qdd2;
        }

        // ---- forward dynamics step (v, a hold the parent's state on entry, the body's on exit) ----------
        {
            SV6 vn, an;
            if (ident) rnea_forward_identity(S.body[b], v, a, vn, an, f);
            else rnea_forward(S.body[b], s0, c0, s1, c1, s2, c2, qd0, qd1, qd2, qdd0, qdd1, qdd2, v, a, vn, an, f);
            v = vn; a = an;
        }

        // ---- forward kinematics: p_b and the world affine handed to b's children --------------------------
        if (FK) {
            const double o0 = S.offset[b][0], o1 = S.offset[b][1], o2 = S.offset[b][2];
            const double p0 = L00 * o0 + L01 * o1 + L02 * o2 + W0;
            const double p1 = L10 * o0 + L11 * o1 + L12 * o2 + W1;
            const double p2 = L20 * o0 + L21 * o1 + L22 * o2 + W2;
            if (P.p_orig) {
                const double d0 = p0 - __ldg(P.p_orig + (size_t)(3 * b) * N + s), d1 = p1 - __ldg(P.p_orig + (size_t)(3 * b + 1) * N + s),
                             d2 = p2 - __ldg(P.p_orig + (size_t)(3 * b + 2) * N + s);
                dtr += d0 * d0 + d1 * d1 + d2 * d2;
            }
            if (OUT && owned && P.o_p) { P.o_p[(size_t)s * 3 * J + 3 * b] = p0; P.o_p[(size_t)s * 3 * J + 3 * b + 1] = p1; P.o_p[(size_t)s * 3 * J + 3 * b + 2] = p2; }
            if (S.n_child[b] > 0) {
                // R = Rz(q2) Ry(q1) Rx(q0)
                const double R00 = c2 * c1, R01 = c2 * s1 * s0 - s2 * c0, R02 = c2 * s1 * c0 + s2 * s0;
                const double R10 = s2 * c1, R11 = s2 * s1 * s0 + c2 * c0, R12 = s2 * s1 * c0 - c2 * s0;
                const double R20 = -s1, R21 = c1 * s0, R22 = c1 * c0;
                const double n00 = L00 * R00 + L01 * R10 + L02 * R20, n01 = L00 * R01 + L01 * R11 + L02 * R21, n02 = L00 * R02 + L01 * R12 + L02 * R22;
                const double n10 = L10 * R00 + L11 * R10 + L12 * R20, n11 = L10 * R01 + L11 * R11 + L12 * R21, n12 = L10 * R02 + L11 * R12 + L12 * R22;
                const double n20 = L20 * R00 + L21 * R10 + L22 * R20, n21 = L20 * R01 + L21 * R11 + L22 * R21, n22 = L20 * R02 + L21 * R12 + L22 * R22;
                L00 = n00; L01 = n01; L02 = n02; L10 = n10; L11 = n11; L12 = n12; L20 = n20; L21 = n21; L22 = n22;
                W0 = p0; W1 = p1; W2 = p2;
            }
        }
    }

    // ---- fold everything that is still open back to the root -------------------------------------------
    if (J > 1) {
        SV6 g;
        double t0, t1, t2;
        int k = J - 1;
        if (S.ident[k]) rnea_backward_identity(S.body[k].r[0], S.body[k].r[1], S.body[k].r[2], f, g, t0, t1, t2);
            else rnea_backward(S.body[k].r[0], S.body[k].r[1], S.body[k].r[2], s0, c0, s1, c1, s2, c2, f, g, t0, t1, t2);
        tsq += t0 * t0 + t1 * t1 + t2 * t2;
        if (OUT && owned) { P.o_tau[(size_t)s * 3 * J + 3 * k] = t0; P.o_tau[(size_t)s * 3 * J + 3 * k + 1] = t1; P.o_tau[(size_t)s * 3 * J + 3 * k + 2] = t2; }
        k = S.parent[k];
#pragma unroll 1
        while (k != 0) {
            const int fs = S.fs_slot[k];
            SV6 fk_;
            fk_.w0 = OMO_STK(fs, 0) + g.w0; fk_.w1 = OMO_STK(fs, 1) + g.w1; fk_.w2 = OMO_STK(fs, 2) + g.w2;
            fk_.l0 = OMO_STK(fs, 3) + g.l0; fk_.l1 = OMO_STK(fs, 4) + g.l1; fk_.l2 = OMO_STK(fs, 5) + g.l2;
            if (S.ident[k]) rnea_backward_identity(S.body[k].r[0], S.body[k].r[1], S.body[k].r[2], fk_, g, t0, t1, t2);
                else rnea_backward(S.body[k].r[0], S.body[k].r[1], S.body[k].r[2], OMO_STK(fs, 6), OMO_STK(fs, 7), OMO_STK(fs, 8),
                                   OMO_STK(fs, 9), OMO_STK(fs, 10), OMO_STK(fs, 11), fk_, g, t0, t1, t2);
            tsq += t0 * t0 + t1 * t1 + t2 * t2;
            if (OUT && owned) { P.o_tau[(size_t)s * 3 * J + 3 * k] = t0; P.o_tau[(size_t)s * 3 * J + 3 * k + 1] = t1; P.o_tau[(size_t)s * 3 * J + 3 * k + 2] = t2; }
            k = S.parent[k];
        }
        if (OUT) { groot.w0 += g.w0; groot.w1 += g.w1; groot.w2 += g.w2; groot.l0 += g.l0; groot.l1 += g.l1; groot.l2 += g.l2; }
    }

    // ---- cost of this sample and block reduction (local_cost = |tau[3:]|^2 dt, core_rbdl.cpp:331-334) ------
    const double local = tsq * P.dt;
    if (OUT && owned) {
        // tau[0:3] of the TranslationXYZ root = linear part of the force folded into body 1 (Dynamics.cc:91-92)
        P.o_tau[(size_t)s * 3 * J] = groot.l0; P.o_tau[(size_t)s * 3 * J + 1] = groot.l1; P.o_tau[(size_t)s * 3 * J + 2] = groot.l2;
        P.o_local[s] = local;
    }
    double wc = owned ? __ldg(P.wcost + s) * local : 0.0;
    double wt = (FK && owned) ? P.w_traj * dtr * P.dt : 0.0;
    __syncthreads();  // everyone is done with smem
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
    if (P.counters) {
        unsigned cb = owned ? n_bez : 0u, cn = owned ? n_newton : 0u, ci = owned ? n_iksolve : 0u, cj = owned ? n_ikiter : 0u;
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
