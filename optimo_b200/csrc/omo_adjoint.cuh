// Analytic gradient of the objective: the derivative of
//     f(x) = sum_i w_i dt |tau_i[3:]|^2  +  D_param  +  sum_i w_traj dt |p_i - p0_i|^2          (OptiMo/core_rbdl.cpp:381-400)
// with respect to the curve parameters, without the 2F+1 objective evaluations of the finite-difference gradient.
//
//   tau_i = InverseDynamics(q_i, qd_i, qdd_i)      external/rbdl-v2.5.0/src/Dynamics.cc:26-105 (EulerXYZ bodies, Joint.cc:100-139)
//   qd, qdd = stencils of q over the samples       OptiMo/core_rbdl.cpp:28-95
//   q_i[c] = Bezier value of channel c at t_i      OptiMo/curve.cpp:57-93 (u from the time cubic, curve.cpp:43-54)
//
// Chain rule in four kernels (one objective's worth of samples, no batch):
//   1. tangent_kernel   one thread per (sample, keyed channel, kind in {q, qd, qdd}): the RNEA of the bodies that channel can
//                       reach -- the joint's subtree, then the force up its ancestor chain -- in forward-mode (dual number)
//                       arithmetic; gives L[kind][c][i] = d(cost of sample i) / d(kind of channel c at sample i).
//   2. stencil_kernel   G[c][i] = L_q + (stencils)^T L_qd + (stencils)^T L_qdd: the transposed difference operators, including
//                       the reference's lower-order boundary forms.
//   3. bezier_kernel    per (sample, channel): segment, Newton root u (as the reference finds it), the partials of the value with
//                       respect to the ten raw parameters of the segment's two control points (implicit-function theorem for u).
//   4. reduce_kernel    per control-point field: ordered sum over the samples of its two adjacent segments (deterministic, no
//                       atomics), cyclic end tangents shared, + 2 w (x - x0) of D_param.
// Everything here is generic in the scalar type so that the same source is compiled for the host by tests/native (the
// arithmetic is checked against finite differences on the CPU, without a GPU) -- the product only ever runs the CUDA kernels.
#ifndef OMO_ADJOINT_CUH
#define OMO_ADJOINT_CUH

#include "omo_scene.cuh"

#if defined(__CUDACC__)
#define OMO_HD __host__ __device__ __forceinline__
#else
#define OMO_HD inline
#include <cmath>
#endif

namespace omo_ad {

// value + one directional derivative
struct Dual {
    double v, d;
    OMO_HD Dual() : v(0.0), d(0.0) {}
    OMO_HD Dual(double v_) : v(v_), d(0.0) {}
    OMO_HD Dual(double v_, double d_) : v(v_), d(d_) {}
};
OMO_HD Dual operator+(const Dual& a, const Dual& b) { return Dual(a.v + b.v, a.d + b.d); }
OMO_HD Dual operator-(const Dual& a, const Dual& b) { return Dual(a.v - b.v, a.d - b.d); }
OMO_HD Dual operator-(const Dual& a) { return Dual(-a.v, -a.d); }
OMO_HD Dual operator*(const Dual& a, const Dual& b) { return Dual(a.v * b.v, a.v * b.d + a.d * b.v); }
OMO_HD Dual operator*(double a, const Dual& b) { return Dual(a * b.v, a * b.d); }
OMO_HD Dual operator*(const Dual& a, double b) { return Dual(a.v * b, a.d * b); }
OMO_HD Dual operator+(const Dual& a, double b) { return Dual(a.v + b, a.d); }
OMO_HD Dual operator+(double a, const Dual& b) { return Dual(a + b.v, b.d); }
OMO_HD Dual operator-(const Dual& a, double b) { return Dual(a.v - b, a.d); }
OMO_HD double value_of(const Dual& a) { return a.v; }
OMO_HD double value_of(double a) { return a; }
OMO_HD void sincos_of(const Dual& a, Dual& s, Dual& c)
{
    const double sv = sin(a.v), cv = cos(a.v);
    s = Dual(sv, cv * a.d);
    c = Dual(cv, -sv * a.d);
}
OMO_HD void sincos_of(double a, double& s, double& c) { s = sin(a); c = cos(a); }

template <class T> struct V6 { T w0, w1, w2, l0, l1, l2; };
template <class T> OMO_HD void zero(V6<T>& x) { x.w0 = x.w1 = x.w2 = x.l0 = x.l1 = x.l2 = T(0.0); }
template <class T> OMO_HD void add_to(V6<T>& y, const V6<T>& x)
{
    y.w0 = y.w0 + x.w0; y.w1 = y.w1 + x.w1; y.w2 = y.w2 + x.w2; y.l0 = y.l0 + x.l0; y.l1 = y.l1 + x.l1; y.l2 = y.l2 + x.l2;
}

// E of the EulerXYZ joint (Joint.cc:112-116); row-major e[3*r + c]
template <class T> OMO_HD void euler_E(const T s0, const T c0, const T s1, const T c1, const T s2, const T c2, T (&e)[9])
{
    e[0] = c2 * c1;    e[1] = c2 * s1 * s0 + s2 * c0;    e[2] = -(c2 * s1 * c0) + s2 * s0;
    e[3] = -(s2 * c1); e[4] = -(s2 * s1 * s0) + c2 * c0; e[5] = s2 * s1 * c0 + c2 * s0;
    e[6] = s1;         e[7] = -(c1 * s0);                e[8] = c1 * c0;
}

// y = I x (SpatialAlgebraOperators.h:51-65)
template <class T> OMO_HD void rbi_mul(const OmoBody& B, const V6<T>& x, V6<T>& y)
{
    const double h0 = B.h[0], h1 = B.h[1], h2 = B.h[2], m = B.m;
    y.w0 = B.Ixx * x.w0 + B.Iyx * x.w1 + B.Izx * x.w2 + h1 * x.l2 - h2 * x.l1;
    y.w1 = B.Iyx * x.w0 + B.Iyy * x.w1 + B.Izy * x.w2 + h2 * x.l0 - h0 * x.l2;
    y.w2 = B.Izx * x.w0 + B.Izy * x.w1 + B.Izz * x.w2 + h0 * x.l1 - h1 * x.l0;
    y.l0 = m * x.l0 + h2 * x.w1 - h1 * x.w2;
    y.l1 = m * x.l1 + h0 * x.w2 - h2 * x.w0;
    y.l2 = m * x.l2 + h1 * x.w0 - h0 * x.w1;
}

// One body of the forward sweep (Dynamics.cc:39-73 with jcalc EulerXYZ, Joint.cc:100-139): parent state (vp, ap) -> v, a, f.
template <class T>
OMO_HD void forward_body(const OmoBody& B, const T (&e)[9], const T s1, const T c1, const T s2, const T c2, const T (&qd)[3], const T (&qdd)[3],
                         const V6<T>& vp, const V6<T>& ap, V6<T>& v, V6<T>& a, V6<T>& f)
{
    // S = [ (E00, E10, s1) (s2, c2, 0) (0, 0, 1) ] (angular rows)
    const T vJ0 = e[0] * qd[0] + s2 * qd[1], vJ1 = e[3] * qd[0] + c2 * qd[1], vJ2 = s1 * qd[0] + qd[2];
    const T q20 = qd[2] * qd[0], q10 = qd[1] * qd[0], q21 = qd[2] * qd[1];
    const T cJ0 = e[3] * q20 - (c2 * s1) * q10 + c2 * q21;
    const T cJ1 = -(e[0] * q20) + (s2 * s1) * q10 - s2 * q21;
    const T cJ2 = c1 * q10;
    const double r0 = B.r[0], r1 = B.r[1], r2 = B.r[2];
    {
        const T x0 = r2 * vp.w1 - r1 * vp.w2 + vp.l0, x1 = r0 * vp.w2 - r2 * vp.w0 + vp.l1, x2 = r1 * vp.w0 - r0 * vp.w1 + vp.l2;
        v.w0 = e[0] * vp.w0 + e[1] * vp.w1 + e[2] * vp.w2 + vJ0;
        v.w1 = e[3] * vp.w0 + e[4] * vp.w1 + e[5] * vp.w2 + vJ1;
        v.w2 = e[6] * vp.w0 + e[7] * vp.w1 + e[8] * vp.w2 + vJ2;
        v.l0 = e[0] * x0 + e[1] * x1 + e[2] * x2;
        v.l1 = e[3] * x0 + e[4] * x1 + e[5] * x2;
        v.l2 = e[6] * x0 + e[7] * x1 + e[8] * x2;
    }
    {
        const T x0 = r2 * ap.w1 - r1 * ap.w2 + ap.l0, x1 = r0 * ap.w2 - r2 * ap.w0 + ap.l1, x2 = r1 * ap.w0 - r0 * ap.w1 + ap.l2;
        const T cw0 = v.w1 * vJ2 - v.w2 * vJ1 + cJ0, cw1 = v.w2 * vJ0 - v.w0 * vJ2 + cJ1, cw2 = v.w0 * vJ1 - v.w1 * vJ0 + cJ2;
        const T cl0 = v.l1 * vJ2 - v.l2 * vJ1, cl1 = v.l2 * vJ0 - v.l0 * vJ2, cl2 = v.l0 * vJ1 - v.l1 * vJ0;
        a.w0 = e[0] * ap.w0 + e[1] * ap.w1 + e[2] * ap.w2 + e[0] * qdd[0] + s2 * qdd[1] + cw0;
        a.w1 = e[3] * ap.w0 + e[4] * ap.w1 + e[5] * ap.w2 + e[3] * qdd[0] + c2 * qdd[1] + cw1;
        a.w2 = e[6] * ap.w0 + e[7] * ap.w1 + e[8] * ap.w2 + s1 * qdd[0] + qdd[2] + cw2;
        a.l0 = e[0] * x0 + e[1] * x1 + e[2] * x2 + cl0;
        a.l1 = e[3] * x0 + e[4] * x1 + e[5] * x2 + cl1;
        a.l2 = e[6] * x0 + e[7] * x1 + e[8] * x2 + cl2;
    }
    V6<T> Iv, Ia;
    rbi_mul(B, v, Iv);
    rbi_mul(B, a, Ia);
    f.w0 = v.w1 * Iv.w2 - v.w2 * Iv.w1 + v.l1 * Iv.l2 - v.l2 * Iv.l1 + Ia.w0;
    f.w1 = v.w2 * Iv.w0 - v.w0 * Iv.w2 + v.l2 * Iv.l0 - v.l0 * Iv.l2 + Ia.w1;
    f.w2 = v.w0 * Iv.w1 - v.w1 * Iv.w0 + v.l0 * Iv.l1 - v.l1 * Iv.l0 + Ia.w2;
    f.l0 = v.w1 * Iv.l2 - v.w2 * Iv.l1 + Ia.l0;
    f.l1 = v.w2 * Iv.l0 - v.w0 * Iv.l2 + Ia.l1;
    f.l2 = v.w0 * Iv.l1 - v.w1 * Iv.l0 + Ia.l2;
}

// tau = S^T F, g = X_lambda^T F (Dynamics.cc:87-104)
template <class T>
OMO_HD void backward_body(const double r0, const double r1, const double r2, const T (&e)[9], const T s1, const T s2, const T c2, const V6<T>& F,
                          V6<T>& g, T (&tau)[3])
{
    tau[0] = e[0] * F.w0 + e[3] * F.w1 + s1 * F.w2;
    tau[1] = s2 * F.w0 + c2 * F.w1;
    tau[2] = F.w2;
    const T e0 = e[0] * F.l0 + e[3] * F.l1 + e[6] * F.l2, e1 = e[1] * F.l0 + e[4] * F.l1 + e[7] * F.l2, e2 = e[2] * F.l0 + e[5] * F.l1 + e[8] * F.l2;
    g.w0 = e[0] * F.w0 + e[3] * F.w1 + e[6] * F.w2 + r1 * e2 - r2 * e1;
    g.w1 = e[1] * F.w0 + e[4] * F.w1 + e[7] * F.w2 + r2 * e0 - r0 * e2;
    g.w2 = e[2] * F.w0 + e[5] * F.w1 + e[8] * F.w2 + r0 * e1 - r1 * e0;
    g.l0 = e0; g.l1 = e1; g.l2 = e2;
}

// Local rotation of Joint::GetAffineRelativeToParent (OptiMo/joint.cpp:64-74): Rz(q2) Ry(q1) Rx(q0), row-major
template <class T> OMO_HD void local_rotation(const T s0, const T c0, const T s1, const T c1, const T s2, const T c2, T (&R)[9])
{
    R[0] = c2 * c1; R[1] = c2 * s1 * s0 - s2 * c0; R[2] = c2 * s1 * c0 + s2 * s0;
    R[3] = s2 * c1; R[4] = s2 * s1 * s0 + c2 * c0; R[5] = s2 * s1 * c0 - c2 * s0;
    R[6] = -s1;     R[7] = c1 * s0;                R[8] = c1 * c0;
}

constexpr int kMaxDepth = OMO_MAX_JOINTS;

// d(cost of one sample) / d(one scalar input), by forward-mode differentiation restricted to what that input reaches.
//   cost = w_cost dt sum_{k >= 1} |tau_k|^2 + w_traj dt sum_k |p_k - p0_k|^2
// joint j, axis, kind (0: q, 1: qd, 2: qdd) name the input; q / qd / qdd / tau are this sample's [3J] vectors (core.h:120-152
// layout: q[0:3] = root offset); sub_end[j] = one past the last joint of j's subtree (joints are in DFS pre-order);
// p0 = column of the original positions, strided; root_e = the root joint's own Euler angles (constant channels or sampled).
// MAXSUB = capacity of the per-thread subtree arrays.
template <int MAXSUB>
OMO_HD double sample_tangent(const OmoDevScene& S, const int* sub_end, const int j, const int axis, const int kind, const double* q, const double* qd,
                             const double* qdd, const double* tau, const double w_cost, const double w_traj, const double dt, const double* p0,
                             const long long p0_stride, const double* root_e, const bool interior)
{
    // ---- ancestors of j, root side first (the root itself is handled apart: TranslationXYZ, massless) ----
    int path[kMaxDepth];
    int depth = 0;
    for (int k = S.parent[j]; k > 0; k = S.parent[k]) path[depth++] = k;
    // primal state down the path
    V6<double> v, a;
    zero(v); zero(a);
    v.l0 = interior ? S.root_qd[0] : 0.0; v.l1 = interior ? S.root_qd[1] : 0.0; v.l2 = interior ? S.root_qd[2] : 0.0;
    a.l0 = interior ? S.root_qdd[0] : 0.0; a.l1 = interior ? S.root_qdd[1] : 0.0; a.l2 = interior ? S.root_qdd[2] : 0.0;
    // world affine for the trajectory term: L (rotation), W (translation) of the frame the next body hangs in
    double L[9], W[3];
    const bool traj = (w_traj != 0.0) && kind == 0;
    if (traj) {
        double s0, c0, s1, c1, s2, c2;
        sincos_of(root_e[0], s0, c0); sincos_of(root_e[1], s1, c1); sincos_of(root_e[2], s2, c2);
        local_rotation(s0, c0, s1, c1, s2, c2, L);
        W[0] = S.offset[0][0]; W[1] = S.offset[0][1]; W[2] = S.offset[0][2];
    }
    for (int d = depth - 1; d >= 0; --d) {
        const int k = path[d];
        double s0, c0, s1, c1, s2, c2, e[9];
        sincos_of(q[3 * k], s0, c0); sincos_of(q[3 * k + 1], s1, c1); sincos_of(q[3 * k + 2], s2, c2);
        euler_E(s0, c0, s1, c1, s2, c2, e);
        const double qdk[3] = {qd[3 * k], qd[3 * k + 1], qd[3 * k + 2]}, qddk[3] = {qdd[3 * k], qdd[3 * k + 1], qdd[3 * k + 2]};
        V6<double> vn, an, fn;
        forward_body(S.body[k], e, s1, c1, s2, c2, qdk, qddk, v, a, vn, an, fn);
        v = vn; a = an;
        if (traj) {
            const double o0 = S.offset[k][0], o1 = S.offset[k][1], o2 = S.offset[k][2];
            const double p_0 = L[0] * o0 + L[1] * o1 + L[2] * o2 + W[0], p_1 = L[3] * o0 + L[4] * o1 + L[5] * o2 + W[1], p_2 = L[6] * o0 + L[7] * o1 + L[8] * o2 + W[2];
            double R[9], n[9];
            local_rotation(s0, c0, s1, c1, s2, c2, R);
            for (int r = 0; r < 3; ++r)
                for (int c = 0; c < 3; ++c) n[3 * r + c] = L[3 * r] * R[c] + L[3 * r + 1] * R[3 + c] + L[3 * r + 2] * R[6 + c];
            for (int c = 0; c < 9; ++c) L[c] = n[c];
            W[0] = p_0; W[1] = p_1; W[2] = p_2;
        }
    }
    // ---- subtree of j in dual arithmetic ----
    const int end = sub_end[j], n_sub = end - j;
    if (n_sub > MAXSUB) return 0.0;   // the host picks MAXSUB >= the largest subtree
    V6<Dual> sv[MAXSUB], sa[MAXSUB], sF[MAXSUB];
    Dual sL[MAXSUB][9], sW[MAXSUB][3];   // world affine handed to the children of each subtree body (trajectory term only)
    double dcost = 0.0, dtraj = 0.0;
    for (int k = j; k < end; ++k) {
        const int i = k - j;
        V6<Dual> vp, ap;
        if (k == j) {
            vp.w0 = v.w0; vp.w1 = v.w1; vp.w2 = v.w2; vp.l0 = v.l0; vp.l1 = v.l1; vp.l2 = v.l2;
            ap.w0 = a.w0; ap.w1 = a.w1; ap.w2 = a.w2; ap.l0 = a.l0; ap.l1 = a.l1; ap.l2 = a.l2;
        } else {
            vp = sv[S.parent[k] - j]; ap = sa[S.parent[k] - j];
        }
        Dual qk[3], qdk[3], qddk[3];
        for (int c = 0; c < 3; ++c) { qk[c] = Dual(q[3 * k + c]); qdk[c] = Dual(qd[3 * k + c]); qddk[c] = Dual(qdd[3 * k + c]); }
        if (k == j) { if (kind == 0) qk[axis].d = 1.0; else if (kind == 1) qdk[axis].d = 1.0; else qddk[axis].d = 1.0; }
        Dual s0, c0, s1, c1, s2, c2, e[9];
        sincos_of(qk[0], s0, c0); sincos_of(qk[1], s1, c1); sincos_of(qk[2], s2, c2);
        euler_E(s0, c0, s1, c1, s2, c2, e);
        V6<Dual> fk;
        forward_body(S.body[k], e, s1, c1, s2, c2, qdk, qddk, vp, ap, sv[i], sa[i], fk);
        sF[i] = fk;
        if (traj) {
            // p_k = W_parent + L_parent offset_k; only strict descendants of j move with q_j
            Dual Lp[9], Wp[3];
            if (k == j) { for (int c = 0; c < 9; ++c) Lp[c] = Dual(L[c]); for (int c = 0; c < 3; ++c) Wp[c] = Dual(W[c]); }
            else { for (int c = 0; c < 9; ++c) Lp[c] = sL[S.parent[k] - j][c]; for (int c = 0; c < 3; ++c) Wp[c] = sW[S.parent[k] - j][c]; }
            const double o0 = S.offset[k][0], o1 = S.offset[k][1], o2 = S.offset[k][2];
            Dual pk[3];
            for (int r = 0; r < 3; ++r) pk[r] = Lp[3 * r] * o0 + Lp[3 * r + 1] * o1 + Lp[3 * r + 2] * o2 + Wp[r];
            for (int r = 0; r < 3; ++r) dtraj += 2.0 * (pk[r].v - p0[(long long)(3 * k + r) * p0_stride]) * pk[r].d;
            Dual R[9];
            local_rotation(s0, c0, s1, c1, s2, c2, R);
            for (int r = 0; r < 3; ++r)
                for (int c = 0; c < 3; ++c) sL[i][3 * r + c] = Lp[3 * r] * R[c] + Lp[3 * r + 1] * R[3 + c] + Lp[3 * r + 2] * R[6 + c];
            for (int r = 0; r < 3; ++r) sW[i][r] = pk[r];
        }
    }
    // backward over the subtree (children before parents: reverse pre-order)
    V6<Dual> gj;
    zero(gj);
    for (int k = end - 1; k >= j; --k) {
        const int i = k - j;
        Dual s0, c0, s1, c1, s2, c2, e[9];
        Dual qk[3];
        for (int c = 0; c < 3; ++c) qk[c] = Dual(q[3 * k + c]);
        if (k == j && kind == 0) qk[axis].d = 1.0;
        sincos_of(qk[0], s0, c0); sincos_of(qk[1], s1, c1); sincos_of(qk[2], s2, c2);
        euler_E(s0, c0, s1, c1, s2, c2, e);
        V6<Dual> g;
        Dual t[3];
        backward_body(S.body[k].r[0], S.body[k].r[1], S.body[k].r[2], e, s1, s2, c2, sF[i], g, t);
        dcost += 2.0 * (t[0].v * t[0].d + t[1].v * t[1].d + t[2].v * t[2].d);
        if (k > j) add_to(sF[S.parent[k] - j], g); else gj = g;
    }
    // the change of force travels up the ancestor chain: d tau_k = S_k^T dF, dF_parent = X_k^T dF (primal S, X)
    V6<double> dg;
    dg.w0 = gj.w0.d; dg.w1 = gj.w1.d; dg.w2 = gj.w2.d; dg.l0 = gj.l0.d; dg.l1 = gj.l1.d; dg.l2 = gj.l2.d;
    for (int d = 0; d < depth; ++d) {
        const int k = path[d];
        double s0, c0, s1, c1, s2, c2, e[9];
        sincos_of(q[3 * k], s0, c0); sincos_of(q[3 * k + 1], s1, c1); sincos_of(q[3 * k + 2], s2, c2);
        euler_E(s0, c0, s1, c1, s2, c2, e);
        V6<double> g;
        double dt3[3];
        backward_body(S.body[k].r[0], S.body[k].r[1], S.body[k].r[2], e, s1, s2, c2, dg, g, dt3);
        dcost += 2.0 * (tau[3 * k] * dt3[0] + tau[3 * k + 1] * dt3[1] + tau[3 * k + 2] * dt3[2]);
        dg = g;
    }
    return w_cost * dt * dcost + (traj ? w_traj * dt * dtraj : 0.0);
}

// ---------------------------------------------------------------------------------------------------------------------------
// REVERSE mode: d(cost of one sample) / d(q, qd, qdd) of EVERY joint from one forward sweep, one backward sweep and their adjoints
// -- O(1) RNEA passes per sample where sample_tangent pays one restricted pass per (channel, kind). Same cost function:
//     cost = w_cost dt sum_{k >= 1} |tau_k|^2 + w_traj dt sum_k |p_k - p0_k|^2.
// Adjoint rules used throughout (bars = d cost / d .):   y = a x b:  abar += b x ybar,  bbar += ybar x a;
//     y = E x:  xbar += E^T ybar,  Ebar += ybar x^T;      y = E^T x:  xbar += E ybar,  Ebar += x ybar^T;      y = M x, M symmetric
//     (the spatial inertia of rbi_mul):  xbar += M ybar.
// Order: (A) adjoint of the force fold-back, parents before children: tau_k = S_k^T F_k, F_parent += X_k^T F_k give
//     Fbar_k = S_k taubar_k + X_k Fbar_parent (+ the E, s, c bars of S_k and X_k); (B) adjoint of the forward sweep, children before
//     parents: forward_body's outputs (v, a, f) back to (vp, ap, qd, qdd, E, s, c); (C) with w_traj, the forward kinematics:
//     p_k = L_parent o_k + p_parent, L_k = L_parent R_k.
// Work memory: NF doubles per body in `ws`, field f of body k at ws[(k * NF + f) * stride] (stride = number of samples on the
// GPU, where consecutive threads are consecutive samples; 1 on the host).
// ---------------------------------------------------------------------------------------------------------------------------
constexpr int kRevFields = 36;        // v 0-5, a 6-11, F (then Fbar) 12-17, sin/cos 18-23, vbar 24-29, abar 30-35
constexpr int kRevFieldsTraj = 60;    // + L 36-44, Lbar 45-53, p 54-56, pbar collected from the children 57-59

struct Rev3 { double x, y, z; };
OMO_HD Rev3 rcross(const Rev3& a, const Rev3& b) { return Rev3{a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x}; }
OMO_HD Rev3 radd(const Rev3& a, const Rev3& b) { return Rev3{a.x + b.x, a.y + b.y, a.z + b.z}; }
OMO_HD Rev3 rmul(const double (&e)[9], const Rev3& v) { return Rev3{e[0] * v.x + e[1] * v.y + e[2] * v.z, e[3] * v.x + e[4] * v.y + e[5] * v.z, e[6] * v.x + e[7] * v.y + e[8] * v.z}; }
OMO_HD Rev3 rmulT(const double (&e)[9], const Rev3& v) { return Rev3{e[0] * v.x + e[3] * v.y + e[6] * v.z, e[1] * v.x + e[4] * v.y + e[7] * v.z, e[2] * v.x + e[5] * v.y + e[8] * v.z}; }
// Ebar += y x^T
OMO_HD void router(double (&eb)[9], const Rev3& y, const Rev3& x)
{
    eb[0] += y.x * x.x; eb[1] += y.x * x.y; eb[2] += y.x * x.z;
    eb[3] += y.y * x.x; eb[4] += y.y * x.y; eb[5] += y.y * x.z;
    eb[6] += y.z * x.x; eb[7] += y.z * x.y; eb[8] += y.z * x.z;
}

template <bool TRAJ>
OMO_HD void sample_reverse(const OmoDevScene& S, const double* q, const double* qd, const double* qdd, const double w_cost, const double w_traj,
                           const double dt, const double* p0, const long long p0_stride, const double* root_e, const bool interior, double* ws,
                           const long long stride, double* gq, double* gqd, double* gqdd, const long long gstride)
{
    constexpr int NF = TRAJ ? kRevFieldsTraj : kRevFields;
    const int J = S.J;
#define OMO_WS(k, f) ws[((long long)(k) * NF + (f)) * stride]
    auto ld3 = [&](const int k, const int f) { return Rev3{OMO_WS(k, f), OMO_WS(k, f + 1), OMO_WS(k, f + 2)}; };
    auto st3 = [&](const int k, const int f, const Rev3& v) { OMO_WS(k, f) = v.x; OMO_WS(k, f + 1) = v.y; OMO_WS(k, f + 2) = v.z; };
    auto ld6 = [&](const int k, const int f) { V6<double> v; v.w0 = OMO_WS(k, f); v.w1 = OMO_WS(k, f + 1); v.w2 = OMO_WS(k, f + 2); v.l0 = OMO_WS(k, f + 3); v.l1 = OMO_WS(k, f + 4); v.l2 = OMO_WS(k, f + 5); return v; };
    auto st6 = [&](const int k, const int f, const V6<double>& v) { OMO_WS(k, f) = v.w0; OMO_WS(k, f + 1) = v.w1; OMO_WS(k, f + 2) = v.w2; OMO_WS(k, f + 3) = v.l0; OMO_WS(k, f + 4) = v.l1; OMO_WS(k, f + 5) = v.l2; };
    const double wc2 = 2.0 * w_cost * dt, wt2 = 2.0 * w_traj * dt;
    // ---- primal: forward sweep (v, a, own force), forward kinematics, force fold-back ---------------------------------------
    {
        V6<double> z; zero(z);
        V6<double> v0 = z, a0 = z;
        if (interior) { v0.l0 = S.root_qd[0]; v0.l1 = S.root_qd[1]; v0.l2 = S.root_qd[2]; a0.l0 = S.root_qdd[0]; a0.l1 = S.root_qdd[1]; a0.l2 = S.root_qdd[2]; }
        st6(0, 0, v0); st6(0, 6, a0); st6(0, 12, z); st6(0, 24, z); st6(0, 30, z);
        if (TRAJ) {
            double s0, c0, s1, c1, s2, c2, L0[9];
            sincos_of(root_e[0], s0, c0); sincos_of(root_e[1], s1, c1); sincos_of(root_e[2], s2, c2);
            local_rotation(s0, c0, s1, c1, s2, c2, L0);
            for (int c = 0; c < 9; ++c) { OMO_WS(0, 36 + c) = L0[c]; OMO_WS(0, 45 + c) = 0.0; }
            for (int c = 0; c < 3; ++c) { OMO_WS(0, 54 + c) = S.offset[0][c]; OMO_WS(0, 57 + c) = 0.0; }
        }
    }
    for (int k = 1; k < J; ++k) {
        const int p = S.parent[k];
        double s0, c0, s1, c1, s2, c2, e[9];
        sincos_of(q[3 * k], s0, c0); sincos_of(q[3 * k + 1], s1, c1); sincos_of(q[3 * k + 2], s2, c2);
        OMO_WS(k, 18) = s0; OMO_WS(k, 19) = c0; OMO_WS(k, 20) = s1; OMO_WS(k, 21) = c1; OMO_WS(k, 22) = s2; OMO_WS(k, 23) = c2;
        euler_E(s0, c0, s1, c1, s2, c2, e);
        const double qdk[3] = {qd[3 * k], qd[3 * k + 1], qd[3 * k + 2]}, qddk[3] = {qdd[3 * k], qdd[3 * k + 1], qdd[3 * k + 2]};
        V6<double> vk, ak, fk, z; zero(z);
        forward_body(S.body[k], e, s1, c1, s2, c2, qdk, qddk, ld6(p, 0), ld6(p, 6), vk, ak, fk);
        st6(k, 0, vk); st6(k, 6, ak); st6(k, 12, fk); st6(k, 24, z); st6(k, 30, z);
        if (TRAJ) {
            double Lp[9], R[9];
            for (int c = 0; c < 9; ++c) Lp[c] = OMO_WS(p, 36 + c);
            const Rev3 o{S.offset[k][0], S.offset[k][1], S.offset[k][2]};
            const Rev3 pk = radd(rmul(Lp, o), ld3(p, 54));
            st3(k, 54, pk); st3(k, 57, Rev3{0.0, 0.0, 0.0});
            local_rotation(s0, c0, s1, c1, s2, c2, R);
            for (int r = 0; r < 3; ++r)
                for (int c = 0; c < 3; ++c) { OMO_WS(k, 36 + 3 * r + c) = Lp[3 * r] * R[c] + Lp[3 * r + 1] * R[3 + c] + Lp[3 * r + 2] * R[6 + c]; OMO_WS(k, 45 + 3 * r + c) = 0.0; }
        }
    }
    for (int k = J - 1; k >= 1; --k) {   // F_parent += X_k^T F_k (children before parents); the root's own force is never used
        const int p = S.parent[k];
        if (p <= 0) continue;
        double e[9];
        euler_E(OMO_WS(k, 18), OMO_WS(k, 19), OMO_WS(k, 20), OMO_WS(k, 21), OMO_WS(k, 22), OMO_WS(k, 23), e);
        V6<double> g;
        double t[3];
        backward_body(S.body[k].r[0], S.body[k].r[1], S.body[k].r[2], e, (double)OMO_WS(k, 20), (double)OMO_WS(k, 22), (double)OMO_WS(k, 23), ld6(k, 12), g, t);
        V6<double> Fp = ld6(p, 12);
        add_to(Fp, g);
        st6(p, 12, Fp);
    }
    // ---- (A) adjoint of the fold-back, parents before children: slot F becomes Fbar; the E / s / c bars go straight into gq --------
    for (int k = 1; k < J; ++k) {
        const int p = S.parent[k];
        const double s0 = OMO_WS(k, 18), c0 = OMO_WS(k, 19), s1 = OMO_WS(k, 20), c1 = OMO_WS(k, 21), s2 = OMO_WS(k, 22), c2 = OMO_WS(k, 23);
        double e[9], eb[9] = {0, 0, 0, 0, 0, 0, 0, 0, 0};
        euler_E(s0, c0, s1, c1, s2, c2, e);
        const Rev3 Fw = ld3(k, 12), Fl = ld3(k, 15);
        // tau_k = S^T F.w (backward_body), taubar = 2 w dt tau
        const Rev3 tb{wc2 * (e[0] * Fw.x + e[3] * Fw.y + s1 * Fw.z), wc2 * (s2 * Fw.x + c2 * Fw.y), wc2 * Fw.z};
        Rev3 gw{0, 0, 0}, gl{0, 0, 0};
        if (p > 0) { gw = ld3(p, 12); gl = ld3(p, 15); }        // Fbar of the parent (already converted)
        const Rev3 r{S.body[k].r[0], S.body[k].r[1], S.body[k].r[2]};
        const Rev3 evb = radd(gl, rcross(gw, r));                // e_vec = E^T F.l, g.w = E^T F.w + r x e_vec, g.l = e_vec
        const Rev3 Fbw = radd(Rev3{e[0] * tb.x + s2 * tb.y, e[3] * tb.x + c2 * tb.y, s1 * tb.x + tb.z}, rmul(e, gw));
        const Rev3 Fbl = rmul(e, evb);
        router(eb, Fw, gw); router(eb, Fl, evb);
        double s1b = Fw.z * tb.x, s2b = Fw.x * tb.y, c2b = Fw.y * tb.y, s0b = 0.0, c0b = 0.0, c1b = 0.0;
        eb[0] += Fw.x * tb.x; eb[3] += Fw.y * tb.x;
        st3(k, 12, Fbw); st3(k, 15, Fbl);
        // euler_E adjoint
        c2b += eb[0] * c1 + eb[1] * s1 * s0 - eb[2] * s1 * c0 + eb[4] * c0 + eb[5] * s0;
        c1b += eb[0] * c2 - eb[3] * s2 - eb[7] * s0 + eb[8] * c0;
        s1b += eb[1] * c2 * s0 - eb[2] * c2 * c0 - eb[4] * s2 * s0 + eb[5] * s2 * c0 + eb[6];
        s0b += eb[1] * c2 * s1 + eb[2] * s2 - eb[4] * s2 * s1 + eb[5] * c2 - eb[7] * c1;
        s2b += eb[1] * c0 + eb[2] * s0 - eb[3] * c1 - eb[4] * s1 * s0 + eb[5] * s1 * c0;
        c0b += eb[1] * s2 - eb[2] * c2 * s1 + eb[4] * c2 + eb[5] * s2 * s1 + eb[8] * c1;
        gq[(long long)(3 * k) * gstride] = s0b * c0 - c0b * s0;
        gq[(long long)(3 * k + 1) * gstride] = s1b * c1 - c1b * s1;
        gq[(long long)(3 * k + 2) * gstride] = s2b * c2 - c2b * s2;
    }
    // ---- (B) adjoint of the forward sweep + (C) of the forward kinematics, children before parents ------------------------------
    for (int k = J - 1; k >= 1; --k) {
        const int p = S.parent[k];
        const OmoBody& B = S.body[k];
        const double s0 = OMO_WS(k, 18), c0 = OMO_WS(k, 19), s1 = OMO_WS(k, 20), c1 = OMO_WS(k, 21), s2 = OMO_WS(k, 22), c2 = OMO_WS(k, 23);
        double e[9], eb[9] = {0, 0, 0, 0, 0, 0, 0, 0, 0};
        euler_E(s0, c0, s1, c1, s2, c2, e);
        const Rev3 r{B.r[0], B.r[1], B.r[2]};
        const Rev3 w = ld3(k, 0), l = ld3(k, 3);
        const Rev3 wp = ld3(p, 0), lp = ld3(p, 3), alp = ld3(p, 6), bep = ld3(p, 9);
        const Rev3 qdk{qd[3 * k], qd[3 * k + 1], qd[3 * k + 2]}, qddk{qdd[3 * k], qdd[3 * k + 1], qdd[3 * k + 2]};
        const Rev3 vJ{e[0] * qdk.x + s2 * qdk.y, e[3] * qdk.x + c2 * qdk.y, s1 * qdk.x + qdk.z};
        const Rev3 xv = radd(lp, rcross(wp, r)), xa = radd(bep, rcross(alp, r));
        V6<double> vk, Iv; vk.w0 = w.x; vk.w1 = w.y; vk.w2 = w.z; vk.l0 = l.x; vk.l1 = l.y; vk.l2 = l.z;
        rbi_mul(B, vk, Iv);
        const Rev3 Ivw{Iv.w0, Iv.w1, Iv.w2}, Ivl{Iv.l0, Iv.l1, Iv.l2};
        const Rev3 fbw = ld3(k, 12), fbl = ld3(k, 15);          // fbar = Fbar_k
        // f = I a + v x* (I v)
        Rev3 wb = radd(ld3(k, 24), radd(rcross(Ivw, fbw), rcross(Ivl, fbl)));
        Rev3 lb = radd(ld3(k, 27), rcross(Ivl, fbw));
        V6<double> Ivb, Iab, t6;
        { const Rev3 a1 = rcross(fbw, w), a2 = radd(rcross(fbw, l), rcross(fbl, w)); Ivb.w0 = a1.x; Ivb.w1 = a1.y; Ivb.w2 = a1.z; Ivb.l0 = a2.x; Ivb.l1 = a2.y; Ivb.l2 = a2.z; }
        rbi_mul(B, Ivb, t6);
        wb = radd(wb, Rev3{t6.w0, t6.w1, t6.w2}); lb = radd(lb, Rev3{t6.l0, t6.l1, t6.l2});
        Iab.w0 = fbw.x; Iab.w1 = fbw.y; Iab.w2 = fbw.z; Iab.l0 = fbl.x; Iab.l1 = fbl.y; Iab.l2 = fbl.z;
        rbi_mul(B, Iab, t6);
        const Rev3 alb = radd(ld3(k, 30), Rev3{t6.w0, t6.w1, t6.w2}), beb = radd(ld3(k, 33), Rev3{t6.l0, t6.l1, t6.l2});
        // a = X ap + S qdd + cJ + v x vJ
        Rev3 apbw = rmulT(e, alb);
        router(eb, alb, alp);
        const Rev3 qddb{e[0] * alb.x + e[3] * alb.y + s1 * alb.z, s2 * alb.x + c2 * alb.y, alb.z};
        const Rev3 xab = rmulT(e, beb);
        router(eb, beb, xa);
        apbw = radd(apbw, rcross(r, xab));
        wb = radd(wb, rcross(vJ, alb));
        Rev3 vJb = rcross(alb, w);
        lb = radd(lb, rcross(vJ, beb));
        vJb = radd(vJb, rcross(beb, l));
        // v = X vp + vJ
        Rev3 vpbw = rmulT(e, wb);
        router(eb, wb, wp);
        vJb = radd(vJb, wb);
        const Rev3 xvb = rmulT(e, lb);
        router(eb, lb, xv);
        vpbw = radd(vpbw, rcross(r, xvb));
        // vJ = S qd, S qdd inside a
        Rev3 qdb{e[0] * vJb.x + e[3] * vJb.y + s1 * vJb.z, s2 * vJb.x + c2 * vJb.y, vJb.z};
        double e0b = alb.x * qddk.x + vJb.x * qdk.x, e3b = alb.y * qddk.x + vJb.y * qdk.x, s1b = alb.z * qddk.x + vJb.z * qdk.x;
        double s2b = alb.x * qddk.y + vJb.x * qdk.y, c2b = alb.y * qddk.y + vJb.y * qdk.y, s0b = 0.0, c0b = 0.0, c1b = 0.0;
        // cJ (cJbar = alb)
        {
            const double q20 = qdk.z * qdk.x, q10 = qdk.y * qdk.x, q21 = qdk.z * qdk.y;
            e3b += alb.x * q20; e0b -= alb.y * q20;
            const double q20b = alb.x * e[3] - alb.y * e[0];
            const double q10b = -alb.x * c2 * s1 + alb.y * s2 * s1 + alb.z * c1;
            const double q21b = alb.x * c2 - alb.y * s2;
            c2b += alb.x * (-s1 * q10 + q21);
            s1b += -alb.x * c2 * q10 + alb.y * s2 * q10;
            s2b += alb.y * (s1 * q10 - q21);
            c1b += alb.z * q10;
            qdb.x += q20b * qdk.z + q10b * qdk.y;
            qdb.y += q10b * qdk.x + q21b * qdk.z;
            qdb.z += q20b * qdk.x + q21b * qdk.y;
        }
        eb[0] += e0b; eb[3] += e3b;
        // parents collect (the root's state is constant)
        if (p > 0) {
            st3(p, 24, radd(ld3(p, 24), vpbw)); st3(p, 27, radd(ld3(p, 27), xvb));
            st3(p, 30, radd(ld3(p, 30), apbw)); st3(p, 33, radd(ld3(p, 33), xab));
        }
        // euler_E adjoint
        c2b += eb[0] * c1 + eb[1] * s1 * s0 - eb[2] * s1 * c0 + eb[4] * c0 + eb[5] * s0;
        c1b += eb[0] * c2 - eb[3] * s2 - eb[7] * s0 + eb[8] * c0;
        s1b += eb[1] * c2 * s0 - eb[2] * c2 * c0 - eb[4] * s2 * s0 + eb[5] * s2 * c0 + eb[6];
        s0b += eb[1] * c2 * s1 + eb[2] * s2 - eb[4] * s2 * s1 + eb[5] * c2 - eb[7] * c1;
        s2b += eb[1] * c0 + eb[2] * s0 - eb[3] * c1 - eb[4] * s1 * s0 + eb[5] * s1 * c0;
        c0b += eb[1] * s2 - eb[2] * c2 * s1 + eb[4] * c2 + eb[5] * s2 * s1 + eb[8] * c1;
        if (TRAJ) {
            // p_k = L_p o_k + p_p and L_k = L_p R_k (R_k = Rz Ry Rx of this joint's angles): every child of k has been here already
            double Lp[9], Lbk[9], R[9], Rb[9];
            for (int c = 0; c < 9; ++c) { Lp[c] = OMO_WS(p, 36 + c); Lbk[c] = OMO_WS(k, 45 + c); }
            const Rev3 pk = ld3(k, 54);
            const Rev3 pb = radd(ld3(k, 57), Rev3{wt2 * (pk.x - p0[(long long)(3 * k) * p0_stride]), wt2 * (pk.y - p0[(long long)(3 * k + 1) * p0_stride]),
                                                  wt2 * (pk.z - p0[(long long)(3 * k + 2) * p0_stride])});
            local_rotation(s0, c0, s1, c1, s2, c2, R);
            for (int m = 0; m < 3; ++m)
                for (int c = 0; c < 3; ++c) Rb[3 * m + c] = Lp[m] * Lbk[c] + Lp[3 + m] * Lbk[3 + c] + Lp[6 + m] * Lbk[6 + c];
            if (p > 0) {   // the root's frame and position are constants
                const double o[3] = {S.offset[k][0], S.offset[k][1], S.offset[k][2]};
                const double pbv[3] = {pb.x, pb.y, pb.z};
                for (int rr = 0; rr < 3; ++rr)
                    for (int m = 0; m < 3; ++m)
                        OMO_WS(p, 45 + 3 * rr + m) += pbv[rr] * o[m] + Lbk[3 * rr] * R[3 * m] + Lbk[3 * rr + 1] * R[3 * m + 1] + Lbk[3 * rr + 2] * R[3 * m + 2];
                st3(p, 57, radd(ld3(p, 57), pb));
            }
            c2b += Rb[0] * c1 + Rb[1] * s1 * s0 + Rb[2] * s1 * c0 + Rb[4] * c0 - Rb[5] * s0;
            c1b += Rb[0] * c2 + Rb[3] * s2 + Rb[7] * s0 + Rb[8] * c0;
            s1b += Rb[1] * c2 * s0 + Rb[2] * c2 * c0 + Rb[4] * s2 * s0 + Rb[5] * s2 * c0 - Rb[6];
            s0b += Rb[1] * c2 * s1 + Rb[2] * s2 + Rb[4] * s2 * s1 - Rb[5] * c2 + Rb[7] * c1;
            s2b += -Rb[1] * c0 + Rb[2] * s0 + Rb[3] * c1 + Rb[4] * s1 * s0 + Rb[5] * s1 * c0;
            c0b += -Rb[1] * s2 + Rb[2] * c2 * s1 + Rb[4] * c2 + Rb[5] * s2 * s1 + Rb[8] * c1;
        }
        gq[(long long)(3 * k) * gstride] += s0b * c0 - c0b * s0;
        gq[(long long)(3 * k + 1) * gstride] += s1b * c1 - c1b * s1;
        gq[(long long)(3 * k + 2) * gstride] += s2b * c2 - c2b * s2;
        gqd[(long long)(3 * k) * gstride] = qdb.x; gqd[(long long)(3 * k + 1) * gstride] = qdb.y; gqd[(long long)(3 * k + 2) * gstride] = qdb.z;
        gqdd[(long long)(3 * k) * gstride] = qddb.x; gqdd[(long long)(3 * k + 1) * gstride] = qddb.y; gqdd[(long long)(3 * k + 2) * gstride] = qddb.z;
    }
#undef OMO_WS
}

// Transposed stencils (core_rbdl.cpp:28-95, SECOND_ORDER_ACCURATE off): coefficient of q[i] in qd[m] / qdd[m], N samples.
OMO_HD double d1_coef(const int m, const int i, const int N, const double dt)
{
    const int o = i - m;
    if (m >= 2 && m <= N - 3) return (o == 2 ? -1.0 : o == 1 ? 8.0 : o == -1 ? -8.0 : o == -2 ? 1.0 : 0.0) / (12.0 * dt);
    if (m == 0) return (o == 1 ? 1.0 : o == 0 ? -1.0 : 0.0) / dt;
    if (m == N - 1) return (o == 0 ? 1.0 : o == -1 ? -1.0 : 0.0) / dt;
    return (o == 1 ? 1.0 : o == -1 ? -1.0 : 0.0) / (2.0 * dt);
}
OMO_HD double d2_coef(const int m, const int i, const int N, const double dt)
{
    const int o = i - m;
    if (m >= 2 && m <= N - 3) return (o == 2 || o == -2 ? -1.0 : o == 1 || o == -1 ? 16.0 : o == 0 ? -30.0 : 0.0) / (12.0 * dt * dt);
    if (m == 0) return (o == 2 ? 1.0 : o == 1 ? -2.0 : o == 0 ? 1.0 : 0.0) / (2.0 * dt * dt);          // the reference's / (2 dx^2), core_rbdl.cpp:92-93
    if (m == N - 1) return (o == 0 ? 1.0 : o == -1 ? -2.0 : o == -2 ? 1.0 : 0.0) / (2.0 * dt * dt);
    return (o == 1 || o == -1 ? 1.0 : o == 0 ? -2.0 : 0.0) / (dt * dt);
}

// Value of one Bezier segment at time t and its partials with respect to the raw parameters (t, v, theta, t_b, t_f) of its start
// point s (out[0..4]) and end point e (out[5..9]); curve.cpp:57-93. u is found exactly as the reference finds it (Newton from
// 0.5, stop when |f| < 1e-9, at most 50 iterations), du/d(time control points) by the implicit-function theorem.
OMO_HD double bezier_partials(const double* ps, const double* pe, const double ths, const double the, const double t, double (&out)[10])
{
    const double p0 = ps[0], p1 = ps[0] + ps[4], p2 = pe[0] + pe[3], p3 = pe[0];
    const double tns = tan(ths), tne = tan(the);
    const double v0 = ps[1], v1 = ps[1] + ps[4] * tns, v2 = pe[1] + pe[3] * tne, v3 = pe[1];
    const double a = -p0 + 3.0 * p1 - 3.0 * p2 + p3, b = 3.0 * (p0 - 2.0 * p1 + p2), c = -3.0 * p0 + 3.0 * p1, d = p0 - t;
    double u = 0.5;
    for (int it = 0; it < 50; ++it) {
        const double f = u * (u * (a * u + b) + c) + d, fp = u * (3.0 * a * u + 2.0 * b) + c;
        if (fabs(f) < 1e-09) break;
        u = u - f / fp;
    }
    const double w = 1.0 - u;
    const double b0 = w * w * w, b1 = 3.0 * w * w * u, b2 = 3.0 * w * u * u, b3 = u * u * u;
    const double val = v0 * b0 + v1 * b1 + v2 * b2 + v3 * b3;
    const double dB = 3.0 * ((v1 - v0) * w * w + 2.0 * (v2 - v1) * w * u + (v3 - v2) * u * u);   // d value / du
    const double dT = 3.0 * ((p1 - p0) * w * w + 2.0 * (p2 - p1) * w * u + (p3 - p2) * u * u);   // d time / du
    const double k = -dB / dT;                                                                   // d value / d p_j = k b_j (through u)
    const double sec2s = 1.0 + tns * tns, sec2e = 1.0 + tne * tne;
    out[0] = k * (b0 + b1);                 // t_s moves p0 and p1
    out[1] = b0 + b1;                       // v_s moves v0 and v1
    out[2] = b1 * ps[4] * sec2s;            // theta_s moves v1
    out[3] = 0.0;                           // t_b of the start point belongs to the previous segment
    out[4] = k * b1 + b1 * tns;             // t_f of the start point moves p1 and v1
    out[5] = k * (b2 + b3);                 // t_e
    out[6] = b2 + b3;                       // v_e
    out[7] = b2 * pe[3] * sec2e;            // theta_e moves v2
    out[8] = k * b2 + b2 * tne;             // t_b of the end point moves p2 and v2
    out[9] = 0.0;
    return val;
}

}  // namespace omo_ad
#endif
