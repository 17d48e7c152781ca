// TEST INFRASTRUCTURE -- host build of the analytic gradient's arithmetic (optimo_b200/csrc/omo_adjoint.cuh, the same source the
// CUDA kernels compile) so that it can be checked against finite differences on the CPU, without a GPU:
//   g++ -O2 -std=c++17 -shared -fPIC -I optimo_b200/csrc -o tests/native/_build/libadjoint_check.so tests/native/adjoint_check.cpp
// Nothing in the product links or loads this.
#include <cstring>

struct alignas(16) int4 { int x, y, z, w; };   // the CUDA vector type OmoDevScene uses for its packed walk metadata
#include "omo_adjoint.cuh"

using namespace omo_ad;

extern "C" {

unsigned long long adj_sizeof_scene() { return sizeof(OmoDevScene); }

// cost of one sample from (q, qd, qdd) over the FULL tree in plain doubles: w_cost dt sum |tau[3:]|^2 + w_traj dt sum |p - p0|^2
double adj_sample_cost(const void* scene, const double* q, const double* qd, const double* qdd, double w_cost, double w_traj, double dt, const double* p0,
                       long long p0_stride, const double* root_e, int interior, double* tau_out)
{
    const OmoDevScene& S = *static_cast<const OmoDevScene*>(scene);
    const int J = S.J;
    V6<double> v[OMO_MAX_JOINTS], a[OMO_MAX_JOINTS], F[OMO_MAX_JOINTS];
    double L[OMO_MAX_JOINTS][9], W[OMO_MAX_JOINTS][3];
    zero(v[0]); zero(a[0]); zero(F[0]);
    if (interior) { v[0].l0 = S.root_qd[0]; v[0].l1 = S.root_qd[1]; v[0].l2 = S.root_qd[2]; a[0].l0 = S.root_qdd[0]; a[0].l1 = S.root_qdd[1]; a[0].l2 = S.root_qdd[2]; }
    double s0, c0, s1, c1, s2, c2;
    sincos_of(root_e[0], s0, c0); sincos_of(root_e[1], s1, c1); sincos_of(root_e[2], s2, c2);
    local_rotation(s0, c0, s1, c1, s2, c2, L[0]);
    for (int k = 0; k < 3; ++k) W[0][k] = S.offset[0][k];
    double traj = 0.0;
    if (w_traj != 0.0) for (int r = 0; r < 3; ++r) { const double d = W[0][r] - p0[(long long)r * p0_stride]; traj += d * d; }
    for (int k = 1; k < J; ++k) {
        const int p = S.parent[k];
        double e[9];
        sincos_of(q[3 * k], s0, c0); sincos_of(q[3 * k + 1], s1, c1); sincos_of(q[3 * k + 2], s2, c2);
        euler_E(s0, c0, s1, c1, s2, c2, e);
        const double qdk[3] = {qd[3 * k], qd[3 * k + 1], qd[3 * k + 2]}, qddk[3] = {qdd[3 * k], qdd[3 * k + 1], qdd[3 * k + 2]};
        forward_body(S.body[k], e, s1, c1, s2, c2, qdk, qddk, v[p], a[p], v[k], a[k], F[k]);
        double pk[3];
        for (int r = 0; r < 3; ++r) pk[r] = L[p][3 * r] * S.offset[k][0] + L[p][3 * r + 1] * S.offset[k][1] + L[p][3 * r + 2] * S.offset[k][2] + W[p][r];
        if (w_traj != 0.0) for (int r = 0; r < 3; ++r) { const double d = pk[r] - p0[(long long)(3 * k + r) * p0_stride]; traj += d * d; }
        double R[9];
        local_rotation(s0, c0, s1, c1, s2, c2, R);
        for (int r = 0; r < 3; ++r) for (int c = 0; c < 3; ++c) L[k][3 * r + c] = L[p][3 * r] * R[c] + L[p][3 * r + 1] * R[3 + c] + L[p][3 * r + 2] * R[6 + c];
        for (int r = 0; r < 3; ++r) W[k][r] = pk[r];
    }
    double tsq = 0.0;
    for (int k = J - 1; k >= 1; --k) {
        double e[9];
        sincos_of(q[3 * k], s0, c0); sincos_of(q[3 * k + 1], s1, c1); sincos_of(q[3 * k + 2], s2, c2);
        euler_E(s0, c0, s1, c1, s2, c2, e);
        V6<double> g;
        double t[3];
        backward_body(S.body[k].r[0], S.body[k].r[1], S.body[k].r[2], e, s1, s2, c2, F[k], g, t);
        if (tau_out) { tau_out[3 * k] = t[0]; tau_out[3 * k + 1] = t[1]; tau_out[3 * k + 2] = t[2]; }
        tsq += t[0] * t[0] + t[1] * t[1] + t[2] * t[2];
        if (S.parent[k] > 0) add_to(F[S.parent[k]], g);
    }
    return w_cost * dt * tsq + w_traj * dt * traj;
}

double adj_sample_tangent(const void* scene, const int* sub_end, int j, int axis, int kind, const double* q, const double* qd, const double* qdd,
                          const double* tau, double w_cost, double w_traj, double dt, const double* p0, long long p0_stride, const double* root_e, int interior)
{
    return sample_tangent<OMO_MAX_JOINTS>(*static_cast<const OmoDevScene*>(scene), sub_end, j, axis, kind, q, qd, qdd, tau, w_cost, w_traj, dt, p0, p0_stride,
                                          root_e, interior != 0);
}

// reverse mode: all 3 x 3J partials of one sample's cost at once (gq, gqd, gqdd: [3J], entries of the root untouched)
void adj_sample_reverse(const void* scene, const double* q, const double* qd, const double* qdd, double w_cost, double w_traj, double dt, const double* p0,
                        long long p0_stride, const double* root_e, int interior, double* gq, double* gqd, double* gqdd)
{
    const OmoDevScene& S = *static_cast<const OmoDevScene*>(scene);
    static double ws[OMO_MAX_JOINTS * kRevFieldsTraj];
    if (w_traj != 0.0) sample_reverse<true>(S, q, qd, qdd, w_cost, w_traj, dt, p0, p0_stride, root_e, interior != 0, ws, 1, gq, gqd, gqdd, 1);
    else sample_reverse<false>(S, q, qd, qdd, w_cost, 0.0, dt, nullptr, 0, root_e, interior != 0, ws, 1, gq, gqd, gqdd, 1);
}

double adj_bezier(const double* ps, const double* pe, double ths, double the, double t, double* out10)
{
    double o[10];
    const double v = bezier_partials(ps, pe, ths, the, t, o);
    std::memcpy(out10, o, sizeof(o));
    return v;
}

double adj_d1(int m, int i, int N, double dt) { return d1_coef(m, i, N, dt); }
double adj_d2(int m, int i, int N, double dt) { return d2_coef(m, i, N, dt); }

}  // extern "C"
