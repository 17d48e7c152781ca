// TEST INFRASTRUCTURE -- part of the CPU oracle, never linked into or called by the product path.
//
// Plain-C++ restatement of the subset of RBDL v2.5.0 (bundled under the reference's
// external/rbdl-v2.5.0, SimpleMath backend) that OptiMo's objective path executes. All paths are
// relative to /root/reference/external/rbdl-v2.5.0/.
//
//   SpatialTransform apply / applyTranspose / product / inverse / toMatrix
//                                        include/rbdl/SpatialAlgebraOperators.h:159-194,233-249,297-304
//   SpatialRigidBodyInertia::operator*   include/rbdl/SpatialAlgebraOperators.h:51-65
//   createFromMassComInertiaC            include/rbdl/SpatialAlgebraOperators.h:116-128
//   crossm / crossf                      include/rbdl/SpatialAlgebraOperators.h:414-445
//   jcalc (EulerXYZ, EulerZYX, TranslationXYZ)   src/Joint.cc:61-139,179-209
//   Model::AddBody (3-DoF joints only)   src/Model.cc:258-464
//   InverseDynamics                      src/Dynamics.cc:26-105
//   UpdateKinematicsCustom(Q only)       src/Kinematics.cc:88-113
//   CalcBodyToBaseCoordinates            src/Kinematics.cc:168-199
//   CalcPointJacobian                    src/Kinematics.cc:255-317
//   InverseKinematics                    src/Kinematics.cc:615-709
//   LinSolveGaussElimPivot               src/rbdl_mathutils.cc:54-127
//
// It is validated against the real RBDL objects (oracle/_ref build) by tests/test_oracle_pin.py.
#ifndef ORACLE_OWN_RBDL_HPP
#define ORACLE_OWN_RBDL_HPP

#include <array>
#include <cmath>
#include <limits>
#include <vector>

#include "omath.hpp"

namespace orc {
namespace own {

using SV = std::array<double, 6>;  // spatial vector (angular; linear)

enum JointKind { kTranslationXYZ = 0, kEulerXYZ = 1, kEulerZYX = 2 };

struct Xform {  // SpatialTransform {E, r}
    M3 E; V3 r;
    Xform() : E(M3::identity()), r() {}
    Xform(const M3& e, const V3& t) : E(e), r(t) {}

    SV apply(const SV& s) const
    {
        const V3 w(s[0], s[1], s[2]);
        const V3 rxw(s[3] - r[1] * s[2] + r[2] * s[1],
                     s[4] - r[2] * s[0] + r[0] * s[2],
                     s[5] - r[0] * s[1] + r[1] * s[0]);
        const V3 a = E * w, b = E * rxw;
        return SV{a[0], a[1], a[2], b[0], b[1], b[2]};
    }
    SV applyTranspose(const SV& f) const
    {
        const V3 etf(E(0, 0) * f[3] + E(1, 0) * f[4] + E(2, 0) * f[5],
                     E(0, 1) * f[3] + E(1, 1) * f[4] + E(2, 1) * f[5],
                     E(0, 2) * f[3] + E(1, 2) * f[4] + E(2, 2) * f[5]);
        return SV{E(0, 0) * f[0] + E(1, 0) * f[1] + E(2, 0) * f[2] - r[2] * etf[1] + r[1] * etf[2],
                  E(0, 1) * f[0] + E(1, 1) * f[1] + E(2, 1) * f[2] + r[2] * etf[0] - r[0] * etf[2],
                  E(0, 2) * f[0] + E(1, 2) * f[1] + E(2, 2) * f[2] - r[1] * etf[0] + r[0] * etf[1],
                  etf[0], etf[1], etf[2]};
    }
    Xform inverse() const
    {
        const V3 er = E * r;
        return Xform(transpose(E), V3(-er[0], -er[1], -er[2]));
    }
    /// (*this) * XT == SpatialTransform(E * XT.E, XT.r + XT.E^T * r)
    Xform operator*(const Xform& xt) const { return Xform(E * xt.E, xt.r + transpose(xt.E) * r); }
};

struct RBI {  // SpatialRigidBodyInertia
    double m = 0.0; V3 h;
    double Ixx = 0.0, Iyx = 0.0, Iyy = 0.0, Izx = 0.0, Izy = 0.0, Izz = 0.0;

    SV mul(const SV& mv) const
    {
        const V3 lo(mv[3], mv[4], mv[5]), up(mv[0], mv[1], mv[2]);
        const V3 hx = cross(h, lo);
        const V3 ru(Ixx * mv[0] + Iyx * mv[1] + Izx * mv[2] + hx[0],
                    Iyx * mv[0] + Iyy * mv[1] + Izy * mv[2] + hx[1],
                    Izx * mv[0] + Izy * mv[1] + Izz * mv[2] + hx[2]);
        const V3 rl = m * lo - cross(h, up);
        return SV{ru[0], ru[1], ru[2], rl[0], rl[1], rl[2]};
    }
    static RBI from_mass_com_inertia_c(double mass, const V3& com, const M3& inertia_c)
    {
        RBI r;
        r.m = mass;
        r.h = V3(com[0] * mass, com[1] * mass, com[2] * mass);
        M3 cx;  // VectorCrossMatrix(com)
        cx(0, 1) = -com[2]; cx(0, 2) = com[1];
        cx(1, 0) = com[2];  cx(1, 2) = -com[0];
        cx(2, 0) = -com[1]; cx(2, 1) = com[0];
        const M3 cc = cx * transpose(cx);
        M3 I;
        for (int i = 0; i < 3; ++i) for (int j = 0; j < 3; ++j) I(i, j) = inertia_c(i, j) + cc(i, j) * mass;
        r.Ixx = I(0, 0); r.Iyx = I(1, 0); r.Iyy = I(1, 1); r.Izx = I(2, 0); r.Izy = I(2, 1); r.Izz = I(2, 2);
        return r;
    }
};

inline SV crossm(const SV& a, const SV& b)
{
    return SV{-a[2] * b[1] + a[1] * b[2],
              a[2] * b[0] - a[0] * b[2],
              -a[1] * b[0] + a[0] * b[1],
              -a[5] * b[1] + a[4] * b[2] - a[2] * b[4] + a[1] * b[5],
              a[5] * b[0] - a[3] * b[2] + a[2] * b[3] - a[0] * b[5],
              -a[4] * b[0] + a[3] * b[1] - a[1] * b[3] + a[0] * b[4]};
}
inline SV crossf(const SV& a, const SV& b)
{
    return SV{-a[2] * b[1] + a[1] * b[2] - a[5] * b[4] + a[4] * b[5],
              a[2] * b[0] - a[0] * b[2] + a[5] * b[3] - a[3] * b[5],
              -a[1] * b[0] + a[0] * b[1] - a[4] * b[3] + a[3] * b[4],
              -a[2] * b[4] + a[1] * b[5],
              a[2] * b[3] - a[0] * b[5],
              -a[1] * b[3] + a[0] * b[4]};
}
inline SV add(const SV& a, const SV& b)
{
    SV r; for (int i = 0; i < 6; ++i) r[i] = a[i] + b[i]; return r;
}

using S63 = std::array<std::array<double, 3>, 6>;  // multdof3_S

inline SV s_times(const S63& S, double a, double b, double c)
{
    SV r; for (int i = 0; i < 6; ++i) r[i] = S[i][0] * a + S[i][1] * b + S[i][2] * c; return r;
}

struct Model {
    // index 0 is the fixed base, bodies start at 1 (Model::Model, src/Model.cc:23-86)
    std::vector<unsigned> lambda;
    std::vector<int> kind;
    std::vector<unsigned> q_index;
    std::vector<Xform> X_T, X_J, X_lambda, X_base;
    std::vector<RBI> I;
    std::vector<SV> v, a, c, f, v_J, c_J;
    std::vector<S63> S;
    unsigned dof_count = 0;
    V3 gravity = V3(0.0, -9.81, 0.0);

    Model() { push_slot(0, -1, Xform(), RBI()); q_index[0] = 0; }

    unsigned add_body(unsigned parent, const Xform& joint_frame, int joint_kind, double mass, const V3& com,
                      const M3& inertia_c)
    {
        const unsigned prev = static_cast<unsigned>(lambda.size()) - 1;
        const unsigned prev_dof = (prev == 0) ? 0u : 3u;
        push_slot(parent, joint_kind, joint_frame, RBI::from_mass_com_inertia_c(mass, com, inertia_c));
        q_index.back() = q_index[prev] + prev_dof;
        dof_count += 3;
        return static_cast<unsigned>(lambda.size()) - 1;
    }

    void jcalc(unsigned i, const VecN& q, const double* qd /* 3 values or null => zero */)
    {
        const unsigned qi = q_index[i];
        const double q0 = q[qi], q1 = q[qi + 1], q2 = q[qi + 2];
        const double d0 = qd ? qd[0] : 0.0, d1 = qd ? qd[1] : 0.0, d2 = qd ? qd[2] : 0.0;
        S63& s = S[i];
        if (kind[i] == kEulerZYX) {
            const double s0 = std::sin(q0), c0 = std::cos(q0), s1 = std::sin(q1), c1 = std::cos(q1),
                         s2 = std::sin(q2), c2 = std::cos(q2);
            M3& E = X_J[i].E;
            E(0, 0) = c0 * c1;                E(0, 1) = s0 * c1;                E(0, 2) = -s1;
            E(1, 0) = c0 * s1 * s2 - s0 * c2; E(1, 1) = s0 * s1 * s2 + c0 * c2; E(1, 2) = c1 * s2;
            E(2, 0) = c0 * s1 * c2 + s0 * s2; E(2, 1) = s0 * s1 * c2 - c0 * s2; E(2, 2) = c1 * c2;
            s[0][0] = -s1;     s[0][2] = 1.0;
            s[1][0] = c1 * s2; s[1][1] = c2;
            s[2][0] = c1 * c2; s[2][1] = -s2;
            v_J[i] = s_times(s, d0, d1, d2);
            c_J[i] = SV{-c1 * d0 * d1,
                        -s1 * s2 * d0 * d1 + c1 * c2 * d0 * d2 - s2 * d1 * d2,
                        -s1 * c2 * d0 * d1 - c1 * s2 * d0 * d2 - c2 * d1 * d2, 0.0, 0.0, 0.0};
        } else if (kind[i] == kEulerXYZ) {
            const double s0 = std::sin(q0), c0 = std::cos(q0), s1 = std::sin(q1), c1 = std::cos(q1),
                         s2 = std::sin(q2), c2 = std::cos(q2);
            M3& E = X_J[i].E;
            E(0, 0) = c2 * c1;  E(0, 1) = s2 * c0 + c2 * s1 * s0; E(0, 2) = s2 * s0 - c2 * s1 * c0;
            E(1, 0) = -s2 * c1; E(1, 1) = c2 * c0 - s2 * s1 * s0; E(1, 2) = c2 * s0 + s2 * s1 * c0;
            E(2, 0) = s1;       E(2, 1) = -c1 * s0;               E(2, 2) = c1 * c0;
            s[0][0] = c2 * c1;  s[0][1] = s2;
            s[1][0] = -s2 * c1; s[1][1] = c2;
            s[2][0] = s1;       s[2][2] = 1.0;
            v_J[i] = s_times(s, d0, d1, d2);
            c_J[i] = SV{-s2 * c1 * d2 * d0 - c2 * s1 * d1 * d0 + c2 * d2 * d1,
                        -c2 * c1 * d2 * d0 + s2 * s1 * d1 * d0 - s2 * d2 * d1,
                        c1 * d1 * d0, 0.0, 0.0, 0.0};
        } else {  // TranslationXYZ
            X_J[i].E = M3::identity();
            X_J[i].r = V3(q0, q1, q2);
            s[3][0] = 1.0; s[4][1] = 1.0; s[5][2] = 1.0;
            v_J[i] = s_times(s, d0, d1, d2);
            c_J[i] = SV{0.0, 0.0, 0.0, 0.0, 0.0, 0.0};
        }
        X_lambda[i] = X_J[i] * X_T[i];
    }

    void inverse_dynamics(const VecN& Q, const VecN& QDot, const VecN& QDDot, VecN& Tau)
    {
        v[0] = SV{0, 0, 0, 0, 0, 0};
        a[0] = SV{0.0, 0.0, 0.0, -gravity[0], -gravity[1], -gravity[2]};
        const unsigned n = static_cast<unsigned>(lambda.size());
        for (unsigned i = 1; i < n; ++i) {
            const unsigned qi = q_index[i], lam = lambda[i];
            jcalc(i, Q, &QDot[qi]);
            v[i] = add(X_lambda[i].apply(v[lam]), v_J[i]);
            c[i] = add(c_J[i], crossm(v[i], v_J[i]));
            a[i] = add(add(X_lambda[i].apply(a[lam]), c[i]), s_times(S[i], QDDot[qi], QDDot[qi + 1], QDDot[qi + 2]));
            f[i] = add(I[i].mul(a[i]), crossf(v[i], I[i].mul(v[i])));
        }
        for (unsigned i = n - 1; i > 0; --i) {
            const unsigned qi = q_index[i];
            for (int k = 0; k < 3; ++k) {
                double t = 0.0;
                for (int r = 0; r < 6; ++r) t += S[i][r][k] * f[i][r];
                Tau[qi + k] = t;
            }
            if (lambda[i] != 0) f[lambda[i]] = add(f[lambda[i]], X_lambda[i].applyTranspose(f[i]));
        }
    }

    void update_kinematics_q(const VecN& Q)
    {
        const unsigned n = static_cast<unsigned>(lambda.size());
        for (unsigned i = 1; i < n; ++i) {
            jcalc(i, Q, nullptr);
            X_base[i] = (lambda[i] != 0) ? X_lambda[i] * X_base[lambda[i]] : X_lambda[i];
        }
    }

    V3 body_to_base(unsigned body, const V3& p) const
    {
        return X_base[body].r + transpose(X_base[body].E) * p;
    }

    /// G is 3 x dof_count, row-major, zero-initialised by the caller.
    void point_jacobian(unsigned body, const V3& point, std::vector<double>& G) const
    {
        const Xform point_trans(M3::identity(), body_to_base(body, point));
        unsigned j = body;
        while (j != 0) {
            const Xform x = point_trans * X_base[j].inverse();
            // lower three rows of x.toMatrix(): [ -E*rx , E ]
            M3 rx;
            rx(0, 1) = -x.r[2]; rx(0, 2) = x.r[1];
            rx(1, 0) = x.r[2];  rx(1, 2) = -x.r[0];
            rx(2, 0) = -x.r[1]; rx(2, 1) = x.r[0];
            const M3 erx = x.E * rx;
            for (int r = 0; r < 3; ++r)
                for (int k = 0; k < 3; ++k) {
                    double t = 0.0;
                    for (int m = 0; m < 3; ++m) t += (-erx(r, m)) * S[j][m][k];
                    for (int m = 0; m < 3; ++m) t += x.E(r, m) * S[j][3 + m][k];
                    G[r * dof_count + q_index[j] + k] = t;
                }
            j = lambda[j];
        }
    }

    /// Gauss elimination with the column pivoting scheme of rbdl_mathutils.cc:54-127 (A, b by value).
    static bool lin_solve_gauss_pivot(std::vector<double> A, std::vector<double> b, std::vector<double>& x, unsigned n)
    {
        std::vector<size_t> pivot(n);
        std::vector<double> px(n, 0.0);
        for (unsigned i = 0; i < n; ++i) pivot[i] = i;
        auto at = [&](unsigned r, size_t c) -> double& { return A[r * n + c]; };
        for (unsigned j = 0; j < n; ++j) {
            unsigned pi = j;
            double pv = std::fabs(at(j, pivot[j]));
            for (unsigned k = j; k < n; ++k) {
                const double pt = std::fabs(at(j, pivot[k]));
                if (pt > pv) {
                    pv = pt; pi = k;
                    const size_t sw = pivot[j]; pivot[j] = pivot[pi]; pivot[pi] = sw;
                }
            }
            for (unsigned i = j + 1; i < n; ++i) {
                const double d = at(i, pivot[j]) / at(j, pivot[j]);
                b[i] -= b[j] * d;
                for (unsigned k = j; k < n; ++k) at(i, pivot[k]) -= at(j, pivot[k]) * d;
            }
        }
        unsigned i = n;
        do {
            --i;
            for (unsigned j = i + 1; j < n; ++j) px[i] += at(i, pivot[j]) * px[j];
            px[i] = (b[i] - px[i]) / at(i, pivot[i]);
        } while (i > 0);
        x.assign(n, 0.0);
        for (unsigned k = 0; k < n; ++k) x[pivot[k]] = px[k];
        return true;
    }

    /// Single-target damped-least-squares IK; returns the executed iteration count through *iters.
    bool inverse_kinematics(const VecN& Qinit, unsigned body, const V3& body_point, const V3& target, VecN& Qres,
                            double step_tol, double lambda_damp, unsigned max_iter, int* iters = nullptr)
    {
        const unsigned nd = dof_count;
        std::vector<double> J(3 * nd, 0.0), e(3, 0.0);
        Qres = Qinit;
        for (unsigned it = 0; it < max_iter; ++it) {
            if (iters) *iters = static_cast<int>(it) + 1;
            update_kinematics_q(Qres);
            std::vector<double> G(3 * nd, 0.0);
            point_jacobian(body, body_point, G);
            const V3 pb = body_to_base(body, body_point);
            J = G;
            for (int i = 0; i < 3; ++i) e[i] = target[i] - pb[i];
            if (std::sqrt(e[0] * e[0] + e[1] * e[1] + e[2] * e[2]) < step_tol) return true;
            std::vector<double> A(9, 0.0);
            for (int r = 0; r < 3; ++r)
                for (int c2 = 0; c2 < 3; ++c2) {
                    double t = 0.0;
                    for (unsigned k = 0; k < nd; ++k) t += J[r * nd + k] * J[c2 * nd + k];
                    A[r * 3 + c2] = t + lambda_damp * lambda_damp * (r == c2 ? 1.0 : 0.0);
                }
            std::vector<double> z;
            lin_solve_gauss_pivot(A, e, z, 3);
            double dn = 0.0;
            for (unsigned k = 0; k < nd; ++k) {
                double t = 0.0;
                for (int r = 0; r < 3; ++r) t += J[r * nd + k] * z[r];
                Qres[k] = Qres[k] + t;
                dn += t * t;
            }
            if (std::sqrt(dn) < step_tol) return true;
        }
        return false;
    }

private:
    void push_slot(unsigned parent, int joint_kind, const Xform& xt, const RBI& rbi)
    {
        lambda.push_back(parent);
        kind.push_back(joint_kind);
        q_index.push_back(0);
        X_T.push_back(xt); X_J.push_back(Xform()); X_lambda.push_back(Xform()); X_base.push_back(Xform());
        I.push_back(rbi);
        const SV z{0, 0, 0, 0, 0, 0};
        v.push_back(z); a.push_back(z); c.push_back(z); f.push_back(z); v_J.push_back(z); c_J.push_back(z);
        S63 s; for (auto& row : s) row = {0.0, 0.0, 0.0};
        S.push_back(s);
    }
};

}  // namespace own
}  // namespace orc

#endif
