// RNEA body steps (FP64, registers only). Restates, per body, RBDL v2.5.0:
//   jcalc EulerXYZ                 external/rbdl-v2.5.0/src/Joint.cc:100-139, X_lambda = X_J X_T :209
//   InverseDynamics forward sweep  src/Dynamics.cc:39-73
//   InverseDynamics backward sweep src/Dynamics.cc:87-104
//   SpatialTransform::apply / applyTranspose   include/rbdl/SpatialAlgebraOperators.h:159-194
//   crossm / crossf                            SpatialAlgebraOperators.h:414-445
//   SpatialRigidBodyInertia::operator*         SpatialAlgebraOperators.h:51-65
// Structural zeros are exploited (X_T.E = I, X_J.r = 0, S angular-only) and every sum of products is written
// as an explicit FMA chain: the FP64 pipe issues one DFMA per slot, a separate DMUL + DADD costs two.
// The *_identity variants are the same formulas for a joint whose three angles are the constant 0
// (E = I, S = I3, qd = qdd = 0 exactly): the results are bit-identical to the general path on those inputs.
#ifndef OMO_RNEA_CUH
#define OMO_RNEA_CUH

#include "omo_scene.cuh"

namespace omo {

struct SV6 { double w0, w1, w2, l0, l1, l2; };  // spatial vector: angular ; linear

struct Rot9 { double e00, e01, e02, e10, e11, e12, e20, e21, e22; };

// E of the EulerXYZ joint (Joint.cc:112-116) from (s0,c0,s1,c1,s2,c2)
__device__ __forceinline__ void euler_xyz_E(const double s0, const double c0, const double s1, const double c1, const double s2,
                                            const double c2, Rot9& E)
{
    const double c2s1 = c2 * s1, s2s1 = s2 * s1;
    E.e00 = c2 * c1;
    E.e01 = fma(c2s1, s0, s2 * c0);
    E.e02 = fma(-c2s1, c0, s2 * s0);
    E.e10 = -(s2 * c1);
    E.e11 = fma(-s2s1, s0, c2 * c0);
    E.e12 = fma(s2s1, c0, c2 * s0);
    E.e20 = s1;
    E.e21 = -(c1 * s0);
    E.e22 = c1 * c0;
}

// I * x for a spatial motion vector x (SpatialAlgebraOperators.h:51-65)
__device__ __forceinline__ void rbi_mul(const OmoBody& B, const SV6& x, SV6& y)
{
    const double h0 = B.h[0], h1 = B.h[1], h2 = B.h[2], m = B.m;
    y.w0 = fma(B.Ixx, x.w0, fma(B.Iyx, x.w1, fma(B.Izx, x.w2, fma(h1, x.l2, -(h2 * x.l1)))));
    y.w1 = fma(B.Iyx, x.w0, fma(B.Iyy, x.w1, fma(B.Izy, x.w2, fma(h2, x.l0, -(h0 * x.l2)))));
    y.w2 = fma(B.Izx, x.w0, fma(B.Izy, x.w1, fma(B.Izz, x.w2, fma(h0, x.l1, -(h1 * x.l0)))));
    y.l0 = fma(m, x.l0, fma(h2, x.w1, -(h1 * x.w2)));
    y.l1 = fma(m, x.l1, fma(h0, x.w2, -(h2 * x.w0)));
    y.l2 = fma(m, x.l2, fma(h1, x.w0, -(h0 * x.w1)));
}

// f = I a + v x* (I v)
__device__ __forceinline__ void body_force(const OmoBody& B, const SV6& v, const SV6& a, SV6& f)
{
    SV6 Iv, Ia;
    rbi_mul(B, v, Iv);
    rbi_mul(B, a, Ia);
    f.w0 = fma(v.w1, Iv.w2, fma(-v.w2, Iv.w1, fma(v.l1, Iv.l2, fma(-v.l2, Iv.l1, Ia.w0))));
    f.w1 = fma(v.w2, Iv.w0, fma(-v.w0, Iv.w2, fma(v.l2, Iv.l0, fma(-v.l0, Iv.l2, Ia.w1))));
    f.w2 = fma(v.w0, Iv.w1, fma(-v.w1, Iv.w0, fma(v.l0, Iv.l1, fma(-v.l1, Iv.l0, Ia.w2))));
    f.l0 = fma(v.w1, Iv.l2, fma(-v.w2, Iv.l1, Ia.l0));
    f.l1 = fma(v.w2, Iv.l0, fma(-v.w0, Iv.l2, Ia.l1));
    f.l2 = fma(v.w0, Iv.l1, fma(-v.w1, Iv.l0, Ia.l2));
}

// General EulerXYZ body: (vp, ap) parent state -> (v, a, f) of this body.
__device__ __forceinline__ void rnea_forward(const OmoBody& B, const double s0, const double c0, const double s1, const double c1,
                                             const double s2, const double c2, const double qd0, const double qd1, const double qd2,
                                             const double qdd0, const double qdd1, const double qdd2, const SV6& vp, const SV6& ap,
                                             SV6& v, SV6& a, SV6& f)
{
    Rot9 E;
    euler_xyz_E(s0, c0, s1, c1, s2, c2, E);
    // v_J = S qd with S = [ (E00,E10,s1) (s2,c2,0) (0,0,1) ] (angular rows; Joint.cc:118-132)
    const double vJ0 = fma(E.e00, qd0, s2 * qd1);
    const double vJ1 = fma(E.e10, qd0, c2 * qd1);
    const double vJ2 = fma(s1, qd0, qd2);
    // c_J (Joint.cc:134-139)
    const double q20 = qd2 * qd0, q10 = qd1 * qd0, q21 = qd2 * qd1;
    const double cJ0 = fma(E.e10, q20, fma(-(c2 * s1), q10, c2 * q21));
    const double cJ1 = fma(-E.e00, q20, fma(s2 * s1, q10, -(s2 * q21)));
    const double cJ2 = c1 * q10;
    const double r0 = B.r[0], r1 = B.r[1], r2 = B.r[2];
    // v = X_lambda.apply(v_parent) + v_J
    {
        const double x0 = fma(r2, vp.w1, fma(-r1, vp.w2, vp.l0));
        const double x1 = fma(r0, vp.w2, fma(-r2, vp.w0, vp.l1));
        const double x2 = fma(r1, vp.w0, fma(-r0, vp.w1, vp.l2));
        v.w0 = fma(E.e00, vp.w0, fma(E.e01, vp.w1, fma(E.e02, vp.w2, vJ0)));
        v.w1 = fma(E.e10, vp.w0, fma(E.e11, vp.w1, fma(E.e12, vp.w2, vJ1)));
        v.w2 = fma(E.e20, vp.w0, fma(E.e21, vp.w1, fma(E.e22, vp.w2, vJ2)));
        v.l0 = fma(E.e00, x0, fma(E.e01, x1, E.e02 * x2));
        v.l1 = fma(E.e10, x0, fma(E.e11, x1, E.e12 * x2));
        v.l2 = fma(E.e20, x0, fma(E.e21, x1, E.e22 * x2));
    }
    // a = X_lambda.apply(a_parent) + (c_J + v x v_J) + S qdd
    {
        const double x0 = fma(r2, ap.w1, fma(-r1, ap.w2, ap.l0));
        const double x1 = fma(r0, ap.w2, fma(-r2, ap.w0, ap.l1));
        const double x2 = fma(r1, ap.w0, fma(-r0, ap.w1, ap.l2));
        const double cw0 = fma(v.w1, vJ2, fma(-v.w2, vJ1, cJ0));
        const double cw1 = fma(v.w2, vJ0, fma(-v.w0, vJ2, cJ1));
        const double cw2 = fma(v.w0, vJ1, fma(-v.w1, vJ0, cJ2));
        const double cl0 = fma(v.l1, vJ2, -(v.l2 * vJ1));
        const double cl1 = fma(v.l2, vJ0, -(v.l0 * vJ2));
        const double cl2 = fma(v.l0, vJ1, -(v.l1 * vJ0));
        a.w0 = fma(E.e00, ap.w0, fma(E.e01, ap.w1, fma(E.e02, ap.w2, fma(E.e00, qdd0, fma(s2, qdd1, cw0)))));
        a.w1 = fma(E.e10, ap.w0, fma(E.e11, ap.w1, fma(E.e12, ap.w2, fma(E.e10, qdd0, fma(c2, qdd1, cw1)))));
        a.w2 = fma(E.e20, ap.w0, fma(E.e21, ap.w1, fma(E.e22, ap.w2, fma(s1, qdd0, qdd2 + cw2))));
        a.l0 = fma(E.e00, x0, fma(E.e01, x1, fma(E.e02, x2, cl0)));
        a.l1 = fma(E.e10, x0, fma(E.e11, x1, fma(E.e12, x2, cl1)));
        a.l2 = fma(E.e20, x0, fma(E.e21, x1, fma(E.e22, x2, cl2)));
    }
    body_force(B, v, a, f);
}

// Same for a joint frozen at zero angles: E = I, v_J = c_J = S qdd = 0.
__device__ __forceinline__ void rnea_forward_identity(const OmoBody& B, const SV6& vp, const SV6& ap, SV6& v, SV6& a, SV6& f)
{
    const double r0 = B.r[0], r1 = B.r[1], r2 = B.r[2];
    v.w0 = vp.w0; v.w1 = vp.w1; v.w2 = vp.w2;
    v.l0 = fma(r2, vp.w1, fma(-r1, vp.w2, vp.l0));
    v.l1 = fma(r0, vp.w2, fma(-r2, vp.w0, vp.l1));
    v.l2 = fma(r1, vp.w0, fma(-r0, vp.w1, vp.l2));
    a.w0 = ap.w0; a.w1 = ap.w1; a.w2 = ap.w2;
    a.l0 = fma(r2, ap.w1, fma(-r1, ap.w2, ap.l0));
    a.l1 = fma(r0, ap.w2, fma(-r2, ap.w0, ap.l1));
    a.l2 = fma(r1, ap.w0, fma(-r0, ap.w1, ap.l2));
    body_force(B, v, a, f);
}

// Backward step: tau = S^T f (returned as |tau|^2 accumulated into tsq and as t0..t2), g = X_lambda^T f.
__device__ __forceinline__ void rnea_backward(const double r0, const double r1, const double r2, const double s0, const double c0,
                                              const double s1, const double c1, const double s2, const double c2, const SV6& f,
                                              SV6& g, double& t0, double& t1, double& t2)
{
    Rot9 E;
    euler_xyz_E(s0, c0, s1, c1, s2, c2, E);
    t0 = fma(E.e00, f.w0, fma(E.e10, f.w1, s1 * f.w2));
    t1 = fma(s2, f.w0, c2 * f.w1);
    t2 = f.w2;
    const double e0 = fma(E.e00, f.l0, fma(E.e10, f.l1, E.e20 * f.l2));
    const double e1 = fma(E.e01, f.l0, fma(E.e11, f.l1, E.e21 * f.l2));
    const double e2 = fma(E.e02, f.l0, fma(E.e12, f.l1, E.e22 * f.l2));
    g.w0 = fma(E.e00, f.w0, fma(E.e10, f.w1, fma(E.e20, f.w2, fma(r1, e2, -(r2 * e1)))));
    g.w1 = fma(E.e01, f.w0, fma(E.e11, f.w1, fma(E.e21, f.w2, fma(r2, e0, -(r0 * e2)))));
    g.w2 = fma(E.e02, f.w0, fma(E.e12, f.w1, fma(E.e22, f.w2, fma(r0, e1, -(r1 * e0)))));
    g.l0 = e0; g.l1 = e1; g.l2 = e2;
}

__device__ __forceinline__ void rnea_backward_identity(const double r0, const double r1, const double r2, const SV6& f, SV6& g,
                                                       double& t0, double& t1, double& t2)
{
    t0 = f.w0; t1 = f.w1; t2 = f.w2;
    g.w0 = fma(r1, f.l2, fma(-r2, f.l1, f.w0));
    g.w1 = fma(r2, f.l0, fma(-r0, f.l2, f.w1));
    g.w2 = fma(r0, f.l1, fma(-r1, f.l0, f.w2));
    g.l0 = f.l0; g.l1 = f.l1; g.l2 = f.l2;
}

}  // namespace omo
#endif
