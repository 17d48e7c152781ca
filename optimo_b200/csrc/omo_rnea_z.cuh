// RNEA body steps with STRUCTURAL ZEROS resolved at compile time (scene-specialised kernels only).
//
// The formulas are those of omo_rnea.cuh (RBDL v2.5.0: jcalc EulerXYZ src/Joint.cc:100-139, InverseDynamics
// src/Dynamics.cc:39-104, SpatialAlgebraOperators.h:51-65,159-194,414-445) written ONCE over a tiny symbolic algebra:
// an operand is a double, or the type ZeroT (exactly 0.0) or OneT (exactly 1.0). zmul / zfma / zadd fold away every
// product with a ZeroT and every multiplication by a OneT while the C++ is compiled, so a joint whose x and y angles
// are the constant 0 (sin = ZeroT, cos = OneT, qd = qdd = ZeroT) or a body whose offset lies on an axis (r, h, the
// inertia's off-diagonals ZeroT) executes only the DFMAs that can change the result. With every operand a double the
// expansion is operation for operation the code of omo_rnea.cuh; with ZeroT / OneT operands the results differ from
// it at most in the sign of a zero (x*0 + c == c). On B200 a DFMA holds the issue port for two cycles and the
// specialised kernel is instruction-fetch bound (profiles/micro/icache_mix), so every instruction removed is time.
#ifndef OMO_RNEA_Z_CUH
#define OMO_RNEA_Z_CUH

#include "omo_rnea.cuh"

namespace omo {

struct ZeroT {};
struct OneT {};

template <class T> struct zis_zero { static constexpr bool value = false; };
template <> struct zis_zero<ZeroT> { static constexpr bool value = true; };
template <class T> struct zis_one { static constexpr bool value = false; };
template <> struct zis_one<OneT> { static constexpr bool value = true; };

__device__ __forceinline__ double zval(const double x) { return x; }
__device__ __forceinline__ double zval(ZeroT) { return 0.0; }
__device__ __forceinline__ double zval(OneT) { return 1.0; }

template <class A> __device__ __forceinline__ auto zneg(const A a)
{
    if constexpr (zis_zero<A>::value) return ZeroT{};
    else return -zval(a);
}
template <class A, class B> __device__ __forceinline__ auto zmul(const A a, const B b)
{
    if constexpr (zis_zero<A>::value || zis_zero<B>::value) return ZeroT{};
    else if constexpr (zis_one<A>::value) return b;
    else if constexpr (zis_one<B>::value) return a;
    else return a * b;
}
template <class A, class B> __device__ __forceinline__ auto zadd(const A a, const B b)
{
    if constexpr (zis_zero<A>::value) return b;
    else if constexpr (zis_zero<B>::value) return a;
    else return zval(a) + zval(b);
}
// a * b + c
template <class A, class B, class C> __device__ __forceinline__ auto zfma(const A a, const B b, const C c)
{
    if constexpr (zis_zero<A>::value || zis_zero<B>::value) return c;
    else if constexpr (zis_zero<C>::value) return zmul(a, b);
    else if constexpr (zis_one<A>::value) return zval(b) + zval(c);
    else if constexpr (zis_one<B>::value) return zval(a) + zval(c);
    else return fma(a, b, zval(c));
}

// sin / cos of the three joint angles, each a double or a structural constant
template <class S0, class C0, class S1, class C1, class S2, class C2> struct SinCosZ {
    S0 s0; C0 c0; S1 s1; C1 c1; S2 s2; C2 c2;
};
template <class T0, class T1, class T2> struct Vec3Z { T0 x; T1 y; T2 z; };

// Body constants with the zero pattern ZB resolved at compile time.
//   bit 0-2: r[0..2] == 0   bit 3-5: h[0..2] == 0   bit 6: Iyx == 0   bit 7: Izx == 0   bit 8: Izy == 0
template <int ZB> struct BodyZ {
    const OmoBody& B;
    __device__ __forceinline__ explicit BodyZ(const OmoBody& b) : B(b) {}
#define OMO_BZ(name, bit, expr) \
    __device__ __forceinline__ auto name() const { if constexpr ((ZB >> bit) & 1) return ZeroT{}; else return expr; }
    OMO_BZ(r0, 0, B.r[0]) OMO_BZ(r1, 1, B.r[1]) OMO_BZ(r2, 2, B.r[2])
    OMO_BZ(h0, 3, B.h[0]) OMO_BZ(h1, 4, B.h[1]) OMO_BZ(h2, 5, B.h[2])
    OMO_BZ(Iyx, 6, B.Iyx) OMO_BZ(Izx, 7, B.Izx) OMO_BZ(Izy, 8, B.Izy)
#undef OMO_BZ
};

// E of the EulerXYZ joint (Joint.cc:112-116), entries typed
template <class SC> struct RotZ {
    const SC& q;
    __device__ __forceinline__ explicit RotZ(const SC& sc) : q(sc) {}
    __device__ __forceinline__ auto c2s1() const { return zmul(q.c2, q.s1); }
    __device__ __forceinline__ auto s2s1() const { return zmul(q.s2, q.s1); }
    __device__ __forceinline__ auto e00() const { return zmul(q.c2, q.c1); }
    __device__ __forceinline__ auto e01() const { return zfma(c2s1(), q.s0, zmul(q.s2, q.c0)); }
    __device__ __forceinline__ auto e02() const { return zfma(zneg(c2s1()), q.c0, zmul(q.s2, q.s0)); }
    __device__ __forceinline__ auto e10() const { return zneg(zmul(q.s2, q.c1)); }
    __device__ __forceinline__ auto e11() const { return zfma(zneg(s2s1()), q.s0, zmul(q.c2, q.c0)); }
    __device__ __forceinline__ auto e12() const { return zfma(s2s1(), q.c0, zmul(q.c2, q.s0)); }
    __device__ __forceinline__ auto e20() const { return q.s1; }
    __device__ __forceinline__ auto e21() const { return zneg(zmul(q.c1, q.s0)); }
    __device__ __forceinline__ auto e22() const { return zmul(q.c1, q.c0); }
};

// I * x (SpatialAlgebraOperators.h:51-65)
template <int ZB> __device__ __forceinline__ void rbi_mul_z(const BodyZ<ZB>& K, const SV6& x, SV6& y)
{
    const OmoBody& B = K.B;
    const auto h0 = K.h0(); const auto h1 = K.h1(); const auto h2 = K.h2();
    y.w0 = zval(zfma(B.Ixx, x.w0, zfma(K.Iyx(), x.w1, zfma(K.Izx(), x.w2, zfma(h1, x.l2, zneg(zmul(h2, x.l1)))))));
    y.w1 = zval(zfma(K.Iyx(), x.w0, zfma(B.Iyy, x.w1, zfma(K.Izy(), x.w2, zfma(h2, x.l0, zneg(zmul(h0, x.l2)))))));
    y.w2 = zval(zfma(K.Izx(), x.w0, zfma(K.Izy(), x.w1, zfma(B.Izz, x.w2, zfma(h0, x.l1, zneg(zmul(h1, x.l0)))))));
    y.l0 = zval(zfma(B.m, x.l0, zfma(h2, x.w1, zneg(zmul(h1, x.w2)))));
    y.l1 = zval(zfma(B.m, x.l1, zfma(h0, x.w2, zneg(zmul(h2, x.w0)))));
    y.l2 = zval(zfma(B.m, x.l2, zfma(h1, x.w0, zneg(zmul(h0, x.w1)))));
}

// f = I a + v x* (I v)
template <int ZB> __device__ __forceinline__ void body_force_z(const BodyZ<ZB>& K, const SV6& v, const SV6& a, SV6& f)
{
    SV6 Iv, Ia;
    rbi_mul_z(K, v, Iv);
    rbi_mul_z(K, a, Ia);
    f.w0 = fma(v.w1, Iv.w2, fma(-v.w2, Iv.w1, fma(v.l1, Iv.l2, fma(-v.l2, Iv.l1, Ia.w0))));
    f.w1 = fma(v.w2, Iv.w0, fma(-v.w0, Iv.w2, fma(v.l2, Iv.l0, fma(-v.l0, Iv.l2, Ia.w1))));
    f.w2 = fma(v.w0, Iv.w1, fma(-v.w1, Iv.w0, fma(v.l0, Iv.l1, fma(-v.l1, Iv.l0, Ia.w2))));
    f.l0 = fma(v.w1, Iv.l2, fma(-v.w2, Iv.l1, Ia.l0));
    f.l1 = fma(v.w2, Iv.l0, fma(-v.w0, Iv.l2, Ia.l1));
    f.l2 = fma(v.w0, Iv.l1, fma(-v.w1, Iv.l0, Ia.l2));
}

// General EulerXYZ body: (vp, ap) parent state -> (v, a, f) of this body (omo_rnea.cuh: rnea_forward).
template <int ZB, class SC, class QD, class QDD>
__device__ __forceinline__ void rnea_forward_z(const BodyZ<ZB>& K, const SC& sc, const QD& qd, const QDD& qdd, const SV6& vp, const SV6& ap,
                                               SV6& v, SV6& a, SV6& f)
{
    const RotZ<SC> E(sc);
    const auto e00 = E.e00(); const auto e01 = E.e01(); const auto e02 = E.e02();
    const auto e10 = E.e10(); const auto e11 = E.e11(); const auto e12 = E.e12();
    const auto e20 = E.e20(); const auto e21 = E.e21(); const auto e22 = E.e22();
    // v_J = S qd with S = [ (E00,E10,s1) (s2,c2,0) (0,0,1) ] (angular rows; Joint.cc:118-132)
    const auto vJ0 = zfma(e00, qd.x, zmul(sc.s2, qd.y));
    const auto vJ1 = zfma(e10, qd.x, zmul(sc.c2, qd.y));
    const auto vJ2 = zfma(sc.s1, qd.x, qd.z);
    // c_J (Joint.cc:134-139)
    const auto q20 = zmul(qd.z, qd.x); const auto q10 = zmul(qd.y, qd.x); const auto q21 = zmul(qd.z, qd.y);
    const auto cJ0 = zfma(e10, q20, zfma(zneg(zmul(sc.c2, sc.s1)), q10, zmul(sc.c2, q21)));
    const auto cJ1 = zfma(zneg(e00), q20, zfma(zmul(sc.s2, sc.s1), q10, zneg(zmul(sc.s2, q21))));
    const auto cJ2 = zmul(sc.c1, q10);
    const auto r0 = K.r0(); const auto r1 = K.r1(); const auto r2 = K.r2();
    // v = X_lambda.apply(v_parent) + v_J
    {
        const auto x0 = zfma(r2, vp.w1, zfma(zneg(r1), vp.w2, vp.l0));
        const auto x1 = zfma(r0, vp.w2, zfma(zneg(r2), vp.w0, vp.l1));
        const auto x2 = zfma(r1, vp.w0, zfma(zneg(r0), vp.w1, vp.l2));
        v.w0 = zval(zfma(e00, vp.w0, zfma(e01, vp.w1, zfma(e02, vp.w2, vJ0))));
        v.w1 = zval(zfma(e10, vp.w0, zfma(e11, vp.w1, zfma(e12, vp.w2, vJ1))));
        v.w2 = zval(zfma(e20, vp.w0, zfma(e21, vp.w1, zfma(e22, vp.w2, vJ2))));
        v.l0 = zval(zfma(e00, x0, zfma(e01, x1, zmul(e02, x2))));
        v.l1 = zval(zfma(e10, x0, zfma(e11, x1, zmul(e12, x2))));
        v.l2 = zval(zfma(e20, x0, zfma(e21, x1, zmul(e22, x2))));
    }
    // a = X_lambda.apply(a_parent) + (c_J + v x v_J) + S qdd
    {
        const auto x0 = zfma(r2, ap.w1, zfma(zneg(r1), ap.w2, ap.l0));
        const auto x1 = zfma(r0, ap.w2, zfma(zneg(r2), ap.w0, ap.l1));
        const auto x2 = zfma(r1, ap.w0, zfma(zneg(r0), ap.w1, ap.l2));
        const auto cw0 = zfma(v.w1, vJ2, zfma(-v.w2, vJ1, cJ0));
        const auto cw1 = zfma(v.w2, vJ0, zfma(-v.w0, vJ2, cJ1));
        const auto cw2 = zfma(v.w0, vJ1, zfma(-v.w1, vJ0, cJ2));
        const auto cl0 = zfma(v.l1, vJ2, zneg(zmul(v.l2, vJ1)));
        const auto cl1 = zfma(v.l2, vJ0, zneg(zmul(v.l0, vJ2)));
        const auto cl2 = zfma(v.l0, vJ1, zneg(zmul(v.l1, vJ0)));
        a.w0 = zval(zfma(e00, ap.w0, zfma(e01, ap.w1, zfma(e02, ap.w2, zfma(e00, qdd.x, zfma(sc.s2, qdd.y, cw0))))));
        a.w1 = zval(zfma(e10, ap.w0, zfma(e11, ap.w1, zfma(e12, ap.w2, zfma(e10, qdd.x, zfma(sc.c2, qdd.y, cw1))))));
        a.w2 = zval(zfma(e20, ap.w0, zfma(e21, ap.w1, zfma(e22, ap.w2, zfma(sc.s1, qdd.x, zadd(qdd.z, cw2))))));
        a.l0 = zval(zfma(e00, x0, zfma(e01, x1, zfma(e02, x2, cl0))));
        a.l1 = zval(zfma(e10, x0, zfma(e11, x1, zfma(e12, x2, cl1))));
        a.l2 = zval(zfma(e20, x0, zfma(e21, x1, zfma(e22, x2, cl2))));
    }
    body_force_z(K, v, a, f);
}

// Backward step: tau = S^T f, g = X_lambda^T f (omo_rnea.cuh: rnea_backward).
template <int ZB, class SC>
__device__ __forceinline__ void rnea_backward_z(const BodyZ<ZB>& K, const SC& sc, const SV6& f, SV6& g, double& t0, double& t1, double& t2)
{
    const RotZ<SC> E(sc);
    const auto e00 = E.e00(); const auto e01 = E.e01(); const auto e02 = E.e02();
    const auto e10 = E.e10(); const auto e11 = E.e11(); const auto e12 = E.e12();
    const auto e20 = E.e20(); const auto e21 = E.e21(); const auto e22 = E.e22();
    const auto r0 = K.r0(); const auto r1 = K.r1(); const auto r2 = K.r2();
    t0 = zval(zfma(e00, f.w0, zfma(e10, f.w1, zmul(sc.s1, f.w2))));
    t1 = zval(zfma(sc.s2, f.w0, zmul(sc.c2, f.w1)));
    t2 = f.w2;
    const auto e0 = zfma(e00, f.l0, zfma(e10, f.l1, zmul(e20, f.l2)));
    const auto e1 = zfma(e01, f.l0, zfma(e11, f.l1, zmul(e21, f.l2)));
    const auto e2 = zfma(e02, f.l0, zfma(e12, f.l1, zmul(e22, f.l2)));
    g.w0 = zval(zfma(e00, f.w0, zfma(e10, f.w1, zfma(e20, f.w2, zfma(r1, e2, zneg(zmul(r2, e1)))))));
    g.w1 = zval(zfma(e01, f.w0, zfma(e11, f.w1, zfma(e21, f.w2, zfma(r2, e0, zneg(zmul(r0, e2)))))));
    g.w2 = zval(zfma(e02, f.w0, zfma(e12, f.w1, zfma(e22, f.w2, zfma(r0, e1, zneg(zmul(r1, e0)))))));
    g.l0 = zval(e0); g.l1 = zval(e1); g.l2 = zval(e2);
}

}  // namespace omo
#endif
