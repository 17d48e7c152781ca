// FP64 sin/cos pair for joint angles (sm_100a). The reference calls libm sin/cos inside RBDL's jcalc
// (external/rbdl-v2.5.0/src/Joint.cc:100-110) and Eigen's AngleAxis (OptiMo/joint.cpp:66-73); any
// implementation within ~1 ulp is inside the 1e-9 parity bar by seven orders of magnitude.
//
// CUDA's sincos() costs ~95 SASS instructions per call here (generic quadrant / huge-argument handling);
// this one is ~40: Cody-Waite reduction by pi/2 with three FMA steps (exact products), the classic
// minimax kernels on [-pi/4, pi/4] (coefficients of the fdlibm/Sun kernels, public domain constants),
// quadrant fix-up with integer ops on the low bits of k. |x| > 1e5 (never a joint angle in practice)
// falls back to CUDA's sincos so the full double range stays correct.
#ifndef OMO_MATH_CUH
#define OMO_MATH_CUH

#ifndef __CUDACC_RTC__
#include <cuda_runtime.h>
#include <math.h>
#else
#ifndef INFINITY
#define INFINITY __longlong_as_double(0x7ff0000000000000LL)
#endif
#endif

namespace omo {

// The polynomial / reduction constants live in constant memory: a DFMA takes a constant-bank operand directly, whereas a
// 64-bit literal costs two UMOVs per use -- ~200 extra issue slots per thread on the stick-figure walk.
__constant__ double c_sincos[18] = {
    6.36619772367581382433e-01, 6755399441055744.0,                                        // 0: 2/pi  1: rounding magic
    1.57079632679489655800e+00, 6.12323399573676603587e-17, -1.49738490485916983692e-33,   // 2-4: pi/2 in three pieces
    1.58969099521155010221e-10, -2.50507602534068634195e-08, 2.75573137070700676789e-06,   // 5-10: sin kernel S6..S1
    -1.98412698298579493134e-04, 8.33333333332248946124e-03, -1.66666666666666324348e-01,
    -1.13596475577881948265e-11, 2.08757232129817482790e-09, -2.75573143513906633035e-07,  // 11-16: cos kernel C6..C1
    2.48015872894767294178e-05, -1.38888888888741095749e-03, 4.16666666666666019037e-02,
    1.0e5};

// out of line: the full-range routine is ~90 instructions that no joint angle ever executes -- inlined at each of a walk's
// sin/cos sites it would be a seventh of the specialised kernel's code
__device__ __noinline__ double2 sincos_full_range(const double x)
{
    double2 r;
    sincos(x, &r.x, &r.y);
    return r;   // by value: the callers' sin / cos stay in registers
}

__device__ __forceinline__ void sincos_angle(const double x, double* __restrict__ s_out, double* __restrict__ c_out)
{
    if (!(fabs(x) <= c_sincos[17])) { const double2 r = sincos_full_range(x); *s_out = r.x; *c_out = r.y; return; }   // also catches NaN / inf
    // k = nearest integer to x * 2/pi (magic-number rounding), r = x - k*pi/2 in three exact-product steps
    const double t = fma(x, c_sincos[0], c_sincos[1]);
    const int k = __double2loint(t);
    const double kd = t - c_sincos[1];
    double r = fma(-kd, c_sincos[2], x);     // pi/2 rounded to 53 bits
    r = fma(-kd, c_sincos[3], r);            // next 53 bits
    r = fma(-kd, c_sincos[4], r);           // and the next
    const double z = r * r;
    // sin kernel: r + r*z*(S1 + z*(S2 + ... ))
    double ps = fma(z, c_sincos[5], c_sincos[6]);
    ps = fma(z, ps, c_sincos[7]);
    ps = fma(z, ps, c_sincos[8]);
    ps = fma(z, ps, c_sincos[9]);
    ps = fma(z, ps, c_sincos[10]);
    const double sn = fma(r * z, ps, r);
    // cos kernel: 1 - z/2 + z*z*(C1 + z*(C2 + ...)), evaluated as a + (((1 - a) - hz) + z*rc) with a = 1 - hz
    double pc = fma(z, c_sincos[11], c_sincos[12]);
    pc = fma(z, pc, c_sincos[13]);
    pc = fma(z, pc, c_sincos[14]);
    pc = fma(z, pc, c_sincos[15]);
    pc = fma(z, pc, c_sincos[16]);
    const double hz = 0.5 * z;
    const double a = 1.0 - hz;
    const double cs = a + (((1.0 - a) - hz) + z * (z * pc));
    // quadrant: k mod 4 = 0: (s, c)  1: (c, -s)  2: (-s, -c)  3: (-c, s)
    double so = (k & 1) ? cs : sn;
    double co = (k & 1) ? sn : cs;
    if (k & 2) so = -so;
    if ((k + 1) & 2) co = -co;
    *s_out = so;
    *c_out = co;
}

// Out-of-line forms (one copy of the code per kernel instead of one per keyed angle; OMO_SINCOS_CALL in omo_walk.cuh): the same
// operations on the same operands -- same bits as the inline form. Six doubles come back in registers.
struct SinCos3 { double s0, c0, s1, c1, s2, c2; };
__device__ __noinline__ SinCos3 sincos_angle3(const double x0, const double x1, const double x2)
{
    SinCos3 r;
    sincos_angle(x0, &r.s0, &r.c0);
    sincos_angle(x1, &r.s1, &r.c1);
    sincos_angle(x2, &r.s2, &r.c2);
    return r;
}
__device__ __noinline__ double2 sincos_angle1(const double x)
{
    double2 r;
    sincos_angle(x, &r.x, &r.y);
    return r;
}

}  // namespace omo
#endif
