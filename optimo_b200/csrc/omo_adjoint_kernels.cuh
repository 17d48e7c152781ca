// CUDA kernels of the analytic gradient (omo_adjoint.cuh explains the chain rule and holds the arithmetic).
// One objective's worth of samples; latency-sized launches, plain table-driven code (the scene table is c_scene).
#ifndef OMO_ADJOINT_KERNELS_CUH
#define OMO_ADJOINT_KERNELS_CUH

#include "omo_adjoint.cuh"
#include "omo_kernels.cuh"

namespace omo {

struct OmoAdjParams {
    const int* sub_end;            // [J] one past the last joint of each joint's subtree (DFS pre-order)
    const double *q, *qd, *qdd, *tau;   // [N_s][3J] of the evaluated x (the OptimizationCache arrays of the per-sample-output kernel)
    const double* root_e;          // [N_s][3] Euler angles of the root joint's own channels (null: its constants)
    const double* wcost;           // [N_s]
    const double* p_orig;          // [3J][N_s] or null
    const double* tgrid;           // [N_s]
    double w_traj, dt;
    int N_s, n_keyed, n_cp, D, cyclic;
    // curves
    const double* x;               // [D] the evaluated descriptor
    const int *cp_of_pos, *first_of_pos, *last_of_pos, *pos_of_cp, *ci_of_pos;
    // work arrays
    double* L;                     // [3][L_rows][N_s]   d cost_i / d (q, qd, qdd)_c at sample i; rows = keyed channels (forward mode: L_rows =
                                   //                    n_keyed) or ALL joint channels 3j + axis (reverse mode: L_rows = 3J, L_by_channel = 1)
    int L_rows, L_by_channel;
    double* ws;                    // reverse mode: [J][fields][N_s] work memory of omo_ad::sample_reverse
    double* G;                     // [n_keyed][N_s]     d f / d q_c(t_i)
    double* contrib;               // [n_keyed][N_s][10] G times the value's partials w.r.t. the segment's two control points
    int* seg_m;                    // [n_keyed][N_s]     number of keys with time <= t_i
    double* gfull;                 // [D]
    // output
    const int* free_idx; int F;
    const double *x0, *wparam;
    double* gout;                  // [F]
};

// 1. one thread per (kind, keyed channel, sample)
template <int MAXSUB>
__global__ void __launch_bounds__(128) adj_tangent_kernel(const OmoAdjParams P)
{
    const long long idx = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    const long long total = 3LL * P.n_keyed * P.N_s;
    if (idx >= total) return;
    const int i = (int)(idx % P.N_s);
    const int ci = (int)((idx / P.N_s) % P.n_keyed);
    const int kind = (int)(idx / ((long long)P.N_s * P.n_keyed));
    const OmoDevScene& S = c_scene;
    const int ch = S.grp_list[ci], j = ch / 3, axis = ch - 3 * j;
    const size_t dof = (size_t)3 * S.J;
    const bool interior = (i >= 2) && (i <= P.N_s - 3);
    double re[3] = {S.chan_const[0], S.chan_const[1], S.chan_const[2]};
    if (P.root_e) { re[0] = P.root_e[3 * (size_t)i]; re[1] = P.root_e[3 * (size_t)i + 1]; re[2] = P.root_e[3 * (size_t)i + 2]; }
    P.L[idx] = omo_ad::sample_tangent<MAXSUB>(S, P.sub_end, j, axis, kind, P.q + i * dof, P.qd + i * dof, P.qdd + i * dof, P.tau + i * dof,
                                              P.wcost[i], P.p_orig ? P.w_traj : 0.0, P.dt, P.p_orig ? P.p_orig + i : nullptr, P.N_s, re, interior);
}

// 1'. reverse mode: one thread per sample, every joint channel's three partials from one forward sweep, one fold-back and their
// adjoints (omo_ad::sample_reverse); consecutive threads are consecutive samples in every array it touches
template <bool TRAJ>
__global__ void __launch_bounds__(128) adj_reverse_kernel(const OmoAdjParams P)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= P.N_s) return;
    const OmoDevScene& S = c_scene;
    const size_t dof = (size_t)3 * S.J, N = (size_t)P.N_s;
    const bool interior = (i >= 2) && (i <= P.N_s - 3);
    double re[3] = {S.chan_const[0], S.chan_const[1], S.chan_const[2]};
    if (P.root_e) { re[0] = P.root_e[3 * (size_t)i]; re[1] = P.root_e[3 * (size_t)i + 1]; re[2] = P.root_e[3 * (size_t)i + 2]; }
    omo_ad::sample_reverse<TRAJ>(S, P.q + i * dof, P.qd + i * dof, P.qdd + i * dof, P.wcost[i], TRAJ ? P.w_traj : 0.0, P.dt, TRAJ ? P.p_orig + i : nullptr,
                                 (long long)N, re, interior, P.ws + i, (long long)N, P.L + i, P.L + dof * N + i, P.L + 2 * dof * N + i, (long long)N);
}

// 2. transposed stencils: one thread per (keyed channel, sample)
__global__ void adj_stencil_kernel(const OmoAdjParams P)
{
    const long long idx = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= (long long)P.n_keyed * P.N_s) return;
    const int i = (int)(idx % P.N_s), ci = (int)(idx / P.N_s), N = P.N_s;
    const size_t row = P.L_by_channel ? (size_t)c_scene.grp_list[ci] : (size_t)ci;
    const double* __restrict__ Lq = P.L + row * N;
    const double* __restrict__ Lv = P.L + ((size_t)P.L_rows + row) * N;
    const double* __restrict__ La = P.L + ((size_t)2 * P.L_rows + row) * N;
    double g = Lq[i];
    for (int m = max(i - 2, 0); m <= min(i + 2, N - 1); ++m)
        g += omo_ad::d1_coef(m, i, N, P.dt) * Lv[m] + omo_ad::d2_coef(m, i, N, P.dt) * La[m];
    P.G[idx] = g;
}

// 3. curve partials: one thread per (keyed channel, sample)
__global__ void adj_bezier_kernel(const OmoAdjParams P)
{
    const long long idx = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= (long long)P.n_keyed * P.N_s) return;
    const int i = (int)(idx % P.N_s), ci = (int)(idx / P.N_s);
    const OmoDevScene& S = c_scene;
    const int ch = S.grp_list[ci], pos0 = S.chan_pos[ch], n = S.chan_n[ch];
    const double t = P.tgrid[i];
    int lo = 0, hi = n;   // number of keys with time <= t (chains are time-sorted here)
    while (lo < hi) {
        const int mid = (lo + hi) >> 1;
        if (P.x[5 * P.cp_of_pos[pos0 + mid]] <= t) lo = mid + 1; else hi = mid;
    }
    const int m = lo;
    P.seg_m[idx] = m;
    double* __restrict__ out = P.contrib + (size_t)idx * 10;
    const double g = P.G[idx];
    for (int k = 0; k < 10; ++k) out[k] = 0.0;
    if (m == 0) { out[6] = g; return; }           // before the first key: the first key's value (curve.cpp:60-64)
    if (m == n) { out[1] = g; return; }           // after the last key: the last key's value (curve.cpp:66-71)
    const int cps = P.cp_of_pos[pos0 + m - 1], cpe = P.cp_of_pos[pos0 + m];
    auto theta = [&](const int cp, const int pos) {
        double th = P.x[5 * cp + 2];
        if (P.cyclic) {
            const int fi = P.first_of_pos[pos], la = P.last_of_pos[pos];
            if (cp == fi || cp == la) th = 0.5 * (P.x[5 * fi + 2] + P.x[5 * la + 2]);
        }
        return th;
    };
    double part[10];
    omo_ad::bezier_partials(P.x + 5 * cps, P.x + 5 * cpe, theta(cps, pos0 + m - 1), theta(cpe, pos0 + m), t, part);
    for (int k = 0; k < 10; ++k) out[k] = g * part[k];
}

// 4. one thread per (chain position, field): ordered sum over the samples of the two adjacent segments
__global__ void adj_reduce_kernel(const OmoAdjParams P)
{
    const int idx = blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= 5 * P.n_cp) return;
    const int pos = idx / 5, field = idx - 5 * pos;
    const int ci = P.ci_of_pos[pos];
    const int cp = P.cp_of_pos[pos];
    if (ci < 0) { P.gfull[5 * cp + field] = 0.0; return; }   // not a joint channel (IK target curves are not covered)
    const OmoDevScene& S = c_scene;
    const int ch = S.grp_list[ci], p = pos - S.chan_pos[ch], N = P.N_s;
    const int* __restrict__ M = P.seg_m + (size_t)ci * N;
    auto lower = [&](const int val) {   // first sample with m >= val
        int lo = 0, hi = N;
        while (lo < hi) { const int mid = (lo + hi) >> 1; if (M[mid] < val) lo = mid + 1; else hi = mid; }
        return lo;
    };
    const double* __restrict__ C = P.contrib + (size_t)ci * N * 10;
    double acc = 0.0;
    // end point of segment p-1 -> p: samples with m == p
    for (int i = lower(p), e = lower(p + 1); i < e; ++i) acc += C[(size_t)i * 10 + 5 + field];
    // start point of segment p -> p+1: samples with m == p + 1
    for (int i = lower(p + 1), e = lower(p + 2); i < e; ++i) acc += C[(size_t)i * 10 + field];
    P.gfull[5 * cp + field] = acc;
}

// 5. free entries: cyclic end tangents are one shared parameter (curve.cpp:179-188), D_param added (core_rbdl.cpp:385)
__global__ void adj_output_kernel(const OmoAdjParams P)
{
    const int k = blockIdx.x * blockDim.x + threadIdx.x;
    if (k >= P.F) return;
    const int i = P.free_idx[k], cp = i / 5, field = i - 5 * cp;
    double g = P.gfull[i];
    if (P.cyclic && field == 2) {
        const int pos = P.pos_of_cp[cp], fi = P.first_of_pos[pos], la = P.last_of_pos[pos];
        if (cp == fi || cp == la) g = 0.5 * (P.gfull[5 * fi + 2] + P.gfull[5 * la + 2]);
    }
    g += 2.0 * P.wparam[i] * (P.x[i] - P.x0[i]);
    P.gout[k] = g;
}

}  // namespace omo
#endif
