// objective_kernel: the fused sampling -> stencil -> RNEA -> cost kernel, table-driven form. It interprets the scene
// tables in constant memory and works for any scene; the steps themselves live in omo_walk.cuh and are shared with
// the scene-specialised (NVRTC) kernel of omo_jit.cpp, which emits the same step sequence as straight-line code.
#ifndef OMO_OBJECTIVE_CUH
#define OMO_OBJECTIVE_CUH

#include "omo_walk.cuh"

namespace omo {

__device__ __forceinline__ int meta_byte(const int w, const int sh) { const int v = (w >> sh) & 0xff; return v == 0xff ? -1 : v; }

template <bool FK, bool IK, bool OUT, bool COUNT, bool DEEP = false>
__global__ void __launch_bounds__(OMO_NT, DEEP ? 1 : 4) objective_kernel(const OmoEvalParams P)
{
    extern __shared__ double smem[];
    const OmoDevScene& S = c_scene;
    typedef Walker<FK, IK, OUT, COUNT, DEEP ? OMO_DEEP_SLOTS : 0, 0, false, DEEP> W;
    constexpr int TREE = W::TREE;
    W w(P);
    w.begin(smem, DEEP ? P.n_smem_slots : (FK ? S.n_slots_tree[TREE] : S.n_slots_nofk_tree[TREE]));

    const int n_act = S.n_act[TREE];
    int bp = 0;                               // the body visited last (starts at the root)
    int4 mp = S.meta[TREE][0];
#pragma unroll 1
    for (int ai = 0; ai < n_act; ++ai) {
        const int b = S.act[TREE][ai];
        const int4 m = S.meta[TREE][b];
        const int par = m.x & 0xff;
        const bool ident = (m.y >> 3) & 1;
        // ---- leave the previous body: park it (descending) or fold the finished subtree(s) back ----------------
        if (par == bp) {
            const int fs = meta_byte(mp.x, 8), va = meta_byte(mp.x, 16), fk = meta_byte(mp.x, 24);
            if (fs >= 0) w.park_fs(fs, (mp.y >> 3) & 1);
            if (va >= 0) w.park_va(va);
            if (FK && fk >= 0) w.park_fk(fk);
        } else {
            w.fold_leaf(bp, (mp.y >> 3) & 1);
            int k = mp.x & 0xff;
#pragma unroll 1
            while (k != par) {
                const int4 mk = S.meta[TREE][k];
                w.fold_up(k, (mk.x >> 8) & 0xff, (mk.y >> 3) & 1);
                k = mk.x & 0xff;
            }
            const int4 ms = S.meta[TREE][par];
            if (par != 0) { w.fold_into((ms.x >> 8) & 0xff); w.load_va((ms.x >> 16) & 0xff); }
            else { w.fold_root(); w.set_root(); }
            if (FK) w.load_fk((ms.x >> 24) & 0xff);
        }
        // ---- IK chain rooted here, curve group starting here ------------------------------------------------------
        if (IK) {
            const int h = ((m.y >> 24) & 0xf) - 1;
            if (h >= 0) w.ik_chain(h);
        }
        const int grp = ((m.y >> 8) & 0xff) - 1;
        if (grp >= 0) {
            const int g0 = S.grp_begin[grp], g1 = S.grp_begin[grp + 1];
#pragma unroll 1
            for (int c0 = g0; c0 < g1; c0 += 3)
                w.curve_triple(c0, S.grp_list[c0], c0 + 1 < g1 ? S.grp_list[c0 + 1] : -1, c0 + 2 < g1 ? S.grp_list[c0 + 2] : -1);
            w.group_sync();
        }
        // ---- the body itself ------------------------------------------------------------------------------------------
        if (ident) w.body_ident(b);
        else w.body_general(b, m.y & 7, m.z & 0xff, (m.z >> 8) & 0xff, (m.z >> 16) & 0xff, IK ? ((m.y >> 16) & 0xf) - 1 : -1, (m.y >> 20) & 0xf);
        if (FK) w.fk_step(b, (m.y >> 4) & 1, ident);
        mp = m; bp = b;
    }

    // ---- fold everything that is still open back to the root ------------------------------------------------------
    if (n_act > 0) {
        w.fold_leaf(bp, (mp.y >> 3) & 1);
        int k = mp.x & 0xff;
#pragma unroll 1
        while (k != 0) {
            const int4 mk = S.meta[TREE][k];
            w.fold_up(k, (mk.x >> 8) & 0xff, (mk.y >> 3) & 1);
            k = mk.x & 0xff;
        }
        w.fold_root();
    }
    w.finish(smem);
}

}  // namespace omo
#endif
