// Loop form of the scene-specialised kernel: the tree walk of omo_objective.cuh as ONE loop over the bodies, with what the
// straight-line form gets from specialisation kept where it does not cost code size -- staged curve groups (cp.async
// windows, looped triples), the compact / hybrid stack placement (slot -> shared-memory row or per-thread slot through the
// `place` table the generator emits), compile-time instance flags.
//
// Why a second form: a DFMA holds the issue port for two cycles and code that does not fit the instruction caches is
// supplied at about half an instruction per cycle per SM sub-partition (profiles/micro/icache_mix). The straight-line
// form of a 64-joint tree is 1.1 MB of SASS that every warp streams through once; this loop body is a few KB that stays
// resident, and the per-body table lookups it pays for are a small fraction of a 63-body RNEA step.
#ifndef OMO_LOOP_CUH
#define OMO_LOOP_CUH

#include "omo_walk.cuh"

namespace omo {

__device__ __forceinline__ int loop_meta_byte(const int w, const int sh) { const int v = (w >> sh) & 0xff; return v == 0xff ? -1 : v; }

// W: a Walker instance with CPT = true. place[slot of the scene's plan] = first shared-memory row, or 4096 + per-thread slot.
template <class W, bool FK, bool IK>
__device__ __forceinline__ void walk_loop(W& w, const short* __restrict__ place)
{
    const OmoDevScene& S = c_scene;
    constexpr int TREE = 1;
    const int n_act = S.n_act[TREE];
    int bp = 0;                               // the body visited last (starts at the root)
    int4 mp = S.meta[TREE][0];
    if (S.n_groups > 0) {
        w.template stage_issue<true>(0, S.grp_begin[0], S.grp_begin[1] - S.grp_begin[0]);
        w.stage_wait();
        w.group_sync();
    }
#pragma unroll 1
    for (int ai = 0; ai < n_act; ++ai) {
        const int b = S.act[TREE][ai];
        const int4 m = S.meta[TREE][b];
        const int par = m.x & 0xff;
        const bool ident = (m.y >> 3) & 1;
        if (par == bp) {
            const int fs = loop_meta_byte(mp.x, 8), va = loop_meta_byte(mp.x, 16), fk = loop_meta_byte(mp.x, 24);
            if (fs >= 0) w.park_fs(place[fs], (mp.y >> 3) & 1);
            if (va >= 0) w.park_va(place[va]);
            if (FK && fk >= 0) w.park_fk(place[fk]);
        } else {
            w.fold_leaf(bp, (mp.y >> 3) & 1);
            int k = mp.x & 0xff;
#pragma unroll 1
            while (k != par) {
                const int4 mk = S.meta[TREE][k];
                w.fold_up(k, place[(mk.x >> 8) & 0xff], (mk.y >> 3) & 1);
                k = mk.x & 0xff;
            }
            const int4 ms = S.meta[TREE][par];
            if (par != 0) { w.fold_into(place[(ms.x >> 8) & 0xff]); w.load_va(place[(ms.x >> 16) & 0xff]); }
            else { w.fold_root(); w.set_root(); }
            if (FK) w.load_fk(place[(ms.x >> 24) & 0xff]);
        }
        if (IK) {
            const int h = ((m.y >> 24) & 0xf) - 1;
            if (h >= 0) w.ik_chain(h);
        }
        const int grp = ((m.y >> 8) & 0xff) - 1;
        if (grp >= 0) {
            const int g0 = S.grp_begin[grp], g1 = S.grp_begin[grp + 1];
            if (grp + 1 < S.n_groups) w.stage_issue((grp + 1) & 1, g1, S.grp_begin[grp + 2] - g1);
            const int n3 = (g1 - g0) / 3, rem = (g1 - g0) - 3 * n3;
            if (rem == 0) w.template curve_group<0>(grp & 1, g0, n3);
            else if (rem == 1) w.template curve_group<1>(grp & 1, g0, n3);
            else w.template curve_group<2>(grp & 1, g0, n3);
            w.stage_wait();
            w.group_sync();
        }
        if (ident) w.body_ident(b);
        else w.body_general(b, m.y & 7, m.z & 0xff, (m.z >> 8) & 0xff, (m.z >> 16) & 0xff, IK ? ((m.y >> 16) & 0xf) - 1 : -1, (m.y >> 20) & 0xf);
        if (FK) w.fk_step(b, (m.y >> 4) & 1, ident);
        mp = m; bp = b;
    }
    if (n_act > 0) {
        w.fold_leaf(bp, (mp.y >> 3) & 1);
        int k = mp.x & 0xff;
#pragma unroll 1
        while (k != 0) {
            const int4 mk = S.meta[TREE][k];
            w.fold_up(k, place[(mk.x >> 8) & 0xff], (mk.y >> 3) & 1);
            k = mk.x & 0xff;
        }
        w.fold_root();
    }
}

}  // namespace omo
#endif
