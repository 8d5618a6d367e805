// niftyb200 -- host-side dispatch of the persistent TMA-fed strided pass (template definition).
#pragma once

#include "nb2_fast_impl.cuh"
#include "nb2_tma.cuh"

namespace nb2 {

// Persistent TMA-fed variant (nb2_tma.cuh) for the plain (un-remapped) passes of 256..1024 rows when
// there are enough 128-byte-wide tiles to keep every SM busy.  Developer knobs, read ONCE when the plan is
// created (nb2_plan::knobs): NB2_TMA=0 falls back to StridedPass, NB2_TMA_SCATTER=1, NB2_TMA_FX=0.
template <typename T>
int run_strided_tma(const nb2_plan* pl, int axis, int logn, long long inner, long long outer, const void* in, void* out,
                    const nb2_pointop* pro, rt_stream st, bool* done, const RowRemap* rm) {
    *done = false;
    if (!pl->knob_tma) return 0;
    constexpr int W = sizeof(T) == 4 ? 16 : 8;
    if (logn < 8 || logn > 10 || inner % W || !aligned16(in) || !tma_available()) return 0;
    if (pro && pro->amp_mode == 1) return 0;   // table gather: the batched loads of StridedPass are faster
    if (rm && rm->out_coloff && (1ll << pl->g.logM) < W) return 0;   // a tile must not straddle column blocks
    if (rm && rm->out_rowoff) {
        // Scattered (peer-memory) stores: measured at 2 and 8 GPUs the many small CTAs of StridedPass hide the
        // NVLink store latency better than one persistent CTA per SM (0.69 vs 0.75 ms per sending pass at
        // 8 GPUs, profiles/r1_prof_slab_n8_p2p*.log), so the TMA kernel takes them only on request
        if (!pl->knob_tma_scatter) return 0;
    }
    const long long ntiles = outer * (inner / W);
    if (ntiles < 2ll * rt_sm_count()) return 0;
    // symmetric streamed amplitude field: staged through shared memory by TMA (NB2_TMA_FX=0: register loads)
    bool fx = pro && pro->amp_mode == 2 && (pro->amp_flags & NB2_AMP_SYMMETRIC) && aligned16(pro->amp) &&
              !(axis == 0 ? pl->g.shift0 : pl->g.shift1);
    if (!pl->knob_tma_fx) fx = false;
    // fused direction update: the field (if any) must come through X, the registers stream r
    if (pro && pro->pre_add && pro->amp_mode == 2 && !fx) return 0;
    int rc = 0;
#define NB2_RUN_TMA(A, C)                                        \
    {                                                            \
        typedef StridedTma<T, A, C> K;                           \
        typename K::Params p;                                    \
        p.in = (const cplx<T>*)in;                               \
        p.out = (cplx<T>*)out;                                   \
        p.tw = (const cplx<T>*)(axis == 0 ? pl->tw0 : pl->tw1);  \
        p.inner = inner;                                         \
        p.tiles_per_outer = inner / W;                           \
        p.axis = axis;                                           \
        p.has_pro = pro ? 1 : 0;                                 \
        p.pro = to_pointop<T>(pro, nullptr);                     \
        p.g = pl->g;                                             \
        p.out_rowoff = rm ? rm->out_rowoff : nullptr;            \
        p.out_ooff = rm ? rm->out_ooff : nullptr;                \
        p.out_coloff = rm ? rm->out_coloff : nullptr;            \
        p.out_ostride = rm ? rm->out_ostride : 0;                \
        p.fx = fx ? 1 : 0;                                       \
        p.f_outer_mask = axis == 0 ? 0 : (1ll << pl->g.logN0) - 1; \
        rc = rt_launch_tma<K, T>(ntiles, outer, st, p);          \
    }
    switch (logn) {
        case 8: NB2_RUN_TMA(16, 16) break;
        case 9: NB2_RUN_TMA(32, 16) break;
        default: NB2_RUN_TMA(32, 32) break;
    }
#undef NB2_RUN_TMA
    *done = rc == 0;
    return rc;
}

}  // namespace nb2
