// niftyb200 -- host-side dispatch of the cluster strided pass (nb2_cluster.cuh) for 2048- and 4096-point axes.
#pragma once

#include "nb2_cluster.cuh"
#include "nb2_fast_impl.cuh"

namespace nb2 {

// Plain (un-remapped) passes over all columns of axes of 2048 (2 CTAs) and 4096 (4 CTAs) points.  Developer
// knob NB2_CLUSTER=0 (read once at plan creation) falls back to StridedPass.
template <typename T>
int run_strided_cluster(const nb2_plan* pl, int axis, int logn, long long inner, long long outer, const void* in,
                        void* out, const nb2_pointop* pro, rt_stream st, bool* done) {
    *done = false;
    if (!pl->knob_cluster) return 0;
    constexpr int W = sizeof(T) == 4 ? 16 : 8;
    if (logn < 11 || logn > 12 || inner % W || !aligned16(in) || !tma_available()) return 0;
    if (pro && (pro->amp_mode == 1 || pro->bin_x)) return 0;   // table gather: StridedPass
    const long long ntiles = outer * (inner / W);
    if (ntiles < rt_sm_count() / 2) return 0;
    int rc = 0;
#define NB2_RUN_CLUSTER(CL, A, C)                                \
    {                                                            \
        typedef StridedCluster<T, CL, A, C> K;                   \
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
        rc = rt_launch_cluster<K, T, CL>(ntiles, outer, st, p);  \
    }
    if (logn == 11) NB2_RUN_CLUSTER(2, 32, 32)
    else NB2_RUN_CLUSTER(4, 32, 32)
#undef NB2_RUN_CLUSTER
    *done = rc == 0;
    return rc;
}

}  // namespace nb2
