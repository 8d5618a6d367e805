// niftyb200 -- host-side dispatch of the contiguous-axis c2c kernel (template definition).
#pragma once

#include "nb2_c2c.cuh"
#include "nb2_fast_impl.cuh"

namespace nb2 {

// rows of L = 2^logl complex numbers: out = scale * FFT(in) along the rows (conj_in / conj_out: see nb2_c2c.cuh)
template <typename T>
int run_rows_c2c(const void* in, void* out, const void* tw, int logl, long long nrows, int conj_in, int conj_out,
                 double scale, rt_stream st) {
    int rc = 0;
#define NB2_RUN_C2C(A, B, C)                                                                 \
    {                                                                                        \
        typedef RowC2C<T, A, B, C> K;                                                        \
        int G = 1;                                                                           \
        while (G < 256 && 2ll * G <= nrows && K::smem_bytes(2 * G) <= 36 * 1024 &&           \
               (long long)G * (B * C) < 2 * K::MAX_THREADS)                                  \
            G *= 2;                                                                          \
        typename K::Params p;                                                                \
        p.in = (const cplx<T>*)in;                                                           \
        p.out = (cplx<T>*)out;                                                               \
        p.tw = (const cplx<T>*)tw;                                                           \
        p.nrows = nrows;                                                                     \
        p.G = G;                                                                             \
        p.conj_in = conj_in;                                                                 \
        p.conj_out = conj_out;                                                               \
        p.scale = (T)scale;                                                                  \
        const size_t smem = K::smem_bytes(G);                                                \
        if (smem > kSmemLimit) {                                                             \
            rc = NB2_ERR_UNSUPPORTED;                                                        \
        } else {                                                                             \
            const int nthreads = std::min((int)K::MAX_THREADS, round32((long long)G * (B * C))); \
            rc = rt_launch<K>((nrows + G - 1) / G, nthreads, smem, st, p);                   \
        }                                                                                    \
    }
    NB2_PLAN_SWITCH(logl, NB2_RUN_C2C)
#undef NB2_RUN_C2C
    return rc;
}

}  // namespace nb2
