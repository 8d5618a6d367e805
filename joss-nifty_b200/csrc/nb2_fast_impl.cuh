// niftyb200 -- host-side dispatch of the two fast-path pass kernels (template definitions).
#pragma once

#include "nb2_passes.cuh"
#include "nb2_plan.h"
#include "nb2_vector.cuh"

namespace nb2 {

// ---------------------------------------------------------------- radix plans
// log2(n) -> (R1, R2, R3)
#define NB2_PLAN_SWITCH(LOGN, CALL)                 \
    switch (LOGN) {                                 \
        case 1: CALL(2, 1, 1); break;               \
        case 2: CALL(4, 1, 1); break;               \
        case 3: CALL(8, 1, 1); break;               \
        case 4: CALL(16, 1, 1); break;              \
        case 5: CALL(32, 1, 1); break;              \
        case 6: CALL(8, 1, 8); break;               \
        case 7: CALL(16, 1, 8); break;              \
        case 8: CALL(16, 1, 16); break;             \
        case 9: CALL(32, 1, 16); break;             \
        case 10: CALL(32, 1, 32); break;            \
        case 11: CALL(16, 8, 16); break;            \
        case 12: CALL(16, 16, 16); break;           \
        case 13: CALL(32, 16, 16); break;           \
        default: rc = NB2_ERR_UNSUPPORTED; break;   \
    }

// choose tile width for a strided pass over rows of `inner` complex elements
template <typename T>
int choose_logw(int logn, long long inner, bool remote_out = false) {
    const size_t col = size_t(pitch_of(logn)) * sizeof(cplx<T>);  // bytes per tile column
    int logw = 0;
    while ((1ll << (logw + 1)) <= inner && (col << (logw + 1)) <= 72 * 1024) ++logw;
    // row segments of at least 64 bytes; 128 bytes when consecutive rows are >= 1 MiB apart
    // (measured on B200, 1024^3 fp32: axis-0 pass 2.26 ms with 128 B vs 2.89 ms with 64 B segments,
    //  axis-1 pass 1.82 ms with 64 B / two CTAs per SM vs 2.14 ms with 128 B / one CTA per SM)
    // stores that cross NVLink into a peer's memory (slab exchange): 128 B segments
    // (2 GPUs, 1024^3 fp32, axis-1 pass with half of the rows remote: 1.52 ms vs 2.42 ms with 64 B)
    const size_t seg = (remote_out || inner * sizeof(cplx<T>) >= (size_t(1) << 20)) ? 128 : 64;
    const int want = ilog2(seg / sizeof(cplx<T>));
    while (logw < want && (1ll << (logw + 1)) <= inner && (col << (logw + 1)) <= kSmemLimit) ++logw;
    return logw;
}

// Persistent TMA-fed variant (nb2_tma.cuh); defined in nb2_fast_tma_impl.cuh and instantiated in its own
// translation units (nb2_fast_tma_f32.cu / _f64.cu) so that the build compiles it in parallel.
template <typename T>
int run_strided_tma(const nb2_plan* pl, int axis, int logn, long long inner, long long outer, const void* in, void* out,
                    const nb2_pointop* pro, rt_stream st, bool* done, const RowRemap* rm = nullptr);

template <typename T>
int run_strided(const nb2_plan* pl, int axis, const void* in, void* out, const nb2_pointop* pro, rt_stream st,
                const RowRemap* rm) {
    const int logn = axis == 0 ? pl->g.logN0 : pl->g.logN1;
    if (logn == 0) {  // nothing to transform; still apply prologue / copy
        return NB2_ERR_ARG;
    }
    const long long inner = axis == 0 ? (1ll << (pl->g.logN1 + pl->g.logM)) : (1ll << pl->g.logM);
    // natural input, all columns: the TMA kernel (its stores may be scattered through the out tables)
    if (!rm || (!rm->in_rowoff && !rm->ncols && !rm->col0)) {
        bool done = false;
        const int rc = run_strided_tma<T>(pl, axis, logn, inner, axis == 0 ? pl->g.batch : (pl->g.batch << pl->g.logN0),
                                          in, out, pro, st, &done, rm);
        if (rc || done) return rc;
    }
    const long long outer = axis == 0 ? pl->g.batch : (pl->g.batch << pl->g.logN0);
    const long long ncols = (rm && rm->ncols) ? rm->ncols : inner;
    const long long run = (rm && rm->run) ? rm->run : ncols;
    if (ncols > inner || (ncols & (ncols - 1)) || (run & (run - 1)) || run > ncols) {
        last_error() = "strided pass: column subset must consist of power-of-two runs";
        return NB2_ERR_ARG;
    }
    int logw = choose_logw<T>(logn, inner, rm && rm->out_rowoff);
    while ((1ll << logw) > run) --logw;
    if (rm && rm->out_coloff && logw > pl->g.logM) logw = pl->g.logM;   // a tile must not straddle column blocks
    int rc = 0;
#define NB2_RUN_STRIDED(A, B, C)                                                              \
    {                                                                                         \
        typedef StridedPass<T, A, B, C> K;                                                    \
        typename K::Params p;                                                                 \
        p.in = (const cplx<T>*)in;                                                            \
        p.out = (cplx<T>*)out;                                                                \
        p.tw = (const cplx<T>*)(axis == 0 ? pl->tw0 : pl->tw1);                               \
        p.inner = inner;                                                                      \
        p.logw = logw;                                                                        \
        p.tiles_per_outer = ncols >> logw;                                                    \
        p.axis = axis;                                                                        \
        p.has_pro = pro ? 1 : 0;                                                              \
        p.pro = to_pointop<T>(pro, nullptr);                                                  \
        p.g = pl->g;                                                                          \
        p.in_rowoff = rm ? rm->in_rowoff : nullptr;                                           \
        p.out_rowoff = rm ? rm->out_rowoff : nullptr;                                         \
        p.out_ooff = rm ? rm->out_ooff : nullptr;                                             \
        p.out_coloff = rm ? rm->out_coloff : nullptr;                                         \
        p.in_ostride = rm ? rm->in_ostride : 0;                                               \
        p.out_ostride = rm ? rm->out_ostride : 0;                                             \
        p.col0 = rm ? rm->col0 : 0;                                                           \
        p.run_stride = rm ? rm->run_stride : 0;                                               \
        p.logrun = ilog2(run);                                                                \
        const size_t smem = K::smem_bytes(logw);                                              \
        if (smem > kSmemLimit) {                                                              \
            rc = NB2_ERR_UNSUPPORTED;                                                         \
        } else {                                                                              \
            const long long nb1 = (long long)(B * C) << logw;                                 \
            const int nthreads = std::min((int)K::MAX_THREADS, round32(nb1));                 \
            rc = rt_launch<K>(outer * p.tiles_per_outer, nthreads, smem, st, p);              \
        }                                                                                     \
    }
    NB2_PLAN_SWITCH(logn, NB2_RUN_STRIDED)
#undef NB2_RUN_STRIDED
    return rc;
}

template <typename T>
int choose_G(int logm, long long Q) {
    const size_t row = size_t(pitch_of(logm)) * sizeof(cplx<T>);
    int G = 1;
    while (G < 16 && 2 * G <= Q && size_t(4 * G) * row <= 36 * 1024) G *= 2;
    return G;
}

template <typename T>
int rows_pairs_per_cta(const nb2_plan* pl) {
    const RowPairs rp = row_pairs(pl->g);
    return choose_G<T>(pl->g.logM, rp.Q);
}

// The row pass is compiled in three parts (separate translation units, so that the build runs them in
// parallel): PART 0 = rows of up to 256 scalars, PART 1 = 512..2048 scalars (with the compile-time
// epilogues), PART 2 = longer rows (three-step plans).
template <typename T, int PART>
int run_rows_part(nb2_plan* pl, const void* in, void* out, int conv, const nb2_pointop* pro, const nb2_pointop* epi,
                  rt_stream st);

template <typename T>
int run_rows(nb2_plan* pl, const void* in, void* out, int conv, const nb2_pointop* pro, const nb2_pointop* epi,
             rt_stream st) {
    const int logm = pl->g.logM;
    if (logm <= 7) return run_rows_part<T, 0>(pl, in, out, conv, pro, epi, st);
    if (logm <= 10) return run_rows_part<T, 1>(pl, in, out, conv, pro, epi, st);
    return run_rows_part<T, 2>(pl, in, out, conv, pro, epi, st);
}

template <typename T, int PART>
int run_rows_part(nb2_plan* pl, const void* in, void* out, int conv, const nb2_pointop* pro, const nb2_pointop* epi,
                  rt_stream st) {
    const int logm = pl->g.logM;
    const RowPairs rp = row_pairs(pl->g);
    const int G = rows_pairs_per_cta<T>(pl);
    const long long ctas_per_item = (rp.Q + G - 1) / G;
    const long long grid = ctas_per_item * pl->g.batch;
    const bool want_dot = epi && epi->dot;
    if (want_dot) {
        if (!epi->dot_out) {
            last_error() = "nb2_pointop.dot set without dot_out";
            return NB2_ERR_ARG;
        }
        if (grid > pl->npartials) {
            rt_sync(st);
            rt_free(pl->partials);
            pl->partials = nullptr;
            pl->npartials = 0;
            if (rt_malloc((void**)&pl->partials, grid * sizeof(double))) return NB2_ERR_CUDA;
            pl->npartials = grid;
        }
    }
    int rc = 0;
#define NB2_RUN_ROWS(A, B, C) NB2_RUN_ROWS_E(A, B, C, -1)
#define NB2_RUN_ROWS_E(A, B, C, E)                                                            \
    {                                                                                         \
        typedef RowPass<T, A, B, C, E> K;                                                     \
        typename K::Params p;                                                                 \
        p.in = (const cplx<T>*)in;                                                            \
        p.out = (T*)out;                                                                      \
        p.tw = (const cplx<T>*)pl->twr;                                                       \
        p.g = pl->g;                                                                          \
        p.ctas_per_item = ctas_per_item;                                                      \
        p.G = G;                                                                              \
        p.conv = conv;                                                                        \
        p.has_pro = pro ? 1 : 0;                                                              \
        p.pro = to_pointop<T>(pro, nullptr);                                                  \
        p.epi = to_pointop<T>(epi, pl->partials);                                             \
        const long long nb1 = (long long)2 * G * B * C;                                       \
        const int nthreads = std::min((int)K::MAX_THREADS, round32(nb1));                     \
        const size_t smem = (K::smem_bytes(G, nthreads) + 15) & ~size_t(15);                  \
        if (smem > kSmemLimit) {                                                              \
            rc = NB2_ERR_UNSUPPORTED;                                                         \
        } else {                                                                              \
            rc = rt_launch<K>(grid, nthreads, smem, st, p);                                   \
        }                                                                                     \
    }
    // compile-time epilogues for the three operand sets of the hot path (plain transform, metric forward
    // `* weight`, metric adjoint `* field + add, dot`) on rows of 512..2048 scalars
    int mask = -1;
    if (logm >= 8 && logm <= 10 && !(conv & NB2_PRODUCT_LAST)) {
        if (!epi) {
            mask = 0;
        } else if (!epi->bin_x && epi->amp_mode != 1 && (!epi->dot || epi->dot == epi->add)) {
            mask = (epi->amp_mode == 2 ? 1 : 0) | (epi->mul ? 2 : 0) | (epi->add ? 4 : 0) | (epi->dot ? 8 : 0);
        }
        if (mask != 0 && mask != 2 && mask != 13) mask = -1;
    }
#define NB2_ROWS_MASKS(A, B, C)                                                  \
    if (mask == 0) NB2_RUN_ROWS_E(A, B, C, 0)                                    \
    else if (mask == 2) NB2_RUN_ROWS_E(A, B, C, 2)                               \
    else NB2_RUN_ROWS_E(A, B, C, 13)
    if constexpr (PART == 0) {
        switch (logm) {
            case 1: NB2_RUN_ROWS(2, 1, 1) break;
            case 2: NB2_RUN_ROWS(4, 1, 1) break;
            case 3: NB2_RUN_ROWS(8, 1, 1) break;
            case 4: NB2_RUN_ROWS(16, 1, 1) break;
            case 5: NB2_RUN_ROWS(32, 1, 1) break;
            case 6: NB2_RUN_ROWS(8, 1, 8) break;
            case 7: NB2_RUN_ROWS(16, 1, 8) break;
            default: rc = NB2_ERR_UNSUPPORTED; break;
        }
    } else if constexpr (PART == 1) {
        if (mask >= 0) {
            if (logm == 8) { NB2_ROWS_MASKS(16, 1, 16) }
            else if (logm == 9) { NB2_ROWS_MASKS(32, 1, 16) }
            else { NB2_ROWS_MASKS(32, 1, 32) }
        } else {
            switch (logm) {
                case 8: NB2_RUN_ROWS(16, 1, 16) break;
                case 9: NB2_RUN_ROWS(32, 1, 16) break;
                case 10: NB2_RUN_ROWS(32, 1, 32) break;
                default: rc = NB2_ERR_UNSUPPORTED; break;
            }
        }
    } else {
        switch (logm) {
            case 11: NB2_RUN_ROWS(16, 8, 16) break;
            case 12: NB2_RUN_ROWS(16, 16, 16) break;
            case 13: NB2_RUN_ROWS(32, 16, 16) break;
            default: rc = NB2_ERR_UNSUPPORTED; break;
        }
    }
#undef NB2_ROWS_MASKS
#undef NB2_RUN_ROWS_E
#undef NB2_RUN_ROWS
    if (rc) return rc;
    if (want_dot) {
        rc = launch_final_reduce(pl->partials, grid, 1, epi->dot_out, st);
    }
    return rc;
}


}  // namespace nb2
