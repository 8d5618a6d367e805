// niftyb200 -- the two FFT pass kernels of the power-of-two fast path.
//
// Genuine multi-D Hartley transform of a real (batch, N0, N1, N2) array
// (reference semantics: src/ducc_dispatch.py:46-50,75-78  H = Re F +/- Im F of the JOINT FFT):
//
//   1. the real array is *viewed* as complex with M = N2/2 elements per row
//      (z[m] = x[2m] + i x[2m+1]; no data movement),
//   2. StridedPass: complex FFT along axis 0, then along axis 1 (tiles of `w` contiguous
//      complex columns staged through shared memory, in place),
//   3. RowPass: complex FFT along the contiguous axis for a row (k0,k1) together with its
//      mirror row (-k0,-k1), then the "untangle" step recovers the spectrum of the real
//      data (even/odd split) and writes H[k] = Re F[k] +/- Im F[k] for the full rows.
//
// Every axis therefore costs exactly one read and one write of the field (2*N*s bytes), the
// nominal-sweep model of SURVEY.md section 8(d).  All passes may run in place.
//
// Each pass is an N = R1*R2*R3 point FFT done in up to three register-radix steps
// (digits n = n3 + R3*(n2 + R2*n1), k = k1 + R1*(k2 + R2*k3)); R2 == 1 means two steps,
// R2 == R3 == 1 a single step.
#pragma once

#include "nb2_common.cuh"
#include "nb2_pointwise.cuh"
#include "nb2_radix.cuh"

// occupancy targets (CTAs per SM) handed to ptxas through __launch_bounds__
#ifndef NB2_MINB256_F32
#define NB2_MINB256_F32 1   // radix-32 kernels: compiled for up to 512 threads -> <= 128 registers
#endif
#ifndef NB2_MINB_SHORT_F32
#define NB2_MINB_SHORT_F32 3   // row pass, generic epilogue, CTAs of <= 256 threads: x 3 -> <= 85 registers
#endif
#ifndef NB2_MINB_SHORT_F64
#define NB2_MINB_SHORT_F64 2   // <= 128 registers
#endif
#ifndef NB2_MINB512_F32
#define NB2_MINB512_F32 2   // radix <= 16 kernels, 512 threads: <= 64 registers
#endif
#ifndef NB2_MINB256_F64
#define NB2_MINB256_F64 1
#endif
#ifndef NB2_MINB512_F64
#define NB2_MINB512_F64 1
#endif
#ifndef NB2_TWPOW
#define NB2_TWPOW 1         // 1: inter-step twiddles by binary powering of W^r instead of table loads
#endif

namespace nb2 {

// shared-memory slot of logical row L (padded so that every step is bank-conflict free)
template <int R2, int R3> NB2_HD int slot_of(int L) {
    return L + L / R3 + (R2 > 1 ? L / (R2 * R3) : 0);
}
template <int R1, int R2, int R3> constexpr int slot_pitch() {
    return R1 * R2 * R3 + R1 * R2 + (R2 > 1 ? R1 : 0);
}

// w[k] = w1^k for k = 1..R-1 by binary powering (multiplication depth <= log2 R)
template <int R, typename T> NB2_HD void tw_powers(cplx<T> w1, cplx<T> (&w)[R]) {
    w[0] = mk<T>(T(1), T(0));
    if (R > 1) w[1] = w1;
    NB2_UNROLL
    for (int k = 2; k < R; ++k) {
        // largest power of two <= k
        int hb = 1;
        while (2 * hb <= k) hb *= 2;
        w[k] = (k == hb) ? w[hb / 2] * w[hb / 2] : w[hb] * w[k - hb];
    }
}

// Rows of a streamed amplitude field (amp_mode 2) along the transformed axis of a strided pass.
// sym (NB2_AMP_SYMMETRIC): row j > N/2 equals row N - j, so the canonical row is read instead -- with
// the default L2 policy, because its other reader (same CTA, a little earlier or later) finds it there.
template <typename T>
struct FieldRows {
    const cplx<T>* base;   // element (row 0, this thread's column)
    long long inner;
    int n;
    bool sym;
    NB2_HD FieldRows(const cplx<T>* b, long long in, int nn, bool s) : base(b), inner(in), n(nn), sym(s) {}
    NB2_HD cplx<T> load(int row) const {
        if (sym) return ldc_keep(base + (long long)(row <= n - row ? row : n - row) * inner);
        return ldc_stream(base + (long long)row * inner);
    }
};

// ============================================================================ StridedPass
template <typename T, int R1, int R2, int R3>
struct StridedPass {
    static constexpr int N = R1 * R2 * R3;
    static constexpr int MM = R2 * R3;
    static constexpr bool SINGLE = (MM == 1);
    static constexpr int TPITCH = slot_pitch<R1, R2, R3>();
    static constexpr int NPHASES = 3;
    static constexpr bool BIG = (R1 >= 32 || R3 >= 32);
    static constexpr int MAX_THREADS = (BIG && sizeof(T) == 8) ? 256 : 512;
    // three-step plans (2048 rows and more): a tile of useful width fills the shared memory of an SM, so a second
    // resident CTA is out of reach anyway -- compile for one (128 registers: the streamed operands of the fused
    // prologues are then fetched in batches instead of one dependent load after the other)
    static constexpr bool LONG = (R2 > 1);
    static constexpr int MIN_BLOCKS = sizeof(T) == 4 ? ((BIG || LONG) ? NB2_MINB256_F32 : NB2_MINB512_F32)
                                                     : (MAX_THREADS == 256 ? NB2_MINB256_F64 : NB2_MINB512_F64);

    struct Params {
        const cplx<T>* in;
        cplx<T>* out;
        const cplx<T>* tw;          // tw[j] = exp(-2 pi i j / N), j in [0, N)
        long long inner;            // complex elements between consecutive rows
        long long tiles_per_outer;  // inner >> logw
        int logw;                   // log2(tile width)
        int axis;                   // 0 or 1: which coordinate the transformed row index is
        int has_pro;
        PointOp<T> pro;
        Geom g;
        // Optional row remapping (slab decomposition: the pack / unpack of the all-to-all is folded
        // into the pass on either side of the exchange).  With a table, row j of outer item o
        // starts at  o * ostride + rowoff[j]  instead of  (o * N + j) * inner.
        // Column subset of the natural side (tiles_per_outer then counts the tiles of the subset;
        // remapped sides are dense in it): dense column j is natural column
        //   col0 + (j >> logrun) * run_stride + (j & (run - 1)),  run = 1 << logrun.
        const long long* in_rowoff;
        const long long* out_rowoff;
        // scatter form of the output remap (nb2_strided_pass_scatter): with out_rowoff, outer item o
        // starts at out_ooff[o] (instead of o * out_ostride) and natural column col lies at
        // out_coloff[col >> logM] + (col & (M - 1)) (instead of col); NULL = the default rule
        const long long* out_ooff;
        const long long* out_coloff;
        long long in_ostride, out_ostride, col0, run_stride;
        int logrun;
    };

    static size_t smem_bytes(int logw) {
        return SINGLE ? 0 : (size_t(slot_pitch<R1, R2, R3>()) << logw) * sizeof(cplx<T>);
    }
    NB2_HDC static bool enabled(int ph) { return ph == 0 || (ph == 1 && R2 > 1) || (ph == 2 && R3 > 1); }

    NB2_HD static long long tile_base(const Params& p, long long bid) {
        const long long o = bid / p.tiles_per_outer;
        const long long tile = bid - o * p.tiles_per_outer;
        const long long tc = tile << p.logw;
        return o * N * p.inner + p.col0 + (tc >> p.logrun) * p.run_stride + (tc & ((1ll << p.logrun) - 1));
    }
    NB2_HD static long long out_origin(const Params& p, long long o, long long tcol) {
        const long long oo = p.out_ooff ? ld_ro(p.out_ooff + o) : o * p.out_ostride;
        const long long cc = p.out_coloff ? ld_ro(p.out_coloff + (tcol >> p.g.logM)) + (tcol & ((1ll << p.g.logM) - 1)) : tcol;
        return oo + cc;
    }
    NB2_HD static void store_out(const Params&, cplx<T>* dst, cplx<T> v) { stc_stream(dst, v); }

    template <int PH>
    NB2_HD static void phase(const Params& p, long long bid, int tid, int nthreads, char* smem_raw) {
        cplx<T>* smem = reinterpret_cast<cplx<T>*>(smem_raw);
        const int w = 1 << p.logw;
        const long long base = tile_base(p, bid);
        const long long tcol = (bid % p.tiles_per_outer) << p.logw;
        const long long obase = p.out_rowoff ? out_origin(p, bid / p.tiles_per_outer, tcol) : base;
        if (PH == 0) {
            const int nb = MM << p.logw;
            for (int b = tid; b < nb; b += nthreads) {
                const int c = b & (w - 1), r = b >> p.logw;
                cplx<T> x[R1];
                if (p.in_rowoff) {
                    const cplx<T>* src = p.in + (bid / p.tiles_per_outer) * p.in_ostride + tcol + c;
                    const long long* ro = p.in_rowoff + r;
                    NB2_UNROLL
                    for (int n1 = 0; n1 < R1; ++n1) x[n1] = ldc_stream(src + ld_ro(ro + n1 * MM));
                } else {
                    const cplx<T>* src = p.in + base + c + (long long)r * p.inner;
                    NB2_UNROLL
                    for (int n1 = 0; n1 < R1; ++n1) x[n1] = ldc_stream(src + (long long)n1 * MM * p.inner);
                }
                if (p.has_pro && p.pro.pre_add) {
                    // fused direction update d = r + beta d (natural rows only): same element, same thread
                    const long long e0 = base + c + (long long)r * p.inner;
                    const cplx<T>* rs = reinterpret_cast<const cplx<T>*>(p.pro.pre_add) + e0;
                    cplx<T>* dd = p.pro.pre_out ? reinterpret_cast<cplx<T>*>(p.pro.pre_out) + e0 : nullptr;
                    const T beta = pre_scale_of(p.pro);
                    NB2_UNROLL
                    for (int n1 = 0; n1 < R1; ++n1) {
                        const cplx<T> rr = ldc_stream(rs + (long long)n1 * MM * p.inner);
                        x[n1].re = rr.re + beta * x[n1].re;
                        x[n1].im = rr.im + beta * x[n1].im;
                        if (dd) stc_stream(dd + (long long)n1 * MM * p.inner, x[n1]);
                    }
                }
                if (p.has_pro) {
                    // amplitude factors are fetched in batches of PB rows, two batches in flight,
                    // so that their latency overlaps instead of serialising per row
                    constexpr int PB = R1 >= 8 ? 8 : R1;
                    constexpr int NBATCH = R1 / PB;
                    if (p.pro.amp_mode == 2) {
                        const long long itemmask = (1ll << (p.g.logN0 + p.g.logN1 + p.g.logM)) - 1;
                        const FieldRows<T> fr(reinterpret_cast<const cplx<T>*>(p.pro.amp) + ((base + c) & itemmask), p.inner,
                                              N, p.pro.amp_sym && !(p.axis == 0 ? p.g.shift0 : p.g.shift1));
                        cplx<T> f[2][PB];
                        NB2_UNROLL
                        for (int i = 0; i < PB; ++i) f[0][i] = fr.load(r + i * MM);
                        NB2_UNROLL
                        for (int g = 0; g < NBATCH; ++g) {
                            if (g + 1 < NBATCH) {
                                NB2_UNROLL
                                for (int i = 0; i < PB; ++i) f[(g + 1) & 1][i] = fr.load(r + ((g + 1) * PB + i) * MM);
                            }
                            NB2_UNROLL
                            for (int i = 0; i < PB; ++i) {
                                x[g * PB + i].re *= f[g & 1][i].re * p.pro.scale;
                                x[g * PB + i].im *= f[g & 1][i].im * p.pro.scale;
                            }
                        }
                    } else if (p.pro.amp_mode == 1) {
                        long long k0, k1, m;
                        decode(p.g, base + c, k0, k1, m);
                        // fixed part of ksq: the non-transformed coordinates
                        const int fx = int(p.axis == 0 ? sq_mirror(k1, p.g.logN1) : sq_mirror(k0, p.g.logN0));
                        const T* te = p.pro.amp + fx + int(sq_mirror(2 * m, p.g.logM + 1));
                        const T* to = p.pro.amp + fx + int(sq_mirror(2 * m + 1, p.g.logM + 1));
                        const int nrow = 1 << (p.axis == 0 ? p.g.logN0 : p.g.logN1);
                        T fe[2][PB], fo[2][PB];
                        NB2_UNROLL
                        for (int i = 0; i < PB; ++i) {
                            const int row = r + i * MM, km = row < nrow - row ? row : nrow - row;
                            fe[0][i] = ld_ro(te + km * km);
                            fo[0][i] = ld_ro(to + km * km);
                        }
                        NB2_UNROLL
                        for (int g = 0; g < NBATCH; ++g) {
                            if (g + 1 < NBATCH) {
                                NB2_UNROLL
                                for (int i = 0; i < PB; ++i) {
                                    const int row = r + ((g + 1) * PB + i) * MM, km = row < nrow - row ? row : nrow - row;
                                    fe[(g + 1) & 1][i] = ld_ro(te + km * km);
                                    fo[(g + 1) & 1][i] = ld_ro(to + km * km);
                                }
                            }
                            NB2_UNROLL
                            for (int i = 0; i < PB; ++i) {
                                x[g * PB + i].re *= fe[g & 1][i] * p.pro.scale;
                                x[g * PB + i].im *= fo[g & 1][i] * p.pro.scale;
                            }
                        }
                    } else {
                        NB2_UNROLL
                        for (int n1 = 0; n1 < R1; ++n1) x[n1] = x[n1] * p.pro.scale;
                    }
                }
                fft_reg<R1, true>(x);
#if NB2_TWPOW
                cplx<T> wp[R1];
                if (!SINGLE) tw_powers<R1>(ldc_ro(p.tw + r), wp);
#endif
                NB2_UNROLL
                for (int k1 = 0; k1 < R1; ++k1) {
                    cplx<T> v = x[brev<R1>(k1)];
#if NB2_TWPOW
                    if (!SINGLE && k1 != 0) v = v * wp[k1];
#else
                    if (!SINGLE && k1 != 0) v = v * ldc_ro(p.tw + r * k1);
#endif
                    if (SINGLE) {
                        store_out(p, p.out + obase + c + (p.out_rowoff ? ld_ro(p.out_rowoff + k1) : (long long)k1 * p.inner), v);
                    } else {
                        smem[(slot_of<R2, R3>(k1 * MM + r) << p.logw) + c] = v;
                    }
                }
            }
        } else if (PH == 1) {
            if (R2 > 1) {
                const int nb = (R1 * R3) << p.logw;
                for (int b = tid; b < nb; b += nthreads) {
                    const int c = b & (w - 1), t = b >> p.logw;
                    const int n3 = t % R3, k1 = t / R3;
                    cplx<T> x[R2];
                    NB2_UNROLL
                    for (int n2 = 0; n2 < R2; ++n2)
                        x[n2] = smem[(slot_of<R2, R3>(k1 * MM + n2 * R3 + n3) << p.logw) + c];
                    fft_reg<R2, true>(x);
                    NB2_UNROLL
                    for (int k2 = 0; k2 < R2; ++k2) {
                        cplx<T> v = x[brev<R2>(k2)];
                        if (k2 != 0) v = v * ldc_ro(p.tw + R1 * n3 * k2);
                        smem[(slot_of<R2, R3>(k1 * MM + k2 * R3 + n3) << p.logw) + c] = v;
                    }
                }
            }
        } else {
            if (R3 > 1) {
                const int nb = (R1 * R2) << p.logw;
                for (int b = tid; b < nb; b += nthreads) {
                    const int c = b & (w - 1), t = b >> p.logw;
                    const int k2 = t % R2, k1 = t / R2;
                    cplx<T> x[R3];
                    NB2_UNROLL
                    for (int n3 = 0; n3 < R3; ++n3)
                        x[n3] = smem[(slot_of<R2, R3>(k1 * MM + k2 * R3 + n3) << p.logw) + c];
                    fft_reg<R3, true>(x);
                    if (p.out_rowoff) {
                        cplx<T>* dst = p.out + obase + c;
                        const long long* ro = p.out_rowoff + (k1 + R1 * k2);
                        NB2_UNROLL
                        for (int k3 = 0; k3 < R3; ++k3) stc_stream(dst + ld_ro(ro + k3 * (R1 * R2)), x[brev<R3>(k3)]);
                    } else {
                        cplx<T>* dst = p.out + obase + c + (long long)(k1 + R1 * k2) * p.inner;
                        NB2_UNROLL
                        for (int k3 = 0; k3 < R3; ++k3)
                            store_out(p, dst + (long long)k3 * (R1 * R2) * p.inner, x[brev<R3>(k3)]);
                    }
                }
            }
        }
    }
};

// ============================================================================ RowPass
// Row enumeration: canonical representatives of the mirror pairs {(k0,k1), (-k0,-k1)}.
struct RowPairs {
    long long nmid, h, Q;  // Q pairs per batch item
};
NB2_HD RowPairs row_pairs(const Geom& g) {
    const long long N0 = 1ll << g.logN0, N1 = 1ll << g.logN1;
    RowPairs rp;
    if (g.shift0) {   // no self-mirrored k0: canonical k0 in [0, N0/2), every k1
        rp.h = 0;
        rp.nmid = (N0 / 2) * N1;
        rp.Q = rp.nmid;
        return rp;
    }
    rp.h = g.shift1 ? N1 / 2 : N1 / 2 + 1;
    rp.nmid = N0 >= 2 ? (N0 / 2 - 1) * N1 : 0;
    rp.Q = rp.nmid + (N0 >= 2 ? 2 : 1) * rp.h;
    return rp;
}

struct RowSlot {
    bool valid, dup;
    long long k0, k1, row;  // row: linear row index over (batch, N0, N1)
};

// EPI >= 0: the epilogue operands are known at compile time (bit 0 streamed amplitude field, bit 1 mul,
// bit 2 add, bit 3 dot with dot == add; no ksq table, no binned reduction), which removes the per-element
// operand tests of the untangle phase -- the phase that executes 57 % of the kernel's instructions.
// EPI < 0: everything decided at run time.
#if defined(__CUDA_ARCH__)
#define NB2_ASSUME(c) __builtin_assume(c)
#else
#define NB2_ASSUME(c) do { if (!(c)) __builtin_unreachable(); } while (0)
#endif
template <typename T, int R1, int R2, int R3, int EPI = -1>
struct RowPass {
    static constexpr int M = R1 * R2 * R3;
    static constexpr int MM = R2 * R3;
    static constexpr int PITCH = slot_pitch<R1, R2, R3>();
    static constexpr int NPHASES = 7;
    static constexpr bool BIG = (R1 >= 32 || R3 >= 32);
    // radix <= 16 plans whose CTAs never have more than 256 threads (2 G R2 R3: rows of up to 512 and of 4096
    // scalars) are compiled for that: the generic epilogue keeps its state in registers (at 512 x 2 it spilled
    // 10 % of its instructions through the LSU pipe that bounds these kernels)
    static constexpr int gmax_of(size_t row_bytes) {     // == choose_G (nb2_fast_impl.cuh) without the row-count cap
        int g = 1;
        while (g < 16 && size_t(4 * g) * row_bytes <= 36 * 1024) g *= 2;
        return g;
    }
    static constexpr int THREADS_NEEDED = 2 * gmax_of(size_t(PITCH) * sizeof(cplx<T>)) * R2 * R3;
    static constexpr bool SHORT = (!BIG && EPI < 0 && THREADS_NEEDED <= 256);
    static constexpr int MAX_THREADS = SHORT ? 256 : ((BIG && sizeof(T) == 8) ? 256 : 512);
    static constexpr int MIN_BLOCKS = SHORT ? (sizeof(T) == 4 ? NB2_MINB_SHORT_F32 : NB2_MINB_SHORT_F64)
                                     : sizeof(T) == 4 ? (BIG ? NB2_MINB256_F32 : NB2_MINB512_F32)
                                                      : (MAX_THREADS == 256 ? NB2_MINB256_F64 : NB2_MINB512_F64);

    struct Params {
        const cplx<T>* in;
        T* out;
        const cplx<T>* tw;  // tw[j] = exp(-2 pi i j / (2M)), j in [0, 2M)
        Geom g;
        long long ctas_per_item;
        int G;     // mirror pairs per CTA (2G rows)
        int conv;  // NB2_NONCANONICAL / NB2_CANONICAL, | NB2_PRODUCT_LAST: Hartley over the last axis
                   // composed with the genuine Hartley over the leading axes (a product of two spaces)
        int has_pro;
        PointOp<T> pro;
        PointOp<T> epi;
    };

    // per-slot data computed once per tile (phase 0) and reused by the untangle phase
    struct SlotCache {
        long long row;  // linear row index over (batch, N0, N1)
        long long frow;    // row (inside one item) whose amplitude-field values this slot uses
        int f01;        // ksq contribution of (k0, k1)
        int flags;      // bit 0: valid, bit 1: duplicate of its partner (self-mirrored row)
    };
    static size_t smem_bytes(int G, int nthreads) {
        return size_t(2 * G) * PITCH * sizeof(cplx<T>) + size_t(nthreads) * sizeof(double) +
               size_t(2 * G) * sizeof(SlotCache);
    }
    NB2_HD static SlotCache* slot_cache(const Params& p, char* smem_raw, int nthreads) {
        return reinterpret_cast<SlotCache*>(smem_raw + size_t(2 * p.G) * PITCH * sizeof(cplx<T>) +
                                            size_t(nthreads) * sizeof(double));
    }
    NB2_HDC static bool enabled(int ph) {
        return ph <= 1 || (ph == 2 && R2 > 1) || (ph == 3 && MM > 1) || ph >= 4;
    }

    NB2_HD static RowSlot slot_info(const Params& p, long long bid, int s) {
        const RowPairs rp = row_pairs(p.g);
        const long long item = p.g.batch == 1 ? 0 : bid / p.ctas_per_item;
        const long long q = (bid - item * p.ctas_per_item) * p.G + (s >> 1);
        const long long N0 = 1ll << p.g.logN0, N1 = 1ll << p.g.logN1;
        RowSlot rs;
        rs.valid = q < rp.Q;
        rs.dup = false;
        rs.k0 = rs.k1 = rs.row = 0;
        if (!rs.valid) return rs;
        long long k0, k1;
        if (q < rp.nmid) {
            k0 = (p.g.shift0 ? 0 : 1) + (q >> p.g.logN1);
            k1 = q & (N1 - 1);
        } else {
            const long long qq = q - rp.nmid;
            k0 = (qq / rp.h) ? N0 / 2 : 0;
            k1 = qq % rp.h;
        }
        const long long k0m = (N0 - k0 - p.g.shift0) & (N0 - 1), k1m = (N1 - k1 - p.g.shift1) & (N1 - 1);
        if (s & 1) {
            rs.dup = (k0m == k0 && k1m == k1);
            k0 = k0m;
            k1 = k1m;
        }
        rs.k0 = k0;
        rs.k1 = k1;
        rs.row = (item * N0 + k0) * N1 + k1;
        return rs;
    }

    // slot of output index k = k1 + R1 (k2 + R2 k3): with L = k1 MM + k2 R3 + k3 and k3 < R3, k2 < R2
    // the padding terms are L / R3 = k1 R2 + k2 and L / (R2 R3) = k1, i.e. a fixed stride per digit
    static constexpr int KS1 = MM + R2 + (R2 > 1 ? 1 : 0);   // slot stride of k1 (consecutive k)
    NB2_HD static int slot_of_k(int k) {
        return (k & (R1 - 1)) * KS1 + ((k / R1) & (R2 - 1)) * (R3 + 1) + k / (R1 * R2);
    }

    // one contribution to the spectrum cotangent: fp64 atomics, or the order-independent fixed-point bins
    template <typename OP>
    NB2_HD static void bin_add(const OP& e, long long bin, double v) {
        if (e.bin_fixed) {
            atomic_add_fixed(e.bin_fixed + 2, bin, e.bin_fixed, v, e.bin_fixed_scale);   // entry 0: the flag
        } else {
            atomic_add_f64(e.bin_out + bin, v);
        }
    }

    template <typename OP>
    NB2_HD static T apply_epi(const OP& e, const Geom& g, T h, long long k0, long long k1, long long k2,
                              long long item, long long idx, double& acc) {
        if (e.bin_x) {
            const long long ksq = sq_mirror(k0, g.logN0) + sq_mirror(k1, g.logN1) + sq_mirror(k2, g.logM + 1);
            bin_add(e, ksq, double(ld_stream(e.bin_x + idx)) * double(h * e.scale));
        }
        T v = h * amp_factor(e, g, k0, k1, k2, item);
        if (e.mul) v *= ld_stream(e.mul + idx);
        v += e.add_scalar;
        if (e.add) v += ld_stream(e.add + idx);
        if (e.dot) acc += double(ld_stream(e.dot + idx)) * double(v);
        return v;
    }

    // epilogue for four consecutive outputs.  `kq`: |k2| of the first element counted from the
    // nearer end of the last axis -- first half: k2 = kq + i  (mirror length kq + i),
    // second half: k2 = M + k + i (mirror length M - k - i = kq - i), selected by `down`.
    // Epilogue for four consecutive outputs, split in two so that every global load is issued
    // before the shared-memory work of the untangle step and only consumed after it.
    struct EpiLoads {
        vec4<T> f, m, a, d, bx;
    };
    // `kq`/`down`: mode length along the last axis of element i is kq + i (first half of the
    // row) or kq - i (second half).
    template <typename OP>
    NB2_HD static void epi_fetch(const OP& e, EpiLoads& L, int f01, int kq, bool down, long long item,
                                 long long idx) {
        if (e.amp_mode == 1) {
            NB2_UNROLL
            for (int i = 0; i < 4; ++i) {
                const int km = down ? kq - i : kq + i;
                L.f.v[i] = ld_ro(e.amp + f01 + km * km);
            }
        } else if (e.amp_mode == 2) {
            L.f = e.amp_sym ? ld4_keep(e.amp + item) : ld4_stream(e.amp + item);
        }
        if (e.mul) L.m = ld4_stream(e.mul + idx);
        if (e.add) L.a = ld4_stream(e.add + idx);
        if (e.dot && e.dot != e.add) L.d = ld4_stream(e.dot + idx);   // the CG curvature reads d once
        if (e.bin_x) L.bx = ld4_stream(e.bin_x + idx);
    }
    template <typename OP>
    NB2_HD static void epi_apply(const OP& e, const EpiLoads& L, vec4<T>& h, double& acc, int f01, int kq, bool down) {
        NB2_UNROLL
        for (int i = 0; i < 4; ++i) {
            T v = h.v[i] * e.scale;
            if (e.bin_x) {
                const int km = down ? kq - i : kq + i;
                bin_add(e, f01 + km * km, double(L.bx.v[i]) * double(v));
            }
            if (e.amp_mode != 0) v *= L.f.v[i];
            if (e.mul) v *= L.m.v[i];
            v += e.add_scalar;
            if (e.add) v += L.a.v[i];
            if (e.dot) acc += double(e.dot != e.add ? L.d.v[i] : L.a.v[i]) * double(v);
            h.v[i] = v;
        }
    }

    template <int PH>
    NB2_HD static void phase(const Params& p, long long bid, int tid, int nthreads, char* smem_raw) {
        cplx<T>* smem = reinterpret_cast<cplx<T>*>(smem_raw);
        double* red = reinterpret_cast<double*>(smem_raw + size_t(2 * p.G) * PITCH * sizeof(cplx<T>));
        const int nslots = 2 * p.G;
        if (PH == 0) {
            // row bookkeeping of the tile, once per slot
            for (int s = tid; s < nslots; s += nthreads) {
                const RowSlot rs = slot_info(p, bid, s);
                SlotCache sc;
                sc.row = rs.row;
                sc.f01 = int(sq_mirror(rs.k0, p.g.logN0) + sq_mirror(rs.k1, p.g.logN1));
                {
                    // NB2_AMP_SYMMETRIC: the field rows (+-k0, +-k1) are equal -- every one of them reads
                    // the canonical row (min(k0, N0 - k0), min(k1, N1 - k1)); its repeats come from L2
                    const long long N0 = 1ll << p.g.logN0, N1 = 1ll << p.g.logN1;
                    long long f0 = rs.k0, f1 = rs.k1;
                    if (p.epi.amp_sym && !p.g.shift0 && !p.g.shift1) {
                        f0 = f0 <= N0 - f0 ? f0 : N0 - f0;
                        f1 = f1 <= N1 - f1 ? f1 : N1 - f1;
                    }
                    sc.frow = (f0 << p.g.logN1) + f1;
                }
                sc.flags = (rs.valid ? 1 : 0) | (rs.dup ? 2 : 0);
                slot_cache(p, smem_raw, nthreads)[s] = sc;
            }
        } else if (PH == 1) {
            const int nb = nslots * MM;
            const SlotCache* cache = slot_cache(p, smem_raw, nthreads);
            for (int b = tid; b < nb; b += nthreads) {
                const int r = b % MM, s = b / MM;
                const SlotCache sc = cache[s];
                if (!(sc.flags & 1)) continue;
                cplx<T> x[R1];
                {
                    const cplx<T>* src = p.in + sc.row * M + r;
                    NB2_UNROLL
                    for (int n1 = 0; n1 < R1; ++n1) x[n1] = ldc_stream(src + n1 * MM);
                }
                if (p.has_pro) {
                    const long long N0N1 = 1ll << (p.g.logN0 + p.g.logN1);
                    const long long rowitem = sc.row & (N0N1 - 1);
                    const long long f01 = sc.f01;
                    NB2_UNROLL
                    for (int n1 = 0; n1 < R1; ++n1) {
                        const long long m = r + n1 * MM;
                        T ae, ao;
                        if (p.pro.amp_mode == 1) {
                            ae = ld_ro(p.pro.amp + f01 + sq_mirror(2 * m, p.g.logM + 1)) * p.pro.scale;
                            ao = ld_ro(p.pro.amp + f01 + sq_mirror(2 * m + 1, p.g.logM + 1)) * p.pro.scale;
                        } else if (p.pro.amp_mode == 2) {
                            const cplx<T> f =
                                ldc_stream(reinterpret_cast<const cplx<T>*>(p.pro.amp) + rowitem * M + m);
                            ae = f.re * p.pro.scale;
                            ao = f.im * p.pro.scale;
                        } else {
                            ae = ao = p.pro.scale;
                        }
                        x[n1].re *= ae;
                        x[n1].im *= ao;
                    }
                }
                fft_reg<R1, true>(x);
                cplx<T>* dst = smem + s * PITCH;
#if NB2_TWPOW
                cplx<T> wp[R1];
                if (MM > 1) tw_powers<R1>(ldc_ro(p.tw + 2 * r), wp);
#endif
                NB2_UNROLL
                for (int k1 = 0; k1 < R1; ++k1) {
                    cplx<T> v = x[brev<R1>(k1)];
                    // W_M^(r k1) = W_2M^(2 r k1)
#if NB2_TWPOW
                    if (MM > 1 && k1 != 0) v = v * wp[k1];
#else
                    if (MM > 1 && k1 != 0) v = v * ldc_ro(p.tw + 2 * r * k1);
#endif
                    dst[slot_of<R2, R3>(k1 * MM + r)] = v;
                }
            }
        } else if (PH == 2) {
            if (R2 > 1) {
                const int nb = nslots * R1 * R3;
                for (int b = tid; b < nb; b += nthreads) {
                    const int n3 = b % R3, k1 = (b / R3) % R1, s = b / (R3 * R1);
                    cplx<T>* row = smem + s * PITCH;
                    cplx<T> x[R2];
                    NB2_UNROLL
                    for (int n2 = 0; n2 < R2; ++n2) x[n2] = row[slot_of<R2, R3>(k1 * MM + n2 * R3 + n3)];
                    fft_reg<R2, true>(x);
                    NB2_UNROLL
                    for (int k2 = 0; k2 < R2; ++k2) {
                        cplx<T> v = x[brev<R2>(k2)];
                        // W_MM^(n3 k2) = W_2M^(2 R1 n3 k2)
                        if (k2 != 0) v = v * ldc_ro(p.tw + 2 * R1 * n3 * k2);
                        row[slot_of<R2, R3>(k1 * MM + k2 * R3 + n3)] = v;
                    }
                }
            }
        } else if (PH == 3) {
            if (MM > 1) {
                const int nb = nslots * R1 * R2;
                for (int b = tid; b < nb; b += nthreads) {
                    const int k2 = b % R2, k1 = (b / R2) % R1, s = b / (R2 * R1);
                    cplx<T>* row = smem + s * PITCH;
                    cplx<T> x[R3];
                    NB2_UNROLL
                    for (int n3 = 0; n3 < R3; ++n3) x[n3] = row[slot_of<R2, R3>(k1 * MM + k2 * R3 + n3)];
                    fft_reg<R3, true>(x);
                    NB2_UNROLL
                    for (int k3 = 0; k3 < R3; ++k3) row[slot_of<R2, R3>(k1 * MM + k2 * R3 + k3)] = x[brev<R3>(k3)];
                }
            }
        } else if (PH == 4) {
            // untangle + Hartley combine + epilogue
            double acc = 0.0;
            const long long itemrows = 1ll << (p.g.logN0 + p.g.logN1);
            const T sgn = (p.conv & NB2_CANONICAL) ? T(-1) : T(1);
            // Product of a leading-axes space with the last axis: with F the joint spectrum of the real
            // field, H_last H_lead x [k, k2] = Re F[k, -k2] + sgn Im F[k, k2].  F[k, -k2] comes from the
            // same two rows: index (M - j) of this row and index j of the mirror row.
            const bool prod = EPI < 0 && (p.conv & NB2_PRODUCT_LAST);
            int prev_s = -1;
            RowSlot rs{};
            if (M >= 4) {
                // four consecutive k per thread: 16/32-byte global accesses
                constexpr int Q4 = M >= 4 ? M / 4 : 1;
                const int nb = nslots * Q4;
                const SlotCache* cache = slot_cache(p, smem_raw, nthreads);
                PointOp<T> e = p.epi;
                e.scale = p.epi.scale * T(0.5);          // the 1/2 of the untangle step rides on the scale
                if (EPI >= 0) {
                    e.amp_mode = (EPI & 1) ? 2 : 0;
                    if (EPI & 2) NB2_ASSUME(e.mul != nullptr); else e.mul = nullptr;
                    if (EPI & 4) NB2_ASSUME(e.add != nullptr); else e.add = nullptr;
                    e.dot = (EPI & 8) ? e.add : nullptr;
                    e.bin_x = nullptr;
                }
                for (int b = tid; b < nb; b += nthreads) {
                    const int k = (b % Q4) * 4, s = b / Q4;
                    const SlotCache sc = cache[s];
                    if (sc.flags != 1) continue;          // invalid, or the duplicate of a self-mirrored row
                    const long long rowbase = sc.row * (2 * M);
                    const long long itembase = sc.frow * (2 * M);   // == the slot's own row unless amp_sym
                    const int f01 = sc.f01;
                    EpiLoads L1, L2;
                    epi_fetch(e, L1, f01, k, false, itembase + k, rowbase + k);
                    epi_fetch(e, L2, f01, M - k, true, itembase + M + k, rowbase + M + k);
                    // a_i = Z[k + i] (this row), b_i = conj Z'[(M - k - i) mod M] (mirror row): consecutive
                    // k are KS1 slots apart; i = 1..3 of the mirror index lie in the quad below M - k
                    const cplx<T>* za = smem + s * PITCH + slot_of_k(k);
                    const cplx<T>* zb = smem + (s ^ 1) * PITCH;
                    const cplx<T>* zb0 = zb + slot_of_k((M - k) & (M - 1));
                    const cplx<T>* zb4 = zb + slot_of_k(M - k - 4);
                    const cplx<T>* zr = smem + s * PITCH;
                    const cplx<T>* za0p = zr + slot_of_k((M - k) & (M - 1));
                    const cplx<T>* za4p = zr + slot_of_k(M - k - 4);
                    const cplx<T>* zbk = zb + slot_of_k(k);
                    const vec4<T> w01 = ld4_ro(reinterpret_cast<const T*>(p.tw + k));
                    const vec4<T> w23 = ld4_ro(reinterpret_cast<const T*>(p.tw + k + 2));
                    vec4<T> h1, h2;
                    NB2_UNROLL
                    for (int i = 0; i < 4; ++i) {
                        const cplx<T> a = za[i * KS1];
                        const cplx<T> bz = i == 0 ? zb0[0] : zb4[(4 - i) * KS1];
                        const T wr = i < 2 ? w01.v[2 * i] : w23.v[2 * i - 4];
                        const T wi = i < 2 ? w01.v[2 * i + 1] : w23.v[2 * i - 3];
                        // E = a + conj b, O = -i (a - conj b); 2 H[k] = cas(E) + cas(W O), 2 H[k+M] = cas(E) - cas(W O)
                        // with cas(z) = Re z + sgn Im z
                        const T Er = a.re + bz.re, Ei = a.im - bz.im;
                        const T Or = a.im + bz.im, Oi = bz.re - a.re;
                        T ce = Er + sgn * Ei;
                        T co = Or * (wr + sgn * wi) + Oi * (sgn * wr - wi);
                        if (prod) {
                            // a' = Z[(M - j) mod M] of this row, b' = Z'[j] of the mirror row; their twiddle
                            // is W^(M - j) = -conj W^j (also right for j = 0, where the halves swap)
                            const cplx<T> ap = i == 0 ? za0p[0] : za4p[(4 - i) * KS1];
                            const cplx<T> bp = zbk[i * KS1];
                            const T Epr = ap.re + bp.re, Opr = ap.im + bp.im, Opi = bp.re - ap.re;
                            ce = Epr + sgn * Ei;
                            co = wr * (Opr + sgn * Oi) + wi * (Opi + sgn * Or);
                        }
                        h1.v[i] = ce + co;
                        h2.v[i] = ce - co;
                    }
                    epi_apply(e, L1, h1, acc, f01, k, false);
                    epi_apply(e, L2, h2, acc, f01, M - k, true);
                    st4_stream(p.out + rowbase + k, h1);
                    st4_stream(p.out + rowbase + M + k, h2);
                }
            } else {
                const int nb = nslots * M;
                for (int b = tid; b < nb; b += nthreads) {
                    const int k = b % M, s = b / M;
                    if (s != prev_s) {
                        rs = slot_info(p, bid, s);
                        prev_s = s;
                    }
                    if (!rs.valid || rs.dup) continue;
                    const cplx<T> a = smem[s * PITCH + slot_of_k(k)];
                    const cplx<T> bb = conj(smem[(s ^ 1) * PITCH + slot_of_k((M - k) & (M - 1))]);
                    const cplx<T> E = a + bb;              // 2 * even-sample spectrum
                    const cplx<T> O = mul_mi(a - bb);      // 2 * odd-sample spectrum
                    const cplx<T> WO = ldc_ro(p.tw + k) * O;
                    const cplx<T> F1 = E + WO, F2 = E - WO;  // 2 F[k], 2 F[k + M]
                    T h1 = T(0.5) * (F1.re + sgn * F1.im), h2 = T(0.5) * (F2.re + sgn * F2.im);
                    if (prod) {
                        const cplx<T> ap = smem[s * PITCH + slot_of_k((M - k) & (M - 1))];
                        const cplx<T> bp = conj(smem[(s ^ 1) * PITCH + slot_of_k(k)]);
                        const cplx<T> Ep = ap + bp, Op = mul_mi(ap - bp);
                        const cplx<T> w = ldc_ro(p.tw + k);
                        const cplx<T> wp = {-w.re, w.im};   // W^(M - k) = -conj W^k
                        const cplx<T> WOp = wp * Op;
                        // 2 F[M - k] = Ep + WOp and 2 F[2M - k] = Ep - WOp
                        h1 = T(0.5) * ((Ep.re - WOp.re) + sgn * F1.im);
                        h2 = T(0.5) * ((Ep.re + WOp.re) + sgn * F2.im);
                    }
                    const long long idx = rs.row * (2 * M) + k;
                    const long long item = (rs.row & (itemrows - 1)) * (2 * M) + k;
                    const T o1 = apply_epi(p.epi, p.g, h1, rs.k0, rs.k1, k, item, idx, acc);
                    const T o2 = apply_epi(p.epi, p.g, h2, rs.k0, rs.k1, k + M, item + M, idx + M, acc);
                    st_stream(p.out + idx, o1);
                    st_stream(p.out + idx + M, o2);
                }
            }
            if (p.epi.dot) red[tid] = acc;
        } else if (PH == 5) {
            if (p.epi.dot && tid < 32) {
                double s = 0.0;
                for (int j = tid; j < nthreads; j += 32) s += red[j];
#if defined(__CUDA_ARCH__)
                NB2_UNROLL
                for (int off = 16; off > 0; off >>= 1) s += __shfl_down_sync(0xffffffffu, s, off);
                if (tid == 0) p.epi.dot_partials[bid] = s;
#else
                red[tid] = s;
#endif
            }
        } else {
#if !defined(__CUDA_ARCH__)
            if (p.epi.dot && tid == 0) {
                // same pairing tree as the shuffle reduction above
                double v[32];
                for (int j = 0; j < 32; ++j) v[j] = j < nthreads ? red[j] : 0.0;
                for (int off = 16; off > 0; off >>= 1)
                    for (int j = 0; j < off; ++j) v[j] += v[j + off];
                p.epi.dot_partials[bid] = v[0];
            }
#endif
        }
    }
};

// ============================================================================ MirrorCombine
// Transforms over LEADING axes with a contiguous inner batch (product spaces: the spatial
// sub-space of a space x frequency field, src/library/correlated_fields.py:790-793).  Two
// adjacent batch entries a, b are packed as one complex number z = x_a + i x_b (a free
// reinterpretation of the real array); after the complex FFT over the transformed axes
//   F_a[k] = (Z[k] + conj Z[-k]) / 2,   F_b[k] = -i (Z[k] - conj Z[-k]) / 2,
// and H = Re F +- Im F.  This kernel does that combine for a row k and its mirror row -k at
// once (each element read once, written once; in place is fine) and applies the epilogue.
template <typename T>
struct MirrorCombine {
    static constexpr int NPHASES = 3;
    static constexpr int MAX_THREADS = 256;
    static constexpr int MIN_BLOCKS = 1;
    struct Params {
        const cplx<T>* in;
        T* out;
        Geom g;
        int conv;
        PointOp<T> epi;
        long long nctas;
    };
    static size_t smem_bytes(int nthreads) { return size_t(nthreads) * sizeof(double); }
    NB2_HDC static bool enabled(int) { return true; }

    NB2_HD static T epi1(const PointOp<T>& e, T h, long long item, long long idx, double& acc) {
        T v = h * e.scale;
        if (e.amp_mode == 2) v *= ld_stream(e.amp + item);
        if (e.mul) v *= ld_stream(e.mul + idx);
        v += e.add_scalar;
        if (e.add) v += ld_stream(e.add + idx);
        if (e.dot) acc += double(ld_stream(e.dot + idx)) * double(v);
        return v;
    }

    template <int PH>
    NB2_HD static void phase(const Params& p, long long bid, int tid, int nthreads, char* smem_raw) {
        double* red = reinterpret_cast<double*>(smem_raw);
        if (PH == 0) {
            double acc = 0.0;
            const RowPairs rp = row_pairs(p.g);
            const long long N0 = 1ll << p.g.logN0, N1 = 1ll << p.g.logN1, M = 1ll << p.g.logM;
            const long long total = p.g.batch * rp.Q * M, stride = p.nctas * nthreads;
            const T sgn = (p.conv & NB2_CANONICAL) ? T(-1) : T(1);
            for (long long i = bid * nthreads + tid; i < total; i += stride) {
                const long long m = i & (M - 1), pr = i >> p.g.logM;
                const long long item = pr / rp.Q, q = pr - item * rp.Q;
                long long k0, k1;
                if (q < rp.nmid) {
                    k0 = (p.g.shift0 ? 0 : 1) + (q >> p.g.logN1);
                    k1 = q & (N1 - 1);
                } else {
                    const long long qq = q - rp.nmid;
                    k0 = (qq / rp.h) ? N0 / 2 : 0;
                    k1 = qq % rp.h;
                }
                const long long k0m = (N0 - k0 - p.g.shift0) & (N0 - 1), k1m = (N1 - k1 - p.g.shift1) & (N1 - 1);
                const long long ra = (k0 << p.g.logN1) + k1, rb = (k0m << p.g.logN1) + k1m;   // rows inside the item
                const long long rows_item = N0 * N1;
                const long long ea = (item * rows_item + ra) * M + m, eb = (item * rows_item + rb) * M + m;
                const cplx<T> za = ldc_stream(p.in + ea), zb = ldc_stream(p.in + eb);
                // row k:   F_a = (za + conj zb)/2, F_b = -i (za - conj zb)/2
                const T ha = T(0.5) * ((za.re + zb.re) + sgn * (za.im - zb.im));
                const T hb = T(0.5) * ((za.im + zb.im) - sgn * (za.re - zb.re));
                cplx<T> oa;
                oa.re = epi1(p.epi, ha, (ra * M + m) * 2, ea * 2, acc);
                oa.im = epi1(p.epi, hb, (ra * M + m) * 2 + 1, ea * 2 + 1, acc);
                cplx<T> ob;
                if (ra != rb) {  // row -k: the roles of za and zb exchanged
                    const T ga = T(0.5) * ((zb.re + za.re) + sgn * (zb.im - za.im));
                    const T gb = T(0.5) * ((zb.im + za.im) - sgn * (zb.re - za.re));
                    ob.re = epi1(p.epi, ga, (rb * M + m) * 2, eb * 2, acc);
                    ob.im = epi1(p.epi, gb, (rb * M + m) * 2 + 1, eb * 2 + 1, acc);
                }
                stc_stream(reinterpret_cast<cplx<T>*>(p.out) + ea, oa);
                if (ra != rb) stc_stream(reinterpret_cast<cplx<T>*>(p.out) + eb, ob);
            }
            red[tid] = acc;
        } else if (PH == 1) {
            if (p.epi.dot && tid < 32) {
                double s = 0.0;
                for (int j = tid; j < nthreads; j += 32) s += red[j];
#if defined(__CUDA_ARCH__)
                NB2_UNROLL
                for (int off = 16; off > 0; off >>= 1) s += __shfl_down_sync(0xffffffffu, s, off);
                if (tid == 0) p.epi.dot_partials[bid] = s;
#else
                red[tid] = s;
#endif
            }
        } else {
#if !defined(__CUDA_ARCH__)
            if (p.epi.dot && tid == 0) {
                double v[32];
                for (int j = 0; j < 32; ++j) v[j] = j < nthreads ? red[j] : 0.0;
                for (int off = 16; off > 0; off >>= 1)
                    for (int j = 0; j < off; ++j) v[j] += v[j + off];
                p.epi.dot_partials[bid] = v[0];
            }
#endif
        }
    }
};

}  // namespace nb2
