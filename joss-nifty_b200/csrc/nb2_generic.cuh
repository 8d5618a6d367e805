// niftyb200 -- generic-size path (correctness-first): any length, any subset of axes.
//
// Used for everything the power-of-two fast path does not cover (odd lengths such as the
// reference tests' (10,), (12,9), (15,27); transforms over non-trailing axes; complex c2c
// for ducc_dispatch.fftn/ifftn, src/ducc_dispatch.py:66-72).  One O(n) direct DFT sum per
// output element and axis, accumulated in fp64 with an fp64 root table.  Not
// roofline-graded.
#pragma once

#include <cmath>

#include "nb2_common.cuh"

namespace nb2 {

// out[o, k, i] = sum_j in[o, j, i] * root[(j k) mod n]^(+-1)
template <typename T>
struct NaiveDft {
    static constexpr int NPHASES = 1;
    static constexpr int MAX_THREADS = 256;
    static constexpr int MIN_BLOCKS = 1;
    struct Params {
        const cplx<T>* in;
        cplx<T>* out;
        const cplx<double>* root;  // root[j] = exp(-2 pi i j / n)
        long long outer, n, inner;
        int forward;
        long long nctas;
    };
    static size_t smem_bytes(int) { return 0; }
    NB2_HDC static bool enabled(int) { return true; }
    template <int PH>
    NB2_HD static void phase(const Params& p, long long bid, int tid, int nthreads, char*) {
        const long long total = p.outer * p.n * p.inner, stride = p.nctas * nthreads;
        for (long long e = bid * nthreads + tid; e < total; e += stride) {
            const long long i = e % p.inner, k = (e / p.inner) % p.n, o = e / (p.inner * p.n);
            const cplx<T>* src = p.in + o * p.n * p.inner + i;
            double sr = 0, si = 0;
            long long ph = 0;
            for (long long j = 0; j < p.n; ++j) {
                const cplx<T> v = src[j * p.inner];
                cplx<double> w = p.root[ph];
                if (!p.forward) w.im = -w.im;
                sr += double(v.re) * w.re - double(v.im) * w.im;
                si += double(v.re) * w.im + double(v.im) * w.re;
                ph += k;
                if (ph >= p.n) ph -= p.n;
            }
            p.out[e] = mk<T>(T(sr), T(si));
        }
    }
};

// real -> complex (imag = 0), with optional scale
template <typename T>
struct RealToCplx {
    static constexpr int NPHASES = 1;
    static constexpr int MAX_THREADS = 256;
    static constexpr int MIN_BLOCKS = 1;
    struct Params {
        const T* in;
        cplx<T>* out;
        long long n, nctas;
    };
    static size_t smem_bytes(int) { return 0; }
    NB2_HDC static bool enabled(int) { return true; }
    template <int PH>
    NB2_HD static void phase(const Params& p, long long bid, int tid, int nthreads, char*) {
        const long long stride = p.nctas * nthreads;
        for (long long e = bid * nthreads + tid; e < p.n; e += stride) p.out[e] = mk<T>(p.in[e], T(0));
    }
};

// complex -> Hartley combine:  out = (re +/- im) * scale     (src/ducc_dispatch.py:46-50)
template <typename T>
struct CplxToHartley {
    static constexpr int NPHASES = 1;
    static constexpr int MAX_THREADS = 256;
    static constexpr int MIN_BLOCKS = 1;
    struct Params {
        const cplx<T>* in;
        T* out;
        T scale;
        int conv;
        long long n, nctas;
    };
    static size_t smem_bytes(int) { return 0; }
    NB2_HDC static bool enabled(int) { return true; }
    template <int PH>
    NB2_HD static void phase(const Params& p, long long bid, int tid, int nthreads, char*) {
        const long long stride = p.nctas * nthreads;
        for (long long e = bid * nthreads + tid; e < p.n; e += stride) {
            const cplx<T> v = p.in[e];
            p.out[e] = (p.conv == NB2_CANONICAL ? v.re - v.im : v.re + v.im) * p.scale;
        }
    }
};

// complex scale / copy: out = in * scale
template <typename T>
struct CplxScale {
    static constexpr int NPHASES = 1;
    static constexpr int MAX_THREADS = 256;
    static constexpr int MIN_BLOCKS = 1;
    struct Params {
        const cplx<T>* in;
        cplx<T>* out;
        T scale;
        long long n, nctas;
    };
    static size_t smem_bytes(int) { return 0; }
    NB2_HDC static bool enabled(int) { return true; }
    template <int PH>
    NB2_HD static void phase(const Params& p, long long bid, int tid, int nthreads, char*) {
        const long long stride = p.nctas * nthreads;
        for (long long e = bid * nthreads + tid; e < p.n; e += stride) p.out[e] = p.in[e] * p.scale;
    }
};

// ---------------------------------------------------------------- Bluestein (chirp-z) for one axis of any length n
// X[k] = w[k] * sum_j (x[j] w[j]) conj(w[k - j]),  w[j] = exp(-+ i pi j^2 / n): a circular convolution of length
// L = 2^ceil(log2(2n - 1)) done with the power-of-two contiguous-axis kernel (RowC2C): O(n log n) per line instead
// of the O(n^2) direct sum.  Lines are gathered into a padded contiguous buffer and scattered back.

// buf[line, j] = in[o, j, i] * w[j] (j < n), 0 (n <= j < L); line = o * inner + i
template <typename T>
struct BlueGather {
    static constexpr int NPHASES = 1;
    static constexpr int MAX_THREADS = 256;
    static constexpr int MIN_BLOCKS = 1;
    struct Params {
        const cplx<T>* in;
        cplx<T>* buf;
        const cplx<T>* w;      // forward chirp exp(-i pi j^2 / n)
        long long outer, n, inner, L;
        int forward;
        long long nctas;
    };
    static size_t smem_bytes(int) { return 0; }
    NB2_HDC static bool enabled(int) { return true; }
    template <int PH>
    NB2_HD static void phase(const Params& p, long long bid, int tid, int nthreads, char*) {
        const long long total = p.outer * p.inner * p.L, stride = p.nctas * nthreads;
        for (long long e = bid * nthreads + tid; e < total; e += stride) {
            const long long j = e % p.L, line = e / p.L, i = line % p.inner, o = line / p.inner;
            cplx<T> v = mk<T>(T(0), T(0));
            if (j < p.n) {
                v = p.in[(o * p.n + j) * p.inner + i];
                if (p.w) {                         // nullptr: power-of-two axis, plain gather (L == n)
                    cplx<T> w = p.w[j];
                    if (!p.forward) w.im = -w.im;
                    v = v * w;
                }
            }
            p.buf[e] = v;
        }
    }
};

// The same gather through 32 x 32 shared-memory tiles for axes with inner > 1: reads run along `i` (contiguous in
// the source), writes along `j` (contiguous in the line buffer).  One CTA per (o, j-tile, i-tile), 256 threads.
template <typename T>
struct BlueGatherTiled {
    static constexpr int NPHASES = 2;
    static constexpr int MAX_THREADS = 256;
    static constexpr int MIN_BLOCKS = 1;
    typedef typename BlueGather<T>::Params Params;
    static size_t smem_bytes(int) { return 32 * 33 * sizeof(cplx<T>); }
    NB2_HDC static bool enabled(int) { return true; }
    static long long grid(const Params& p) { return p.outer * ((p.L + 31) / 32) * ((p.inner + 31) / 32); }
    template <int PH>
    NB2_HD static void phase(const Params& p, long long bid, int tid, int nthreads, char* smem_raw) {
        cplx<T>* tile = reinterpret_cast<cplx<T>*>(smem_raw);
        const long long ti = (p.inner + 31) / 32, tj = (p.L + 31) / 32;
        const long long it = bid % ti, jt = (bid / ti) % tj, o = bid / (ti * tj);
        const int tx = tid & 31;
        for (int r = tid >> 5; r < 32; r += nthreads >> 5) {
            if (PH == 0) {
                const long long j = jt * 32 + r, i = it * 32 + tx;
                cplx<T> v = mk<T>(T(0), T(0));
                if (j < p.n && i < p.inner) {
                    v = p.in[(o * p.n + j) * p.inner + i];
                    if (p.w) {
                        cplx<T> w = p.w[j];
                        if (!p.forward) w.im = -w.im;
                        v = v * w;
                    }
                }
                tile[r * 33 + tx] = v;
            } else {
                const long long i = it * 32 + r, j = jt * 32 + tx;
                if (i < p.inner && j < p.L) p.buf[(o * p.inner + i) * p.L + j] = tile[tx * 33 + r];
            }
        }
    }
};

// buf[line, m] = conj(buf[line, m] * B[m])   (B of the other direction is conj(B): the filter is symmetric);
// the conjugate turns the following FORWARD transform into the inverse one
template <typename T>
struct BlueMulConj {
    static constexpr int NPHASES = 1;
    static constexpr int MAX_THREADS = 256;
    static constexpr int MIN_BLOCKS = 1;
    struct Params {
        cplx<T>* buf;
        const cplx<T>* B;
        long long nlines, L;
        int forward;
        long long nctas;
    };
    static size_t smem_bytes(int) { return 0; }
    NB2_HDC static bool enabled(int) { return true; }
    template <int PH>
    NB2_HD static void phase(const Params& p, long long bid, int tid, int nthreads, char*) {
        const long long total = p.nlines * p.L, stride = p.nctas * nthreads;
        for (long long e = bid * nthreads + tid; e < total; e += stride) {
            cplx<T> b = p.B[e % p.L];
            if (!p.forward) b.im = -b.im;
            const cplx<T> v = p.buf[e] * b;
            p.buf[e] = mk<T>(v.re, -v.im);
        }
    }
};

// out[o, k, i] = w[k] * conj(buf[line, k]) / L
template <typename T>
struct BlueScatter {
    static constexpr int NPHASES = 1;
    static constexpr int MAX_THREADS = 256;
    static constexpr int MIN_BLOCKS = 1;
    struct Params {
        const cplx<T>* buf;
        cplx<T>* out;
        const cplx<T>* w;
        long long outer, n, inner, L;
        int forward;
        long long nctas;
    };
    static size_t smem_bytes(int) { return 0; }
    NB2_HDC static bool enabled(int) { return true; }
    template <int PH>
    NB2_HD static void phase(const Params& p, long long bid, int tid, int nthreads, char*) {
        const long long total = p.outer * p.n * p.inner, stride = p.nctas * nthreads;
        const T s = T(1) / T(p.L);
        for (long long e = bid * nthreads + tid; e < total; e += stride) {
            const long long i = e % p.inner, k = (e / p.inner) % p.n, o = e / (p.inner * p.n);
            const cplx<T> y = p.buf[(o * p.inner + i) * p.L + k];
            if (!p.w) {                            // power-of-two axis: the row kernel's result as it is
                p.out[e] = y;
                continue;
            }
            cplx<T> w = p.w[k];
            if (!p.forward) w.im = -w.im;
            p.out[e] = (w * mk<T>(y.re, -y.im)) * s;
        }
    }
};

NB2_HD void sincospi_hd(double x, double* sn, double* cs) {
#if defined(__CUDA_ARCH__)
    sincospi(x, sn, cs);
#else
    *sn = std::sin(3.14159265358979323846 * x);
    *cs = std::cos(3.14159265358979323846 * x);
#endif
}

// Per line: (R x C) row-major matrix -> (C x R), optionally times the four-step twiddle W_n^(r c), n = R C
// (sign by direction) -- the transpositions of the four-step transform of lines longer than the row kernel takes.
// 32 x 32 shared-memory tiles, one CTA per (line, row-tile, column-tile), 256 threads.
template <typename T>
struct LineTranspose {
    static constexpr int NPHASES = 2;
    static constexpr int MAX_THREADS = 256;
    static constexpr int MIN_BLOCKS = 1;
    struct Params {
        const cplx<T>* in;
        cplx<T>* out;
        long long nlines, R, C;
        int twiddle;     // 0: plain; 1: multiply by exp(-+ 2 pi i r c / (R C))
        int forward;
    };
    static size_t smem_bytes(int) { return 32 * 33 * sizeof(cplx<T>); }
    NB2_HDC static bool enabled(int) { return true; }
    static long long grid(const Params& p) { return p.nlines * ((p.R + 31) / 32) * ((p.C + 31) / 32); }
    template <int PH>
    NB2_HD static void phase(const Params& p, long long bid, int tid, int nthreads, char* smem_raw) {
        cplx<T>* tile = reinterpret_cast<cplx<T>*>(smem_raw);
        const long long tc = (p.C + 31) / 32, tr = (p.R + 31) / 32;
        const long long ct = bid % tc, rt = (bid / tc) % tr, line = bid / (tc * tr);
        const long long base = line * p.R * p.C;
        const int tx = tid & 31;
        for (int y = tid >> 5; y < 32; y += nthreads >> 5) {
            if (PH == 0) {
                const long long r = rt * 32 + y, c = ct * 32 + tx;
                if (r < p.R && c < p.C) {
                    cplx<T> v = p.in[base + r * p.C + c];
                    if (p.twiddle) {
                        const long long n = p.R * p.C, q = (r * c) % n;
                        double sn, cs;
                        sincospi_hd(-2.0 * double(q) / double(n), &sn, &cs);
                        if (!p.forward) sn = -sn;
                        v = v * mk<T>(T(cs), T(sn));
                    }
                    tile[y * 33 + tx] = v;
                }
            } else {
                const long long c = ct * 32 + y, r = rt * 32 + tx;
                if (r < p.R && c < p.C) p.out[base + c * p.R + r] = tile[tx * 33 + y];
            }
        }
    }
};

// tiled twin of BlueScatter: reads along `k` (contiguous in the line buffer), writes along `i`
template <typename T>
struct BlueScatterTiled {
    static constexpr int NPHASES = 2;
    static constexpr int MAX_THREADS = 256;
    static constexpr int MIN_BLOCKS = 1;
    typedef typename BlueScatter<T>::Params Params;
    static size_t smem_bytes(int) { return 32 * 33 * sizeof(cplx<T>); }
    NB2_HDC static bool enabled(int) { return true; }
    static long long grid(const Params& p) { return p.outer * ((p.n + 31) / 32) * ((p.inner + 31) / 32); }
    template <int PH>
    NB2_HD static void phase(const Params& p, long long bid, int tid, int nthreads, char* smem_raw) {
        cplx<T>* tile = reinterpret_cast<cplx<T>*>(smem_raw);
        const long long ti = (p.inner + 31) / 32, tk = (p.n + 31) / 32;
        const long long it = bid % ti, kt = (bid / ti) % tk, o = bid / (ti * tk);
        const int tx = tid & 31;
        const T s = T(1) / T(p.L);
        for (int r = tid >> 5; r < 32; r += nthreads >> 5) {
            if (PH == 0) {
                const long long i = it * 32 + r, k = kt * 32 + tx;
                if (i < p.inner && k < p.n) tile[r * 33 + tx] = p.buf[(o * p.inner + i) * p.L + k];
            } else {
                const long long k = kt * 32 + r, i = it * 32 + tx;
                if (k < p.n && i < p.inner) {
                    const cplx<T> y = tile[tx * 33 + r];
                    cplx<T> res = y;
                    if (p.w) {
                        cplx<T> w = p.w[k];
                        if (!p.forward) w.im = -w.im;
                        res = (w * mk<T>(y.re, -y.im)) * s;
                    }
                    p.out[(o * p.n + k) * p.inner + i] = res;
                }
            }
        }
    }
};

}  // namespace nb2
