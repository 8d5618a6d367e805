// niftyb200 -- complex-to-complex FFT along the contiguous axis (ducc_dispatch.fftn / ifftn,
// src/ducc_dispatch.py:66-72; FFTOperator, src/operators/harmonic_operators.py:106-123).
//
// The strided axes of a c2c transform are served by the strided passes of nb2_passes.cuh / nb2_tma.cuh
// as they are (those are complex FFTs over rows of complex columns); this kernel is the missing piece:
// rows of L = R1*R2*R3 complex numbers, G rows per CTA, the same register-radix steps through padded
// shared memory as RowPass, without the pairing / untangle of the real-data transform.
//
// Backward transforms use  ifft(x) = conj(fft(conj(x))):  `conj_in` conjugates on load, `conj_out` on store
// (for a multi-axis backward transform the caller conjugates once before the first and once after the last
// pass); `scale` rides on the store.
#pragma once

#include "nb2_common.cuh"
#include "nb2_passes.cuh"
#include "nb2_radix.cuh"

namespace nb2 {

template <typename T, int R1, int R2, int R3>
struct RowC2C {
    static constexpr int L = R1 * R2 * R3;
    static constexpr int MM = R2 * R3;
    static constexpr int PITCH = slot_pitch<R1, R2, R3>();
    static constexpr int NPHASES = 4;
    static constexpr bool BIG = (R1 >= 32 || R3 >= 32);
    static constexpr int MAX_THREADS = (BIG && sizeof(T) == 8) ? 256 : 512;
    static constexpr int MIN_BLOCKS = 1;
    static constexpr int KS1 = MM + R2 + (R2 > 1 ? 1 : 0);   // slot stride of consecutive output indices

    struct Params {
        const cplx<T>* in;
        cplx<T>* out;
        const cplx<T>* tw;   // tw[j] = exp(-2 pi i j / L), j in [0, L)
        long long nrows;
        int G;               // rows per CTA
        int conj_in, conj_out;
        T scale;
    };
    static size_t smem_bytes(int G) { return size_t(G) * PITCH * sizeof(cplx<T>); }
    NB2_HDC static bool enabled(int ph) { return ph == 0 || (ph == 1 && R2 > 1) || (ph == 2 && MM > 1) || ph == 3; }
    NB2_HD static int slot_of_k(int k) {
        return (k & (R1 - 1)) * KS1 + ((k / R1) & (R2 - 1)) * (R3 + 1) + k / (R1 * R2);
    }

    template <int PH>
    NB2_HD static void phase(const Params& p, long long bid, int tid, int nthreads, char* smem_raw) {
        cplx<T>* smem = reinterpret_cast<cplx<T>*>(smem_raw);
        const long long row0 = bid * p.G;
        const int nrows = int(p.nrows - row0 < p.G ? p.nrows - row0 : p.G);
        if (PH == 0) {
            const int nb = nrows * MM;
            for (int b = tid; b < nb; b += nthreads) {
                const int r = b % MM, s = b / MM;
                const cplx<T>* src = p.in + (row0 + s) * L + r;
                cplx<T> x[R1];
                NB2_UNROLL
                for (int n1 = 0; n1 < R1; ++n1) x[n1] = ldc_stream(src + n1 * MM);
                if (p.conj_in) {
                    NB2_UNROLL
                    for (int n1 = 0; n1 < R1; ++n1) x[n1].im = -x[n1].im;
                }
                fft_reg<R1, true>(x);
                cplx<T>* dst = smem + s * PITCH;
                cplx<T> wp[R1];
                if (MM > 1) tw_powers<R1>(ldc_ro(p.tw + r), wp);          // W_L^(r k1)
                NB2_UNROLL
                for (int k1 = 0; k1 < R1; ++k1) {
                    cplx<T> v = x[brev<R1>(k1)];
                    if (MM > 1 && k1 != 0) v = v * wp[k1];
                    dst[slot_of<R2, R3>(k1 * MM + r)] = v;
                }
            }
        } else if (PH == 1) {
            if (R2 > 1) {
                const int nb = nrows * R1 * R3;
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
                        if (k2 != 0) v = v * ldc_ro(p.tw + R1 * n3 * k2);   // W_MM^(n3 k2) = W_L^(R1 n3 k2)
                        row[slot_of<R2, R3>(k1 * MM + k2 * R3 + n3)] = v;
                    }
                }
            }
        } else if (PH == 2) {
            if (MM > 1) {
                const int nb = nrows * R1 * R2;
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
        } else {
            const int nb = nrows * L;
            const T sc = p.scale, si = p.conj_out ? -p.scale : p.scale;
            for (int b = tid; b < nb; b += nthreads) {
                const int k = b % L, s = b / L;
                // single-step plans (MM == 1) leave X[k] at slot_of(k) = k of the first phase
                const cplx<T> v = smem[s * PITCH + (MM > 1 ? slot_of_k(k) : slot_of<R2, R3>(k))];
                stc_stream(p.out + (row0 + s) * L + k, mk<T>(v.re * sc, v.im * si));
            }
        }
    }
};

// out = conj(in) * scale over n complex numbers (the conjugations around a multi-axis backward transform)
template <typename T>
struct CplxConjScale {
    static constexpr int NPHASES = 1;
    static constexpr int MAX_THREADS = 256;
    static constexpr int MIN_BLOCKS = 1;
    struct Params {
        const cplx<T>* in;
        cplx<T>* out;
        T scale;
        int conj;
        long long n, nctas;
    };
    static size_t smem_bytes(int) { return 0; }
    NB2_HDC static bool enabled(int) { return true; }
    template <int PH>
    NB2_HD static void phase(const Params& p, long long bid, int tid, int nthreads, char*) {
        const long long stride = p.nctas * nthreads;
        const T si = p.conj ? -p.scale : p.scale;
        for (long long e = bid * nthreads + tid; e < p.n; e += stride) {
            const cplx<T> v = ldc_stream(p.in + e);
            stc_stream(p.out + e, mk<T>(v.re * p.scale, v.im * si));
        }
    }
};

}  // namespace nb2
