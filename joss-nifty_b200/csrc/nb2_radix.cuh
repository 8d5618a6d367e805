// niftyb200 -- in-register power-of-two FFT butterflies (radix 2 .. 32).
//
// fft_reg<R, FWD>(x): decimation-in-frequency radix-2 network on R complex values
// held in registers, every loop fully unrolled so all twiddles are compile-time
// constants (trivial ones are folded away by the special cases in twmul32).  The
// result is left in bit-reversed order: X[k] == x[brev<R>(k)].
#pragma once

#include "nb2_common.cuh"

namespace nb2 {

// cos(2*pi*j/32), j = 0..8 (fp64-rounded literals); everything else by symmetry.
template <typename T> NB2_HD T cos32_q(int j) {
    switch (j) {
        case 0: return T(1.0);
        case 1: return T(0.98078528040323044912618223613424);
        case 2: return T(0.92387953251128675612818318939679);
        case 3: return T(0.83146961230254523707878837761791);
        case 4: return T(0.70710678118654752440084436210485);
        case 5: return T(0.55557023301960222474283081394853);
        case 6: return T(0.38268343236508977172845998403040);
        case 7: return T(0.19509032201612826784828486847702);
        default: return T(0.0);
    }
}
// cos/sin(2*pi*j/32) for j in [0, 16]
template <typename T> NB2_HD T cos32(int j) { return j <= 8 ? cos32_q<T>(j) : -cos32_q<T>(16 - j); }
template <typename T> NB2_HD T sin32(int j) { return j <= 8 ? cos32_q<T>(8 - j) : cos32_q<T>(j - 8); }

// d * exp(-/+ 2*pi*i*q/32), q in [0,16)  (FWD: minus sign)
template <bool FWD, typename T> NB2_HD cplx<T> twmul32(cplx<T> d, int q) {
    if (q == 0) return d;
    if (q == 8) return FWD ? mul_mi(d) : mul_pi(d);
    if (q == 4) {
        const T c = cos32_q<T>(4);
        return FWD ? mk<T>((d.re + d.im) * c, (d.im - d.re) * c) : mk<T>((d.re - d.im) * c, (d.im + d.re) * c);
    }
    if (q == 12) {
        const T c = cos32_q<T>(4);
        return FWD ? mk<T>((d.im - d.re) * c, -(d.re + d.im) * c) : mk<T>(-(d.re + d.im) * c, (d.re - d.im) * c);
    }
    const T c = cos32<T>(q), s = sin32<T>(q);
    return FWD ? mk<T>(d.re * c + d.im * s, d.im * c - d.re * s) : mk<T>(d.re * c - d.im * s, d.im * c + d.re * s);
}

template <int R> NB2_HD constexpr int brev(int k) {
    int r = 0;
    for (int b = 1; b < R; b <<= 1) { r = (r << 1) | (k & 1); k >>= 1; }
    return r;
}

template <int R> struct log2c { static constexpr int value = 1 + log2c<R / 2>::value; };
template <> struct log2c<1> { static constexpr int value = 0; };

template <int R, bool FWD, typename T> NB2_HD void fft_reg(cplx<T> (&x)[R]) {
    static_assert(R >= 1 && R <= 32 && (R & (R - 1)) == 0, "radix must be a power of two <= 32");
    NB2_UNROLL
    for (int s = 0; s < log2c<R>::value; ++s) {
        const int len = R >> s, half = len >> 1, tstep = 32 / len;
        NB2_UNROLL
        for (int blk = 0; blk < R; blk += len) {
            NB2_UNROLL
            for (int j = 0; j < half; ++j) {
                const cplx<T> a = x[blk + j], b = x[blk + j + half];
                x[blk + j] = a + b;
                x[blk + j + half] = twmul32<FWD>(a - b, j * tstep);
            }
        }
    }
}

}  // namespace nb2
