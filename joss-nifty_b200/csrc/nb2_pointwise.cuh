// niftyb200 -- pointwise prologue / epilogue operators fused into the FFT passes.
//
// The correlated-field hot tail (reference: src/library/correlated_fields.py:801-818,
// src/re/correlated_field.py:690-693) multiplies the harmonic-space excitation by the
// amplitude gathered through the mode-length index (PowerDistributor,
// src/operators/distributors.py:114-119) before the Hartley transform, and scales by the
// harmonic volume factor / adds the offset after it.  Here these ride on the first load and
// the last store of the transform.
//
// The reference's int64 `pindex` array (src/domains/power_space.py:170-198) is replaced, for
// isotropic grids, by a table indexed with the integer squared mode length
//   ksq = sum_j min(k_j, N_j - k_j)^2        (src/domains/rg_space.py:105-116,130-141)
// (amp_mode 1); anything else uses a precomputed full-size factor field (amp_mode 2).
#pragma once

#include "nb2_common.cuh"

namespace nb2 {

template <typename T>
struct PointOp {
    int amp_mode;        // 0: none, 1: table[ksq], 2: full-size field (one batch item, broadcast over batch)
    int amp_sym;         // amp_mode 2 only: field[k] == field[-k] along the transformed axes (NB2_AMP_SYMMETRIC)
    const T* amp;        // table or field
    T scale;             // multiplies the value (volume factor, azm, ...)
    const T* mul;        // optional element-wise multiplier (full array incl. batch), epilogue only
    const T* add;        // optional element-wise addend, epilogue only
    T add_scalar;        // epilogue only (offset_mean)
    const T* dot;        // optional: accumulate sum(dot[i] * out[i]) in fp64, epilogue only
    double* dot_partials;  // one slot per CTA (deterministic two-level reduction)
    // optional spectrum cotangent (epilogue of the row pass only): bin_out[ksq(i)] += bin_x[i] * value
    // with value = transform * scale BEFORE the amplitude factor -- the binned reduction of
    // PowerDistributor.adjoint_times (src/operators/distributors.py:105-112, utilities.py:223-240)
    // in the ksq parametrisation of the bins; fp64 atomics
    const T* bin_x;
    double* bin_out;
    // bit-reproducible variant: two-limb fixed-point integers, integer atomics (see niftyb200.h)
    long long* bin_fixed;
    double bin_fixed_scale;
    // prologue of a strided pass only: in' = pre_add + pre_scale * in, stored to pre_out if set
    const T* pre_add;
    T pre_scale;
    T* pre_out;
    const double* pre_scale_dev;   // device-resident CG: beta read on the device when not NULL
};

template <typename T> NB2_HD T pre_scale_of(const PointOp<T>& op) {
    return op.pre_scale_dev ? T(*op.pre_scale_dev) : op.pre_scale;
}

NB2_HD void atomic_add_f64(double* p, double v) {
#if defined(__CUDA_ARCH__)
    atomicAdd(p, v);
#else
    *p += v;
#endif
}

// bins[2b] += hi, bins[2b + 1] += lo with v * S = hi + lo * 2^-40: integer additions commute exactly, so the
// result does not depend on the order of the atomics.  `flag`: set when v * S is not representable.
NB2_HD void atomic_add_fixed(long long* bins, long long b, long long* flag, double v, double S) {
    const double q = v * S;                       // exact: S is a power of two
    if (!(q > -4.6e18 && q < 4.6e18)) {           // also catches NaN
#if defined(__CUDA_ARCH__)
        atomicExch(reinterpret_cast<unsigned long long*>(flag), 1ull);
#else
        *flag = 1;
#endif
        return;
    }
    const double hi = rint(q);
    const long long ihi = (long long)hi;
    const long long ilo = (long long)rint((q - hi) * 1099511627776.0);   // 2^40; q - hi is exact
#if defined(__CUDA_ARCH__)
    atomicAdd(reinterpret_cast<unsigned long long*>(bins + 2 * b), (unsigned long long)ihi);
    if (ilo) atomicAdd(reinterpret_cast<unsigned long long*>(bins + 2 * b + 1), (unsigned long long)ilo);
#else
    bins[2 * b] += ihi;
    bins[2 * b + 1] += ilo;
#endif
}

// Geometry of the (batch, N0, N1, N2) real array; N2 = 2*M, all powers of two.
struct Geom {
    int logN0, logN1, logM;   // log2 of N0, N1 and M = N2/2
    long long batch;
    // Mirror rule of the row pass: the partner of row (k0, k1) is ((N0 - k0 - shift0) mod N0,
    // (N1 - k1 - shift1) mod N1).  shift = 0: ordinary index negation (a full axis);
    // shift = 1: pure reflection without fixed points -- a mirror-symmetric LOCAL part of a
    // slab-decomposed axis (niftyb200/distributed.py), whose two halves are each other's mirrors.
    int shift0, shift1;
};

NB2_HD long long sq_mirror(long long k, int logN) {
    const long long N = 1ll << logN;
    const long long km = k < N - k ? k : N - k;
    return km * km;
}

// complex linear index -> (k0, k1, m)
NB2_HD void decode(const Geom& g, long long lin, long long& k0, long long& k1, long long& m) {
    m = lin & ((1ll << g.logM) - 1);
    lin >>= g.logM;
    k1 = lin & ((1ll << g.logN1) - 1);
    lin >>= g.logN1;
    k0 = lin & ((1ll << g.logN0) - 1);
}

// amplitude factor for the real element with coordinates (k0, k1, k2); `item` is the real
// linear index inside one batch item (used by amp_mode 2).
template <typename T>
NB2_HD T amp_factor(const PointOp<T>& op, const Geom& g, long long k0, long long k1, long long k2,
                    long long item) {
    if (op.amp_mode == 1) {
        const long long ksq = sq_mirror(k0, g.logN0) + sq_mirror(k1, g.logN1) + sq_mirror(k2, g.logM + 1);
        return ld_ro(op.amp + ksq) * op.scale;
    }
    if (op.amp_mode == 2) return ld_ro(op.amp + item) * op.scale;
    return op.scale;
}

}  // namespace nb2
