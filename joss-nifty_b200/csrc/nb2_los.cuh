// niftyb200 -- line-of-sight response of the tomography configuration.
//
// Reference: SamplingCartesianGridLOS / _los (src/re/extra/sampling_los.py:15-91): every line is sampled at
// n points  pp_a = start_i + (a + 1/2) (end_i - start_i) / n  in index coordinates (loc * (shape-1)/shape / dist),
// the field is read there by multilinear interpolation (jax.scipy.ndimage.map_coordinates, order 1, constant
// mode with cval = NaN: BOTH corners of every axis are visited, a corner outside the grid contributes NaN even
// with weight 0), and the sum is scaled by |end - start| / n.  Forward: gather (one CTA per line, fp64 block
// reduction).  Adjoint: the transpose scatter-add into the field (fp64 / fp32 atomics; corners outside the grid
// are skipped -- the adjoint of a constant contribution is nothing).
#pragma once

#include "nb2_common.cuh"
#include "nb2_vector.cuh"

namespace nb2 {

constexpr int kLosMaxDim = 3;

struct LosGeom {
    int ndim;
    long long shape[kLosMaxDim];
    double l2i[kLosMaxDim];       // ((shape - 1) / shape) / distances
    long long n_los;
    int n_sampling;
    const double* start;          // (n_los, ndim) or (ndim) when start_stride == 0
    const double* end;
    long long start_stride, end_stride;
};

// sample point `a` of line `l`: lower corner indices, upper weights; returns false if the line is degenerate
NB2_HD void los_point(const LosGeom& g, long long l, int a, long long (&lo)[kLosMaxDim], double (&wu)[kLosMaxDim]) {
    for (int i = 0; i < g.ndim; ++i) {
        const double s = g.start[l * g.start_stride + i] * g.l2i[i], e = g.end[l * g.end_stride + i] * g.l2i[i];
        const double c = s + ((e - s) / g.n_sampling) * (a + 0.5);
        const double f = floor(c);
        lo[i] = (long long)f;
        wu[i] = c - f;
    }
}
NB2_HD double los_length(const LosGeom& g, long long l) {
    double d2 = 0.0;
    for (int i = 0; i < g.ndim; ++i) {
        const double d = g.end[l * g.end_stride + i] - g.start[l * g.start_stride + i];
        d2 += d * d;
    }
    return sqrt(d2);
}

template <typename T>
struct LosForward {
    static constexpr int NPHASES = 3;
    static constexpr int MAX_THREADS = 256;
    static constexpr int MIN_BLOCKS = 1;
    struct Params {
        LosGeom g;
        const T* x;
        T* out;        // (n_los)
    };
    static size_t smem_bytes(int nthreads) { return size_t(nthreads) * sizeof(double); }
    NB2_HDC static bool enabled(int) { return true; }
    template <int PH>
    NB2_HD static void phase(const Params& p, long long bid, int tid, int nthreads, char* smem_raw) {
        double* red = reinterpret_cast<double*>(smem_raw);
        const LosGeom& g = p.g;
        if (PH == 0) {
            double acc = 0.0;
            for (int a = tid; a < g.n_sampling; a += nthreads) {
                long long lo[kLosMaxDim];
                double wu[kLosMaxDim];
                los_point(g, bid, a, lo, wu);
                double v = 0.0;
                for (int corner = 0; corner < (1 << g.ndim); ++corner) {
                    double w = 1.0;
                    long long lin = 0;
                    bool inside = true;
                    for (int i = 0; i < g.ndim; ++i) {
                        const int up = (corner >> i) & 1;
                        const long long idx = lo[i] + up;
                        w *= up ? wu[i] : 1.0 - wu[i];
                        inside = inside && idx >= 0 && idx < g.shape[i];
                        lin = lin * g.shape[i] + (inside ? idx : 0);
                    }
                    v += w * (inside ? double(p.x[lin]) : NAN);
                }
                acc += v;
            }
            red[tid] = acc;
        } else if (PH == 1) {
            double s = 0.0;
            block_reduce_B<1>(red, tid, nthreads, &s);
#if defined(__CUDA_ARCH__)
            if (tid == 0) p.out[bid] = T(s * (los_length(g, bid) / g.n_sampling));
#endif
        } else {
#if !defined(__CUDA_ARCH__)
            double s = 0.0;
            block_reduce_C<1>(red, tid, nthreads, &s);
            if (tid == 0) p.out[bid] = T(s * (los_length(g, bid) / g.n_sampling));
#endif
        }
    }
};

template <typename T> NB2_HD void atomic_add_t(T* p, T v) {
#if defined(__CUDA_ARCH__)
    atomicAdd(p, v);
#else
    *p += v;
#endif
}

// xbar += R^T ybar (the caller zeroes xbar)
template <typename T>
struct LosAdjoint {
    static constexpr int NPHASES = 1;
    static constexpr int MAX_THREADS = 256;
    static constexpr int MIN_BLOCKS = 1;
    struct Params {
        LosGeom g;
        const T* ybar;   // (n_los)
        T* xbar;         // field
    };
    static size_t smem_bytes(int) { return 0; }
    NB2_HDC static bool enabled(int) { return true; }
    template <int PH>
    NB2_HD static void phase(const Params& p, long long bid, int tid, int nthreads, char*) {
        const LosGeom& g = p.g;
        const double yl = double(p.ybar[bid]) * (los_length(g, bid) / g.n_sampling);
        for (int a = tid; a < g.n_sampling; a += nthreads) {
            long long lo[kLosMaxDim];
            double wu[kLosMaxDim];
            los_point(g, bid, a, lo, wu);
            for (int corner = 0; corner < (1 << g.ndim); ++corner) {
                double w = 1.0;
                long long lin = 0;
                bool inside = true;
                for (int i = 0; i < g.ndim; ++i) {
                    const int up = (corner >> i) & 1;
                    const long long idx = lo[i] + up;
                    w *= up ? wu[i] : 1.0 - wu[i];
                    inside = inside && idx >= 0 && idx < g.shape[i];
                    lin = lin * g.shape[i] + (inside ? idx : 0);
                }
                if (inside) atomic_add_t(p.xbar + lin, T(w * yl));
            }
        }
    }
};

}  // namespace nb2
