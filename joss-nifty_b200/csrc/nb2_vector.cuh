// niftyb200 -- streaming vector kernels: fused conjugate-gradient updates with
// deterministic fp64 dot-product reductions, and the element-wise helpers of the
// generic (non power-of-two) path.
//
// Reference recurrences: ConjugateGradient.__call__ (src/minimization/conjugate_gradient.py:
// 84-124), QuadraticEnergy (src/minimization/quadratic_energy.py:35-37), re `_cg`
// (src/re/conjugate_gradient.py:146-205).  Per iteration the reference sweeps the vectors
// ~10 times (every Field op allocates); here one kernel does  x -= a d, r -= a q  and the
// three sums  r.r, x.r, x.b  from which both stacks' energies follow:
//   numpy stack  E = 0.5 x.(r + b) - b.x      re stack  E = ((r - j)/2).x
//
// Reductions: per-thread fp64 accumulators -> warp shuffle -> one partial per CTA ->
// a single-CTA final pass.  No floating-point atomics, so results are bit-reproducible
// run to run (the property the reference's pairwise-tree allreduce provides,
// src/utilities.py:331-390).
#pragma once

#include "../../include/niftyb200.h"
#include "nb2_common.cuh"
#include "nb2_pointwise.cuh"

namespace nb2 {

NB2_HD bool aligned16(const void* p) { return (reinterpret_cast<uintptr_t>(p) & 15) == 0; }
template <typename... P> NB2_HD bool all_aligned16(P... ps) {
    bool ok = true;
    const void* a[] = {static_cast<const void*>(ps)...};
    for (const void* q : a) ok = ok && (q == nullptr || aligned16(q));
    return ok;
}

// Block reduction of NV fp64 values per thread.  `red` holds NV * nthreads doubles.
// phase A (end of the accumulating phase): red[v * nthreads + tid] = acc[v]
// phase B: warp 0 folds, phase C: (emulator only) final tree.
template <int NV>
NB2_HD void block_reduce_B(double* red, int tid, int nthreads, double* out /* NV values */) {
    if (tid < 32) {
        for (int v = 0; v < NV; ++v) {
            double s = 0.0;
            for (int j = tid; j < nthreads; j += 32) s += red[v * nthreads + j];
#if defined(__CUDA_ARCH__)
            NB2_UNROLL
            for (int off = 16; off > 0; off >>= 1) s += __shfl_down_sync(0xffffffffu, s, off);
            if (tid == 0) out[v] = s;
#else
            red[v * nthreads + tid] = s;
#endif
        }
    }
}
template <int NV>
NB2_HD void block_reduce_C(double* red, int tid, int nthreads, double* out) {
#if !defined(__CUDA_ARCH__)
    if (tid == 0) {
        for (int v = 0; v < NV; ++v) {
            double t[32];
            for (int j = 0; j < 32; ++j) t[j] = j < nthreads ? red[v * nthreads + j] : 0.0;
            for (int off = 16; off > 0; off >>= 1)
                for (int j = 0; j < off; ++j) t[j] += t[j + off];
            out[v] = t[0];
        }
    }
#endif
}

// ---------------------------------------------------------------- fused CG update
// mode 0:  x -= alpha d ; r -= alpha q ; sums          (conjugate_gradient.py:100-101)
// mode 1:  r  = q - b ; sums   (exact residual after a reset step, q = A x; :103-105)
// sums[0] = r.r, sums[1] = x.r, sums[2] = x.b      (after the update)
template <typename T>
struct CgUpdate {
    static constexpr int NPHASES = 3;
    static constexpr int MAX_THREADS = 256;
    static constexpr int MIN_BLOCKS = 1;
    struct Params {
        T* x;
        T* r;
        const T* d;
        const T* q;
        const T* b;
        T alpha;
        int mode;
        long long n;
        double* partials;  // 3 per CTA
        long long nctas;
        const double* alpha_dev;   // device-resident CG: alpha read on the device when not NULL
    };
    static size_t smem_bytes(int nthreads) { return size_t(3) * nthreads * sizeof(double); }
    NB2_HDC static bool enabled(int) { return true; }

    template <int PH>
    NB2_HD static void phase(const Params& p, long long bid, int tid, int nthreads, char* smem_raw) {
        double* red = reinterpret_cast<double*>(smem_raw);
        if (PH == 0) {
            double s0 = 0, s1 = 0, s2 = 0;
            const T alpha = p.alpha_dev ? T(*p.alpha_dev) : p.alpha;
            const long long stride = p.nctas * nthreads;
            const long long n4 = all_aligned16(p.x, p.r, p.d, p.q, p.b) ? p.n / 4 : 0;
            for (long long v = bid * nthreads + tid; v < n4; v += stride) {
                const long long i = 4 * v;
                vec4<T> xv, rv;
                if (p.mode == 0) {
                    xv = ld4_stream(p.x + i);
                    rv = ld4_stream(p.r + i);
                    const vec4<T> dv = ld4_stream(p.d + i), qv = ld4_stream(p.q + i);
                    NB2_UNROLL
                    for (int e = 0; e < 4; ++e) {
                        xv.v[e] -= alpha * dv.v[e];
                        rv.v[e] -= alpha * qv.v[e];
                    }
                    st4_stream(p.x + i, xv);
                } else {
                    xv = ld4_stream(p.x + i);
                    rv = ld4_stream(p.q + i);
                }
                const vec4<T> bv = ld4_stream(p.b + i);
                if (p.mode != 0) {
                    NB2_UNROLL
                    for (int e = 0; e < 4; ++e) rv.v[e] -= bv.v[e];
                }
                st4_stream(p.r + i, rv);
                NB2_UNROLL
                for (int e = 0; e < 4; ++e) {
                    s0 += double(rv.v[e]) * double(rv.v[e]);
                    s1 += double(xv.v[e]) * double(rv.v[e]);
                    s2 += double(xv.v[e]) * double(bv.v[e]);
                }
            }
            for (long long i = 4 * n4 + bid * nthreads + tid; i < p.n; i += stride) {
                T xv, rv;
                if (p.mode == 0) {
                    xv = ld_stream(p.x + i) - alpha * ld_stream(p.d + i);
                    rv = ld_stream(p.r + i) - alpha * ld_stream(p.q + i);
                    st_stream(p.x + i, xv);
                } else {
                    xv = ld_stream(p.x + i);
                    rv = ld_stream(p.q + i) - ld_stream(p.b + i);
                }
                st_stream(p.r + i, rv);
                const T bv = ld_stream(p.b + i);
                s0 += double(rv) * double(rv);
                s1 += double(xv) * double(rv);
                s2 += double(xv) * double(bv);
            }
            red[tid] = s0;
            red[nthreads + tid] = s1;
            red[2 * nthreads + tid] = s2;
        } else if (PH == 1) {
            block_reduce_B<3>(red, tid, nthreads, p.partials + 3 * bid);
        } else {
            block_reduce_C<3>(red, tid, nthreads, p.partials + 3 * bid);
        }
    }
};

// ---------------------------------------------------------------- y = a*x + b*y  (+ optional dot(y_new, z))
// covers  d = beta d + r  (conjugate_gradient.py:124),  x -= alpha d  and plain copies/scales.
template <typename T>
struct Axpby {
    static constexpr int NPHASES = 1;
    static constexpr int MAX_THREADS = 256;
    static constexpr int MIN_BLOCKS = 1;
    struct Params {
        const T* x;
        T* y;
        T a, b;
        long long n;
        long long nctas;
        const double* a_dev;   // optional device factors (device-resident CG): a *= *a_dev, b *= *b_dev
        const double* b_dev;
    };
    static size_t smem_bytes(int) { return 0; }
    NB2_HDC static bool enabled(int) { return true; }
    template <int PH>
    NB2_HD static void phase(const Params& q, long long bid, int tid, int nthreads, char*) {
        Params p = q;
        if (q.a_dev) p.a = T(double(q.a) * *q.a_dev);
        if (q.b_dev) p.b = T(double(q.b) * *q.b_dev);
        const long long stride = p.nctas * nthreads;
        const long long n4 = all_aligned16(p.x, p.y) ? p.n / 4 : 0;
        for (long long v = bid * nthreads + tid; v < n4; v += stride) {
            vec4<T> xv = ld4_stream(p.x + 4 * v);
            if (p.b != T(0)) {
                const vec4<T> yv = ld4_stream(p.y + 4 * v);
                NB2_UNROLL
                for (int e = 0; e < 4; ++e) xv.v[e] = p.a * xv.v[e] + p.b * yv.v[e];
            } else {
                NB2_UNROLL
                for (int e = 0; e < 4; ++e) xv.v[e] = p.a * xv.v[e];
            }
            st4_stream(p.y + 4 * v, xv);
        }
        for (long long i = 4 * n4 + bid * nthreads + tid; i < p.n; i += stride) {
            const T xv = ld_stream(p.x + i);
            const T yv = p.b == T(0) ? T(0) : p.b * ld_stream(p.y + i);
            st_stream(p.y + i, p.a * xv + yv);
        }
    }
};

// ---------------------------------------------------------------- dot(a, b) partials (a, b real; complex data are passed as 2n reals)
template <typename T>
struct Dot {
    static constexpr int NPHASES = 3;
    static constexpr int MAX_THREADS = 256;
    static constexpr int MIN_BLOCKS = 1;
    struct Params {
        const T* a;
        const T* b;
        long long n;
        double* partials;  // 1 per CTA
        long long nctas;
    };
    static size_t smem_bytes(int nthreads) { return size_t(nthreads) * sizeof(double); }
    NB2_HDC static bool enabled(int) { return true; }
    template <int PH>
    NB2_HD static void phase(const Params& p, long long bid, int tid, int nthreads, char* smem_raw) {
        double* red = reinterpret_cast<double*>(smem_raw);
        if (PH == 0) {
            double s = 0;
            const long long stride = p.nctas * nthreads;
            const long long n4 = all_aligned16(p.a, p.b) ? p.n / 4 : 0;
            for (long long v = bid * nthreads + tid; v < n4; v += stride) {
                const vec4<T> av = ld4_stream(p.a + 4 * v), bv = ld4_stream(p.b + 4 * v);
                NB2_UNROLL
                for (int e = 0; e < 4; ++e) s += double(av.v[e]) * double(bv.v[e]);
            }
            for (long long i = 4 * n4 + bid * nthreads + tid; i < p.n; i += stride)
                s += double(ld_stream(p.a + i)) * double(ld_stream(p.b + i));
            red[tid] = s;
        } else if (PH == 1) {
            block_reduce_B<1>(red, tid, nthreads, p.partials + bid);
        } else {
            block_reduce_C<1>(red, tid, nthreads, p.partials + bid);
        }
    }
};

// ---------------------------------------------------------------- final reduction of per-CTA partials (single CTA)
// ---------------------------------------------------------------- device-resident CG: the scalar logic of one step
// cg_single_step of the reference (src/re/conjugate_gradient.py:266-313), evaluated by one thread on the
// scalars in `state` (layout: NB2_CG_* of include/niftyb200.h).  Comparisons involving NaN are false, as
// with jnp.where.
NB2_HD void cg_decide_step(double* s, int* gate, int stage) {
    double info = s[NB2_CG_INFO];
    if (stage == 0) {
        const double curv = s[NB2_CG_CURV];
        if (curv == 0.0) info = -1;                                   // :272-273
        const double alpha = s[NB2_CG_GAMMA] / curv;                  // :274
        if (alpha < 0.0) info = -1;                                   // :275-276
        s[NB2_CG_ALPHA] = alpha;
        s[NB2_CG_INFO] = info;
        return;
    }
    const double i = s[NB2_CG_ITER] + 1.0;                            // :269
    const double previous_gamma = s[NB2_CG_GAMMA], previous_energy = s[NB2_CG_ENERGY];
    const double gamma = s[NB2_CG_SUMS], xr = s[NB2_CG_SUMS + 1], xb = s[NB2_CG_SUMS + 2];
    const double miniter = s[NB2_CG_MINITER], maxiter = s[NB2_CG_MAXITER];
    if (gamma >= 0.0 && gamma <= s[NB2_CG_TINY] && info != -1) info = 0;            // :292-294
    if (s[NB2_CG_RESNORM] >= 0.0) {                                                  // :295-299 (norm_ord 2)
        if (sqrt(gamma) < s[NB2_CG_RESNORM] && i >= miniter && info != -1) info = 0;
    }
    const double energy = 0.5 * (xr - xb);                                           // :303
    const double energy_diff = previous_energy - energy;
    if (energy_diff < -s[NB2_CG_EPS] * fabs(energy)) info = -1;                      // :305-307
    const double absdelta = s[NB2_CG_ABSDELTA];
    if (absdelta == absdelta) {                                                      // :308-312
        if (energy_diff < absdelta && i >= miniter && info != -1) info = 0;
    }
    if (i >= maxiter && info != -1) info = i;                                        // :313
    const double ratio = gamma / previous_gamma;
    s[NB2_CG_BETA] = ratio > 0.0 ? ratio : (ratio != ratio ? ratio : 0.0);           // jnp.maximum(0, .) :315
    s[NB2_CG_GAMMA] = gamma;
    s[NB2_CG_ENERGY] = energy;
    s[NB2_CG_ITER] = i;
    s[NB2_CG_INFO] = info;
    *gate = info >= -1.0 ? 1 : 0;                                                    // continue_condition :263-264
}

struct CgDecide {
    static constexpr int NPHASES = 1;
    static constexpr int MAX_THREADS = 32;
    static constexpr int MIN_BLOCKS = 1;
    struct Params {
        double* state;
        int* gate;
        int stage;
    };
    static size_t smem_bytes(int) { return 0; }
    NB2_HDC static bool enabled(int) { return true; }
    template <int PH>
    NB2_HD static void phase(const Params& p, long long bid, int tid, int, char*) {
        if (bid != 0 || tid != 0) return;
        cg_decide_step(p.state, p.gate, p.stage);
    }
};

// out[v] = sum_c partials[c * nv + v], fixed order.
struct FinalReduce {
    static constexpr int NPHASES = 3;
    static constexpr int MAX_THREADS = 256;
    static constexpr int MIN_BLOCKS = 1;
    struct Params {
        const double* partials;
        long long nparts;
        int nv;  // <= 4
        double* out;
        // device-resident CG: the scalar step that consumes these sums runs in the same launch (stage 0 after the
        // curvature, stage 1 after the sums of CgUpdate); nullptr = off
        double* cg_state;
        int* cg_gate;
        int cg_stage;
    };
    static size_t smem_bytes(int nthreads) { return size_t(4) * nthreads * sizeof(double); }
    NB2_HDC static bool enabled(int) { return true; }
    template <int PH>
    NB2_HD static void phase(const Params& p, long long, int tid, int nthreads, char* smem_raw) {
        double* red = reinterpret_cast<double*>(smem_raw);
        if (PH == 0) {
            for (int v = 0; v < 4; ++v) {
                double s = 0;
                if (v < p.nv)
                    for (long long c = tid; c < p.nparts; c += nthreads) s += p.partials[c * p.nv + v];
                red[v * nthreads + tid] = s;
            }
        } else if (PH == 1) {
            double tmp[4];
            block_reduce_B<4>(red, tid, nthreads, tmp);
#if defined(__CUDA_ARCH__)
            if (tid == 0) {
                for (int v = 0; v < p.nv; ++v) p.out[v] = tmp[v];
                if (p.cg_state) cg_decide_step(p.cg_state, p.cg_gate, p.cg_stage);
            }
#endif
        } else {
#if !defined(__CUDA_ARCH__)
            double tmp[4];
            block_reduce_C<4>(red, tid, nthreads, tmp);
            if (tid == 0) {
                for (int v = 0; v < p.nv; ++v) p.out[v] = tmp[v];
                if (p.cg_state) cg_decide_step(p.cg_state, p.cg_gate, p.cg_stage);
            }
#endif
        }
    }
};

// ---------------------------------------------------------------- generic element-wise op (generic path pro/epilogue)
// out[i] = in[i] * amp * [mul[i]] + add_scalar + [add[i]] ; optional dot partials.
// Only amp_mode 0 and 2 are meaningful here (no coordinates).
template <typename T>
struct Eltwise {
    static constexpr int NPHASES = 3;
    static constexpr int MAX_THREADS = 256;
    static constexpr int MIN_BLOCKS = 1;
    struct Params {
        const T* in;
        T* out;
        PointOp<T> op;
        long long n, item;  // item: elements of one batch item (amp field size)
        long long nctas;
    };
    static size_t smem_bytes(int nthreads) { return size_t(nthreads) * sizeof(double); }
    NB2_HDC static bool enabled(int) { return true; }
    template <int PH>
    NB2_HD static void phase(const Params& p, long long bid, int tid, int nthreads, char* smem_raw) {
        double* red = reinterpret_cast<double*>(smem_raw);
        if (PH == 0) {
            double s = 0;
            const long long stride = p.nctas * nthreads;
            for (long long i = bid * nthreads + tid; i < p.n; i += stride) {
                T v = p.in[i] * p.op.scale;
                if (p.op.amp_mode == 2) v *= p.op.amp[i % p.item];
                if (p.op.mul) v *= p.op.mul[i];
                v += p.op.add_scalar;
                if (p.op.add) v += p.op.add[i];
                if (p.op.dot) s += double(p.op.dot[i]) * double(v);
                p.out[i] = v;
            }
            red[tid] = s;
        } else if (PH == 1) {
            if (p.op.dot) block_reduce_B<1>(red, tid, nthreads, p.op.dot_partials + bid);
        } else {
            if (p.op.dot) block_reduce_C<1>(red, tid, nthreads, p.op.dot_partials + bid);
        }
    }
};

// ---------------------------------------------------------------- amplitude table -> full-size factor field
// out[k0,k1,k2] = table[ksq(k0,k1,k2)] * scale  -- the PowerDistributor gather
// (src/operators/distributors.py:114-119) evaluated once per expansion point, so that the
// transform passes can stream the factor (amp_mode 2) instead of gathering it per element.
template <typename T>
struct ExpandTable {
    static constexpr int NPHASES = 1;
    static constexpr int MAX_THREADS = 256;
    static constexpr int MIN_BLOCKS = 1;
    struct Params {
        const T* table;
        T* out;
        T scale;
        Geom g;
        long long n;  // elements of one item (N0*N1*N2)
        long long nctas;
    };
    static size_t smem_bytes(int) { return 0; }
    NB2_HDC static bool enabled(int) { return true; }
    template <int PH>
    NB2_HD static void phase(const Params& p, long long bid, int tid, int nthreads, char*) {
        const long long stride = p.nctas * nthreads;
        const int logN2 = p.g.logM + 1;
        for (long long i = bid * nthreads + tid; i < p.n; i += stride) {
            const long long k2 = i & ((1ll << logN2) - 1);
            const long long k1 = (i >> logN2) & ((1ll << p.g.logN1) - 1);
            const long long k0 = (i >> (logN2 + p.g.logN1)) & ((1ll << p.g.logN0) - 1);
            const long long ksq = sq_mirror(k0, p.g.logN0) + sq_mirror(k1, p.g.logN1) + sq_mirror(k2, logN2);
            p.out[i] = ld_ro(p.table + ksq) * p.scale;
        }
    }
};

}  // namespace nb2
