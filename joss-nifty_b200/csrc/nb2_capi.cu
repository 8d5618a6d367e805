// niftyb200 -- C ABI implementation (see include/niftyb200.h).
//
// Built with nvcc for sm_100a into libniftyb200.so.  With -DNB2_EMU (host compiler) the
// same file builds the sequential thread emulator used by the CPU-only index tests.
#include <complex>
#include <mutex>

#include "nb2_amplitude.cuh"
#include "nb2_c2c.cuh"
#include "nb2_generic.cuh"
#include "nb2_los.cuh"
#include "nb2_plan.h"
#include "nb2_passes.cuh"
#include "nb2_vector.cuh"

using namespace nb2;

namespace {

// Reduction scratch of the vector kernels (nb2_vdot, nb2_cg_update): one buffer per DEVICE, shared by all
// host threads and streams of the process -- the vector kernels and the plans are single-stream objects:
// two reductions must not be in flight on different streams of one device at the same time.
struct Ctx {
    double* partials = nullptr;
    long long npartials = 0;
};
Ctx& ctx() {
    static Ctx c[kMaxDevices];
    return c[rt_device() % kMaxDevices];
}
int ensure_ctx() {
    static std::mutex m;
    Ctx& c = ctx();
    if (!c.partials) {
        std::lock_guard<std::mutex> lock(m);
        if (!c.partials) {
            const long long n = (long long)rt_sm_count() * 8 * 4;
            double* p = nullptr;
            if (rt_malloc((void**)&p, n * sizeof(double))) return NB2_ERR_CUDA;
            c.npartials = n;
            c.partials = p;
        }
    }
    return 0;
}
long long vec_grid(long long n) {
    const long long want = (n + kVecThreads - 1) / kVecThreads;
    return std::max(1ll, std::min(want, (long long)rt_sm_count() * 8));
}

}  // namespace


namespace nb2 {
PlanDims plan_dims(int logn) {
    switch (logn) {
        case 1: return {2, 1, 1};
        case 2: return {4, 1, 1};
        case 3: return {8, 1, 1};
        case 4: return {16, 1, 1};
        case 5: return {32, 1, 1};
        case 6: return {8, 1, 8};
        case 7: return {16, 1, 8};
        case 8: return {16, 1, 16};
        case 9: return {32, 1, 16};
        case 10: return {32, 1, 32};
        case 11: return {16, 8, 16};
        case 12: return {16, 16, 16};
        case 13: return {32, 16, 16};
        default: return {0, 0, 0};
    }
}
int pitch_of(int logn) {
    const PlanDims d = plan_dims(logn);
    if (!d.r1) return 0;
    return d.r1 * d.r2 * d.r3 + d.r1 * d.r2 + (d.r2 > 1 ? d.r1 : 0);
}
// nb2_set_cg_fusion: reductions that land in the CG state of the calling thread also run the scalar step
struct CgFusion {
    double* state = nullptr;
    int* gate = nullptr;
};
CgFusion& cg_fusion() {
    static thread_local CgFusion f;
    return f;
}
int launch_final_reduce(const double* partials, long long nparts, int nv, double* out, rt_stream st) {
    FinalReduce::Params fp{partials, nparts, nv, out, nullptr, nullptr, 0};
    const CgFusion& f = cg_fusion();
    if (f.state && out == f.state + NB2_CG_CURV && nv == 1) {
        fp.cg_state = f.state, fp.cg_gate = f.gate, fp.cg_stage = 0;
    } else if (f.state && out == f.state + NB2_CG_SUMS && nv == 3) {
        fp.cg_state = f.state, fp.cg_gate = f.gate, fp.cg_stage = 1;
    }
    return rt_launch<FinalReduce>(1, kVecThreads, FinalReduce::smem_bytes(kVecThreads), st, fp);
}
}  // namespace nb2
namespace {

template <typename T>
int upload_table(void** dev, long long n, long long denom) {
    std::vector<cplx<T>> h(n);
    const long double tau = 6.283185307179586476925286766559005768394L;
    for (long long j = 0; j < n; ++j) {
        // exact octant reduction keeps the table accurate to the last bit of T
        const long double a = tau * (long double)j / (long double)denom;
        h[j].re = (T)cosl(a);
        h[j].im = (T)(-sinl(a));
    }
    if (rt_malloc(dev, n * sizeof(cplx<T>))) return NB2_ERR_CUDA;
    return rt_h2d(*dev, h.data(), n * sizeof(cplx<T>), nullptr);
}

template <typename T>
int hartley_fast(nb2_plan* pl, const void* in, void* out, int conv, const nb2_pointop* pro, const nb2_pointop* epi,
                 rt_stream st) {
    const void* cur = in;
    const nb2_pointop* p = pro;
    int rc;
    if (pl->g.logN0 > 0) {
        if ((rc = run_strided<T>(pl, 0, cur, out, p, st))) return rc;
        cur = out;
        p = nullptr;
    }
    if (pl->g.logN1 > 0) {
        if ((rc = run_strided<T>(pl, 1, cur, out, p, st))) return rc;
        cur = out;
        p = nullptr;
    }
    return run_rows<T>(pl, cur, out, conv, p, epi, st);
}

// leading axes + inner batch (product spaces): strided passes, then the mirror combine
template <typename T>
int hartley_inner(nb2_plan* pl, const void* in, void* out, int conv, const nb2_pointop* pro, const nb2_pointop* epi,
                  rt_stream st) {
    if ((pro && pro->amp_mode == 1) || (epi && epi->amp_mode == 1)) {
        last_error() = "amp_mode 1 (ksq table) needs a transform over the trailing axes";
        return NB2_ERR_UNSUPPORTED;
    }
    const void* cur = in;
    const nb2_pointop* p = pro;
    int rc;
    if (pl->g.logN0 > 0) {
        if ((rc = run_strided<T>(pl, 0, cur, out, p, st))) return rc;
        cur = out;
        p = nullptr;
    }
    if (pl->g.logN1 > 0) {
        if ((rc = run_strided<T>(pl, 1, cur, out, p, st))) return rc;
        cur = out;
        p = nullptr;
    }
    if (p) {
        last_error() = "prologue without a transformed axis";
        return NB2_ERR_UNSUPPORTED;
    }
    const RowPairs rp = row_pairs(pl->g);
    const long long work = pl->g.batch * rp.Q << pl->g.logM;
    const long long grid = vec_grid(work);
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
    typename MirrorCombine<T>::Params mp;
    mp.in = (const cplx<T>*)cur;
    mp.out = (T*)out;
    mp.g = pl->g;
    mp.conv = conv;
    mp.epi = to_pointop<T>(epi, pl->partials);
    mp.nctas = grid;
    rc = rt_launch<MirrorCombine<T>>(grid, kVecThreads, MirrorCombine<T>::smem_bytes(kVecThreads), st, mp);
    if (rc) return rc;
    if (want_dot) rc = launch_final_reduce(pl->partials, grid, 1, epi->dot_out, st);
    return rc;
}

// ---------------------------------------------------------------- generic path
template <typename T>
int ensure_work(nb2_plan* pl) {
    if (!pl->work0) {
        if (rt_malloc(&pl->work0, pl->total * sizeof(cplx<T>))) return NB2_ERR_CUDA;
        if (rt_malloc(&pl->work1, pl->total * sizeof(cplx<T>))) return NB2_ERR_CUDA;
    }
    return 0;
}

// The generic path treats one axis at a time: lines are gathered into a contiguous buffer, transformed with the
// power-of-two contiguous-axis kernel and scattered back.  Axes of kBlueMinN points and more take this route (below
// it the direct sum is as fast and exact to the last bit): power-of-two lengths up to 2^26 directly, any other
// length up to 2^23 by the chirp-z (Bluestein) convolution with padded lines of up to 2^24 points.
constexpr long long kBlueMinN = 32;
constexpr int kRowMaxLog = 13, kLineMaxLog = 26, kBlueMaxLogL = 24;

inline int blue_logl(long long n) {
    int l = 1;
    while ((1ll << l) < 2 * n - 1) ++l;
    return l;
}
inline int ilog2ll(long long n) {
    int l = 0;
    while ((1ll << l) < n) ++l;
    return l;
}

// in-place radix-2 transform of a power-of-two host vector (plan creation only)
inline void host_fft(std::vector<std::complex<long double>>& a) {
    const size_t n = a.size();
    for (size_t i = 1, j = 0; i < n; ++i) {
        size_t bit = n >> 1;
        for (; j & bit; bit >>= 1) j ^= bit;
        j ^= bit;
        if (i < j) std::swap(a[i], a[j]);
    }
    const long double pi = 3.141592653589793238462643383279502884L;
    for (size_t len = 2; len <= n; len <<= 1) {
        for (size_t i = 0; i < n; i += len)
            for (size_t k = 0; k < len / 2; ++k) {
                const long double ang = -2 * pi * (long double)k / (long double)len;
                const std::complex<long double> w(cosl(ang), sinl(ang));
                const std::complex<long double> u = a[i + k], v = a[i + k + len / 2] * w;
                a[i + k] = u + v;
                a[i + k + len / 2] = u - v;
            }
    }
}

template <typename T>
int grow(void** buf, long long* have, long long want, rt_stream st) {
    if (want <= *have) return 0;
    rt_sync(st);
    rt_free(*buf);
    *buf = nullptr;
    *have = 0;
    if (rt_malloc(buf, want * sizeof(cplx<T>))) return NB2_ERR_CUDA;
    *have = want;
    return 0;
}

// roots of the contiguous-axis kernel for lines of 2^logl points (cached per length)
template <typename T>
int line_table(nb2_plan* pl, int logl, const void** tw) {
    if (!pl->line_tw[logl] && upload_table<T>(&pl->line_tw[logl], 1ll << logl, 1ll << logl)) return NB2_ERR_CUDA;
    *tw = pl->line_tw[logl];
    return 0;
}

// In-place unnormalised transform (forward or backward sign) of `nlines` contiguous lines of 2^logl points.
// Up to 8192 points: one pass of the row kernel.  Longer lines: four-step -- view the line as (n1, n2); transpose,
// n1-point rows, transpose back with the twiddles W_n^(j2 k1), n2-point rows, transpose into natural order.
template <typename T>
int lines_fft(nb2_plan* pl, cplx<T>* buf, long long nlines, int logl, int forward, rt_stream st) {
    int rc;
    const void* tw;
    const int cj = forward ? 0 : 1;          // backward = conj(fft(conj(.)))
    if (logl <= kRowMaxLog) {
        if ((rc = line_table<T>(pl, logl, &tw))) return rc;
        return run_rows_c2c<T>(buf, buf, tw, logl, nlines, cj, cj, 1.0, st);
    }
    const int l1 = logl / 2, l2 = logl - l1;
    const long long n = 1ll << logl, n1 = 1ll << l1, n2 = 1ll << l2;
    if ((rc = grow<T>(&pl->fs_buf, &pl->fs_buf_elems, nlines * n, st))) return rc;
    cplx<T>* tmp = (cplx<T>*)pl->fs_buf;
    typedef LineTranspose<T> LT;
    const size_t sm = LT::smem_bytes(256);
    typename LT::Params a{buf, tmp, nlines, n1, n2, 0, forward};                   // (n1, n2) -> (n2, n1)
    if ((rc = rt_launch<LT>(LT::grid(a), 256, sm, st, a))) return rc;
    if ((rc = line_table<T>(pl, l1, &tw))) return rc;
    if ((rc = run_rows_c2c<T>(tmp, tmp, tw, l1, nlines * n2, cj, cj, 1.0, st))) return rc;   // over j1 -> [j2][k1]
    typename LT::Params b{tmp, buf, nlines, n2, n1, 1, forward};                   // -> [k1][j2] * W_n^(j2 k1)
    if ((rc = rt_launch<LT>(LT::grid(b), 256, sm, st, b))) return rc;
    if ((rc = line_table<T>(pl, l2, &tw))) return rc;
    if ((rc = run_rows_c2c<T>(buf, buf, tw, l2, nlines * n1, cj, cj, 1.0, st))) return rc;   // over j2 -> [k1][k2]
    typename LT::Params c{buf, tmp, nlines, n1, n2, 0, forward};                   // -> [k2][k1]: k = k1 + n1 k2
    if ((rc = rt_launch<LT>(LT::grid(c), 256, sm, st, c))) return rc;
    return rt_d2d(buf, tmp, nlines * n * sizeof(cplx<T>), st);
}

// chirp and filter spectrum of axis t (first use)
template <typename T>
int blue_prepare(nb2_plan* pl, int t, long long n, rt_stream st) {
    if (pl->blue_w[t]) return 0;
    const int logL = blue_logl(n);
    const long long L = 1ll << logL;
    const long double pi = 3.141592653589793238462643383279502884L;
    std::vector<cplx<T>> hw(n);
    std::vector<std::complex<long double>> bt;
    const bool on_host = logL <= kRowMaxLog;       // long filters are transformed on the device
    if (on_host) bt.assign(L, std::complex<long double>(0, 0));
    std::vector<cplx<T>> hb(on_host ? 0 : L, mk<T>(T(0), T(0)));
    for (long long j = 0; j < n; ++j) {
        const long long q = (long long)(((unsigned __int128)j * (unsigned __int128)j) % (unsigned __int128)(2 * n));
        const long double ang = -pi * (long double)q / (long double)n;   // exact phase reduction: j^2 mod 2n
        const long double c = cosl(ang), sn = sinl(ang);
        hw[j] = mk<T>((T)c, (T)sn);
        if (on_host) {
            bt[j] = std::complex<long double>(c, -sn);
            if (j) bt[L - j] = bt[j];
        } else {
            hb[j] = mk<T>((T)c, (T)-sn);
            if (j) hb[L - j] = hb[j];
        }
    }
    if (rt_malloc(&pl->blue_w[t], n * sizeof(cplx<T>)) || rt_malloc(&pl->blue_B[t], L * sizeof(cplx<T>)))
        return NB2_ERR_CUDA;
    if (rt_h2d(pl->blue_w[t], hw.data(), n * sizeof(cplx<T>), nullptr)) return NB2_ERR_CUDA;
    if (on_host) {
        host_fft(bt);
        std::vector<cplx<T>> hB(L);
        for (long long m = 0; m < L; ++m) hB[m] = mk<T>((T)bt[m].real(), (T)bt[m].imag());
        if (rt_h2d(pl->blue_B[t], hB.data(), L * sizeof(cplx<T>), nullptr)) return NB2_ERR_CUDA;
    } else {
        if (rt_h2d(pl->blue_B[t], hb.data(), L * sizeof(cplx<T>), nullptr)) return NB2_ERR_CUDA;
        rt_sync(nullptr);
        if (int rc = lines_fft<T>(pl, (cplx<T>*)pl->blue_B[t], 1, logL, 1, st)) return rc;
    }
    pl->blue_logL[t] = logL;
    return 0;
}

// gather / scatter of the lines of one axis: through shared-memory tiles when the lines are strided (inner > 1)
template <typename T>
int blue_gather(const typename BlueGather<T>::Params& gp, rt_stream st) {
    if (gp.inner >= 8)
        return rt_launch<BlueGatherTiled<T>>(BlueGatherTiled<T>::grid(gp), 256, BlueGatherTiled<T>::smem_bytes(256), st, gp);
    return rt_launch<BlueGather<T>>(gp.nctas, kVecThreads, 0, st, gp);
}
template <typename T>
int blue_scatter(const typename BlueScatter<T>::Params& sp, rt_stream st) {
    if (sp.inner >= 8)
        return rt_launch<BlueScatterTiled<T>>(BlueScatterTiled<T>::grid(sp), 256, BlueScatterTiled<T>::smem_bytes(256), st, sp);
    return rt_launch<BlueScatter<T>>(sp.nctas, kVecThreads, 0, st, sp);
}

// one axis of a generic plan: power-of-two lengths = gather, transform the lines, scatter; other lengths = the same
// around the chirp-z convolution (gather * chirp into zero-padded lines, transform, * filter spectrum, transform
// back, * chirp on the scatter)
template <typename T>
int line_axis(nb2_plan* pl, int t, const cplx<T>* in, cplx<T>* out, long long outer, long long n, long long inner,
              int forward, rt_stream st) {
    const long long nlines = outer * inner;
    const bool pow2 = (n & (n - 1)) == 0;
    int rc;
    if (!pow2 && (rc = blue_prepare<T>(pl, t, n, st))) return rc;
    const int logL = pow2 ? ilog2ll(n) : pl->blue_logL[t];
    const long long L = 1ll << logL;
    if (pow2 && inner == 1 && logL <= kRowMaxLog) {     // contiguous power-of-two lines: the row kernel as it is
        const void* tw;
        if ((rc = line_table<T>(pl, logL, &tw))) return rc;
        return run_rows_c2c<T>(in, out, tw, logL, nlines, forward ? 0 : 1, forward ? 0 : 1, 1.0, st);
    }
    if ((rc = grow<T>(&pl->blue_buf, &pl->blue_buf_elems, nlines * L, st))) return rc;
    cplx<T>* buf = (cplx<T>*)pl->blue_buf;
    const cplx<T>* w = pow2 ? nullptr : (const cplx<T>*)pl->blue_w[t];
    const long long gridL = vec_grid(nlines * L), gridN = vec_grid(nlines * n);
    typename BlueGather<T>::Params gp{in, buf, w, outer, n, inner, L, forward, gridL};
    if ((rc = blue_gather<T>(gp, st))) return rc;
    if (pow2) {
        if ((rc = lines_fft<T>(pl, buf, nlines, logL, forward, st))) return rc;
    } else {
        if ((rc = lines_fft<T>(pl, buf, nlines, logL, 1, st))) return rc;
        typename BlueMulConj<T>::Params mp{buf, (const cplx<T>*)pl->blue_B[t], nlines, L, forward, gridL};
        if ((rc = rt_launch<BlueMulConj<T>>(gridL, kVecThreads, 0, st, mp))) return rc;
        if ((rc = lines_fft<T>(pl, buf, nlines, logL, 1, st))) return rc;
    }
    typename BlueScatter<T>::Params sp{buf, out, w, outer, n, inner, L, forward, gridN};
    return blue_scatter<T>(sp, st);
}

inline bool line_axis_takes(long long n) {
    if (n < kBlueMinN) return false;
    if ((n & (n - 1)) == 0) return ilog2ll(n) <= kLineMaxLog;
    return blue_logl(n) <= kBlueMaxLogL;
}

// transform work buffer `a` over all plan axes; result pointer returned in *res (a or b)
template <typename T>
int generic_axes(nb2_plan* pl, cplx<T>* a, cplx<T>* b, int forward, rt_stream st, cplx<T>** res) {
    for (int t = 0; t < pl->naxes; ++t) {
        const int ax = pl->axes[t];
        {
            const long long n = pl->shape[ax];
            if (line_axis_takes(n) && pl->knob_bluestein) {
                long long outer = 1, inner = 1;
                for (int i = 0; i < ax; ++i) outer *= pl->shape[i];
                for (int i = ax + 1; i < pl->ndim; ++i) inner *= pl->shape[i];
                const int rc = line_axis<T>(pl, t, a, b, outer, n, inner, forward, st);
                if (rc) return rc;
                std::swap(a, b);
                continue;
            }
        }
        typename NaiveDft<T>::Params p;
        p.in = a;
        p.out = b;
        p.root = (const cplx<double>*)pl->roots[t];
        p.n = pl->shape[ax];
        p.outer = 1;
        p.inner = 1;
        for (int i = 0; i < ax; ++i) p.outer *= pl->shape[i];
        for (int i = ax + 1; i < pl->ndim; ++i) p.inner *= pl->shape[i];
        p.forward = forward;
        p.nctas = vec_grid(pl->total);
        if (p.n > 1) {
            int rc = rt_launch<NaiveDft<T>>(p.nctas, kVecThreads, 0, st, p);
            if (rc) return rc;
            std::swap(a, b);
        }
    }
    *res = a;
    return 0;
}

template <typename T>
int eltwise(nb2_plan* pl, const T* in, T* out, const nb2_pointop* op, rt_stream st) {
    if (op->amp_mode == 1) {
        last_error() = "amp_mode 1 (ksq table) requires the power-of-two fast path";
        return NB2_ERR_UNSUPPORTED;
    }
    const long long grid = vec_grid(pl->total);
    if (op->dot && grid > pl->npartials) {
        rt_sync(st);
        rt_free(pl->partials);
        pl->partials = nullptr;
        if (rt_malloc((void**)&pl->partials, grid * sizeof(double))) return NB2_ERR_CUDA;
        pl->npartials = grid;
    }
    typename Eltwise<T>::Params p;
    p.in = in;
    p.out = out;
    p.op = to_pointop<T>(op, pl->partials);
    p.n = pl->total;
    // amp field spans the trailing block that contains all transformed axes
    p.item = 1;
    for (int i = pl->axes[0]; i < pl->ndim; ++i) p.item *= pl->shape[i];
    p.nctas = grid;
    int rc = rt_launch<Eltwise<T>>(grid, kVecThreads, Eltwise<T>::smem_bytes(kVecThreads), st, p);
    if (rc) return rc;
    if (op->dot) {
        if (!op->dot_out) {
            last_error() = "nb2_pointop.dot set without dot_out";
            return NB2_ERR_ARG;
        }
        rc = launch_final_reduce(pl->partials, grid, 1, op->dot_out, st);
    }
    return rc;
}

template <typename T>
int hartley_generic(nb2_plan* pl, const void* in, void* out, int conv, const nb2_pointop* pro,
                    const nb2_pointop* epi, rt_stream st) {
    int rc;
    if ((rc = ensure_work<T>(pl))) return rc;
    const long long grid = vec_grid(pl->total);
    const T* src = (const T*)in;
    if (pro) {
        if ((rc = eltwise<T>(pl, (const T*)in, (T*)out, pro, st))) return rc;
        src = (const T*)out;
    }
    typename RealToCplx<T>::Params rp;
    rp.in = src;
    rp.out = (cplx<T>*)pl->work0;
    rp.n = pl->total;
    rp.nctas = grid;
    if ((rc = rt_launch<RealToCplx<T>>(grid, kVecThreads, 0, st, rp))) return rc;
    cplx<T>* res = nullptr;
    if ((rc = generic_axes<T>(pl, (cplx<T>*)pl->work0, (cplx<T>*)pl->work1, 1, st, &res))) return rc;
    typename CplxToHartley<T>::Params hp;
    hp.in = res;
    hp.out = (T*)out;
    hp.scale = T(1);
    hp.conv = conv;
    hp.n = pl->total;
    hp.nctas = grid;
    if ((rc = rt_launch<CplxToHartley<T>>(grid, kVecThreads, 0, st, hp))) return rc;
    if (epi) rc = eltwise<T>(pl, (const T*)out, (T*)out, epi, st);
    return rc;
}

// c2c over trailing power-of-two axes: the strided passes as they are (complex FFTs over rows of complex
// columns) + the contiguous-axis kernel of nb2_c2c.cuh.  Backward: conj(fft(conj(x))) -- one conjugation
// sweep in front when a strided pass comes first, the last conjugation and the scale on the last store.
template <typename T>
int c2c_fast(nb2_plan* pl, const void* in, void* out, int forward, double scale, rt_stream st) {
    int rc;
    const Geom& g = pl->gc;
    if (!pl->ctwr) {
        if (g.logN0 && (rc = upload_table<T>(&pl->ctw0, 1ll << g.logN0, 1ll << g.logN0))) return rc;
        if (g.logN1 && (rc = upload_table<T>(&pl->ctw1, 1ll << g.logN1, 1ll << g.logN1))) return rc;
        if ((rc = upload_table<T>(&pl->ctwr, 1ll << g.logM, 1ll << g.logM))) return rc;
    }
    nb2_plan view = *pl;          // the strided passes read geometry and tables from the plan: complex-row view
    view.g = g;
    view.tw0 = pl->ctw0;
    view.tw1 = pl->ctw1;
    view.fast = true;
    const bool strided = g.logN0 > 0 || g.logN1 > 0;
    const void* cur = in;
    const long long total = g.batch << (g.logN0 + g.logN1 + g.logM);
    if (!forward && strided) {
        const long long grid = vec_grid(total);
        typename CplxConjScale<T>::Params cp{(const cplx<T>*)in, (cplx<T>*)out, T(1), 1, total, grid};
        if ((rc = rt_launch<CplxConjScale<T>>(grid, kVecThreads, 0, st, cp))) return rc;
        cur = out;
    }
    if (g.logN0 > 0) {
        if ((rc = run_strided<T>(&view, 0, cur, out, nullptr, st))) return rc;
        cur = out;
    }
    if (g.logN1 > 0) {
        if ((rc = run_strided<T>(&view, 1, cur, out, nullptr, st))) return rc;
        cur = out;
    }
    return run_rows_c2c<T>(cur, out, pl->ctwr, g.logM, total >> g.logM, (!forward && !strided) ? 1 : 0, forward ? 0 : 1,
                           scale, st);
}

template <typename T>
int c2c_impl(nb2_plan* pl, const void* in, void* out, int forward, double scale, rt_stream st) {
    int rc;
    if (pl->cfast) return c2c_fast<T>(pl, in, out, forward, scale, st);
    if ((rc = ensure_work<T>(pl))) return rc;
    if ((rc = rt_d2d(pl->work0, in, pl->total * sizeof(cplx<T>), st))) return rc;
    cplx<T>* res = nullptr;
    if ((rc = generic_axes<T>(pl, (cplx<T>*)pl->work0, (cplx<T>*)pl->work1, forward, st, &res))) return rc;
    typename CplxScale<T>::Params sp;
    sp.in = res;
    sp.out = (cplx<T>*)out;
    sp.scale = (T)scale;
    sp.n = pl->total;
    sp.nctas = vec_grid(pl->total);
    return rt_launch<CplxScale<T>>(sp.nctas, kVecThreads, 0, st, sp);
}

}  // namespace

// ================================================================== extern "C"
extern "C" {

int nb2_version(void) { return 100; }
const char* nb2_last_error(void) { return last_error().c_str(); }
int nb2_sm_count(void) { return rt_sm_count(); }

int nb2_device_alloc(int64_t bytes, void** ptr) {
    if (!ptr || bytes < 0) {
        last_error() = "nb2_device_alloc: invalid argument";
        return NB2_ERR_ARG;
    }
    if (ensure_ctx()) return NB2_ERR_CUDA;
    return rt_malloc(ptr, (size_t)bytes) ? NB2_ERR_CUDA : NB2_OK;
}
int nb2_device_free(void* ptr) { return rt_free(ptr) ? NB2_ERR_CUDA : NB2_OK; }
int nb2_ipc_export(void* ptr, void* handle64) {
    if (!ptr || !handle64) {
        last_error() = "nb2_ipc_export: invalid argument";
        return NB2_ERR_ARG;
    }
    return rt_ipc_export(ptr, handle64);
}
int nb2_ipc_open(const void* handle64, void** ptr) {
    if (!ptr || !handle64) {
        last_error() = "nb2_ipc_open: invalid argument";
        return NB2_ERR_ARG;
    }
    if (ensure_ctx()) return NB2_ERR_CUDA;
    return rt_ipc_open(handle64, ptr);
}
int nb2_ipc_close(void* ptr) { return ptr ? rt_ipc_close(ptr) : NB2_OK; }

int nb2_plan_create(int ndim, const int64_t* shape, int naxes, const int32_t* axes, int dtype, nb2_plan** out) {
    if (!out || !shape || ndim < 1 || ndim > kMaxDim || naxes < 0 || naxes > ndim ||
        (dtype != NB2_F32 && dtype != NB2_F64)) {
        last_error() = "nb2_plan_create: invalid argument";
        return NB2_ERR_ARG;
    }
    if (ensure_ctx()) return NB2_ERR_CUDA;
    nb2_plan* pl = new (std::nothrow) nb2_plan();
    if (!pl) return NB2_ERR_ARG;
    pl->ndim = ndim;
    pl->dtype = dtype;
    {
        const char* e;
        if ((e = std::getenv("NB2_TMA"))) pl->knob_tma = e[0] != '0';
        if ((e = std::getenv("NB2_TMA_SCATTER"))) pl->knob_tma_scatter = e[0] == '1';
        if ((e = std::getenv("NB2_TMA_FX"))) pl->knob_tma_fx = e[0] != '0';
        if ((e = std::getenv("NB2_BLUESTEIN"))) pl->knob_bluestein = e[0] != '0';
    }
    for (int i = 0; i < ndim; ++i) {
        if (shape[i] < 1) {
            delete pl;
            last_error() = "nb2_plan_create: non-positive extent";
            return NB2_ERR_ARG;
        }
        pl->shape[i] = shape[i];
        pl->total *= shape[i];
    }
    std::vector<int> ax;
    if (!axes || naxes == 0) {
        for (int i = 0; i < ndim; ++i) ax.push_back(i);
    } else {
        for (int i = 0; i < naxes; ++i) {
            int a = axes[i] < 0 ? axes[i] + ndim : axes[i];
            if (a < 0 || a >= ndim || std::find(ax.begin(), ax.end(), a) != ax.end()) {
                delete pl;
                last_error() = "nb2_plan_create: invalid or repeated axis";
                return NB2_ERR_ARG;
            }
            ax.push_back(a);
        }
    }
    std::sort(ax.begin(), ax.end());
    pl->naxes = (int)ax.size();
    for (int i = 0; i < pl->naxes; ++i) pl->axes[i] = ax[i];

    // ---- fast path eligibility: trailing 1..3 axes, powers of two, last >= 4
    bool fast = pl->naxes >= 1 && pl->naxes <= 3;
    for (int i = 0; fast && i < pl->naxes; ++i) {
        if (pl->axes[i] != ndim - pl->naxes + i) fast = false;
        if (!is_pow2(pl->shape[pl->axes[i]])) fast = false;
    }
    if (fast && pl->shape[ndim - 1] < 4) fast = false;
    if (fast) {
        long long n[3] = {1, 1, 1};
        for (int i = 0; i < pl->naxes; ++i) n[3 - pl->naxes + i] = pl->shape[pl->axes[i]];
        pl->g.logN0 = ilog2(n[0]);
        pl->g.logN1 = ilog2(n[1]);
        pl->g.logM = ilog2(n[2]) - 1;
        pl->g.batch = 1;
        pl->g.shift0 = pl->g.shift1 = 0;
        for (int i = 0; i < ndim - pl->naxes; ++i) pl->g.batch *= pl->shape[i];
        const size_t es = dtype == NB2_F32 ? sizeof(cplx<float>) : sizeof(cplx<double>);
        if (pl->g.logN0 > 13 || pl->g.logN1 > 13 || pl->g.logM > 13 || pl->g.logM < 1) fast = false;
        if (fast && pl->g.logN0 && size_t(pitch_of(pl->g.logN0)) * es > kSmemLimit) fast = false;
        if (fast && pl->g.logN1 && size_t(pitch_of(pl->g.logN1)) * es > kSmemLimit) fast = false;
        if (fast && size_t(2 * pitch_of(pl->g.logM)) * es + 2048 > kSmemLimit) fast = false;
    }
    pl->fast = fast;
    // ---- complex-to-complex fast path: trailing 1..3 power-of-two axes, rows of 2..8192 complex numbers
    {
        bool cf = pl->naxes >= 1 && pl->naxes <= 3;
        for (int i = 0; cf && i < pl->naxes; ++i)
            if (pl->axes[i] != ndim - pl->naxes + i || !is_pow2(pl->shape[pl->axes[i]])) cf = false;
        if (cf) {
            long long n[3] = {1, 1, 1};
            for (int i = 0; i < pl->naxes; ++i) n[3 - pl->naxes + i] = pl->shape[pl->axes[i]];
            Geom& gc = pl->gc;
            gc.logN0 = ilog2(n[0]);
            gc.logN1 = ilog2(n[1]);
            gc.logM = ilog2(n[2]);
            gc.batch = 1;
            gc.shift0 = gc.shift1 = 0;
            for (int i = 0; i < ndim - pl->naxes; ++i) gc.batch *= pl->shape[i];
            const size_t es = dtype == NB2_F32 ? sizeof(cplx<float>) : sizeof(cplx<double>);
            if (gc.logM < 1 || gc.logM > 13 || gc.logN0 > 13 || gc.logN1 > 13) cf = false;
            if (cf && gc.logN0 && size_t(pitch_of(gc.logN0)) * es > kSmemLimit) cf = false;
            if (cf && gc.logN1 && size_t(pitch_of(gc.logN1)) * es > kSmemLimit) cf = false;
            if (cf && size_t(pitch_of(gc.logM)) * es > kSmemLimit) cf = false;
        }
        pl->cfast = cf;
    }
    // ---- leading 1..2 contiguous axes with an even power-of-two-pairs inner batch
    if (!fast && pl->naxes >= 1 && pl->naxes <= 2 && pl->axes[pl->naxes - 1] < ndim - 1 &&
        (pl->naxes == 1 || pl->axes[1] == pl->axes[0] + 1)) {
        long long innerB = 1, outerB = 1;
        for (int i = pl->axes[pl->naxes - 1] + 1; i < ndim; ++i) innerB *= pl->shape[i];
        for (int i = 0; i < pl->axes[0]; ++i) outerB *= pl->shape[i];
        bool ok = innerB >= 2 && innerB % 2 == 0 && is_pow2(innerB / 2);
        long long n[2] = {1, 1};
        for (int i = 0; i < pl->naxes; ++i) {
            n[2 - pl->naxes + i] = pl->shape[pl->axes[i]];
            if (!is_pow2(pl->shape[pl->axes[i]])) ok = false;
        }
        if (ok) {
            pl->g.logN0 = ilog2(n[0]);
            pl->g.logN1 = ilog2(n[1]);
            pl->g.logM = ilog2(innerB / 2);
            pl->g.batch = outerB;
            pl->g.shift0 = pl->g.shift1 = 0;
            const size_t es = dtype == NB2_F32 ? sizeof(cplx<float>) : sizeof(cplx<double>);
            if (pl->g.logN0 > 13 || pl->g.logN1 > 13 || pl->g.logM > 40) ok = false;
            if (ok && pl->g.logN0 && size_t(pitch_of(pl->g.logN0)) * es > kSmemLimit) ok = false;
            if (ok && pl->g.logN1 && size_t(pitch_of(pl->g.logN1)) * es > kSmemLimit) ok = false;
            if (ok && !(pl->g.logN0 || pl->g.logN1)) ok = false;
        }
        pl->inner = ok;
    }
    int rc = 0;
    if (pl->inner) {
        if (dtype == NB2_F32) {
            if (pl->g.logN0) rc |= upload_table<float>(&pl->tw0, 1ll << pl->g.logN0, 1ll << pl->g.logN0);
            if (pl->g.logN1) rc |= upload_table<float>(&pl->tw1, 1ll << pl->g.logN1, 1ll << pl->g.logN1);
        } else {
            if (pl->g.logN0) rc |= upload_table<double>(&pl->tw0, 1ll << pl->g.logN0, 1ll << pl->g.logN0);
            if (pl->g.logN1) rc |= upload_table<double>(&pl->tw1, 1ll << pl->g.logN1, 1ll << pl->g.logN1);
        }
    }
    if (fast) {
        if (dtype == NB2_F32) {
            if (pl->g.logN0) rc |= upload_table<float>(&pl->tw0, 1ll << pl->g.logN0, 1ll << pl->g.logN0);
            if (pl->g.logN1) rc |= upload_table<float>(&pl->tw1, 1ll << pl->g.logN1, 1ll << pl->g.logN1);
            rc |= upload_table<float>(&pl->twr, 2ll << pl->g.logM, 2ll << pl->g.logM);
        } else {
            if (pl->g.logN0) rc |= upload_table<double>(&pl->tw0, 1ll << pl->g.logN0, 1ll << pl->g.logN0);
            if (pl->g.logN1) rc |= upload_table<double>(&pl->tw1, 1ll << pl->g.logN1, 1ll << pl->g.logN1);
            rc |= upload_table<double>(&pl->twr, 2ll << pl->g.logM, 2ll << pl->g.logM);
        }
    }
    for (int t = 0; t < pl->naxes && !rc; ++t) {
        const long long n = pl->shape[pl->axes[t]];
        rc |= upload_table<double>(&pl->roots[t], n, n);
    }
    if (rc) {
        nb2_plan_destroy(pl);
        return NB2_ERR_CUDA;
    }
    *out = pl;
    return NB2_OK;
}

int nb2_plan_destroy(nb2_plan* pl) {
    if (!pl) return NB2_OK;
    rt_free(pl->tw0);
    rt_free(pl->tw1);
    rt_free(pl->twr);
    rt_free(pl->ctw0);
    rt_free(pl->ctw1);
    rt_free(pl->ctwr);
    rt_free(pl->partials);
    rt_free(pl->work0);
    rt_free(pl->work1);
    for (int t = 0; t < kMaxDim; ++t) {
        rt_free(pl->roots[t]);
        rt_free(pl->blue_w[t]);
        rt_free(pl->blue_B[t]);
    }
    for (int l = 0; l < 16; ++l) rt_free(pl->line_tw[l]);
    rt_free(pl->blue_buf);
    rt_free(pl->fs_buf);
    delete pl;
    return NB2_OK;
}

int nb2_plan_set_mirror(nb2_plan* pl, int shift0, int shift1) {
    if (!pl || !pl->fast || (shift0 & ~1) || (shift1 & ~1) || (shift0 && !pl->g.logN0) || (shift1 && !pl->g.logN1)) {
        last_error() = "nb2_plan_set_mirror: invalid argument";
        return NB2_ERR_ARG;
    }
    pl->g.shift0 = shift0;
    pl->g.shift1 = shift1;
    return NB2_OK;
}

int nb2_plan_is_fast(const nb2_plan* pl) { return pl && pl->fast ? 1 : 0; }
int nb2_plan_is_c2c_fast(const nb2_plan* pl) { return pl && pl->cfast ? 1 : 0; }
int nb2_plan_kind(const nb2_plan* pl) { return !pl ? -1 : (pl->fast ? 1 : (pl->inner ? 2 : 0)); }

int64_t nb2_plan_ksq_size(const nb2_plan* pl) {
    if (!pl || !pl->fast) return 0;
    const long long h0 = (1ll << pl->g.logN0) / 2, h1 = (1ll << pl->g.logN1) / 2, h2 = 1ll << pl->g.logM;
    return 1 + h0 * h0 + h1 * h1 + h2 * h2;
}

// pre_add / pre_out (fused direction update) belong to the prologue of a transform that starts with a
// strided pass; `strided_first`: such a pass exists in what is about to run
static int check_pre_update(const nb2_plan* pl, const nb2_pointop* pro, const nb2_pointop* epi, const void* out,
                            bool strided_first) {
    if (epi && (epi->pre_add || epi->pre_out)) {
        last_error() = "nb2_pointop.pre_add / pre_out are prologue fields";
        return NB2_ERR_ARG;
    }
    if (!pro || (!pro->pre_add && !pro->pre_out)) return NB2_OK;
    if (!pro->pre_add || pro->pre_out == out) {
        last_error() = "nb2_pointop.pre_out needs pre_add and must not alias the output of the transform";
        return NB2_ERR_ARG;
    }
    if (!strided_first || !(pl->fast || pl->inner)) {
        last_error() = "nb2_pointop.pre_add: the transform must start with a strided pass of the fast path";
        return NB2_ERR_UNSUPPORTED;
    }
    return NB2_OK;
}

int nb2_hartley(nb2_plan* pl, const void* in, void* out, int conv, const nb2_pointop* pro, const nb2_pointop* epi,
                void* stream) {
    if (!pl || !in || !out || conv < 0 || conv > 3) {
        last_error() = "nb2_hartley: invalid argument";
        return NB2_ERR_ARG;
    }
    if (conv & NB2_PRODUCT_LAST) {
        if (!pl->fast) {
            last_error() = "nb2_hartley: NB2_HARTLEY_PRODUCT_LAST needs a fast-path plan (see nb2_plan_is_fast)";
            return NB2_ERR_UNSUPPORTED;
        }
        if ((pro && pro->amp_mode == 1) || (epi && (epi->amp_mode == 1 || epi->bin_x))) {
            last_error() = "nb2_hartley: |k|^2 tables and the binned reduction belong to one isotropic space, "
                           "not to a product of spaces (use amp_mode 2)";
            return NB2_ERR_ARG;
        }
    }
    if (int rc = check_pre_update(pl, pro, epi, out, pl->g.logN0 > 0 || pl->g.logN1 > 0)) return rc;
    rt_stream st = (rt_stream)stream;
    if (epi && epi->bin_x && epi->bin_fixed && !(epi->bin_fixed_scale > 0.0)) {
        last_error() = "nb2_hartley: bin_fixed needs a positive power-of-two bin_fixed_scale";
        return NB2_ERR_ARG;
    }
    if ((pro && pro->bin_x) || (epi && epi->bin_x && (!pl->fast || !(epi->bin_out || epi->bin_fixed)))) {
        last_error() = "nb2_hartley: the binned reduction (bin_x/bin_out) is an epilogue of fast-path plans";
        return NB2_ERR_UNSUPPORTED;
    }
    if (pl->fast)
        return pl->dtype == NB2_F32 ? hartley_fast<float>(pl, in, out, conv, pro, epi, st)
                                    : hartley_fast<double>(pl, in, out, conv, pro, epi, st);
    if (pl->inner)
        return pl->dtype == NB2_F32 ? hartley_inner<float>(pl, in, out, conv, pro, epi, st)
                                    : hartley_inner<double>(pl, in, out, conv, pro, epi, st);
    return pl->dtype == NB2_F32 ? hartley_generic<float>(pl, in, out, conv, pro, epi, st)
                                : hartley_generic<double>(pl, in, out, conv, pro, epi, st);
}

int nb2_expand_table(nb2_plan* pl, const void* table, double scale, void* out, void* stream) {
    if (!pl || !table || !out) {
        last_error() = "nb2_expand_table: invalid argument";
        return NB2_ERR_ARG;
    }
    if (!pl->fast) {
        last_error() = "nb2_expand_table: fast-path plans only";
        return NB2_ERR_UNSUPPORTED;
    }
    const long long n = 1ll << (pl->g.logN0 + pl->g.logN1 + pl->g.logM + 1);
    const long long grid = vec_grid(n);
    rt_stream st = (rt_stream)stream;
    if (pl->dtype == NB2_F32) {
        ExpandTable<float>::Params p{(const float*)table, (float*)out, (float)scale, pl->g, n, grid};
        return rt_launch<ExpandTable<float>>(grid, kVecThreads, 0, st, p);
    }
    ExpandTable<double>::Params p{(const double*)table, (double*)out, scale, pl->g, n, grid};
    return rt_launch<ExpandTable<double>>(grid, kVecThreads, 0, st, p);
}

int nb2_hartley_pass(nb2_plan* pl, int which, const void* in, void* out, int conv, const nb2_pointop* pro,
                     const nb2_pointop* epi, void* stream) {
    if (!pl || !in || !out || which < 0 || which > 2 || conv < 0 || conv > 3) {
        last_error() = "nb2_hartley_pass: invalid argument";
        return NB2_ERR_ARG;
    }
    if (!pl->fast || (which == 0 && !pl->g.logN0) || (which == 1 && !pl->g.logN1)) {
        last_error() = "nb2_hartley_pass: pass not part of this plan";
        return NB2_ERR_UNSUPPORTED;
    }
    if (int rc = check_pre_update(pl, pro, epi, out, which < 2)) return rc;
    rt_stream st = (rt_stream)stream;
    if (which < 2)
        return pl->dtype == NB2_F32 ? run_strided<float>(pl, which, in, out, pro, st)
                                    : run_strided<double>(pl, which, in, out, pro, st);
    return pl->dtype == NB2_F32 ? run_rows<float>(pl, in, out, conv, pro, epi, st)
                                : run_rows<double>(pl, in, out, conv, pro, epi, st);
}

int nb2_strided_pass_remap(nb2_plan* pl, int which, const void* in, void* out, const nb2_pointop* pro,
                           const int64_t* in_rowoff, int64_t in_outer_stride, const int64_t* out_rowoff,
                           int64_t out_outer_stride, int64_t col0, int64_t ncols, int64_t run, int64_t run_stride,
                           void* stream) {
    if (!pl || !in || !out || which < 0 || which > 1) {
        last_error() = "nb2_strided_pass_remap: invalid argument";
        return NB2_ERR_ARG;
    }
    if (!pl->fast || (which == 0 && !pl->g.logN0) || (which == 1 && !pl->g.logN1)) {
        last_error() = "nb2_strided_pass_remap: pass not part of this plan";
        return NB2_ERR_UNSUPPORTED;
    }
    if (col0 < 0 || ncols < 0 || run < 0 || run_stride < 0) {
        last_error() = "nb2_strided_pass_remap: negative column range";
        return NB2_ERR_ARG;
    }
    if (pro && (pro->pre_add || pro->pre_out)) {
        if (in_rowoff) {
            last_error() = "nb2_strided_pass_remap: the fused direction update needs natural input rows";
            return NB2_ERR_UNSUPPORTED;
        }
        if (int rc = check_pre_update(pl, pro, nullptr, out, true)) return rc;
    }
    if ((in_rowoff || out_rowoff) && in == out) {
        last_error() = "nb2_strided_pass_remap: a remapped pass cannot run in place";
        return NB2_ERR_ARG;
    }
    RowRemap rm;
    rm.in_rowoff = (const long long*)in_rowoff;
    rm.in_ostride = in_outer_stride;
    rm.out_rowoff = (const long long*)out_rowoff;
    rm.out_ostride = out_outer_stride;
    rm.col0 = col0;
    rm.ncols = ncols;
    rm.run = run;
    rm.run_stride = run_stride;
    rt_stream st = (rt_stream)stream;
    return pl->dtype == NB2_F32 ? run_strided<float>(pl, which, in, out, pro, st, &rm)
                                : run_strided<double>(pl, which, in, out, pro, st, &rm);
}

int nb2_strided_pass_scatter(nb2_plan* pl, int which, const void* in, void* out, const nb2_pointop* pro,
                             const int64_t* out_rowoff, const int64_t* out_outer_off, const int64_t* out_col_off,
                             void* stream) {
    if (!pl || !in || !out || !out_rowoff || which < 0 || which > 1 || in == out) {
        last_error() = "nb2_strided_pass_scatter: invalid argument";
        return NB2_ERR_ARG;
    }
    if (!pl->fast || (which == 0 && !pl->g.logN0) || (which == 1 && !pl->g.logN1)) {
        last_error() = "nb2_strided_pass_scatter: pass not part of this plan";
        return NB2_ERR_UNSUPPORTED;
    }
    if (int rc = check_pre_update(pl, pro, nullptr, out, true)) return rc;
    RowRemap rm;
    rm.out_rowoff = (const long long*)out_rowoff;
    rm.out_ooff = (const long long*)out_outer_off;
    rm.out_coloff = (const long long*)out_col_off;
    rt_stream st = (rt_stream)stream;
    return pl->dtype == NB2_F32 ? run_strided<float>(pl, which, in, out, pro, st, &rm)
                                : run_strided<double>(pl, which, in, out, pro, st, &rm);
}

int64_t nb2_launch_count(void) { return launch_counter(); }
int64_t nb2_tma_launch_count(void) { return tma_launch_counter(); }

int nb2_c2c(nb2_plan* pl, const void* in, void* out, int forward, double scale, void* stream) {
    if (!pl || !in || !out) {
        last_error() = "nb2_c2c: invalid argument";
        return NB2_ERR_ARG;
    }
    rt_stream st = (rt_stream)stream;
    return pl->dtype == NB2_F32 ? c2c_impl<float>(pl, in, out, forward, scale, st)
                                : c2c_impl<double>(pl, in, out, forward, scale, st);
}

int nb2_vdot(int dtype, int64_t n, const void* a, const void* b, double* out_dev, void* stream) {
    if (n < 0 || !out_dev || (n > 0 && (!a || !b))) {
        last_error() = "nb2_vdot: invalid argument";
        return NB2_ERR_ARG;
    }
    if (ensure_ctx()) return NB2_ERR_CUDA;
    rt_stream st = (rt_stream)stream;
    const long long grid = vec_grid(n);
    int rc;
    if (dtype == NB2_F32) {
        Dot<float>::Params p{(const float*)a, (const float*)b, n, ctx().partials, grid};
        rc = rt_launch<Dot<float>>(grid, kVecThreads, Dot<float>::smem_bytes(kVecThreads), st, p);
    } else if (dtype == NB2_F64) {
        Dot<double>::Params p{(const double*)a, (const double*)b, n, ctx().partials, grid};
        rc = rt_launch<Dot<double>>(grid, kVecThreads, Dot<double>::smem_bytes(kVecThreads), st, p);
    } else {
        return NB2_ERR_ARG;
    }
    if (rc) return rc;
    FinalReduce::Params fp{ctx().partials, grid, 1, out_dev};
    return rt_launch<FinalReduce>(1, kVecThreads, FinalReduce::smem_bytes(kVecThreads), st, fp);
}

// ---- nb2_vdot_noalloc: the dot-product reduction with its scratch in a static device array (see the header)
#if !defined(NB2_EMU)
static __device__ __align__(256) double nb2_noalloc_partials[NB2_VDOT_NOALLOC_MAX_PARTIALS];
#else
alignas(256) static double nb2_noalloc_partials[NB2_VDOT_NOALLOC_MAX_PARTIALS];
#endif

// Device address of the array on the CURRENT device.  In host code `nb2_noalloc_partials` is only a
// stand-in for the symbol, not a device address: the address comes from cudaGetSymbolAddress, once per
// device, with that device current.
static int noalloc_scratch(double** out) {
#if !defined(NB2_EMU)
    static double* addr[kMaxDevices] = {nullptr};
    static std::mutex m;
    const int dev = rt_device();
    if (dev < 0 || dev >= kMaxDevices) {
        last_error() = "nb2_vdot_noalloc: device ordinal out of range";
        return NB2_ERR_UNSUPPORTED;
    }
    std::lock_guard<std::mutex> lock(m);
    if (!addr[dev]) {
        double* p = nullptr;
        const cudaError_t err = cudaGetSymbolAddress((void**)&p, nb2_noalloc_partials);
        if (err != cudaSuccess || !p) {
            last_error() = std::string("nb2_vdot_noalloc: cudaGetSymbolAddress: ") + cudaGetErrorString(err);
            return NB2_ERR_CUDA;
        }
        addr[dev] = p;
    }
    *out = addr[dev];
#else
    *out = nb2_noalloc_partials;
#endif
    return NB2_OK;
}

int nb2_vdot_noalloc(int dtype, int64_t n, const void* a, const void* b, double* out_dev, void* stream) {
    if (n < 0 || !out_dev || (n > 0 && (!a || !b)) || (dtype != NB2_F32 && dtype != NB2_F64)) {
        last_error() = "nb2_vdot_noalloc: invalid argument";
        return NB2_ERR_ARG;
    }
    if (n > NB2_VDOT_NOALLOC_MAX_N) {
        last_error() = "nb2_vdot_noalloc: n = " + std::to_string((long long)n) + " exceeds the maximum of " +
                       std::to_string((long long)NB2_VDOT_NOALLOC_MAX_N) +
                       " scalars that the static scratch array covers (use nb2_vdot, or split the vectors)";
        return NB2_ERR_ARG;
    }
    double* partials = nullptr;
    if (int rc = noalloc_scratch(&partials)) return rc;
    rt_stream st = (rt_stream)stream;
    // one CTA per chunk of NB2_VDOT_NOALLOC_CHUNK elements: the number of partials follows n
    const long long grid = std::max<long long>(1, (n + NB2_VDOT_NOALLOC_CHUNK - 1) / NB2_VDOT_NOALLOC_CHUNK);
    int rc;
    if (dtype == NB2_F32) {
        Dot<float>::Params p{(const float*)a, (const float*)b, n, partials, grid};
        rc = rt_launch<Dot<float>>(grid, kVecThreads, Dot<float>::smem_bytes(kVecThreads), st, p);
    } else {
        Dot<double>::Params p{(const double*)a, (const double*)b, n, partials, grid};
        rc = rt_launch<Dot<double>>(grid, kVecThreads, Dot<double>::smem_bytes(kVecThreads), st, p);
    }
    if (rc) return rc;
    FinalReduce::Params fp{partials, grid, 1, out_dev};
    return rt_launch<FinalReduce>(1, kVecThreads, FinalReduce::smem_bytes(kVecThreads), st, fp);
}

int nb2_axpby_dev(int dtype, int64_t n, double a, const double* a_dev, const void* x, double b, const double* b_dev,
                  void* y, void* stream) {
    if (n < 0 || (n > 0 && (!x || !y))) {
        last_error() = "nb2_axpby: invalid argument";
        return NB2_ERR_ARG;
    }
    rt_stream st = (rt_stream)stream;
    const long long grid = vec_grid(n);
    if (dtype == NB2_F32) {
        Axpby<float>::Params p{(const float*)x, (float*)y, (float)a, (float)b, n, grid, a_dev, b_dev};
        return rt_launch<Axpby<float>>(grid, kVecThreads, 0, st, p);
    } else if (dtype == NB2_F64) {
        Axpby<double>::Params p{(const double*)x, (double*)y, a, b, n, grid, a_dev, b_dev};
        return rt_launch<Axpby<double>>(grid, kVecThreads, 0, st, p);
    }
    last_error() = "nb2_axpby: invalid dtype";
    return NB2_ERR_ARG;
}

int nb2_axpby(int dtype, int64_t n, double a, const void* x, double b, void* y, void* stream) {
    return nb2_axpby_dev(dtype, n, a, nullptr, x, b, nullptr, y, stream);
}

// x -= alpha d, r -= alpha q (mode 0) / r = q - b (mode 1), then the three sums into sums_dev
static int cg_update_impl(int dtype, int64_t n, int mode, double alpha, const double* alpha_dev, const void* d,
                          const void* q, const void* b, void* x, void* r, double* sums_dev, void* stream) {
    if (n < 0 || !sums_dev || (mode != 0 && mode != 1) || !x || !r || !q || !b || (mode == 0 && !d)) {
        last_error() = "nb2_cg_update: invalid argument";
        return NB2_ERR_ARG;
    }
    if (ensure_ctx()) return NB2_ERR_CUDA;
    rt_stream st = (rt_stream)stream;
    const long long grid = vec_grid(n);
    int rc;
    if (dtype == NB2_F32) {
        CgUpdate<float>::Params p{(float*)x, (float*)r, (const float*)d, (const float*)q, (const float*)b,
                                  (float)alpha, mode, n, ctx().partials, grid, alpha_dev};
        rc = rt_launch<CgUpdate<float>>(grid, kVecThreads, CgUpdate<float>::smem_bytes(kVecThreads), st, p);
    } else if (dtype == NB2_F64) {
        CgUpdate<double>::Params p{(double*)x, (double*)r, (const double*)d, (const double*)q, (const double*)b,
                                   alpha, mode, n, ctx().partials, grid, alpha_dev};
        rc = rt_launch<CgUpdate<double>>(grid, kVecThreads, CgUpdate<double>::smem_bytes(kVecThreads), st, p);
    } else {
        last_error() = "nb2_cg_update: invalid dtype";
        return NB2_ERR_ARG;
    }
    if (rc) return rc;
    return launch_final_reduce(ctx().partials, grid, 3, sums_dev, st);
}

int nb2_cg_update(int dtype, int64_t n, int mode, double alpha, const void* d, const void* q, const void* b,
                  void* x, void* r, double* sums_dev, void* stream) {
    return cg_update_impl(dtype, n, mode, alpha, nullptr, d, q, b, x, r, sums_dev, stream);
}

int nb2_cg_update_dev(int dtype, int64_t n, int mode, double* state, const void* d, const void* q, const void* b,
                      void* x, void* r, void* stream) {
    if (!state) {
        last_error() = "nb2_cg_update_dev: invalid argument";
        return NB2_ERR_ARG;
    }
    return cg_update_impl(dtype, n, mode, 0.0, state + NB2_CG_ALPHA, d, q, b, x, r, state + NB2_CG_SUMS, stream);
}

int nb2_cg_decide(double* state, int* gate, int stage, void* stream) {
    if (!state || !gate || (stage != 0 && stage != 1)) {
        last_error() = "nb2_cg_decide: invalid argument";
        return NB2_ERR_ARG;
    }
    CgDecide::Params p{state, gate, stage};
    return rt_launch<CgDecide>(1, 32, 0, (rt_stream)stream, p);
}

int nb2_amplitude_forward(int64_t nb, const double* rel, const double* dt, const double* mult, double totvol,
                          const double* xi, double flu, double slope, double flx, double asp, double azm, int power_kind,
                          double* scratch, double* amp, void* table, const int64_t* table_idx, int table_f32, void* stream) {
    if (nb < 1 || !rel || !mult || !scratch || !amp || (nb > 2 && !dt) || (table && !table_idx) || !(azm > 0.0)) {
        last_error() = "nb2_amplitude_forward: invalid argument";
        return NB2_ERR_ARG;
    }
    AmplitudeForward::Params p{{nb, rel, dt, mult, totvol}, {flu, slope, flx, asp, azm, power_kind}, xi,
                               scratch, scratch + nb, scratch + 2 * nb, amp, table, (const long long*)table_idx, table_f32};
    return rt_launch<AmplitudeForward>(1, kAmpThreads, AmplitudeForward::smem_bytes(kAmpThreads), (rt_stream)stream, p);
}

int nb2_amplitude_jvp(int64_t nb, const double* rel, const double* dt, const double* mult, double totvol,
                      const double* xi, double flu, double slope, double flx, double asp, double azm, int power_kind,
                      const double* scratch, const double* amp, const double* dxi, const double* dscalars_host,
                      double* work, double* damp, void* stream) {
    if (nb < 1 || !rel || !mult || !scratch || !amp || !dscalars_host || !work || !damp || (nb > 2 && !dt)) {
        last_error() = "nb2_amplitude_jvp: invalid argument";
        return NB2_ERR_ARG;
    }
    const double* d = dscalars_host;
    AmplitudeJvp::Params p{{nb, rel, dt, mult, totvol}, {flu, slope, flx, asp, azm, power_kind},
                           {d[0], d[1], d[2], d[3], d[4], power_kind}, xi, dxi, scratch + nb, amp, scratch + 2 * nb,
                           work, work + nb, damp};
    return rt_launch<AmplitudeJvp>(1, kAmpThreads, AmplitudeJvp::smem_bytes(kAmpThreads), (rt_stream)stream, p);
}

int nb2_amplitude_vjp(int64_t nb, const double* rel, const double* dt, const double* mult, double totvol,
                      const double* xi, double flu, double slope, double flx, double asp, double azm, int power_kind,
                      const double* scratch, const double* amp, const double* ampbar, double* work, double* xibar,
                      double* scalars_bar, void* stream) {
    if (nb < 1 || !rel || !mult || !scratch || !amp || !ampbar || !work || !scalars_bar || (nb > 2 && !dt) ||
        (flx >= 0.0 && nb > 2 && xi && !xibar)) {
        last_error() = "nb2_amplitude_vjp: invalid argument";
        return NB2_ERR_ARG;
    }
    AmplitudeVjp::Params p{{nb, rel, dt, mult, totvol}, {flu, slope, flx, asp, azm, power_kind}, xi, amp,
                           scratch + 2 * nb, ampbar, scratch + nb, work + nb, work, xibar, scalars_bar};
    return rt_launch<AmplitudeVjp>(1, kAmpThreads, AmplitudeVjp::smem_bytes(kAmpThreads), (rt_stream)stream, p);
}

static int los_geom(LosGeom& g, int ndim, const int64_t* shape, const double* distances, int64_t n_los, int n_sampling,
                    const double* start, int64_t start_stride, const double* end, int64_t end_stride) {
    if (ndim < 1 || ndim > kLosMaxDim || !shape || !distances || n_los < 0 || n_sampling < 1 || !start || !end ||
        (start_stride != 0 && start_stride != ndim) || (end_stride != 0 && end_stride != ndim)) {
        last_error() = "nb2_los: invalid argument";
        return NB2_ERR_ARG;
    }
    g.ndim = ndim;
    for (int i = 0; i < ndim; ++i) {
        if (shape[i] < 1 || !(distances[i] > 0.0)) {
            last_error() = "nb2_los: invalid shape / distances";
            return NB2_ERR_ARG;
        }
        g.shape[i] = shape[i];
        g.l2i[i] = (double(shape[i] - 1) / double(shape[i])) / distances[i];
    }
    g.n_los = n_los;
    g.n_sampling = n_sampling;
    g.start = start;
    g.end = end;
    g.start_stride = start_stride;
    g.end_stride = end_stride;
    return NB2_OK;
}

int nb2_los_forward(int dtype, int ndim, const int64_t* shape, const double* distances, int64_t n_los, int n_sampling,
                    const double* start, int64_t start_stride, const double* end, int64_t end_stride, const void* x,
                    void* out, void* stream) {
    LosGeom g{};
    if (int rc = los_geom(g, ndim, shape, distances, n_los, n_sampling, start, start_stride, end, end_stride)) return rc;
    if (!x || !out || (dtype != NB2_F32 && dtype != NB2_F64)) {
        last_error() = "nb2_los_forward: invalid argument";
        return NB2_ERR_ARG;
    }
    if (dtype == NB2_F32) {
        LosForward<float>::Params p{g, (const float*)x, (float*)out};
        return rt_launch<LosForward<float>>(n_los, 128, LosForward<float>::smem_bytes(128), (rt_stream)stream, p);
    }
    LosForward<double>::Params p{g, (const double*)x, (double*)out};
    return rt_launch<LosForward<double>>(n_los, 128, LosForward<double>::smem_bytes(128), (rt_stream)stream, p);
}

int nb2_los_adjoint(int dtype, int ndim, const int64_t* shape, const double* distances, int64_t n_los, int n_sampling,
                    const double* start, int64_t start_stride, const double* end, int64_t end_stride, const void* ybar,
                    void* xbar, void* stream) {
    LosGeom g{};
    if (int rc = los_geom(g, ndim, shape, distances, n_los, n_sampling, start, start_stride, end, end_stride)) return rc;
    if (!ybar || !xbar || (dtype != NB2_F32 && dtype != NB2_F64)) {
        last_error() = "nb2_los_adjoint: invalid argument";
        return NB2_ERR_ARG;
    }
    if (dtype == NB2_F32) {
        LosAdjoint<float>::Params p{g, (const float*)ybar, (float*)xbar};
        return rt_launch<LosAdjoint<float>>(n_los, 128, 0, (rt_stream)stream, p);
    }
    LosAdjoint<double>::Params p{g, (const double*)ybar, (double*)xbar};
    return rt_launch<LosAdjoint<double>>(n_los, 128, 0, (rt_stream)stream, p);
}

int nb2_set_launch_gate(const int* gate) {
    launch_gate() = gate;
    return NB2_OK;
}

int nb2_set_cg_fusion(double* state, int* gate) {
    if ((state == nullptr) != (gate == nullptr)) {
        last_error() = "nb2_set_cg_fusion: state and gate must both be set or both be NULL";
        return NB2_ERR_ARG;
    }
    cg_fusion().state = state;
    cg_fusion().gate = gate;
    return NB2_OK;
}

int nb2_hartley_host(nb2_plan* pl, const void* in_host, void* out_host, int conv, double scale) {
    if (!pl || !in_host || !out_host || (conv != 0 && conv != 1)) {
        last_error() = "nb2_hartley_host: invalid argument";
        return NB2_ERR_ARG;
    }
    const size_t es = pl->dtype == NB2_F32 ? 4 : 8;
    void* buf = nullptr;
    if (rt_malloc(&buf, pl->total * es)) return NB2_ERR_CUDA;
    int rc = rt_h2d(buf, in_host, pl->total * es, nullptr);
    nb2_pointop epi{};
    epi.scale = scale;
    if (!rc) rc = nb2_hartley(pl, buf, buf, conv, nullptr, scale == 1.0 ? nullptr : &epi, nullptr);
    if (!rc) rc = rt_d2h(out_host, buf, pl->total * es, nullptr);
    rt_free(buf);
    return rc;
}

int nb2_c2c_host(nb2_plan* pl, const void* in_host, void* out_host, int forward, double scale) {
    if (!pl || !in_host || !out_host) {
        last_error() = "nb2_c2c_host: invalid argument";
        return NB2_ERR_ARG;
    }
    const size_t es = pl->dtype == NB2_F32 ? 8 : 16;
    void* buf = nullptr;
    if (rt_malloc(&buf, pl->total * es)) return NB2_ERR_CUDA;
    int rc = rt_h2d(buf, in_host, pl->total * es, nullptr);
    if (!rc) rc = nb2_c2c(pl, buf, buf, forward, scale, nullptr);
    if (!rc) rc = rt_d2h(out_host, buf, pl->total * es, nullptr);
    rt_free(buf);
    return rc;
}

int nb2_vdot_host(int dtype, int64_t n, const void* a_host, const void* b_host, double* out_host) {
    if (n < 0 || !out_host || (n > 0 && (!a_host || !b_host)) || (dtype != NB2_F32 && dtype != NB2_F64)) {
        last_error() = "nb2_vdot_host: invalid argument";
        return NB2_ERR_ARG;
    }
    const size_t bytes = (size_t)n * (dtype == NB2_F32 ? 4 : 8);
    void *a = nullptr, *b = nullptr;
    double* out = nullptr;
    int rc = rt_malloc(&a, bytes) || rt_malloc(&b, bytes) || rt_malloc((void**)&out, sizeof(double)) ? NB2_ERR_CUDA : 0;
    if (!rc) rc = rt_h2d(a, a_host, bytes, nullptr);
    if (!rc) rc = rt_h2d(b, b_host, bytes, nullptr);
    if (!rc) rc = nb2_vdot(dtype, n, a, b, out, nullptr);
    if (!rc) rc = rt_d2h(out_host, out, sizeof(double), nullptr);
    rt_free(a);
    rt_free(b);
    rt_free(out);
    return rc;
}

}  // extern "C"
