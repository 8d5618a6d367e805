// niftyb200 -- plan object and host-side helpers shared by the translation units.
#pragma once

#include <algorithm>
#include <cmath>
#include <new>

#include "../../include/niftyb200.h"
#include "nb2_launch.cuh"
#include "nb2_pointwise.cuh"

namespace nb2 {
constexpr int kMaxDim = 8;
constexpr size_t kSmemLimit = 200 * 1024;
constexpr int kVecThreads = 256;
}  // namespace nb2

struct nb2_plan {
    int ndim = 0, naxes = 0, dtype = 0;
    long long shape[nb2::kMaxDim] = {0};
    int axes[nb2::kMaxDim] = {0};
    long long total = 1;
    bool fast = false;
    bool inner = false;  // leading transformed axes + contiguous inner batch: strided passes + MirrorCombine
    // fast path geometry: (batch, N0, N1, N2)
    nb2::Geom g{};
    void* tw0 = nullptr;  // strided axis-0 table (N0 entries)
    void* tw1 = nullptr;  // strided axis-1 table (N1 entries)
    void* twr = nullptr;  // row table (N2 entries)
    // complex-to-complex fast path over the trailing axes (nb2_c2c): geometry with M = N2 complex numbers per
    // row and its own twiddle tables (created at the first nb2_c2c call)
    bool cfast = false;
    nb2::Geom gc{};
    void* ctw0 = nullptr;
    void* ctw1 = nullptr;
    void* ctwr = nullptr;
    double* partials = nullptr;
    long long npartials = 0;
    // developer knobs, read from the environment once at plan creation (never on the dispatch path)
    bool knob_tma = true, knob_tma_scatter = false, knob_tma_fx = true, knob_bluestein = true;
    // generic path
    void* roots[nb2::kMaxDim] = {nullptr};  // cplx<double>[n] per transformed axis
    void* work0 = nullptr;
    void* work1 = nullptr;  // complex scratch
    // Bluestein (chirp-z) tables of the generic path, per transformed axis, built on first use: chirp w[n], filter
    // spectrum B[L]; one padded line buffer shared by the axes
    void* blue_w[nb2::kMaxDim] = {nullptr};
    void* blue_B[nb2::kMaxDim] = {nullptr};
    int blue_logL[nb2::kMaxDim] = {0};
    void* line_tw[16] = {nullptr};      // roots of the contiguous-axis kernel per log2(line length)
    void* blue_buf = nullptr;           // gathered (padded) lines
    long long blue_buf_elems = 0;
    void* fs_buf = nullptr;             // four-step scratch for lines longer than the row kernel takes
    long long fs_buf_elems = 0;
};


namespace nb2 {

// Row remapping of a strided pass (see StridedPass::Params); offsets in complex elements.
struct RowRemap {
    const long long* in_rowoff = nullptr;
    long long in_ostride = 0;
    const long long* out_rowoff = nullptr;
    long long out_ostride = 0;
    const long long* out_ooff = nullptr;     // scatter form: per-outer-item and per-column-block origins
    const long long* out_coloff = nullptr;
    // column subset of the natural side: ncols dense columns (0 = every column) in runs of `run`
    // contiguous columns (0 = one run) whose starts are `run_stride` apart, the first at col0
    long long col0 = 0, ncols = 0, run = 0, run_stride = 0;
};

struct PlanDims { int r1, r2, r3; };
PlanDims plan_dims(int logn);
int pitch_of(int logn);
inline int round32(long long v) { return (int)std::max(32ll, ((v + 31) / 32) * 32); }

template <typename T>
PointOp<T> to_pointop(const nb2_pointop* o, double* partials) {
    PointOp<T> r;
    r.amp_mode = o ? o->amp_mode : 0;
    r.amp_sym = (o && o->amp_mode == 2 && (o->amp_flags & NB2_AMP_SYMMETRIC)) ? 1 : 0;
    r.amp = o ? (const T*)o->amp : nullptr;
    r.scale = o ? (T)o->scale : T(1);
    r.mul = o ? (const T*)o->mul : nullptr;
    r.add = o ? (const T*)o->add : nullptr;
    r.add_scalar = o ? (T)o->add_scalar : T(0);
    r.dot = o ? (const T*)o->dot : nullptr;
    r.dot_partials = partials;
    r.bin_x = o ? (const T*)o->bin_x : nullptr;
    r.bin_out = o ? o->bin_out : nullptr;
    r.bin_fixed = o ? (long long*)o->bin_fixed : nullptr;
    r.bin_fixed_scale = o ? o->bin_fixed_scale : 0.0;
    r.pre_add = o ? (const T*)o->pre_add : nullptr;
    r.pre_scale = o ? (T)o->pre_scale : T(0);
    r.pre_out = o ? (T*)o->pre_out : nullptr;
    r.pre_scale_dev = o ? o->pre_scale_dev : nullptr;
    return r;
}

// fast-path passes (explicitly instantiated for float and double in nb2_fast_*.cu)
template <typename T>
int run_strided(const nb2_plan* pl, int axis, const void* in, void* out, const nb2_pointop* pro, rt_stream st,
                const RowRemap* rm = nullptr);
template <typename T>
int run_rows(nb2_plan* pl, const void* in, void* out, int conv, const nb2_pointop* pro, const nb2_pointop* epi,
             rt_stream st);
template <typename T> int rows_pairs_per_cta(const nb2_plan* pl);
int launch_final_reduce(const double* partials, long long nparts, int nv, double* out, rt_stream st);
template <typename T>
int run_rows_c2c(const void* in, void* out, const void* tw, int logl, long long nrows, int conj_in, int conj_out,
                 double scale, rt_stream st);

}  // namespace nb2
