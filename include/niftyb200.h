/* niftyb200 -- C ABI of the B200-native correlated-field / metric-CG hot path.
 *
 * This is the drop-in boundary.  The reference (NIFTy 8.5,
 * pure Python) reaches its native arithmetic through exactly four callables,
 *   ducc_dispatch.fftn / ifftn / hartley / vdot       (src/ducc_dispatch.py:66-86)
 * bound to ducc0's pybind11 API (src/ducc_dispatch.py:61-64):
 *   ducc0.fft.c2c(a, axes=, forward=, inorm=, nthreads=)
 *   ducc0.fft.genuine_hartley(a, axes=, nthreads=) / genuine_fht(...)
 *   ducc0.misc.vdot(a, b)
 * Each entry point below names the reference call it replaces.  Everything is
 * extern "C", plain pointers and sizes; all array pointers are DEVICE pointers unless
 * the name ends in `_host`.  Functions return 0 on success, a negative code otherwise;
 * nb2_last_error() gives the message of the last failure on the calling thread.
 * `stream` is a cudaStream_t passed as void* (NULL = default stream); calls are
 * asynchronous with respect to the host unless stated otherwise.
 */
#ifndef NIFTYB200_H
#define NIFTYB200_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define NB2_F32 0
#define NB2_F64 1

/* Hartley conventions of nifty8.config "hartley_convention" (src/config.py:3-5,
 * src/ducc_dispatch.py:46-50,75-78): 0 = "non_canonical_hartley" (Re F + Im F, default),
 * 1 = "canonical_hartley" (Re F - Im F). */
#define NB2_HARTLEY_NONCANONICAL 0
#define NB2_HARTLEY_CANONICAL 1
/* flag OR-ed into the convention of nb2_hartley: product of (leading axes) x (last axis), see there */
#define NB2_HARTLEY_PRODUCT_LAST 2

#define NB2_OK 0
#define NB2_ERR_ARG (-1)
#define NB2_ERR_CUDA (-2)
#define NB2_ERR_UNSUPPORTED (-3)

typedef struct nb2_plan nb2_plan;

/* Point-wise operator fused into the first load (prologue) or last store (epilogue) of a
 * transform.  value' = value * amp(k) * scale [* mul[i]] + add_scalar [+ add[i]];
 * if `dot` is set, sum_i dot[i]*value'[i] is reduced in fp64 into *dot_out.
 * mul/add/add_scalar/dot are honoured in the epilogue only.
 *   amp_mode 0: no amplitude
 *   amp_mode 1: amp = table indexed by ksq = sum_j min(k_j, N_j-k_j)^2 over the transformed
 *               axes (isotropic grids; replaces PowerDistributor's gather through the int64
 *               pindex, src/operators/distributors.py:114-119, src/domains/power_space.py:170-198);
 *               fast path only
 *   amp_mode 2: amp = full-size factor field of one batch item (any grid).
 *               amp_flags & NB2_AMP_SYMMETRIC: the caller guarantees field[k] == field[-k mod N] along
 *               every transformed axis (true for every amplitude that is a function of the per-axis
 *               mode lengths, i.e. every PowerDistributor output, rg_space.py:105-116).  The fast-path
 *               kernels then read the canonical member of each mirror pair of rows, so the second
 *               read of a pair is served by L2 instead of HBM.  Results are bit-identical. */
#define NB2_AMP_SYMMETRIC 1
typedef struct nb2_pointop {
    int32_t amp_mode;
    int32_t amp_flags;   /* NB2_AMP_SYMMETRIC or 0 */
    const void* amp;
    double scale;
    const void* mul;
    const void* add;
    double add_scalar;
    const void* dot;
    double* dot_out;
    /* Spectrum cotangent of the correlated field (epilogue of a fast-path nb2_hartley only):
     * bin_out[ksq(i)] += bin_x[i] * (transform[i] * scale), i.e. the value BEFORE the amplitude
     * factor, binned by the integer squared mode length -- PowerDistributor.adjoint_times
     * (src/operators/distributors.py:105-112 via special_add_at, src/utilities.py:223-240) in the
     * ksq parametrisation of the bins (nb2_plan_ksq_size entries, fp64, accumulated with atomics:
     * zero it first).  NULL = off. */
    const void* bin_x;
    double* bin_out;
    /* Order-independent (bit-reproducible) variant of the binned reduction: if bin_fixed != NULL the
     * contributions are accumulated as two-limb fixed-point integers with integer atomics (exact, hence
     * independent of the order in which CTAs arrive) instead of the fp64 atomics of bin_out, which is then
     * ignored.  Layout: int64 [2 + 2 * nbins], zeroed by the caller: entry 0 is set non-zero if a
     * contribution was not finite or exceeded 2^62 / bin_fixed_scale (entry 1: unused); entry 2+2b / 3+2b =
     * high / low limb of bin b, value = (hi + lo * 2^-40) / bin_fixed_scale.  bin_fixed_scale: a power of two with
     * |bin_x[i] * transform[i] * scale| * bin_fixed_scale <= 2^42 for every i (e.g. from
     * |transform[i]| <= sqrt(N) * ||in||_2), which leaves room for 2^20 contributions per bin. */
    int64_t* bin_fixed;
    double bin_fixed_scale;
    /* Fused direction update of the conjugate gradient (PROLOGUE of a fast-path transform whose first
     * pass is a strided pass, i.e. >= 2 transformed axes; NULL = off): before anything else
     *   in'[i] = pre_add[i] + pre_scale * in[i]        (d = r + beta d, conjugate_gradient.py:124,
     *                                                    re/conjugate_gradient.py:205)
     * and, if pre_out != NULL, in'[i] is also stored to pre_out[i] (pre_out may be `in` itself: the
     * update then happens in place while the transform reads it).  pre_add / pre_out are full-size
     * arrays like `in`.  Saves the separate axpby sweep (one read of d). */
    const void* pre_add;
    double pre_scale;
    void* pre_out;
    /* Device-resident conjugate gradient: if not NULL, pre_scale is read from this DEVICE double when the
     * kernel runs (beta = state[NB2_CG_BETA], see nb2_cg_decide) instead of the host value above. */
    const double* pre_scale_dev;
} nb2_pointop;

int nb2_version(void);
const char* nb2_last_error(void);
/* number of SMs of the current device (148 on B200) */
int nb2_sm_count(void);

/* Plan for transforms over `axes` of a C-contiguous array of `shape`.
 * Replaces the per-call planning inside ducc0.fft.* (src/ducc_dispatch.py:62-78). */
int nb2_plan_create(int ndim, const int64_t* shape, int naxes, const int32_t* axes, int dtype, nb2_plan** plan);
int nb2_plan_destroy(nb2_plan* plan);
/* 1 if the plan runs on the power-of-two fast path (3 fused passes), 0 for the generic path */
int nb2_plan_is_fast(const nb2_plan* plan);
/* 0: generic path; 1: fast path over the trailing axes (3 fused passes); 2: fast path over 1-2
 * leading axes with a contiguous inner batch (product spaces, e.g. the spatial sub-space of a
 * space x frequency field: HarmonicTransformOperator(..., space=0),
 * src/library/correlated_fields.py:790-793): strided passes + one mirror-combine pass. */
int nb2_plan_kind(const nb2_plan* plan);
/* 1 if nb2_c2c runs on the fast path: trailing 1..3 power-of-two axes with rows of 2..8192 complex numbers
 * (the strided passes + the contiguous-axis c2c kernel); 0: the generic O(n^2)-per-axis path */
int nb2_plan_is_c2c_fast(const nb2_plan* plan);
/* Mirror rule of the contiguous (row) pass for slab-decomposed grids: the Hartley partner of
 * row (k0,k1) is ((N0-k0-shift0) mod N0, (N1-k1-shift1) mod N1).  0 = index negation (a full
 * axis, default); 1 = pure reflection (the mirror-symmetric local part of a distributed axis,
 * whose two halves are each other's mirror images).  Fast-path plans only. */
int nb2_plan_set_mirror(nb2_plan* plan, int shift0, int shift1);
/* number of entries an amp_mode-1 table must have: 1 + sum_j (N_j/2)^2 ; 0 if not fast */
int64_t nb2_plan_ksq_size(const nb2_plan* plan);

/* Genuine (non-separable) Hartley transform, unnormalised, with optional fused point-wise
 * operators.  Replaces ducc_dispatch.hartley (src/ducc_dispatch.py:75-78,46-50) and, with
 * pro/epi, the correlated-field hot tail ht(azm*corr*xi)+offset
 * (src/library/correlated_fields.py:801-818, src/re/correlated_field.py:690-693) and its
 * adjoint.  `in` is not modified unless in == out (in-place is allowed).
 *
 * convention | NB2_HARTLEY_PRODUCT_LAST (fast-path plans): the transformed axes are a product of two
 * spaces, (leading axes) x (last axis); the result is the 1-D Hartley transform over the last axis
 * composed with the genuine Hartley transform over the leading axes -- what the reference computes with
 * one HarmonicTransformOperator per sub-space (src/library/correlated_fields.py:790-793,
 * src/re/correlated_field.py:671-679) -- out of the SAME three passes: with F the joint spectrum of the
 * real input, H[k, k2] = Re F[k, -k2] + sgn Im F[k, k2], and both terms come from the two rows of a
 * mirror pair.  amp_mode 1 and bin_x (|k|^2 of ONE isotropic space) are rejected with it. */
int nb2_hartley(nb2_plan* plan, const void* in, void* out, int convention, const nb2_pointop* pro,
                const nb2_pointop* epi, void* stream);

/* Expand an amp_mode-1 table into the full-size factor field of one batch item,
 * out[k] = table[ksq(k)] * scale: PowerDistributor.times (src/operators/distributors.py:114-119)
 * evaluated once per expansion point, for use as an amp_mode-2 field.  Fast-path plans only. */
int nb2_expand_table(nb2_plan* plan, const void* table, double scale, void* out, void* stream);

/* Profiling / tuning hook: run ONE pass of the fast path (which = 0: strided pass over axis 0,
 * 1: strided pass over axis 1, 2: contiguous row pass with the Hartley combine).  `pro` is
 * honoured by passes 0/1/2, `epi` by pass 2 only.  Fast-path plans only. */
int nb2_hartley_pass(nb2_plan* plan, int which, const void* in, void* out, int convention,
                     const nb2_pointop* pro, const nb2_pointop* epi, void* stream);

/* Slab decomposition (no counterpart in the reference, which shards samples only:
 * src/utilities.py:295-390): strided pass `which` (0 or 1) whose input and/or output rows are
 * remapped so that the pack / unpack around the all-to-all costs no extra sweep.  With a table,
 * row j (0 <= j < N_which) of outer item o starts at complex offset  o * outer_stride + rowoff[j]
 * (tables are device int64 arrays, offsets in complex elements = pairs of real scalars);
 * NULL table = natural layout on that side.  `ncols` > 0 restricts the pass to a subset of the
 * complex columns of the natural side, so that the exchange can be pipelined in chunks of the
 * last axis: dense column j (0 <= j < ncols) is natural column
 * col0 + (j / run) * run_stride + j % run  (run = 0: one run of ncols columns; powers of two);
 * a remapped side is dense in j.  Cannot run in place when a table is given. */
int nb2_strided_pass_remap(nb2_plan* plan, int which, const void* in, void* out, const nb2_pointop* pro,
                           const int64_t* in_rowoff, int64_t in_outer_stride, const int64_t* out_rowoff,
                           int64_t out_outer_stride, int64_t col0, int64_t ncols, int64_t run, int64_t run_stride,
                           void* stream);

/* Scatter form of the same idea (peer-memory exchange): natural input; output row j of outer item o,
 * natural complex column col (0 <= col < inner, M = complex length of the last axis) is stored at
 *   out + out_rowoff[j] + (out_outer_off ? out_outer_off[o] : 0)
 *       + (out_col_off ? out_col_off[col / M] + col % M : col)
 * so that a rank can store every row straight into its owner's buffer IN THE OWNER'S NATURAL LAYOUT
 * (the pass after the exchange then needs no table).  Tables are device int64 arrays of complex
 * offsets; out_rowoff is mandatory.  Cannot run in place. */
int nb2_strided_pass_scatter(nb2_plan* plan, int which, const void* in, void* out, const nb2_pointop* pro,
                             const int64_t* out_rowoff, const int64_t* out_outer_off, const int64_t* out_col_off,
                             void* stream);

/* Peer-visible device buffers for the slab exchange: with row-offset tables that point into the
 * peers' mapped buffers (offsets relative to `out`, unified virtual addressing), the strided pass
 * before the exchange stores its output straight into the other GPUs' memory over NVLink -- the
 * all-to-all becomes part of the pass.  nb2_device_alloc: plain device allocation whose handle
 * (64 bytes, CUDA IPC) nb2_ipc_export writes; nb2_ipc_open maps a peer process's buffer into this
 * process (peer access is enabled on demand) and nb2_ipc_close unmaps it. */
int nb2_device_alloc(int64_t bytes, void** ptr);
int nb2_device_free(void* ptr);
int nb2_ipc_export(void* ptr, void* handle64);
int nb2_ipc_open(const void* handle64, void** ptr);
int nb2_ipc_close(void* ptr);

/* Total number of kernels this library has launched in the calling process. */
int64_t nb2_launch_count(void);
/* How many of them were launches of the persistent TMA-fed strided pass (nb2_tma.cuh). */
int64_t nb2_tma_launch_count(void);

/* Complex-to-complex FFT, out = scale * FFT(in); forward != 0 -> exp(-i..), else exp(+i..).
 * Replaces ducc_dispatch.fftn / ifftn (src/ducc_dispatch.py:66-72; ifftn = backward, scale 1/N).
 * `in`/`out`: interleaved complex of the plan's precision. */
int nb2_c2c(nb2_plan* plan, const void* in, void* out, int forward, double scale, void* stream);

/* sum_i a[i]*b[i] over n real scalars (complex data: pass 2n and conjugate on the caller's
 * side), fp64 accumulation, deterministic.  Replaces ducc_dispatch.vdot
 * (src/ducc_dispatch.py:81-86).  *out_dev is a device double. */
int nb2_vdot(int dtype, int64_t n, const void* a, const void* b, double* out_dev, void* stream);

/* nb2_vdot WITHOUT any allocation.  No counterpart in the reference: added
 * for callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving loops,
 * harnesses that forbid it).  Same result contract as nb2_vdot (fp64 accumulation, deterministic; the
 * summation order differs from nb2_vdot's, so the two agree to rounding, not bit for bit).  The per-CTA
 * partial sums live in a file-scope `static __device__` array of NB2_VDOT_NOALLOC_MAX_PARTIALS doubles
 * (512 KiB, held by every GPU that loads the library until exit); one CTA reduces
 * NB2_VDOT_NOALLOC_CHUNK elements, hence n <= NB2_VDOT_NOALLOC_MAX_N (= 2^30 scalars, one 1024^3 field);
 * larger n is rejected with NB2_ERR_ARG and a message.  The call makes no cudaMalloc / cudaFree / allocating
 * library call (the first call on a device resolves the array's address with cudaGetSymbolAddress).
 * Every call on a device shares the one array: it keeps the previous call's partials (every slot that is
 * read is overwritten first, nothing needs to start at zero), and two calls that overlap on the GPU
 * (different streams, or threads) RACE -- serialise them. */
#define NB2_VDOT_NOALLOC_CHUNK 16384
#define NB2_VDOT_NOALLOC_MAX_PARTIALS 65536
#define NB2_VDOT_NOALLOC_MAX_N ((int64_t)NB2_VDOT_NOALLOC_CHUNK * NB2_VDOT_NOALLOC_MAX_PARTIALS)
int nb2_vdot_noalloc(int dtype, int64_t n, const void* a, const void* b, double* out_dev, void* stream);

/* y = a*x + b*y.  The vector updates of ConjugateGradient (src/minimization/
 * conjugate_gradient.py:101,124) and re `_cg` (src/re/conjugate_gradient.py:169,205). */
int nb2_axpby(int dtype, int64_t n, double a, const void* x, double b, void* y, void* stream);

/* Fused CG step.  mode 0: x -= alpha d; r -= alpha q.  mode 1: r = q - b (exact residual of a
 * reset step, q = A x).  Then sums_dev[0..2] = (r.r, x.r, x.b) in fp64.
 * (src/minimization/conjugate_gradient.py:100-109, src/minimization/quadratic_energy.py:35-37,
 *  src/re/conjugate_gradient.py:169-191) */
int nb2_cg_update(int dtype, int64_t n, int mode, double alpha, const void* d, const void* q, const void* b,
                  void* x, void* r, double* sums_dev, void* stream);

/* ---- The O(n_bins) amplitude (power-spectrum) chain of the correlated field on the device, and its VJP
 * (SURVEY.md 8f N1).  Replaces non_parametric_amplitude.correlate (src/re/correlated_field.py:277-304) with the
 * integrated Wiener process (src/re/gauss_markov.py:105-121) and amp.at[1:].mul(1/azm) (:605-611); numpy twin
 * _Amplitude (src/library/correlated_fields.py:276-385).  All arrays are DEVICE fp64 arrays:
 *   rel (nb) relative_log_mode_lengths, dt (nb-2) log_volume, mult (nb) mode multiplicity, xi (nb-2, 2) the
 *   white excitations of the Wiener process (NULL / flx < 0: no flexibility; asp < 0: no asperity);
 *   flu, slope, flx, asp, azm are the prior-transformed scalars; power_kind 1 = "power", 0 = "amplitude".
 * forward: amp (nb) = normalised amplitude (amp[0] = totvol); optionally also scattered into a table
 *   (table[table_idx[k]] = amp[k], float if table_f32 else double: the ksq table of amp_mode 1);
 *   scratch (2 nb + 2 doubles) keeps what the VJP needs.
 * vjp: given ampbar (nb, cotangent of amp; entry 0 ignored) -> xibar (nb-2, 2) and
 *   scalars_bar[5] = cotangents of (flu, slope, flx, asp, azm) through the chain.  Needs `work` (2 nb doubles);
 *   the forward's scratch is only read, so jvp / vjp may be called any number of times per forward. */
int nb2_amplitude_forward(int64_t nb, const double* rel, const double* dt, const double* mult, double totvol,
                          const double* xi, double flu, double slope, double flx, double asp, double azm, int power_kind,
                          double* scratch, double* amp, void* table, const int64_t* table_idx, int table_f32, void* stream);
/* jvp: tangents dxi (nb-2, 2; NULL = 0) and dscalars[5] = tangents of (flu, slope, flx, asp, azm) -> damp (nb),
 *   the tangent of amp (damp[0] = 0).  Reads the forward's scratch, needs `work` (2 nb doubles) of its own. */
int nb2_amplitude_jvp(int64_t nb, const double* rel, const double* dt, const double* mult, double totvol,
                      const double* xi, double flu, double slope, double flx, double asp, double azm, int power_kind,
                      const double* scratch, const double* amp, const double* dxi, const double* dscalars_host,
                      double* work, double* damp, void* stream);
int nb2_amplitude_vjp(int64_t nb, const double* rel, const double* dt, const double* mult, double totvol,
                      const double* xi, double flu, double slope, double flx, double asp, double azm, int power_kind,
                      const double* scratch, const double* amp, const double* ampbar, double* work, double* xibar,
                      double* scalars_bar, void* stream);

/* ---- Line-of-sight response of the tomography configuration.  Replaces
 * SamplingCartesianGridLOS.__call__ / _los (src/re/extra/sampling_los.py:15-91; multilinear interpolation of
 * jax.scipy.ndimage.map_coordinates, order 1, constant mode, cval = NaN) and its linear transpose.
 * shape / distances: HOST arrays of ndim (1..3) entries; start / end: DEVICE fp64 arrays (n_los, ndim), or (ndim)
 * with the stride argument 0 (one point shared by every line); x / out / ybar / xbar in the precision `dtype`.
 * forward: out[l] = |end_l - start_l| / n * sum_a interp(x, point_a).  adjoint: xbar += R^T ybar (zero xbar first;
 * accumulated with atomics, so the summation order -- not the result to rounding -- varies between runs). */
int nb2_los_forward(int dtype, int ndim, const int64_t* shape, const double* distances, int64_t n_los, int n_sampling,
                    const double* start, int64_t start_stride, const double* end, int64_t end_stride, const void* x,
                    void* out, void* stream);
int nb2_los_adjoint(int dtype, int ndim, const int64_t* shape, const double* distances, int64_t n_los, int n_sampling,
                    const double* start, int64_t start_stride, const double* end, int64_t end_stride, const void* ybar,
                    void* xbar, void* stream);

/* ---- Device-resident conjugate gradient (static_cg, src/re/conjugate_gradient.py:219-389: the reference keeps
 * the whole loop on the device inside lax.while_loop, :382).  The scalars of the recurrence live in a DEVICE
 * array `state` of NB2_CG_STATE_SIZE doubles (integers stored as doubles); the kernels of an iteration read
 * alpha / beta from it and write their reductions into it, and nb2_cg_decide evaluates the reference's
 * stopping tests on the device (cg_single_step, :266-313).  While `*gate != 0` ("stopped") every kernel this
 * library launches on the calling thread returns at once (nb2_set_launch_gate), so the host may enqueue
 * iterations ahead of the decision and poll state[NB2_CG_INFO] only now and then. */
#define NB2_CG_GAMMA 0        /* r.r of the last completed iteration (:291) */
#define NB2_CG_ENERGY 1       /* vdot((r - j)/2, pos) (:303) */
#define NB2_CG_CURV 2         /* d.q of the running iteration: the metric apply's dot_out (:271) */
#define NB2_CG_ALPHA 3        /* gamma / curv (:274), written by stage 0 */
#define NB2_CG_BETA 4         /* max(0, gamma_new / gamma_old) (:315), written by stage 1 */
#define NB2_CG_SUMS 5         /* 5,6,7: r.r, x.r, x.b of nb2_cg_update_dev */
#define NB2_CG_ITER 8         /* completed iterations i */
#define NB2_CG_INFO 9         /* -2 running, -1 error, 0 converged, i > 0 iteration limit (:276-313) */
#define NB2_CG_MINITER 10
#define NB2_CG_MAXITER 11
#define NB2_CG_RESNORM 12     /* threshold on sqrt(r.r) (norm_ord 2); < 0: criterion off */
#define NB2_CG_ABSDELTA 13    /* threshold on the energy decrease; NaN: criterion off */
#define NB2_CG_EPS 14         /* 6 * finfo.eps  (:259) */
#define NB2_CG_TINY 15        /* 6 * finfo.tiny (:260) */
#define NB2_CG_STATE_SIZE 16
/* stage 0 (after the metric apply): info = -1 if curv == 0; alpha = gamma / curv; info = -1 if alpha < 0.
 * stage 1 (after nb2_cg_update_dev [and the all-reduce of the sums]): the tests of :292-313 in the reference's
 * order, beta, iteration count; *gate = (info >= -1).  One thread; asynchronous on `stream`. */
int nb2_cg_decide(double* state, int* gate, int stage, void* stream);
/* Every later launch of this library on the calling host thread checks `*gate` on the device first and
 * returns at once if it is non-zero.  NULL switches the check off.  The pointer must stay valid. */
int nb2_set_launch_gate(const int* gate);
/* Single-process loops: while set, a reduction of this library (calling host thread) whose result lands in
 * state[NB2_CG_CURV] (the curvature dot of a transform epilogue) or state[NB2_CG_SUMS..+2] (nb2_cg_update_dev) runs
 * the matching nb2_cg_decide stage in the SAME launch, so an iteration needs two launches less.  Do not use it when
 * the sums are all-reduced between ranks before the decision.  (NULL, NULL) switches it off. */
int nb2_set_cg_fusion(double* state, int* gate);
/* nb2_cg_update with alpha = state[NB2_CG_ALPHA] read on the device; the sums go to state[NB2_CG_SUMS..]. */
int nb2_cg_update_dev(int dtype, int64_t n, int mode, double* state, const void* d, const void* q, const void* b,
                      void* x, void* r, void* stream);
/* y = (a * *a_dev) x + (b * *b_dev) y with the device factors optional (NULL = 1): `pos -= alpha d` of a reset
 * step (:277, a = -1, a_dev = &state[NB2_CG_ALPHA]) and the unfused `d = r + beta d` (:315). */
int nb2_axpby_dev(int dtype, int64_t n, double a, const double* a_dev, const void* x, double b, const double* b_dev,
                  void* y, void* stream);

/* Host-buffer convenience wrappers (numpy drop-in path): copy in, transform, copy out,
 * synchronous. */
int nb2_hartley_host(nb2_plan* plan, const void* in_host, void* out_host, int convention, double scale);
int nb2_c2c_host(nb2_plan* plan, const void* in_host, void* out_host, int forward, double scale);
/* sum_i a[i]*b[i] of two HOST arrays of n real scalars into *out_host (fp64 accumulation on the device):
 * the numpy drop-in form of ducc_dispatch.vdot (src/ducc_dispatch.py:81-86).  Synchronous. */
int nb2_vdot_host(int dtype, int64_t n, const void* a_host, const void* b_host, double* out_host);

#ifdef __cplusplus
}
#endif
#endif /* NIFTYB200_H */
