// niftyb200 -- StridedTma: the strided FFT pass as a persistent, TMA-fed kernel.
//
// Same arithmetic as StridedPass<T, R1, 1, R3> (two register-radix steps with a transposition
// through shared memory) for a tile of W complex columns (W * sizeof(cplx) = 128 bytes per row
// segment) x N = R1 * R3 rows, but organised for the sm_100a memory system:
//
//   * one persistent CTA per SM walks the tiles blockIdx.x, blockIdx.x + gridDim.x, ...;
//   * the tile is fetched by the TMA engine (cp.async.bulk.tensor, boxes of 256 rows x 128 bytes)
//     into the dense buffer P and signalled through an mbarrier.  As soon as every thread has
//     copied its R1 rows from P into registers the copy of the NEXT tile is issued, so the HBM
//     reads of tile t+1 overlap the butterflies, the transposition and the stores of tile t;
//   * P cannot double as the transposition buffer (it is being refilled), and 2 x 128 KB does not
//     fit in 227 KB, so the transposition goes through a half-size buffer X in two rounds: first
//     the threads of the lower W/2 columns write and read back, then the upper half;
//   * results leave the registers with streaming 8/16-byte stores (a row segment of 128 bytes
//     per 16 (fp32) / 8 (fp64) lanes).
//
// Thread (c, r): c = tid % W column, r = tid / W in [0, R3).  Step 1 works on rows r + n1 * R3,
// step 2 on the I2 = R1 / R3 row groups k1 = r + i * R3.
#pragma once

#include "nb2_launch.cuh"
#include "nb2_passes.cuh"

#if !defined(NB2_EMU)
#include <cuda.h>
#endif

namespace nb2 {

template <typename T, int R1, int R3>
struct StridedTma {
    static_assert(R1 >= R3 && R3 > 1, "two-step plans with R1 >= R3");
    static constexpr int N = R1 * R3;
    static constexpr int MM = R3;
    static constexpr int LOGW = sizeof(T) == 4 ? 4 : 3;
    static constexpr int W = 1 << LOGW;
    static constexpr int HW = W / 2;
    static constexpr int NTHREADS = W * R3;
    static constexpr int I2 = R1 / R3;
    static constexpr int BOX_ROWS = N < 256 ? N : 256;
    static constexpr int NBOX = N / BOX_ROWS;
    static constexpr int ROW_BYTES = W * int(sizeof(cplx<T>));       // 128
    static constexpr int BOX_ELEMS = ROW_BYTES / int(sizeof(T));     // innermost box extent in scalars
    static constexpr size_t P_BYTES = size_t(N) * ROW_BYTES;
    static constexpr int XROWS = N + R1;                              // one pad row per R3 rows
    static constexpr size_t X_BYTES = size_t(XROWS) * (ROW_BYTES / 2);
    static constexpr size_t SMEM = P_BYTES + X_BYTES + 16;            // + the two mbarriers
    static constexpr int PB = R1 >= 8 ? 8 : R1;                       // prologue batch (rows)
    static constexpr int NBATCH = R1 / PB;
    static constexpr int MIN_BLOCKS = (SMEM <= 110 * 1024 && NTHREADS <= 256) ? 2 : 1;
    // Mirror-symmetric amplitude field through shared memory ("fx"): only rows 0..N/2 of the field tile are
    // needed (row j > N/2 equals row N - j); these N/2 + 1 rows of 128 bytes fit exactly into the
    // transposition buffer X, which is idle between the last read of one tile and the first write of the
    // next.  They are fetched by TMA (N/2 rows in boxes of FBOX_ROWS + a tail box of FTAIL_ROWS rows).
    static constexpr int HALF = N / 2;
    static constexpr int FBOX_ROWS = HALF < 256 ? HALF : 256;
    static constexpr int NFBOX = HALF / FBOX_ROWS;
    static constexpr int FTAIL_ROWS = 8;
    static constexpr size_t F_BYTES = size_t(HALF + FTAIL_ROWS) * ROW_BYTES;
    static_assert(F_BYTES <= X_BYTES, "the half field tile must fit into the transposition buffer");

    struct Params {
        const cplx<T>* in;
        cplx<T>* out;
        const cplx<T>* tw;          // tw[j] = exp(-2 pi i j / N)
        long long inner;            // complex elements between consecutive rows
        long long tiles_per_outer;  // inner / W
        int axis;
        int has_pro;
        PointOp<T> pro;
        Geom g;
        // scatter of the output rows (slab exchange, see StridedPass::Params): row k of outer item o and
        // column col goes to out_rowoff[k] + out_ooff[o] + out_coloff[col >> logM] + (col & (M - 1))
        const long long* out_rowoff;
        const long long* out_ooff;
        const long long* out_coloff;
        long long out_ostride;
        int fx;                     // amplitude field (amp_mode 2, symmetric) staged in X by TMA / the emulator
        long long f_outer_mask;     // outer index of the field tile = o & f_outer_mask (the field has no batch)
    };

    // registers that live across the barriers of one tile
    struct Regs {
        cplx<T> x[R1];
        cplx<T> f0[PB];   // first batch of streamed amplitude factors, fetched before the tile wait
    };

    NB2_HD static long long tile_base(const Params& p, long long tile, long long& o, long long& tc) {
        o = tile / p.tiles_per_outer;
        tc = (tile - o * p.tiles_per_outer) << LOGW;
        return o * N * p.inner + tc;
    }
    NB2_HD static int xslot(int L) { return L + L / MM; }

    // ---- stage 0a: amplitude factors of the first row batch (global loads issued before the wait)
    NB2_HD static void pro_begin(const Params& p, long long tile, int tid, Regs& R) {
        if (p.has_pro && p.pro.pre_add) {
            // fused direction update: first batch of r (the field, if any, comes through X: fx)
            const int c = tid & (W - 1), r = tid >> LOGW;
            long long o, tc;
            const long long base = tile_base(p, tile, o, tc);
            const cplx<T>* rs = reinterpret_cast<const cplx<T>*>(p.pro.pre_add) + base + c + (long long)r * p.inner;
            NB2_UNROLL
            for (int i = 0; i < PB; ++i) R.f0[i] = ldc_stream(rs + (long long)i * MM * p.inner);
            return;
        }
        if (!p.has_pro || p.pro.amp_mode != 2 || p.fx) return;
        const int c = tid & (W - 1), r = tid >> LOGW;
        long long o, tc;
        const long long base = tile_base(p, tile, o, tc);
        const long long itemmask = (1ll << (p.g.logN0 + p.g.logN1 + p.g.logM)) - 1;
        const FieldRows<T> fr(reinterpret_cast<const cplx<T>*>(p.pro.amp) + ((base + c) & itemmask), p.inner, N,
                              p.pro.amp_sym && !(p.axis == 0 ? p.g.shift0 : p.g.shift1));
        NB2_UNROLL
        for (int i = 0; i < PB; ++i) R.f0[i] = fr.load(r + i * MM);
    }

    // ---- stage 0b: tile rows from P into registers
    NB2_HD static void load_tile(const Params& p, long long tile, int tid, const char* Praw, const char* Fraw, Regs& R) {
        const cplx<T>* P = reinterpret_cast<const cplx<T>*>(Praw);
        const int c = tid & (W - 1), r = tid >> LOGW;
        if (p.has_pro && p.pro.pre_add) {
            // d' = r + beta d for the tile (d from P, r streamed in batches of PB rows, two in flight), stored
            // back to pre_out; then the field factor from X (fx) -- a plain scale is left to step1
            long long o, tc;
            const long long base = tile_base(p, tile, o, tc);
            const long long e0 = base + c + (long long)r * p.inner, rstride = (long long)MM * p.inner;
            const cplx<T>* rs = reinterpret_cast<const cplx<T>*>(p.pro.pre_add) + e0;
            cplx<T>* dd = p.pro.pre_out ? reinterpret_cast<cplx<T>*>(p.pro.pre_out) + e0 : nullptr;
            const cplx<T>* F = reinterpret_cast<const cplx<T>*>(Fraw);
            cplx<T> rr[2][PB];
            const T beta = pre_scale_of(p.pro);
            NB2_UNROLL
            for (int i = 0; i < PB; ++i) rr[0][i] = R.f0[i];
            NB2_UNROLL
            for (int g = 0; g < NBATCH; ++g) {
                if (g + 1 < NBATCH) {
                    NB2_UNROLL
                    for (int i = 0; i < PB; ++i) rr[(g + 1) & 1][i] = ldc_stream(rs + ((g + 1) * PB + i) * rstride);
                }
                NB2_UNROLL
                for (int i = 0; i < PB; ++i) {
                    const int n1 = g * PB + i, row = r + n1 * MM;
                    cplx<T> v = P[(row << LOGW) + c];
                    v.re = rr[g & 1][i].re + beta * v.re;
                    v.im = rr[g & 1][i].im + beta * v.im;
                    if (dd) stc_stream(dd + n1 * rstride, v);
                    if (p.fx) {
                        const cplx<T> f = F[((row <= N - row ? row : N - row) << LOGW) + c];
                        v.re *= f.re * p.pro.scale;
                        v.im *= f.im * p.pro.scale;
                    }
                    R.x[n1] = v;
                }
            }
            return;
        }
        if (p.fx) {
            const cplx<T>* F = reinterpret_cast<const cplx<T>*>(Fraw);   // rows 0..N/2 of the field tile, dense
            NB2_UNROLL
            for (int n1 = 0; n1 < R1; ++n1) {
                const int row = r + n1 * MM, fr = row <= N - row ? row : N - row;
                const cplx<T> v = P[(row << LOGW) + c], f = F[(fr << LOGW) + c];
                R.x[n1].re = v.re * (f.re * p.pro.scale);
                R.x[n1].im = v.im * (f.im * p.pro.scale);
            }
        } else {
            NB2_UNROLL
            for (int n1 = 0; n1 < R1; ++n1) R.x[n1] = P[((r + n1 * MM) << LOGW) + c];
        }
    }

    // ---- stage 1: prologue, radix-R1 step, inter-step twiddles (in place, bit-reversed order)
    NB2_HD static void step1(const Params& p, long long tile, int tid, Regs& R) {
        const int c = tid & (W - 1), r = tid >> LOGW;
        cplx<T>(&x)[R1] = R.x;
        if (p.has_pro) {
            long long o, tc;
            const long long base = tile_base(p, tile, o, tc);
            if (p.pro.amp_mode == 2 && p.fx) {
                // applied in load_tile
            } else if (p.pro.amp_mode == 2) {
                const long long itemmask = (1ll << (p.g.logN0 + p.g.logN1 + p.g.logM)) - 1;
                const FieldRows<T> fr(reinterpret_cast<const cplx<T>*>(p.pro.amp) + ((base + c) & itemmask), p.inner, N,
                                      p.pro.amp_sym && !(p.axis == 0 ? p.g.shift0 : p.g.shift1));
                cplx<T> f[2][PB];
                NB2_UNROLL
                for (int i = 0; i < PB; ++i) f[0][i] = R.f0[i];
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
                const int fx = int(p.axis == 0 ? sq_mirror(k1, p.g.logN1) : sq_mirror(k0, p.g.logN0));
                const T* te = p.pro.amp + fx + int(sq_mirror(2 * m, p.g.logM + 1));
                const T* to = p.pro.amp + fx + int(sq_mirror(2 * m + 1, p.g.logM + 1));
                NB2_UNROLL
                for (int n1 = 0; n1 < R1; ++n1) {
                    const int row = r + n1 * MM, km = row < N - row ? row : N - row;
                    x[n1].re *= ld_ro(te + km * km) * p.pro.scale;
                    x[n1].im *= ld_ro(to + km * km) * p.pro.scale;
                }
            } else {
                NB2_UNROLL
                for (int n1 = 0; n1 < R1; ++n1) x[n1] = x[n1] * p.pro.scale;
            }
        }
        fft_reg<R1, true>(x);
        cplx<T> wp[R1];
        tw_powers<R1>(ldc_ro(p.tw + r), wp);
        NB2_UNROLL
        for (int k1 = 1; k1 < R1; ++k1) x[brev<R1>(k1)] = x[brev<R1>(k1)] * wp[k1];
    }

    // ---- stages 2..5: transposition through X, one half of the columns per round
    NB2_HD static void xwrite(int tid, char* Xraw, const Regs& R) {
        cplx<T>* X = reinterpret_cast<cplx<T>*>(Xraw);
        const int ch = tid & (HW - 1), r = tid >> LOGW;
        NB2_UNROLL
        for (int k1 = 0; k1 < R1; ++k1) X[xslot(k1 * MM + r) * HW + ch] = R.x[brev<R1>(k1)];
    }
    NB2_HD static void xread(int tid, const char* Xraw, Regs& R) {
        const cplx<T>* X = reinterpret_cast<const cplx<T>*>(Xraw);
        const int ch = tid & (HW - 1), r = tid >> LOGW;
        NB2_UNROLL
        for (int i = 0; i < I2; ++i) {
            const int k1 = r + i * R3;
            NB2_UNROLL
            for (int n3 = 0; n3 < R3; ++n3) R.x[i * R3 + n3] = X[xslot(k1 * MM + n3) * HW + ch];
        }
    }

    // ---- stage 6: radix-R3 step and the stores
    NB2_HD static void step2(const Params& p, long long tile, int tid, Regs& R) {
        const int c = tid & (W - 1), r = tid >> LOGW;
        long long o, tc;
        const long long base = tile_base(p, tile, o, tc);
        NB2_UNROLL
        for (int i = 0; i < I2; ++i) {
            const int k1 = r + i * R3;
            cplx<T>(&y)[R3] = reinterpret_cast<cplx<T>(&)[R3]>(R.x[i * R3]);
            fft_reg<R3, true>(y);
            if (p.out_rowoff) {
                const long long oo = p.out_ooff ? ld_ro(p.out_ooff + o) : o * p.out_ostride;
                const long long cc =
                    p.out_coloff ? ld_ro(p.out_coloff + (tc >> p.g.logM)) + (tc & ((1ll << p.g.logM) - 1)) : tc;
                cplx<T>* dst = p.out + oo + cc + c;
                const long long* ro = p.out_rowoff + k1;
                NB2_UNROLL
                for (int k3 = 0; k3 < R3; ++k3) stc_stream(dst + ld_ro(ro + k3 * R1), y[brev<R3>(k3)]);
            } else {
                cplx<T>* dst = p.out + base + c + (long long)k1 * p.inner;
                NB2_UNROLL
                for (int k3 = 0; k3 < R3; ++k3) stc_stream(dst + (long long)k3 * R1 * p.inner, y[brev<R3>(k3)]);
            }
        }
    }
};

#if !defined(NB2_EMU)
// ------------------------------------------------------------------ device side
__device__ __forceinline__ unsigned smem_u32(const void* p) { return (unsigned)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(void* bar, unsigned count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_expect_tx(void* bar, unsigned bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait(void* bar, unsigned parity) {
    asm volatile(
        "{\n\t"
        ".reg .pred P1;\n\t"
        "NB2_WAIT_%=:\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 P1, [%0], %1;\n\t"
        "@P1 bra NB2_DONE_%=;\n\t"
        "bra NB2_WAIT_%=;\n\t"
        "NB2_DONE_%=:\n\t"
        "}" ::"r"(smem_u32(bar)), "r"(parity)
        : "memory");
}
__device__ __forceinline__ void tma_load_3d(void* dst, const CUtensorMap* map, void* bar, int c0, int c1, int c2) {
    asm volatile(
        "cp.async.bulk.tensor.3d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, %5}], [%2];" ::"r"(
            smem_u32(dst)),
        "l"(reinterpret_cast<unsigned long long>(map)), "r"(smem_u32(bar)), "r"(c0), "r"(c1), "r"(c2)
        : "memory");
}

template <class K>
__global__ void __launch_bounds__(K::NTHREADS, K::MIN_BLOCKS)
    nb2_kernel_tma(const __grid_constant__ CUtensorMap map, const __grid_constant__ CUtensorMap mapf,
                   const __grid_constant__ CUtensorMap mapft, const typename K::Params p, long long ntiles,
                   const int* gate) {
    extern __shared__ __align__(128) char nb2_smem[];
    if (gate != nullptr && *gate != 0) return;   // device-resident CG: the solver has stopped (nb2_launch.cuh)
    char* P = nb2_smem;
    char* X = nb2_smem + K::P_BYTES;
    void* bar = nb2_smem + K::P_BYTES + K::X_BYTES;
    void* barf = nb2_smem + K::P_BYTES + K::X_BYTES + 8;
    const int tid = (int)threadIdx.x;
    const bool lower = (tid & (K::W - 1)) < K::HW;
    auto issue = [&](long long tile) {
        const long long o = tile / p.tiles_per_outer;
        const int c0 = int((tile - o * p.tiles_per_outer) * K::BOX_ELEMS);
        mbar_expect_tx(bar, (unsigned)K::P_BYTES);
#pragma unroll
        for (int b = 0; b < K::NBOX; ++b)
            tma_load_3d(P + size_t(b) * K::BOX_ROWS * K::ROW_BYTES, &map, bar, c0, b * K::BOX_ROWS, (int)o);
    };
    // rows 0 .. N/2 of the amplitude-field tile into X (fx)
    auto issue_f = [&](long long tile) {
        const long long o = tile / p.tiles_per_outer;
        const int c0 = int((tile - o * p.tiles_per_outer) * K::BOX_ELEMS);
        const int of = int(o & p.f_outer_mask);
        mbar_expect_tx(barf, (unsigned)K::F_BYTES);
#pragma unroll
        for (int b = 0; b < K::NFBOX; ++b)
            tma_load_3d(X + size_t(b) * K::FBOX_ROWS * K::ROW_BYTES, &mapf, barf, c0, b * K::FBOX_ROWS, of);
        tma_load_3d(X + size_t(K::HALF) * K::ROW_BYTES, &mapft, barf, c0, K::HALF, of);
    };
    if (tid == 0) {
        mbar_init(bar, 1);
        mbar_init(barf, 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncthreads();
    long long t = blockIdx.x;
    if (tid == 0 && t < ntiles) {
        issue(t);
        if (p.fx) issue_f(t);
    }
    unsigned parity = 0;
    for (; t < ntiles; t += gridDim.x) {
        typename K::Regs R;
        K::pro_begin(p, t, tid, R);
        mbar_wait(bar, parity);
        if (p.fx) mbar_wait(barf, parity);
        parity ^= 1u;
        K::load_tile(p, t, tid, P, X, R);
        __syncthreads();   // P and X are free (the previous tile's readers of X are done as well)
        if (tid == 0 && t + gridDim.x < ntiles) {
            asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
            issue(t + gridDim.x);
        }
        K::step1(p, t, tid, R);
        if (lower) K::xwrite(tid, X, R);
        __syncthreads();
        if (lower) K::xread(tid, X, R);
        __syncthreads();
        if (!lower) K::xwrite(tid, X, R);
        __syncthreads();
        if (!lower) K::xread(tid, X, R);
        if (p.fx) {
            __syncthreads();   // X is free until the next tile's transposition: fetch its field rows now
            if (tid == 0 && t + gridDim.x < ntiles) {
                asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
                issue_f(t + gridDim.x);
            }
        }
        K::step2(p, t, tid, R);
    }
}

typedef CUresult (*nb2_encode_tiled_fn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                        const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                        CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
inline nb2_encode_tiled_fn tma_encoder() {
    static nb2_encode_tiled_fn fn = [] {
        void* f = nullptr;
        cudaDriverEntryPointQueryResult q;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &f, cudaEnableDefault, &q) != cudaSuccess ||
            q != cudaDriverEntryPointSuccess)
            f = nullptr;
        return (nb2_encode_tiled_fn)f;
    }();
    return fn;
}

// 3-D tensor map over the scalars of a complex array: (2 * inner, N, outer), box (128 bytes, box_rows, 1)
template <class K, typename T>
inline int tma_make_map(CUtensorMap* map, const void* in, long long inner, long long outer, int box_rows) {
    // the CG loop re-launches the same few (buffer, geometry) pairs every iteration: keep the encoded maps
    struct Entry {
        const void* in;
        long long inner, outer;
        int box_rows;
        CUtensorMap map;
    };
    constexpr int kEntries = 8;
    static thread_local Entry cache[kEntries] = {};
    static thread_local int next = 0;
    for (int i = 0; i < kEntries; ++i)
        if (cache[i].in == in && cache[i].inner == inner && cache[i].outer == outer && cache[i].box_rows == box_rows) {
            *map = cache[i].map;
            return 0;
        }
    nb2_encode_tiled_fn enc = tma_encoder();
    if (!enc) {
        last_error() = "cuTensorMapEncodeTiled is not available";
        return -2;
    }
    const cuuint64_t dims[3] = {cuuint64_t(2 * inner), cuuint64_t(K::N), cuuint64_t(outer)};
    const cuuint64_t strides[2] = {cuuint64_t(2 * inner) * sizeof(T), cuuint64_t(2 * inner) * sizeof(T) * K::N};
    const cuuint32_t box[3] = {cuuint32_t(K::BOX_ELEMS), cuuint32_t(box_rows), 1};
    const cuuint32_t estr[3] = {1, 1, 1};
    const CUresult r = enc(map, sizeof(T) == 4 ? CU_TENSOR_MAP_DATA_TYPE_FLOAT32 : CU_TENSOR_MAP_DATA_TYPE_FLOAT64, 3,
                           const_cast<void*>(in), dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                           CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_NONE, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) {
        last_error() = "cuTensorMapEncodeTiled failed (code " + std::to_string((int)r) + ")";
        return -2;
    }
    cache[next] = Entry{in, inner, outer, box_rows, *map};
    next = (next + 1) % kEntries;
    return 0;
}

template <class K, typename T>
inline int rt_launch_tma(long long ntiles, long long outer, rt_stream stream, const typename K::Params& p) {
    if (ntiles <= 0) return 0;
    CUtensorMap map, mapf, mapft;
    if (tma_make_map<K, T>(&map, p.in, p.inner, outer, K::BOX_ROWS)) return -2;
    if (p.fx) {
        if (tma_make_map<K, T>(&mapf, p.pro.amp, p.inner, p.f_outer_mask + 1, K::FBOX_ROWS)) return -2;
        if (tma_make_map<K, T>(&mapft, p.pro.amp, p.inner, p.f_outer_mask + 1, K::FTAIL_ROWS)) return -2;
    } else {
        mapf = mapft = map;
    }
    static int per_sm_dev[kMaxDevices] = {0};   // per kernel instantiation and device; 0 = not configured yet
    int& per_sm = per_sm_dev[rt_device() % kMaxDevices];
    if (per_sm == 0) {
        NB2_CUDA_OK(cudaFuncSetAttribute(nb2_kernel_tma<K>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)K::SMEM));
        int v = 0;
        NB2_CUDA_OK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&v, nb2_kernel_tma<K>, K::NTHREADS, K::SMEM));
        per_sm = v < 1 ? 1 : v;
    }
    long long grid = (long long)per_sm * rt_sm_count();
    if (grid > ntiles) grid = ntiles;
    nb2_kernel_tma<K><<<(unsigned)grid, K::NTHREADS, K::SMEM, stream>>>(map, mapf, mapft, p, ntiles, launch_gate());
    NB2_CUDA_OK(cudaGetLastError());
    ++launch_counter();
    ++tma_launch_counter();
    return 0;
}
inline bool tma_available() { return tma_encoder() != nullptr; }

#else
// ------------------------------------------------------------------ sequential emulator (tests only)
// The TMA copy is a plain row copy into the dense buffer P; the registers that live across the
// barriers of a tile are kept per emulated thread.
template <class K, typename T>
inline int rt_launch_tma(long long ntiles, long long, rt_stream, const typename K::Params& p) {
    if (launch_gate() != nullptr && *launch_gate() != 0) {
        ++launch_counter();
        return 0;
    }
    std::vector<char> Pbuf(K::P_BYTES), Xbuf(K::X_BYTES);
    std::vector<typename K::Regs> regs(K::NTHREADS);
    for (long long tile = 0; tile < ntiles; ++tile) {
        std::memset(Pbuf.data(), 0xAB, K::P_BYTES);
        std::memset(Xbuf.data(), 0xAB, K::X_BYTES);
        long long o, tc;
        const long long base = K::tile_base(p, tile, o, tc);
        for (int row = 0; row < K::N; ++row)
            std::memcpy(Pbuf.data() + size_t(row) * K::ROW_BYTES, p.in + base + (long long)row * p.inner, K::ROW_BYTES);
        if (p.fx) {   // what the field TMA does: rows 0 .. N/2 of the field tile, dense, into X
            const long long itemmask = (1ll << (p.g.logN0 + p.g.logN1 + p.g.logM)) - 1;
            const cplx<T>* f = reinterpret_cast<const cplx<T>*>(p.pro.amp) + (base & itemmask);
            for (int row = 0; row <= K::HALF; ++row)
                std::memcpy(Xbuf.data() + size_t(row) * K::ROW_BYTES, f + (long long)row * p.inner, K::ROW_BYTES);
        }
        auto lower = [](int tid) { return (tid & (K::W - 1)) < K::HW; };
        for (int tid = 0; tid < K::NTHREADS; ++tid) {
            K::pro_begin(p, tile, tid, regs[tid]);
            K::load_tile(p, tile, tid, Pbuf.data(), Xbuf.data(), regs[tid]);
        }
        std::memset(Xbuf.data(), 0xAB, K::X_BYTES);
        for (int tid = 0; tid < K::NTHREADS; ++tid) {
            K::step1(p, tile, tid, regs[tid]);
            if (lower(tid)) K::xwrite(tid, Xbuf.data(), regs[tid]);
        }
        for (int tid = 0; tid < K::NTHREADS; ++tid)
            if (lower(tid)) K::xread(tid, Xbuf.data(), regs[tid]);
        std::memset(Xbuf.data(), 0xAB, K::X_BYTES);
        for (int tid = 0; tid < K::NTHREADS; ++tid)
            if (!lower(tid)) K::xwrite(tid, Xbuf.data(), regs[tid]);
        for (int tid = 0; tid < K::NTHREADS; ++tid) {
            if (!lower(tid)) K::xread(tid, Xbuf.data(), regs[tid]);
            K::step2(p, tile, tid, regs[tid]);
        }
    }
    ++launch_counter();
    ++tma_launch_counter();
    return 0;
}
inline bool tma_available() { return true; }
#endif

}  // namespace nb2
