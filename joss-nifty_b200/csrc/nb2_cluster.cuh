// niftyb200 -- StridedCluster: the strided FFT pass for axes whose tile does not fit one SM.
//
// A tile of N = CL * NL rows x W complex columns (W * sizeof(cplx) = 128 bytes per row segment) is held by a
// thread-block CLUSTER of CL CTAs: CTA j fetches rows [j NL, (j+1) NL) with TMA into its own shared memory
// (NL x 128 bytes = 128 KB for NL = 1024 fp32).  The length-N transform is split by decimation in frequency,
//
//   X[CL k' + q] = sum_i  W_NL^(i k')  *  W_N^(i q)  *  sum_j W_CL^(j q) x[j NL + i] ,
//
// so CTA q forms  z_q[i] = W_N^(i q) * sum_j W_CL^(j q) x_j[i]  by reading row i of EVERY CTA of the cluster
// through distributed shared memory, then runs the NL-point transform of StridedTma on it (two register-radix
// steps with the transposition through the half-size buffer X) and stores the rows q, q + CL, q + 2 CL, ...
// Rows of 2048 (CL = 2) and 4096 (CL = 4) points keep 128-byte row segments this way; the one-SM StridedPass
// has to fall back to 64- and 32-byte segments for them (128 KB of shared memory per tile).
//
// Point-wise prologues (amplitude field, scale, fused direction update d <- r + beta d) are applied by each CTA
// to ITS OWN rows in shared memory before the cluster barrier, so no operand is read twice.
//
// Per tile: wait for the own TMA -> local prologue -> cluster barrier (every CTA's rows are final) ->
// cross-CTA butterfly into registers -> cluster barrier (every CTA is done reading) -> the next tile's TMA is
// issued and overlaps the butterflies, the transposition and the stores of this tile.
#pragma once

#include "nb2_tma.cuh"

#if !defined(NB2_EMU)
#include <cooperative_groups.h>
#endif

namespace nb2 {

template <typename T, int CL, int R1, int R3, int LW = (sizeof(T) == 4 ? 4 : 3)>
struct StridedCluster {
    static_assert(CL == 2 || CL == 4, "clusters of 2 or 4 CTAs");
    static_assert(R1 >= R3 && R3 > 1, "two-step local plans with R1 >= R3");
    static constexpr int NL = R1 * R3;        // rows per CTA
    static constexpr int N = CL * NL;         // transform length
    static constexpr int MM = R3;
    static constexpr int LOGW = LW;
    static constexpr int W = 1 << LOGW;
    static constexpr int HW = W / 2;
    static constexpr int NTHREADS = W * R3;
    static constexpr int I2 = R1 / R3;
    static constexpr int BOX_ROWS = NL < 256 ? NL : 256;
    static constexpr int NBOX = NL / BOX_ROWS;
    static constexpr int ROW_BYTES = W * int(sizeof(cplx<T>));
    static constexpr int BOX_ELEMS = ROW_BYTES / int(sizeof(T));
    static constexpr size_t P_BYTES = size_t(NL) * ROW_BYTES;
    static constexpr int XROWS = NL + R1;
    static constexpr size_t X_BYTES = size_t(XROWS) * (ROW_BYTES / 2);
    static constexpr size_t SMEM = P_BYTES + X_BYTES + 16;
    static constexpr int PB = R1 >= 8 ? 8 : R1;
    static constexpr int NBATCH = R1 / PB;

    struct Params {
        const cplx<T>* in;
        cplx<T>* out;
        const cplx<T>* tw;          // tw[j] = exp(-2 pi i j / N), j in [0, N)
        long long inner;            // complex elements between consecutive rows
        long long tiles_per_outer;  // inner / W
        int axis;
        int has_pro;
        PointOp<T> pro;
        Geom g;
    };
    struct Regs {
        cplx<T> x[R1];
    };

    NB2_HD static long long tile_base(const Params& p, long long tile, long long& o, long long& tc) {
        o = tile / p.tiles_per_outer;
        tc = (tile - o * p.tiles_per_outer) << LOGW;
        return o * N * p.inner + tc;
    }
    NB2_HD static int xslot(int L) { return L + L / MM; }

    // ---- local prologue: CTA `rank` applies the point-wise operators to its own NL rows, in place in P
    NB2_HD static void prologue(const Params& p, long long tile, int rank, int tid, char* Praw) {
        if (!p.has_pro) return;
        cplx<T>* P = reinterpret_cast<cplx<T>*>(Praw);
        const int c = tid & (W - 1), r = tid >> LOGW;
        long long o, tc;
        const long long base = tile_base(p, tile, o, tc);
        const long long row0 = (long long)rank * NL;
        const long long e0 = base + c + (row0 + r) * p.inner, rstride = (long long)MM * p.inner;
        const bool upd = p.pro.pre_add != nullptr;
        const cplx<T>* rs = upd ? reinterpret_cast<const cplx<T>*>(p.pro.pre_add) + e0 : nullptr;
        cplx<T>* dd = (upd && p.pro.pre_out) ? reinterpret_cast<cplx<T>*>(p.pro.pre_out) + e0 : nullptr;
        const T beta = upd ? pre_scale_of(p.pro) : T(0);
        const bool field = p.pro.amp_mode == 2;
        const long long itemmask = (1ll << (p.g.logN0 + p.g.logN1 + p.g.logM)) - 1;
        const FieldRows<T> fr(reinterpret_cast<const cplx<T>*>(p.pro.amp) + ((base + c) & itemmask), p.inner, N,
                              p.pro.amp_sym && !(p.axis == 0 ? p.g.shift0 : p.g.shift1));
        NB2_UNROLL
        for (int g = 0; g < NBATCH; ++g) {
            cplx<T> rr[PB], ff[PB];
            NB2_UNROLL
            for (int i = 0; i < PB; ++i) {
                const int n1 = g * PB + i;
                if (upd) rr[i] = ldc_stream(rs + n1 * rstride);
                if (field) ff[i] = fr.load(int(row0) + r + n1 * MM);
            }
            NB2_UNROLL
            for (int i = 0; i < PB; ++i) {
                const int n1 = g * PB + i, row = r + n1 * MM;
                cplx<T> v = P[(row << LOGW) + c];
                if (upd) {
                    v.re = rr[i].re + beta * v.re;
                    v.im = rr[i].im + beta * v.im;
                    if (dd) stc_stream(dd + n1 * rstride, v);
                }
                if (field) {
                    v.re *= ff[i].re * p.pro.scale;
                    v.im *= ff[i].im * p.pro.scale;
                } else {
                    v = v * p.pro.scale;
                }
                P[(row << LOGW) + c] = v;
            }
        }
    }

    // ---- cross-CTA butterfly: z_rank[i] = W_N^(i rank) sum_j W_CL^(j rank) x_j[i], rows i = r + n1 R3
    NB2_HD static void load_tile(const Params& p, int rank, int tid, const char* const* Ps, Regs& R) {
        const int c = tid & (W - 1), r = tid >> LOGW;
        NB2_UNROLL
        for (int n1 = 0; n1 < R1; ++n1) {
            const int i = r + n1 * MM;
            cplx<T> v[CL];
            NB2_UNROLL
            for (int j = 0; j < CL; ++j) v[j] = reinterpret_cast<const cplx<T>*>(Ps[j])[(i << LOGW) + c];
            cplx<T> y;
            if (CL == 2) {
                y = rank == 0 ? v[0] + v[1] : v[0] - v[1];
            } else {
                // W_4 = -i:  q = 0: a + b + c + d;  1: (a - c) - i (b - d);  2: a - b + c - d;  3: (a - c) + i (b - d)
                const cplx<T> s02 = (rank & 1) ? v[0] - v[2 % CL] : v[0] + v[2 % CL];
                const cplx<T> s13 = (rank & 1) ? v[1] - v[3 % CL] : v[1] + v[3 % CL];
                if (rank == 0) y = s02 + s13;
                else if (rank == 2) y = s02 - s13;
                else if (rank == 1) y = s02 + mul_mi(s13);
                else y = s02 - mul_mi(s13);
            }
            if (rank != 0) y = y * ldc_ro(p.tw + (long long)i * rank);
            R.x[n1] = y;
        }
    }

    // ---- local NL-point transform: radix-R1 step + inter-step twiddles (W_NL = W_N^CL)
    NB2_HD static void step1(const Params& p, int tid, Regs& R) {
        const int r = tid >> LOGW;
        cplx<T>(&x)[R1] = R.x;
        fft_reg<R1, true>(x);
        cplx<T> wp[R1];
        tw_powers<R1>(ldc_ro(p.tw + CL * r), wp);
        NB2_UNROLL
        for (int k1 = 1; k1 < R1; ++k1) x[brev<R1>(k1)] = x[brev<R1>(k1)] * wp[k1];
    }
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
    // ---- radix-R3 step; local output index k' = k1 + k3 R1 is row CL k' + rank of the transform
    NB2_HD static void step2(const Params& p, long long tile, int rank, int tid, Regs& R) {
        const int c = tid & (W - 1), r = tid >> LOGW;
        long long o, tc;
        const long long base = tile_base(p, tile, o, tc);
        NB2_UNROLL
        for (int i = 0; i < I2; ++i) {
            const int k1 = r + i * R3;
            cplx<T>(&y)[R3] = reinterpret_cast<cplx<T>(&)[R3]>(R.x[i * R3]);
            fft_reg<R3, true>(y);
            cplx<T>* dst = p.out + base + c + ((long long)CL * k1 + rank) * p.inner;
            NB2_UNROLL
            for (int k3 = 0; k3 < R3; ++k3) stc_stream(dst + (long long)CL * k3 * R1 * p.inner, y[brev<R3>(k3)]);
        }
    }
};

#if !defined(NB2_EMU)
// ------------------------------------------------------------------ device side
template <class K, int CL>
__global__ void __cluster_dims__(CL, 1, 1) __launch_bounds__(K::NTHREADS, 1)
    nb2_kernel_cluster(const __grid_constant__ CUtensorMap map, const typename K::Params p, long long ntiles,
                       const int* gate) {
    namespace cg = cooperative_groups;
    extern __shared__ __align__(128) char nb2_smem[];
    if (gate != nullptr && *gate != 0) return;   // device-resident CG: every CTA of every cluster takes this exit
    cg::cluster_group cluster = cg::this_cluster();
    const int rank = (int)cluster.block_rank();
    const long long nclusters = gridDim.x / CL;
    char* P = nb2_smem;
    char* X = nb2_smem + K::P_BYTES;
    void* bar = nb2_smem + K::P_BYTES + K::X_BYTES;
    const int tid = (int)threadIdx.x;
    const bool lower = (tid & (K::W - 1)) < K::HW;
    const char* Ps[CL];
#pragma unroll
    for (int j = 0; j < CL; ++j) Ps[j] = cluster.map_shared_rank(P, j);
    auto issue = [&](long long tile) {
        const long long o = tile / p.tiles_per_outer;
        const int c0 = int((tile - o * p.tiles_per_outer) * K::BOX_ELEMS);
        mbar_expect_tx(bar, (unsigned)K::P_BYTES);
#pragma unroll
        for (int b = 0; b < K::NBOX; ++b)
            tma_load_3d(P + size_t(b) * K::BOX_ROWS * K::ROW_BYTES, &map, bar, c0, rank * K::NL + b * K::BOX_ROWS, (int)o);
    };
    if (tid == 0) {
        mbar_init(bar, 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncthreads();
    long long t = blockIdx.x / CL;
    if (tid == 0 && t < ntiles) issue(t);
    unsigned parity = 0;
    for (; t < ntiles; t += nclusters) {
        typename K::Regs R;
        mbar_wait(bar, parity);
        parity ^= 1u;
        K::prologue(p, t, rank, tid, P);
        cluster.sync();                 // the rows of every CTA of the cluster have landed (and are scaled)
        K::load_tile(p, rank, tid, Ps, R);
        cluster.sync();                 // every CTA is done reading every P: refill
        if (tid == 0 && t + nclusters < ntiles) {
            asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
            issue(t + nclusters);
        }
        K::step1(p, tid, R);
        if (lower) K::xwrite(tid, X, R);
        __syncthreads();
        if (lower) K::xread(tid, X, R);
        __syncthreads();
        if (!lower) K::xwrite(tid, X, R);
        __syncthreads();
        if (!lower) K::xread(tid, X, R);
        K::step2(p, t, rank, tid, R);
        __syncthreads();                // X is reused by the next tile's lower half
    }
}

template <class K, typename T, int CL>
inline int rt_launch_cluster(long long ntiles, long long outer, rt_stream stream, const typename K::Params& p) {
    if (ntiles <= 0) return 0;
    CUtensorMap map;
    if (tma_make_map<K, T>(&map, p.in, p.inner, outer, K::BOX_ROWS)) return -2;
    static int nclusters_dev[kMaxDevices] = {0};   // per kernel instantiation and device; 0 = not configured yet
    int& ncl = nclusters_dev[rt_device() % kMaxDevices];
    if (ncl == 0) {
        NB2_CUDA_OK(cudaFuncSetAttribute(nb2_kernel_cluster<K, CL>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                         (int)K::SMEM));
        cudaLaunchConfig_t cfg = {};
        cfg.gridDim = dim3((unsigned)(CL * rt_sm_count()), 1, 1);
        cfg.blockDim = dim3(K::NTHREADS, 1, 1);
        cfg.dynamicSmemBytes = K::SMEM;
        cudaLaunchAttribute attr;
        attr.id = cudaLaunchAttributeClusterDimension;
        attr.val.clusterDim.x = CL;
        attr.val.clusterDim.y = attr.val.clusterDim.z = 1;
        cfg.attrs = &attr;
        cfg.numAttrs = 1;
        int v = 0;
        NB2_CUDA_OK(cudaOccupancyMaxActiveClusters(&v, nb2_kernel_cluster<K, CL>, &cfg));
        ncl = v < 1 ? 1 : v;
    }
    long long n = ncl;
    if (n > ntiles) n = ntiles;
    nb2_kernel_cluster<K, CL><<<(unsigned)(n * CL), K::NTHREADS, K::SMEM, stream>>>(map, p, ntiles, launch_gate());
    NB2_CUDA_OK(cudaGetLastError());
    ++launch_counter();
    ++tma_launch_counter();
    return 0;
}

#else
// ------------------------------------------------------------------ sequential emulator (tests only)
template <class K, typename T, int CL>
inline int rt_launch_cluster(long long ntiles, long long, rt_stream, const typename K::Params& p) {
    if (launch_gate() != nullptr && *launch_gate() != 0) {
        ++launch_counter();
        return 0;
    }
    std::vector<std::vector<char>> Pb(CL, std::vector<char>(K::P_BYTES));
    std::vector<char> Xbuf(K::X_BYTES);
    std::vector<typename K::Regs> regs(K::NTHREADS);
    const char* Ps[CL];
    for (int j = 0; j < CL; ++j) Ps[j] = Pb[j].data();
    auto lower = [](int tid) { return (tid & (K::W - 1)) < K::HW; };
    for (long long tile = 0; tile < ntiles; ++tile) {
        long long o, tc;
        const long long base = K::tile_base(p, tile, o, tc);
        for (int j = 0; j < CL; ++j) {       // what each CTA's TMA does, then its local prologue
            for (int row = 0; row < K::NL; ++row)
                std::memcpy(Pb[j].data() + size_t(row) * K::ROW_BYTES,
                            p.in + base + ((long long)j * K::NL + row) * p.inner, K::ROW_BYTES);
            for (int tid = 0; tid < K::NTHREADS; ++tid) K::prologue(p, tile, j, tid, Pb[j].data());
        }
        for (int rank = 0; rank < CL; ++rank) {
            for (int tid = 0; tid < K::NTHREADS; ++tid) {
                K::load_tile(p, rank, tid, Ps, regs[tid]);
                K::step1(p, tid, regs[tid]);
            }
            std::memset(Xbuf.data(), 0xAB, K::X_BYTES);
            for (int tid = 0; tid < K::NTHREADS; ++tid)
                if (lower(tid)) K::xwrite(tid, Xbuf.data(), regs[tid]);
            for (int tid = 0; tid < K::NTHREADS; ++tid)
                if (lower(tid)) K::xread(tid, Xbuf.data(), regs[tid]);
            std::memset(Xbuf.data(), 0xAB, K::X_BYTES);
            for (int tid = 0; tid < K::NTHREADS; ++tid)
                if (!lower(tid)) K::xwrite(tid, Xbuf.data(), regs[tid]);
            for (int tid = 0; tid < K::NTHREADS; ++tid) {
                if (!lower(tid)) K::xread(tid, Xbuf.data(), regs[tid]);
                K::step2(p, tile, rank, tid, regs[tid]);
            }
        }
    }
    ++launch_counter();
    ++tma_launch_counter();
    return 0;
}
#endif

}  // namespace nb2
