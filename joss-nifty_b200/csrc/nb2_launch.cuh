// niftyb200 -- kernel launcher and the tiny runtime shim.
//
// Device build (nvcc): every kernel struct K (Params, NPHASES, enabled(ph), phase<PH>(),
// smem_bytes) is run by `nb2_kernel<K>`: its phases in order with __syncthreads() between
// them.  Emulator build (-DNB2_EMU, host compiler, TEST INFRASTRUCTURE ONLY -- see
// tests/emu): the same phases are executed sequentially for every (block, thread).
#pragma once

#include <algorithm>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <vector>

#include "nb2_common.cuh"

#if !defined(NB2_EMU)
#include <cuda_runtime.h>
#endif

namespace nb2 {

inline std::string& last_error() {
    static thread_local std::string e;
    return e;
}
// number of kernels launched by this library so far (bench.py reports it as gpu_launches)
inline long long& launch_counter() {
    static long long n = 0;
    return n;
}
inline long long& tma_launch_counter() {
    static long long n = 0;
    return n;
}

#if !defined(NB2_EMU)
// ------------------------------------------------------------------ device runtime
#define NB2_CUDA_OK(call)                                                                      \
    do {                                                                                       \
        cudaError_t err__ = (call);                                                            \
        if (err__ != cudaSuccess) {                                                            \
            nb2::last_error() = std::string(#call) + ": " + cudaGetErrorString(err__);        \
            return -2;                                                                         \
        }                                                                                      \
    } while (0)

typedef cudaStream_t rt_stream;
inline int rt_sm_count();

template <class K, int PH, bool IN_RANGE = (PH < K::NPHASES)>
struct PhaseStep {
    __device__ __forceinline__ static void run(const typename K::Params& p, char* smem, bool first) {
        bool f = first;
        if (K::enabled(PH)) {
            if (!f) __syncthreads();
            K::template phase<PH>(p, (long long)blockIdx.x, (int)threadIdx.x, (int)blockDim.x, smem);
            f = false;
        }
        PhaseStep<K, PH + 1>::run(p, smem, f);
    }
};
template <class K, int PH>
struct PhaseStep<K, PH, false> {
    __device__ __forceinline__ static void run(const typename K::Params&, char*, bool) {}
};

template <class K>
__global__ void __launch_bounds__(K::MAX_THREADS, K::MIN_BLOCKS) nb2_kernel(const typename K::Params p) {
    extern __shared__ __align__(16) char nb2_smem[];
    PhaseStep<K, 0>::run(p, nb2_smem, true);
}

// Persistent, software-pipelined variant: each CTA walks the tiles blockIdx.x, blockIdx.x + gridDim.x,
// ... ; K::prefetch issues the asynchronous copies (cp.async) of a tile's input into one of
// STAGES shared-memory stage buffers, STAGES-1 tiles ahead of the tile whose phases are running.
template <class K, int PH, bool IN_RANGE = (PH < K::NPHASES)>
struct PipeStep {
    __device__ __forceinline__ static void run(const typename K::Params& p, long long tile, char* smem) {
        if (K::enabled(PH)) {
            if (PH != 0) __syncthreads();
            K::template phase<PH>(p, tile, (int)threadIdx.x, (int)blockDim.x, smem);
        }
        PipeStep<K, PH + 1>::run(p, tile, smem);
    }
};
template <class K, int PH>
struct PipeStep<K, PH, false> {
    __device__ __forceinline__ static void run(const typename K::Params&, long long, char*) {}
};

template <class K, int STAGES>
__global__ void __launch_bounds__(K::MAX_THREADS, K::MIN_BLOCKS)
    nb2_kernel_pipe(const typename K::Params p, long long ntiles, unsigned stage_bytes) {
    extern __shared__ __align__(16) char nb2_smem[];
    const int tid = (int)threadIdx.x, nt = (int)blockDim.x;
    long long t = blockIdx.x;
    NB2_UNROLL
    for (int s = 0; s < STAGES - 1; ++s) {
        const long long tt = t + (long long)s * gridDim.x;
        if (tt < ntiles) K::prefetch(p, tt, tid, nt, nb2_smem + (size_t)s * stage_bytes);
        cp_async_commit();
    }
    int st = 0;
    for (; t < ntiles; t += gridDim.x) {
        const long long tn = t + (long long)(STAGES - 1) * gridDim.x;
        const int sn = st == 0 ? STAGES - 1 : st - 1;   // the stage the previous tile has just left
        if (tn < ntiles) K::prefetch(p, tn, tid, nt, nb2_smem + (size_t)sn * stage_bytes);
        cp_async_commit();
        cp_async_wait<STAGES - 1>();
        __syncthreads();
        PipeStep<K, 0>::run(p, t, nb2_smem + (size_t)st * stage_bytes);
        __syncthreads();
        st = st + 1 == STAGES ? 0 : st + 1;
    }
}

template <class K, int STAGES>
inline int rt_launch_pipe_n(long long ntiles, long long grid, int nthreads, size_t stage_bytes, rt_stream stream,
                            const typename K::Params& p) {
    const size_t smem = stage_bytes * STAGES;
    static thread_local size_t configured = 0;  // per kernel instantiation
    if (smem > 48 * 1024 && smem > configured) {
        NB2_CUDA_OK(cudaFuncSetAttribute(nb2_kernel_pipe<K, STAGES>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        configured = smem;
    }
    if (grid <= 0) {  // as many CTAs as can be resident at once
        static thread_local int per_sm = 0, per_sm_threads = 0;
        static thread_local size_t per_sm_smem = 0;
        if (per_sm == 0 || per_sm_threads != nthreads || per_sm_smem != smem) {
            NB2_CUDA_OK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, nb2_kernel_pipe<K, STAGES>, nthreads, smem));
            per_sm_threads = nthreads;
            per_sm_smem = smem;
            if (per_sm < 1) per_sm = 1;
        }
        int cap = per_sm;
        if (const char* e = std::getenv("NB2_PIPE_CTAS")) cap = std::max(1, std::min(per_sm, std::atoi(e)));
        grid = (long long)cap * rt_sm_count();
    }
    if (grid > ntiles) grid = ntiles;
    nb2_kernel_pipe<K, STAGES><<<(unsigned)grid, nthreads, smem, stream>>>(p, ntiles, (unsigned)stage_bytes);
    NB2_CUDA_OK(cudaGetLastError());
    ++launch_counter();
    return 0;
}
template <class K>
inline int rt_launch_pipe(long long ntiles, long long grid, int nthreads, size_t stage_bytes, int stages, rt_stream stream,
                          const typename K::Params& p) {
    if (ntiles <= 0) return 0;
    if (stage_bytes % 16) {
        last_error() = "stage size must be a multiple of 16 bytes";
        return -1;
    }
    switch (stages) {
        case 2: return rt_launch_pipe_n<K, 2>(ntiles, grid, nthreads, stage_bytes, stream, p);
        case 3: return rt_launch_pipe_n<K, 3>(ntiles, grid, nthreads, stage_bytes, stream, p);
        default: last_error() = "unsupported number of pipeline stages"; return -1;
    }
}

template <class K>
inline int rt_launch(long long grid, int nthreads, size_t smem, rt_stream stream, const typename K::Params& p) {
    if (grid <= 0) return 0;
    if (grid > 2147483647ll) {
        last_error() = "grid too large";
        return -1;
    }
    if (smem > 48 * 1024) {
        static thread_local size_t configured = 0;  // per kernel instantiation
        if (smem > configured) {
            NB2_CUDA_OK(cudaFuncSetAttribute(nb2_kernel<K>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
            configured = smem;
        }
    }
    nb2_kernel<K><<<(unsigned)grid, nthreads, smem, stream>>>(p);
    NB2_CUDA_OK(cudaGetLastError());
    ++launch_counter();
    return 0;
}

inline int rt_malloc(void** p, size_t bytes) {
    NB2_CUDA_OK(cudaMalloc(p, bytes ? bytes : 1));
    return 0;
}
inline int rt_free(void* p) {
    if (p) NB2_CUDA_OK(cudaFree(p));
    return 0;
}
inline int rt_h2d(void* dst, const void* src, size_t bytes, rt_stream s) {
    NB2_CUDA_OK(cudaMemcpyAsync(dst, src, bytes, cudaMemcpyHostToDevice, s));
    NB2_CUDA_OK(cudaStreamSynchronize(s));
    return 0;
}
inline int rt_d2h(void* dst, const void* src, size_t bytes, rt_stream s) {
    NB2_CUDA_OK(cudaMemcpyAsync(dst, src, bytes, cudaMemcpyDeviceToHost, s));
    NB2_CUDA_OK(cudaStreamSynchronize(s));
    return 0;
}
inline int rt_d2d(void* dst, const void* src, size_t bytes, rt_stream s) {
    NB2_CUDA_OK(cudaMemcpyAsync(dst, src, bytes, cudaMemcpyDeviceToDevice, s));
    return 0;
}
inline int rt_sync(rt_stream s) {
    NB2_CUDA_OK(cudaStreamSynchronize(s));
    return 0;
}
// auxiliary streams + fork/join events (the L2-blocked tail runs its plane groups on several streams
// so that the tail of one group's kernel overlaps the head of the next group's)
typedef cudaEvent_t rt_event;
inline int rt_stream_create(rt_stream* s) {
    NB2_CUDA_OK(cudaStreamCreateWithFlags(s, cudaStreamNonBlocking));
    return 0;
}
inline int rt_stream_destroy(rt_stream s) {
    if (s) NB2_CUDA_OK(cudaStreamDestroy(s));
    return 0;
}
inline int rt_event_create(rt_event* e) {
    NB2_CUDA_OK(cudaEventCreateWithFlags(e, cudaEventDisableTiming));
    return 0;
}
inline int rt_event_destroy(rt_event e) {
    if (e) NB2_CUDA_OK(cudaEventDestroy(e));
    return 0;
}
inline int rt_event_record(rt_event e, rt_stream s) {
    NB2_CUDA_OK(cudaEventRecord(e, s));
    return 0;
}
inline int rt_stream_wait(rt_stream s, rt_event e) {
    NB2_CUDA_OK(cudaStreamWaitEvent(s, e, 0));
    return 0;
}
// peer-visible allocations: CUDA IPC handles (64 bytes) let the other processes of the box map the
// buffer, so that a pass can store its output straight into the peers' memory over NVLink
inline int rt_ipc_export(void* p, void* handle64) {
    static_assert(sizeof(cudaIpcMemHandle_t) == 64, "handle size");
    NB2_CUDA_OK(cudaIpcGetMemHandle((cudaIpcMemHandle_t*)handle64, p));
    return 0;
}
inline int rt_ipc_open(const void* handle64, void** p) {
    cudaIpcMemHandle_t h;
    std::memcpy(&h, handle64, sizeof(h));
    NB2_CUDA_OK(cudaIpcOpenMemHandle(p, h, cudaIpcMemLazyEnablePeerAccess));
    return 0;
}
inline int rt_ipc_close(void* p) {
    NB2_CUDA_OK(cudaIpcCloseMemHandle(p));
    return 0;
}
inline int rt_sm_count() {
    int dev = 0, n = 148;
    if (cudaGetDevice(&dev) == cudaSuccess) cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev);
    return n;
}

#else
// ------------------------------------------------------------------ sequential emulator (tests only)
typedef void* rt_stream;

template <class K, int PH, bool IN_RANGE = (PH < K::NPHASES)>
struct EmuStep {
    static void run(const typename K::Params& p, long long bid, int nthreads, char* smem) {
        if (K::enabled(PH))
            for (int tid = 0; tid < nthreads; ++tid) K::template phase<PH>(p, bid, tid, nthreads, smem);
        EmuStep<K, PH + 1>::run(p, bid, nthreads, smem);
    }
};
template <class K, int PH>
struct EmuStep<K, PH, false> {
    static void run(const typename K::Params&, long long, int, char*) {}
};

template <class K>
inline int rt_launch(long long grid, int nthreads, size_t smem, rt_stream, const typename K::Params& p) {
    std::vector<char> buf(smem + 64);
    char* base = buf.data();
    base += (64 - (reinterpret_cast<uintptr_t>(base) & 63)) & 63;
    for (long long bid = 0; bid < grid; ++bid) {
        std::memset(base, 0xAB, smem);  // poison: catches reads of unwritten shared memory
        EmuStep<K, 0>::run(p, bid, nthreads, base);
    }
    ++launch_counter();
    return 0;
}
// pipelined kernels: the emulator runs prefetch (plain copies) and then the phases, tile by tile
template <class K>
inline int rt_launch_pipe(long long ntiles, long long, int nthreads, size_t stage_bytes, int, rt_stream,
                          const typename K::Params& p) {
    std::vector<char> buf(stage_bytes + 64);
    char* base = buf.data();
    base += (64 - (reinterpret_cast<uintptr_t>(base) & 63)) & 63;
    for (long long tile = 0; tile < ntiles; ++tile) {
        std::memset(base, 0xAB, stage_bytes);
        for (int tid = 0; tid < nthreads; ++tid) K::prefetch(p, tile, tid, nthreads, base);
        EmuStep<K, 0>::run(p, tile, nthreads, base);
    }
    ++launch_counter();
    return 0;
}
inline int rt_malloc(void** p, size_t bytes) {
    *p = std::malloc(bytes ? bytes : 1);
    return *p ? 0 : -2;
}
inline int rt_free(void* p) {
    std::free(p);
    return 0;
}
inline int rt_h2d(void* d, const void* s, size_t n, rt_stream) { std::memcpy(d, s, n); return 0; }
inline int rt_d2h(void* d, const void* s, size_t n, rt_stream) { std::memcpy(d, s, n); return 0; }
inline int rt_d2d(void* d, const void* s, size_t n, rt_stream) { std::memmove(d, s, n); return 0; }
inline int rt_sync(rt_stream) { return 0; }
typedef void* rt_event;
inline int rt_stream_create(rt_stream* s) { *s = nullptr; return 0; }
inline int rt_stream_destroy(rt_stream) { return 0; }
inline int rt_event_create(rt_event* e) { *e = nullptr; return 0; }
inline int rt_event_destroy(rt_event) { return 0; }
inline int rt_event_record(rt_event, rt_stream) { return 0; }
inline int rt_stream_wait(rt_stream, rt_event) { return 0; }
inline int rt_ipc_export(void*, void*) { last_error() = "peer memory needs the CUDA build"; return -3; }
inline int rt_ipc_open(const void*, void**) { last_error() = "peer memory needs the CUDA build"; return -3; }
inline int rt_ipc_close(void*) { return -3; }
inline int rt_sm_count() { return 4; }
#endif

}  // namespace nb2
