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

// `gate`: device-resident conjugate gradient (nb2_cg_* in nb2_capi.cu).  While the solver's state says
// "stopped" (*gate != 0) every kernel of an iteration that was launched ahead returns at once, so the
// host may enqueue iterations without waiting for the stopping decision of the previous one.
template <class K>
__global__ void __launch_bounds__(K::MAX_THREADS, K::MIN_BLOCKS) nb2_kernel(const typename K::Params p, const int* gate) {
    extern __shared__ __align__(16) char nb2_smem[];
    if (gate != nullptr && *gate != 0) return;
    PhaseStep<K, 0>::run(p, nb2_smem, true);
}

// gate handed to every launch of the calling thread (nullptr: none)
inline const int*& launch_gate() {
    static thread_local const int* g = nullptr;
    return g;
}
inline int rt_device() {
    int dev = 0;
    cudaGetDevice(&dev);
    return dev;
}
constexpr int kMaxDevices = 64;

template <class K>
inline int rt_launch(long long grid, int nthreads, size_t smem, rt_stream stream, const typename K::Params& p) {
    if (grid <= 0) return 0;
    if (grid > 2147483647ll) {
        last_error() = "grid too large";
        return -1;
    }
    if (smem > 48 * 1024) {
        static size_t configured[kMaxDevices] = {0};  // per kernel instantiation and device
        size_t& c = configured[rt_device() % kMaxDevices];
        if (smem > c) {
            NB2_CUDA_OK(cudaFuncSetAttribute(nb2_kernel<K>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
            c = smem;
        }
    }
    nb2_kernel<K><<<(unsigned)grid, nthreads, smem, stream>>>(p, launch_gate());
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
    static int cached[kMaxDevices] = {0};
    int& n = cached[rt_device() % kMaxDevices];
    if (n == 0) {
        int v = 148;
        cudaDeviceGetAttribute(&v, cudaDevAttrMultiProcessorCount, rt_device());
        n = v;
    }
    return n;
}

#else
// ------------------------------------------------------------------ sequential emulator (tests only)
typedef void* rt_stream;
inline const int*& launch_gate() {
    static thread_local const int* g = nullptr;
    return g;
}
inline int rt_device() { return 0; }
constexpr int kMaxDevices = 64;

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
    if (launch_gate() != nullptr && *launch_gate() != 0) {
        ++launch_counter();
        return 0;
    }
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
