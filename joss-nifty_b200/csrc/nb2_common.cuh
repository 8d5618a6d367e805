// niftyb200 -- common definitions shared by every kernel.
//
// All kernels are written as a sequence of *phases* separated by block barriers.
// A phase is a plain `NB2_HD` function of (params, block id, thread id, #threads, smem).
// On the device the phases of a kernel are executed by the generic launcher in
// nb2_launch.cuh with `__syncthreads()` between them.  The same phase functions can be
// compiled by a host C++ compiler (-DNB2_EMU) and executed by a sequential thread
// emulator; this is TEST INFRASTRUCTURE used to check the index arithmetic on machines
// without a GPU (tests/emu) and is never part of the shipped library.
#pragma once

#include <cstddef>
#include <cstdint>
#include <cmath>

#if defined(__CUDACC__)
#define NB2_HD __host__ __device__ __forceinline__
#define NB2_D __device__ __forceinline__
#define NB2_HDC __host__ __device__ constexpr
#else
#define NB2_HD inline
#define NB2_D inline
#define NB2_HDC constexpr
#endif

#if defined(__CUDA_ARCH__)
#define NB2_UNROLL _Pragma("unroll")
#else
#define NB2_UNROLL
#endif

namespace nb2 {

// ---------------------------------------------------------------- complex numbers
template <typename T>
struct alignas(2 * sizeof(T)) cplx {
    T re, im;
};

template <typename T> NB2_HD cplx<T> mk(T a, T b) { cplx<T> r; r.re = a; r.im = b; return r; }
template <typename T> NB2_HD cplx<T> operator+(cplx<T> a, cplx<T> b) { return mk<T>(a.re + b.re, a.im + b.im); }
template <typename T> NB2_HD cplx<T> operator-(cplx<T> a, cplx<T> b) { return mk<T>(a.re - b.re, a.im - b.im); }
template <typename T> NB2_HD cplx<T> operator*(cplx<T> a, cplx<T> b) {
    return mk<T>(a.re * b.re - a.im * b.im, a.re * b.im + a.im * b.re);
}
template <typename T> NB2_HD cplx<T> operator*(cplx<T> a, T s) { return mk<T>(a.re * s, a.im * s); }
template <typename T> NB2_HD cplx<T> conj(cplx<T> a) { return mk<T>(a.re, -a.im); }
// multiply by -i (forward quarter turn) / +i
template <typename T> NB2_HD cplx<T> mul_mi(cplx<T> a) { return mk<T>(a.im, -a.re); }
template <typename T> NB2_HD cplx<T> mul_pi(cplx<T> a) { return mk<T>(-a.im, a.re); }

// ---------------------------------------------------------------- small helpers
NB2_HD int ilog2(unsigned long long v) { int l = 0; while (v > 1) { v >>= 1; ++l; } return l; }
NB2_HD bool is_pow2(long long v) { return v > 0 && (v & (v - 1)) == 0; }

// Hartley sign: +1 -> "non_canonical_hartley" (Re F + Im F, the reference default,
// src/ducc_dispatch.py:46-50), -1 -> "canonical_hartley" (Re F - Im F).
enum : int { NB2_NONCANONICAL = 0, NB2_CANONICAL = 1, NB2_PRODUCT_LAST = 2 };

// ---------------------------------------------------------------- cache-hinted global access
// Streaming data is touched exactly once per pass: keep it out of L1.
template <typename V> NB2_HD V ld_stream(const V* p) {
#if defined(__CUDA_ARCH__)
    return __ldcs(p);
#else
    return *p;
#endif
}
template <typename V> NB2_HD void st_stream(V* p, V v) {
#if defined(__CUDA_ARCH__)
    __stcs(p, v);
#else
    *p = v;
#endif
}
template <typename V> NB2_HD V ld_ro(const V* p) {
#if defined(__CUDA_ARCH__)
    return __ldg(p);
#else
    return *p;
#endif
}

// ---------------------------------------------------------------- asynchronous global -> shared copies
// cp.async (LDGSTS): the copy engine of the load/store unit moves BYTES (8 or 16) from global
// to shared memory without passing through registers; commit/wait group the copies so that a
// persistent CTA can fetch its next tile while it computes on the current one.
template <int BYTES> NB2_HD void cp_async(void* smem_dst, const void* gmem_src) {
#if defined(__CUDA_ARCH__)
    const unsigned d = (unsigned)__cvta_generic_to_shared(smem_dst);
    if (BYTES == 16)
        asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(d), "l"(gmem_src) : "memory");
    else
        asm volatile("cp.async.ca.shared.global [%0], [%1], 8;" ::"r"(d), "l"(gmem_src) : "memory");
#else
    for (int i = 0; i < BYTES; ++i) ((char*)smem_dst)[i] = ((const char*)gmem_src)[i];
#endif
}
// Bulk L2 prefetch (cp.async.bulk.prefetch.L2): asks the memory system to bring `bytes` (multiple of
// 16, 16-byte aligned) into L2 without occupying registers, shared memory or a load slot.
NB2_HD void prefetch_l2(const void* gmem_src, unsigned bytes) {
#if defined(__CUDA_ARCH__)
    asm volatile("cp.async.bulk.prefetch.L2.global [%0], %1;" ::"l"(gmem_src), "r"(bytes) : "memory");
#else
    (void)gmem_src; (void)bytes;
#endif
}
NB2_HD void cp_async_commit() {
#if defined(__CUDA_ARCH__)
    asm volatile("cp.async.commit_group;" ::: "memory");
#endif
}
template <int PENDING> NB2_HD void cp_async_wait() {
#if defined(__CUDA_ARCH__)
    asm volatile("cp.async.wait_group %0;" ::"n"(PENDING) : "memory");
#endif
}

#if defined(__CUDACC__)
template <typename T> struct vec2_of;
template <> struct vec2_of<float> { using type = float2; };
template <> struct vec2_of<double> { using type = double2; };
#endif

template <typename T> NB2_HD cplx<T> ldc_stream(const cplx<T>* p) {
#if defined(__CUDA_ARCH__)
    typename vec2_of<T>::type v = __ldcs(reinterpret_cast<const typename vec2_of<T>::type*>(p));
    return mk<T>(v.x, v.y);
#else
    return *p;
#endif
}
template <typename T> NB2_HD void stc_stream(cplx<T>* p, cplx<T> v) {
#if defined(__CUDA_ARCH__)
    typename vec2_of<T>::type w; w.x = v.re; w.y = v.im;
    __stcs(reinterpret_cast<typename vec2_of<T>::type*>(p), w);
#else
    *p = v;
#endif
}
// L2-resident intermediates (the group buffer of the L2-blocked tail): cached in L2 with the
// default policy, bypassing L1
template <typename T> NB2_HD cplx<T> ldc_keep(const cplx<T>* p) {
#if defined(__CUDA_ARCH__)
    typename vec2_of<T>::type v = __ldcg(reinterpret_cast<const typename vec2_of<T>::type*>(p));
    return mk<T>(v.x, v.y);
#else
    return *p;
#endif
}
template <typename T> NB2_HD void stc_keep(cplx<T>* p, cplx<T> v) {
#if defined(__CUDA_ARCH__)
    typename vec2_of<T>::type w; w.x = v.re; w.y = v.im;
    __stcg(reinterpret_cast<typename vec2_of<T>::type*>(p), w);
#else
    *p = v;
#endif
}
template <typename T> NB2_HD cplx<T> ldc_ro(const cplx<T>* p) {
#if defined(__CUDA_ARCH__)
    typename vec2_of<T>::type v = __ldg(reinterpret_cast<const typename vec2_of<T>::type*>(p));
    return mk<T>(v.x, v.y);
#else
    return *p;
#endif
}

// ---------------------------------------------------------------- 4-wide vector access (16 B / 32 B)
template <typename T> struct alignas(4 * sizeof(T) > 16 ? 16 : 4 * sizeof(T)) vec4 { T v[4]; };

template <typename T> NB2_HD vec4<T> ld4_stream(const T* p) {
    vec4<T> r;
#if defined(__CUDA_ARCH__)
    if (sizeof(T) == 4) {
        const float4 t = __ldcs(reinterpret_cast<const float4*>(p));
        r.v[0] = T(t.x); r.v[1] = T(t.y); r.v[2] = T(t.z); r.v[3] = T(t.w);
    } else {
        const double2 t0 = __ldcs(reinterpret_cast<const double2*>(p));
        const double2 t1 = __ldcs(reinterpret_cast<const double2*>(p) + 1);
        r.v[0] = T(t0.x); r.v[1] = T(t0.y); r.v[2] = T(t1.x); r.v[3] = T(t1.y);
    }
#else
    for (int i = 0; i < 4; ++i) r.v[i] = p[i];
#endif
    return r;
}
template <typename T> NB2_HD vec4<T> ld4_ro(const T* p) {
    vec4<T> r;
#if defined(__CUDA_ARCH__)
    if (sizeof(T) == 4) {
        const float4 t = __ldg(reinterpret_cast<const float4*>(p));
        r.v[0] = T(t.x); r.v[1] = T(t.y); r.v[2] = T(t.z); r.v[3] = T(t.w);
    } else {
        const double2 t0 = __ldg(reinterpret_cast<const double2*>(p));
        const double2 t1 = __ldg(reinterpret_cast<const double2*>(p) + 1);
        r.v[0] = T(t0.x); r.v[1] = T(t0.y); r.v[2] = T(t1.x); r.v[3] = T(t1.y);
    }
#else
    for (int i = 0; i < 4; ++i) r.v[i] = p[i];
#endif
    return r;
}
// default L2 policy, bypassing L1 (data that another CTA will ask for again soon)
template <typename T> NB2_HD vec4<T> ld4_keep(const T* p) {
    vec4<T> r;
#if defined(__CUDA_ARCH__)
    if (sizeof(T) == 4) {
        const float4 t = __ldcg(reinterpret_cast<const float4*>(p));
        r.v[0] = T(t.x); r.v[1] = T(t.y); r.v[2] = T(t.z); r.v[3] = T(t.w);
    } else {
        const double2 t0 = __ldcg(reinterpret_cast<const double2*>(p));
        const double2 t1 = __ldcg(reinterpret_cast<const double2*>(p) + 1);
        r.v[0] = T(t0.x); r.v[1] = T(t0.y); r.v[2] = T(t1.x); r.v[3] = T(t1.y);
    }
#else
    for (int i = 0; i < 4; ++i) r.v[i] = p[i];
#endif
    return r;
}
template <typename T> NB2_HD void st4_stream(T* p, const vec4<T>& r) {
#if defined(__CUDA_ARCH__)
    if (sizeof(T) == 4) {
        __stcs(reinterpret_cast<float4*>(p), make_float4(float(r.v[0]), float(r.v[1]), float(r.v[2]), float(r.v[3])));
    } else {
        __stcs(reinterpret_cast<double2*>(p), make_double2(double(r.v[0]), double(r.v[1])));
        __stcs(reinterpret_cast<double2*>(p) + 1, make_double2(double(r.v[2]), double(r.v[3])));
    }
#else
    for (int i = 0; i < 4; ++i) p[i] = r.v[i];
#endif
}

}  // namespace nb2
