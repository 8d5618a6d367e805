#include <cstdio>
#include <cuda_runtime.h>
#define CK(x) do { cudaError_t e = (x); printf("%-60s -> %s\n", #x, cudaGetErrorString(e)); } while (0)
__global__ void k(float* p) { p[threadIdx.x] = 1.f; }
int main() {
    cudaStream_t aux; cudaEvent_t ev, ej; float* buf = nullptr;
    CK(cudaMalloc(&buf, 1 << 20));
    CK(cudaEventCreateWithFlags(&ev, cudaEventDisableTiming));
    CK(cudaEventCreateWithFlags(&ej, cudaEventDisableTiming));
    CK(cudaStreamCreateWithFlags(&aux, cudaStreamNonBlocking));
    for (int it = 0; it < 3; ++it) {
        if (it == 1) {
            CK(cudaStreamSynchronize(0)); CK(cudaStreamSynchronize(aux));
            CK(cudaFree(buf)); CK(cudaMalloc(&buf, 4 << 20));
        }
        k<<<1, 32, 0, 0>>>(buf);
        CK(cudaEventRecord(ev, 0));
        CK(cudaStreamWaitEvent(aux, ev, 0));
        k<<<1, 32, 0, aux>>>(buf);
        CK(cudaEventRecord(ej, aux));
        CK(cudaStreamWaitEvent(0, ej, 0));
        CK(cudaDeviceSynchronize());
    }
    return 0;
}
