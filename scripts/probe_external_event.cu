// Does cudaStreamWaitEvent on an event recorded by an EXTERNAL record node of a just-launched graph wait for that launch's node?
#include <cstdio>
#include <cuda_runtime.h>
__global__ void spin(long long cycles, int* flag, int v) { long long t0 = clock64(); while (clock64() - t0 < cycles) {} if (flag) atomicExch(flag, v); }
__global__ void observe(const int* flag, int* out) { *out = *flag; }
int main() {
    cudaStream_t s1, s2; cudaStreamCreateWithFlags(&s1, cudaStreamNonBlocking); cudaStreamCreateWithFlags(&s2, cudaStreamNonBlocking);
    cudaEvent_t E; cudaEventCreateWithFlags(&E, cudaEventDisableTiming);
    int *flag, *out; cudaMalloc(&flag, 4); cudaMalloc(&out, 4);
    cudaGraph_t g; cudaGraphExec_t ge;
    cudaStreamBeginCapture(s1, cudaStreamCaptureModeThreadLocal);
    spin<<<1, 1, 0, s1>>>(20000000, flag, 1);               // A: ~10 ms, then flag = 1
    cudaError_t e = cudaEventRecordWithFlags(E, s1, cudaEventRecordExternal);
    spin<<<1, 1, 0, s1>>>(20000000, flag, 2);               // B: ~10 ms, then flag = 2
    cudaStreamEndCapture(s1, &g);
    printf("record external: %s\n", cudaGetErrorString(e));
    e = cudaGraphInstantiate(&ge, g, 0); printf("instantiate: %s\n", cudaGetErrorString(e));
    for (int rep = 0; rep < 3; ++rep) {
        cudaMemset(flag, 0, 4); cudaMemset(out, 0xff, 4); cudaDeviceSynchronize();
        cudaGraphLaunch(ge, s1);
        cudaStreamWaitEvent(s2, E, 0);                       // enqueued at once, long before A finishes
        observe<<<1, 1, 0, s2>>>(flag, out);
        cudaDeviceSynchronize();
        int h = -1; cudaMemcpy(&h, out, 4, cudaMemcpyDeviceToHost);
        printf("rep %d: observer saw flag = %d (1 = waited for A of THIS launch and ran before B ended; 0 = did not wait; 2 = waited for all)\n", rep, h);
    }
    return 0;
}
