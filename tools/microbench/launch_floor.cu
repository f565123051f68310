// dev microbenchmark: the floor of one host-driven round on this box: launch a kernel with a 2 KB parameter block that
// writes one word to mapped host memory; the host waits (a) with cudaEventSynchronize, (b) by polling the word.
#include <chrono>
#include <cstdio>
#include <cuda_runtime.h>
struct Blob { int x[512]; };
__global__ void tiny(const __grid_constant__ Blob b, volatile long long *out, long long v) { if (threadIdx.x == 0) *out = v + b.x[0]; }
__global__ void chain(const __grid_constant__ Blob b, volatile long long *out, long long v, const uint4 *rows, int steps)
{ // a dependent chain of `steps` 16-byte loads (one warp), like a narrow proposal
    uint4 acc = make_uint4(0, 0, 0, 0);
    unsigned idx = threadIdx.x;
    for (int i = 0; i < steps; i++) { uint4 r = rows[idx]; acc.x ^= r.x; idx = (idx + 32 + (acc.x & 0)) & 4095; }
    if (threadIdx.x == 0) *out = v + b.x[0] + (acc.x & 0);
}
int main()
{
    long long *h, *d; cudaHostAlloc(&h, 8, cudaHostAllocMapped); cudaHostGetDevicePointer(&d, h, 0);
    uint4 *rows; cudaMalloc(&rows, 4096 * 16); cudaMemset(rows, 0, 4096 * 16);
    cudaStream_t s; cudaStreamCreateWithFlags(&s, cudaStreamNonBlocking);
    cudaEvent_t e; cudaEventCreateWithFlags(&e, cudaEventDisableTiming);
    Blob b = {};
    const int N = 20000;
    for (int mode = 0; mode < 4; mode++)
    {
        *h = 0;
        auto t0 = std::chrono::steady_clock::now();
        for (long long i = 1; i <= N; i++)
        {
            if (mode < 2) tiny<<<1, 64, 0, s>>>(b, d, i); else chain<<<4, 32, 0, s>>>(b, d, i, rows, 40);
            if (mode % 2 == 0) { cudaEventRecord(e, s); cudaEventSynchronize(e); }
            else while (*(volatile long long *)h != i) {}
        }
        cudaStreamSynchronize(s);
        double us = std::chrono::duration<double, std::micro>(std::chrono::steady_clock::now() - t0).count() / N;
        printf("%s kernel, wait by %s: %.2f us per round\n", mode < 2 ? "empty" : "40-step chain", mode % 2 ? "polling mapped memory" : "event", us);
    }
    return 0;
}
