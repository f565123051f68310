// dev microbenchmark: HBM bandwidth of a copy made of row SEGMENTS (what the store kernel's traffic looks like)
// against a linear copy: rows of `stride` bytes, a work item copies `seg` bytes of one row to the same columns of
// another row.  order 0: consecutive items = adjacent segments of the same row pair (piece-major);
// order 1: consecutive items = the same columns of different row pairs (column-block-major); order 2: random.
#include <cstdio>
#include <cstdlib>
#include <vector>
#include <algorithm>
#include <random>
#include <cuda_runtime.h>
#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("%s: %s\n", #x, cudaGetErrorString(e)); exit(1); } } while (0)

__global__ void __launch_bounds__(256) seg_copy(const uint4 *pool, uint4 *out, const int2 *items, int n_items, long long stride_u4,
                                                int seg_u4, unsigned *ticket)
{
    __shared__ unsigned s_t;
    for (;;)
    {
        __syncthreads();
        if (threadIdx.x == 0)
            s_t = atomicAdd(ticket, 1u);
        __syncthreads();
        const unsigned t = s_t;
        if (t >= (unsigned)n_items)
            return;
        const int2 it = items[t]; // x: src row | dst row << 16 ... kept simple: x = src row, y = (dst row << 12) | seg index
        const long long src = (long long)it.x * stride_u4 + (long long)(it.y & 0xfff) * seg_u4;
        const long long dst = (long long)(it.y >> 12) * stride_u4 + (long long)(it.y & 0xfff) * seg_u4;
        for (int i = threadIdx.x; i < seg_u4; i += 4 * blockDim.x)
        {
            uint4 v[4];
#pragma unroll
            for (int k = 0; k < 4; k++)
                if (i + k * (int)blockDim.x < seg_u4)
                    v[k] = __ldcs(pool + src + i + k * blockDim.x);
#pragma unroll
            for (int k = 0; k < 4; k++)
                if (i + k * (int)blockDim.x < seg_u4)
                    out[dst + i + k * blockDim.x] = v[k];
        }
    }
}

int main(int argc, char **argv)
{
    const int rows = 2000;
    const long long stride_u4 = 31256; // 500,096 bytes
    uint4 *pool, *out;
    CK(cudaMalloc(&pool, rows * stride_u4 * 16));
    CK(cudaMalloc(&out, rows * stride_u4 * 16));
    CK(cudaMemset(pool, 1, rows * stride_u4 * 16));
    CK(cudaMemset(out, 0, rows * stride_u4 * 16));
    unsigned *ticket;
    CK(cudaMalloc(&ticket, 4));
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0);
    cudaEventCreate(&e1);
    std::mt19937 rng(1);
    int sms = 148;
    for (int threads : {128, 256})
        for (int seg_kb : {1, 2, 4, 8, 16, 64})
            for (int order = 0; order < 3; order++)
            {
                const int seg_u4 = seg_kb * 64;
                const int nseg = (int)(stride_u4 / seg_u4); // whole segments only
                std::vector<int> src(rows), dst(rows);
                for (int i = 0; i < rows; i++)
                    src[i] = dst[i] = i;
                std::shuffle(src.begin(), src.end(), rng);
                std::shuffle(dst.begin(), dst.end(), rng);
                std::vector<int2> items;
                if (order == 0)
                    for (int r = 0; r < rows; r++)
                        for (int s = 0; s < nseg; s++)
                            items.push_back(make_int2(src[r], (dst[r] << 12) | s));
                else
                    for (int s = 0; s < nseg; s++)
                        for (int r = 0; r < rows; r++)
                            items.push_back(make_int2(src[r], (dst[r] << 12) | s));
                if (nseg > 4095)
                    continue;
                if (order == 2)
                    std::shuffle(items.begin(), items.end(), rng);
                int2 *d_items;
                CK(cudaMalloc(&d_items, items.size() * sizeof(int2)));
                CK(cudaMemcpy(d_items, items.data(), items.size() * sizeof(int2), cudaMemcpyHostToDevice));
                int per_sm = 0;
                CK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, seg_copy, threads, 0));
                float best = 1e9;
                for (int rep = 0; rep < 5; rep++)
                {
                    CK(cudaMemset(ticket, 0, 4));
                    cudaEventRecord(e0);
                    seg_copy<<<sms * per_sm, threads>>>(pool, out, d_items, (int)items.size(), stride_u4, seg_u4, ticket);
                    cudaEventRecord(e1);
                    CK(cudaEventSynchronize(e1));
                    float ms;
                    cudaEventElapsedTime(&ms, e0, e1);
                    best = std::min(best, ms);
                }
                const double bytes = 2.0 * items.size() * seg_u4 * 16;
                printf("threads %3d seg %2d KB order %d: %.3f ms, %.0f GB/s (%d CTAs/SM)\n", threads, seg_kb, order, best, bytes / best / 1e6, per_sm);
                fflush(stdout);
                cudaFree(d_items);
            }
    // linear copy baseline
    {
        float best = 1e9;
        for (int rep = 0; rep < 5; rep++)
        {
            cudaEventRecord(e0);
            CK(cudaMemcpyAsync(out, pool, rows * stride_u4 * 16, cudaMemcpyDeviceToDevice));
            cudaEventRecord(e1);
            CK(cudaEventSynchronize(e1));
            float ms;
            cudaEventElapsedTime(&ms, e0, e1);
            best = std::min(best, ms);
        }
        printf("cudaMemcpy D2D: %.3f ms, %.0f GB/s\n", best, 2.0 * rows * stride_u4 * 16 / best / 1e6);
    }
    return 0;
}
