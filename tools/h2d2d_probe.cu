// h2d2d_probe.cu — host->device rate of a column strip of a row-major pinned matrix (development aid):
// cudaMemcpy2DAsync versus a gather kernel reading the pinned host memory directly, by strip width.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o tools/h2d2d_probe tools/h2d2d_probe.cu
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>
typedef unsigned long long u64;
// one warp per (row, 32 x 16-byte chunk): dst[row*w + c] = src[row*ld + c]
__global__ void gather_kernel(ulonglong2 *__restrict__ dst, const ulonglong2 *__restrict__ src, u64 rows, u64 w2, u64 ld2)
{
    const u64 total = rows * w2;
    for (u64 i = (u64)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (u64)gridDim.x * blockDim.x)
    {
        const u64 r = i / w2, c = i - r * w2;
        dst[i] = src[r * ld2 + c];
    }
}
int main()
{
    const u64 rows = 1ull << 20, ncols = 664; // even column count so that 16-byte accesses stay aligned
    u64 *h, *d;
    cudaHostAlloc(&h, rows * ncols * 8, cudaHostAllocDefault);
    for (u64 i = 0; i < rows * ncols; i += 512) h[i] = i;
    cudaMalloc(&d, rows * ncols * 8);
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0);
    cudaEventCreate(&e1);
    const int widths[] = {16, 32, 64, 112, 128, 192, 256, 332, 664};
    for (int w : widths)
    {
        float ms;
        cudaMemcpy2DAsync(d, w * 8, h, ncols * 8, w * 8, rows, cudaMemcpyHostToDevice, 0);
        cudaEventRecord(e0);
        cudaMemcpy2DAsync(d, w * 8, h, ncols * 8, w * 8, rows, cudaMemcpyHostToDevice, 0);
        cudaEventRecord(e1);
        cudaEventSynchronize(e1);
        cudaEventElapsedTime(&ms, e0, e1);
        printf("w=%3d cols (%4d B/row)  memcpy2D %.1f GB/s", w, w * 8, rows * w * 8 / ms / 1e6);
        for (int ctas : {8, 16, 32, 64, 148})
        {
            gather_kernel<<<ctas, 512>>>((ulonglong2 *)d, (const ulonglong2 *)h, rows, w / 2, ncols / 2);
            cudaEventRecord(e0);
            gather_kernel<<<ctas, 512>>>((ulonglong2 *)d, (const ulonglong2 *)h, rows, w / 2, ncols / 2);
            cudaEventRecord(e1);
            cudaEventSynchronize(e1);
            cudaEventElapsedTime(&ms, e0, e1);
            printf("  gather@%d %.1f", ctas, rows * w * 8 / ms / 1e6);
        }
        printf("\n");
    }
    printf("status: %s\n", cudaGetErrorString(cudaGetLastError()));
    return 0;
}
