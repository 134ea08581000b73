// PCIe probe for the e2e pipeline design: how fast can a k-slice of the row-major
// C (m x k doubles) leave the device?  (a) contiguous D2H, (b) cudaMemcpy2DAsync with a narrow
// width, (c) SM stores straight into mapped pinned host memory, (d) H2D and D2H at the same time.
// build on the GPU box: nvcc -O2 -gencode arch=compute_100a,code=sm_100a -o /tmp/pcie_probe scripts/pcie_probe.cu
#include <cstdio>
#include <cstdlib>
#include <cuda_runtime.h>
#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("%s:%d %s\n", __FILE__, __LINE__, cudaGetErrorString(e)); exit(1); } } while (0)

// every thread copies 16 B; a row of `w16` 16-byte words out of a pitch of `p16` words
__global__ void store_slice(const double2* __restrict__ src, double2* __restrict__ dst, long rows, int w16, int p16, int off16)
{
    long t = blockIdx.x * (long)blockDim.x + threadIdx.x;
    long n = rows * w16;
    for (; t < n; t += (long)gridDim.x * blockDim.x) {
        long r = t / w16; int j = (int)(t % w16);
        dst[r * p16 + off16 + j] = src[r * p16 + off16 + j];
    }
}

int main()
{
    const long m = 1000000; const int k = 32;
    const size_t bytes = (size_t)m * k * 8;
    double *h_x, *h_y, *d_x, *d_y;
    CK(cudaHostAlloc(&h_x, bytes, cudaHostAllocMapped));
    CK(cudaHostAlloc(&h_y, bytes, cudaHostAllocMapped));
    CK(cudaMalloc(&d_x, bytes)); CK(cudaMalloc(&d_y, bytes));
    CK(cudaMemset(d_y, 1, bytes));
    for (size_t i = 0; i < bytes / 8; i += 512) h_x[i] = 1.0, h_y[i] = 0.0;
    cudaStream_t s1, s2; CK(cudaStreamCreate(&s1)); CK(cudaStreamCreate(&s2));
    cudaEvent_t e0, e1; CK(cudaEventCreate(&e0)); CK(cudaEventCreate(&e1));
    float ms;
    double* hy_dev; CK(cudaHostGetDevicePointer(&hy_dev, h_y, 0));

    for (int rep = 0; rep < 2; rep++) {
        CK(cudaEventRecord(e0, s1));
        CK(cudaMemcpyAsync(h_y, d_y, bytes, cudaMemcpyDeviceToHost, s1));
        CK(cudaEventRecord(e1, s1)); CK(cudaEventSynchronize(e1)); CK(cudaEventElapsedTime(&ms, e0, e1));
        printf("contiguous D2H        %7.3f ms  %6.1f GB/s\n", ms, bytes / ms / 1e6);
        CK(cudaEventRecord(e0, s1));
        CK(cudaMemcpyAsync(d_x, h_x, bytes, cudaMemcpyHostToDevice, s1));
        CK(cudaEventRecord(e1, s1)); CK(cudaEventSynchronize(e1)); CK(cudaEventElapsedTime(&ms, e0, e1));
        printf("contiguous H2D        %7.3f ms  %6.1f GB/s\n", ms, bytes / ms / 1e6);
        CK(cudaEventRecord(e0, s1));
        CK(cudaMemcpyAsync(d_x, h_x, bytes, cudaMemcpyHostToDevice, s1));
        CK(cudaMemcpyAsync(h_y, d_y, bytes, cudaMemcpyDeviceToHost, s2));
        CK(cudaStreamSynchronize(s2));
        CK(cudaEventRecord(e1, s1)); CK(cudaEventSynchronize(e1)); CK(cudaEventElapsedTime(&ms, e0, e1));
        printf("H2D || D2H            %7.3f ms  %6.1f GB/s each way\n", ms, bytes / ms / 1e6);
        for (int w = 64; w <= 256; w *= 2) {
            CK(cudaEventRecord(e0, s1));
            for (int off = 0; off < k * 8; off += w)
                CK(cudaMemcpy2DAsync((char*)h_y + off, k * 8, (char*)d_y + off, k * 8, w, m, cudaMemcpyDeviceToHost, s1));
            CK(cudaEventRecord(e1, s1)); CK(cudaEventSynchronize(e1)); CK(cudaEventElapsedTime(&ms, e0, e1));
            printf("2D D2H width %3d B     %7.3f ms  %6.1f GB/s (all slices)\n", w, ms, bytes / ms / 1e6);
        }
        for (int w = 64; w <= 256; w *= 2) {
            for (int grid = 148; grid <= 148 * 8; grid *= 4) {
                CK(cudaEventRecord(e0, s1));
                for (int off = 0; off < k * 8; off += w)
                    store_slice<<<grid, 256, 0, s1>>>((const double2*)d_y, (double2*)hy_dev, m, w / 16, k * 8 / 16, off / 16);
                CK(cudaEventRecord(e1, s1)); CK(cudaEventSynchronize(e1)); CK(cudaEventElapsedTime(&ms, e0, e1));
                printf("SM store to mapped host, slice %3d B, grid %4d  %7.3f ms  %6.1f GB/s\n", w, grid, ms, bytes / ms / 1e6);
            }
        }
        // slice store overlapped with an H2D of the same size
        CK(cudaEventRecord(e0, s1));
        CK(cudaMemcpyAsync(d_x, h_x, bytes, cudaMemcpyHostToDevice, s2));
        for (int off = 0; off < k * 8; off += 64)
            store_slice<<<592, 256, 0, s1>>>((const double2*)d_y, (double2*)hy_dev, m, 4, k * 8 / 16, off / 16);
        CK(cudaStreamSynchronize(s2));
        CK(cudaEventRecord(e1, s1)); CK(cudaEventSynchronize(e1)); CK(cudaEventElapsedTime(&ms, e0, e1));
        printf("H2D || SM store 64 B slices  %7.3f ms  %6.1f GB/s each way\n", ms, bytes / ms / 1e6);
    }
    CK(cudaGetLastError());
    printf("check %g\n", h_y[12345]);
    return 0;
}
