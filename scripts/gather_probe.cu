// L2 -> SM gather ceiling probe: every lane group of G lanes reads random ROWB-byte rows of a
// window of W rows with UNROLL independent loads in flight and nothing else to do (indices from an LCG in
// registers, one add per loaded value).  Reports achieved bytes/s for several windows and row widths --
// the ceiling the row-split SpMM kernel (one 8k-byte B row per nonzero) can be compared with.
// build: nvcc -O3 -gencode arch=compute_100a,code=sm_100a -o /tmp/gather_probe scripts/gather_probe.cu
#include <cstdio>
#include <cstdlib>
#include <cuda_runtime.h>
#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("%s:%d %s\n", __FILE__, __LINE__, cudaGetErrorString(e)); exit(1); } } while (0)

template <int G, int UNROLL>
__global__ void __launch_bounds__(256) gather(const double2* __restrict__ B, unsigned rows, int iters, double* out)
{
    const int lig = threadIdx.x % G;
    const unsigned gid = (blockIdx.x * 256u + threadIdx.x) / G;
    unsigned s = gid * 2654435761u + 12345u;
    double acc = 0.0;
    for (int it = 0; it < iters; it++) {
        double2 v[UNROLL];
        #pragma unroll
        for (int u = 0; u < UNROLL; u++) {
            s = s * 1664525u + 1013904223u;
            const unsigned r = (unsigned) (((unsigned long long) s * rows) >> 32);
            v[u] = __ldg(B + (size_t) r * G + lig);
        }
        #pragma unroll
        for (int u = 0; u < UNROLL; u++) acc += v[u].x + v[u].y;
    }
    if (acc == 123.456) out[0] = acc;
}

template <int G, int UNROLL>
void run(const double2* B, size_t window_bytes, double* out, int ctas_per_sm)
{
    const unsigned rows = (unsigned) (window_bytes / (G * 16));
    const int grid = 148 * ctas_per_sm, iters = 4096 / UNROLL;
    cudaEvent_t e0, e1; CK(cudaEventCreate(&e0)); CK(cudaEventCreate(&e1));
    gather<G, UNROLL><<<grid, 256>>>(B, rows, iters, out);
    CK(cudaEventRecord(e0));
    gather<G, UNROLL><<<grid, 256>>>(B, rows, iters, out);
    CK(cudaEventRecord(e1)); CK(cudaEventSynchronize(e1));
    float ms; CK(cudaEventElapsedTime(&ms, e0, e1));
    const double bytes = (double) grid * 256 / G * iters * UNROLL * (G * 16.0);
    printf("row %4d B  window %7.1f MB  unroll %d  ctas/SM %2d : %7.3f ms  %6.2f TB/s\n", G * 16, window_bytes / 1e6, UNROLL, ctas_per_sm, ms, bytes / ms / 1e9);
}

int main()
{
    const size_t cap = (size_t) 1 << 30;
    double2* B; double* out;
    CK(cudaMalloc(&B, cap)); CK(cudaMemset(B, 0, cap)); CK(cudaMalloc(&out, 8));
    const double windows_mb[] = {0.1, 1.6, 12.8, 50, 100, 256, 1000};
    for (double w : windows_mb) {
        const size_t wb = (size_t) (w * 1e6);
        run<16, 4>(B, wb, out, 8);
        run<16, 8>(B, wb, out, 8);
        run<16, 8>(B, wb, out, 4);
        run<8, 8>(B, wb, out, 8);
        run<32, 4>(B, wb, out, 8);
        run<4, 8>(B, wb, out, 8);
    }
    return 0;
}
