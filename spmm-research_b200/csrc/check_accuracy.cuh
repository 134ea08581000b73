// check_accuracy.cuh -- device twin of the reference's CheckAccuracy()
// (benchmark_code/CPU/AMD/spmv_code_bench/spmv_bench.cpp:121-206) and of the error metrics it prints
// (lib/array_metrics.c: mae, max_ae, mse, mape, smape, lnQ_error, mlare, gmare).
//
// Gold result: the reference forms each product a[j]*x[..] in fp64 (csr_a_ref and x_ref are double*,
// spmv_bench.cpp:152) and accumulates the products in __float128 with Kahan compensation.  sm_100a has no
// fp128, so the sum runs in double-double: the same fp64 product, a (hi, lo) running sum updated with
// TwoSum -- about 106 significant bits, far below the rounding of the final conversion to fp64, which is
// what the metrics consume (reference_to_double).
#pragma once
#include <stdint.h>
#include <float.h>
#include <cuda_runtime.h>

namespace sb200 {

struct dd { double hi, lo; };

__device__ __forceinline__ dd dd_add_prod(dd s, double a, double x)
{
	const double p = __dmul_rn(a, x);             // fp64 product like the reference (never contracted into an FMA)
	const double t = s.hi + p;                    // TwoSum(s.hi, p)
	const double bb = t - s.hi;
	const double err = (s.hi - (t - bb)) + (p - bb);
	const double lo = s.lo + err;
	dd r;
	r.hi = t + lo;                                // FastTwoSum(t, lo)
	r.lo = lo - (r.hi - t);
	return r;
}

// One thread per C entry (row, n), n fastest: B is row-major (c x k) so a row's k threads read one
// contiguous B row per nonzero.
__global__ void gold_spmm_kernel(const int32_t* __restrict__ ia, const int32_t* __restrict__ ja, const double* __restrict__ a,
                                 const double* __restrict__ B, double* __restrict__ gold, double* __restrict__ gold_lo, int64_t m, int k)
{
	const int64_t t = (int64_t) blockIdx.x * blockDim.x + threadIdx.x;
	if (t >= m * k) return;
	const int64_t row = t / k;
	const int n = (int) (t % k);
	dd s; s.hi = 0.0; s.lo = 0.0;
	for (int j = ia[row]; j < ia[row + 1]; j++)
		s = dd_add_prod(s, __ldg(a + j), __ldg(B + (size_t) __ldg(ja + j) * k + n));
	const double g = s.hi + s.lo;                 // (double) gold, what the metrics consume
	gold[t] = g;
	gold_lo[t] = (s.hi - g) + s.lo;               // gold - (double) gold, for the fp128-style max relative difference
}

constexpr int ACC_NPART = 7;       // max_rel, sum_ae, max_ae, sum_se, sum_are, sum_sare, sum_lnq

// Per-block partials of the metrics, combined in a fixed order (deterministic).
template <int THREADS>
__global__ void __launch_bounds__(THREADS)
accuracy_stats_kernel(const double* __restrict__ gold, const double* __restrict__ gold_lo, const double* __restrict__ test, int64_t n, double eps,
                      double* __restrict__ partials)
{
	double v[ACC_NPART];
	#pragma unroll
	for (int i = 0; i < ACC_NPART; i++) v[i] = 0.0;
	for (int64_t i = (int64_t) blockIdx.x * THREADS + threadIdx.x; i < n; i += (int64_t) gridDim.x * THREADS) {
		const double g = gold[i], f = test[i];
		const double ae = fabs(g - f);
		// spmv_bench.cpp:165-172: |gold - y| / |gold| in the reference's extended type, only where gold > eps (signed)
		if (g > eps) v[0] = fmax(v[0], fabs((g - f) + gold_lo[i]) / fabs(g));
		v[1] += ae;                                                                  // mae
		v[2] = fmax(v[2], ae);                                                       // max_ae
		v[3] += (g - f) * (g - f);                                                   // mse
		v[4] += ae / fmax(fabs(g), DBL_EPSILON);                                     // mare -> mape
		v[5] += ae / fmax(fabs(g) + fabs(f), DBL_EPSILON);                           // smare -> smape
		v[6] += log10(fmax(fabs(f), DBL_EPSILON)) - log10(fmax(fabs(g), DBL_EPSILON));   // lnQ_error
	}
	__shared__ double sh[ACC_NPART][THREADS];
	#pragma unroll
	for (int i = 0; i < ACC_NPART; i++) sh[i][threadIdx.x] = v[i];
	__syncthreads();
	for (int off = THREADS / 2; off > 0; off >>= 1) {
		if (threadIdx.x < off) {
			#pragma unroll
			for (int i = 0; i < ACC_NPART; i++) {
				const double x = sh[i][threadIdx.x], y = sh[i][threadIdx.x + off];
				sh[i][threadIdx.x] = (i == 0 || i == 2) ? fmax(x, y) : x + y;
			}
		}
		__syncthreads();
	}
	if (threadIdx.x < ACC_NPART) partials[(size_t) blockIdx.x * ACC_NPART + threadIdx.x] = sh[threadIdx.x][0];
}

} // namespace sb200
