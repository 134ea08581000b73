// umma_probe.cu -- isolates the tcgen05 part of spmm_tf32_blocked.cuh: one CTA multiplies a dense 128 x 32 TF32 tile by a
// 32 x N tile with the SAME shared-memory layouts, descriptors and wrappers the product kernel uses, reads TMEM back and
// compares with the host.  Variants swap the descriptor strides so that one run tells which convention the hardware uses.
// build: nvcc -O2 -std=c++17 -gencode arch=compute_100a,code=sm_100a -I spmm-research_b200/csrc -o /tmp/umma_probe scripts/umma_probe.cu
#include <cstdio>
#include <cstdlib>
#include <cmath>
#include <vector>
#include <cuda_runtime.h>
#include "spmm_tf32_blocked.cuh"
using namespace sb200;
#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("%s:%d %s\n", __FILE__, __LINE__, cudaGetErrorString(e)); exit(1); } } while (0)

// variant bits: 1 = swap LBO/SBO in the A descriptor, 2 = swap in the B descriptor, 4 = B stored MN-major (b_major = 1,
// element (n, kk) at (n/4)*512 + (kk/8)*128 + (kk%8)*16 + (n%4)*4 -- the form cp.async could fill without a transpose),
// 8 = no K advance inside the stage (one MMA only, K = 8).  Variant 0 is the product kernel's layout.
// Result on B200 (profiles/r2_tf32_probe.md): K-major B exact; every MN-major variant returns all zeros.
template <int N>
__global__ void __launch_bounds__(128) probe(const float* __restrict__ A, const float* __restrict__ B, float* __restrict__ D, int variant)
{
	using Cfg = Tf32Cfg<N, 1>;
	constexpr int KT = TF32_KT;
	extern __shared__ __align__(128) unsigned char sm[];
	const int tid = threadIdx.x, warp = tid >> 5;
	unsigned char* ctl = sm + Cfg::STAGE_BYTES;
	const uint32_t sA_u = smem_u32(sm), sB_u = sA_u + Cfg::A_BYTES, bar = smem_u32(ctl);
	volatile uint32_t* slot = reinterpret_cast<volatile uint32_t*>(ctl + 8);
	if (warp == 0) tmem_alloc(smem_u32(ctl + 8), N);
	if (tid == 0) { mbar_init(bar, 1); asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory"); }
	tc_fence_before(); __syncthreads(); tc_fence_after();
	const uint32_t tmem = *slot;
	const bool b_kmajor = !(variant & 4);
	constexpr int MN_LBO = 128, MN_SBO = (TF32_KT / 8) * 128;
	// A: row tid
	for (int kk = 0; kk < KT; kk++) {
		uint32_t bits;
		asm("cvt.rna.tf32.f32 %0, %1;" : "=r"(bits) : "f"(A[tid * KT + kk]));
		*reinterpret_cast<uint32_t*>(sm + (tid >> 3) * Cfg::SBO_A + (kk >> 2) * Cfg::LBO + (tid & 7) * 16 + (kk & 3) * 4) = bits;
	}
	// B: element (kk, n)
	for (int e = tid; e < KT * N; e += 128) {
		const int kk = e / N, n = e % N;
		uint32_t off;
		if (!b_kmajor) off = (n >> 2) * MN_SBO + (kk >> 3) * MN_LBO + (kk & 7) * 16 + (n & 3) * 4;
		else off = (n >> 3) * Cfg::SBO_B + (kk >> 2) * Cfg::LBO + (n & 7) * 16 + (kk & 3) * 4;       // K-major like A
		*reinterpret_cast<float*>(sm + Cfg::A_BYTES + off) = B[e];
	}
	asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
	tc_fence_before(); __syncthreads();
	if (tid == 0) {
		tc_fence_after();
		uint32_t idesc = Cfg::IDESC;
		if (!b_kmajor) idesc |= 1u << 16;
		const int nk = (variant & 8) ? 1 : KT / 8;
		for (int ks = 0; ks < nk; ks++) {
			const uint32_t a_lbo = (variant & 1) ? Cfg::SBO_A : Cfg::LBO, a_sbo = (variant & 1) ? Cfg::LBO : Cfg::SBO_A;
			uint64_t ad = umma_desc(sA_u + ks * 2 * Cfg::LBO, a_lbo, a_sbo), bd;
			if (!b_kmajor) {
				bd = umma_desc(sB_u + ks * MN_LBO, (variant & 2) ? MN_SBO : MN_LBO, (variant & 2) ? MN_LBO : MN_SBO);
			} else {
				bd = umma_desc(sB_u + ks * 2 * Cfg::LBO, (variant & 2) ? Cfg::SBO_B : Cfg::LBO, (variant & 2) ? Cfg::LBO : Cfg::SBO_B);
			}
			tc_mma_tf32(tmem, ad, bd, idesc, ks > 0 ? 1u : 0u);
		}
		tc_commit(bar);
	}
	mbar_wait_bounded(bar, 0);
	tc_fence_after();
	for (int c0 = 0; c0 < N; c0 += 16) {
		float v[16];
		tmem_ld16(tmem + ((uint32_t) (warp * 32) << 16) + c0, v);
		for (int q = 0; q < 16; q++) D[tid * N + c0 + q] = v[q];
	}
	tc_fence_before(); __syncthreads();
	if (warp == 0) tmem_dealloc(tmem, N);
}

template <int N>
void run(int variant)
{
	const int KT = TF32_KT;
	std::vector<float> A(128 * KT), B(KT * N), D(128 * N), R(128 * N, 0.f);
	for (int i = 0; i < 128 * KT; i++) A[i] = (float) ((i * 7 + 3) % 11 - 5);           // small integers: exact in TF32
	for (int i = 0; i < KT * N; i++) B[i] = (float) ((i * 5 + 1) % 13 - 6);
	const int nk = (variant & 8) ? 8 : KT;
	for (int r = 0; r < 128; r++) for (int n = 0; n < N; n++) { float s = 0; for (int k = 0; k < nk; k++) s += A[r * KT + k] * B[k * N + n]; R[r * N + n] = s; }
	float *dA, *dB, *dD;
	CK(cudaMalloc(&dA, A.size() * 4)); CK(cudaMalloc(&dB, B.size() * 4)); CK(cudaMalloc(&dD, D.size() * 4));
	CK(cudaMemcpy(dA, A.data(), A.size() * 4, cudaMemcpyHostToDevice)); CK(cudaMemcpy(dB, B.data(), B.size() * 4, cudaMemcpyHostToDevice));
	CK(cudaMemset(dD, 0xff, D.size() * 4));
	const size_t smem = Tf32Cfg<N, 1>::SMEM;
	CK(cudaFuncSetAttribute(probe<N>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int) smem));
	probe<N><<<1, 128, smem>>>(dA, dB, dD, variant);
	cudaError_t e = cudaDeviceSynchronize();
	if (e != cudaSuccess) { printf("N=%d variant=%d: kernel failed: %s\n", N, variant, cudaGetErrorString(e)); exit(2); }
	CK(cudaMemcpy(D.data(), dD, D.size() * 4, cudaMemcpyDeviceToHost));
	int bad = 0, zero = 0; double worst = 0;
	for (size_t i = 0; i < D.size(); i++) { if (D[i] != R[i]) bad++; if (D[i] == 0.f) zero++; worst = fmax(worst, fabs((double) D[i] - R[i])); }
	printf("N=%3d variant=%2d (a_swap=%d b_swap=%d b_mnmajor=%d oneK=%d): mismatches %5d / %d, zeros %d, max abs diff %.1f | D[0..3]=%.0f %.0f %.0f %.0f ref %.0f %.0f %.0f %.0f | D[row1]=%.0f ref %.0f\n",
	       N, variant, variant & 1, (variant >> 1) & 1, (variant >> 2) & 1, (variant >> 3) & 1, bad, (int) D.size(), zero, worst,
	       D[0], D[1], D[2], D[3], R[0], R[1], R[2], R[3], D[N], R[N]);
	cudaFree(dA); cudaFree(dB); cudaFree(dD);
}

int main()
{
	for (int v = 0; v < 16; v++) run<128>(v);
	run<32>(0); run<32>(4);
	run<64>(0);
	return 0;
}
