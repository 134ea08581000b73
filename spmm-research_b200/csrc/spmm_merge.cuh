// spmm_merge.cuh -- CSR SpMM, merge-path variant (placeholder until the kernel lands).
#pragma once
#include <stdint.h>
namespace sb200 {
struct MergePlan {
	int ntiles = 0;
};
} // namespace sb200
