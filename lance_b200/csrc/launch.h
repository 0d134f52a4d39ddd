// launch.h -- the scan kernels live in their own translation units (scan12_tu.cu, scan3_tu.cu: ~40 large
// template instantiations) so that the host pipeline (api.cu) rebuilds without them.
#pragma once
#include <cuda_runtime.h>

#include "scan.cuh"
#include "scan3.cuh"

namespace lb200 {

struct ScanConfig {
  int variant, ql, dsub_t;
  uint32_t capw;
  size_t smem;
  int minb;  // variant 3: CTAs per SM the kernel is compiled for
  int nw = 8;  // variant 3: warps per CTA
};

constexpr int SCAN_NW = 8;

// variants 1 / 2 (tile16 layout), scan12_tu.cu
cudaError_t launch_scan(const ScanConfig& c, int grid, cudaStream_t st, const ScanParams& sp,
                        const uint32_t* total_items);
// variant 3 (skew streams), scan3_tu.cu
cudaError_t launch_scan3(const ScanConfig& c, bool boot, int sm_count, cudaStream_t st, const Scan3Params& sp,
                         const uint32_t* total_items);

void preload_scan3(int ql, int dsub_t);  // module-load every variant-3 instantiation of this (QL, dsub)

}  // namespace lb200
