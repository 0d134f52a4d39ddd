// scan12_tu.cu -- instantiations + launchers of scan_kernel (variants 1 and 2).
#include "launch.h"

namespace lb200 {

template <int VARIANT, int QL>
static cudaError_t launch_scan_d(int dsub_t, int grid, size_t smem, cudaStream_t st, const ScanParams& sp,
                                 const uint32_t* total_items) {
#define LB200_LAUNCH(D)                                                                           \
  do {                                                                                            \
    auto kern = scan_kernel<VARIANT, QL, SCAN_NW, D>;                                             \
    cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem); \
    if (e != cudaSuccess) return e;                                                               \
    kern<<<grid, SCAN_NW * 32, smem, st>>>(sp, total_items);                                      \
    return cudaGetLastError();                                                                    \
  } while (0)
  if (dsub_t == 8) LB200_LAUNCH(8);
  if (dsub_t == 16) LB200_LAUNCH(16);
  LB200_LAUNCH(0);
#undef LB200_LAUNCH
}

cudaError_t launch_scan(const ScanConfig& c, int grid, cudaStream_t st, const ScanParams& sp,
                               const uint32_t* total_items) {
  if (c.variant == 1) return launch_scan_d<1, 1>(c.dsub_t, grid, c.smem, st, sp, total_items);
  switch (c.ql) {
    case 8: return launch_scan_d<2, 8>(c.dsub_t, grid, c.smem, st, sp, total_items);
    case 4: return launch_scan_d<2, 4>(c.dsub_t, grid, c.smem, st, sp, total_items);
    case 2: return launch_scan_d<2, 2>(c.dsub_t, grid, c.smem, st, sp, total_items);
    default: return launch_scan_d<2, 1>(c.dsub_t, grid, c.smem, st, sp, total_items);
  }
}

}  // namespace lb200
