// scan3_tu.cu -- instantiations + launchers of scan3_kernel (variant 3).
#include "launch.h"

namespace lb200 {

template <int QL, int D, int NW, int MINB, int DS, bool BOOT>
static auto scan3_entry() {
  return scan3_kernel<QL, D, NW, MINB, DS, BOOT>;
}

template <int QL, int D, int MINB, bool BOOT, int NW = SCAN_NW>
static cudaError_t launch_scan3_d(int dsub_t, int sm_count, size_t smem, cudaStream_t st, const Scan3Params& sp,
                                  const uint32_t* total_items) {
#define LB200_LAUNCH3(DS)                                                                         \
  do {                                                                                            \
    auto kern = scan3_entry<QL, D, NW, MINB, DS, BOOT>();                                         \
    cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem); \
    if (e != cudaSuccess) return e;                                                               \
    int nb = 0;                                                                                   \
    e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&nb, kern, NW * 32, smem);                  \
    if (e != cudaSuccess) return e;                                                               \
    if (nb < 1) return cudaErrorLaunchOutOfResources;                                             \
    kern<<<sm_count * nb, NW * 32, smem, st>>>(sp, total_items);                                  \
    return cudaGetLastError();                                                                    \
  } while (0)
  if (dsub_t == 8) LB200_LAUNCH3(8);
  if (dsub_t == 16) LB200_LAUNCH3(16);
  if constexpr (NW == SCAN_NW) LB200_LAUNCH3(0);
  return cudaErrorInvalidValue;
#undef LB200_LAUNCH3
}

template <bool BOOT>
static cudaError_t launch_scan3_b(const ScanConfig& c, int sm_count, cudaStream_t st, const Scan3Params& sp,
                                  const uint32_t* total_items) {
  if (c.ql == 4) {
    if (c.minb == 2) return launch_scan3_d<4, 8, 2, BOOT>(c.dsub_t, sm_count, c.smem, st, sp, total_items);
    return launch_scan3_d<4, 8, 1, BOOT>(c.dsub_t, sm_count, c.smem, st, sp, total_items);
  }
  if (c.minb == 2) return launch_scan3_d<2, 16, 2, BOOT>(c.dsub_t, sm_count, c.smem, st, sp, total_items);
  return launch_scan3_d<2, 16, 1, BOOT>(c.dsub_t, sm_count, c.smem, st, sp, total_items);
}
cudaError_t launch_scan3(const ScanConfig& c, bool boot, int sm_count, cudaStream_t st, const Scan3Params& sp,
                                const uint32_t* total_items) {
  if (c.nw == 16) {  // wide CTA: one CTA per SM, tables pre-built (see search_device)
    if (c.dsub_t == 0) return cudaErrorInvalidValue;
    if (c.ql == 4) return launch_scan3_d<4, 8, 1, true, 16>(c.dsub_t, sm_count, c.smem, st, sp, total_items);
    return launch_scan3_d<2, 16, 1, true, 16>(c.dsub_t, sm_count, c.smem, st, sp, total_items);
  }
  return boot ? launch_scan3_b<true>(c, sm_count, st, sp, total_items)
              : launch_scan3_b<false>(c, sm_count, st, sp, total_items);
}


// Force the module load of every instantiation a handle with this (QL, dsub) may launch.  Under CUDA's lazy
// module loading the FIRST launch of a kernel may have to synchronise with running kernels; a shard whose
// stream holds a kernel waiting for a peer (api_group.inl) must never meet that on its way to the kernels
// the peer is waiting for.
template <int QL, int D, int DS>
static void preload_qd() {
  cudaFuncAttributes a;
  cudaFuncGetAttributes(&a, scan3_entry<QL, D, 8, 1, DS, true>());
  cudaFuncGetAttributes(&a, scan3_entry<QL, D, 8, 2, DS, true>());
  cudaFuncGetAttributes(&a, scan3_entry<QL, D, 8, 1, DS, false>());
  cudaFuncGetAttributes(&a, scan3_entry<QL, D, 8, 2, DS, false>());
  if constexpr (DS != 0) cudaFuncGetAttributes(&a, scan3_entry<QL, D, 16, 1, DS, true>());
}
void preload_scan3(int ql, int dsub_t) {
  if (ql == 4) {
    if (dsub_t == 8) preload_qd<4, 8, 8>();
    else if (dsub_t == 16) preload_qd<4, 8, 16>();
    else preload_qd<4, 8, 0>();
  } else {
    if (dsub_t == 8) preload_qd<2, 16, 8>();
    else if (dsub_t == 16) preload_qd<2, 16, 16>();
    else preload_qd<2, 16, 0>();
  }
  cudaGetLastError();
}

}  // namespace lb200
