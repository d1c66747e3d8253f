// srsran-5g_b200/csrc/fec_kernel_packed.cuh -- packed (SIMD-in-register) decode kernel; placeholder until written.
#pragma once
#include "fec_device.cuh"
static inline bool fec_packed_supported(uint32_t) { return false; }
static inline int  fec_packed_launch(cudaStream_t, const fec_cb_desc*, uint32_t, const int8_t*, uint8_t*, fec_cb_result*,
                                     int8_t*, const fec_dev_tables*, uint32_t, int)
{
  return FEC_ERR_BAD_ARG;
}
