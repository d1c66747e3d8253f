// srsran-5g_b200/csrc/fec_device.cuh -- device-side tables and helpers shared by the FEC kernels (sm_100a).
//
// Arithmetic contract (bit-exact with the reference's generic kernels; file:line relative to the reference tree):
//   saturating / promoting LLR sums   lib/phy/upper/log_likelihood_ratio.cpp:40-88
//   v2c / check analysis / c2v / soft lib/phy/upper/channel_coding/ldpc/ldpc_decoder_generic.cpp:30-122
//   layer loop, layer trimming        lib/phy/upper/channel_coding/ldpc/ldpc_decoder_impl.cpp:60-150
//   hard decision + zero test         lib/phy/upper/log_likelihood_ratio.cpp:350-376
//   CRC remainder                     lib/phy/upper/channel_coding/crc_calculator_generic_impl.cpp:111-133
//   rate dematcher                    lib/phy/upper/channel_coding/ldpc/ldpc_rate_dematcher_impl.cpp:46-257
#pragma once
#include "../../include/fec_b200.h"
#include <cuda_runtime.h>
#include <stdint.h>

#define FEC_MAX_EDGES 316
#define FEC_MAX_ROWS 46
#define FEC_NOF_GRAPHS 102
#define FEC_CRC_X32_ENTRIES 272 // >= 8448/32 + 1
#define FEC_CRC_X8_ENTRIES 1064 // >= 8448/8 + 1

// All graph data in one HBM-resident blob (uploaded once at engine creation, L2-resident afterwards).
struct fec_dev_tables {
  uint16_t shift[FEC_NOF_GRAPHS][FEC_MAX_EDGES]; // already reduced mod Z
  uint8_t  col[2][FEC_MAX_EDGES];
  uint8_t  deg[2][FEC_MAX_ROWS + 2];
  uint16_t row_start[2][FEC_MAX_ROWS + 2];
  uint8_t  scale_lut[128];                        // lut[x] = round(x * sf) for x <= 120, x otherwise
  uint32_t crc_x32[3][FEC_CRC_X32_ENTRIES];       // x^(32 j) mod g, indexed by FEC_CRC16/24B/24A
  uint32_t crc_x8[3][FEC_CRC_X8_ENTRIES];         // x^(8 j) mod g (byte-granular variant used by the packed kernel)
  uint32_t crc_poly[3];                           // generator with the top bit
  uint32_t crc_len[3];
};

__device__ __forceinline__ int fec_clampi(int v, int lo, int hi)
{
  return max(lo, min(v, hi));
}

// x + y with the reference's special cases (x is the accumulator, y the new LLR). Branch-free: later rules override
// earlier ones in the order of log_likelihood_ratio.cpp:40-73 (equal-and-opposite first, then y infinite, then x infinite).
__device__ __forceinline__ int fec_sat_add(int x, int y)
{
  int r = fec_clampi(x + y, -120, 120);
  r     = (x == 127 || x == -127) ? x : r;
  r     = (y == 127 || y == -127) ? y : r;
  return (y == -x) ? 0 : r;
}

// One step of polynomial long division: r <- (r * x) mod g, r of degree < len.
__device__ __forceinline__ uint32_t fec_crc_shift1(uint32_t r, uint32_t poly, uint32_t top)
{
  r <<= 1;
  return (r & top) ? (r ^ poly) : r;
}

// (a(x) * w(x)) mod g for a 32-bit w (MSB = highest degree), a of degree < len.
__device__ __forceinline__ uint32_t fec_crc_mulmod32(uint32_t a, uint32_t w, uint32_t poly, uint32_t top)
{
  uint32_t acc = 0;
#pragma unroll 8
  for (int i = 0; i < 32; ++i) {
    acc ^= (w & 1u) ? a : 0u;
    w >>= 1;
    a = fec_crc_shift1(a, poly, top);
  }
  return acc;
}
