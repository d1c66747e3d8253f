// fec_demod.cuh -- the step in front of the decode path: soft demodulation of equalised symbols into
// int8 log-likelihood ratios and descrambling, on the device, so that the LLRs never exist on the host.
//
// One descriptor = one demodulation_mapper::demodulate_soft call of the reference (pusch_demodulator_impl.cpp:355-375: one
// call per OFDM symbol), followed by revert_scrambling with the running TS 38.211 5.2.1 sequence (ibid. 39-207, 377-382).
// The reference computes an LLR with one of three instruction sequences, chosen when it is COMPILED (AVX-512, AVX2 or plain
// scalar code; lib/phy/upper/channel_modulation/demodulation_mapper_{qpsk,qam16,qam64,qam256}.cpp) and by the position of
// the symbol in the call (SIMD loops first, scalar tail). They round differently, so the kernel carries all three and a
// `flavour` argument says which build of the reference it stands in for. Every arithmetic step is a single IEEE-754
// operation spelled with an intrinsic (__fmul_rn, __fmaf_rn, __fdiv_rn ...): nvcc must neither fuse nor split anything.
//
// Work layout: one CTA per tile of 256 consecutive symbols of one call, one thread per symbol.
//   1. warp 0 jumps the two 31-bit LFSRs to the first bit of the tile (matrix powers A^(2^k), one lane per state bit),
//   2. all threads expand the tile's scrambling bits (row vectors B[q] = e0^T A^q, one parity per bit, ballot per word),
//   3. each thread demodulates its symbol, flips signs, and parks the Qm bytes in shared memory,
//   4. the tile is written out in 16-byte rows (shared memory staged at the destination's 16-byte phase).
// HBM traffic: 12 bytes in (symbol + noise variance) and Qm bytes out per symbol.
#pragma once
#include "../../include/fec_b200.h"
#include <cstdint>
#include <cuda_runtime.h>

#define FEC_DEMOD_THREADS 256
#define FEC_DEMOD_TILE_WORDS 64 /* 256 symbols x 8 bits */

struct fec_demod_tables {
  // piecewise-linear max-log LLR, rows 0-2: 64QAM bit pairs (0,1) (2,3) (4,5); rows 3-6: 256QAM bit pairs
  float    slope[7][16];
  float    icpt[7][16];
  float    width[7];
  float    rcp_width[7];
  int32_t  nof[7];
  float    gain16;  // 4 / sqrt(10)
  float    thr16;   // 2 / sqrt(10)
  uint32_t A[2][32][32];                       // [lfsr][k][state bit]: row of the 2^k-step transition matrix
  uint32_t B[2][FEC_DEMOD_TILE_WORDS * 32];    // [lfsr][q]: mask whose parity with the state gives x(p + q)
};

// ---- host side: tables derived from the constellation (TS 38.211 5.1.4-5.1.6) and the generator polynomials (5.2.1) -----
inline int fec_demod_pam_level(const int* bits, int m)
{
  // d = (1-2b0)[2^(m-1) - (1-2b1)[2^(m-2) - ... [2 - (1-2b_(m-1))]]]
  int acc = 1 - 2 * bits[m - 1];
  for (int i = m - 2; i >= 0; --i) {
    acc = (1 - 2 * bits[i]) * ((1 << (m - 1 - i)) - acc);
  }
  return acc;
}

inline void fec_demod_build_tables(fec_demod_tables* t)
{
  memset(t, 0, sizeof(*t));
  const float m10 = 1.0f / sqrtf(10.0f);
  t->gain16       = 4 * m10;
  t->thr16        = 2 * m10;
  int row         = 0;
  for (int m = 3; m <= 4; ++m) {
    const float norm = 1.0f / sqrtf(m == 3 ? 42.0f : 170.0f);
    const float den  = (m == 3) ? 21.0f : 85.0f; // (k1^2 - k0^2) / (2 * 42 or 170 / 2)
    for (int bit = 0; bit < m; ++bit, ++row) {
      // In every interval the max-log LLR is ((x - a1)^2 - (x - a0)^2) / noise with a0 / a1 the nearest levels whose bit is
      // 0 / 1: slope 2 (a0 - a1), intercept a1^2 - a0^2. The last bit changes its nearest pair every 4 levels, the others
      // are tabulated every 2.
      const int wmult = (bit == m - 1) ? 4 : 2;
      const int n     = (2 << m) / wmult;
      t->nof[row]     = n;
      t->width[row]   = (float)wmult * norm;
      t->rcp_width[row] = 1.0f / t->width[row];
      for (int j = 0; j < n; ++j) {
        const float mid = ((float)j - (float)n / 2 + 0.5f) * (float)wmult;
        int         k[2] = {0, 0};
        float       best[2] = {1e9f, 1e9f};
        for (int v = 0; v < (1 << m); ++v) {
          int bits[4];
          for (int i = 0; i < m; ++i) {
            bits[i] = (v >> (m - 1 - i)) & 1;
          }
          const int   lvl = fec_demod_pam_level(bits, m);
          const float d   = fabsf((float)lvl - mid);
          if (d < best[bits[bit]]) {
            best[bits[bit]] = d;
            k[bits[bit]]    = lvl;
          }
        }
        t->slope[row][j] = (float)(2 * (k[0] - k[1])) * norm;
        t->icpt[row][j]  = (float)((k[1] * k[1] - k[0] * k[0]) / 2) / den;
      }
    }
  }
  // x1(n+31) = x1(n+3) + x1(n), x2(n+31) = x2(n+3) + x2(n+2) + x2(n+1) + x2(n); state bit i = x(p + i).
  const uint32_t fb[2] = {0x9u, 0xfu};
  for (int g = 0; g < 2; ++g) {
    uint32_t(*A)[32] = t->A[g];
    for (int i = 0; i < 30; ++i) {
      A[0][i] = 1u << (i + 1);
    }
    A[0][30] = fb[g];
    for (int k = 1; k < 32; ++k) {
      for (int i = 0; i < 31; ++i) {
        uint32_t r = 0, row_i = A[k - 1][i];
        for (int j = 0; j < 31; ++j) {
          if ((row_i >> j) & 1u) {
            r ^= A[k - 1][j];
          }
        }
        A[k][i] = r;
      }
    }
    uint32_t r = 1u; // e0^T A^0
    for (int q = 0; q < FEC_DEMOD_TILE_WORDS * 32; ++q) {
      t->B[g][q] = r;
      uint32_t nx = 0;
      for (int j = 0; j < 31; ++j) {
        if ((r >> j) & 1u) {
          nx ^= A[0][j];
        }
      }
      r = nx;
    }
  }
}

// ---- device side ------------------------------------------------------------------------------------------------------------
#define FEC_NEAR_ZERO 1e-9f /* include/srsran/support/math/math_utils.h:39 */

// x86 float -> int32: out of range and NaN give 0x80000000 (CUDA's own conversion saturates instead).
__device__ __forceinline__ int fec_cvt_i32(float f)
{
  return (f >= -2147483648.0f && f < 2147483648.0f) ? (int)f : (int)0x80000000;
}

__device__ __forceinline__ int fec_dm_clamp(int v, int lo, int hi)
{
  return v < lo ? lo : (v > hi ? hi : v);
}

__device__ __forceinline__ float fec_safe_rcp(float n)
{
  return (n > 0) ? __fdiv_rn(1.0f, n) : 0.0f;
}

// log_likelihood_ratio::quantize (log_likelihood_ratio.cpp:90-99).
__device__ __forceinline__ int fec_quant_scalar(float v, float range)
{
  float c = v;
  if (fabsf(v) > range) {
    c = copysignf(range, v);
  }
  float q = __fdiv_rn(c, range);
  q       = __fmul_rn(q, 120.0f);
  q       = roundf(q);
  return (q != q) ? 0 : (int)q;
}

// mm256::quantize_ps (avx2_helpers.h:123-166). scale = 120 / range_limit (6 or 5, exact).
__device__ __forceinline__ int fec_quant_256(float v, float scale)
{
  float q = __fmul_rn(v, scale);
  if (q > 120.0f) {
    q = 120.0f;
  }
  if (q < -120.0f) {
    q = -120.0f;
  }
  int i = fec_cvt_i32(rintf(q));
  return (i > 120 || i < -120) ? 0 : i;
}

// mm512::quantize_ps (avx512_helpers.h:48-107).
__device__ __forceinline__ int fec_quant_512(float v, float scale)
{
  if (v != v) {
    return 0;
  }
  float q = __fmul_rn(scale, v);
  q       = q < 120.0f ? q : 120.0f;
  q       = q > -120.0f ? q : -120.0f;
  return fec_cvt_i32(rintf(q));
}

enum { FEC_PATH_SCALAR = 0, FEC_PATH_256 = 1, FEC_PATH_512 = 2 };

// interval_function in its three spellings (demodulation_mapper_intervals.h:36-64, avx2_helpers.h:168-249,
// avx512_helpers.h:156-241), then the matching quantiser.
__device__ __forceinline__ int fec_interval_llr(const fec_demod_tables* __restrict__ T, int row, int path, bool fma, float x,
                                                float rcp)
{
  const int n = T->nof[row];
  int       idx;
  if (path == FEC_PATH_512) {
    idx = fec_dm_clamp(fec_cvt_i32(floorf(__fmaf_rn(x, T->rcp_width[row], (float)(n / 2)))), 0, n - 1);
  } else if (path == FEC_PATH_256) {
    idx = fec_dm_clamp((int)((uint32_t)fec_cvt_i32(floorf(__fmul_rn(x, T->rcp_width[row]))) + (uint32_t)(n / 2)), 0, n - 1);
  } else {
    idx = fec_dm_clamp(fec_cvt_i32(floorf(__fdiv_rn(x, T->width[row]))) + n / 2, 0, n - 1);
  }
  const float s = T->slope[row][idx], c = T->icpt[row][idx];
  float       l = (fma || path == FEC_PATH_512) ? __fmaf_rn(s, x, c) : __fadd_rn(__fmul_rn(s, x), c);
  l             = __fmul_rn(l, rcp);
  if (path == FEC_PATH_512) {
    return fec_quant_512((fabsf(x) > FEC_NEAR_ZERO) ? l : 0.0f, 6.0f);
  }
  if (path == FEC_PATH_256) {
    return fec_quant_256((fabsf(x) <= FEC_NEAR_ZERO) ? 0.0f : l, 6.0f);
  }
  return fec_quant_scalar(l, 20.0f);
}

// All Qm LLRs of one symbol (before descrambling) into out[0 .. qm).
__device__ __forceinline__ void fec_demod_symbol(const fec_demod_tables* __restrict__ T, int qm, bool pi2_odd, int path, bool fma,
                                                 float re, float im, float nv, int* out)
{
  const float gain2 = 2.0f * 1.41421356237309504880f; // 2.0F * M_SQRT2f32
  if (qm == 1) {
    // demodulation_mapper_impl.cpp:33-75
    const float s = pi2_odd ? __fadd_rn(im, -re) : __fadd_rn(re, im);
    out[0]        = !(nv > 0) ? 0 : fec_quant_scalar(__fdiv_rn(__fmul_rn(gain2, s), nv), 24.0f);
    return;
  }
  const float xs[2] = {re, im};
  if (qm == 2) {
    // demodulation_mapper_qpsk.cpp:38-80 (AVX2), 119-127 (scalar)
#pragma unroll
    for (int c = 0; c < 2; ++c) {
      if (path == FEC_PATH_256) {
        out[c] = fec_quant_256(__fmul_rn(__fmul_rn(gain2, xs[c]), fec_safe_rcp(nv)), 5.0f);
      } else {
        out[c] = !(nv > 0) ? 0 : fec_quant_scalar(__fdiv_rn(__fmul_rn(gain2, xs[c]), nv), 24.0f);
      }
    }
    return;
  }
  const bool near_zero =
      (fma ? __fmaf_rn(re, re, __fmul_rn(im, im)) : __fadd_rn(__fmul_rn(re, re), __fmul_rn(im, im))) < FEC_NEAR_ZERO;
  if (qm == 4) {
    // demodulation_mapper_qam16.cpp:42-121 (AVX2), 200-275 (scalar)
    if (path == FEC_PATH_256) {
      const float rcp = fec_safe_rcp(nv);
#pragma unroll
      for (int c = 0; c < 2; ++c) {
        const float x     = xs[c];
        const float first = __fmul_rn(T->gain16, x);
        float       l01   = first;
        if (fabsf(x) > T->thr16) {
          l01 = __fsub_rn(__fmul_rn(2.0f, first), copysignf(0.8f, x));
        }
        float l23 = __fsub_rn(0.8f, fabsf(first));
        l01       = __fmul_rn(l01, rcp);
        l23       = __fmul_rn(l23, rcp);
        if (fabsf(x) <= FEC_NEAR_ZERO) {
          l01 = 0.0f;
          l23 = 0.0f;
        }
        out[c]     = fec_quant_256(l01, 6.0f);
        out[2 + c] = fec_quant_256(l23, 6.0f);
      }
    } else {
#pragma unroll
      for (int c = 0; c < 2; ++c) {
        const float x = xs[c];
        if (near_zero || !(nv > 0)) {
          out[c] = out[2 + c] = 0;
          continue;
        }
        float l = __fmul_rn(T->gain16, x);
        if (fabsf(x) > T->thr16) {
          l = __fsub_rn(__fmul_rn(2.0f, l), copysignf(0.8f, x));
        }
        out[c]        = fec_quant_scalar(__fdiv_rn(l, nv), 20.0f);
        const float m = fma ? __fmaf_rn(-T->gain16, fabsf(x), 0.8f) : __fsub_rn(0.8f, __fmul_rn(T->gain16, fabsf(x)));
        out[2 + c]    = fec_quant_scalar(__fdiv_rn(m, nv), 20.0f);
      }
    }
    return;
  }
  // 64QAM / 256QAM: demodulation_mapper_qam64.cpp:80-182, 184-279, 360-463; _qam256.cpp:163-209, 211-261, 319-425
  const int row0 = (qm == 6) ? 0 : 3, nt = qm / 2;
  if (path == FEC_PATH_SCALAR && near_zero) {
    for (int j = 0; j < qm; ++j) {
      out[j] = 0;
    }
    return;
  }
  const float rcp = fec_safe_rcp(nv);
  for (int t = 0; t < nt; ++t) {
    out[2 * t]     = fec_interval_llr(T, row0 + t, path, fma, re, rcp);
    out[2 * t + 1] = fec_interval_llr(T, row0 + t, path, fma, im, rcp);
  }
}

// Symbols per iteration of the reference's AVX-512 / AVX2 loops, by flavour (see the header comment).
__host__ __device__ __forceinline__ void fec_demod_blocks(uint32_t flavour, uint32_t qm, uint32_t* v512, uint32_t* v256)
{
  *v256 = (flavour >= 1) ? ((qm == 2) ? 16u : (qm == 4) ? 8u : (qm == 6) ? 16u : (qm == 8) ? 4u : 0u) : 0u;
  *v512 = (flavour >= 2) ? ((qm == 6) ? 32u : (qm == 8) ? 8u : 0u) : 0u;
}

__global__ void __launch_bounds__(FEC_DEMOD_THREADS)
    fec_demod_kernel(const fec_demod_desc* __restrict__ descs, const float2* __restrict__ symbols, const float* __restrict__ noise,
                     int8_t* __restrict__ llr, const fec_demod_tables* __restrict__ T, uint32_t flavour)
{
  __shared__ uint32_t seq[FEC_DEMOD_TILE_WORDS];
  __shared__ uint32_t state[2];
  __shared__ __align__(16) uint8_t stage[FEC_DEMOD_THREADS * 8 + 16];

  const fec_demod_desc d    = descs[blockIdx.x];
  const uint32_t       tile = blockIdx.y * FEC_DEMOD_THREADS;
  if (tile >= d.nof_symbols) {
    return;
  }
  const uint32_t qm = d.qm, tid = threadIdx.x, lane = tid & 31u;
  const uint32_t nsym  = min(d.nof_symbols - tile, (uint32_t)FEC_DEMOD_THREADS);
  const uint32_t nbits = nsym * qm;
  const bool     scramble = !(d.flags & FEC_DEMOD_FLAG_NO_DESCRAMBLE);
  if (scramble) {
    if (tid < 32) {
      // c(n) = x1(n + 1600) + x2(n + 1600); x1 starts from 1, x2 from c_init.
      const uint32_t p = 1600u + d.seq_offset + tile * qm;
      uint32_t s1 = 1u, s2 = d.c_init & 0x7fffffffu;
      for (uint32_t k = 0; k < 32; ++k) {
        if ((p >> k) & 1u) {
          s1 = __ballot_sync(0xffffffffu, lane < 31 && (__popc(T->A[0][k][lane] & s1) & 1)) & 0x7fffffffu;
          s2 = __ballot_sync(0xffffffffu, lane < 31 && (__popc(T->A[1][k][lane] & s2) & 1)) & 0x7fffffffu;
        }
      }
      if (tid == 0) {
        state[0] = s1;
        state[1] = s2;
      }
    }
    __syncthreads();
    const uint32_t s1 = state[0], s2 = state[1], nwords = (nbits + 31) / 32;
    for (uint32_t w = tid >> 5; w < nwords; w += FEC_DEMOD_THREADS / 32) {
      const uint32_t q   = w * 32 + lane;
      const uint32_t bit = (__popc(T->B[0][q] & s1) ^ __popc(T->B[1][q] & s2)) & 1u;
      const uint32_t wd  = __ballot_sync(0xffffffffu, bit);
      if (lane == 0) {
        seq[w] = wd;
      }
    }
    __syncthreads();
  }
  int8_t*        dst   = llr + d.llr_offset + (uint64_t)tile * qm;
  const uint32_t phase = (uint32_t)(reinterpret_cast<uintptr_t>(dst) & 15u);
  if (tid < nsym) {
    const uint32_t i = tile + tid; // position in the demodulate_soft call
    uint32_t       v512, v256;
    fec_demod_blocks(flavour, qm, &v512, &v256);
    const uint32_t end512 = v512 ? (d.nof_symbols / v512) * v512 : 0u;
    const uint32_t end256 = v256 ? (d.nof_symbols / v256) * v256 : 0u;
    const int      path   = (i < end512) ? FEC_PATH_512 : (i < end256) ? FEC_PATH_256 : FEC_PATH_SCALAR;
    const float2   s      = symbols[d.sym_offset + i];
    const float    nv     = noise[d.sym_offset + i];
    int            out[8];
    fec_demod_symbol(T, (int)qm, (d.flags & FEC_DEMOD_FLAG_PI2_BPSK) && (i & 1u), path, flavour >= 1, s.x, s.y, nv, out);
#pragma unroll
    for (uint32_t j = 0; j < 8; ++j) {
      if (j < qm) {
        const uint32_t b = tid * qm + j;
        int            v = out[j];
        if (scramble && ((seq[b >> 5] >> (b & 31u)) & 1u)) {
          v = -v;
        }
        stage[phase + b] = (uint8_t)(int8_t)v;
      }
    }
  }
  __syncthreads();
  // Rows of 16 bytes at the destination's alignment; the first and the last row may be partial.
  const uint32_t total = phase + nbits;
  for (uint32_t r = tid; r * 16 < total; r += FEC_DEMOD_THREADS) {
    const uint32_t lo = r * 16, hi = lo + 16;
    if (lo >= phase && hi <= total) {
      *reinterpret_cast<uint4*>(dst - phase + lo) = *reinterpret_cast<const uint4*>(stage + lo);
    } else {
      for (uint32_t b = max(lo, phase); b < min(hi, total); ++b) {
        dst[(int64_t)b - phase] = (int8_t)stage[b];
      }
    }
  }
}
