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
//   1. warp 0 jumps the two 31-bit LFSRs from the call's first soft bit (states prepared on the host, in the descriptor) to
//      the first bit of the tile (matrix powers A^(2^k), one lane per state bit, one product per set bit of the distance),
//   2. threads 0 .. 8 Qm - 1 expand the tile's scrambling bits, one 32-bit word each: the XOR of 16 table rows selected by the
//      nibbles of the combined 62-bit LFSR state (the table holds e0^T A^q, q < 2048, summed per nibble value),
//   3. each thread demodulates its symbol (code specialised per modulation and flavour, tables in shared memory), flips
//      signs four LLRs at a time, and parks the Qm bytes in shared memory,
//   4. the tile is written out in 16-byte rows (shared memory staged at the destination's 16-byte phase).
// HBM traffic: 12 bytes in (symbol + noise variance) and Qm bytes out per symbol.
#pragma once
#include "../../include/fec_b200.h"
#include <cstdint>
#include <cuda_runtime.h>

#define FEC_DEMOD_THREADS 256
#define FEC_DEMOD_TILE_WORDS 64 /* 256 symbols x 8 bits */
static_assert(sizeof(fec_demod_desc) == 40, "fec_demod_desc layout");

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
  // Sequence expansion: word w of the tile (bits c(p + 32 w ...)) = XOR over the 16 nibbles of the 62-bit state (x1 | x2 << 31)
  // of L[nibble position][nibble value][w].
  uint32_t L[16][16][FEC_DEMOD_TILE_WORDS];
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
    // Row vectors e0^T A^q: x(p + q) = parity(row_q & state). Column j of the 32-bit output word w collects bit q = 32 w + i.
    uint32_t r = 1u;
    for (int q = 0; q < FEC_DEMOD_TILE_WORDS * 32; ++q) {
      for (int j = 0; j < 31; ++j) {
        if ((r >> j) & 1u) {
          // state bit j of LFSR g is bit (31 g + j) of the combined state: nibble (31 g + j) / 4, value bit (31 g + j) % 4
          const int sb = 31 * g + j;
          for (int v = 0; v < 16; ++v) {
            if ((v >> (sb & 3)) & 1) {
              t->L[sb >> 2][v][q >> 5] ^= 1u << (q & 31);
            }
          }
        }
      }
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

// LFSR states (x1, x2) at sequence position `pos` are prepared on the host: the same 2^k-step matrices by COLUMN, so that a
// jump is one XOR per set state bit.
struct fec_demod_host_jump {
  uint32_t col[2][32][32];
};

inline void fec_demod_build_jump(const fec_demod_tables* t, fec_demod_host_jump* j)
{
  memset(j, 0, sizeof(*j));
  for (int g = 0; g < 2; ++g) {
    for (int k = 0; k < 32; ++k) {
      for (int r = 0; r < 31; ++r) {
        for (int c = 0; c < 31; ++c) {
          j->col[g][k][c] |= ((t->A[g][k][r] >> c) & 1u) << r;
        }
      }
    }
  }
}

// c(pos) = bit 0 of x1 ^ x2: 1600 + pos steps from the initial states (x1 = 1, x2 = c_init).
inline void fec_demod_lfsr_states(const fec_demod_host_jump* j, uint32_t c_init, uint32_t pos, uint32_t* x1, uint32_t* x2)
{
  uint32_t s[2] = {1u, c_init & 0x7fffffffu};
  for (uint64_t p = 1600ull + pos, k = 0; p != 0; p >>= 1, ++k) {
    if (p & 1u) {
      for (int g = 0; g < 2; ++g) {
        uint32_t nx = 0;
        for (uint32_t v = s[g]; v != 0; v &= v - 1) {
          nx ^= j->col[g][k][__builtin_ctz(v)];
        }
        s[g] = nx;
      }
    }
  }
  *x1 = s[0];
  *x2 = s[1];
}

// ---- device side ------------------------------------------------------------------------------------------------------------
#define FEC_NEAR_ZERO 1e-9f /* include/srsran/support/math/math_utils.h:39 */

// x86 float -> int32 of an already floored value: out of range and NaN give 0x80000000 (CUDA's conversion saturates).
__device__ __forceinline__ int fec_x86_cvt(float f)
{
  return (f >= -2147483648.0f && f < 2147483648.0f) ? __float2int_rd(f) : (int)0x80000000;
}

__device__ __forceinline__ int fec_dm_clamp(int v, int lo, int hi)
{
  return max(lo, min(v, hi));
}

__device__ __forceinline__ float fec_safe_rcp(float n)
{
  return (n > 0) ? __fdiv_rn(1.0f, n) : 0.0f;
}

// log_likelihood_ratio::quantize (log_likelihood_ratio.cpp:90-99): clip, divide, multiply, round half away from zero.
__device__ __forceinline__ int fec_quant_scalar(float v, float range)
{
  float c = v;
  if (fabsf(v) > range) {
    c = copysignf(range, v);
  }
  const float q = roundf(__fmul_rn(__fdiv_rn(c, range), 120.0f));
  return (q != q) ? 0 : (int)q;
}

// mm256::quantize_ps (avx2_helpers.h:123-166) and mm512::quantize_ps (avx512_helpers.h:48-107) are the same function of one
// value: a NaN gives 0 (bounds check after the conversion / masked multiply), otherwise scale, clamp to +-120, round to
// nearest even. scale = 120 / range_limit (6 or 5, exact).
// CHK = false: the caller knows v cannot be a NaN (finite symbol, finite 1 / noise variance).
template <bool CHK>
__device__ __forceinline__ int fec_quant_simd(float v, float scale)
{
  const float q = __fmul_rn(v, scale);
  const int   i = __float2int_rn(fminf(fmaxf(q, -120.0f), 120.0f));
  return (CHK && q != q) ? 0 : i;
}

// Four soft bits (low bytes of a..d) into one word, a in byte 0.
__device__ __forceinline__ uint32_t fec_pack4(int a, int b, int c, int d)
{
  return __byte_perm(__byte_perm((uint32_t)a, (uint32_t)b, 0x0040), __byte_perm((uint32_t)c, (uint32_t)d, 0x0040), 0x5410);
}

enum { FEC_PATH_SCALAR = 0, FEC_PATH_256 = 1, FEC_PATH_512 = 2 };

// Per-CTA copy of the tables a tile needs: [row][interval] = (slope, intercept), then width and 1 / width of each row.
struct fec_demod_smem_tab {
  float2 lin[7][16];
  float  width[7];
  float  rcp_width[7];
};

// compute_interval_idx in its three spellings (demodulation_mapper_intervals.h:36-43, avx2_helpers.h:168-187,
// avx512_helpers.h:156-171). Rows with the same interval width and count share the index. TAME = |x| < 16384: the
// conversion cannot leave the int32 range, so the x86 out-of-range rule needs no emulation.
template <int PATH, int ROW, int N, bool TAME>
__device__ __forceinline__ int fec_interval_idx(const fec_demod_smem_tab& T, float x)
{
  if (PATH == FEC_PATH_512) {
    const float t = __fmaf_rn(x, T.rcp_width[ROW], (float)(N / 2));
    return fec_dm_clamp(TAME ? __float2int_rd(t) : fec_x86_cvt(floorf(t)), 0, N - 1);
  }
  if (PATH == FEC_PATH_256) {
    const float t = __fmul_rn(x, T.rcp_width[ROW]);
    return fec_dm_clamp((int)((uint32_t)(TAME ? __float2int_rd(t) : fec_x86_cvt(floorf(t))) + (uint32_t)(N / 2)), 0, N - 1);
  }
  const float t = __fdiv_rn(x, T.width[ROW]);
  return fec_dm_clamp((TAME ? __float2int_rd(t) : fec_x86_cvt(floorf(t))) + N / 2, 0, N - 1);
}

// interval_function (demodulation_mapper_intervals.h:45-64, avx2_helpers.h:217-249, avx512_helpers.h:214-241) for a known
// interval, then the matching quantiser. `live` = |x| > near_zero (the SIMD paths force 0 otherwise).
template <int PATH, bool FMA, int ROW, bool TAME>
__device__ __forceinline__ int fec_interval_llr(const fec_demod_smem_tab& T, int idx, float x, float rcp, bool live)
{
  const float2 sc = T.lin[ROW][idx];
  float        l  = (FMA || PATH == FEC_PATH_512) ? __fmaf_rn(sc.x, x, sc.y) : __fadd_rn(__fmul_rn(sc.x, x), sc.y);
  l               = __fmul_rn(l, rcp);
  if (PATH == FEC_PATH_SCALAR) {
    return fec_quant_scalar(l, 20.0f);
  }
  return fec_quant_simd<!TAME>(live ? l : 0.0f, 6.0f);
}

// All QM LLRs of one symbol (before descrambling), packed little-endian into two 32-bit words (LLR j = byte j).
// rcp = safe 1 / nv (used by the SIMD paths and the interval functions). TAME: both components below 16384 in magnitude
// and rcp finite, so neither the x86 conversion rule nor a NaN can come into play on the SIMD paths.
template <int QM, int PATH, bool FMA, bool TAME>
__device__ __forceinline__ uint2 fec_demod_symbol(const fec_demod_smem_tab& T, float gain16, float thr16, bool pi2_odd, float re,
                                                  float im, float nv, float rcp)
{
  const float gain2 = 2.0f * 1.41421356237309504880f; // 2.0F * M_SQRT2f32
  int         o[8]  = {0, 0, 0, 0, 0, 0, 0, 0};
  if (QM == 1) {
    // demodulation_mapper_impl.cpp:33-75
    const float s = pi2_odd ? __fadd_rn(im, -re) : __fadd_rn(re, im);
    o[0]          = !(nv > 0) ? 0 : fec_quant_scalar(__fdiv_rn(__fmul_rn(gain2, s), nv), 24.0f);
  } else if (QM == 2) {
    // demodulation_mapper_qpsk.cpp:38-80 (AVX2), 119-127 (scalar)
    if (PATH == FEC_PATH_256) {
      o[0] = fec_quant_simd<!TAME>(__fmul_rn(__fmul_rn(gain2, re), rcp), 5.0f);
      o[1] = fec_quant_simd<!TAME>(__fmul_rn(__fmul_rn(gain2, im), rcp), 5.0f);
    } else if (nv > 0) {
      o[0] = fec_quant_scalar(__fdiv_rn(__fmul_rn(gain2, re), nv), 24.0f);
      o[1] = fec_quant_scalar(__fdiv_rn(__fmul_rn(gain2, im), nv), 24.0f);
    }
  } else {
    const bool near_zero =
        (FMA ? __fmaf_rn(re, re, __fmul_rn(im, im)) : __fadd_rn(__fmul_rn(re, re), __fmul_rn(im, im))) < FEC_NEAR_ZERO;
    const float xs[2] = {re, im};
    if (QM == 4) {
      // demodulation_mapper_qam16.cpp:42-121 (AVX2), 200-275 (scalar)
      if (PATH == FEC_PATH_256) {
#pragma unroll
        for (int c = 0; c < 2; ++c) {
          const float x     = xs[c];
          const float first = __fmul_rn(gain16, x);
          float       l01   = first;
          if (fabsf(x) > thr16) {
            l01 = __fsub_rn(__fmul_rn(2.0f, first), copysignf(0.8f, x));
          }
          float l23 = __fsub_rn(0.8f, fabsf(first));
          l01       = __fmul_rn(l01, rcp);
          l23       = __fmul_rn(l23, rcp);
          if (fabsf(x) <= FEC_NEAR_ZERO) {
            l01 = 0.0f;
            l23 = 0.0f;
          }
          o[c]     = fec_quant_simd<!TAME>(l01, 6.0f);
          o[2 + c] = fec_quant_simd<!TAME>(l23, 6.0f);
        }
      } else if (!near_zero && nv > 0) {
#pragma unroll
        for (int c = 0; c < 2; ++c) {
          const float x = xs[c];
          float       l = __fmul_rn(gain16, x);
          if (fabsf(x) > thr16) {
            l = __fsub_rn(__fmul_rn(2.0f, l), copysignf(0.8f, x));
          }
          o[c]          = fec_quant_scalar(__fdiv_rn(l, nv), 20.0f);
          const float m = FMA ? __fmaf_rn(-gain16, fabsf(x), 0.8f) : __fsub_rn(0.8f, __fmul_rn(gain16, fabsf(x)));
          o[2 + c]      = fec_quant_scalar(__fdiv_rn(m, nv), 20.0f);
        }
      }
    } else if (!(PATH == FEC_PATH_SCALAR && near_zero)) {
      // 64QAM / 256QAM: demodulation_mapper_qam64.cpp:80-182, 184-279, 360-463; _qam256.cpp:163-209, 211-261, 319-425
#pragma unroll
      for (int c = 0; c < 2; ++c) {
        const float x = xs[c];
        // avx512: !(|x| > near_zero) is forced to 0; avx2: |x| <= near_zero is (a NaN passes and ends as 0 in the quantiser)
        const bool live = (PATH == FEC_PATH_512) ? (fabsf(x) > FEC_NEAR_ZERO) : !(fabsf(x) <= FEC_NEAR_ZERO);
        if (QM == 6) {
          const int i2 = fec_interval_idx<PATH, 0, 8, TAME>(T, x), i4 = fec_interval_idx<PATH, 2, 4, TAME>(T, x);
          o[c]         = fec_interval_llr<PATH, FMA, 0, TAME>(T, i2, x, rcp, live);
          o[2 + c]     = fec_interval_llr<PATH, FMA, 1, TAME>(T, i2, x, rcp, live);
          o[4 + c]     = fec_interval_llr<PATH, FMA, 2, TAME>(T, i4, x, rcp, live);
        } else {
          const int i2 = fec_interval_idx<PATH, 3, 16, TAME>(T, x), i4 = fec_interval_idx<PATH, 6, 8, TAME>(T, x);
          o[c]         = fec_interval_llr<PATH, FMA, 3, TAME>(T, i2, x, rcp, live);
          o[2 + c]     = fec_interval_llr<PATH, FMA, 4, TAME>(T, i2, x, rcp, live);
          o[4 + c]     = fec_interval_llr<PATH, FMA, 5, TAME>(T, i2, x, rcp, live);
          o[6 + c]     = fec_interval_llr<PATH, FMA, 6, TAME>(T, i4, x, rcp, live);
        }
      }
    }
  }
  uint2 r;
  r.x = fec_pack4(o[0], o[1], o[2], o[3]);
  r.y = (QM > 4) ? fec_pack4(o[4], o[5], o[6], o[7]) : 0u;
  return r;
}

// One symbol on one path: 1 / noise variance where the path uses it, then the common case (TAME) without the emulation of
// x86's out-of-range conversions and without NaN checks, anything else through the fully general instantiation.
template <int QM, int PATH, bool FMA>
__device__ __forceinline__ uint2 fec_demod_dispatch(const fec_demod_smem_tab& T, float gain16, float thr16, bool pi2_odd, float re,
                                                    float im, float nv)
{
  if (QM == 1 || (PATH == FEC_PATH_SCALAR && QM <= 4)) {
    return fec_demod_symbol<QM, PATH, FMA, false>(T, gain16, thr16, pi2_odd, re, im, nv, 0.0f); // these divide by nv itself
  }
  const float rcp = fec_safe_rcp(nv);
  if (fabsf(re) < 16384.0f && fabsf(im) < 16384.0f && rcp <= 3.402823466e+38f) {
    return fec_demod_symbol<QM, PATH, FMA, true>(T, gain16, thr16, pi2_odd, re, im, nv, rcp);
  }
  return fec_demod_symbol<QM, PATH, FMA, false>(T, gain16, thr16, pi2_odd, re, im, nv, rcp);
}

// Negates the bytes of w whose mask bit is set (two's complement per byte, |LLR| <= 120 so nothing overflows).
__device__ __forceinline__ uint32_t fec_negate_bytes(uint32_t w, uint32_t bits4)
{
  const uint32_t m = ((bits4 & 1u) * 0xffu) | ((bits4 & 2u) * (0xff00u >> 1)) | ((bits4 & 4u) * (0xff0000u >> 2)) |
                     ((bits4 & 8u) * (0xff000000u >> 3));
  return __vsub4(w ^ m, m); // (w ^ m) - m per byte: -w where m = 0xff
}

template <int QM, int FLAVOUR>
__device__ __forceinline__ void fec_demod_tile(const fec_demod_desc& d, uint32_t tile, uint32_t nsym, float2 s, float nv,
                                               const fec_demod_smem_tab& T, float gain16, float thr16, const uint32_t* seq,
                                               bool scramble, uint8_t* stage_at_phase)
{
  const uint32_t tid = threadIdx.x;
  if (tid >= nsym) {
    return;
  }
  const uint32_t     i    = tile + tid; // position in the demodulate_soft call
  constexpr uint32_t V256 = (FLAVOUR >= 1) ? ((QM == 2) ? 16u : (QM == 4) ? 8u : (QM == 6) ? 16u : (QM == 8) ? 4u : 0u) : 0u;
  constexpr uint32_t V512 = (FLAVOUR >= 2) ? ((QM == 6) ? 32u : (QM == 8) ? 8u : 0u) : 0u;
  constexpr bool     FMA  = FLAVOUR >= 1;
  const bool         odd  = (d.flags & FEC_DEMOD_FLAG_PI2_BPSK) && (i & 1u);
  uint2              r;
  if (V512 != 0 && i < d.nof_symbols / (V512 ? V512 : 1u) * V512) {
    r = fec_demod_dispatch<QM, FEC_PATH_512, FMA>(T, gain16, thr16, odd, s.x, s.y, nv);
  } else if (V256 != 0 && i < d.nof_symbols / (V256 ? V256 : 1u) * V256) {
    r = fec_demod_dispatch<QM, FEC_PATH_256, FMA>(T, gain16, thr16, odd, s.x, s.y, nv);
  } else {
    r = fec_demod_dispatch<QM, FEC_PATH_SCALAR, FMA>(T, gain16, thr16, odd, s.x, s.y, nv);
  }
  if (scramble) {
    const uint32_t b    = tid * QM; // QM <= 8 bits from bit b of the tile's sequence
    const uint32_t bits = (uint32_t)((((uint64_t)seq[(b >> 5) + 1] << 32) | seq[b >> 5]) >> (b & 31u));
    r.x                 = fec_negate_bytes(r.x, bits & 0xfu);
    if (QM > 4) {
      r.y = fec_negate_bytes(r.y, (bits >> 4) & 0xfu);
    }
  }
  uint8_t* dst = stage_at_phase + tid * QM;
#pragma unroll
  for (int j = 0; j < QM; ++j) {
    dst[j] = (uint8_t)(((j < 4) ? r.x : r.y) >> (8 * (j & 3)));
  }
}

__global__ void __launch_bounds__(FEC_DEMOD_THREADS)
    fec_demod_kernel(const fec_demod_desc* __restrict__ descs, const float2* __restrict__ symbols, const float* __restrict__ noise,
                     int8_t* __restrict__ llr, const fec_demod_tables* __restrict__ T, uint32_t flavour)
{
  __shared__ uint32_t             seq[FEC_DEMOD_TILE_WORDS + 1];
  __shared__ uint32_t             state[2];
  __shared__ fec_demod_smem_tab   tab;
  __shared__ __align__(16) uint8_t stage[FEC_DEMOD_THREADS * 8 + 16];

  const fec_demod_desc d    = descs[blockIdx.x];
  const uint32_t       tile = blockIdx.y * FEC_DEMOD_THREADS;
  if (tile >= d.nof_symbols) {
    return;
  }
  const uint32_t qm = d.qm, tid = threadIdx.x, lane = tid & 31u;
  const uint32_t nsym     = min(d.nof_symbols - tile, (uint32_t)FEC_DEMOD_THREADS);
  const uint32_t nbits    = nsym * qm;
  const bool     scramble = !(d.flags & FEC_DEMOD_FLAG_NO_DESCRAMBLE);
  // The thread's symbol is requested first: its trip to HBM overlaps the sequence work below.
  float2 sym = make_float2(0.0f, 0.0f);
  float  nv  = 0.0f;
  if (tid < nsym) {
    sym = symbols[d.sym_offset + tile + tid];
    nv  = noise[d.sym_offset + tile + tid];
  }
  if (qm >= 6 && tid < 7 * 16) {
    const uint32_t row = tid >> 4, k = tid & 15u;
    tab.lin[row][k]    = make_float2(T->slope[row][k], T->icpt[row][k]);
    if (k == 0) {
      tab.width[row]     = T->width[row];
      tab.rcp_width[row] = T->rcp_width[row];
    }
  }
  if (scramble) {
    if (tid < 32) {
      // The descriptor carries the LFSR states at the call's first soft bit (fec_demod_desc_prepare); the tile starts
      // tile * qm bits further on: one matrix-vector product per set bit of that distance, one lane per state bit.
      uint32_t s1 = d.x1_state, s2 = d.x2_state;
      for (uint32_t p = tile * qm; p != 0; p &= p - 1) {
        const uint32_t k = __ffs(p) - 1;
        s1 = __ballot_sync(0xffffffffu, lane < 31 && (__popc(T->A[0][k][lane] & s1) & 1)) & 0x7fffffffu;
        s2 = __ballot_sync(0xffffffffu, lane < 31 && (__popc(T->A[1][k][lane] & s2) & 1)) & 0x7fffffffu;
      }
      if (tid == 0) {
        state[0]                  = s1;
        state[1]                  = s2;
        seq[FEC_DEMOD_TILE_WORDS] = 0;
      }
    }
    __syncthreads();
    // Word w of the tile's sequence = XOR of one table row per nibble of the combined state: threads 0 .. 8 qm - 1, one word
    // each (a full tile holds 256 * qm bits; where the call ends earlier the surplus is never read).
    if (tid < 8 * qm) {
      const uint64_t S  = (uint64_t)state[0] | ((uint64_t)state[1] << 31);
      uint32_t       wd = 0;
#pragma unroll
      for (int n = 0; n < 16; ++n) {
        wd ^= T->L[n][(uint32_t)(S >> (4 * n)) & 15u][tid];
      }
      seq[tid] = wd;
    }
  }
  __syncthreads();
  int8_t*        dst   = llr + d.llr_offset + (uint64_t)tile * qm;
  const uint32_t phase = (uint32_t)(reinterpret_cast<uintptr_t>(dst) & 15u);
  const float    g16 = T->gain16, t16 = T->thr16;
#define FEC_DEMOD_CASE(Q, F)                                                                                                  \
  case (Q) * 4 + (F):                                                                                                          \
    fec_demod_tile<Q, F>(d, tile, nsym, sym, nv, tab, g16, t16, seq, scramble, stage + phase);                          \
    break;
  switch (qm * 4 + flavour) {
    FEC_DEMOD_CASE(1, 0) FEC_DEMOD_CASE(1, 1) FEC_DEMOD_CASE(1, 2)
    FEC_DEMOD_CASE(2, 0) FEC_DEMOD_CASE(2, 1) FEC_DEMOD_CASE(2, 2)
    FEC_DEMOD_CASE(4, 0) FEC_DEMOD_CASE(4, 1) FEC_DEMOD_CASE(4, 2)
    FEC_DEMOD_CASE(6, 0) FEC_DEMOD_CASE(6, 1) FEC_DEMOD_CASE(6, 2)
    FEC_DEMOD_CASE(8, 0) FEC_DEMOD_CASE(8, 1) FEC_DEMOD_CASE(8, 2)
    default:
      break;
  }
#undef FEC_DEMOD_CASE
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
