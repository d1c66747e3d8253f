/* oracle/fec_demod.c -- TEST INFRASTRUCTURE, NOT PRODUCT CODE.
 *
 * CPU restatement of the step in front of the PUSCH decode path: soft demodulation of equalised
 * symbols into int8 log-likelihood ratios and descrambling with the TS 38.211 5.2.1 Gold sequence. Plain scalar C, one
 * IEEE-754 operation per line of arithmetic (compiled with -ffp-contract=off; every fused multiply-add is an explicit
 * fmaf), so that it can be compared bit for bit with the reference compiled here (oracle/_ref/libref_demod_*.so, pinned by
 * tests/test_demod_cpu.py).
 *
 * The reference computes the same LLR with up to three different instruction sequences, chosen at COMPILE time and by the
 * POSITION of the symbol inside one demodulate_soft call: an AVX-512 loop, an AVX2 loop for what is left, and a scalar loop
 * for the tail. They round differently (reciprocal-multiply vs divide, fused vs separate multiply-add, round-half-even vs
 * round-half-away), so each is restated separately:
 *   path 512   lib/phy/upper/channel_modulation/avx512_helpers.h:48-264 (quantize_ps, compute_interval_idx, interval_function)
 *              demodulation_mapper_qam64.cpp:80-182, demodulation_mapper_qam256.cpp:163-209
 *   path 256   lib/phy/upper/channel_modulation/avx2_helpers.h:123-274
 *              demodulation_mapper_qpsk.cpp:38-80, _qam16.cpp:42-121, _qam64.cpp:184-279, _qam256.cpp:211-261
 *   scalar     demodulation_mapper_intervals.h:36-64, log_likelihood_ratio.cpp:90-99 (quantize),
 *              demodulation_mapper_impl.cpp:33-75 (BPSK, pi/2-BPSK), _qpsk.cpp:119-127, _qam16.cpp:200-224,
 *              _qam64.cpp:360-376, _qam256.cpp:319-341, and the tail loops of the four demodulate_soft_* functions
 * Which positions take which path is the `flavour` (what the reference was compiled for):
 *   0 generic : scalar everywhere, no fused multiply-add
 *   1 avx2    : QPSK 16 / QAM16 8 / QAM64 16 / QAM256 4 symbols per AVX2 iteration, scalar tail; g++ contracts a*b+c into
 *               FMA in the scalar code and in the AVX2 helper (the intrinsics are plain vector operators in GCC's headers)
 *   2 avx512  : QAM64 32 / QAM256 8 symbols per AVX-512 iteration first, then as avx2
 */
#include <math.h>
#include <stdint.h>
#include <string.h>

#define NEAR_ZERO 1e-9f /* include/srsran/support/math/math_utils.h:39 */
#define LLR_MAX 120

typedef struct {
  int   n;        /* number of intervals */
  float width;    /* interval width */
  float slope[16];
  float icpt[16];
} itab;

static itab  T64[3], T256[4];
static float M10, M42, M170;
static int   tables_ready;

static void build(itab* t, int n, float width, const int* sl, float m, const int* ic, float den)
{
  memset(t, 0, sizeof(*t));
  t->n     = n;
  t->width = width;
  for (int i = 0; i < n; ++i) {
    t->slope[i] = (float)sl[i] * m;
    t->icpt[i]  = (float)ic[i] / den;
  }
}

static void tables(void)
{
  if (tables_ready) {
    return;
  }
  M10  = 1.0f / sqrtf(10.0f);
  M42  = 1.0f / sqrtf(42.0f);
  M170 = 1.0f / sqrtf(170.0f);
  /* demodulation_mapper_qam64.cpp:48-78 */
  static const int s64_01[8] = {16, 12, 8, 4, 4, 8, 12, 16}, i64_01[8] = {24, 12, 4, 0, 0, -4, -12, -24};
  static const int s64_23[8] = {8, 4, 4, 8, -8, -4, -4, -8}, i64_23[8] = {20, 8, 8, 12, 12, 8, 8, 20};
  static const int s64_45[4] = {4, -4, 4, -4}, i64_45[4] = {12, -4, -4, 12};
  build(&T64[0], 8, 2 * M42, s64_01, M42, i64_01, 21);
  build(&T64[1], 8, 2 * M42, s64_23, M42, i64_23, 21);
  build(&T64[2], 4, 4 * M42, s64_45, M42, i64_45, 21);
  /* demodulation_mapper_qam256.cpp:48-161 */
  static const int s01[16] = {32, 28, 24, 20, 16, 12, 8, 4, 4, 8, 12, 16, 20, 24, 28, 32};
  static const int i01[16] = {112, 84, 60, 40, 24, 12, 4, 0, 0, -4, -12, -24, -40, -60, -84, -112};
  static const int s23[16] = {16, 12, 8, 4, 4, 8, 12, 16, -16, -12, -8, -4, -4, -8, -12, -16};
  static const int i23[16] = {88, 60, 36, 16, 16, 28, 36, 40, 40, 36, 28, 16, 16, 36, 60, 88};
  static const int s45[16] = {8, 4, 4, 8, -8, -4, -4, -8, 8, 4, 4, 8, -8, -4, -4, -8};
  static const int i45[16] = {52, 24, 24, 44, -20, -8, -8, -12, -12, -8, -8, -20, 44, 24, 24, 52};
  static const int s67[8]  = {4, -4, 4, -4, 4, -4, 4, -4};
  static const int i67[8]  = {28, -20, 12, -4, -4, 12, -20, 28};
  build(&T256[0], 16, 2 * M170, s01, M170, i01, 85);
  build(&T256[1], 16, 2 * M170, s23, M170, i23, 85);
  build(&T256[2], 16, 2 * M170, s45, M170, i45, 85);
  build(&T256[3], 8, 4 * M170, s67, M170, i67, 85);
  tables_ready = 1;
}

/* The tables as the device wants them (row = modulation table, 16 slopes then 16 intercepts, then width, 1/width, n). */
void orc_demod_tables(float* out /* [7][36] */)
{
  tables();
  const itab* all[7] = {&T64[0], &T64[1], &T64[2], &T256[0], &T256[1], &T256[2], &T256[3]};
  for (int t = 0; t < 7; ++t) {
    memcpy(out + t * 36, all[t]->slope, 64);
    memcpy(out + t * 36 + 16, all[t]->icpt, 64);
    out[t * 36 + 32] = all[t]->width;
    out[t * 36 + 33] = 1.0f / all[t]->width;
    out[t * 36 + 34] = (float)all[t]->n;
    out[t * 36 + 35] = 0;
  }
}

/* x86 float -> int32 conversion: out-of-range and NaN give the "integer indefinite" value. */
static int32_t cvt_i32(float integral)
{
  if (!(integral >= -2147483648.0f && integral < 2147483648.0f)) {
    return INT32_MIN;
  }
  return (int32_t)integral;
}

static int clampi(int64_t v, int lo, int hi)
{
  return v < lo ? lo : (v > hi ? hi : (int)v);
}

static float safe_rcp(float n)
{
  return (n > 0) ? 1.0f / n : 0.0f;
}

/* ---- quantisers ------------------------------------------------------------------------------------------------------ */
/* log_likelihood_ratio::quantize: clip, divide, multiply, round half away from zero. */
static int8_t quant_scalar(float v, float range)
{
  float c = v;
  if (fabsf(v) > range) {
    c = copysignf(range, v);
  }
  float q = c / range;
  q       = q * (float)LLR_MAX;
  q       = roundf(q);
  if (q != q) {
    return 0; /* float -> int8 of a NaN goes through cvttss2si: 0x80000000, low byte 0 */
  }
  return (int8_t)q;
}

/* mm256::quantize_ps: scale, clip (ordered compares keep a NaN), round to nearest even, convert, zero what is out of bounds. */
static int8_t quant_256(float v, float range)
{
  float scale = (float)LLR_MAX / range;
  float q     = v * scale;
  if (q > (float)LLR_MAX) {
    q = (float)LLR_MAX;
  }
  if (q < -(float)LLR_MAX) {
    q = -(float)LLR_MAX;
  }
  q = nearbyintf(q);
  int32_t i = cvt_i32(q);
  if (i > LLR_MAX || i < -LLR_MAX) {
    i = 0;
  }
  return (int8_t)i;
}

/* mm512::quantize_ps: NaN -> 0, scale, min, max, convert with round to nearest even. */
static int8_t quant_512(float v, float range)
{
  if (v != v) {
    return 0;
  }
  float scale = (float)LLR_MAX / range;
  float q     = scale * v;
  q           = q < (float)LLR_MAX ? q : (float)LLR_MAX;
  q           = q > -(float)LLR_MAX ? q : -(float)LLR_MAX;
  return (int8_t)cvt_i32(nearbyintf(q));
}

/* ---- interval functions ---------------------------------------------------------------------------------------------- */
static float interval_scalar(const itab* t, float x, float rcp, int fma)
{
  int32_t f   = cvt_i32(floorf(x / t->width));
  int     idx = clampi((int64_t)f + t->n / 2, 0, t->n - 1);
  float   l   = fma ? fmaf(t->slope[idx], x, t->icpt[idx]) : (t->slope[idx] * x + t->icpt[idx]);
  return l * rcp;
}

static float interval_256(const itab* t, float x, float rcp, int fma)
{
  float   s   = x * (1.0f / t->width);
  int32_t f   = cvt_i32(floorf(s));
  int32_t w   = (int32_t)((uint32_t)f + (uint32_t)(t->n / 2)); /* _mm256_add_epi32 wraps */
  int     idx = clampi(w, 0, t->n - 1);
  float   l   = fma ? fmaf(t->slope[idx], x, t->icpt[idx]) : (t->slope[idx] * x + t->icpt[idx]);
  l           = l * rcp;
  if (fabsf(x) <= NEAR_ZERO) {
    l = 0;
  }
  return l;
}

static float interval_512(const itab* t, float x, float rcp)
{
  float   s   = fmaf(x, 1.0f / t->width, (float)(t->n / 2));
  int32_t f   = cvt_i32(floorf(s));
  int     idx = clampi(f, 0, t->n - 1);
  float   l   = fmaf(t->slope[idx], x, t->icpt[idx]);
  l           = l * rcp;
  if (!(fabsf(x) > NEAR_ZERO)) {
    l = 0;
  }
  return l;
}

/* ---- one symbol, by path ----------------------------------------------------------------------------------------------- */
enum { PATH_SCALAR = 0, PATH_256 = 1, PATH_512 = 2 };

static int near_zero_symbol(float re, float im, int fma)
{
  float a = fma ? fmaf(re, re, im * im) : (re * re + im * im);
  return a < NEAR_ZERO;
}

static void demod_symbol(int qm, int path, int fma, float re, float im, float nv, int8_t* out)
{
  const float xs[2] = {re, im};
  if (qm == 2) {
    const float gain = 2.0f * 1.41421356237309504880f; /* 2.0F * M_SQRT2f32 */
    for (int c = 0; c < 2; ++c) {
      if (path == PATH_256) {
        out[c] = quant_256((gain * xs[c]) * safe_rcp(nv), 24.0f);
      } else {
        out[c] = !(nv > 0) ? 0 : quant_scalar(gain * xs[c] / nv, 24.0f);
      }
    }
    return;
  }
  if (qm == 4) {
    if (path == PATH_256) {
      float rcp = safe_rcp(nv);
      for (int c = 0; c < 2; ++c) {
        float x     = xs[c];
        float first = (4.0f * M10) * x;
        float l01   = first;
        if (fabsf(x) > 2 * M10) {
          l01 = 2.0f * first - copysignf(0.8f, x);
        }
        float l23 = 0.8f - fabsf(first);
        l01       = l01 * rcp;
        l23       = l23 * rcp;
        if (fabsf(x) <= NEAR_ZERO) {
          l01 = 0;
          l23 = 0;
        }
        out[c]     = quant_256(l01, 20.0f);
        out[2 + c] = quant_256(l23, 20.0f);
      }
    } else {
      if (near_zero_symbol(re, im, fma)) {
        memset(out, 0, 4);
        return;
      }
      for (int c = 0; c < 2; ++c) {
        float x = xs[c];
        if (!(nv > 0)) {
          out[c] = out[2 + c] = 0;
          continue;
        }
        float l = (4 * M10) * x;
        if (fabsf(x) > 2 * M10) {
          l = 2 * l - copysignf(0.8f, x);
        }
        out[c]  = quant_scalar(l / nv, 20.0f);
        float m = fma ? fmaf(-(4 * M10), fabsf(x), 0.8f) : (0.8f - (4 * M10) * fabsf(x));
        out[2 + c] = quant_scalar(m / nv, 20.0f);
      }
    }
    return;
  }
  const itab* tabs = (qm == 6) ? T64 : T256;
  const int   nt   = qm / 2;
  if (path == PATH_SCALAR && near_zero_symbol(re, im, fma)) {
    memset(out, 0, (size_t)qm);
    return;
  }
  float rcp = safe_rcp(nv);
  for (int t = 0; t < nt; ++t) {
    for (int c = 0; c < 2; ++c) {
      float x = xs[c];
      if (path == PATH_512) {
        out[2 * t + c] = quant_512(interval_512(&tabs[t], x, rcp), 20.0f);
      } else if (path == PATH_256) {
        out[2 * t + c] = quant_256(interval_256(&tabs[t], x, rcp, fma), 20.0f);
      } else {
        out[2 * t + c] = quant_scalar(interval_scalar(&tabs[t], x, rcp, fma), 20.0f);
      }
    }
  }
}

/* Symbols per iteration of the AVX-512 and AVX2 loops of demodulate_soft_<mod>, by flavour. */
void orc_demod_blocks(int flavour, int qm, unsigned* v512, unsigned* v256)
{
  *v512 = 0;
  *v256 = 0;
  if (flavour >= 1) {
    *v256 = (qm == 2) ? 16 : (qm == 4) ? 8 : (qm == 6) ? 16 : (qm == 8) ? 4 : 0;
  }
  if (flavour >= 2) {
    *v512 = (qm == 6) ? 32 : (qm == 8) ? 8 : 0;
  }
}

/* One demodulation_mapper::demodulate_soft call (demodulation_mapper_impl.cpp:77-106). */
void orc_demodulate_soft(int flavour, int qm, int pi2, unsigned n, const float* sym, const float* noise, int8_t* llr)
{
  tables();
  const int fma = flavour >= 1;
  if (qm == 1) {
    const float gain = 2.0f * 1.41421356237309504880f;
    for (unsigned i = 0; i < n; ++i) {
      float re = sym[2 * i], im = sym[2 * i + 1];
      float s  = (pi2 && (i & 1)) ? (im + (-re)) : (re + im);
      llr[i]   = !(noise[i] > 0) ? 0 : quant_scalar(gain * s / noise[i], 24.0f);
    }
    return;
  }
  unsigned v512, v256;
  orc_demod_blocks(flavour, qm, &v512, &v256);
  const unsigned end512 = v512 ? (n / v512) * v512 : 0;
  const unsigned end256 = v256 ? (n / v256) * v256 : 0;
  for (unsigned i = 0; i < n; ++i) {
    int path = (i < end512) ? PATH_512 : (i < end256) ? PATH_256 : PATH_SCALAR;
    demod_symbol(qm, path, fma, sym[2 * i], sym[2 * i + 1], noise[i], llr + (size_t)i * qm);
  }
}

/* ---- TS 38.211 5.2.1 pseudo-random sequence (pseudo_random_generator_impl.cpp: init / advance / generate) ---------------- */
void orc_gold_sequence(uint32_t c_init, uint32_t offset, uint32_t nbits, uint8_t* packed)
{
  uint32_t x1 = 1, x2 = c_init & 0x7fffffffu; /* 31-bit states, bit k = x(n + k) */
  memset(packed, 0, (nbits + 7) / 8);
  const uint64_t total = 1600ull + offset + nbits;
  for (uint64_t n = 0; n < total; ++n) {
    if (n >= 1600ull + offset) {
      uint32_t c = (x1 ^ x2) & 1u;
      uint64_t i = n - 1600ull - offset;
      packed[i / 8] |= (uint8_t)(c << (7 - (i % 8)));
    }
    uint32_t f1 = ((x1 >> 3) ^ x1) & 1u;
    uint32_t f2 = ((x2 >> 3) ^ (x2 >> 2) ^ (x2 >> 1) ^ x2) & 1u;
    x1          = (x1 >> 1) | (f1 << 30);
    x2          = (x2 >> 1) | (f2 << 30);
  }
}

/* revert_scrambling (pusch_demodulator_impl.cpp:39-207): the soft bit changes sign where the sequence bit is one. */
void orc_descramble(const uint8_t* packed_seq, uint32_t n, const int8_t* in, int8_t* out)
{
  for (uint32_t i = 0; i < n; ++i) {
    int bit = (packed_seq[i / 8] >> (7 - (i % 8))) & 1;
    out[i]  = bit ? (int8_t)(-in[i]) : in[i];
  }
}
