/* oracle/fec_tx.c -- TEST INFRASTRUCTURE: transmit side + channel, used only to GENERATE inputs for the decode path.
 *
 * Restates (from 3GPP TS 38.212 5.1 / 5.3.2 / 5.4.2 and the reference behaviour, paths relative to /root/reference):
 *   CRC attachment        lib/phy/upper/channel_coding/crc_calculator_generic_impl.cpp:32-133
 *   LDPC encoding         lib/phy/upper/channel_coding/ldpc/ldpc_encoder_impl.cpp, ldpc_encoder_generic.cpp (systematic bits,
 *                         4 core parity blocks by the double-diagonal structure, extension parity blocks)
 *   rate matching         lib/phy/upper/channel_coding/ldpc/ldpc_rate_matcher_impl.cpp:36-160 (bit selection from k0 skipping
 *                         the filler range, row-column interleaver)
 *   LLR quantisation      lib/phy/upper/log_likelihood_ratio.cpp:90-99
 * Checked bit-for-bit against the reference encoder / rate matcher in tests/test_oracle_vs_ref.py.
 */
#include "fec_oracle.h"
#include <math.h>
#include <stdlib.h>
#include <string.h>

static const int CRC_LEN[3] = {24, 24, 16};

int orc_crc_attach_bits(int poly, uint8_t* bits, uint32_t nbits)
{
  uint32_t nbytes = (nbits + 7) / 8;
  uint8_t* packed = (uint8_t*)calloc(nbytes + 1, 1);
  for (uint32_t i = 0; i < nbits; ++i) {
    if (bits[i] & 1) {
      packed[i >> 3] |= (uint8_t)(0x80 >> (i & 7));
    }
  }
  uint32_t crc = orc_crc(poly, packed, nbits);
  int      len = CRC_LEN[poly];
  for (int i = 0; i < len; ++i) {
    bits[nbits + (uint32_t)i] = (uint8_t)((crc >> (len - 1 - i)) & 1u);
  }
  free(packed);
  return len;
}

/* acc[c] ^= x[(c + s) mod Z] */
static void xor_rot(uint8_t* acc, const uint8_t* x, uint32_t s, uint32_t Z)
{
  for (uint32_t c = 0; c < Z; ++c) {
    uint32_t j = c + s;
    if (j >= Z) {
      j -= Z;
    }
    acc[c] ^= x[j];
  }
}

int orc_encode(int bg, int Z, const uint8_t* msg, uint8_t* cb)
{
  orc_graph g;
  if (orc_graph_init(&g, bg, Z) != 0) {
    return -1;
  }
  const uint32_t uZ = (uint32_t)Z;
  const uint32_t Kb = (uint32_t)g.Kb;
  uint8_t*       x  = (uint8_t*)calloc((size_t)g.n_full * uZ, 1); /* full codeword, block per base column */
  for (uint32_t i = 0; i < Kb * uZ; ++i) {
    x[i] = msg[i] & 1;
  }
  /* lambda_i = sum over systematic edges of row i, i < 4 */
  uint8_t* lam = (uint8_t*)calloc(4 * uZ, 1);
  uint32_t p0_shift[4];
  int      p0_n = 0;
  for (uint32_t i = 0; i < 4; ++i) {
    for (uint32_t e = g.row_start[i]; e < g.row_start[i + 1]; ++e) {
      if (g.col[e] < Kb) {
        xor_rot(lam + i * uZ, x + (uint32_t)g.col[e] * uZ, g.shift[e], uZ);
      } else if (g.col[e] == Kb) {
        p0_shift[p0_n++] = g.shift[e];
      }
    }
  }
  /* Sum of the four core rows: the double-diagonal columns cancel, column Kb leaves P^r p0 = sum(lambda). */
  if (p0_n != 3) {
    free(lam);
    free(x);
    return -2;
  }
  uint32_t r;
  if (p0_shift[0] == p0_shift[1]) {
    r = p0_shift[2];
  } else if (p0_shift[0] == p0_shift[2]) {
    r = p0_shift[1];
  } else if (p0_shift[1] == p0_shift[2]) {
    r = p0_shift[0];
  } else {
    free(lam);
    free(x);
    return -3;
  }
  uint8_t* sum = (uint8_t*)calloc(uZ, 1);
  for (uint32_t i = 0; i < 4; ++i) {
    for (uint32_t c = 0; c < uZ; ++c) {
      sum[c] ^= lam[i * uZ + c];
    }
  }
  uint8_t* p0 = x + Kb * uZ;
  for (uint32_t c = 0; c < uZ; ++c) {
    p0[(c + r) % uZ] = sum[c];
  }
  /* Rows 0..2: the one still-unknown core parity block of row i is column Kb+1+i. */
  for (uint32_t i = 0; i < 3; ++i) {
    uint8_t* acc = (uint8_t*)calloc(uZ, 1);
    memcpy(acc, lam + i * uZ, uZ);
    uint32_t s_unknown = 0;
    for (uint32_t e = g.row_start[i]; e < g.row_start[i + 1]; ++e) {
      uint32_t col = g.col[e];
      if (col < Kb) {
        continue;
      }
      if (col == Kb + 1 + i) {
        s_unknown = g.shift[e];
      } else {
        xor_rot(acc, x + col * uZ, g.shift[e], uZ);
      }
    }
    uint8_t* pi = x + (Kb + 1 + i) * uZ;
    for (uint32_t c = 0; c < uZ; ++c) {
      pi[(c + s_unknown) % uZ] = acc[c];
    }
    free(acc);
  }
  /* Extension rows: the last edge of row i >= 4 is the degree-1 column Kb+i (shift 0). */
  for (uint32_t i = 4; i < (uint32_t)g.M; ++i) {
    uint8_t* acc = (uint8_t*)calloc(uZ, 1);
    uint32_t s_unknown = 0;
    for (uint32_t e = g.row_start[i]; e < g.row_start[i + 1]; ++e) {
      uint32_t col = g.col[e];
      if (col == Kb + i) {
        s_unknown = g.shift[e];
      } else {
        xor_rot(acc, x + col * uZ, g.shift[e], uZ);
      }
    }
    uint8_t* pi = x + (Kb + i) * uZ;
    for (uint32_t c = 0; c < uZ; ++c) {
      pi[(c + s_unknown) % uZ] = acc[c];
    }
    free(acc);
  }
  memcpy(cb, x + 2 * uZ, (size_t)(g.n_full - 2) * uZ);
  free(sum);
  free(lam);
  free(x);
  return 0;
}

int orc_rate_match(int bg, int Z, const uint8_t* cb, uint8_t* out, uint32_t E, uint32_t rv, uint32_t qm, uint32_t Nref,
                   uint32_t fillers)
{
  static const double F1[4] = {0, 17, 33, 56};
  static const double F2[4] = {0, 13, 25, 43};
  const uint32_t      N     = (uint32_t)((bg == 1) ? 66 : 50) * (uint32_t)Z;
  const uint32_t      Kb    = (bg == 1) ? 22 : 10;
  const uint32_t      Ncb   = (Nref > 0) ? (Nref < N ? Nref : N) : N;
  const uint32_t      sys   = (Kb - 2) * (uint32_t)Z;
  const uint32_t      f_lo  = sys - fillers;
  const double        f     = (bg == 1) ? F1[rv] : F2[rv];
  const uint32_t      k0    = (uint32_t)(uint16_t)floor(f * (double)Ncb / (double)N) * (uint32_t)Z;
  if (E % qm) {
    return -1;
  }
  uint8_t* sel = (uint8_t*)malloc(E);
  uint32_t k   = k0 % Ncb;
  for (uint32_t i = 0; i < E;) {
    if (!(k >= f_lo && k < sys)) {
      sel[i++] = cb[k];
    }
    k = (k + 1) % Ncb;
  }
  uint32_t K = E / qm;
  for (uint32_t i = 0; i < K; ++i) {
    for (uint32_t j = 0; j < qm; ++j) {
      out[i * qm + j] = sel[j * K + i];
    }
  }
  free(sel);
  return 0;
}

int8_t orc_quantize(float v, float range)
{
  float c = v;
  if (fabsf(v) > range) {
    c = copysignf(range, v);
  }
  return (int8_t)roundf(c / range * 120.0f);
}

static uint64_t xs_next(uint64_t* s)
{
  uint64_t x = *s;
  x ^= x >> 12;
  x ^= x << 25;
  x ^= x >> 27;
  *s = x;
  return x * 0x2545F4914F6CDD1DULL;
}

void orc_random_bits(uint8_t* bits, uint32_t n, uint64_t seed)
{
  uint64_t s = seed * 0x9E3779B97F4A7C15ULL + 0x1234567ULL;
  if (s == 0) {
    s = 1;
  }
  for (uint32_t i = 0; i < n; i += 64) {
    uint64_t w = xs_next(&s);
    for (uint32_t j = 0; j < 64 && i + j < n; ++j) {
      bits[i + j] = (uint8_t)((w >> j) & 1);
    }
  }
}

void orc_awgn_llr(const uint8_t* bits, int8_t* llr, uint32_t E, float esn0_db, uint64_t seed)
{
  uint64_t s = seed * 0xD1B54A32D192ED03ULL + 0x7654321ULL;
  if (s == 0) {
    s = 1;
  }
  double esn0   = pow(10.0, (double)esn0_db / 10.0);
  double sigma2 = 1.0 / (2.0 * esn0);
  double sigma  = sqrt(sigma2);
  for (uint32_t i = 0; i < E; i += 2) {
    double u1 = ((double)(xs_next(&s) >> 11) + 1.0) / 9007199254740993.0;
    double u2 = (double)(xs_next(&s) >> 11) / 9007199254740992.0;
    double rr = sqrt(-2.0 * log(u1));
    double n0 = rr * cos(6.283185307179586 * u2);
    double n1 = rr * sin(6.283185307179586 * u2);
    double y0 = (bits[i] ? -1.0 : 1.0) + sigma * n0;
    llr[i]    = orc_quantize((float)(2.0 * y0 / sigma2), 20.0f);
    if (i + 1 < E) {
      double y1  = (bits[i + 1] ? -1.0 : 1.0) + sigma * n1;
      llr[i + 1] = orc_quantize((float)(2.0 * y1 / sigma2), 20.0f);
    }
  }
}
