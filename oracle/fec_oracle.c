/* oracle/fec_oracle.c -- TEST INFRASTRUCTURE (the parity oracle), NOT PRODUCT CODE. See fec_oracle.h.
 *
 * Scalar restatement of the reference algorithm (all citations relative to /root/reference):
 *   LLR algebra      include/srsran/phy/upper/log_likelihood_ratio.h:45-256, lib/phy/upper/log_likelihood_ratio.cpp:40-99
 *   rate dematcher   lib/phy/upper/channel_coding/ldpc/ldpc_rate_dematcher_impl.cpp:46-257
 *   decoder driver   lib/phy/upper/channel_coding/ldpc/ldpc_decoder_impl.cpp:33-352
 *   decoder kernels  lib/phy/upper/channel_coding/ldpc/ldpc_decoder_generic.cpp:30-122
 *   CRC              lib/phy/upper/channel_coding/crc_calculator_generic_impl.cpp:32-133
 *   per-CB chain     lib/phy/upper/channel_processors/pusch/pusch_codeblock_decoder.cpp:34-71
 * It treats every edge uniformly through the adjacency list; the reference's split between the high-rate region and the
 * extension column (ldpc_decoder_impl.cpp:200-237) is storage layout only.
 */
#include "fec_oracle.h"
#include "../srsran-5g_b200/csrc/ldpc_tables.h"
#include <math.h>
#include <stdlib.h>
#include <string.h>

/* ------------------------------------------------------------------ graphs */
int orc_graph_init(orc_graph* g, int bg, int Z)
{
  int set = -1;
  for (int i = 0; i < 51; ++i) {
    if (FEC_ALL_Z[i] == Z) {
      set = FEC_Z_SET[i];
    }
  }
  if (set < 0 || (bg != 1 && bg != 2)) {
    return -1;
  }
  const uint8_t*  deg = (bg == 1) ? FEC_BG1_DEG : FEC_BG2_DEG;
  const uint8_t*  col = (bg == 1) ? FEC_BG1_COL : FEC_BG2_COL;
  const uint16_t* sh  = (bg == 1) ? FEC_BG1_SHIFT[set] : FEC_BG2_SHIFT[set];
  g->bg               = bg;
  g->Z                = Z;
  g->M                = (bg == 1) ? FEC_BG1_M : FEC_BG2_M;
  g->Kb               = (bg == 1) ? 22 : 10;
  g->n_full           = (bg == 1) ? 68 : 52;
  g->n_edges          = (bg == 1) ? FEC_BG1_EDGES : FEC_BG2_EDGES;
  int e               = 0;
  for (int m = 0; m < g->M; ++m) {
    g->deg[m]       = deg[m];
    g->row_start[m] = (uint16_t)e;
    e += deg[m];
  }
  g->row_start[g->M] = (uint16_t)e;
  for (int i = 0; i < g->n_edges; ++i) {
    g->col[i]   = col[i];
    g->shift[i] = (uint16_t)(sh[i] % Z); /* ldpc_luts_impl.cpp:4536-4541 */
  }
  return 0;
}

/* ------------------------------------------------------------------ LLR algebra */
/* x + y as the reference evaluates it: operator+ is "rhs += lhs" (log_likelihood_ratio.h:98-102), and operator+=
 * tests equal-and-opposite first, then infinities (rhs first), then clips to +-120 (log_likelihood_ratio.cpp:40-73). */
int8_t orc_sat_add(int8_t x, int8_t y)
{
  if (y == -x) {
    return 0;
  }
  if (y == 127 || y == -127) {
    return y;
  }
  if (x == 127 || x == -127) {
    return x;
  }
  int t = (int)x + (int)y;
  if (t > 120) {
    t = 120;
  }
  if (t < -120) {
    t = -120;
  }
  return (int8_t)t;
}

/* promotion_sum(c2v, v2c) (log_likelihood_ratio.cpp:75-88): overflow promotes to +-127. */
int8_t orc_promo(int8_t c, int8_t v)
{
  if (c == -v) {
    return 0;
  }
  if (c == 127 || c == -127) {
    return c;
  }
  if (v == 127 || v == -127) {
    return v;
  }
  int t = (int)c + (int)v;
  if (t > 120) {
    return 127;
  }
  if (t < -120) {
    return -127;
  }
  return (int8_t)t;
}

/* ldpc_decoder_generic.cpp:70-79: round(float(x) * sf), +-127 passes through; x in 0..127 here. */
void orc_make_scale_lut(float sf, uint8_t lut[128])
{
  for (int x = 0; x < 128; ++x) {
    lut[x] = (x > 120) ? (uint8_t)x : (uint8_t)(int)roundf((float)x * sf);
  }
}

/* ------------------------------------------------------------------ CRC */
static const uint32_t CRC_POLY[3] = {0x1864CFB, 0x1800063, 0x11021};
static const int      CRC_LEN[3]  = {24, 24, 16};

/* Bitwise long division, MSB first, zero initial state, no final xor (crc_calculator_generic_impl.cpp:111-133). */
uint32_t orc_crc(int poly, const uint8_t* bytes, uint32_t nbits)
{
  const int      len  = CRC_LEN[poly];
  const uint32_t g    = CRC_POLY[poly];
  const uint32_t top  = 1u << len;
  uint32_t       r    = 0;
  for (uint32_t i = 0; i < nbits; ++i) {
    uint32_t b = (bytes[i >> 3] >> (7 - (i & 7))) & 1u;
    r          = (r << 1) | b;
    if (r & top) {
      r ^= g;
    }
  }
  /* multiply by x^len */
  for (int i = 0; i < len; ++i) {
    r <<= 1;
    if (r & top) {
      r ^= g;
    }
  }
  return r & (top - 1);
}

/* ------------------------------------------------------------------ rate dematcher */
int orc_rate_dematch(int bg, int Z, int8_t* out, const int8_t* in, uint32_t E, int new_data, uint32_t rv, uint32_t qm,
                     uint32_t Nref, uint32_t fillers)
{
  static const double F1[4] = {0, 17, 33, 56};
  static const double F2[4] = {0, 13, 25, 43};
  if ((bg != 1 && bg != 2) || rv > 3 || E == 0 || (qm != 1 && qm != 2 && qm != 4 && qm != 6 && qm != 8) || E % qm) {
    return -1;
  }
  const uint32_t N    = (uint32_t)((bg == 1) ? 66 : 50) * (uint32_t)Z;
  const uint32_t Kb   = (bg == 1) ? 22 : 10;
  const uint32_t Ncb  = (Nref > 0) ? (Nref < N ? Nref : N) : N;       /* :62-66 */
  const uint32_t sys  = (Kb - 2) * (uint32_t)Z;                       /* :99 */
  if (fillers >= sys) {
    return -1;
  }
  const uint32_t info = sys - fillers;
  const double   f    = (bg == 1) ? F1[rv] : F2[rv];
  const uint32_t k0   = (uint32_t)(uint16_t)floor(f * (double)Ncb / (double)N) * (uint32_t)Z; /* :104-105 */

  /* de-interleave (:203-257): d[(E/Qm)*j + i] = in[i*Qm + j] */
  int8_t* d = (int8_t*)malloc(E);
  if (qm == 1) {
    memcpy(d, in, E);
  } else {
    uint32_t K = E / qm;
    for (uint32_t i = 0; i < K; ++i) {
      for (uint32_t j = 0; j < qm; ++j) {
        d[K * j + i] = in[i * qm + j];
      }
    }
  }

  /* allot_llrs (:128-201) */
  uint32_t p    = 0;
  uint32_t idx  = k0;
  int      copy = new_data;
#define PUT(o, n)                                                                                                      \
  do {                                                                                                                 \
    for (uint32_t ii = 0; ii < (n); ++ii) {                                                                            \
      out[(o) + ii] = copy ? d[p + ii] : orc_sat_add(out[(o) + ii], d[p + ii]);                                        \
    }                                                                                                                  \
    p += (n);                                                                                                          \
  } while (0)
  while (p < E) {
    if (idx < info) {
      uint32_t n = info - idx;
      if (n > E - p) {
        n = E - p;
      }
      if (copy) {
        memset(out, 0, idx);
      }
      PUT(idx, n);
      idx += n;
    } else if (copy) {
      memset(out, 0, info);
    }
    if (copy) {
      memset(out + info, 127, fillers);
    }
    if (idx < sys) {
      idx = sys;
    }
    uint32_t n = Ncb - idx;
    if (n > E - p) {
      n = E - p;
    }
    PUT(idx, n);
    idx = (idx + n) % Ncb;
    if (p < E) {
      copy = 0;
    }
  }
#undef PUT
  if (copy && idx != 0) {
    uint32_t n = Ncb - idx;
    memset(out + N - n, 0, n); /* measured from the end of the N-sized span (:198-200) */
  }
  free(d);
  return 0;
}

/* ------------------------------------------------------------------ decoder */
static int8_t clamp8(int v, int lo, int hi)
{
  return (int8_t)(v < lo ? lo : (v > hi ? hi : v));
}

int orc_decode(int bg, int Z, uint8_t* out, const int8_t* llr, uint32_t n_llr, uint32_t fillers, int crc_poly,
               uint32_t max_iter, float scaling)
{
  orc_graph g;
  if (orc_graph_init(&g, bg, Z) != 0 || max_iter == 0) {
    return -1;
  }
  uint8_t lut[128];
  orc_make_scale_lut(scaling, lut);
  const uint32_t uZ      = (uint32_t)Z;
  const uint32_t K       = (uint32_t)g.Kb * uZ;
  const uint32_t n_short = (uint32_t)(g.n_full - 2) * uZ;
  const uint32_t kbytes  = (K + 7) / 8;
  if (n_llr < (uint32_t)(g.Kb + 2) * uZ || n_llr > n_short) {
    return -1;
  }
  /* trim trailing zeros (ldpc_decoder_impl.cpp:86-90) */
  uint32_t n_sig = n_llr;
  while (n_sig > 0 && llr[n_sig - 1] == 0) {
    --n_sig;
  }
  if (n_sig < K) { /* :93-99 */
    if (crc_poly < 0) {
      memset(out, 0xff, kbytes);
      if (K % 8) {
        out[kbytes - 1] = (uint8_t)(0xff << (8 - K % 8));
      }
    }
    return -1;
  }

  /* load_soft_bits (:152-193) */
  const uint32_t n_soft = (uint32_t)g.n_full * uZ;
  int8_t*        soft   = (int8_t*)calloc(n_soft, 1);
  {
    /* The reference walks whole nodes while "significant" LLRs remain, clamping to +-64, then copies an unaligned
     * tail unclamped. Restated: positions < floor(n_llr/Z)*Z that are < ceil-to-node(n_sig) are clamped copies. */
    uint32_t full_nodes = n_llr / uZ;
    for (uint32_t node = 0; node < full_nodes; ++node) {
      if (node * uZ < n_sig) {
        for (uint32_t j = 0; j < uZ; ++j) {
          soft[(node + 2) * uZ + j] = clamp8(llr[node * uZ + j], -64, 64);
        }
      }
    }
    uint32_t tail = n_llr % uZ;
    if (tail != 0 && full_nodes * uZ < n_sig) {
      for (uint32_t j = 0; j < tail; ++j) {
        soft[(full_nodes + 2) * uZ + j] = llr[full_nodes * uZ + j];
      }
    }
  }

  /* codeblock length / layers actually processed (:108-117) */
  uint32_t cbl = n_sig + 2 * uZ;
  if (cbl < K + 4 * uZ) {
    cbl = K + 4 * uZ;
  }
  cbl              = ((cbl + uZ - 1) / uZ) * uZ;
  const uint32_t L = cbl / uZ - (uint32_t)g.Kb;

  int8_t*  c2v   = (int8_t*)calloc((size_t)g.n_edges * uZ, 1);
  int8_t*  v2c   = (int8_t*)malloc(20u * uZ);
  uint8_t* m1    = (uint8_t*)malloc(uZ);
  uint8_t* m2    = (uint8_t*)malloc(uZ);
  uint8_t* mi    = (uint8_t*)malloc(uZ);
  uint8_t* sg    = (uint8_t*)malloc(uZ);
  uint8_t  init[46];
  memset(init, 0, sizeof(init));
  int      ret   = -1;
  uint8_t* bits  = (uint8_t*)malloc(kbytes);

  for (uint32_t it = 0; it < max_iter && ret < 0; ++it) {
    for (uint32_t l = 0; l < L; ++l) {
      const uint32_t e0 = g.row_start[l];
      const uint32_t d  = g.deg[l];
      /* variable-to-check (impl.cpp:195-238, generic.cpp:30-44) */
      for (uint32_t e = 0; e < d; ++e) {
        const int8_t* s = soft + (uint32_t)g.col[e0 + e] * uZ;
        const int8_t* c = c2v + (size_t)(e0 + e) * uZ;
        for (uint32_t j = 0; j < uZ; ++j) {
          v2c[e * uZ + j] = init[l] ? orc_sat_add(s[j], (int8_t)-c[j]) : s[j];
        }
      }
      /* check node analysis (impl.cpp:255-331, generic.cpp:46-68) */
      for (uint32_t c = 0; c < uZ; ++c) {
        uint8_t a1 = 120, a2 = 120, ai = 0, s = 0;
        for (uint32_t e = 0; e < d; ++e) {
          int8_t  m = v2c[e * uZ + (c + g.shift[e0 + e]) % uZ];
          uint8_t a = (uint8_t)(m < 0 ? -m : m);
          if (a < a1) {
            a2 = a1;
            a1 = a;
            ai = (uint8_t)e;
          } else if (a < a2) {
            a2 = a;
          }
          s ^= (uint8_t)(m < 0);
        }
        m1[c] = a1;
        m2[c] = a2;
        mi[c] = ai;
        sg[c] = s;
      }
      /* check-to-variable (generic.cpp:85-108) + soft-bit update (impl.cpp:240-253, generic.cpp:110-122) */
      for (uint32_t e = 0; e < d; ++e) {
        int8_t*  s  = soft + (uint32_t)g.col[e0 + e] * uZ;
        int8_t*  cc = c2v + (size_t)(e0 + e) * uZ;
        uint32_t sh = g.shift[e0 + e];
        for (uint32_t j = 0; j < uZ; ++j) {
          uint32_t c   = (j + uZ - sh) % uZ;
          uint8_t  mag = lut[(mi[c] == e) ? m2[c] : m1[c]];
          int8_t   v   = v2c[e * uZ + j];
          int      neg = sg[c] ^ (v < 0);
          cc[j]        = (int8_t)(neg ? -(int)mag : (int)mag);
          s[j]         = orc_promo(cc[j], v);
        }
      }
      init[l] = 1;
    }
    if (crc_poly >= 0) { /* impl.cpp:129-137, 333-352; log_likelihood_ratio.cpp:350-376 */
      int clean = 1;
      memset(bits, 0, kbytes);
      for (uint32_t i = 0; i < K; ++i) {
        if (soft[i] <= 0) {
          bits[i >> 3] |= (uint8_t)(0x80 >> (i & 7));
        }
        if (soft[i] == 0) {
          clean = 0;
        }
      }
      memcpy(out, bits, kbytes);
      if (clean && orc_crc(crc_poly, bits, K - fillers) == 0) {
        ret = (int)it + 1;
      }
    }
  }
  if (crc_poly < 0) { /* :148-149 */
    memset(out, 0, kbytes);
    for (uint32_t i = 0; i < K; ++i) {
      if (soft[i] <= 0) {
        out[i >> 3] |= (uint8_t)(0x80 >> (i & 7));
      }
    }
  }
  free(bits);
  free(sg);
  free(mi);
  free(m2);
  free(m1);
  free(v2c);
  free(c2v);
  free(soft);
  return ret;
}

/* ------------------------------------------------------------------ per-CB chain */
int orc_cb_chain(int bg, int Z, int8_t* soft, const int8_t* in, uint32_t E, int new_data, uint32_t rv, uint32_t qm,
                 uint32_t Nref, uint32_t fillers, int crc_poly, uint32_t max_iter, int early_stop, uint8_t* out)
{
  const uint32_t N = (uint32_t)((bg == 1) ? 66 : 50) * (uint32_t)Z;
  const uint32_t K = (uint32_t)((bg == 1) ? 22 : 10) * (uint32_t)Z;
  if (orc_rate_dematch(bg, Z, soft, in, E, new_data, rv, qm, Nref, fillers) != 0) {
    return -2;
  }
  if (early_stop) {
    return orc_decode(bg, Z, out, soft, N, fillers, crc_poly, max_iter, 0.8f);
  }
  orc_decode(bg, Z, out, soft, N, fillers, -1, max_iter, 0.8f);
  /* NB when the decoder bails out (too few LLRs) with crc == nullptr it sets all ones, which then fails the CRC. */
  return (orc_crc(crc_poly, out, K - fillers) == 0) ? (int)max_iter : -1;
}
