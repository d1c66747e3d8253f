/* oracle/fec_ulsch.c -- CPU restatement of the reference's UL-SCH demultiplexer (TS 38.212 6.2.7), TEST INFRASTRUCTURE.
 * Only tests/, __graft_entry__.smoke() and bench.py's CPU legs may use it; the product never links or loads it.
 *
 * Follows lib/phy/upper/channel_processors/pusch/ulsch_demultiplex_impl.cpp of the reference:
 *   re_set_select                          :74-97    (every d-th member of a set, the first m of them)
 *   on_uci_placeholder_1bit / _2bit        :106-189  (placeholder bits x / y: undo the scrambling of the repeated bits)
 *   demultiplex (derived parameters)       :191-238
 *   configure_current_ofdm_symbol          :330-451  (steps 1, 2, 3, 3bis, 5 of 6.2.7 for one OFDM symbol)
 *   configure_csi_part2_current_ofdm_symbol:453-477
 *   demux_current_ofdm_symbol              :479-590  (order: HARQ-ACK, CSI Part 1, CSI Part 2, SCH; 1/2-bit HARQ-ACK zeroes
 *                                                     the soft bits it took, which punctures whatever else sits there)
 * Parity: PINNED against the reference itself (oracle/_ref/libref_fec.so: the unmodified ulsch_demultiplex_impl.cpp behind
 * oracle/ref_ulsch_shim.cpp) on random configurations -- tests/test_ulsch_cpu.py -- because the reference's golden files for
 * this block (ulsch_demultiplex_test_data.tar.gz) are not in the checkout; vectors generated from that build are committed
 * as tests/golden/ulsch_golden.npz.
 *
 * cfg[16]: see oracle/ref_ulsch_shim.cpp. cfg[15] != 0: CSI Part 2 becomes known when CSI Part 1 ends (pusch_processor_impl's
 * flow) instead of at configuration time (the unit test's flow). */
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

#define MAX_RE (275 * 12)

typedef struct {
  uint8_t ulsch[MAX_RE], uci[MAX_RE], rvd[MAX_RE], harq[MAX_RE], csi1[MAX_RE], csi2[MAX_RE];
  unsigned n;
} sets_t;

static unsigned count(const uint8_t* s, unsigned n)
{
  unsigned c = 0;
  for (unsigned i = 0; i < n; ++i) {
    c += s[i];
  }
  return c;
}

/* ulsch_demultiplex_impl.cpp:74-97 */
static void re_select(uint8_t* out, const uint8_t* set, unsigned n, unsigned d, unsigned m)
{
  memset(out, 0, n);
  unsigned taken = 0, seen = 0;
  for (unsigned i = 0; i < n && taken != m; ++i) {
    if (!set[i]) {
      continue;
    }
    if (seen % d == 0) {
      out[i] = 1;
      ++taken;
    }
    ++seen;
  }
}

static void spacing(unsigned avail, unsigned remainder, unsigned* d, unsigned* m)
{
  *d = 1;
  *m = avail;
  if (remainder < avail) {
    *d = avail / remainder;
    *m = remainder;
  }
}

static int8_t neg(int8_t v)
{
  return (int8_t)(-v);
}

/* :106-189. `seq` is the unpacked scrambling sequence of the codeword, `pos` the position of data[0] in it. */
static void placeholder(int8_t* out, const int8_t* data, const uint8_t* seq, unsigned pos, unsigned qm, unsigned nbits_re, int two)
{
  if (qm == 1) {
    memcpy(out, data, nbits_re);
    return;
  }
  for (unsigned s = 0; s < nbits_re / qm; ++s) {
    const int8_t*  x = data + s * qm;
    const uint8_t* c = seq + pos + s * qm;
    int8_t*        o = out + s * qm;
    o[0]             = x[0];
    o[1]             = (!two && ((c[0] ^ c[1]) == 1)) ? neg(x[1]) : x[1];
    for (unsigned i = 2; i < qm; ++i) {
      o[i] = c[i] ? neg(x[i]) : x[i];
    }
  }
}

static void csi2_step(sets_t* S, unsigned l, unsigned l1_csi, unsigned B, unsigned enc, unsigned* m_count)
{ /* :453-477 */
  const unsigned M_uci = count(S->uci, S->n);
  const unsigned rem   = (enc - *m_count) / B;
  if (l >= l1_csi && M_uci > 0 && rem > 0) {
    unsigned d, m;
    spacing(M_uci, rem, &d, &m);
    re_select(S->csi2, S->uci, S->n, d, m);
    for (unsigned i = 0; i < S->n; ++i) {
      if (S->csi2[i]) {
        S->ulsch[i] = 0;
        S->uci[i]   = 0;
      }
    }
    *m_count += m * B;
  }
}

/* Same interface as ref_ulsch_demux (without the block size: the result does not depend on it). codeword is modified only in
 * a private copy. Returns 0, or -1 on an invalid configuration. */
int orc_ulsch_demux(const uint32_t* cfg,
                    const int8_t*   codeword,
                    uint32_t        n,
                    const uint8_t*  scrambling,
                    int8_t*         o_sch,
                    int8_t*         o_harq,
                    int8_t*         o_csi1,
                    int8_t*         o_csi2,
                    uint32_t        cap,
                    uint32_t*       counts)
{
  const unsigned qm = cfg[0], nl = cfg[1], nprb = cfg[2], start = cfg[3], nsym = cfg[4], nof_rvd = cfg[5];
  const unsigned dmrs_type = cfg[6], dmrs_mask = cfg[7], ncdm = cfg[8];
  const unsigned harq_bits = cfg[9], harq_enc = cfg[10], csi1_bits = cfg[11], csi1_enc = cfg[12];
  const unsigned csi2_bits = cfg[13], csi2_enc = cfg[14], dynamic = cfg[15];
  if (qm == 0 || qm > 8 || nl == 0 || nl > 4 || nprb == 0 || nprb > 275 || start + nsym > 14 || (dmrs_mask & 0x3fff) == 0 ||
      ncdm == 0 || ncdm > (dmrs_type == 1 ? 2u : 3u)) {
    return -1;
  }
  const unsigned B = qm * nl; /* soft bits per resource element */
  /* :33-55 */
  unsigned l_dmrs = 0;
  while (!((dmrs_mask >> l_dmrs) & 1u)) {
    ++l_dmrs;
  }
  unsigned l1 = l_dmrs;
  while (l1 < 14 && ((dmrs_mask >> l1) & 1u)) {
    ++l1;
  }
  unsigned l1_csi = 0;
  while (l1_csi < 14 && ((dmrs_mask >> l1_csi) & 1u)) {
    ++l1_csi;
  }
  /* :57-72 */
  const unsigned nof_re_dmrs = (12 - ncdm * (dmrs_type == 1 ? 6u : 4u)) * nprb;

  sets_t* S = (sets_t*)malloc(sizeof(sets_t));
  int8_t* sym = (int8_t*)malloc((size_t)MAX_RE * 32);
  if (!S || !sym) {
    free(S);
    free(sym);
    return -1;
  }
  unsigned m_rvd = 0, m_harq = 0, m_csi1 = 0, m_csi2 = 0, pos = 0;
  unsigned n_sch = 0, n_harq = 0, n_csi1 = 0, n_csi2 = 0;
  int      csi2_known = (csi2_enc != 0) && !dynamic, harq_open = harq_bits != 0, csi1_open = csi1_bits != 0, csi2_open = 0;
  int      end_harq = 0, end_csi1 = 0, end_csi2 = 0;
  int8_t   tmp[32];
  for (unsigned l = start; l < start + nsym; ++l) {
    /* ---- configure_current_ofdm_symbol, :330-451 ---- */
    const int contains_dmrs = (dmrs_mask >> l) & 1u;
    const unsigned M        = contains_dmrs ? nof_re_dmrs : nprb * 12;
    if (M == 0) {
      continue;
    }
    S->n = M;
    memset(S->ulsch, 1, M);
    memset(S->uci, contains_dmrs ? 0 : 1, M);
    memset(S->rvd, 0, M);
    memset(S->harq, 0, M);
    memset(S->csi1, 0, M);
    memset(S->csi2, 0, M);
    unsigned M_uci = count(S->uci, M), d, m;
    /* step 1 */
    const unsigned rem_rvd = (nof_rvd - m_rvd) / B;
    if (l >= l1 && M_uci > 0 && rem_rvd > 0) {
      spacing(M_uci, rem_rvd, &d, &m);
      re_select(S->rvd, S->ulsch, M, d, m);
      m_rvd += m * B;
    }
    /* step 2 */
    const unsigned rem_harq = (harq_enc - m_harq) / B;
    if (l >= l1 && M_uci > 0 && harq_bits > 2 && rem_harq > 0) {
      spacing(M_uci, rem_harq, &d, &m);
      re_select(S->harq, S->uci, M, d, m);
      for (unsigned i = 0; i < M; ++i) {
        if (S->harq[i]) {
          S->ulsch[i] = 0;
          S->uci[i]   = 0;
        }
      }
      M_uci = count(S->uci, M);
      m_harq += m * B;
    }
    /* step 3 */
    const unsigned rem_csi1 = (csi1_enc - m_csi1) / B;
    const unsigned M_rvd    = count(S->rvd, M);
    if (l >= l1_csi && (M_uci - M_rvd) > 0 && rem_csi1 > 0) {
      spacing(M_uci - M_rvd, rem_csi1, &d, &m);
      uint8_t* cand = S->csi2; /* scratch: csi2 is still empty here */
      for (unsigned i = 0; i < M; ++i) {
        cand[i] = (uint8_t)(!S->rvd[i] && S->uci[i]);
      }
      re_select(S->csi1, cand, M, d, m);
      memset(S->csi2, 0, M);
      for (unsigned i = 0; i < M; ++i) {
        if (S->csi1[i]) {
          S->ulsch[i] = 0;
          S->uci[i]   = 0;
        }
      }
      m_csi1 += m * B;
    }
    /* step 3bis */
    if (csi2_known) {
      csi2_step(S, l, l1_csi, B, csi2_enc, &m_csi2);
      csi2_open = 1;
    }
    /* step 5 */
    if (M_rvd > 0 && harq_bits <= 2 && rem_harq > 0) {
      spacing(M_rvd, rem_harq, &d, &m);
      re_select(S->harq, S->rvd, M, d, m);
      m_harq += m * B;
    }

    /* ---- demux_current_ofdm_symbol, :479-590 ---- */
    if (pos + M * B > n) {
      free(S);
      free(sym);
      return -1;
    }
    memcpy(sym, codeword + pos, (size_t)M * B);
    if (harq_open) {
      for (unsigned i = 0; i < M; ++i) {
        if (!S->harq[i]) {
          continue;
        }
        int8_t* x = sym + i * B;
        if (harq_bits == 1 || harq_bits == 2) {
          placeholder(tmp, x, scrambling, pos + i * B, qm, B, harq_bits == 2);
          memset(x, 0, B);
          x = tmp;
        }
        for (unsigned b = 0; b < B; ++b, ++n_harq) {
          if (n_harq < cap) {
            o_harq[n_harq] = x[b];
          }
        }
      }
      if (count(S->harq, M) && m_harq == harq_enc) {
        end_harq  = 1;
        harq_open = 0;
      }
    }
    if (csi1_open) {
      int any = 0;
      for (unsigned i = 0; i < M; ++i) {
        if (!S->csi1[i]) {
          continue;
        }
        any             = 1;
        const int8_t* x = sym + i * B;
        if (csi1_bits == 1 || csi1_bits == 2) {
          placeholder(tmp, x, scrambling, pos + i * B, qm, B, csi1_bits == 2);
          x = tmp;
        }
        for (unsigned b = 0; b < B; ++b, ++n_csi1) {
          if (n_csi1 < cap) {
            o_csi1[n_csi1] = x[b];
          }
        }
      }
      if (any && m_csi1 == csi1_enc) {
        end_csi1  = 1;
        csi1_open = 0;
        if (dynamic && csi2_enc != 0 && !csi2_known) { /* pusch_processor_impl.cpp:73-101 -> set_csi_part2, :241-252 */
          csi2_known = 1;
          csi2_open  = 1;
          csi2_step(S, l, l1_csi, B, csi2_enc, &m_csi2);
        }
      }
    }
    if (csi2_open) {
      int any = 0;
      for (unsigned i = 0; i < M; ++i) {
        if (!S->csi2[i]) {
          continue;
        }
        any             = 1;
        const int8_t* x = sym + i * B;
        if (csi2_bits == 1 || csi2_bits == 2) {
          placeholder(tmp, x, scrambling, pos + i * B, qm, B, csi2_bits == 2);
          x = tmp;
        }
        for (unsigned b = 0; b < B; ++b, ++n_csi2) {
          if (n_csi2 < cap) {
            o_csi2[n_csi2] = x[b];
          }
        }
      }
      if (any && m_csi2 == csi2_enc) {
        end_csi2  = 1;
        csi2_open = 0;
        csi2_known = 0; /* the reference drops its csi_part2 pointer: no further allocation is extracted */
      }
    }
    for (unsigned i = 0; i < M; ++i) {
      if (!S->ulsch[i]) {
        continue;
      }
      for (unsigned b = 0; b < B; ++b, ++n_sch) {
        if (n_sch < cap) {
          o_sch[n_sch] = sym[i * B + b];
        }
      }
    }
    pos += M * B;
  }
  counts[0] = n_sch;
  counts[1] = n_harq;
  counts[2] = n_csi1;
  counts[3] = n_csi2;
  counts[4] = 1;
  counts[5] = (uint32_t)end_harq;
  counts[6] = (uint32_t)end_csi1;
  counts[7] = (uint32_t)end_csi2;
  free(S);
  free(sym);
  return (pos == n) ? 0 : 1; /* 1: the codeword is longer than the allocation (the reference would keep waiting) */
}
