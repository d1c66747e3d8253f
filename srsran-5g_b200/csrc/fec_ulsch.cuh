// fec_ulsch.cuh -- UL-SCH demultiplexing on the device (last pointer): the soft bits of a PUSCH codeword
// (demodulated and descrambled, resource element by resource element, OFDM symbol by OFDM symbol) are sorted into the SCH
// data stream the decoder reads and the HARQ-ACK / CSI Part 1 / CSI Part 2 streams the UCI decoders read, as
// ulsch_demultiplex_impl does on the CPU (lib/phy/upper/channel_processors/pusch/ulsch_demultiplex_impl.cpp; TS 38.212 6.2.7).
//
// Split of the work:
//   * WHICH resource element goes where depends on the configuration only, and on a chain of "every d-th member of the set
//     that is left" selections that run through the OFDM symbols in order (ibid. :330-477). The host walks that chain once per
//     configuration (fec_ulsch_plan_build: a few tens of microseconds, cached by the engine) and writes one flag byte per
//     resource element plus, per segment of 512 resource elements, how many elements of each stream precede it.
//   * MOVING the soft bits is the device's part: one CTA per segment, one thread per resource element; the rank of an element
//     inside its segment comes from warp ballots; Qm x layers soft bits per element are copied to each stream the element
//     belongs to. HARQ-ACK of one or two bits sits on reserved elements ON TOP of whatever is mapped there: its soft bits are
//     extracted first and the element then reads as zeros for the stream underneath (:497-510). One- and two-bit payloads of
//     any UCI field have the scrambling of their placeholder bits undone (:106-189), for which the caller provides the
//     scrambling sequence of the codeword as packed bits (MSB first, like bit_buffer).
#pragma once
#include "../../include/fec_b200.h"
#include <cstdint>
#include <cstring>
#include <vector>

#define FEC_ULSCH_SEG 512u // resource elements per CTA
#define FEC_ULSCH_F_SCH 1u
#define FEC_ULSCH_F_HARQ 2u
#define FEC_ULSCH_F_CSI1 4u
#define FEC_ULSCH_F_CSI2 8u

struct fec_ulsch_seg {
  uint32_t first_re;  // index of the segment's first resource element in the flag array (= in the codeword, in elements)
  uint32_t nof_re;    // <= FEC_ULSCH_SEG, never across an OFDM symbol boundary
  uint32_t base[4];   // resource elements of {SCH, HARQ-ACK, CSI Part 1, CSI Part 2} before this segment
};

struct fec_ulsch_plan {
  std::vector<uint8_t>       flags; // one byte per resource element of the allocation, in codeword order
  std::vector<fec_ulsch_seg> segs;
  uint32_t                   nof_re[4] = {0, 0, 0, 0}; // elements per stream
  uint32_t                   bits_per_re = 0;
};

namespace fec_ulsch_detail {

// Every d-th member of `set` (members in ascending order), the first m of them (ulsch_demultiplex_impl.cpp:74-97).
inline void select(std::vector<uint8_t>& out, const std::vector<uint8_t>& set, uint32_t d, uint32_t m)
{
  std::fill(out.begin(), out.end(), 0);
  uint32_t taken = 0, seen = 0;
  for (size_t i = 0; i < set.size() && taken != m; ++i) {
    if (set[i]) {
      if (seen % d == 0) {
        out[i] = 1;
        ++taken;
      }
      ++seen;
    }
  }
}

inline uint32_t members(const std::vector<uint8_t>& s)
{
  uint32_t c = 0;
  for (uint8_t v : s) {
    c += v;
  }
  return c;
}

// The "all of them, or every d-th so that `want` fit" rule that every step of 6.2.7 uses.
inline void spacing(uint32_t avail, uint32_t want, uint32_t& d, uint32_t& m)
{
  d = 1;
  m = avail;
  if (want < avail) {
    d = avail / want;
    m = want;
  }
}

} // namespace fec_ulsch_detail

inline bool fec_ulsch_cfg_valid(const fec_ulsch_cfg& c)
{
  const uint32_t max_cdm = (c.dmrs_type == 1) ? 2u : 3u;
  return (c.qm == 1 || c.qm == 2 || c.qm == 4 || c.qm == 6 || c.qm == 8) && c.nof_layers >= 1 && c.nof_layers <= 4 &&
         c.nof_prb >= 1 && c.nof_prb <= 275 && c.nof_symbols >= 1 && c.start_symbol_index + c.nof_symbols <= 14 &&
         (c.dmrs_type == 1 || c.dmrs_type == 2) && c.nof_cdm_groups_without_data >= 1 &&
         c.nof_cdm_groups_without_data <= max_cdm && (c.dmrs_symbol_mask & 0x3fffu) != 0;
}

// Walks the OFDM symbols of the allocation as the reference does when it configures each of them, and records the outcome.
inline bool fec_ulsch_plan_build(const fec_ulsch_cfg& c, fec_ulsch_plan& P)
{
  using namespace fec_ulsch_detail;
  if (!fec_ulsch_cfg_valid(c)) {
    return false;
  }
  const uint32_t B = (uint32_t)c.qm * c.nof_layers;
  // first symbol after the first run of DM-RS symbols; first symbol without DM-RS (:33-55)
  uint32_t l1 = 0;
  while (!((c.dmrs_symbol_mask >> l1) & 1u)) {
    ++l1;
  }
  while (l1 < 14 && ((c.dmrs_symbol_mask >> l1) & 1u)) {
    ++l1;
  }
  uint32_t l1_csi = 0;
  while (l1_csi < 14 && ((c.dmrs_symbol_mask >> l1_csi) & 1u)) {
    ++l1_csi;
  }
  const uint32_t re_dmrs_symbol = (12u - c.nof_cdm_groups_without_data * ((c.dmrs_type == 1) ? 6u : 4u)) * c.nof_prb;

  P.flags.clear();
  P.segs.clear();
  P.bits_per_re = B;
  for (uint32_t& v : P.nof_re) {
    v = 0;
  }
  uint32_t done_rvd = 0, done_harq = 0, done_csi1 = 0, done_csi2 = 0; // soft bits allotted so far
  bool     csi2_known = (c.nof_enc_csi_part2_bits != 0) && (c.csi2_after_csi1 == 0);
  std::vector<uint8_t> sch, uci, rvd, harq, csi1, csi2, cand;
  for (uint32_t l = c.start_symbol_index; l < (uint32_t)c.start_symbol_index + c.nof_symbols; ++l) {
    const bool     dmrs = ((c.dmrs_symbol_mask >> l) & 1u) != 0;
    const uint32_t M    = dmrs ? re_dmrs_symbol : (uint32_t)c.nof_prb * 12u;
    if (M == 0) {
      continue;
    }
    sch.assign(M, 1);
    uci.assign(M, dmrs ? 0 : 1);
    rvd.assign(M, 0);
    harq.assign(M, 0);
    csi1.assign(M, 0);
    csi2.assign(M, 0);
    cand.assign(M, 0);
    uint32_t avail = members(uci), d, m;
    // elements reserved for a HARQ-ACK of up to two bits
    const uint32_t want_rvd = (c.nof_harq_ack_rvd - done_rvd) / B;
    if (l >= l1 && avail > 0 && want_rvd > 0) {
      spacing(avail, want_rvd, d, m);
      select(rvd, sch, d, m);
      done_rvd += m * B;
    }
    // HARQ-ACK of more than two bits takes its elements away from everything else
    const uint32_t want_harq = (c.nof_enc_harq_ack_bits - done_harq) / B;
    if (l >= l1 && avail > 0 && c.nof_harq_ack_bits > 2 && want_harq > 0) {
      spacing(avail, want_harq, d, m);
      select(harq, uci, d, m);
      for (uint32_t i = 0; i < M; ++i) {
        if (harq[i]) {
          sch[i] = uci[i] = 0;
        }
      }
      avail = members(uci);
      done_harq += m * B;
    }
    // CSI Part 1 avoids the reserved elements
    const uint32_t want_csi1 = (c.nof_enc_csi_part1_bits - done_csi1) / B;
    const uint32_t n_rvd     = members(rvd);
    if (l >= l1_csi && (avail - n_rvd) > 0 && want_csi1 > 0) { // unsigned, as in the reference
      spacing(avail - n_rvd, want_csi1, d, m);
      for (uint32_t i = 0; i < M; ++i) {
        cand[i] = (uint8_t)(uci[i] && !rvd[i]);
      }
      select(csi1, cand, d, m);
      for (uint32_t i = 0; i < M; ++i) {
        if (csi1[i]) {
          sch[i] = uci[i] = 0;
        }
      }
      done_csi1 += m * B;
    }
    // CSI Part 2 may use the reserved elements
    const auto place_csi2 = [&]() {
      const uint32_t left = members(uci), want = (c.nof_enc_csi_part2_bits - done_csi2) / B;
      if (l >= l1_csi && left > 0 && want > 0) {
        spacing(left, want, d, m);
        select(csi2, uci, d, m);
        for (uint32_t i = 0; i < M; ++i) {
          if (csi2[i]) {
            sch[i] = uci[i] = 0;
          }
        }
        done_csi2 += m * B;
      }
    };
    if (csi2_known) {
      place_csi2();
    }
    // HARQ-ACK of one or two bits goes on the reserved elements, on top of what is there
    if (n_rvd > 0 && c.nof_harq_ack_bits <= 2 && c.nof_harq_ack_bits > 0 && want_harq > 0) {
      spacing(n_rvd, want_harq, d, m);
      select(harq, rvd, d, m);
      done_harq += m * B;
    }
    // CSI Part 2 that only becomes known when CSI Part 1 has been received in full (pusch_processor_impl.cpp:73-101): the
    // reference places it from the symbol in which CSI Part 1 ends
    if (!csi2_known && c.csi2_after_csi1 && c.nof_enc_csi_part2_bits != 0 && c.nof_csi_part1_bits != 0 &&
        members(csi1) > 0 && done_csi1 == c.nof_enc_csi_part1_bits) {
      csi2_known = true;
      place_csi2();
    }
    if (done_csi2 == c.nof_enc_csi_part2_bits && members(csi2) > 0) {
      // complete: nothing more is placed (want == 0 from here on)
    }
    // record the symbol, in segments
    const uint32_t first = (uint32_t)P.flags.size();
    for (uint32_t i = 0; i < M; ++i) {
      P.flags.push_back((uint8_t)((sch[i] ? FEC_ULSCH_F_SCH : 0) | (harq[i] ? FEC_ULSCH_F_HARQ : 0) |
                                  (csi1[i] ? FEC_ULSCH_F_CSI1 : 0) | (csi2[i] ? FEC_ULSCH_F_CSI2 : 0)));
    }
    for (uint32_t s0 = 0; s0 < M; s0 += FEC_ULSCH_SEG) {
      fec_ulsch_seg S;
      S.first_re = first + s0;
      S.nof_re   = (M - s0 < FEC_ULSCH_SEG) ? (M - s0) : FEC_ULSCH_SEG;
      for (int k = 0; k < 4; ++k) {
        S.base[k] = P.nof_re[k];
      }
      for (uint32_t i = s0; i < s0 + S.nof_re; ++i) {
        const uint8_t f = P.flags[first + i];
        P.nof_re[0] += (f & FEC_ULSCH_F_SCH) ? 1 : 0;
        P.nof_re[1] += (f & FEC_ULSCH_F_HARQ) ? 1 : 0;
        P.nof_re[2] += (f & FEC_ULSCH_F_CSI1) ? 1 : 0;
        P.nof_re[3] += (f & FEC_ULSCH_F_CSI2) ? 1 : 0;
      }
      P.segs.push_back(S);
    }
  }
  return true;
}

#ifdef __CUDACC__
struct fec_ulsch_launch {
  uint32_t bits_per_re, qm;
  uint32_t mode[4]; // per stream: 0 plain copy, 1 one-bit payload, 2 two-bit payload (placeholder rules); [0] unused
  uint32_t harq_punctures; // HARQ-ACK of one or two bits: the element reads as zeros for the other streams
};

// One soft bit of the packed scrambling sequence (MSB first).
__device__ __forceinline__ uint32_t fec_ulsch_seq_bit(const uint8_t* __restrict__ seq, uint32_t p)
{
  return (seq[p >> 3] >> (7u - (p & 7u))) & 1u;
}

__global__ void __launch_bounds__(FEC_ULSCH_SEG)
fec_ulsch_demux_kernel(const fec_ulsch_seg* __restrict__ segs,
                       const uint8_t* __restrict__ flags,
                       fec_ulsch_launch L,
                       const int8_t* __restrict__ codeword,
                       const uint8_t* __restrict__ seq,
                       int8_t* __restrict__ o_sch,
                       int8_t* __restrict__ o_harq,
                       int8_t* __restrict__ o_csi1,
                       int8_t* __restrict__ o_csi2)
{
  __shared__ uint32_t warp_cnt[4][FEC_ULSCH_SEG / 32];
  const fec_ulsch_seg S    = segs[blockIdx.x];
  const uint32_t      t    = threadIdx.x, lane = t & 31u, warp = t >> 5;
  const uint32_t      f    = (t < S.nof_re) ? flags[S.first_re + t] : 0u;
  uint32_t            rank[4];
#pragma unroll
  for (int k = 0; k < 4; ++k) {
    const uint32_t bal = __ballot_sync(0xffffffffu, (f >> k) & 1u);
    rank[k]            = __popc(bal & ((1u << lane) - 1u));
    if (lane == 0) {
      warp_cnt[k][warp] = __popc(bal);
    }
  }
  __syncthreads();
#pragma unroll
  for (int k = 0; k < 4; ++k) {
    uint32_t before = 0;
    for (uint32_t w = 0; w < warp; ++w) {
      before += warp_cnt[k][w];
    }
    rank[k] += before + S.base[k];
  }
  if (f == 0) {
    return;
  }
  const uint32_t B   = L.bits_per_re;
  const uint32_t pos = (S.first_re + t) * B; // position of the element's first soft bit in the codeword
  int8_t*        out[4] = {o_sch, o_harq, o_csi1, o_csi2};
  // Fast path: an element whose soft bits are whole 32-bit words and that is copied as it is (no placeholder rules for the
  // streams it belongs to). The stream bases come from cudaMalloc-aligned buffers in the engine's own calls; a caller's
  // unaligned pointer takes the byte path.
  const bool punct = L.harq_punctures && (f & FEC_ULSCH_F_HARQ);
  bool       plain = (B % 4u == 0) && ((reinterpret_cast<uintptr_t>(codeword) & 3u) == 0);
#pragma unroll
  for (int k = 0; k < 4; ++k) {
    if ((f >> k) & 1u) {
      plain = plain && ((reinterpret_cast<uintptr_t>(out[k]) & 3u) == 0) && (k == 0 || L.mode[k] == 0 || L.qm == 1);
    }
  }
  if (plain) {
    const uint32_t  W   = B / 4u;
    const uint32_t* src = reinterpret_cast<const uint32_t*>(codeword + pos);
    uint32_t        xw[8];
#pragma unroll
    for (uint32_t w = 0; w < 8; ++w) {
      xw[w] = (w < W) ? src[w] : 0u;
    }
#pragma unroll
    for (int k = 1; k < 4; ++k) { // UCI streams first, then SCH: the order of demux_current_ofdm_symbol
      if ((f >> k) & 1u) {
        uint32_t*  o      = reinterpret_cast<uint32_t*>(out[k] + (size_t)rank[k] * B);
        const bool zeroed = (k != 1) && punct;
#pragma unroll
        for (uint32_t w = 0; w < 8; ++w) {
          if (w < W) {
            o[w] = zeroed ? 0u : xw[w];
          }
        }
      }
    }
    if (f & FEC_ULSCH_F_SCH) {
      uint32_t* o = reinterpret_cast<uint32_t*>(o_sch + (size_t)rank[0] * B);
#pragma unroll
      for (uint32_t w = 0; w < 8; ++w) {
        if (w < W) {
          o[w] = punct ? 0u : xw[w];
        }
      }
    }
    return;
  }
  int8_t x[32];
  for (uint32_t b = 0; b < B; ++b) {
    x[b] = codeword[pos + b];
  }
#pragma unroll
  for (int k = 1; k < 4; ++k) {
    if (!((f >> k) & 1u)) {
      continue;
    }
    int8_t*        o      = out[k] + (size_t)rank[k] * B;
    const bool     zeroed = (k != 1) && punct;
    const uint32_t mode   = L.mode[k];
    for (uint32_t b = 0; b < B; ++b) {
      int v = zeroed ? 0 : (int)x[b];
      if (mode != 0 && L.qm > 1 && !zeroed) {
        const uint32_t j = b % L.qm; // soft bit inside the modulation symbol of its layer
        if (j == 1) {
          if (mode == 1 && (fec_ulsch_seq_bit(seq, pos + b - 1) ^ fec_ulsch_seq_bit(seq, pos + b))) {
            v = -v;
          }
        } else if (j >= 2 && fec_ulsch_seq_bit(seq, pos + b)) {
          v = -v;
        }
      }
      o[b] = (int8_t)v;
    }
  }
  if (f & FEC_ULSCH_F_SCH) {
    int8_t* o = o_sch + (size_t)rank[0] * B;
    for (uint32_t b = 0; b < B; ++b) {
      o[b] = punct ? (int8_t)0 : x[b];
    }
  }
}
#endif
