// srsran-5g_b200/csrc/fec_tb.cuh -- transport-block layer: receive-side segmentation (host), TB assembly + TB CRC (device).
//
// Reference behaviour restated here (file:line relative to the reference tree):
//   number of codeblocks, lifting size, TB CRC size     include/srsran/phy/upper/channel_coding/ldpc/ldpc.h:128-208
//   segment lengths, fillers, rate-matched lengths      lib/phy/upper/channel_coding/ldpc/ldpc_segmenter_rx_impl.cpp:53-106
//                                                       lib/phy/upper/channel_coding/ldpc/ldpc_segmenter_helpers.h:82-118
//   payload bits per codeblock, CRC type per codeblock  lib/phy/upper/channel_processors/pusch/pusch_decoder_hw_impl.cpp:38-98
//   TB assembly, TB CRC, delivery rules                 lib/phy/upper/channel_processors/pusch/pusch_decoder_hw_impl.cpp:328-425
#pragma once
#include "fec_device.cuh"

struct fec_tb_geometry {
  uint32_t nof_cbs;        // C
  uint32_t lifting_size;   // Z
  uint32_t K;              // segment length (22 Z or 10 Z)
  uint32_t N;              // full codeblock length (66 Z or 50 Z)
  uint32_t tb_crc_bits;    // 16 or 24
  uint32_t cb_crc_bits;    // 24 when C > 1, else 0
  uint32_t cb_msg_bits;    // ceil(B' / C): payload + zero padding + codeblock CRC, i.e. K - fillers
  uint32_t nof_filler_bits;
  uint32_t data_bits;      // cb_msg_bits - (C == 1 ? tb_crc_bits : 24): bits a codeblock contributes to the stream
  uint32_t stream_bits;    // tbs (C == 1) or tbs + 24 (C > 1): what the codeblock payloads are concatenated into
  uint32_t crc_type;       // FEC_CRC16 / FEC_CRC24A (C == 1) or FEC_CRC24B
  uint32_t sym_per_layer, nof_short; // rate-matched lengths: floor for r < nof_short, ceil otherwise
};

static const uint16_t FEC_TB_ALL_Z[51] = {2,  3,  4,  5,  6,  7,  8,  9,  10,  11,  12,  13,  14,  15,  16,  18,  20,
                                          22, 24, 26, 28, 30, 32, 36, 40, 44,  48,  52,  56,  60,  64,  72,  80,  88,
                                          96, 104, 112, 120, 128, 144, 160, 176, 192, 208, 224, 240, 256, 288, 320, 352, 384};

// Returns false when the inputs violate the reference's preconditions (ldpc_segmenter_rx_impl.cpp:33-51).
static inline bool fec_tb_geometry_of(const fec_tb_desc& tb, fec_tb_geometry& g)
{
  if ((tb.base_graph != 1 && tb.base_graph != 2) || tb.rv > 3 || tb.nof_layers < 1 || tb.nof_layers > 4 || tb.tbs == 0 ||
      (tb.tbs % 8) != 0 || (tb.qm != 1 && tb.qm != 2 && tb.qm != 4 && tb.qm != 6 && tb.qm != 8) || tb.nof_llrs == 0 ||
      (tb.nof_llrs % tb.qm) != 0 || ((tb.nof_llrs / tb.qm) % tb.nof_layers) != 0 || tb.max_iter == 0) {
    return false;
  }
  const bool     bg1     = (tb.base_graph == 1);
  g.tb_crc_bits          = (tb.tbs <= 3824) ? 16 : 24;
  const uint32_t B       = tb.tbs + g.tb_crc_bits;
  const uint32_t max_seg = bg1 ? 8448 : 3840;
  g.nof_cbs              = (B <= max_seg) ? 1 : (B + (max_seg - 24) - 1) / (max_seg - 24);
  if (g.nof_cbs > FEC_MAX_TB_CBS) {
    return false;
  }
  const uint32_t B_out = B + ((g.nof_cbs > 1) ? 24 * g.nof_cbs : 0);
  uint32_t       ref   = 22;
  if (!bg1) {
    ref = (B > 640) ? 10 : (B > 560) ? 9 : (B > 192) ? 8 : 6;
  }
  g.lifting_size = 0;
  for (int i = 0; i < 51; ++i) {
    if ((uint32_t)FEC_TB_ALL_Z[i] * g.nof_cbs * ref >= B_out) {
      g.lifting_size = FEC_TB_ALL_Z[i];
      break;
    }
  }
  if (g.lifting_size == 0) {
    return false;
  }
  g.K               = (bg1 ? 22 : 10) * g.lifting_size;
  g.N               = (bg1 ? 66 : 50) * g.lifting_size;
  g.cb_crc_bits     = (g.nof_cbs > 1) ? 24 : 0;
  g.cb_msg_bits     = (B_out + g.nof_cbs - 1) / g.nof_cbs;
  g.nof_filler_bits = g.K - g.cb_msg_bits;
  g.data_bits       = g.cb_msg_bits - ((g.nof_cbs == 1) ? g.tb_crc_bits : 24);
  g.stream_bits     = tb.tbs + ((g.nof_cbs > 1) ? 24 : 0);
  g.crc_type        = (g.nof_cbs > 1) ? FEC_CRC24B : ((tb.tbs > 3824) ? FEC_CRC24A : FEC_CRC16);
  g.sym_per_layer   = tb.nof_llrs / tb.qm / tb.nof_layers;
  g.nof_short       = g.nof_cbs - (g.sym_per_layer % g.nof_cbs);
  return true;
}

// Rate-matched length of codeblock r (TS 38.212 5.4.2.1 through ldpc_segmenter_helpers.h:82-94).
static inline uint32_t fec_tb_rm_length(const fec_tb_desc& tb, const fec_tb_geometry& g, uint32_t r)
{
  const uint32_t per = (r < g.nof_short) ? (g.sym_per_layer / g.nof_cbs) : ((g.sym_per_layer + g.nof_cbs - 1) / g.nof_cbs);
  return per * tb.nof_layers * tb.qm;
}

// ---- device side --------------------------------------------------------------------------------------------------------
// One record per transport block of a call.
struct fec_tb_job {
  uint64_t out_offset;   // TB bytes in the output buffer
  uint32_t nof_cbs;
  uint32_t data_bits;    // bits per codeblock in the stream (the last one contributes what is left)
  uint32_t stream_bits;  // tbs (+ 24)
  uint32_t tbs_bytes;
  uint32_t map_first;    // first of nof_cbs entries in the codeblock map
  uint32_t max_iter;
  uint32_t reserved[2];
};

// Codeblock map entry: where the decoded message lives and whether it was decoded in this call.
struct fec_tb_cb_ref {
  uint32_t slot;         // message store slot (= HARQ slot)
  int32_t  result_index; // index into the results of this launch, or -1: CRC passed in an earlier transmission
  uint32_t tb;           // index of the transport block (job) this codeblock belongs to
  uint32_t weight;       // x^(8 * stream bytes after this codeblock) mod g24A (byte-aligned payloads only)
};

// Host side of the weights: (a * b) mod g24A.
static inline uint32_t fec_tb_crc24a_mulmod_host(uint32_t a, uint32_t b)
{
  uint32_t acc = 0;
  for (int i = 0; i < 24; ++i) {
    acc ^= (b & 1u) ? a : 0u;
    b >>= 1;
    a <<= 1;
    a = (a & (1u << 24)) ? (a ^ 0x1864CFBu) : a;
  }
  return acc;
}

// x^(8 n) mod g24A by square and multiply.
static inline uint32_t fec_tb_crc24a_xpow8_host(uint32_t n)
{
  uint32_t w = 1, base = 0x100;
  while (n != 0) {
    if (n & 1u) {
      w = fec_tb_crc24a_mulmod_host(w, base);
    }
    base = fec_tb_crc24a_mulmod_host(base, base);
    n >>= 1;
  }
  return w;
}

// 8 bits of the stream starting at bit b (b + 8 <= stream_bits not required: bits past the end read as zero).
__device__ __forceinline__ uint32_t fec_tb_stream_byte(const uint8_t* __restrict__ msg_store, const fec_tb_cb_ref* __restrict__ map,
                                                       const fec_tb_job& J, uint32_t b)
{
  uint32_t out = 0, got = 0;
  while (got < 8 && b < J.stream_bits) {
    const uint32_t r    = min(b / J.data_bits, J.nof_cbs - 1);
    const uint32_t off  = b - r * J.data_bits;
    const uint32_t lim  = (r == J.nof_cbs - 1) ? (J.stream_bits - r * J.data_bits) : J.data_bits; // bits of this codeblock
    const uint32_t take = min(8u - got, lim - off);
    const uint8_t* m    = msg_store + (size_t)map[J.map_first + r].slot * FEC_MAX_CB_OUT_BYTES;
    // `take` bits starting at bit `off` of the MSB-first message
    const uint32_t two  = ((uint32_t)m[off >> 3] << 8) | (((off >> 3) + 1 < FEC_MAX_CB_OUT_BYTES) ? m[(off >> 3) + 1] : 0u);
    const uint32_t bits = (two >> (16 - (off & 7) - take)) & ((1u << take) - 1u);
    out |= bits << (8 - got - take);
    got += take;
    b += take;
  }
  return out;
}

// (a * b) mod g for polynomials of degree < 24 (CRC24A arithmetic).
__device__ __forceinline__ uint32_t fec_crc24a_mulmod(uint32_t a, uint32_t b)
{
  uint32_t acc = 0;
  for (int i = 0; i < 24; ++i) {
    acc ^= (b & 1u) ? a : 0u;
    b >>= 1;
    a = fec_crc_shift1(a, 0x1864CFBu, 1u << 24);
  }
  return acc;
}

// One CTA per transport block: all-codeblocks-good test, payload concatenation, TB CRC24A, result record.
__global__ void __launch_bounds__(1024) fec_tb_assemble_kernel(const fec_tb_job* __restrict__ jobs,
                                                             const fec_tb_cb_ref* __restrict__ map,
                                                             const fec_cb_result* __restrict__ cb_results,
                                                             const uint8_t* __restrict__ msg_store,
                                                             uint8_t* __restrict__ out_base,
                                                             fec_tb_result* __restrict__ tb_results)
{
  __shared__ uint32_t lut[256];
  __shared__ uint32_t red[32];
  __shared__ uint32_t s_stats[4]; // ok count, decoded count, iteration sum, (unused)
  __shared__ uint32_t s_min, s_max;
  const fec_tb_job J   = jobs[blockIdx.x];
  const uint32_t   tid = threadIdx.x, nth = blockDim.x;
  if (tid < 4) {
    s_stats[tid] = 0;
  }
  if (tid == 0) {
    s_min = 0xffffffffu;
    s_max = 0;
  }
  // CRC24A byte table: lut[v] = (v(x) * x^24) mod g
  for (uint32_t v = tid; v < 256; v += nth) {
    uint32_t r = v << 16;
    for (int i = 0; i < 8; ++i) {
      r = fec_crc_shift1(r, 0x1864CFBu, 1u << 24);
    }
    lut[v] = r;
  }
  __syncthreads();
  for (uint32_t r = tid; r < J.nof_cbs; r += nth) {
    const fec_tb_cb_ref ref = map[J.map_first + r];
    bool                ok  = true;
    if (ref.result_index >= 0) {
      const fec_cb_result cr = cb_results[ref.result_index];
      ok                     = (cr.status & FEC_RES_CRC_PASS) != 0;
      const uint32_t it      = ok ? cr.nof_ldpc_iterations : J.max_iter;
      atomicAdd(&s_stats[1], 1u);
      atomicAdd(&s_stats[2], it);
      atomicMin(&s_min, it);
      atomicMax(&s_max, it);
    }
    if (ok) {
      atomicAdd(&s_stats[0], 1u);
    }
  }
  __syncthreads();
  const bool all_ok = (s_stats[0] == J.nof_cbs);
  bool       tb_ok  = all_ok;
  if (all_ok) {
    uint8_t*       out     = out_base + J.out_offset;
    const uint32_t nbytes  = (J.stream_bits + 7) / 8; // tbs / 8 (+ 3)
    // Contiguous chunk of the stream per thread; remainder by the byte table, then weighted by x^(8 * bytes after it).
    const uint32_t per     = (nbytes + nth - 1) / nth;
    const uint32_t b0      = min(tid * per, nbytes), b1 = min(b0 + per, nbytes);
    uint32_t       rem     = 0;
    if ((J.data_bits & 7u) == 0) {
      // Byte-aligned payloads (every TBS of TS 38.214 5.1.3.2): stream byte i is byte (i - r * dbytes) of codeblock r.
      const uint32_t dbytes = J.data_bits / 8;
      uint32_t       r      = min(b0 / dbytes, J.nof_cbs - 1);
      uint32_t       ob     = b0 - r * dbytes;
      const uint8_t* m      = msg_store + (size_t)map[J.map_first + r].slot * FEC_MAX_CB_OUT_BYTES;
      for (uint32_t i = b0; i < b1; ++i) {
        if (ob == dbytes && r + 1 < J.nof_cbs) {
          ++r;
          ob = 0;
          m  = msg_store + (size_t)map[J.map_first + r].slot * FEC_MAX_CB_OUT_BYTES;
        }
        const uint32_t v = m[ob++];
        if (i < J.tbs_bytes) {
          out[i] = (uint8_t)v;
        }
        rem = ((rem << 8) & 0xffffffu) ^ lut[((rem >> 16) ^ v) & 0xffu];
      }
    } else {
      for (uint32_t i = b0; i < b1; ++i) {
        const uint32_t v = fec_tb_stream_byte(msg_store, map, J, 8 * i);
        if (i < J.tbs_bytes) {
          out[i] = (uint8_t)v;
        }
        rem = ((rem << 8) & 0xffffffu) ^ lut[((rem >> 16) ^ v) & 0xffu];
      }
    }
    if (J.nof_cbs > 1) {
      // rem is (chunk(x) * x^24) mod g; the stream remainder needs chunk(x) * x^(8 * after): the common factor x^24 does not
      // change whether the total is zero (g has a non-zero constant term, so x is invertible mod g).
      uint32_t after = nbytes - b1, w = 1, base = 0x100; // base = x^8
      while (after != 0) {
        if (after & 1u) {
          w = fec_crc24a_mulmod(w, base);
        }
        base = fec_crc24a_mulmod(base, base);
        after >>= 1;
      }
      uint32_t acc = (b1 > b0) ? fec_crc24a_mulmod(rem, w) : 0u;
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) {
        acc ^= __shfl_xor_sync(0xffffffffu, acc, o);
      }
      if ((tid & 31) == 0) {
        red[tid >> 5] = acc;
      }
      __syncthreads();
      uint32_t total = 0;
      for (uint32_t i = 0; i < (nth >> 5); ++i) {
        total ^= red[i];
      }
      tb_ok = (total == 0);
    }
  }
  if (tid == 0) {
    fec_tb_result R;
    R.nof_codeblocks = (uint16_t)J.nof_cbs;
    R.nof_cb_decoded = (uint16_t)s_stats[1];
    R.nof_cb_crc_ok  = (uint16_t)s_stats[0];
    R.tb_crc_ok      = tb_ok ? 1 : 0;
    R.tb_delivered   = all_ok ? 1 : 0;
    R.iter_sum       = s_stats[2];
    R.iter_min       = (uint16_t)(s_stats[1] ? s_min : 0);
    R.iter_max       = (uint16_t)s_max;
    tb_results[blockIdx.x] = R;
  }
}

// Byte-aligned payloads (every TBS of TS 38.214): ONE CTA PER CODEBLOCK copies its payload into the transport block and
// folds the payload's CRC24A remainder, weighted by its position, into a per-block accumulator; the last CTA of a block
// to finish writes the result record. scratch: 2 words per transport block {accumulator, arrival counter}, zeroed by the
// caller before the launch.
__global__ void __launch_bounds__(256) fec_tb_gather_kernel(const fec_tb_job* __restrict__ jobs,
                                                           const fec_tb_cb_ref* __restrict__ map,
                                                           const fec_cb_result* __restrict__ cb_results,
                                                           const uint8_t* __restrict__ msg_store,
                                                           uint8_t* __restrict__ out_base,
                                                           fec_tb_result* __restrict__ tb_results,
                                                           uint32_t* __restrict__ scratch,
                                                           const fec_dev_tables* __restrict__ tab,
                                                           uint32_t first)
{
  __shared__ uint32_t red[8];
  __shared__ uint32_t s_stats[3]; // ok count, decoded count, iteration sum
  __shared__ uint32_t s_min, s_max, s_last;
  const uint32_t      mi  = blockIdx.x + first; // the launch covers the codeblocks [first, first + gridDim.x) of the map
  const fec_tb_cb_ref me  = map[mi];
  const fec_tb_job    J   = jobs[me.tb];
  const uint32_t      r   = mi - J.map_first;
  const uint32_t      tid = threadIdx.x;
  if (tid < 3) {
    s_stats[tid] = 0;
  }
  if (tid == 0) {
    s_min = 0xffffffffu;
    s_max = 0;
  }
  __syncthreads();
  for (uint32_t i = tid; i < J.nof_cbs; i += blockDim.x) {
    const fec_tb_cb_ref ref = map[J.map_first + i];
    bool                ok  = true;
    if (ref.result_index >= 0) {
      const fec_cb_result cr = cb_results[ref.result_index];
      ok                     = (cr.status & FEC_RES_CRC_PASS) != 0;
      const uint32_t it      = ok ? cr.nof_ldpc_iterations : J.max_iter;
      atomicAdd(&s_stats[1], 1u);
      atomicAdd(&s_stats[2], it);
      atomicMin(&s_min, it);
      atomicMax(&s_max, it);
    }
    if (ok) {
      atomicAdd(&s_stats[0], 1u);
    }
  }
  __syncthreads();
  const bool all_ok = (s_stats[0] == J.nof_cbs);
  uint32_t   acc    = 0;
  if (all_ok) {
    const uint32_t dbytes = J.data_bits / 8, nstream = J.stream_bits / 8;
    const uint32_t first  = r * dbytes;                                   // stream position of this payload
    const uint32_t nb     = min(dbytes, nstream - first);                 // the last codeblock contributes what is left
    const uint8_t* m      = msg_store + (size_t)me.slot * FEC_MAX_CB_OUT_BYTES;
    uint8_t*       out    = out_base + J.out_offset;
    const uint32_t poly   = 0x1864CFBu, top = 1u << 24;
    for (uint32_t i = tid; i < nb; i += blockDim.x) {
      const uint32_t v = m[i];
      if (first + i < J.tbs_bytes) {
        out[first + i] = (uint8_t)v;
      }
      if (J.nof_cbs > 1) {
        uint32_t a = tab->crc_x8[FEC_CRC24A][nb - 1 - i];
#pragma unroll
        for (int s = 0; s < 8; ++s) {
          acc ^= ((v >> s) & 1u) ? a : 0u;
          a = fec_crc_shift1(a, poly, top);
        }
      }
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
      acc ^= __shfl_xor_sync(0xffffffffu, acc, o);
    }
    if ((tid & 31) == 0) {
      red[tid >> 5] = acc;
    }
  }
  __syncthreads();
  if (tid == 0) {
    if (all_ok && J.nof_cbs > 1) {
      uint32_t rem = 0;
      for (uint32_t i = 0; i < (blockDim.x >> 5); ++i) {
        rem ^= red[i];
      }
      atomicXor(&scratch[2 * me.tb], fec_crc24a_mulmod(rem, me.weight));
    }
    __threadfence();
    s_last = (atomicAdd(&scratch[2 * me.tb + 1], 1u) == J.nof_cbs - 1) ? 1u : 0u;
    if (s_last) {
      __threadfence();
      const uint32_t total = atomicXor(&scratch[2 * me.tb], 0u);
      fec_tb_result  R;
      R.nof_codeblocks = (uint16_t)J.nof_cbs;
      R.nof_cb_decoded = (uint16_t)s_stats[1];
      R.nof_cb_crc_ok  = (uint16_t)s_stats[0];
      R.tb_crc_ok      = (all_ok && total == 0) ? 1 : 0;
      R.tb_delivered   = all_ok ? 1 : 0;
      R.iter_sum       = s_stats[2];
      R.iter_min       = (uint16_t)(s_stats[1] ? s_min : 0);
      R.iter_max       = (uint16_t)s_max;
      tb_results[me.tb] = R;
    }
  }
}
