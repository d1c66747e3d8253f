// srsran-5g_b200/csrc/fec_encode.cuh -- downlink twin of the decode path: one CTA per PDSCH codeblock
// does LDPC encoding + rate matching (bit selection + bit interleaving) and writes the rate-matched bits, packed or unpacked.
//
// What it computes (bit-exact with the reference; file:line relative to the reference tree):
//   LDPC encoding    lib/phy/upper/channel_coding/ldpc/ldpc_encoder_impl.cpp:64-120 (systematic part, then the parity bits) with
//                    ldpc_encoder_generic.cpp (high-rate region: four core parity blocks from the double-diagonal structure;
//                    extended region: one parity block per remaining base row), TS 38.212 5.3.2
//   rate matching    lib/phy/upper/channel_coding/ldpc/ldpc_rate_matcher_impl.cpp:36-160 (k0 from rv, circular buffer of length
//                    Ncb, filler bits skipped, row-column interleaver with Qm rows), TS 38.212 5.4.2
//
// Representation: every base-graph column of the codeword is a Z-bit vector kept in shared memory as 32-bit words, bit j of
// the column in bit (j & 31) of word (j >> 5). A cyclic shift of a column is then a funnel shift across two neighbouring
// words plus one wrap, i.e. 32 lifted XORs per instruction instead of one; a 316-edge base graph at Z = 384 takes 3 792
// word operations per codeblock. The E output bits are gathered from these vectors in the interleaved order, 32 bits per
// thread, with the position in the circular buffer advanced by additions only (no division per bit).
#pragma once
#include "fec_device.cuh"

#define FEC_ENC_THREADS 256
#define FEC_ENC_MAX_W 12                      // words per column at Z = 384
#define FEC_ENC_STRIDE (FEC_ENC_MAX_W + 1)    // + one zero word so that the funnel shift may read one word past the end

// The same struct is read by the host engine and by the kernel.
static_assert(sizeof(fec_enc_desc) == 40, "fec_enc_desc layout");

// 32 bits of the Z-bit vector v starting at bit `start` (< Z), cyclic. Bits of v at and above Z are zero, v[W] is zero.
__device__ __forceinline__ uint32_t fec_enc_bits32(const uint32_t* v, uint32_t start, uint32_t Z)
{
  if (Z <= 32) {
    unsigned long long t = v[0];
    t |= t << Z;
    return (uint32_t)(t >> start);
  }
  const uint32_t wi = start >> 5, sh = start & 31u;
  uint32_t       x  = __funnelshift_r(v[wi], v[wi + 1], sh);
  const uint32_t n1 = Z - start; // bits before the vector wraps
  if (n1 < 32) {
    x = (x & ((1u << n1) - 1u)) | (v[0] << n1);
  }
  return x;
}

// Word w of the vector y[c] = v[(c + s) mod Z] (the lifted edge of shift s), masked to the bits that exist.
__device__ __forceinline__ uint32_t fec_enc_rot_word(const uint32_t* v, uint32_t s, uint32_t w, uint32_t Z)
{
  uint32_t start = s + 32u * w;
  start          = (start >= Z) ? start - Z : start;
  const uint32_t nv = min(32u, Z - 32u * w);
  const uint32_t x  = fec_enc_bits32(v, start, Z);
  return (nv < 32) ? (x & ((1u << nv) - 1u)) : x;
}

__global__ void __launch_bounds__(FEC_ENC_THREADS)
fec_encode_kernel(const fec_enc_desc* __restrict__ descs,
                  const uint8_t* __restrict__ in_base,
                  uint8_t* __restrict__ out_base,
                  const fec_dev_tables* __restrict__ tab)
{
  __shared__ uint32_t V[68 * FEC_ENC_STRIDE];   // codeword, one Z-bit vector per base column (punctured columns included)
  __shared__ uint32_t T[46 * FEC_ENC_STRIDE];   // per base row: lambda (rows 0..3) / the parity block before its own shift
  __shared__ uint32_t S[FEC_ENC_STRIDE];        // sum of the four lambdas
  __shared__ uint32_t stage[FEC_ENC_THREADS];   // one chunk of packed output words (unpacked outputs are expanded from it)
  __shared__ uint32_t s_p0_shift;

  const fec_enc_desc D   = descs[blockIdx.x];
  const uint32_t     tid = threadIdx.x, nthr = blockDim.x;
  const uint32_t     Z = D.lifting_size, W = (Z + 31) / 32;
  const int          b = (D.base_graph == 1) ? 0 : 1;
  const uint32_t     Kb = b ? 10u : 22u, M = b ? 42u : 46u, nfull = Kb + M;
  const uint32_t     K = Kb * Z, N = (nfull - 2) * Z;
  const uint32_t     fillers = D.nof_filler_bits;
  const uint32_t     Kp      = K - fillers; // payload + CRC bits present in the input
  const uint16_t*    shift   = tab->shift[D.graph_index];
  const uint8_t*     colt    = tab->col[b];
  const uint16_t*    rstart  = tab->row_start[b];

  // ---- rate-matching geometry (same coordinates as the dematcher of the decode path) ---------------------------------------
  const uint32_t Ncb  = (D.Nref > 0) ? min(D.Nref, N) : N;
  const uint32_t sys  = (Kb - 2) * Z;
  const uint32_t info = sys - fillers;
  const uint32_t Lc   = Ncb - fillers;                                           // filler-free length of the circular buffer
  const uint32_t k0c  = (D.k0 < info) ? D.k0 : (max(D.k0, sys) - fillers);       // start, filler-free coordinate
  const uint32_t E    = D.E, qm = D.qm, Kq = E / qm;
  uint32_t       maxpos;
  if (E > Lc - k0c) {
    maxpos = Ncb - 1;
  } else {
    const uint32_t qe = k0c + E - 1;
    maxpos            = (qe < info) ? qe : qe + fillers;
  }
  uint32_t rows_needed = (maxpos + 2 * Z) / Z + 1;             // columns needed ...
  rows_needed          = (rows_needed > Kb) ? rows_needed - Kb : 0; // ... rows that produce them
  rows_needed          = min(max(rows_needed, 4u), M);

  // ---- systematic part: unpack the segment into column vectors; filler bits are zeros for the encoder ------------------
  const uint8_t* in      = in_base + D.in_offset;
  const uint32_t in_len  = (Kp + 7) / 8;
  for (uint32_t i = tid; i < Kb * FEC_ENC_STRIDE; i += nthr) {
    const uint32_t c = i / FEC_ENC_STRIDE, w = i - c * FEC_ENC_STRIDE;
    uint32_t       x = 0;
    if (w < W) {
      const uint32_t g  = c * Z + 32 * w; // first message bit of this word
      const uint32_t nv = min(32u, Z - 32 * w);
      if (g < Kp) {
        const uint32_t     b0 = g >> 3;
        unsigned long long B  = 0;
#pragma unroll
        for (uint32_t k = 0; k < 5; ++k) {
          const uint32_t byte = (b0 + k < in_len) ? in[b0 + k] : 0u;
          B |= (unsigned long long)byte << (32 - 8 * k);
        }
        const uint32_t msb = (uint32_t)((B << (g & 7u)) >> 8); // bit 31 = message bit g
        x                  = __brev(msb);
        const uint32_t nb  = min(nv, Kp - g);                  // bits that are payload (the rest: fillers / next column)
        x                  = (nb < 32) ? (x & ((1u << nb) - 1u)) : x;
      }
    }
    V[i] = x;
  }
  for (uint32_t i = tid; i < (nfull - Kb); i += nthr) {
    V[(Kb + i) * FEC_ENC_STRIDE + W] = 0; // padding word of the parity columns
    T[min(i, 45u) * FEC_ENC_STRIDE + W] = 0;
  }
  if (tid == 0) {
    S[W] = 0;
    // Column Kb appears three times in the four core rows, twice with the same shift: the odd one out is the shift that
    // remains on p0 when the four rows are added up (the double diagonal cancels).
    uint32_t sh[3] = {0, 0, 0};
    uint32_t n     = 0;
    for (uint32_t e = rstart[0]; e < rstart[4]; ++e) {
      if (colt[e] == Kb && n < 3) {
        sh[n++] = shift[e];
      }
    }
    s_p0_shift = (sh[0] == sh[1]) ? sh[2] : ((sh[0] == sh[2]) ? sh[1] : sh[0]);
  }
  __syncthreads();

  // ---- lambda_i = sum of the systematic edges of core row i ---------------------------------------------------------------------
  for (uint32_t i = tid; i < 4 * W; i += nthr) {
    const uint32_t r = i / W, w = i - r * W;
    uint32_t       acc = 0;
    for (uint32_t e = rstart[r]; e < rstart[r + 1]; ++e) {
      const uint32_t c = colt[e];
      if (c < Kb) {
        acc ^= fec_enc_rot_word(&V[c * FEC_ENC_STRIDE], shift[e], w, Z);
      }
    }
    T[r * FEC_ENC_STRIDE + w] = acc;
  }
  __syncthreads();
  for (uint32_t w = tid; w < W; w += nthr) {
    S[w] = T[w] ^ T[FEC_ENC_STRIDE + w] ^ T[2 * FEC_ENC_STRIDE + w] ^ T[3 * FEC_ENC_STRIDE + w];
  }
  __syncthreads();
  // p0[(c + r) mod Z] = S[c]
  for (uint32_t w = tid; w < W; w += nthr) {
    const uint32_t r = s_p0_shift;
    V[Kb * FEC_ENC_STRIDE + w] = fec_enc_rot_word(S, (r == 0) ? 0u : Z - r, w, Z);
  }
  __syncthreads();
  // Rows 0..2: the one unknown block of row i is column Kb + 1 + i.
  for (uint32_t r = 0; r < 3; ++r) {
    uint32_t s_unknown = 0;
    for (uint32_t w = tid; w < W; w += nthr) {
      uint32_t acc = T[r * FEC_ENC_STRIDE + w];
      for (uint32_t e = rstart[r]; e < rstart[r + 1]; ++e) {
        const uint32_t c = colt[e];
        if (c >= Kb && c != Kb + 1 + r) {
          acc ^= fec_enc_rot_word(&V[c * FEC_ENC_STRIDE], shift[e], w, Z);
        }
      }
      S[w] = acc;
    }
    for (uint32_t e = rstart[r]; e < rstart[r + 1]; ++e) {
      if (colt[e] == Kb + 1 + r) {
        s_unknown = shift[e];
      }
    }
    __syncthreads();
    for (uint32_t w = tid; w < W; w += nthr) {
      V[(Kb + 1 + r) * FEC_ENC_STRIDE + w] = fec_enc_rot_word(S, (s_unknown == 0) ? 0u : Z - s_unknown, w, Z);
    }
    __syncthreads();
  }
  // ---- extension rows: the last edge of row i >= 4 is its own degree-1 parity column Kb + i ---------------------------------
  for (uint32_t i = tid; i < (rows_needed - 4) * W; i += nthr) {
    const uint32_t r = 4 + i / W, w = i - (r - 4) * W;
    uint32_t       acc = 0;
    for (uint32_t e = rstart[r]; e < rstart[r + 1]; ++e) {
      const uint32_t c = colt[e];
      if (c != Kb + r) {
        acc ^= fec_enc_rot_word(&V[c * FEC_ENC_STRIDE], shift[e], w, Z);
      }
    }
    T[r * FEC_ENC_STRIDE + w] = acc;
  }
  __syncthreads();
  for (uint32_t i = tid; i < (rows_needed - 4) * W; i += nthr) {
    const uint32_t r = 4 + i / W, w = i - (r - 4) * W;
    uint32_t       s_unknown = 0;
    for (uint32_t e = rstart[r]; e < rstart[r + 1]; ++e) {
      if (colt[e] == Kb + r) {
        s_unknown = shift[e];
      }
    }
    V[(Kb + r) * FEC_ENC_STRIDE + w] = fec_enc_rot_word(&T[r * FEC_ENC_STRIDE], (s_unknown == 0) ? 0u : Z - s_unknown, w, Z);
  }
  __syncthreads();

  // ---- bit selection + interleaving + output ---------------------------------------------------------------------------------
  // Output bit n = i * Qm + j is rate-matched bit p = j * (E / Qm) + i, read at circular-buffer position (k0c + p) mod Lc
  // (filler-free coordinate). Each thread produces 32 consecutive output bits; q follows n by additions only.
  const uint32_t invZ    = 0xffffffffu / Z + 1;
  const uint32_t kq_mod  = Kq % Lc;
  const bool     unpack  = (D.flags & FEC_ENC_FLAG_UNPACKED) != 0;
  uint8_t*       out     = out_base + D.out_offset;
  const uint32_t nwords  = (E + 31) / 32;
  const bool     aligned = (reinterpret_cast<uintptr_t>(out) & 3u) == 0;
  for (uint32_t w0 = 0; w0 < nwords; w0 += nthr) {
    const uint32_t wi = w0 + tid;
    uint32_t       be = 0; // bits MSB first
    if (wi < nwords) {
      const uint32_t n0 = wi * 32;
      uint32_t       i  = n0 / qm, j = n0 - i * qm;
      uint32_t       qi = (k0c + i) % Lc;                                        // position of p = i (row 0)
      uint32_t       q  = (uint32_t)(((unsigned long long)qi + (unsigned long long)j * kq_mod) % Lc);
      const uint32_t nb = min(32u, E - n0);
      for (uint32_t k = 0; k < nb; ++k) {
        const uint32_t pos = (q < info) ? q : q + fillers;
        const uint32_t f   = pos + 2 * Z;
        const uint32_t c   = __umulhi(f, invZ);
        const uint32_t jj  = f - c * Z;
        const uint32_t bit = (V[c * FEC_ENC_STRIDE + (jj >> 5)] >> (jj & 31u)) & 1u;
        be |= bit << (31 - k);
        if (++j == qm) {
          j  = 0;
          qi = (qi + 1 == Lc) ? 0u : qi + 1;
          q  = qi;
        } else {
          q += kq_mod;
          q = (q >= Lc) ? q - Lc : q;
        }
      }
    }
    if (!unpack) {
      if (wi < nwords) {
        const uint32_t nbytes = min(4u, (E - wi * 32 + 7) / 8);
        if (aligned && nbytes == 4) {
          reinterpret_cast<uint32_t*>(out)[wi] = __byte_perm(be, 0, 0x0123);
        } else {
          for (uint32_t k = 0; k < nbytes; ++k) {
            out[wi * 4 + k] = (uint8_t)(be >> (24 - 8 * k));
          }
        }
      }
    } else {
      stage[tid] = be;
      __syncthreads();
      // one bit per byte, coalesced: this chunk covers output bits [w0 * 32, w0 * 32 + 32 * nthr)
      const uint32_t base = w0 * 32;
      const uint32_t cnt  = min(32u * nthr, E - base);
      if (aligned) {
        for (uint32_t g = tid; g < cnt / 4; g += nthr) {
          const uint32_t nib = (stage[g >> 3] >> (28 - 4 * (g & 7u))) & 0xfu; // 4 consecutive bits, MSB first
          reinterpret_cast<uint32_t*>(out + base)[g] =
              ((nib >> 3) & 1u) | (((nib >> 2) & 1u) << 8) | (((nib >> 1) & 1u) << 16) | ((nib & 1u) << 24);
        }
        for (uint32_t g = (cnt & ~3u) + tid; g < cnt; g += nthr) {
          out[base + g] = (uint8_t)((stage[g >> 5] >> (31 - (g & 31u))) & 1u);
        }
      } else {
        for (uint32_t g = tid; g < cnt; g += nthr) {
          out[base + g] = (uint8_t)((stage[g >> 5] >> (31 - (g & 31u))) & 1u);
        }
      }
      __syncthreads();
    }
  }
}
