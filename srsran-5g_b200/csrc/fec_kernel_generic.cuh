// srsran-5g_b200/csrc/fec_kernel_generic.cuh -- one CTA per codeblock, one thread per lifted check node.
//
// Works for every (base graph, lifting size) pair and every mode; it is the correctness anchor and the path taken for
// lifting sizes the packed kernel does not cover. Data layout per CTA (shared memory):
//   soft[n_full * Z]  int8   a-posteriori LLRs, column-major by base column (column b at soft + b*Z)
//   stateA[M * Z]     u32    compressed check-to-variable state of lifted check (row, c):
//                            bits 0-6 scaled min1, 8-14 scaled min2, 16-20 index of the min1 edge,
//                            bits 21-31 per-edge c2v sign bits for rows of degree <= 11
//   stateB[4 * Z]     u32    per-edge c2v sign bits of BG1 rows 0-3 (degree 19)
//   bitsw[K/32 + 1]   u32    hard decisions, big-endian bit order (bit i of the message = bit 31-(i%32) of word i/32)
// All-zero state reproduces the reference's "check-to-variable messages not initialised" case (v2c = soft).
#pragma once
#include "fec_device.cuh"

struct fec_generic_smem_layout {
  uint32_t stateA, stateB, bitsw, red, sh, rs, col, deg, lut, soft, total;
};

__host__ __device__ inline fec_generic_smem_layout fec_generic_layout(uint32_t Z, int bg1)
{
  fec_generic_smem_layout L;
  uint32_t                M = bg1 ? 46u : 42u, nfull = bg1 ? 68u : 52u, Kb = bg1 ? 22u : 10u;
  uint32_t                o = 0;
  L.stateA                  = o;
  o += M * Z * 4;
  L.stateB = o;
  o += (bg1 ? 4u : 0u) * Z * 4;
  L.bitsw = o;
  o += ((Kb * Z + 31) / 32 + 1) * 4;
  L.red = o;
  o += 64 * 4;
  L.sh = o;
  o += 320 * 2;
  L.rs = o;
  o += 48 * 2;
  L.col = o;
  o += 320;
  L.deg = o;
  o += 48;
  L.lut = o;
  o += 128;
  L.soft = o;
  o += (nfull * Z + 15) & ~15u;
  L.total = o;
  return L;
}

// Processes lifted check c of base row `row`: v2c, min1/min2/sign analysis, new c2v, soft-bit update.
template <int MAXDEG>
__device__ __forceinline__ void fec_generic_row(int8_t*         soft,
                                                uint32_t*       stateA,
                                                uint32_t*       stateB,
                                                const uint16_t* sh,
                                                const uint8_t*  col,
                                                const uint8_t*  lut,
                                                uint32_t        row,
                                                uint32_t        e0,
                                                uint32_t        d,
                                                uint32_t        c,
                                                uint32_t        Z)
{
  const uint32_t A     = stateA[row * Z + c];
  const int      m1s   = A & 0x7f;
  const int      m2s   = (A >> 8) & 0x7f;
  const uint32_t oidx  = (A >> 16) & 0x1f;
  const uint32_t osign = (MAXDEG > 11) ? stateB[row * Z + c] : (A >> 21);

  int      v[MAXDEG];
  uint32_t addr[MAXDEG];
  int      min1 = 120, min2 = 120;
  uint32_t nidx = 0, vsign = 0, sg = 0;
#pragma unroll
  for (int e = 0; e < MAXDEG; ++e) {
    if (e < (int)d) {
      uint32_t pos = c + sh[e0 + e];
      pos          = (pos >= Z) ? pos - Z : pos;
      uint32_t a_  = (uint32_t)col[e0 + e] * Z + pos;
      addr[e]      = a_;
      const int s  = soft[a_];
      const int mg = ((uint32_t)e == oidx) ? m2s : m1s;
      const int co = ((osign >> e) & 1u) ? -mg : mg;
      int       t  = fec_clampi(s - co, -120, 120);
      t            = (s == 127 || s == -127) ? s : t;
      v[e]         = t;
      const int a  = abs(t);
      const uint32_t neg = (t < 0) ? 1u : 0u;
      if (a < min1) {
        min2 = min1;
        min1 = a;
        nidx = e;
      } else if (a < min2) {
        min2 = a;
      }
      sg ^= neg;
      vsign |= neg << e;
    }
  }
  const int      n1    = lut[min1];
  const int      n2    = lut[min2];
  const uint32_t nsign = (sg ? ~vsign : vsign) & ((1u << d) - 1u);
#pragma unroll
  for (int e = 0; e < MAXDEG; ++e) {
    if (e < (int)d) {
      const int mg = ((uint32_t)e == nidx) ? n2 : n1;
      const int cn = ((nsign >> e) & 1u) ? -mg : mg;
      const int t  = v[e];
      int       s  = cn + t;
      s            = (s > 120) ? 127 : ((s < -120) ? -127 : s);
      s            = (t == 127 || t == -127) ? t : s;
      soft[addr[e]] = (int8_t)s;
    }
  }
  if (MAXDEG > 11) {
    stateA[row * Z + c] = (uint32_t)n1 | ((uint32_t)n2 << 8) | (nidx << 16);
    stateB[row * Z + c] = nsign;
  } else {
    stateA[row * Z + c] = (uint32_t)n1 | ((uint32_t)n2 << 8) | (nidx << 16) | (nsign << 21);
  }
}

// Rate dematcher for one codeblock, executed by the whole CTA. Writes the updated soft buffer to `harq` (N bytes) and,
// when `soft` is not null, the clamped copy the decoder starts from. Returns through *n_sig_smem (atomicMax) the
// number of LLRs left after trimming trailing zeros.
// Where the decoder's working copy of the dematched soft bits goes: natural int8 layout (generic kernel) or nothing.
struct fec_soft_store_natural {
  int8_t*  soft; // nullptr: dematch only
  uint32_t Z;
  __device__ __forceinline__ void operator()(uint32_t o, int acc) const
  {
    if (soft != nullptr) {
      soft[2 * Z + o] = (int8_t)fec_clampi(acc, -64, 64);
    }
  }
};

template <class SoftStore>
__device__ __forceinline__ void fec_dematch_cta(const fec_cb_desc& D,
                                                const int8_t* __restrict__ in,
                                                int8_t* __restrict__ harq,
                                                const SoftStore& soft_store,
                                                uint32_t* n_sig_smem)
{
  const uint32_t Z       = D.lifting_size;
  const uint32_t Kb      = (D.base_graph == 1) ? 22u : 10u;
  const uint32_t N       = ((D.base_graph == 1) ? 66u : 50u) * Z;
  const uint32_t Ncb     = (D.Nref > 0) ? min(D.Nref, N) : N;
  const uint32_t fillers = D.nof_filler_bits;
  const uint32_t sys     = (Kb - 2) * Z;
  const uint32_t info    = sys - fillers;
  const uint32_t E       = D.E;
  const uint32_t k0      = D.k0;
  const uint32_t qm      = D.qm;
  const uint32_t Kp      = E / qm;
  const bool     newd    = (D.flags & FEC_FLAG_NEW_DATA) != 0;
  const bool     hzero   = (D.flags & FEC_FLAG_HARQ_ZERO) != 0;
  const uint32_t Lc      = Ncb - fillers;                                    // non-filler positions of the circular buffer
  const uint32_t k0c     = (k0 < info) ? k0 : (max(k0, sys) - fillers);      // start, in filler-free coordinates
  const uint32_t zero_lo = min(k0, info);                                    // [0, zero_lo) is zeroed in copy mode
  const uint32_t e_first = Lc - k0c;                                         // LLRs consumed before the first wrap
  const bool     wrapped = E > e_first;
  uint32_t       zero_from = N;                                              // [zero_from, N) zeroed at the end
  if (newd && !wrapped) {
    const uint32_t q_end = k0c + E;
    uint32_t       idxf  = (k0c < info && q_end <= info) ? sys : (q_end + fillers);
    idxf                 = (idxf >= Ncb) ? idxf - Ncb : idxf;
    if (idxf != 0) {
      zero_from = N - (Ncb - idxf);
    }
  }
  uint32_t my_sig = 0;
  for (uint32_t o = threadIdx.x; o < N; o += blockDim.x) {
    int acc;
    const bool is_filler = (o >= info) && (o < sys);
    if (newd && o < zero_lo) {
      acc = 0;
    } else if (newd && is_filler) {
      acc = 127;
    } else {
      acc = hzero ? 0 : harq[o];
    }
    if (!is_filler && o < Ncb) {
      const uint32_t q     = (o < info) ? o : (o - fillers);
      const bool     first = (q >= k0c);
      uint32_t       p     = first ? (q - k0c) : (q + Lc - k0c);
      bool           copy  = newd && first;
      for (; p < E; p += Lc) {
        uint32_t src = p;
        if (qm > 1) {
          const uint32_t j = p / Kp;
          const uint32_t i = p - j * Kp;
          src              = i * qm + j;
        }
        const int dv = in[src];
        acc          = copy ? dv : fec_sat_add(acc, dv);
        copy         = false;
      }
    }
    if (o >= zero_from) {
      acc = 0;
    }
    harq[o] = (int8_t)acc;
    soft_store(o, acc);
    if (acc != 0) {
      my_sig = o + 1; // o increases monotonically per thread
    }
  }
  if (my_sig != 0) {
    atomicMax(n_sig_smem, my_sig);
  }
}

// Hard decision of the first K soft bits into bitsw (big-endian bit order). Returns true if no soft bit is zero.
__device__ __forceinline__ bool fec_hard_bits_cta(const int8_t* soft, uint32_t* bitsw, uint32_t K)
{
  const uint32_t lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nwarps = blockDim.x >> 5;
  const uint32_t nw   = (K + 31) / 32;
  int            zero = 0;
  for (uint32_t w = warp; w < nw; w += nwarps) {
    const uint32_t i = w * 32 + lane;
    int            s = (i < K) ? (int)soft[i] : 1;
    zero |= (s == 0);
    const uint32_t b = __ballot_sync(0xffffffffu, s <= 0);
    if (lane == 0) {
      bitsw[w] = __brev(b);
    }
  }
  return __syncthreads_or(zero) == 0; // also makes bitsw visible
}

// CRC remainder (== 0 test only) of the first T bits of bitsw. red: >= 33 words of scratch.
__device__ __forceinline__ bool fec_crc_ok_cta(const uint32_t* bitsw, uint32_t T, const fec_dev_tables* tab, uint32_t crc_type, uint32_t* red)
{
  const uint32_t poly  = tab->crc_poly[crc_type];
  const uint32_t top   = 1u << tab->crc_len[crc_type];
  const uint32_t nfull = T / 32, r = T % 32;
  uint32_t       acc   = 0;
  for (uint32_t w = threadIdx.x; w < nfull; w += blockDim.x) {
    acc ^= fec_crc_mulmod32(tab->crc_x32[crc_type][nfull - 1 - w], bitsw[w], poly, top);
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    acc ^= __shfl_xor_sync(0xffffffffu, acc, o);
  }
  if ((threadIdx.x & 31) == 0) {
    red[threadIdx.x >> 5] = acc;
  }
  __syncthreads();
  if (threadIdx.x == 0) {
    uint32_t rem = 0;
    for (uint32_t i = 0; i < (blockDim.x >> 5); ++i) {
      rem ^= red[i];
    }
    // rem = (message of the full words) mod g; continue bitwise over the r remaining bits.
    if (r != 0) {
      const uint32_t last = bitsw[nfull];
      for (uint32_t i = 0; i < r; ++i) {
        rem = (rem << 1) | ((last >> (31 - i)) & 1u);
        rem = (rem & top) ? (rem ^ poly) : rem;
      }
    }
    red[32] = rem;
  }
  __syncthreads();
  const bool ok = (red[32] == 0);
  __syncthreads();
  return ok;
}

// One codeblock by the whole CTA (any blockDim that is a multiple of 32): dematch / load, decode, CRC, outputs.
__device__ __noinline__ void fec_generic_cb(const fec_cb_desc& D,
                                            uint32_t           cb,
                                            uint8_t*           smem_raw,
                                            const int8_t* __restrict__ llr_base,
                                            uint8_t* __restrict__ out_base,
                                            fec_cb_result* __restrict__ results,
                                            int8_t* __restrict__ harq_base,
                                            const fec_dev_tables* __restrict__ tab)
{
  __shared__ uint32_t s_nsig;

  const uint32_t    Z   = D.lifting_size;
  const int         bg1 = (D.base_graph == 1);
  const uint32_t    Kb = bg1 ? 22u : 10u, nfull = bg1 ? 68u : 52u, M = bg1 ? 46u : 42u;
  const uint32_t    K = Kb * Z, N = (nfull - 2) * Z;
  const uint32_t    ne = bg1 ? 316u : 197u;

  const fec_generic_smem_layout L = fec_generic_layout(Z, bg1);
  uint32_t* stateA = reinterpret_cast<uint32_t*>(smem_raw + L.stateA);
  uint32_t* stateB = reinterpret_cast<uint32_t*>(smem_raw + L.stateB);
  uint32_t* bitsw  = reinterpret_cast<uint32_t*>(smem_raw + L.bitsw);
  uint32_t* red    = reinterpret_cast<uint32_t*>(smem_raw + L.red);
  uint16_t* sh     = reinterpret_cast<uint16_t*>(smem_raw + L.sh);
  uint16_t* rs     = reinterpret_cast<uint16_t*>(smem_raw + L.rs);
  uint8_t*  col    = smem_raw + L.col;
  uint8_t*  deg    = smem_raw + L.deg;
  uint8_t*  lut    = smem_raw + L.lut;
  int8_t*   soft   = reinterpret_cast<int8_t*>(smem_raw + L.soft);

  // Stage the graph of this codeblock and clear the check state.
  for (uint32_t i = threadIdx.x; i < ne; i += blockDim.x) {
    sh[i]  = tab->shift[D.graph_index][i];
    col[i] = tab->col[bg1 ? 0 : 1][i];
  }
  for (uint32_t i = threadIdx.x; i <= M; i += blockDim.x) {
    rs[i]  = tab->row_start[bg1 ? 0 : 1][i];
    deg[i] = tab->deg[bg1 ? 0 : 1][i];
  }
  for (uint32_t i = threadIdx.x; i < 128; i += blockDim.x) {
    lut[i] = tab->scale_lut[i];
  }
  for (uint32_t i = threadIdx.x; i < (L.bitsw - L.stateA) / 4; i += blockDim.x) {
    stateA[i] = 0; // stateA and stateB are contiguous
  }
  for (uint32_t i = threadIdx.x; i < 2 * Z; i += blockDim.x) {
    soft[i] = 0; // the two punctured columns
  }
  if (threadIdx.x == 0) {
    s_nsig = 0;
  }
  __syncthreads();

  const int8_t* in      = llr_base + D.llr_offset;
  fec_cb_result res;
  res.nof_ldpc_iterations = D.max_iter;
  res.status              = 0;
  res.nof_layers          = 0;
  res.reserved            = 0;

  // ---- stage 1: fill the soft bits ------------------------------------------------------------------------------------
  if (D.mode == FEC_MODE_DECODE) {
    const uint32_t n_in   = D.E;
    const uint32_t n_node = (n_in / Z) * Z;
    uint32_t       my_sig = 0;
    for (uint32_t o = threadIdx.x; o < N; o += blockDim.x) {
      int v = 0;
      if (o < n_in) {
        v = in[o];
        if (v != 0) {
          my_sig = o + 1;
        }
        if (o < n_node) {
          v = fec_clampi(v, -64, 64);
        }
      }
      soft[2 * Z + o] = (int8_t)v;
    }
    if (my_sig != 0) {
      atomicMax(&s_nsig, my_sig);
    }
  } else {
    int8_t* harq = harq_base + (size_t)D.harq_slot * FEC_HARQ_SLOT_BYTES;
    fec_dematch_cta(D, in, harq, fec_soft_store_natural{(D.mode == FEC_MODE_CHAIN) ? soft : nullptr, Z}, &s_nsig);
    if (D.mode == FEC_MODE_DEMATCH) {
      if (threadIdx.x == 0) {
        res.status           = FEC_RES_DONE;
        results[cb] = res;
      }
      return;
    }
  }
  __syncthreads();

  // ---- stage 2: layered scaled min-sum ---------------------------------------------------------------------------------
  const uint32_t n_sig      = s_nsig;
  const bool     early_stop = (D.flags & FEC_FLAG_EARLY_STOP) != 0;
  uint8_t*       out        = out_base + (((size_t)D.reserved1 << 32) | D.out_offset); // reserved1: bits 32.. of the offset
  const uint32_t kbytes     = (K + 7) / 8;
  if (n_sig < K) {
    if (!early_stop) {
      for (uint32_t i = threadIdx.x; i < kbytes; i += blockDim.x) {
        const uint32_t nb = min(8u, K - 8 * i);
        out[i]            = (uint8_t)(0xff00u >> nb);
      }
    }
    if (threadIdx.x == 0) {
      res.status           = FEC_RES_TOO_SHORT | FEC_RES_DONE;
      results[cb] = res;
    }
    return;
  }
  uint32_t cbl = max(n_sig + 2 * Z, K + 4 * Z);
  cbl          = ((cbl + Z - 1) / Z) * Z;
  const uint32_t nlayers = cbl / Z - Kb;
  const uint32_t T       = K - D.nof_filler_bits;

  bool     pass = false;
  uint32_t it   = 0;
  for (; it < D.max_iter; ++it) {
    for (uint32_t l = 0; l < nlayers; ++l) {
      const uint32_t e0 = rs[l], d = deg[l];
      for (uint32_t c = threadIdx.x; c < Z; c += blockDim.x) {
        if (d > 11) {
          fec_generic_row<19>(soft, stateA, stateB, sh, col, lut, l, e0, d, c, Z);
        } else {
          fec_generic_row<11>(soft, stateA, stateB, sh, col, lut, l, e0, d, c, Z);
        }
      }
      __syncthreads();
    }
    if (early_stop) {
      const bool clean = fec_hard_bits_cta(soft, bitsw, K);
      const bool crc   = fec_crc_ok_cta(bitsw, T, tab, D.crc_type, red);
      if (clean && crc) {
        pass = true;
        ++it;
        break;
      }
    }
  }
  if (!early_stop) {
    fec_hard_bits_cta(soft, bitsw, K);
    pass = fec_crc_ok_cta(bitsw, T, tab, D.crc_type, red) && !(D.flags & FEC_FLAG_NO_CRC);
  }
  for (uint32_t i = threadIdx.x; i < kbytes; i += blockDim.x) {
    out[i] = (uint8_t)(bitsw[i >> 2] >> (24 - 8 * (i & 3)));
  }
  if (threadIdx.x == 0) {
    res.nof_ldpc_iterations = (uint16_t)(pass && early_stop ? it : D.max_iter);
    res.status              = (pass ? FEC_RES_CRC_PASS : 0) | FEC_RES_DONE;
    res.nof_layers          = (uint8_t)nlayers;
    results[cb]    = res;
  }
}

// Lifting sizes the packed kernel (fec_kernel_packed.cuh) decodes: Z must split into two interleaved lanes of >= 2 checks.
// Every other codeblock (odd Z, all <= 15, and Z = 2) is left to this kernel.
__host__ __device__ inline bool fec_packed_covers(uint32_t Z)
{
  return (Z % 2 == 0) && (Z >= 4);
}

// leftovers_only: the launch that accompanies the packed kernel. A CTA whose codeblock the packed kernel covers returns at
// once (the two kernels run on the same descriptor array and share the work by lifting size).
__global__ void __launch_bounds__(384, 2)
fec_cb_generic_kernel(const fec_cb_desc* __restrict__ descs,
                      const int8_t* __restrict__ llr_base,
                      uint8_t* __restrict__ out_base,
                      fec_cb_result* __restrict__ results,
                      int8_t* __restrict__ harq_base,
                      const fec_dev_tables* __restrict__ tab,
                      int leftovers_only)
{
  extern __shared__ __align__(16) uint8_t smem_raw[];
  const fec_cb_desc D = descs[blockIdx.x];
  if (leftovers_only && fec_packed_covers(D.lifting_size)) {
    return;
  }
  fec_generic_cb(D, blockIdx.x, smem_raw, llr_base, out_base, results, harq_base, tab);
}
