// srsran-5g_b200/csrc/fec_kernel_packed.cuh -- the hot kernel: one CTA per codeblock, one thread per FOUR lifted check
// nodes, 16-bit SIMD-in-register arithmetic on the native sm_100a half2 pipes.
//
// Number format. Every LLR-domain quantity x (|x| < 2048) lives in a 16-bit lane as sign-magnitude: bit 15 = sign, bits
// 14..0 = |x|. That bit pattern IS the IEEE fp16 number x * 2^-24 (subnormal for |x| < 1024, normal above, no gap), so
// HADD2 / HFMA2 are exact sign-magnitude adders, HMNMX2.XORSIGN is the min-sum check-node primitive
// (min(|a|,|b|) with sign(a)^sign(b)), HSET2.BF gives 1.0/0.0 select masks that feed HFMA2, negation is an XOR and |x| an
// AND. No tensor cores, no floating-point rounding anywhere (every intermediate is an integer below 2048).
//
// Shared-memory layout per codeblock (Z = 4W):
//   soft   nfull*Z bytes. Column b occupies W words; word w holds variables {w, w+W, w+2W, w+3W} of the column in bytes
//          0..3, stored offset-binary (value + 128). Magnitudes 121..127 all mean the reference's +-127 "infinity".
//          With this interleave the cyclic shift s = q*W + r of an edge becomes ONE aligned word load at (t + r) mod W
//          plus a byte rotation by q (+1 on wrap), and the write-back is one aligned word store.
//   state  per base row and thread word: m1 (4 bytes: scaled min1 of 4 checks), m2, idx, and 1-3 words of c2v sign bits
//          (byte k = bits of check lane k for 8 edges). All-zero = "row not initialised" (c2v = 0).
//   edges  per edge {byte address of (col*W + r), W - r, q}: uniform per CTA, read by broadcast.
//
// Bit-exactness with the reference's generic decoder is argued step by step in DESIGN.md ("packed kernel algebra").
#pragma once
#include "fec_device.cuh"
#include "fec_kernel_generic.cuh"
#include <cuda_fp16.h>
#include <string.h>

struct fec_packed_edge {
  uint32_t a;   // (col * W + r) * 4 : byte offset of the word thread 0 reads (before wrap)
  uint32_t thr; // W - r : threads t >= thr wrap around the column
  uint32_t q;   // byte rotation without wrap
  uint32_t pad;
};

// Device-global tables for the packed kernel (built once per engine).
struct fec_packed_tables {
  fec_packed_edge edges[FEC_NOF_GRAPHS][FEC_MAX_EDGES]; // valid for graphs with Z % 4 == 0
  uint16_t        state_off[2][FEC_MAX_ROWS + 2];        // first state word of row r, in units of W words
  uint32_t        scale_mul;                             // scaled = (x * scale_mul + scale_add) >> 8 for x in 0..120
  uint32_t        scale_add;
};

struct fec_packed_smem_layout {
  uint32_t state, edges, soff, rs, deg, bitsw, red, soft, total;
};

__host__ __device__ inline uint32_t fec_packed_state_words(int bg1)
{
  // BG1: 4 rows of degree 19 (6 words), 3 rows of degree 9-10 (5 words), 39 rows of degree <= 8 (4 words).
  // BG2: 2 rows of degree 10 (5 words), 40 rows of degree <= 8 (4 words).
  return bg1 ? (4u * 6u + 3u * 5u + 39u * 4u) : (2u * 5u + 40u * 4u);
}

__host__ __device__ inline fec_packed_smem_layout fec_packed_layout(uint32_t Z, int bg1)
{
  fec_packed_smem_layout L;
  const uint32_t         W = Z / 4, nfull = bg1 ? 68u : 52u, Kb = bg1 ? 22u : 10u, ne = bg1 ? 316u : 197u;
  uint32_t               o = 0;
  L.state                  = o;
  o += fec_packed_state_words(bg1) * W * 4;
  o       = (o + 15) & ~15u;
  L.edges = o;
  o += ne * 16;
  L.soff = o;
  o += 48 * 2;
  L.rs = o;
  o += 48 * 2;
  L.deg = o;
  o += 48;
  L.bitsw = o;
  o += ((Kb * Z + 31) / 32 + 1) * 4;
  L.red = o;
  o += 64 * 4;
  o      = (o + 15) & ~15u;
  L.soft = o;
  o += (nfull * Z + 15) & ~15u;
  L.total = o;
  return L;
}

// ---- 16-bit sign-magnitude SIMD primitives (all single SASS instructions) --------------------------------------------
#define FEC_H2(x) (((uint32_t)(x) & 0xffffu) * 0x10001u)

__device__ __forceinline__ uint32_t h2_add(uint32_t a, uint32_t b)
{
  uint32_t r;
  asm("add.rn.f16x2 %0, %1, %2;" : "=r"(r) : "r"(a), "r"(b));
  return r;
}
__device__ __forceinline__ uint32_t h2_fma(uint32_t a, uint32_t b, uint32_t c)
{
  uint32_t r;
  asm("fma.rn.f16x2 %0, %1, %2, %3;" : "=r"(r) : "r"(a), "r"(b), "r"(c));
  return r;
}
// min(|a|, |b|) with sign(a) ^ sign(b)
__device__ __forceinline__ uint32_t h2_min_xs(uint32_t a, uint32_t b)
{
  uint32_t r;
  asm("min.xorsign.abs.f16x2 %0, %1, %2;" : "=r"(r) : "r"(a), "r"(b));
  return r;
}
__device__ __forceinline__ uint32_t h2_max_xs(uint32_t a, uint32_t b)
{
  uint32_t r;
  asm("max.xorsign.abs.f16x2 %0, %1, %2;" : "=r"(r) : "r"(a), "r"(b));
  return r;
}
// per lane: 1.0 (0x3c00) if a == b else 0.0
__device__ __forceinline__ uint32_t h2_eq_bf(uint32_t a, uint32_t b)
{
  uint32_t r;
  asm("set.eq.f16x2.f16x2 %0, %1, %2;" : "=r"(r) : "r"(a), "r"(b));
  return r;
}
// Generic-mode PRMT with the full selector (the __byte_perm intrinsic strips the sign-replicate bit of each nibble).
__device__ __forceinline__ uint32_t prmt(uint32_t a, uint32_t b, uint32_t c)
{
  uint32_t r;
  asm("prmt.b32 %0, %1, %2, %3;" : "=r"(r) : "r"(a), "r"(b), "r"(c));
  return r;
}
__device__ __forceinline__ uint32_t prmt_f4e(uint32_t a, uint32_t b, uint32_t c)
{
  uint32_t r;
  asm("prmt.b32.f4e %0, %1, %2, %3;" : "=r"(r) : "r"(a), "r"(b), "r"(c));
  return r;
}

// Constants of the number format.
#define H2_SIGN 0x80008000u
#define H2_ABS 0x7fff7fffu
#define H2_P120 FEC_H2(120)
#define H2_P127 FEC_H2(127)
#define H2_P128 FEC_H2(128)
#define H2_M128 FEC_H2(0x8000 | 128)
#define H2_NEG_ONE 0xbc00bc00u  // -1.0 | -1.0
#define H2_F4352 0x6c406c40u    // 4352.0 : |S| * 4352 * 2^-24 crosses 1/32 between |S| = 120 and 121
#define H2_F96 0x56005600u      // 96.0   : ulp 1/16 in [64, 128)
#define H2_FM96 0xd600d600u     // -96.0
#define H2_F2M11 0x10001000u    // 2^-11  : (+-1/16) * 2^-11 = +-512 LLR units

// ---- one base row for the 4 checks {t, t+W, t+2W, t+3W} of this thread -------------------------------------------------
// D = row degree (compile time so that the per-edge values stay in registers). Lanes: half A = checks (t, t+W), half B =
// checks (t+2W, t+3W).
template <int D>
__device__ __forceinline__ void fec_packed_row(uint8_t*               soft,
                                               uint32_t*              st,   // state words of this row for thread t
                                               const fec_packed_edge* ed,   // D edge records (shared memory)
                                               uint32_t               W,
                                               uint32_t               t,
                                               uint32_t               scale_mul,
                                               uint32_t               scale_add)
{
  constexpr int NS = (D + 7) / 8; // sign words
  const uint32_t t4 = t * 4, W4 = W * 4;

  // Issue the soft-word loads of every edge first (independent, one aligned LDS.32 each).
  uint32_t addr[D], rho[D], raw[D];
#pragma unroll
  for (int e = 0; e < D; ++e) {
    const uint4    E    = *reinterpret_cast<const uint4*>(&ed[e]);
    const bool     wrap = t >= E.y;
    addr[e]             = E.x + t4 - (wrap ? W4 : 0u);
    rho[e]              = E.z + (wrap ? 1u : 0u);
    raw[e]              = *reinterpret_cast<const uint32_t*>(soft + addr[e]);
  }

  // Previous check-to-variable state of this row.
  const uint32_t m1w = st[0], m2w = st[W], idw = st[2 * W];
  uint32_t       sgw[NS];
#pragma unroll
  for (int k = 0; k < NS; ++k) {
    sgw[k] = st[(3 + k) * W];
  }
  const uint32_t om1[2] = {__byte_perm(m1w, 0, 0x4140), __byte_perm(m1w, 0, 0x4342)};
  const uint32_t om2[2] = {__byte_perm(m2w, 0, 0x4140), __byte_perm(m2w, 0, 0x4342)};
  const uint32_t oid[2] = {__byte_perm(idw, 0, 0x4140), __byte_perm(idw, 0, 0x4342)};
  uint32_t       odl[2];
#pragma unroll
  for (int h = 0; h < 2; ++h) {
    odl[h] = h2_add(om2[h], om1[h] ^ H2_SIGN); // m2 - m1 >= 0
  }

  // ---- phase A: variable-to-check messages and the min1 / min2 / sign-product analysis -----------------------------------
  uint32_t v[D][2];
  uint32_t m1[2] = {H2_P120, H2_P120}, m2[2] = {H2_P120, H2_P120};
#pragma unroll
  for (int e = 0; e < D; ++e) {
    const uint32_t wr = prmt_f4e(raw[e], raw[e], rho[e]); // lane k <- byte (k + rho) mod 4
    // sign bits of this edge for lanes (0,1) / (2,3): byte k of sgw holds lane k, bit (e & 7)
    const uint32_t sw    = sgw[e >> 3];
    const uint32_t sg[2] = {__byte_perm(sw, 0, 0x1404) << (7 - (e & 7)), __byte_perm(sw, 0, 0x3424) << (7 - (e & 7))};
    const uint32_t up[2] = {__byte_perm(wr, 0, 0x4140), __byte_perm(wr, 0, 0x4342)};
#pragma unroll
    for (int h = 0; h < 2; ++h) {
      const uint32_t S    = h2_add(up[h], H2_M128);                                  // soft value
      const uint32_t ism  = h2_eq_bf(oid[h], FEC_H2(e));                             // edge e was the min1 edge
      const uint32_t mag  = h2_fma(ism, odl[h], om1[h]);                             // |old c2v|
      const uint32_t nsgn = (sg[h] & H2_SIGN) ^ H2_NEG_ONE;                          // -sign(old c2v) as +-1.0
      const uint32_t vraw = h2_fma(mag, nsgn, S);                                    // soft - c2v, |.| <= 223
      const uint32_t y    = h2_add(h2_fma(S, H2_F4352, H2_F96), H2_FM96);            // +-1/16 iff |S| > 120
      const uint32_t vcl  = h2_min_xs(vraw, H2_P120);                                // saturate at +-120
      const uint32_t vv   = h2_fma(y, H2_F2M11, vcl);                                // "infinite" soft bits stay huge
      v[e][h]             = vv;
      const uint32_t mx   = h2_max_xs(m1[h], vv);
      m2[h]               = h2_min_xs(m2[h], mx);
      m1[h]               = h2_min_xs(m1[h], vv);
    }
  }

  // ---- scale the two minima: scaled = (x * mul + add) >> 8, two lanes per 32-bit multiply ---------------------------------
  uint32_t M1s[2], dls[2], a1[2], s1[2], s2[2];
#pragma unroll
  for (int h = 0; h < 2; ++h) {
    const uint32_t sp = m1[h] & H2_SIGN; // product of all v2c signs
    a1[h]             = m1[h] & H2_ABS;
    s1[h]             = (((a1[h]) * scale_mul + scale_add) >> 8) & 0x00ff00ffu;
    s2[h]             = (((m2[h] & H2_ABS) * scale_mul + scale_add) >> 8) & 0x00ff00ffu;
    M1s[h]            = s1[h] | sp;
    dls[h]            = h2_add(s2[h], s1[h] ^ H2_SIGN) | sp;
  }

  // ---- phase B: new check-to-variable messages, soft-bit update, write-back ------------------------------------------------
  uint32_t nid[2] = {0, 0};
  uint32_t nsg[NS];
#pragma unroll
  for (int k = 0; k < NS; ++k) {
    nsg[k] = 0;
  }
#pragma unroll
  for (int e = 0; e < D; ++e) {
    uint32_t o[2];
#pragma unroll
    for (int h = 0; h < 2; ++h) {
      const uint32_t vv  = v[e][h];
      const uint32_t eq  = h2_eq_bf(vv & H2_ABS, a1[h]);       // this edge holds min1 -> it gets min2
      const uint32_t mg  = h2_fma(eq, dls[h], M1s[h]);         // sign product * scaled magnitude
      const uint32_t c   = mg ^ (vv & H2_SIGN);                // new c2v
      nid[h]             = h2_fma(eq, FEC_H2(e), nid[h]);      // index of the min1 edge (any value if min1 == min2)
      const uint32_t tt  = h2_add(c, vv);                      // c2v + v2c
      const uint32_t nw  = h2_min_xs(tt, H2_P127);             // |.| > 120 is "infinity" from here on
      o[h]               = h2_add(nw, H2_P128);                // offset binary in the low byte of each lane
    }
    // sign(v2c) of the 4 lanes as byte masks -> bit (e & 7) of the sign word
    const uint32_t sm = prmt(v[e][0], v[e][1], 0xfdb9);
    nsg[e >> 3] |= sm & (0x01010101u << (e & 7));
    const uint32_t pk = __byte_perm(o[0], o[1], 0x6420);
    *reinterpret_cast<uint32_t*>(soft + addr[e]) = prmt_f4e(pk, pk, 0u - rho[e]); // byte j <- lane (j - rho) mod 4
  }

  // ---- new state: c2v sign = sign product ^ sign(v2c) ----------------------------------------------------------------------
  const uint32_t pm = prmt(m1[0], m1[1], 0xfdb9); // 0xff in the lanes whose sign product is negative
  st[0]             = __byte_perm(s1[0], s1[1], 0x6420);
  st[W]             = __byte_perm(s2[0], s2[1], 0x6420);
  st[2 * W]         = __byte_perm(nid[0], nid[1], 0x6420);
#pragma unroll
  for (int k = 0; k < NS; ++k) {
    st[(3 + k) * W] = nsg[k] ^ pm;
  }
}

// Byte address of variable j (0 <= j < Z) of column col in the interleaved soft layout.
__device__ __forceinline__ uint32_t fec_packed_soft_addr(uint32_t col, uint32_t j, uint32_t Z, uint32_t W, uint32_t invW)
{
  const uint32_t lane = __umulhi(j, invW); // j / W, exact for j < 2^32 / W
  return col * Z + (j - lane * W) * 4 + lane;
}

// Dematcher / loader target: clamp to +-64 (ldpc_decoder_impl.cpp:168-169) and store offset-binary in the packed layout.
struct fec_soft_store_packed {
  uint8_t* soft;
  uint32_t Z, W, invZ, invW;
  __device__ __forceinline__ void operator()(uint32_t o, int acc) const
  {
    const uint32_t op  = o + 2 * Z;
    const uint32_t col = __umulhi(op, invZ);
    soft[fec_packed_soft_addr(col, op - col * Z, Z, W, invW)] = (uint8_t)(fec_clampi(acc, -64, 64) + 128);
  }
};

// Hard decision of the first K soft bits (bit = soft <= 0) into bitsw, big-endian bit order; true if no soft bit is zero.
__device__ __forceinline__ bool fec_hard_bits_packed(const uint8_t* soft, uint32_t* bitsw, uint32_t K, uint32_t Z, uint32_t W,
                                                     uint32_t invZ, uint32_t invW)
{
  const uint32_t lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nwarps = blockDim.x >> 5;
  const uint32_t nw   = (K + 31) / 32;
  int            zero = 0;
  for (uint32_t w = warp; w < nw; w += nwarps) {
    const uint32_t i = w * 32 + lane;
    uint32_t       b = 129; // positive
    if (i < K) {
      const uint32_t col = __umulhi(i, invZ);
      b                  = soft[fec_packed_soft_addr(col, i - col * Z, Z, W, invW)];
    }
    zero |= (b == 128);
    const uint32_t bal = __ballot_sync(0xffffffffu, b <= 128);
    if (lane == 0) {
      bitsw[w] = __brev(bal);
    }
  }
  return __syncthreads_or(zero) == 0;
}

__device__ __forceinline__ bool fec_packed_cb_supported(const fec_cb_desc& D)
{
  // Z must split into 4 interleaved lanes; the unclamped-tail corner (input not a multiple of Z, never produced by
  // PUSCH) stays on the generic path.
  return (D.lifting_size % 4 == 0) && (D.lifting_size >= 8) && (D.lifting_size / 4 <= blockDim.x) &&
         !(D.mode == FEC_MODE_DECODE && (D.E % D.lifting_size) != 0) && D.mode != FEC_MODE_DEMATCH;
}

__device__ __noinline__ void fec_packed_cb(const fec_cb_desc& D,
                                           uint32_t           cb,
                                           uint8_t*           smem_raw,
                                           const int8_t* __restrict__ llr_base,
                                           uint8_t* __restrict__ out_base,
                                           fec_cb_result* __restrict__ results,
                                           int8_t* __restrict__ harq_base,
                                           const fec_dev_tables* __restrict__ tab,
                                           const fec_packed_tables* __restrict__ ptab)
{
  __shared__ uint32_t s_nsig;

  const uint32_t Z = D.lifting_size, W = Z / 4;
  const int      bg1 = (D.base_graph == 1);
  const uint32_t Kb = bg1 ? 22u : 10u, nfull = bg1 ? 68u : 52u, M = bg1 ? 46u : 42u;
  const uint32_t K = Kb * Z, N = (nfull - 2) * Z;
  const uint32_t ne = bg1 ? 316u : 197u;
  const uint32_t invZ = 0xffffffffu / Z + 1, invW = 0xffffffffu / W + 1;
  const uint32_t tid = threadIdx.x;

  const fec_packed_smem_layout L = fec_packed_layout(Z, bg1);
  uint32_t*        state = reinterpret_cast<uint32_t*>(smem_raw + L.state);
  fec_packed_edge* edges = reinterpret_cast<fec_packed_edge*>(smem_raw + L.edges);
  uint16_t*        soff  = reinterpret_cast<uint16_t*>(smem_raw + L.soff);
  uint16_t*        rs    = reinterpret_cast<uint16_t*>(smem_raw + L.rs);
  uint8_t*         deg   = smem_raw + L.deg;
  uint32_t*        bitsw = reinterpret_cast<uint32_t*>(smem_raw + L.bitsw);
  uint32_t*        red   = reinterpret_cast<uint32_t*>(smem_raw + L.red);
  uint8_t*         soft  = smem_raw + L.soft;

  // Stage the graph of this codeblock and clear the check state.
  {
    const uint4* src = reinterpret_cast<const uint4*>(ptab->edges[D.graph_index]);
    uint4*       dst = reinterpret_cast<uint4*>(edges);
    for (uint32_t i = tid; i < ne; i += blockDim.x) {
      dst[i] = src[i];
    }
    for (uint32_t i = tid; i <= M; i += blockDim.x) {
      rs[i]   = tab->row_start[bg1 ? 0 : 1][i];
      deg[i]  = tab->deg[bg1 ? 0 : 1][i];
      soff[i] = ptab->state_off[bg1 ? 0 : 1][i];
    }
    uint4*         s4 = reinterpret_cast<uint4*>(state);
    const uint32_t n4 = (L.edges - L.state) / 16;
    for (uint32_t i = tid; i < n4; i += blockDim.x) {
      s4[i] = make_uint4(0, 0, 0, 0);
    }
    uint32_t* p2 = reinterpret_cast<uint32_t*>(soft);
    for (uint32_t i = tid; i < 2 * W; i += blockDim.x) {
      p2[i] = 0x80808080u; // the two punctured columns: soft = 0
    }
    if (tid == 0) {
      s_nsig = 0;
    }
  }
  __syncthreads();

  const int8_t* in = llr_base + D.llr_offset;
  fec_cb_result res;
  res.nof_ldpc_iterations = D.max_iter;
  res.status              = 0;
  res.nof_layers          = 0;
  res.reserved            = 0;

  // ---- stage 1: soft bits into shared memory -----------------------------------------------------------------------------
  const fec_soft_store_packed store{soft, Z, W, invZ, invW};
  if (D.mode == FEC_MODE_DECODE) {
    const uint32_t n_in   = D.E; // a multiple of Z on this path
    uint32_t       my_sig = 0;
    for (uint32_t o = tid; o < N; o += blockDim.x) {
      int vv = 0;
      if (o < n_in) {
        vv = in[o];
        if (vv != 0) {
          my_sig = o + 1;
        }
      }
      store(o, vv);
    }
    if (my_sig != 0) {
      atomicMax(&s_nsig, my_sig);
    }
  } else {
    int8_t* harq = harq_base + (size_t)D.harq_slot * FEC_HARQ_SLOT_BYTES;
    fec_dematch_cta(D, in, harq, store, &s_nsig);
  }
  __syncthreads();

  // ---- stage 2: layered scaled min-sum -------------------------------------------------------------------------------------
  const uint32_t n_sig      = s_nsig;
  const bool     early_stop = (D.flags & FEC_FLAG_EARLY_STOP) != 0;
  uint8_t*       out        = out_base + D.out_offset;
  const uint32_t kbytes     = (K + 7) / 8;
  if (n_sig < K) {
    if (!early_stop) {
      for (uint32_t i = tid; i < kbytes; i += blockDim.x) {
        const uint32_t nb = min(8u, K - 8 * i);
        out[i]            = (uint8_t)(0xff00u >> nb);
      }
    }
    if (tid == 0) {
      res.status   = FEC_RES_TOO_SHORT | FEC_RES_DONE;
      results[cb] = res;
    }
    return;
  }
  uint32_t cbl = max(n_sig + 2 * Z, K + 4 * Z);
  cbl          = ((cbl + Z - 1) / Z) * Z;
  const uint32_t nlayers = cbl / Z - Kb;
  const uint32_t T       = K - D.nof_filler_bits;
  const uint32_t smul = ptab->scale_mul, sadd = ptab->scale_add;

  bool     pass = false;
  uint32_t it   = 0;
  for (; it < D.max_iter; ++it) {
    for (uint32_t l = 0; l < nlayers; ++l) {
      if (tid < W) {
        uint32_t*              st = state + (uint32_t)soff[l] * W + tid;
        const fec_packed_edge* ed = edges + rs[l];
        switch (deg[l]) {
          case 3: fec_packed_row<3>(soft, st, ed, W, tid, smul, sadd); break;
          case 4: fec_packed_row<4>(soft, st, ed, W, tid, smul, sadd); break;
          case 5: fec_packed_row<5>(soft, st, ed, W, tid, smul, sadd); break;
          case 6: fec_packed_row<6>(soft, st, ed, W, tid, smul, sadd); break;
          case 7: fec_packed_row<7>(soft, st, ed, W, tid, smul, sadd); break;
          case 8: fec_packed_row<8>(soft, st, ed, W, tid, smul, sadd); break;
          case 9: fec_packed_row<9>(soft, st, ed, W, tid, smul, sadd); break;
          case 10: fec_packed_row<10>(soft, st, ed, W, tid, smul, sadd); break;
          default: fec_packed_row<19>(soft, st, ed, W, tid, smul, sadd); break;
        }
      }
      __syncthreads();
    }
    if (early_stop) {
      const bool clean = fec_hard_bits_packed(soft, bitsw, K, Z, W, invZ, invW);
      const bool crc   = fec_crc_ok_cta(bitsw, T, tab, D.crc_type, red);
      if (clean && crc) {
        pass = true;
        ++it;
        break;
      }
    }
  }
  if (!early_stop) {
    fec_hard_bits_packed(soft, bitsw, K, Z, W, invZ, invW);
    pass = fec_crc_ok_cta(bitsw, T, tab, D.crc_type, red) && !(D.flags & FEC_FLAG_NO_CRC);
  }
  for (uint32_t i = tid; i < kbytes; i += blockDim.x) {
    out[i] = (uint8_t)(bitsw[i >> 2] >> (24 - 8 * (i & 3)));
  }
  if (tid == 0) {
    res.nof_ldpc_iterations = (uint16_t)(pass && early_stop ? it : D.max_iter);
    res.status              = (pass ? FEC_RES_CRC_PASS : 0) | FEC_RES_DONE;
    res.nof_layers          = (uint8_t)nlayers;
    results[cb]            = res;
  }
}

// Dispatcher: every codeblock the packed path covers goes there, the rest (Z not a multiple of 4, dematch-only
// operations, unaligned decoder inputs) runs the generic per-check code in the same launch.
__global__ void __launch_bounds__(96, 2)
fec_cb_packed_kernel(const fec_cb_desc* __restrict__ descs,
                     const int8_t* __restrict__ llr_base,
                     uint8_t* __restrict__ out_base,
                     fec_cb_result* __restrict__ results,
                     int8_t* __restrict__ harq_base,
                     const fec_dev_tables* __restrict__ tab,
                     const fec_packed_tables* __restrict__ ptab)
{
  extern __shared__ __align__(16) uint8_t smem_raw[];
  const fec_cb_desc D = descs[blockIdx.x];
  if (fec_packed_cb_supported(D)) {
    fec_packed_cb(D, blockIdx.x, smem_raw, llr_base, out_base, results, harq_base, tab, ptab);
  } else {
    fec_generic_cb(D, blockIdx.x, smem_raw, llr_base, out_base, results, harq_base, tab);
  }
}

// ---- host side ---------------------------------------------------------------------------------------------------------
// Fills the packed tables from the generic ones. Returns false if the scaling LUT cannot be written as
// (x * mul + add) >> 8 on 0..120 (then the engine keeps to the generic kernel).
static inline bool fec_packed_build_tables(fec_packed_tables& pt, const fec_dev_tables& t, const uint16_t* all_z)
{
  memset(&pt, 0, sizeof(pt));
  bool found = false;
  for (uint32_t mul = 1; mul < 256 && !found; ++mul) {
    for (uint32_t add = 0; add < 256 && !found; ++add) {
      bool ok = true;
      for (uint32_t x = 0; x <= 120 && ok; ++x) {
        ok = (((x * mul + add) >> 8) == t.scale_lut[x]);
      }
      if (ok) {
        pt.scale_mul = mul;
        pt.scale_add = add * 0x10001u;
        found        = true;
      }
    }
  }
  for (int bg = 0; bg < 2; ++bg) {
    const int M   = bg ? 42 : 46;
    uint32_t  off = 0;
    for (int m = 0; m < M; ++m) {
      pt.state_off[bg][m] = (uint16_t)off;
      off += 3 + (t.deg[bg][m] + 7) / 8;
    }
    pt.state_off[bg][M] = (uint16_t)off;
    if (off != fec_packed_state_words(bg == 0)) {
      return false;
    }
    const int ne = t.row_start[bg][M];
    for (int zi = 0; zi < 51; ++zi) {
      const uint32_t Z = all_z[zi];
      if (Z % 4 != 0 || Z < 8) {
        continue;
      }
      const uint32_t W = Z / 4;
      for (int i = 0; i < ne; ++i) {
        const uint32_t   s = t.shift[bg * 51 + zi][i], col = t.col[bg][i];
        fec_packed_edge& e = pt.edges[bg * 51 + zi][i];
        e.a                = (col * W + s % W) * 4;
        e.thr              = W - s % W;
        e.q                = s / W;
      }
    }
  }
  return found;
}

static inline int fec_packed_launch(cudaStream_t             st,
                                    const fec_cb_desc*       descs,
                                    uint32_t                 n,
                                    const int8_t*            llr,
                                    uint8_t*                 out,
                                    fec_cb_result*           res,
                                    int8_t*                  harq,
                                    const fec_dev_tables*    tab,
                                    const fec_packed_tables* ptab,
                                    uint32_t                 max_Z,
                                    int                      any_bg1)
{
  uint32_t threads = ((max_Z / 4 + 31) / 32) * 32;
  threads          = threads < 32 ? 32 : (threads > 96 ? 96 : threads);
  const uint32_t a = fec_generic_layout(max_Z, any_bg1).total;
  const uint32_t b = (max_Z % 4 == 0 && max_Z >= 8) ? fec_packed_layout(max_Z, any_bg1).total : 0;
  const uint32_t smem = a > b ? a : b;
  if (cudaFuncSetAttribute(fec_cb_packed_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem) != cudaSuccess) {
    return FEC_ERR_CUDA;
  }
  fec_cb_packed_kernel<<<n, threads, smem, st>>>(descs, llr, out, res, harq, tab, ptab);
  return cudaGetLastError() == cudaSuccess ? FEC_OK : FEC_ERR_CUDA;
}
