// srsran-5g_b200/csrc/fec_kernel_packed.cuh -- the hot kernel: one CTA per codeblock, one thread per TWO lifted check
// nodes, 16-bit SIMD-in-register arithmetic on the native sm_100a half2 instructions.
//
// Number format. Every LLR-domain quantity x (|x| < 2048) lives in a 16-bit lane as sign-magnitude: bit 15 = sign, bits
// 14..0 = |x|. That bit pattern IS the IEEE fp16 number x * 2^-24 (subnormal for |x| < 1024, normal above, no gap), so
// HADD2 / HFMA2 are exact sign-magnitude adders, HMNMX2.XORSIGN is the min-sum check-node primitive
// (min(|a|,|b|) with sign(a)^sign(b)), HSET2.BF gives 1.0/0.0 select masks that feed HFMA2, negation is an XOR and |x| an
// AND. No tensor cores, no floating-point rounding anywhere (every intermediate is an integer below 2048).
//
// Shared-memory layout per codeblock (Z = 2H, H "halfwords" per column):
//   soft   nfull*Z bytes. Column b occupies H halfwords; halfword w holds variables {w, w+H} of the column in bytes 0/1,
//          stored offset-binary (value + 128). Magnitudes 121..127 all mean the reference's +-127 "infinity".
//          With this interleave the cyclic shift s = q*H + r of an edge becomes ONE aligned 16-bit load at (t + r) mod H
//          plus a byte swap when q + wrap is odd, and the write-back is one aligned 16-bit store.
//   state  per base row and thread: word0 = scaled min1 / min2 of both checks; then index of the min1 edge and the c2v
//          sign bits (1 more word for rows of degree <= 8, 2 otherwise). All-zero = "row not initialised" (c2v = 0).
//   edges  per edge {byte address of (col*H + r), H - r, unpack/pack selectors}: uniform per CTA, read by broadcast.
//
// Bit-exactness with the reference's generic decoder is argued step by step in DESIGN.md ("packed kernel algebra").
#pragma once
#include "fec_device.cuh"
#include "fec_kernel_generic.cuh"
#include <cuda_fp16.h>
#include <string.h>

#define FEC_PACKED_MAX_THREADS 192

struct fec_packed_edge {
  uint32_t aN;  // byte offset of lane 0 for thread 0 when it does not wrap: (col * H + r) * 2 + (q & 1)
  uint32_t aW;  // same for threads that wrap around the column: (col * H + r - H) * 2 + ((q + 1) & 1)
  uint32_t thr; // H - r : threads t >= thr wrap
  uint32_t pad; // lane 1 always sits at the lane-0 address ^ 1
};

// Device-global tables for the packed kernel (built once per engine).
struct fec_packed_tables {
  fec_packed_edge edges[FEC_NOF_GRAPHS][FEC_MAX_EDGES]; // valid for graphs with even Z >= 4
  uint16_t        state_off[2][FEC_MAX_ROWS + 2];        // first state word of row r, in units of H words
  uint32_t        scale_mul;                             // scaled = (x * scale_mul + scale_add) >> 8 for x in 0..120
  uint32_t        scale_add;
};

struct fec_packed_smem_layout {
  uint32_t state, edges, soff, rs, deg, bitsw, red, x8, soft, total;
};

__host__ __device__ inline uint32_t fec_packed_row_words(uint32_t d)
{
  return d <= 8 ? 2u : 3u;
}

// Where the check state lives. The LAST FEC_PACKED_TMEM_ROWS base rows of a graph (all of degree <= 8: two words per thread)
// keep their state in TENSOR MEMORY: thread t of warp w owns TMEM lane 32 (w % 4) + (t % 32), columns
// 64 (w / 4) + 2 (row - first_tmem_row) + {0, 1} of the CTA's allocation (tcgen05.alloc / ld / st as a private per-thread
// scratchpad; no MMA is ever issued). Every thread reads back only what it wrote itself. The first rows (the dense core of
// the graph) stay in shared memory. This takes 49 KB of check state out of shared memory at BG1 / Z = 384, which is what lets
// three codeblocks instead of two be resident per SM.
#define FEC_PACKED_TMEM_ROWS 32u
#define FEC_PACKED_TMEM_COLS_PER_WARPSLOT (2u * FEC_PACKED_TMEM_ROWS)

__host__ __device__ inline uint32_t fec_packed_first_tmem_row(int bg1)
{
  return (bg1 ? 46u : 42u) - FEC_PACKED_TMEM_ROWS;
}

// State words per thread kept in SHARED memory (rows before the first TMEM row).
__host__ __device__ inline uint32_t fec_packed_state_words(int bg1)
{
  // BG1 rows 0..13: degrees 19,19,19,19,3,8,9,7,10,9,7,8,7,6 -> 7 rows of 3 words, 7 rows of 2.
  // BG2 rows 0..9: degrees 8,10,8,10,4,6,6,6,4,5 -> 2 rows of 3 words, 8 rows of 2.
  return bg1 ? (7u * 3u + 7u * 2u) : (2u * 3u + 8u * 2u);
}

__host__ __device__ inline fec_packed_smem_layout fec_packed_layout(uint32_t Z, int bg1)
{
  fec_packed_smem_layout L;
  const uint32_t         H = Z / 2, nfull = bg1 ? 68u : 52u, Kb = bg1 ? 22u : 10u, ne = bg1 ? 316u : 197u;
  uint32_t               o = 0;
  L.state                  = o;
  o += fec_packed_state_words(bg1) * H * 4;
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
  L.x8 = o;
  o += (Kb * Z / 8 + 2) * 4;
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
__device__ __forceinline__ uint32_t h2_abs(uint32_t a)
{
  uint32_t r;
  asm("abs.f16x2 %0, %1;" : "=r"(r) : "r"(a));
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

// ---- tensor memory as per-thread scratch (32x32b shape: thread i of the warp <-> lane base + i, consecutive columns) ------
__device__ __forceinline__ void tmem_alloc(uint32_t smem_dst, uint32_t ncols)
{
  asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_dst), "r"(ncols) : "memory");
  asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
}
__device__ __forceinline__ void tmem_dealloc(uint32_t taddr, uint32_t ncols)
{
  asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(taddr), "r"(ncols) : "memory");
}
// Load + wait in ONE asm statement: the registers must not be read before tcgen05.wait::ld.
__device__ __forceinline__ void tmem_ld2(uint32_t taddr, uint32_t& a, uint32_t& b)
{
  asm volatile("tcgen05.ld.sync.aligned.32x32b.x2.b32 {%0, %1}, [%2];\n\ttcgen05.wait::ld.sync.aligned;"
               : "=r"(a), "=r"(b)
               : "r"(taddr)
               : "memory");
}
__device__ __forceinline__ void tmem_st2(uint32_t taddr, uint32_t a, uint32_t b)
{
  asm volatile("tcgen05.st.sync.aligned.32x32b.x2.b32 [%0], {%1, %2};" ::"r"(taddr), "r"(a), "r"(b) : "memory");
}
__device__ __forceinline__ void tmem_st8_zero(uint32_t taddr)
{
  const uint32_t z = 0;
  asm volatile("tcgen05.st.sync.aligned.32x32b.x8.b32 [%0], {%1, %1, %1, %1, %1, %1, %1, %1};" ::"r"(taddr), "r"(z) : "memory");
}
__device__ __forceinline__ void tmem_wait_st()
{
  asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory");
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


// A constant the compiler must keep in a register (LOP3 takes a single immediate).
__device__ __forceinline__ uint32_t fec_reg_const(uint32_t x)
{
  uint32_t r;
  asm("mov.b32 %0, %1;" : "=r"(r) : "r"(x));
  return r;
}

// Shared-memory byte accesses by 32-bit shared-window address (lets the two lanes of a halfword be addressed A and A^1).
// "memory" clobber: ordinary C++ stores to shared memory surround these accesses and must not be moved across them.
__device__ __forceinline__ uint32_t lds_u8(uint32_t a)
{
  uint32_t r;
  asm volatile("ld.shared.u8 %0, [%1];" : "=r"(r) : "r"(a) : "memory");
  return r;
}
__device__ __forceinline__ void sts_u8(uint32_t a, uint32_t v)
{
  asm volatile("st.shared.u8 [%0], %1;" ::"r"(a), "r"(v) : "memory");
}
// Integer helpers pinned to the FMA pipe (IMAD / IMAD.HI), which this kernel under-uses, instead of SHF / PRMT on the ALU.
__device__ __forceinline__ uint32_t imad(uint32_t a, uint32_t b, uint32_t c)
{
  uint32_t r;
  asm("mad.lo.u32 %0, %1, %2, %3;" : "=r"(r) : "r"(a), "r"(b), "r"(c));
  return r;
}
// Right shifts: IMAD.HI runs at a quarter of the issue rate on sm_100a (measured: 0.25 warp instructions per clock and
// scheduler; IMAD, HFMA2, LOP3, SHF, PRMT all 0.5), so shifts right stay on the ALU pipe.
#ifdef FEC_SHR_IMADHI
#define FEC_SHR(x, n) imul_hi((x), 1u << (32 - (n)))
#else
#define FEC_SHR(x, n) ((x) >> (n))
#endif
__device__ __forceinline__ uint32_t imul_hi(uint32_t a, uint32_t b)
{
  uint32_t r;
  asm("mul.hi.u32 %0, %1, %2;" : "=r"(r) : "r"(a), "r"(b));
  return r;
}

// ---- one base row for the 2 checks {t, t+H} of this thread -------------------------------------------------------------
// D = row degree (compile time so that the per-edge values stay in registers). Lane 0 = check t, lane 1 = check t+H.
// Instruction budget per edge and thread: ALU pipe 14 (ISETP, SEL, IADD3, LOP3 for the addresses; 2 HSET2,
// 5 HMNMX2, 3 LOP3), FMA pipe 15 (HFMA2 / HADD2 / IMAD), LSU 5.
template <int D, bool TM>
__device__ __forceinline__ void fec_packed_row(uint32_t               soft_s, // shared-window address of the soft array
                                               uint32_t*              st,     // state words of this row for thread t (!TM)
                                               uint32_t               taddr,  // tensor-memory address of the row's 2 words (TM)
                                               const fec_packed_edge* ed,     // D edge records (shared memory)
                                               uint32_t               H,
                                               uint32_t               t,
                                               uint32_t               scale_mul,
                                               uint32_t               scale_add)
{
  constexpr int  NW = (D <= 8) ? 2 : 3;
  static_assert(!TM || NW == 2, "rows kept in tensor memory have two state words");
  const uint32_t tb = soft_s + t * 2;

  // ---- pre-phase: everything that does not depend on the soft bits (addresses, the previous c2v messages of this row).
  // It runs BEFORE the CTA barrier that orders this row after the previous one, so warps that finished the previous row
  // early do useful work instead of waiting, and the table / state load latency overlaps the other warps' tail.
  uint32_t a0[D], cneg[D];
  {
    uint32_t w0, w1, w2 = 0;
    if (TM) {
      tmem_wait_st(); // this thread's store of the previous tensor-memory row
      tmem_ld2(taddr, w0, w1);
    } else {
      w0 = st[0];
      w1 = st[H];
      if (NW == 3) {
        w2 = st[2 * H];
      }
    }
#pragma unroll
    for (int e = 0; e < D; ++e) {
      const uint4 E = *reinterpret_cast<const uint4*>(&ed[e]);
      a0[e]         = ((t >= E.z) ? E.y : E.x) + tb;
    }
    const uint32_t om1 = __byte_perm(w0, 0, 0x4140), om2 = __byte_perm(w0, 0, 0x4342);
    const uint32_t odl = h2_add(om2, om1 ^ H2_SIGN); // m2 - m1 >= 0
    const uint32_t oid = (D <= 8) ? __byte_perm(w1, 0, 0x4341) : ((D <= 16) ? w2 : __byte_perm(w2, 0, 0x4341));
#pragma unroll
    for (int e = 0; e < D; ++e) {
      const uint32_t sw  = (e < 16) ? w1 : w2;
      const uint32_t ism = h2_eq_bf(oid, FEC_H2(e));                   // edge e was the min1 edge
      const uint32_t mag = h2_fma(ism, odl, om1);                      // |old c2v|
      const uint32_t sgs = imad(sw, 1u << (15 - (e & 15)), 0u);        // its sign bits at 15 / 31
      cneg[e]            = mag ^ (~sgs & H2_SIGN);                     // -(old c2v)
    }
  }
  asm volatile("bar.sync 0;" ::: "memory");

  // ---- phase A: variable-to-check messages and the min1 / min2 / sign-product analysis -----------------------------------
  uint32_t up[D];
#pragma unroll
  for (int e = 0; e < D; ++e) {
    up[e] = imad(lds_u8(a0[e] ^ 1u), 65536u, lds_u8(a0[e])); // lane 0 at A, lane 1 at A ^ 1
  }
  uint32_t v[D];
  uint32_t m1 = H2_P120, m2 = H2_P120;
#pragma unroll
  for (int e = 0; e < D; ++e) {
    const uint32_t S    = h2_add(up[e], H2_M128);                                    // soft value
    const uint32_t vraw = h2_add(S, cneg[e]);                                        // soft - c2v, |.| <= 223
    const uint32_t y    = h2_add(h2_fma(S, H2_F4352, H2_F96), H2_FM96);              // +-1/16 iff |S| > 120
    const uint32_t vcl  = h2_min_xs(vraw, H2_P120);                                  // saturate at +-120
    const uint32_t vv   = h2_fma(y, H2_F2M11, vcl);                                  // "infinite" soft bits stay huge
    v[e]                = vv;
    const uint32_t mx   = h2_max_xs(m1, vv);
    m2                  = h2_min_xs(m2, mx);
    m1                  = h2_min_xs(m1, vv);
  }

  // ---- scale the two minima: scaled = (x * mul + add) >> 8, both lanes in one 32-bit multiply -----------------------------
  const uint32_t sp  = m1 & H2_SIGN; // product of all v2c signs
  const uint32_t a1  = m1 & H2_ABS;
  const uint32_t s1  = FEC_SHR(imad(a1, scale_mul, scale_add), 8) & 0x00ff00ffu;
  const uint32_t s2  = FEC_SHR(imad(m2 & H2_ABS, scale_mul, scale_add), 8) & 0x00ff00ffu;
  const uint32_t M1s = s1 | sp;
  const uint32_t dls = h2_add(s2, s1 ^ H2_SIGN) | sp;

  // ---- phase B: new check-to-variable messages, soft-bit update, write-back ------------------------------------------------
  uint32_t nid = 0, sg1 = 0, sg2 = 0;
#pragma unroll
  for (int e = 0; e < D; ++e) {
    const uint32_t vv = v[e];
    const uint32_t eq = h2_eq_bf(h2_abs(vv), a1);        // this edge holds min1 -> it gets min2
    const uint32_t mg = h2_fma(eq, dls, M1s);            // sign product * scaled magnitude
    const uint32_t c  = mg ^ (vv & H2_SIGN);             // new c2v
    nid               = h2_fma(eq, FEC_H2(e), nid);      // index of the min1 edge (any value if min1 == min2)
    const uint32_t tt = h2_add(c, vv);                   // c2v + v2c
    const uint32_t nw = h2_min_xs(tt, H2_P127);          // |.| > 120 is "infinity" from here on
    const uint32_t o  = h2_add(nw, H2_P128);             // offset binary in the low byte of each lane
    // sign(v2c) of the two lanes -> bit (e & 15) of the sign word
    const uint32_t sh = ((e & 15) == 15) ? vv : FEC_SHR(vv, 15 - (e & 15));
    if (e < 16) {
      sg1 |= sh & (0x00010001u << (e & 15));
    } else {
      sg2 |= sh & (0x00010001u << (e & 15));
    }
    sts_u8(a0[e], o);
    sts_u8(a0[e] ^ 1u, FEC_SHR(o, 16));
  }

  // ---- new state: c2v sign = sign product ^ sign(v2c) ----------------------------------------------------------------------
  const uint32_t pm = prmt(m1, 0u, 0xbb99); // 0xffff in the lanes whose sign product is negative
  if (TM) {
    tmem_st2(taddr, __byte_perm(s1, s2, 0x6420), ((sg1 ^ pm) & 0x00ff00ffu) | (nid << 8));
    return;
  }
  st[0] = __byte_perm(s1, s2, 0x6420);
  if (D <= 8) {
    st[H] = ((sg1 ^ pm) & 0x00ff00ffu) | (nid << 8);
  } else if (D <= 16) {
    st[H]     = sg1 ^ pm;
    st[2 * H] = nid;
  } else {
    st[H]     = sg1 ^ pm;
    st[2 * H] = ((sg2 ^ pm) & 0x00070007u) | (nid << 8);
  }
}

// One decoder iteration over the layers in use. Out of line on purpose: ptxas allocates registers for the row code on its
// own, so changes to the dematcher or the tail cannot perturb the schedule of the hot loop (and vice versa).
__device__ __noinline__ void fec_packed_iteration(uint32_t               soft_s,
                                                  uint32_t*              state,
                                                  const fec_packed_edge* edges,
                                                  const uint16_t*        soff,
                                                  const uint16_t*        rs,
                                                  const uint8_t*         deg,
                                                  uint32_t               H,
                                                  uint32_t               nlayers,
                                                  uint32_t               smul,
                                                  uint32_t               sadd,
                                                  uint32_t               row_tm, // first base row whose state is in tensor memory
                                                  uint32_t               tm_row0) // this warp's tensor-memory address of that row
{
  const uint32_t tid = threadIdx.x;
  for (uint32_t l = 0; l < nlayers; ++l) {
    if ((tid & ~31u) >= H) {
      asm volatile("bar.sync 0;" ::: "memory"); // warps without work still take part in the row barrier
    } else {
      // Lanes past H in the last warp repeat the work of thread H-1 (identical loads, identical stores; their own, identical
      // copy of the tensor-memory state) so that whole warps reach the barrier inside the row function.
      const uint32_t         tq = min(tid, H - 1);
      const fec_packed_edge* ed = edges + rs[l];
      if (l >= row_tm) {
        const uint32_t ta = tm_row0 + 2u * (l - row_tm);
        switch (deg[l]) {
          case 3: fec_packed_row<3, true>(soft_s, nullptr, ta, ed, H, tq, smul, sadd); break;
          case 4: fec_packed_row<4, true>(soft_s, nullptr, ta, ed, H, tq, smul, sadd); break;
          case 5: fec_packed_row<5, true>(soft_s, nullptr, ta, ed, H, tq, smul, sadd); break;
          case 6: fec_packed_row<6, true>(soft_s, nullptr, ta, ed, H, tq, smul, sadd); break;
          case 7: fec_packed_row<7, true>(soft_s, nullptr, ta, ed, H, tq, smul, sadd); break;
          default: fec_packed_row<8, true>(soft_s, nullptr, ta, ed, H, tq, smul, sadd); break;
        }
      } else {
        uint32_t* st = state + (uint32_t)soff[l] * H + tq;
        switch (deg[l]) {
          case 3: fec_packed_row<3, false>(soft_s, st, 0u, ed, H, tq, smul, sadd); break;
          case 4: fec_packed_row<4, false>(soft_s, st, 0u, ed, H, tq, smul, sadd); break;
          case 5: fec_packed_row<5, false>(soft_s, st, 0u, ed, H, tq, smul, sadd); break;
          case 6: fec_packed_row<6, false>(soft_s, st, 0u, ed, H, tq, smul, sadd); break;
          case 7: fec_packed_row<7, false>(soft_s, st, 0u, ed, H, tq, smul, sadd); break;
          case 8: fec_packed_row<8, false>(soft_s, st, 0u, ed, H, tq, smul, sadd); break;
          case 9: fec_packed_row<9, false>(soft_s, st, 0u, ed, H, tq, smul, sadd); break;
          case 10: fec_packed_row<10, false>(soft_s, st, 0u, ed, H, tq, smul, sadd); break;
          default: fec_packed_row<19, false>(soft_s, st, 0u, ed, H, tq, smul, sadd); break;
        }
      }
    }
  }
}

// Byte address of variable j (0 <= j < Z) of column col in the interleaved soft layout.
__device__ __forceinline__ uint32_t fec_packed_soft_addr(uint32_t col, uint32_t j, uint32_t Z, uint32_t H)
{
  const uint32_t lane = (j >= H) ? 1u : 0u;
  return col * Z + (j - lane * H) * 2 + lane;
}

// Dematcher / loader target: clamp to +-64 (ldpc_decoder_impl.cpp:168-169) and store offset-binary in the packed layout.
struct fec_soft_store_packed {
  uint8_t* soft;
  uint32_t Z, H, invZ;
  __device__ __forceinline__ void operator()(uint32_t o, int acc) const
  {
    raw(o, fec_clampi(acc, -64, 64));
  }
  // Without the clamp: the partial last node of a decoder input that is not a multiple of Z (ldpc_decoder_impl.cpp:186-193).
  __device__ __forceinline__ void raw(uint32_t o, int acc) const
  {
    const uint32_t op  = o + 2 * Z;
    const uint32_t col = __umulhi(op, invZ);
    soft[fec_packed_soft_addr(col, op - col * Z, Z, H)] = (uint8_t)(acc + 128);
  }
};

// Hard decision of the first K soft bits (bit = soft <= 0) into bitsw, big-endian bit order; true if no soft bit is zero.
__device__ __forceinline__ bool fec_hard_bits_packed(const uint8_t* soft, uint32_t* bitsw, uint32_t K, uint32_t Z, uint32_t H,
                                                     uint32_t invZ)
{
  const uint32_t lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nwarps = blockDim.x >> 5;
  int            zero = 0;
  if ((H & 31u) == 0) {
    // One column per step: thread t holds variables t and t+H of the column; a ballot yields 32 consecutive bits.
    // Kb is 22 or 10: batches of 11 / 10 loads in flight, then the ballots.
    const uint32_t Kb = K / Z, nb = (Kb == 22) ? 11u : 10u;
    if (threadIdx.x < H) {
      for (uint32_t c0 = 0; c0 < Kb; c0 += nb) {
        uint32_t hw[11];
#pragma unroll
        for (uint32_t c = 0; c < 11; ++c) {
          hw[c] = (c < nb) ? *reinterpret_cast<const uint16_t*>(soft + (c0 + c) * Z + threadIdx.x * 2) : 0x8181u;
        }
#pragma unroll
        for (uint32_t c = 0; c < 11; ++c) {
          if (c < nb) {
            const uint32_t b0 = hw[c] & 0xffu, b1 = hw[c] >> 8;
            zero |= (b0 == 128) | (b1 == 128);
            const uint32_t bal0 = __ballot_sync(0xffffffffu, b0 <= 128);
            const uint32_t bal1 = __ballot_sync(0xffffffffu, b1 <= 128);
            if (lane == 0) {
              bitsw[((c0 + c) * Z + warp * 32) >> 5]     = __brev(bal0);
              bitsw[((c0 + c) * Z + H + warp * 32) >> 5] = __brev(bal1);
            }
          }
        }
      }
    }
    return __syncthreads_or(zero) == 0;
  }
  const uint32_t nw = (K + 31) / 32;
  for (uint32_t w = warp; w < nw; w += nwarps) {
    const uint32_t i = w * 32 + lane;
    uint32_t       b = 129; // positive
    if (i < K) {
      const uint32_t col = __umulhi(i, invZ);
      b                  = soft[fec_packed_soft_addr(col, i - col * Z, Z, H)];
    }
    zero |= (b == 128);
    const uint32_t bal = __ballot_sync(0xffffffffu, b <= 128);
    if (lane == 0) {
      bitsw[w] = __brev(bal);
    }
  }
  return __syncthreads_or(zero) == 0;
}

// CRC remainder == 0 test of the first T bits of bitsw (crc_calculator_generic_impl.cpp:111-133 computes the same
// remainder bytewise). Byte i of the message contributes byte_i(x) * x^(8 (nb-1-i)) mod g; x8[j] = x^(8 j) mod g is staged
// in shared memory once per codeblock. Up to six independent 8-step multiply chains per thread, one barrier.
// Kept out of line: inlined into the codeblock function it changes ptxas' register allocation of the row code for the worse.
__device__ __noinline__ bool fec_crc_ok_packed(const uint32_t* bitsw, uint32_t T, const uint32_t* x8, uint32_t poly,
                                                  uint32_t top, uint32_t* red)
{
  const uint32_t nb = T / 8, r = T % 8;
  uint32_t       acc = 0;
  for (uint32_t i0 = threadIdx.x; i0 < nb; i0 += 6 * blockDim.x) {
    uint32_t a[6], w[6];
#pragma unroll
    for (int k = 0; k < 6; ++k) {
      const uint32_t i = i0 + k * blockDim.x;
      const bool     v = (i < nb);
      a[k]             = v ? x8[nb - 1 - i] : 0u;
      w[k]             = v ? (bitsw[i >> 2] >> (24 - 8 * (i & 3))) : 0u;
    }
#pragma unroll
    for (int s = 0; s < 8; ++s) {
#pragma unroll
      for (int k = 0; k < 6; ++k) {
        acc ^= ((w[k] >> s) & 1u) ? a[k] : 0u;
        a[k] = fec_crc_shift1(a[k], poly, top);
      }
    }
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    acc ^= __shfl_xor_sync(0xffffffffu, acc, o);
  }
  if ((threadIdx.x & 31) == 0) {
    red[threadIdx.x >> 5] = acc;
  }
  __syncthreads();
  uint32_t rem = 0;
  for (uint32_t i = 0; i < (blockDim.x >> 5); ++i) {
    rem ^= red[i];
  }
  if (r != 0) {
    // rem = (the nb full bytes) mod g; continue bitwise over the remaining bits.
    const uint32_t last = bitsw[nb >> 2] >> (24 - 8 * (nb & 3));
    for (uint32_t i = 0; i < r; ++i) {
      rem = (rem << 1) | ((last >> (7 - i)) & 1u);
      rem = (rem & top) ? (rem ^ poly) : rem;
    }
  }
  return rem == 0;
}

// ---- rate dematcher of the packed path ------------------------------------------------------------------------------------
// Same arithmetic as fec_dematch_cta (the per-byte statement of ldpc_rate_dematcher_impl.cpp:46-257), organised for
// throughput:
//  1. the E input LLRs are read once with coalesced 16-byte loads and scattered into shared memory (`stage`: the not yet
//     used check-state area) in RATE-MATCHED order, i.e. the bit de-interleaver of TS 38.212 5.4.2.2 is undone while
//     staging. The staged position of LLR p is p + d, d in {d1, d2} chosen so that a staged 32-bit word lines up with a
//     32-bit word of the soft buffer inside either data run of a non-wrapping transmission;
//  2. the zero tail [zero_from, N) of a first transmission is stored without looking at anything (the decoder's soft bits
//     were preset to zero);
//  3. groups of four soft-buffer bytes inside a data run take one 32-bit shared load, one 32-bit HBM store (plus a 32-bit
//     load and four saturating sums when combining) and four byte stores into the decoder's interleaved layout;
//  4. every other group (region boundaries, fillers, kept parts of the buffer, wrapping transmissions) takes the general
//     path: still four bytes per thread, per-byte rules.
// Inputs that do not fit the staging area (E > 76 KB at Z = 384: heavy repetition) are read from global memory per byte.
struct fec_dematch_geom {
  uint32_t Z, N, Ncb, fillers, sys, info, E, qm, Kp, invKp, Lc, k0c, zero_lo, zero_from, kstart;
  bool     newd, hzero, wrapped;
};

__device__ __forceinline__ fec_dematch_geom fec_dematch_geometry(const fec_cb_desc& D)
{
  fec_dematch_geom G;
  G.Z                = D.lifting_size;
  const uint32_t Kb  = (D.base_graph == 1) ? 22u : 10u;
  G.N                = ((D.base_graph == 1) ? 66u : 50u) * G.Z;
  G.Ncb              = (D.Nref > 0) ? min(D.Nref, G.N) : G.N;
  G.fillers          = D.nof_filler_bits;
  G.sys              = (Kb - 2) * G.Z;
  G.info             = G.sys - G.fillers;
  G.E                = D.E;
  G.qm               = D.qm;
  G.Kp               = G.E / G.qm;
  G.invKp            = 0xffffffffu / G.Kp + 1; // unused when Kp == 1
  G.newd             = (D.flags & FEC_FLAG_NEW_DATA) != 0;
  G.hzero            = (D.flags & FEC_FLAG_HARQ_ZERO) != 0;
  G.Lc               = G.Ncb - G.fillers;                                      // non-filler positions of the circular buffer
  G.k0c              = (D.k0 < G.info) ? D.k0 : (max(D.k0, G.sys) - G.fillers); // start, in filler-free coordinates
  G.kstart           = (D.k0 < G.info) ? D.k0 : max(D.k0, G.sys);              // same, as a buffer position
  G.zero_lo          = min(D.k0, G.info);                                      // [0, zero_lo) is zeroed in copy mode
  const uint32_t e1  = G.Lc - G.k0c;                                           // LLRs consumed before the first wrap
  G.wrapped          = G.E > e1;
  G.zero_from        = G.N;                                                    // [zero_from, N) zeroed at the end
  if (G.newd && !G.wrapped) {
    const uint32_t q_end = G.k0c + G.E;
    uint32_t       idxf  = (G.k0c < G.info && q_end <= G.info) ? G.sys : (q_end + G.fillers);
    idxf                 = (idxf >= G.Ncb) ? idxf - G.Ncb : idxf;
    if (idxf != 0) {
      G.zero_from = G.N - (G.Ncb - idxf);
    }
  }
  return G;
}

__device__ __forceinline__ uint32_t fec_dematch_src_index(const fec_dematch_geom& G, uint32_t p)
{
  if (G.qm == 1) {
    return p;
  }
  const uint32_t j = (G.Kp == 1) ? p : __umulhi(p, G.invKp);
  return (p - j * G.Kp) * G.qm + j;
}

// Where rate-matched LLR number p is read from.
struct fec_dematch_src {
  const int8_t* base;   // staged copy (shared memory, rate-matched order) or the global input (interleaved)
  bool          staged;
  uint32_t      split;  // first p of the second data run (the one after the filler gap)
  uint32_t      d1, d2; // staged position of p is p + (p < split ? d1 : d2)
};

__device__ __forceinline__ uint32_t fec_dematch_stage_pos(const fec_dematch_src& S, uint32_t p)
{
  return p + ((p < S.split) ? S.d1 : S.d2);
}

__device__ __forceinline__ int fec_dematch_fetch(const fec_dematch_geom& G, const fec_dematch_src& S, uint32_t p)
{
  return S.staged ? S.base[fec_dematch_stage_pos(S, p)] : S.base[fec_dematch_src_index(G, p)];
}

// One soft-buffer byte, any case. `old` is the previous buffer value (0 when the slot reads as zeros).
__device__ __forceinline__ int fec_dematch_byte(const fec_dematch_geom& G, const fec_dematch_src& S, uint32_t o, int old)
{
  const bool is_filler = (o >= G.info) && (o < G.sys);
  int        acc;
  if (G.newd && o < G.zero_lo) {
    acc = 0;
  } else if (G.newd && is_filler) {
    acc = 127;
  } else {
    acc = old;
  }
  if (!is_filler && o < G.Ncb) {
    const uint32_t q     = (o < G.info) ? o : (o - G.fillers);
    const bool     first = (q >= G.k0c);
    uint32_t       p     = first ? (q - G.k0c) : (q + G.Lc - G.k0c);
    bool           copy  = G.newd && first;
    for (; p < G.E; p += G.Lc) {
      const int dv = fec_dematch_fetch(G, S, p);
      acc          = copy ? dv : fec_sat_add(acc, dv);
      copy         = false;
    }
  }
  return (o >= G.zero_from) ? 0 : acc;
}

// Four consecutive soft-buffer values (packed little-endian in w, as in the HARQ buffer) into the decoder's soft array.
__device__ __forceinline__ void fec_soft_store4(const fec_soft_store_packed& store, uint32_t o0, uint32_t w)
{
  const uint32_t Z = store.Z, H = store.H;
  const uint32_t op0 = o0 + 2 * Z;
  const uint32_t col = __umulhi(op0, store.invZ);
  const uint32_t j0  = op0 - col * Z;
  int            val[4];
#pragma unroll
  for (int b = 0; b < 4; ++b) {
    val[b] = (int)(int8_t)(w >> (8 * b));
  }
  if (j0 + 3 < H || (j0 >= H && j0 + 3 < Z)) {
    const uint32_t lane = (j0 >= H) ? 1u : 0u;
    uint8_t*       dst  = store.soft + col * Z + (j0 - lane * H) * 2 + lane;
#pragma unroll
    for (int b = 0; b < 4; ++b) {
      dst[2 * b] = (uint8_t)(fec_clampi(val[b], -64, 64) + 128);
    }
  } else {
#pragma unroll
    for (int b = 0; b < 4; ++b) {
      store(o0 + b, val[b]);
    }
  }
}

__device__ __forceinline__ uint32_t fec_pack4(const int* val)
{
  return (uint32_t)(uint8_t)val[0] | ((uint32_t)(uint8_t)val[1] << 8) | ((uint32_t)(uint8_t)val[2] << 16) |
         ((uint32_t)(uint8_t)val[3] << 24);
}

// Position + 1 of the last non-zero byte of group o0 (0 if the group is all zero).
__device__ __forceinline__ uint32_t fec_last_sig(uint32_t o0, uint32_t w)
{
  return (w != 0) ? (o0 + 4 - (__clz(w) >> 3)) : 0u;
}

// The soft buffer of one codeblock, four bytes per thread and step. Returns position + 1 of this thread's last non-zero value.
__device__ __forceinline__ uint32_t fec_dematch_groups(const fec_dematch_geom& G,
                                                       const fec_dematch_src&  S,
                                                       int8_t* __restrict__    harq,
                                                       uint8_t*                soft,
                                                       uint32_t                H,
                                                       uint32_t                invZ)
{
  uint32_t                    my_sig  = 0;
  const uint32_t              ngroups = G.N / 4; // Z is even and N = 66Z or 50Z, so N % 4 == 0
  uint32_t*                   harq4   = reinterpret_cast<uint32_t*>(harq);
  const fec_soft_store_packed store{soft, G.Z, H, invZ};

  // Data runs of a non-wrapping transmission: rate-matched LLR p = o - off lands at buffer position o for o in [lo, hi),
  // once before the filler gap and once after it. hg0 / hg1: the groups entirely inside a run (staged inputs only).
  uint32_t hg0[2] = {1, 1}, hg1[2] = {0, 0}, hoff[2] = {0, 0};
  if (!G.wrapped && S.staged) {
    if (G.k0c < G.info) {
      hg0[0]  = (G.k0c + 3) / 4;
      hg1[0]  = min(G.k0c + G.E, G.info) / 4;
      hoff[0] = G.k0c - S.d1;
    }
    if (G.k0c + G.E > G.info) {
      hg0[1]  = (max(G.k0c, G.info) + G.fillers + 3) / 4;
      hg1[1]  = (G.k0c + G.E + G.fillers) / 4;
      hoff[1] = G.k0c + G.fillers - S.d2;
    }
  }
  const uint32_t gz0 = (G.zero_from + 3) / 4; // groups from here on are all zero

  // (2) zero tail
  for (uint32_t g = gz0 + threadIdx.x; g < ngroups; g += blockDim.x) {
    harq4[g] = 0; // the decoder's soft bits were preset to zero
  }

  // (3) groups inside a data run
#pragma unroll
  for (int r = 0; r < 2; ++r) {
    for (uint32_t g = hg0[r] + threadIdx.x; g < hg1[r]; g += blockDim.x) {
      const uint32_t o0 = g * 4;
      uint32_t       w  = *reinterpret_cast<const uint32_t*>(S.base + (o0 - hoff[r]));
      if (!G.newd) {
        const uint32_t old = G.hzero ? 0u : harq4[g];
        int            val[4];
#pragma unroll
        for (int b = 0; b < 4; ++b) {
          val[b] = fec_sat_add((int)(int8_t)(old >> (8 * b)), (int)(int8_t)(w >> (8 * b)));
        }
        w = fec_pack4(val);
      }
      harq4[g] = w;
      my_sig   = max(my_sig, fec_last_sig(o0, w));
      if (soft != nullptr) {
        fec_soft_store4(store, o0, w);
      }
    }
  }

  // (4) everything else
  for (uint32_t g = threadIdx.x; g < gz0; g += blockDim.x) {
    if ((g >= hg0[0] && g < hg1[0]) || (g >= hg0[1] && g < hg1[1])) {
      continue;
    }
    const uint32_t o0 = g * 4;
    int            val[4];
    // A group is simple when no region boundary falls strictly inside it.
    auto inside = [o0](uint32_t b) { return (b > o0) && (b < o0 + 4); };
    bool simple = !G.wrapped && !inside(G.zero_lo) && !inside(G.info) && !inside(G.sys) && !inside(G.Ncb) &&
                  !inside(G.zero_from) && !inside(G.kstart);
    const bool is_filler = (o0 >= G.info) && (o0 < G.sys);
    const bool live      = !is_filler && (o0 < G.Ncb);
    uint32_t   p0        = 0;
    bool       has_data = false, first = false;
    if (simple && live) {
      const uint32_t q0 = (o0 < G.info) ? o0 : (o0 - G.fillers);
      first             = (q0 >= G.k0c);
      p0                = first ? (q0 - G.k0c) : (q0 + G.Lc - G.k0c);
      has_data          = (p0 + 3 < G.E);
      simple            = has_data || (p0 >= G.E); // the end of the data must not fall inside the group
    }
    if (simple) {
      const bool preset   = G.newd && (o0 < G.zero_lo);
      const bool copy     = G.newd && first && has_data;
      const bool zero_all = (o0 >= G.zero_from);
      const bool uses_old = !zero_all && (is_filler ? !G.newd : (live ? (!copy && !preset) : true));
      uint32_t   old      = 0;
      if (uses_old && !G.hzero) {
        old = harq4[g];
      }
      const bool ld = live && has_data;
#pragma unroll
      for (int b = 0; b < 4; ++b) {
        const int dv   = ld ? fec_dematch_fetch(G, S, p0 + b) : 0;
        const int base = (is_filler && G.newd) ? 127 : (int)(int8_t)(old >> (8 * b));
        const int comb = fec_sat_add(base, dv);
        const int a    = ld ? (copy ? dv : comb) : base;
        val[b]         = zero_all ? 0 : a;
      }
    } else {
      uint32_t old = 0;
      if (!G.hzero) {
        old = harq4[g];
      }
#pragma unroll
      for (int b = 0; b < 4; ++b) {
        val[b] = fec_dematch_byte(G, S, o0 + b, (int)(int8_t)(old >> (8 * b)));
      }
    }
    const uint32_t w = fec_pack4(val);
    harq4[g]         = w;
    my_sig           = max(my_sig, fec_last_sig(o0, w));
    if (soft != nullptr) {
      fec_soft_store4(store, o0, w);
    }
  }
  return my_sig;
}

__device__ __forceinline__ void fec_dematch_packed(const fec_cb_desc& D,
                                                   const int8_t* __restrict__ in,
                                                   int8_t* __restrict__ harq,
                                                   uint8_t*  soft, // interleaved decoder layout, nullptr = do not fill
                                                   uint32_t  H,
                                                   int8_t*   stage,
                                                   uint32_t  stage_cap,
                                                   uint32_t* n_sig_smem)
{
  const fec_dematch_geom G = fec_dematch_geometry(D);
  const uint32_t         Z = G.Z, invZ = 0xffffffffu / Z + 1;

  fec_dematch_src S;
  S.base   = in;
  S.staged = false;
  S.split  = (G.k0c < G.info) ? min(G.E, G.info - G.k0c) : 0u;
  S.d1     = G.k0c & 3u;
  S.d2     = S.d1 + (G.fillers & 3u);
  if (G.E + 8 <= stage_cap) {
    // (1) stage and de-interleave: input byte s = i * Qm + j is rate-matched LLR p = j * (E / Qm) + i.
    const uint32_t inv_qm = 0xffffffffu / G.qm + 1;
    auto           put    = [&](uint32_t s, uint32_t byte) {
      const uint32_t i = (G.qm == 1) ? s : __umulhi(s, inv_qm);
      const uint32_t p = (s - i * G.qm) * G.Kp + i;
      stage[fec_dematch_stage_pos(S, p)] = (int8_t)byte;
    };
    uint32_t done = 0;
    if ((reinterpret_cast<uintptr_t>(in) & 15u) == 0) {
      const uint4*   g4 = reinterpret_cast<const uint4*>(in);
      const uint32_t n4 = G.E / 16;
      for (uint32_t k = threadIdx.x; k < n4; k += blockDim.x) {
        const uint4    v    = __ldg(g4 + k);
        const uint32_t w[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
        for (uint32_t m = 0; m < 16; ++m) {
          put(16 * k + m, w[m / 4] >> (8 * (m % 4)));
        }
      }
      done = n4 * 16;
    }
    for (uint32_t s = done + threadIdx.x; s < G.E; s += blockDim.x) {
      put(s, (uint32_t)(uint8_t)in[s]);
    }
    __syncthreads();
    S.base   = stage;
    S.staged = true;
  }

  const uint32_t my_sig = fec_dematch_groups(G, S, harq, soft, H, invZ);
  if (my_sig != 0) {
    atomicMax(n_sig_smem, my_sig);
  }
}


__device__ __forceinline__ bool fec_packed_cb_supported(const fec_cb_desc& D)
{
  // Z must split into 2 interleaved lanes: every even lifting size, all three modes. Odd Z (<= 15) and Z = 2 take the
  // generic per-check code, whose shared-memory layout is tiny at those sizes.
  return (D.lifting_size % 2 == 0) && (D.lifting_size >= 4) && (D.lifting_size / 2 <= blockDim.x);
}

__device__ __forceinline__ void fec_packed_cb(const fec_cb_desc& D,
                                           uint32_t           cb,
                                           uint8_t*           smem_raw,
                                           const int8_t* __restrict__ llr_base,
                                           uint8_t* __restrict__ out_base,
                                           fec_cb_result* __restrict__ results,
                                           int8_t* __restrict__ harq_base,
                                           const fec_dev_tables* __restrict__ tab,
                                           const fec_packed_tables* __restrict__ ptab,
                                           uint32_t tmem_base)
{
  __shared__ uint32_t s_nsig_arr[2];
#define s_nsig s_nsig_arr[0]

  const uint32_t Z = D.lifting_size, H = Z / 2;
  const int      bg1 = (D.base_graph == 1);
  const uint32_t Kb = bg1 ? 22u : 10u, nfull = bg1 ? 68u : 52u, M = bg1 ? 46u : 42u;
  const uint32_t K = Kb * Z, N = (nfull - 2) * Z;
  const uint32_t ne = bg1 ? 316u : 197u;
  const uint32_t invZ = 0xffffffffu / Z + 1;
  const uint32_t tid = threadIdx.x;

#ifdef FEC_PROFILE
  long long tp0 = clock64(), tp1 = 0, tp2 = 0, tp3 = 0;
#endif
  const fec_packed_smem_layout L = fec_packed_layout(Z, bg1);
  uint32_t*        state = reinterpret_cast<uint32_t*>(smem_raw + L.state);
  fec_packed_edge* edges = reinterpret_cast<fec_packed_edge*>(smem_raw + L.edges);
  uint16_t*        soff  = reinterpret_cast<uint16_t*>(smem_raw + L.soff);
  uint16_t*        rs    = reinterpret_cast<uint16_t*>(smem_raw + L.rs);
  uint8_t*         deg   = smem_raw + L.deg;
  uint32_t*        bitsw = reinterpret_cast<uint32_t*>(smem_raw + L.bitsw);
  uint32_t*        red   = reinterpret_cast<uint32_t*>(smem_raw + L.red);
  uint32_t*        x8    = reinterpret_cast<uint32_t*>(smem_raw + L.x8);
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
    // All soft bits start at zero (offset-binary 128): the two punctured columns, and whatever the dematcher skips.
    uint4*         p4  = reinterpret_cast<uint4*>(soft);
    const uint32_t ns4 = (L.total - L.soft) / 16;
    for (uint32_t i = tid; i < ns4; i += blockDim.x) {
      p4[i] = make_uint4(0x80808080u, 0x80808080u, 0x80808080u, 0x80808080u);
    }
    const uint32_t nx8 = (K - D.nof_filler_bits) / 8;
    for (uint32_t i = tid; i < nx8; i += blockDim.x) {
      x8[i] = tab->crc_x8[D.crc_type][i];
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

#ifdef FEC_PROFILE
  tp1 = clock64();
#endif
  // ---- stage 1: soft bits into shared memory -----------------------------------------------------------------------------
  const fec_soft_store_packed store{soft, Z, H, invZ};
  if (D.mode == FEC_MODE_DECODE) {
    const uint32_t n_in   = D.E;
    const uint32_t n_node = (n_in / Z) * Z; // full nodes are clamped to +-64, a partial last node is copied as it is
    uint32_t       my_sig = 0;
    for (uint32_t o = tid; o < N; o += blockDim.x) {
      int vv = 0;
      if (o < n_in) {
        vv = in[o];
        if (vv != 0) {
          my_sig = o + 1;
        }
      }
      if (o < n_node || o >= n_in) {
        store(o, vv);
      } else {
        store.raw(o, vv);
      }
    }
    if (my_sig != 0) {
      atomicMax(&s_nsig, my_sig);
    }
  } else {
    int8_t* harq = harq_base + (size_t)D.harq_slot * FEC_HARQ_SLOT_BYTES;
    fec_dematch_packed(D, in, harq, (D.mode == FEC_MODE_CHAIN) ? soft : nullptr, H, reinterpret_cast<int8_t*>(state),
                       L.edges - L.state, &s_nsig);
    if (D.mode == FEC_MODE_DEMATCH) {
      if (tid == 0) {
        res.status  = FEC_RES_DONE;
        results[cb] = res;
      }
      return;
    }
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
  // Tensor-memory home of this warp's check state: lane quadrant of the warp, 64 columns per group of four warps.
  const uint32_t row_tm  = fec_packed_first_tmem_row(bg1);
  const uint32_t tm_row0 = tmem_base + ((((tid >> 5) & 3u) * 32u) << 16) + (tid >> 7) * FEC_PACKED_TMEM_COLS_PER_WARPSLOT;
  {
    // The check state doubles as the dematcher's staging area: clear the rows in use ("no c2v messages yet").
    uint4*         s4 = reinterpret_cast<uint4*>(state);
    const uint32_t n4 = ((uint32_t)soff[min(nlayers, row_tm)] * H * 4 + 15) / 16;
    for (uint32_t i = tid; i < n4; i += blockDim.x) {
      s4[i] = make_uint4(0, 0, 0, 0);
    }
    // ... and the tensor-memory rows in use, eight columns per store (every thread clears its own lane)
    if (nlayers > row_tm) {
      for (uint32_t c = 0; c < 2u * (nlayers - row_tm); c += 8) {
        tmem_st8_zero(tm_row0 + c);
      }
      tmem_wait_st();
    }
  }
  const uint32_t crc_poly = tab->crc_poly[D.crc_type], crc_top = 1u << tab->crc_len[D.crc_type];
  __syncthreads();
  const uint32_t smul = ptab->scale_mul, sadd = ptab->scale_add;
  const uint32_t soft_s = (uint32_t)__cvta_generic_to_shared(soft);

#ifdef FEC_PROFILE
  tp2 = clock64();
#endif
  bool     pass = false;
  uint32_t it   = 0;
  for (; it < D.max_iter; ++it) {
    fec_packed_iteration(soft_s, state, edges, soff, rs, deg, H, nlayers, smul, sadd, row_tm, tm_row0);
    __syncthreads(); // the last row's stores, before the hard decision or the next iteration
    if (early_stop) {
      const bool clean = fec_hard_bits_packed(soft, bitsw, K, Z, H, invZ);
      const bool crc   = fec_crc_ok_packed(bitsw, T, x8, crc_poly, crc_top, red);
      if (clean && crc) {
        pass = true;
        ++it;
        break;
      }
    }
  }
#ifdef FEC_PROFILE
  tp3 = clock64();
#endif
  if (!early_stop) {
    fec_hard_bits_packed(soft, bitsw, K, Z, H, invZ);
    pass = fec_crc_ok_packed(bitsw, T, x8, crc_poly, crc_top, red) && !(D.flags & FEC_FLAG_NO_CRC);
  }
  if ((reinterpret_cast<uintptr_t>(out) & 3u) == 0) {
    uint32_t* out4 = reinterpret_cast<uint32_t*>(out);
    for (uint32_t i = tid; i < kbytes / 4; i += blockDim.x) {
      out4[i] = __byte_perm(bitsw[i], 0, 0x0123);
    }
    for (uint32_t i = (kbytes & ~3u) + tid; i < kbytes; i += blockDim.x) {
      out[i] = (uint8_t)(bitsw[i >> 2] >> (24 - 8 * (i & 3)));
    }
  } else {
    for (uint32_t i = tid; i < kbytes; i += blockDim.x) {
      out[i] = (uint8_t)(bitsw[i >> 2] >> (24 - 8 * (i & 3)));
    }
  }
  if (tid == 0) {
    res.nof_ldpc_iterations = (uint16_t)(pass && early_stop ? it : D.max_iter);
    res.status              = (pass ? FEC_RES_CRC_PASS : 0) | FEC_RES_DONE;
    res.nof_layers          = (uint8_t)nlayers;
#ifdef FEC_PROFILE
    // phase clocks / 256: [7:0] setup, [15:8] dematch / 4, [23:16] decode / 16, [31:24] tail
    const long long tp4 = clock64();
    res.reserved = (uint32_t)min(255ll, (tp1 - tp0) >> 8) | ((uint32_t)min(255ll, (tp2 - tp1) >> 10) << 8) |
                   ((uint32_t)min(255ll, (tp3 - tp2) >> 12) << 16) | ((uint32_t)min(255ll, (tp4 - tp3) >> 8) << 24);
#endif
    results[cb]            = res;
  }
}

#undef s_nsig
// Dispatcher: every codeblock the packed path covers goes there, the rest (odd Z, dematch-only
// operations, unaligned decoder inputs) runs the generic per-check code in the same launch.
__global__ void __launch_bounds__(FEC_PACKED_MAX_THREADS, 3)
fec_cb_packed_kernel(const fec_cb_desc* __restrict__ descs,
                     const int8_t* __restrict__ llr_base,
                     uint8_t* __restrict__ out_base,
                     fec_cb_result* __restrict__ results,
                     int8_t* __restrict__ harq_base,
                     const fec_dev_tables* __restrict__ tab,
                     const fec_packed_tables* __restrict__ ptab)
{
  extern __shared__ __align__(16) uint8_t smem_raw[];
  __shared__ uint32_t s_tmem_base;
  const fec_cb_desc D = descs[blockIdx.x];
  if (!fec_packed_cb_supported(D)) {
    fec_generic_cb(D, blockIdx.x, smem_raw, llr_base, out_base, results, harq_base, tab);
    return;
  }
  // Tensor memory for the check state of the last 32 base rows: 64 columns per group of four warps (3 resident CTAs of
  // six warps take 3 x 128 of the SM's 512 columns; smaller CTAs take 64 each). One warp allocates and later frees.
  const uint32_t tm_cols = (blockDim.x > 128) ? 2u * FEC_PACKED_TMEM_COLS_PER_WARPSLOT : FEC_PACKED_TMEM_COLS_PER_WARPSLOT;
  if (threadIdx.x < 32) {
    tmem_alloc((uint32_t)__cvta_generic_to_shared(&s_tmem_base), tm_cols);
  }
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
  const uint32_t tmem_base = s_tmem_base;
  fec_packed_cb(D, blockIdx.x, smem_raw, llr_base, out_base, results, harq_base, tab, ptab, tmem_base);
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  if (threadIdx.x < 32) {
    tmem_dealloc(tmem_base, tm_cols);
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
      off += fec_packed_row_words(t.deg[bg][m]);
    }
    pt.state_off[bg][M] = (uint16_t)off;
    const uint32_t row_tm = fec_packed_first_tmem_row(bg == 0);
    if (pt.state_off[bg][row_tm] != fec_packed_state_words(bg == 0)) {
      return false;
    }
    for (uint32_t m = row_tm; m < (uint32_t)M; ++m) {
      if (t.deg[bg][m] > 8) {
        return false; // tensor-memory rows hold two state words
      }
    }
    const int ne = t.row_start[bg][M];
    for (int zi = 0; zi < 51; ++zi) {
      const uint32_t Z = all_z[zi];
      if (Z % 2 != 0 || Z < 4) {
        continue;
      }
      const uint32_t H = Z / 2;
      for (int i = 0; i < ne; ++i) {
        const uint32_t   s = t.shift[bg * 51 + zi][i], col = t.col[bg][i];
        const uint32_t   q = s / H, r = s % H;
        fec_packed_edge& e = pt.edges[bg * 51 + zi][i];
        e.aN  = (col * H + r) * 2 + (q & 1);
        e.aW  = (col * H + r - H) * 2 + ((q + 1) & 1); // only used by threads t >= H - r, so t + r - H >= 0
        e.thr = H - r;
        e.pad = 0;
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
  uint32_t threads = ((max_Z / 2 + 31) / 32) * 32;
  threads          = threads < 32 ? 32 : (threads > FEC_PACKED_MAX_THREADS ? FEC_PACKED_MAX_THREADS : threads);
  // The generic per-check code only runs for odd lifting sizes and Z = 2 (all <= 15): its layout is sized for those.
  const uint32_t a = fec_generic_layout(max_Z < 15 ? max_Z : 15, any_bg1).total;
  const uint32_t b = (max_Z >= 4) ? fec_packed_layout(max_Z & ~1u, any_bg1).total : 0;
  const uint32_t smem = a > b ? a : b;
  // cudaFuncAttributeMaxDynamicSharedMemorySize is set ONCE, to the largest layout, in fec_engine_create: the attribute is
  // per-function global state and launches from several queues / threads would otherwise race on it.
  fec_cb_packed_kernel<<<n, threads, smem, st>>>(descs, llr, out, res, harq, tab, ptab);
  return cudaGetLastError() == cudaSuccess ? FEC_OK : FEC_ERR_CUDA;
}
