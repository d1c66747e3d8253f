// srsran-5g_b200/csrc/fec_kernel_packed.cuh -- the hot kernel: one CTA per codeblock, one thread per TWO lifted check
// nodes, 16-bit SIMD-in-register arithmetic on the native sm_100a half2 instructions.
//
// Number format. Every LLR-domain quantity x (|x| < 2048) lives in a 16-bit lane as sign-magnitude: bit 15 = sign, bits
// 14..0 = |x|. That bit pattern IS the IEEE fp16 number x * 2^-24 (subnormal for |x| < 1024, normal above, no gap), so
// HADD2 / HFMA2 are exact sign-magnitude adders, HMNMX2.XORSIGN is the min-sum check-node primitive
// (min(|a|,|b|) with sign(a)^sign(b)), HSET2.BF gives 1.0/0.0 select masks that feed HFMA2, negation is an XOR and |x| an
// AND. No tensor cores, no floating-point rounding anywhere (every intermediate is an integer below 2048).
//
// Per-codeblock working set (Z = 2H, H 32-bit "pair words" per column):
//   soft   nfull*H words in SHARED memory. Word w of column b holds variables {w, w+H} of the column as its two 16-bit
//          lanes, already in the sign-magnitude number format above (no conversion on load or store). Finite values are
//          <= 120 in magnitude; anything above 120 is the reference's +-127 "infinity" (sticky: it only grows).
//          With this interleave the cyclic shift s = q*H + r of an edge is ONE aligned 32-bit load at word (t + r) mod H
//          plus a lane swap when q + wrap is odd. The swap bit rides in bit 1 of the edge's byte address, where
//          PRMT.F4E (rotate by address bits 1:0) finds it: one PRMT after the load, one before the store.
//   state  per base row and thread: word0 = scaled min1 / min2 of both checks; then index of the min1 edge and the c2v
//          sign bits (1 more word for rows of degree <= 8, 2 otherwise). All-zero = "row not initialised" (c2v = 0).
//          The dense first rows keep it in shared memory; every later row keeps it in TENSOR MEMORY (see below).
//   edges  per edge {byte address of word (col*H + r) + swap bit, same for threads that wrap, H - r}: uniform per CTA.
//
// Bit-exactness with the reference's generic decoder is argued step by step in DESIGN.md ("packed kernel algebra").
#pragma once
#include "fec_device.cuh"
#include "fec_kernel_generic.cuh"
#include <cuda_fp16.h>
#include <string.h>
#include <algorithm>
#include <cstdlib>
#include <map>
#include <mutex>

#define FEC_PACKED_MAX_THREADS 192
// -DFEC_DEBUG_CHECKS: bounds checks on every shared-memory / tensor-memory access of the hot path (compute-sanitizer is closed
// on the B200 pool used here, so the checks are the kernel's own: a violation prints and traps, which fails the call).
#ifdef FEC_DEBUG_CHECKS
#include <cstdio>
#define FEC_CHECK(cond, what, v)                                                                                        \
  do {                                                                                                                  \
    if (!(cond)) {                                                                                                      \
      printf("FEC_CHECK failed: %s (%u) block %u thread %u\n", what, (unsigned)(v), blockIdx.x, threadIdx.x);           \
      __trap();                                                                                                         \
    }                                                                                                                   \
  } while (0)
#else
#define FEC_CHECK(cond, what, v) ((void)0)
#endif
#define FEC_PACKED_CTAS_PER_SM 3u // resident codeblocks per SM at the largest lifting size (__launch_bounds__ below)

// Edge (shift s = q * H + r into column col): thread t reads pair word (t + r) mod H of the column; its lanes come swapped
// when q + [t + r >= H] is odd. The wrap is ONE unsigned min: with u = 4 t + c0 (negative, i.e. huge, while t + r < H),
//   rel = min(u + c1, u)   is the byte offset of the pair word inside the column, plus 2 if the lanes come swapped,
// (VIADDMNMX.U32) and the byte address is rel + base. Bit 1 of the address is the swap bit PRMT.F4E consumes; the access
// itself uses the address with that bit cleared.
struct fec_packed_edge {
  uint32_t c0;   // 4 (r - H) + 2 ((q + 1) & 1)                      : offset of a thread that wraps
  uint32_t c1;   // 4 H + 2 (q & 1) - 2 ((q + 1) & 1)                : what a thread that does not wrap adds to it
  uint32_t base; // 4 col H (the kernel adds the shared-window address of the soft array while staging the records)
  uint32_t pad;
};

// Device-global tables for the packed kernel (built once per engine).
struct fec_packed_tables {
  fec_packed_edge edges[FEC_NOF_GRAPHS][FEC_MAX_EDGES]; // valid for graphs with even Z >= 4
  uint16_t        state_off[2][FEC_MAX_ROWS + 2];        // first state word of row r, in units of H words
  uint32_t        scale_mul;                             // scaled = (x * scale_mul + scale_add) >> 8 for x in 0..120
  uint32_t        scale_add;
};

struct fec_packed_smem_layout {
  uint32_t state, edges, prog, bitsw, red, soft, total;
};

__host__ __device__ inline uint32_t fec_packed_row_words(uint32_t d)
{
  return d <= 8 ? 2u : 3u;
}

// Where the check state lives. Only the first rows of a graph (BG1: 7, BG2: 3) keep it in shared memory; all later rows
// keep it in TENSOR MEMORY, used as a private per-thread scratchpad (tcgen05.alloc / ld / st in the 32x32b shape; no MMA is
// ever issued): thread t of warp w owns TMEM lane 32 (w % 4) + (t % 32) and reads back only what it wrote itself.
//   region A (64 columns per group of four warps): the last 32 base rows, two words each;
//   region B (16 columns per group of four warps): the 7 rows before them, two or three words each (<= 16 words).
// At BG1 / Z = 384 that takes 61 of the 76 KB of check state out of shared memory, which pays for the 16-bit soft-bit
// format (52 KB) while THREE codeblocks stay resident per SM (3 x 74 KB of shared memory, 3 x 160 of 512 TMEM columns).
#define FEC_PACKED_TMEM_A_ROWS 32u
#define FEC_PACKED_TMEM_B_ROWS 7u
#define FEC_PACKED_TMEM_A_COLS 64u // per warp slot (group of four warps)
#define FEC_PACKED_TMEM_B_COLS 16u

__host__ __device__ inline uint32_t fec_packed_row_a(int bg1) // first row of region A
{
  return (bg1 ? 46u : 42u) - FEC_PACKED_TMEM_A_ROWS;
}
__host__ __device__ inline uint32_t fec_packed_row_b(int bg1) // first row of region B = number of rows in shared memory
{
  return fec_packed_row_a(bg1) - FEC_PACKED_TMEM_B_ROWS;
}

// State words per thread kept in SHARED memory (rows before region B).
__host__ __device__ inline uint32_t fec_packed_state_words(int bg1)
{
  // BG1 rows 0..6: degrees 19,19,19,19,3,8,9 -> 3,3,3,3,2,2,3. BG2 rows 0..2: degrees 8,10,8 -> 2,3,2.
  return bg1 ? 19u : 7u;
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
  L.prog = o; // row program: one 16-byte record per base row (fec_packed_iteration)
  o += 48 * 16;
  L.bitsw = o;
  o += ((Kb * Z + 31) / 32 + 1) * 4;
  L.red = o;
  o += 64 * 4;
  o      = (o + 15) & ~15u;
  L.soft = o;
  o += (nfull * H * 4 + 15) & ~15u;
  L.total = o;
  return L;
}

// ---- FEC_PROFILE build: phase clocks accumulated over all codeblocks (thread 0 of each CTA) ----------------------------------
#ifdef FEC_PROFILE
#define FEC_PROF_PHASES 16
__device__ unsigned long long fec_prof_acc[FEC_PROF_PHASES];
#define FEC_PROF_MARK(k)                                                                                                        \
  do {                                                                                                                          \
    if (threadIdx.x == 0) {                                                                                                     \
      const long long now_ = clock64();                                                                                         \
      atomicAdd(&fec_prof_acc[k], (unsigned long long)(now_ - fec_prof_t));                                                     \
      fec_prof_t = now_;                                                                                                        \
    }                                                                                                                           \
  } while (0)
#else
#define FEC_PROF_MARK(k)
#endif

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
__device__ __forceinline__ uint32_t h2_sub(uint32_t a, uint32_t b)
{
  uint32_t r;
  asm("sub.rn.f16x2 %0, %1, %2;" : "=r"(r) : "r"(a), "r"(b));
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
__device__ __forceinline__ uint32_t h2_min(uint32_t a, uint32_t b)
{
  uint32_t r;
  asm("min.f16x2 %0, %1, %2;" : "=r"(r) : "r"(a), "r"(b));
  return r;
}
__device__ __forceinline__ uint32_t h2_max(uint32_t a, uint32_t b)
{
  uint32_t r;
  asm("max.f16x2 %0, %1, %2;" : "=r"(r) : "r"(a), "r"(b));
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
}
__device__ __forceinline__ void tmem_relinquish()
{
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
// Three consecutive columns (rows of degree 9 / 10): x2 + x1, one wait.
__device__ __forceinline__ void tmem_ld3(uint32_t taddr, uint32_t& a, uint32_t& b, uint32_t& c)
{
  asm volatile("tcgen05.ld.sync.aligned.32x32b.x2.b32 {%0, %1}, [%3];\n\t"
               "tcgen05.ld.sync.aligned.32x32b.x1.b32 {%2}, [%4];\n\t"
               "tcgen05.wait::ld.sync.aligned;"
               : "=r"(a), "=r"(b), "=r"(c)
               : "r"(taddr), "r"(taddr + 2u)
               : "memory");
}
__device__ __forceinline__ void tmem_st3(uint32_t taddr, uint32_t a, uint32_t b, uint32_t c)
{
  asm volatile("tcgen05.st.sync.aligned.32x32b.x2.b32 [%0], {%2, %3};\n\t"
               "tcgen05.st.sync.aligned.32x32b.x1.b32 [%1], {%4};" ::"r"(taddr),
               "r"(taddr + 2u), "r"(a), "r"(b), "r"(c)
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

// Shared-memory pair-word accesses by 32-bit shared-window address.
// "memory" clobber: ordinary C++ stores to shared memory surround these accesses and must not be moved across them.
__device__ __forceinline__ uint32_t lds_u32(uint32_t a)
{
  uint32_t r;
  asm volatile("ld.shared.b32 %0, [%1];" : "=r"(r) : "r"(a) : "memory");
  return r;
}
__device__ __forceinline__ uint4 lds_u128(uint32_t a)
{
  uint4 r;
  asm volatile("ld.shared.v4.b32 {%0, %1, %2, %3}, [%4];" : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w) : "r"(a) : "memory");
  return r;
}
__device__ __forceinline__ void sts_u32(uint32_t a, uint32_t v)
{
  asm volatile("st.shared.b32 [%0], %1;" ::"r"(a), "r"(v) : "memory");
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
// TM: the row's state words live in tensor memory at taddr (2 or 3 consecutive columns of this thread's lane).
template <int D, bool TM>
__device__ __forceinline__ void fec_packed_row(uint32_t st,    // !TM: shared-window address of this thread's first state word
                                               uint32_t h4,    //      distance in bytes between the state words of a row (4 H)
                                               uint32_t taddr, // TM: tensor-memory address of this thread's state words
                                               uint32_t ed,    // shared-window address of the row's D edge records
                                               uint32_t t,
                                               uint32_t scale_mul,
                                               uint32_t scale_add)
{
  constexpr int  NW   = (D <= 8) ? 2 : 3;
  const uint32_t t4   = t * 4;

  // ---- pre-phase: everything that does not depend on the soft bits (addresses, the previous c2v messages of this row).
  // It runs BEFORE the CTA barrier that orders this row after the previous one, so warps that finished the previous row
  // early do useful work instead of waiting, and the table / state load latency overlaps the other warps' tail.
  // The degree-19 rows hold 19 addresses and 19 soft words at once: their old messages are rebuilt one by one in phase A
  // instead (19 registers less at the barrier; the persistent kernel has none to spare and spilt 28 there).
#ifdef FEC_COLD_LATE
  constexpr bool COLD_LATE = (D > 10);
#else
  constexpr bool COLD_LATE = false;
#endif
  uint32_t       a0[D], cold[COLD_LATE ? 1 : D]; // a0: byte address of the pair word, bit 1 = swap bit
  uint32_t       oid, cw1, cw2, codl, com1;
  auto           old_c2v = [&](int e) -> uint32_t {
    const uint32_t sw  = (e < 10) ? cw1 : cw2;
    const uint32_t ism = h2_eq_bf(oid, FEC_H2(e));            // edge e was the min1 edge
    const uint32_t mag = h2_fma(ism, codl, com1);             // |old c2v|
    const uint32_t sgs = imad(sw, 1u << (15 - (e % 10)), 0u); // its sign bits at 15 / 31
    return mag ^ (sgs & H2_SIGN);                             // old c2v (one LOP3; the subtraction negates for free)
  };
  {
    uint32_t w0, w1, w2 = 0;
    if (TM) {
      FEC_CHECK((taddr & 0xffffu) + NW <= 512u && ((taddr >> 16) & 0x1fu) == 0, "row: tensor-memory address", taddr);
      // this thread's store of the previous tensor-memory row (one wait per iteration in front of the row loop would be enough
      // for correctness -- every row has columns of its own -- and was measured: 1.5 % slower)
      tmem_wait_st();
      if (NW == 3) {
        tmem_ld3(taddr, w0, w1, w2);
      } else {
        tmem_ld2(taddr, w0, w1);
      }
    } else {
      w0 = lds_u32(st);
      w1 = lds_u32(st + h4);
      if (NW == 3) {
        w2 = lds_u32(st + 2 * h4);
      }
    }
#pragma unroll
    for (int e = 0; e < D; ++e) {
      const uint4    E = lds_u128(ed + 16 * e);
      const uint32_t u = imad(t4, 1u, E.x);               // the adds on the FMA pipe: the ALU pipe is the busier one
      a0[e]            = imad(min(u + E.y, u), 1u, E.z);  // VIADDMNMX.U32: the cyclic wrap and the swap bit in one
    }
    const uint32_t om1 = __byte_perm(w0, 0, 0x4140), om2 = __byte_perm(w0, 0, 0x4342);
    const uint32_t odl = h2_sub(om2, om1); // m2 - m1 >= 0
    // state words: D <= 8: w1 = sign bits 0..7 | index << 8; D = 9, 10: w1 = sign bits 0..9, w2 = index;
    // D = 19: w1 = sign bits of edges 0..9, w2 = sign bits of edges 10..18 | index << 9 (all per 16-bit lane)
    oid = (D <= 8) ? __byte_perm(w1, 0, 0x4341) : ((D <= 10) ? w2 : ((w2 >> 9) & 0x001f001fu));
    cw1 = w1, cw2 = w2, codl = odl, com1 = om1;
    if (!COLD_LATE) {
#pragma unroll
      for (int e = 0; e < D; ++e) {
        cold[e] = old_c2v(e);
      }
    }
  }
  // Orders this row after the rows that wrote the columns it reads. (In the extension part of both base graphs every other
  // row shares no column with its predecessor and could skip the barrier; measured: 1.5 % SLOWER -- the six warps of a
  // codeblock drift apart and run different row bodies, which costs more in the instruction cache than the barrier does.)
  asm volatile("bar.sync 0;" ::: "memory");

  // ---- phase A: variable-to-check messages and the min1 / min2 / sign-product analysis -----------------------------------
  uint32_t S[D];
#pragma unroll
  for (int e = 0; e < D; ++e) {
    FEC_CHECK((a0[e] & 1u) == 0 && (a0[e] & ~2u) + 4 <= 233472u, "row: soft address", a0[e]);
    const uint32_t x = lds_u32(a0[e] & ~2u);
    S[e]             = prmt_f4e(x, x, a0[e]); // address bit 1 set: the two lanes come swapped
  }
  uint32_t v[D];
  uint32_t m1 = H2_P120, m2 = H2_P120;
#pragma unroll
  for (int e = 0; e < D; ++e) {
    const uint32_t vraw = h2_sub(S[e], COLD_LATE ? old_c2v(e) : cold[e]);            // soft - c2v
    const uint32_t y    = h2_add(h2_fma(S[e], H2_F4352, H2_F96), H2_FM96);           // nonzero, sign of S, iff |S| > 120
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
  const uint32_t s1  = __byte_perm(imad(a1, scale_mul, scale_add), 0, 0x4341); // byte 1 of each lane = (.) >> 8
  const uint32_t s2  = __byte_perm(imad(m2 & H2_ABS, scale_mul, scale_add), 0, 0x4341);
  const uint32_t M1s = s1 | sp;
  const uint32_t dls = h2_sub(s2, s1) | sp;

  // ---- phase B: new check-to-variable messages, soft-bit update, write-back ------------------------------------------------
  uint32_t nid = 0, sg1 = 0, sg2 = 0;
#pragma unroll
  for (int e = 0; e < D; ++e) {
    const uint32_t vv = v[e];
    const uint32_t eq = h2_eq_bf(h2_abs(vv), a1);        // this edge holds min1 -> it gets min2
    const uint32_t mg = h2_fma(eq, dls, M1s);            // sign product * scaled magnitude
    const uint32_t c  = mg ^ (vv & H2_SIGN);             // new c2v
    nid               = h2_fma(eq, FEC_H2(e), nid);      // index of the min1 edge (any value if min1 == min2)
    const uint32_t tt = h2_add(c, vv);                   // c2v + v2c: above 120 in magnitude it IS the "infinity" from here on
    // sign(v2c) of the two lanes -> bit (e % 10) of the sign word (gathering them on the FMA pipe instead -- HSET2 + HFMA2 into
    // a 1024.0-biased accumulator -- was measured: no difference)
    const uint32_t sh = FEC_SHR(vv, 15 - (e % 10));
    if (e < 10) {
      sg1 |= sh & (0x00010001u << (e % 10));
    } else {
      sg2 |= sh & (0x00010001u << (e % 10));
    }
    sts_u32(a0[e] & ~2u, prmt_f4e(tt, tt, a0[e]));
  }

  // ---- new state: c2v sign = sign product ^ sign(v2c) ----------------------------------------------------------------------
  const uint32_t pm  = prmt(m1, 0u, 0xbb99); // 0xffff in the lanes whose sign product is negative
  const uint32_t nw0 = __byte_perm(s1, s2, 0x6420);
  uint32_t       nw1, nw2 = 0;
  if (D <= 8) {
    nw1 = ((sg1 ^ pm) & 0x00ff00ffu) | (nid << 8);
  } else if (D <= 10) {
    nw1 = (sg1 ^ pm) & 0x03ff03ffu;
    nw2 = nid;
  } else {
    nw1 = (sg1 ^ pm) & 0x03ff03ffu;
    nw2 = ((sg2 ^ pm) & 0x01ff01ffu) | (nid << 9);
  }
  if (TM) {
    if (NW == 3) {
      tmem_st3(taddr, nw0, nw1, nw2);
    } else {
      tmem_st2(taddr, nw0, nw1);
    }
  } else {
    sts_u32(st, nw0);
    sts_u32(st + h4, nw1);
    if (NW == 3) {
      sts_u32(st + 2 * h4, nw2);
    }
  }
}

// ---- the row schedule of a base graph -------------------------------------------------------------------------------------
// One word per base row in CONSTANT memory (filled by the engine at creation, read through the uniform datapath):
//   bits 0-8 first edge of the row, bits 9-16 state offset (shared memory: in units of H words; region B: column inside the
//   region; region A: 2 (row - first row of A)), bits 17-20 which instantiation of the row body runs it (below).
// (A fully unrolled schedule -- one straight run of 46 row bodies, 150 KB of code -- was measured: 3 % faster when the three
// codeblocks of an SM run in lockstep, 30 % slower when they do not, as with early stop, because their instruction streams
// then compete for the 32 KB L1.5 instruction cache. The loop keeps the hot code at 13 row bodies.)
__constant__ uint32_t fec_c_rowinfo[2][48];

enum { // instantiation codes: tensor-memory rows of degree 3..10, then the shared-memory rows
  FEC_ROW_T3 = 0, FEC_ROW_T4, FEC_ROW_T5, FEC_ROW_T6, FEC_ROW_T7, FEC_ROW_T8, FEC_ROW_T9, FEC_ROW_T10,
  FEC_ROW_S3, FEC_ROW_S8, FEC_ROW_S9, FEC_ROW_S10, FEC_ROW_S19, FEC_ROW_NONE
};

// One decoder iteration over the layers in use.
__device__ __forceinline__ void fec_packed_iteration(uint32_t prog, // shared-window address of the row program
                                                     uint32_t H,
                                                     uint32_t nlayers,
                                                     uint32_t smul,
                                                     uint32_t sadd,
                                                     uint32_t tm_b, // this warp's address of region B / A
                                                     uint32_t tm_a)
{
  const uint32_t tid = threadIdx.x;
  if ((tid & ~31u) >= H) {
    for (uint32_t l = 0; l < nlayers; ++l) {
      asm volatile("bar.sync 0;" ::: "memory"); // warps without work still take part in the row barriers
    }
    return;
  }
  // Lanes past H in the last warp repeat the work of thread H-1 (identical loads, identical stores; their own, identical
  // copy of the tensor-memory state) so that whole warps reach the barrier inside the row function.
  const uint32_t tq = min(tid, H - 1);
  const uint32_t h4 = 4 * H;
  const uint32_t dab = tm_a - tm_b;
  // Row program record: x = address of the row's edge records, y = where the state is (shared memory: shared-window address
  // of thread 0's first word; tensor memory: column inside the region), z = instantiation, w = 1 for region A.
  // Consecutive rows that run the same body (BG1: 4 x degree 19, 6 x degree 6, 9 x degree 5, ...) stay inside it: the next
  // record is fetched in the body's own loop and the jump table is only consulted when the instantiation changes
  // (+0.8 % forced, +1.7 % with early stop over one dispatch per row).
#define FEC_ROW_CASE(code, D, TM)                                                                                              \
  case code:                                                                                                                     \
    do {                                                                                                                         \
      fec_packed_row<D, TM>(imad(tq, 4u, P.y), h4, imad(P.w, dab, tm_b + P.y), P.x, tq, smul, sadd);                              \
      if (++l >= nlayers) {                                                                                                      \
        return;                                                                                                                  \
      }                                                                                                                          \
      P = lds_u128(prog + 16 * l);                                                                                               \
    } while (P.z == code);                                                                                                       \
    break;
  uint32_t l = 0;
  if (nlayers == 0) {
    return;
  }
  uint4 P = lds_u128(prog);
  for (;;) {
    __builtin_assume(P.z <= FEC_ROW_S19);
    switch (P.z) {
      FEC_ROW_CASE(FEC_ROW_T3, 3, true)
      FEC_ROW_CASE(FEC_ROW_T4, 4, true)
      FEC_ROW_CASE(FEC_ROW_T5, 5, true)
      FEC_ROW_CASE(FEC_ROW_T6, 6, true)
      FEC_ROW_CASE(FEC_ROW_T7, 7, true)
      FEC_ROW_CASE(FEC_ROW_T8, 8, true)
      FEC_ROW_CASE(FEC_ROW_T9, 9, true)
      FEC_ROW_CASE(FEC_ROW_T10, 10, true)
      FEC_ROW_CASE(FEC_ROW_S3, 3, false)
      FEC_ROW_CASE(FEC_ROW_S8, 8, false)
      FEC_ROW_CASE(FEC_ROW_S9, 9, false)
      FEC_ROW_CASE(FEC_ROW_S10, 10, false)
      FEC_ROW_CASE(FEC_ROW_S19, 19, false)
      default:
        return;
    }
  }
#undef FEC_ROW_CASE
}

// Index, in 16-bit units, of variable j (0 <= j < Z) of column col in the pair-word soft layout.
__device__ __forceinline__ uint32_t fec_packed_soft_idx(uint32_t col, uint32_t j, uint32_t H)
{
  const uint32_t lane = (j >= H) ? 1u : 0u;
  return (col * H + (j - lane * H)) * 2 + lane;
}

// An integer LLR in the 16-bit sign-magnitude number format.
__device__ __forceinline__ uint16_t fec_sm16(int v)
{
  return (uint16_t)((v < 0) ? (0x8000 | (-v)) : v);
}

// Loader target: clamp to +-64 (ldpc_decoder_impl.cpp:168-169) and store in the pair-word layout.
struct fec_soft_store_packed {
  uint16_t* soft16;
  uint32_t  Z, H, invZ;
  __device__ __forceinline__ void operator()(uint32_t o, int acc) const
  {
    raw(o, fec_clampi(acc, -64, 64));
  }
  // Without the clamp: the partial last node of a decoder input that is not a multiple of Z (ldpc_decoder_impl.cpp:186-193).
  __device__ __forceinline__ void raw(uint32_t o, int acc) const
  {
    const uint32_t op  = o + 2 * Z;
    const uint32_t col = __umulhi(op, invZ);
    FEC_CHECK(col < 68u && fec_packed_soft_idx(col, op - col * Z, H) < 68u * Z, "soft store: index", op);
    soft16[fec_packed_soft_idx(col, op - col * Z, H)] = fec_sm16(acc);
  }
};

// The decoder's soft bits from the soft buffer the dematcher has just written (natural int8 layout in HBM, L2-hot): the two
// punctured columns are zero, buffer positions [0, npos) are clamped to +-64 and converted; nothing beyond the columns the
// layers in use touch is needed. Four positions per thread and step (npos is a multiple of Z, Z is even: npos % 4 == 0 except
// for Z = 2 (mod 4) with an odd number of columns -- the last two positions are then handled one by one).
// The loads bypass L1 (ld.global.cg): other threads of the CTA wrote these bytes earlier in this kernel.
__device__ __forceinline__ void fec_soft_fill_from_harq(const int8_t* harq, uint16_t* soft16, uint32_t npos, uint32_t Z, uint32_t H,
                                                        uint32_t invZ)
{
  uint32_t* soft32 = reinterpret_cast<uint32_t*>(soft16);
  for (uint32_t i = threadIdx.x; i < Z; i += blockDim.x) {
    soft32[i] = 0; // columns 0 and 1: H pair words each
  }
  const fec_soft_store_packed store{soft16, Z, H, invZ};
  const uint32_t*             h4 = reinterpret_cast<const uint32_t*>(harq);
  const uint32_t              ng = npos / 4;
  for (uint32_t g = threadIdx.x; g < ng; g += blockDim.x) {
    const uint32_t w   = __ldcg(h4 + g);
    const uint32_t o0  = g * 4;
    const uint32_t op0 = o0 + 2 * Z;
    const uint32_t col = __umulhi(op0, invZ);
    const uint32_t j0  = op0 - col * Z;
    if (j0 + 3 < H || (j0 >= H && j0 + 3 < Z)) {
      const uint32_t lane = (j0 >= H) ? 1u : 0u;
      uint16_t*      dst  = soft16 + (col * H + (j0 - lane * H)) * 2 + lane;
#pragma unroll
      for (int b = 0; b < 4; ++b) {
        dst[2 * b] = fec_sm16(fec_clampi((int)(int8_t)(w >> (8 * b)), -64, 64));
      }
    } else {
#pragma unroll
      for (int b = 0; b < 4; ++b) {
        store(o0 + b, (int)(int8_t)(w >> (8 * b)));
      }
    }
  }
  for (uint32_t o = ng * 4 + threadIdx.x; o < npos; o += blockDim.x) {
    store(o, (int)__ldcg(harq + o));
  }
}

// Hard decision of the first K soft bits (bit = soft <= 0) into bitsw, big-endian bit order; true if no soft bit is zero.
__device__ __forceinline__ bool fec_hard_bits_packed(const uint16_t* soft16, uint32_t* bitsw, uint32_t K, uint32_t Z, uint32_t H,
                                                     uint32_t invZ)
{
  const uint32_t lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nwarps = blockDim.x >> 5;
  int            zero = 0;
  if ((H & 31u) == 0) {
    // One column per step: thread t holds variables t and t+H of the column as one f16x2 word. One HSETP2 (x <= 0 on both
    // lanes; -0 and +0 both count) feeds two ballots of 32 consecutive bits each; lane c of the warp keeps the two words of
    // column c and stores them at the end. A zero magnitude anywhere is found with a running f16x2 minimum of |x|.
    const uint32_t* soft32 = reinterpret_cast<const uint32_t*>(soft16);
    const uint32_t  Kb     = K / Z; // 22 or 10
    uint32_t        zacc   = 0x7c007c00u, keep0 = 0, keep1 = 0;
    if (threadIdx.x < H) {
      // (the empty asm keeps the compiler from computing the 22 addresses once, outside the iteration loop, and holding
      // them in registers through every row body)
      uint32_t tix = threadIdx.x;
      asm volatile("" : "+r"(tix));
      uint32_t hw[22];
#pragma unroll
      for (uint32_t c = 0; c < 22; ++c) {
        hw[c] = (c < Kb) ? soft32[c * H + tix] : 0x00010001u;
      }
#pragma unroll
      for (uint32_t c = 0; c < 22; ++c) {
        if (c < Kb) {
          uint32_t b0, b1;
          asm volatile("{\n\t"
                       ".reg .pred p, q;\n\t"
                       "setp.le.f16x2 p|q, %2, %3;\n\t"
                       "vote.sync.ballot.b32 %0, p, 0xffffffff;\n\t"
                       "vote.sync.ballot.b32 %1, q, 0xffffffff;\n\t"
                       "}"
                       : "=r"(b0), "=r"(b1)
                       : "r"(hw[c]), "r"(0u));
          zacc = h2_min(zacc, h2_abs(hw[c]));
          if (lane == c) {
            keep0 = b0, keep1 = b1;
          }
        }
      }
      if (lane < Kb) {
        bitsw[(lane * Z + warp * 32) >> 5]     = __brev(keep0);
        bitsw[(lane * Z + H + warp * 32) >> 5] = __brev(keep1);
      }
    }
    zero = ((zacc & 0x7fffu) == 0) | ((zacc & 0x7fff0000u) == 0);
    return __syncthreads_or(zero) == 0;
  }
  const uint32_t nw = (K + 31) / 32;
  for (uint32_t w = warp; w < nw; w += nwarps) {
    const uint32_t i = w * 32 + lane;
    uint32_t       b = 1; // positive
    if (i < K) {
      const uint32_t col = __umulhi(i, invZ);
      b                  = soft16[fec_packed_soft_idx(col, i - col * Z, H)];
    }
    zero |= ((b & 0x7fffu) == 0);
    const uint32_t bal = __ballot_sync(0xffffffffu, ((b & 0x7fffu) == 0) | ((b & 0x8000u) != 0));
    if (lane == 0) {
      bitsw[w] = __brev(bal);
    }
  }
  return __syncthreads_or(zero) == 0;
}

// CRC remainder == 0 test of the first T bits of bitsw (crc_calculator_generic_impl.cpp:111-133 computes the same
// remainder bytewise). Byte i of the message contributes byte_i(x) * x^(8 (nb-1-i)) mod g; x8[j] = x^(8 j) mod g comes from
// the engine's table in HBM (4 KB per polynomial, L1 / L2 resident; shared memory has no room for it beside the 16-bit
// soft bits). Up to six bytes per thread, eight multiply-and-add steps each, one barrier.
// (Round 1 kept this out of line because an inlined build decoded wrongly. That no longer reproduces: the suspect, a byte
// load in inline asm without a "memory" clobber next to plain C++ shared-memory stores, is gone, every asm access here
// clobbers memory, and the inlined build passes the whole parity suite. Inlined, the kernel is 2 % faster.)
__device__ __forceinline__ bool fec_crc_ok_packed(const uint32_t* bitsw, uint32_t T, const uint32_t* __restrict__ x8, uint32_t poly,
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
      a[k]             = v ? __ldg(x8 + (nb - 1 - i)) : 0u;
      w[k]             = v ? (bitsw[i >> 2] >> (24 - 8 * (i & 3))) : 0u;
    }
    // bit s of the byte contributes x^s * a = a << s: one AND, one multiply by the isolated bit (a << s or 0; a < 2^24 and
    // s <= 7, so no bit is lost) and one XOR per step. The products are NOT reduced mod g here: reduction is linear, so the
    // seven bits above the remainder are folded back once, after the sums of all threads have been added up.
#pragma unroll
    for (int s = 0; s < 8; ++s) {
#pragma unroll
      for (int k = 0; k < 6; ++k) {
        acc ^= (w[k] & (1u << s)) * a[k];
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
#pragma unroll
  for (int b = 6; b >= 0; --b) { // degree < crc_len + 7 -> degree < crc_len
    rem = (rem & (top << b)) ? (rem ^ (poly << b)) : rem;
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
//     used soft-bit area) in RATE-MATCHED order, i.e. the bit de-interleaver of TS 38.212 5.4.2.2 is undone while
//     staging. The staged position of LLR p is p + d, d in {d1, d2} chosen so that a staged 32-bit word lines up with a
//     32-bit word of the soft buffer inside either data run of a non-wrapping transmission;
//  2. the zero tail [zero_from, N) of a first transmission is stored without looking at anything;
//  3. groups of four soft-buffer bytes inside a data run take one 32-bit shared load and one 32-bit HBM store (plus a 32-bit
//     load and four saturating sums when combining);
//  4. every other group (region boundaries, fillers, kept parts of the buffer, wrapping transmissions) takes the general
//     path: still four bytes per thread, per-byte rules.
// The dematcher writes the soft buffer (HBM) only; the decoder's own 16-bit copy is filled from it afterwards
// (fec_soft_fill_from_harq: the lines are L2-hot, and only the columns of the layers in use are needed).
// Inputs that do not fit the staging area (E > 52 KB at Z = 384: heavy repetition) are read from global memory per byte.
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
                                                       int8_t* __restrict__    harq)
{
  uint32_t       my_sig  = 0;
  const uint32_t ngroups = G.N / 4; // Z is even and N = 66Z or 50Z, so N % 4 == 0
  uint32_t*      harq4   = reinterpret_cast<uint32_t*>(harq);

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
  // Repetition (a first transmission from the start of the buffer that wraps around it, E > Ncb - fillers): position q of
  // the circular buffer takes the LLRs p = q, q + Lc, q + 2 Lc, ... < E in that order -- the first is copied, the others are
  // added with the saturating sum. Groups wholly inside [0, info) or [sys, Ncb) do that on words.
  const bool wfast = G.wrapped && G.newd && S.staged && (G.k0c == 0) && (G.kstart == 0);
  if (wfast) {
    hg0[0] = 0;
    hg1[0] = G.info / 4;
    hg0[1] = (G.sys + 3) / 4;
    hg1[1] = G.Ncb / 4;
  }
  const uint32_t gz0 = (G.zero_from + 3) / 4; // groups from here on are all zero

  // (2) zero tail
  for (uint32_t g = gz0 + threadIdx.x; g < ngroups; g += blockDim.x) {
    harq4[g] = 0;
  }

  // (3w) groups inside a data run, repetition
  if (wfast) {
    // four staged LLRs p .. p+3 (rate-matched order) as one word; the staging shift changes at `split`
    auto ldw = [&](uint32_t p) -> uint32_t {
      if (p + 3 < S.split || p >= S.split) {
        const uint32_t x = p + ((p < S.split) ? S.d1 : S.d2), k8 = (x & 3u) * 8u;
        const uint32_t lo = *reinterpret_cast<const uint32_t*>(S.base + (x & ~3u));
        return (k8 == 0) ? lo : __funnelshift_r(lo, *reinterpret_cast<const uint32_t*>(S.base + (x & ~3u) + 4), k8);
      }
      int val[4];
#pragma unroll
      for (int b = 0; b < 4; ++b) {
        val[b] = fec_dematch_fetch(G, S, p + b);
      }
      return fec_pack4(val);
    };
#pragma unroll
    for (int r = 0; r < 2; ++r) {
      for (uint32_t g = hg0[r] + threadIdx.x; g < hg1[r]; g += blockDim.x) {
        const uint32_t o0 = g * 4;
        const uint32_t q0 = (r == 0) ? o0 : (o0 - G.fillers);
        uint32_t       w  = ldw(q0); // E > Lc: the first pass covers every position
        for (uint32_t p = q0 + G.Lc; p < G.E; p += G.Lc) {
          int val[4];
          if (p + 3 < G.E) {
            const uint32_t x = ldw(p);
#pragma unroll
            for (int b = 0; b < 4; ++b) {
              val[b] = fec_sat_add((int)(int8_t)(w >> (8 * b)), (int)(int8_t)(x >> (8 * b)));
            }
          } else { // the transmission ends inside this group
#pragma unroll
            for (int b = 0; b < 4; ++b) {
              const int acc = (int)(int8_t)(w >> (8 * b));
              val[b]        = (p + b < G.E) ? fec_sat_add(acc, fec_dematch_fetch(G, S, p + b)) : acc;
            }
          }
          w = fec_pack4(val);
        }
        harq4[g] = w;
        my_sig   = max(my_sig, fec_last_sig(o0, w));
      }
    }
  }

  // (3) groups inside a data run
#pragma unroll
  for (int r = 0; r < 2 && !wfast; ++r) {
    for (uint32_t g = hg0[r] + threadIdx.x; g < hg1[r]; g += blockDim.x) {
      const uint32_t o0 = g * 4;
      // (aligned by construction after the byte-wise staging, whose shifts d1 / d2 exist for that; the transposed staging
      // has no shifts and the word is put together from two)
      const uint32_t x = o0 - hoff[r], k8 = (x & 3u) * 8u;
      uint32_t       w = *reinterpret_cast<const uint32_t*>(S.base + (x & ~3u));
      if (k8 != 0) {
        w = __funnelshift_r(w, *reinterpret_cast<const uint32_t*>(S.base + (x & ~3u) + 4), k8);
      }
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
  }
  return my_sig;
}

// ---- staging by transposition -------------------------------------------------------------------------------------------
// The bit de-interleaver (TS 38.212 5.4.2.2) undone four symbols at a time: the 4 QM input bytes of symbols i0 .. i0+3 are
// loaded as QM words, and for every bit position j the four bytes {s = i QM + j} are gathered with three PRMTs into one
// word that lands at staged byte j Kp + i0. Needs Kp % 4 == 0, no staging shifts (d1 = d2 = 0) and an 8-byte aligned
// input; 1.3 instructions per LLR instead of the 9 of the byte-wise scatter.
template <uint32_t QM>
__device__ __forceinline__ void fec_stage_transposed(const int8_t* __restrict__ in, int8_t* stage, uint32_t Kp)
{
  constexpr uint32_t U   = (QM <= 4) ? 4u : 2u; // units in flight per thread: 32 bytes of loads (QM 2), 64 (QM 4, 8), 48 (QM 6)
  uint32_t*          st4 = reinterpret_cast<uint32_t*>(stage);
  const uint32_t     ng = Kp / 4, kw = Kp / 4;
  for (uint32_t g0 = threadIdx.x; g0 < ng; g0 += U * blockDim.x) {
    // (every element is defined on every path: in the persistent kernel a register that is only written under a condition
    // counts as live around the whole codeblock loop)
    uint32_t W[U][QM];
#pragma unroll
    for (uint32_t u = 0; u < U; ++u) {
      const uint32_t g = g0 + u * blockDim.x;
#pragma unroll
      for (uint32_t k = 0; k < QM; ++k) {
        W[u][k] = 0;
      }
      if (g < ng) {
        if constexpr (QM % 4 == 0) {
          const uint4* src = reinterpret_cast<const uint4*>(in) + (size_t)g * (QM / 4);
#pragma unroll
          for (uint32_t k = 0; k < QM / 4; ++k) {
            const uint4 v = __ldg(src + k);
            W[u][4 * k] = v.x, W[u][4 * k + 1] = v.y, W[u][4 * k + 2] = v.z, W[u][4 * k + 3] = v.w;
          }
        } else {
          const uint2* src = reinterpret_cast<const uint2*>(in) + (size_t)g * (QM / 2);
#pragma unroll
          for (uint32_t k = 0; k < QM / 2; ++k) {
            const uint2 v = __ldg(src + k);
            W[u][2 * k] = v.x, W[u][2 * k + 1] = v.y;
          }
        }
      }
    }
#pragma unroll
    for (uint32_t u = 0; u < U; ++u) {
      const uint32_t g = g0 + u * blockDim.x;
      if (g < ng) {
#pragma unroll
        for (uint32_t j = 0; j < QM; ++j) {
          const uint32_t s0 = j, s1 = QM + j, s2 = 2 * QM + j, s3 = 3 * QM + j; // input byte of symbol i0 + 0..3
          const uint32_t lo = prmt(W[u][s0 / 4], W[u][s1 / 4], (s0 % 4) | ((4 + s1 % 4) << 4));
          const uint32_t hi = prmt(W[u][s2 / 4], W[u][s3 / 4], (s2 % 4) | ((4 + s3 % 4) << 4));
          st4[j * kw + g]   = prmt(lo, hi, 0x5410);
        }
      }
    }
  }
}

// Four int8 values (one word) -> two f16x2 words in the 16-bit sign-magnitude format, clamped to +-64, without integer
// arithmetic: byte b becomes the half-precision number 1152 + b (bits 0x6400 | (b ^ 0x80)); times 2^-24 (bits 0x0001) minus
// 1152 * 2^-24 (bits 0x0480) leaves b * 2^-24, whose bit pattern is |b| with the sign bit of b. All exact.
__device__ __forceinline__ void fec_cvt4_sm16(uint32_t w, uint32_t& v01, uint32_t& v23)
{
  const uint32_t x = w ^ 0x80808080u;
  v01 = h2_fma(prmt(x, 0x64646464u, 0x4140), 0x00010001u, 0x84808480u);
  v23 = h2_fma(prmt(x, 0x64646464u, 0x4342), 0x00010001u, 0x84808480u);
  v01 = h2_min_xs(v01, 0x00400040u); // min(|v|, 64) with the sign of v
  v23 = h2_min_xs(v23, 0x00400040u);
}

// ---- the direct path: first transmission from the start of the buffer (new data, rv 0, no wrap-around), Z % 8 == 0 --------
// The soft buffer in HBM and the decoder's 16-bit copy in shared memory are written in the same pass: a thread takes the
// 4-byte groups at positions o and o + H of one column (two 32-bit loads from the staged LLRs, two 32-bit HBM stores) and
// stores the four pair words {o + k, o + H + k} with one 128-bit shared store. The staged LLRs live in the same shared
// memory the pair words go to: pair words of column c occupy bytes [2 c Z, 2 (c + 1) Z) and the staged LLRs of that column
// lie below byte (c - 1) Z + 8, so the columns are processed from the top in chunks [c0, c1) with 2 c0 Z >= (c1 - 2) Z + 8
// (nothing a chunk writes is still to be read by it or by a later chunk) and a barrier between chunks: five chunks for 64
// columns. Columns the decoder cannot reach (beyond cbl_max) only get their zeros in HBM.
// Groups that are not entirely inside a data run (the fillers, the run ends, the start of the zero tail) take the per-byte
// rules of fec_dematch_byte.
__device__ __forceinline__ uint32_t fec_dematch_direct(const fec_dematch_geom& G,
                                                       const fec_dematch_src&  S,
                                                       int8_t* __restrict__    harq,
                                                       uint16_t*               soft16,
                                                       uint32_t                K)
{
  const uint32_t Z = G.Z, H = Z / 2, upc = H / 4; // units (pairs of groups) per column
  const uint32_t inv_upc = 0xffffffffu / upc + 1; // unused when upc == 1 (Z = 8)
  uint32_t*      harq4   = reinterpret_cast<uint32_t*>(harq);
  uint4*         soft128 = reinterpret_cast<uint4*>(soft16);
  const uint32_t run0_hi = min(G.E, G.info);                   // data run before the filler gap: positions [0, run0_hi)
  const uint32_t run1_hi = (G.E > G.info) ? G.E + G.fillers : 0u; // after it: [sys, run1_hi), read at position - fillers
  uint32_t       cbl_max = max(G.zero_from + 2 * Z, K + 4 * Z);
  cbl_max                = min(((cbl_max + Z - 1) / Z) * Z, G.N + 2 * Z);
  const uint32_t cproc   = cbl_max / Z; // columns [2, cproc) get pair words
  uint32_t       sig_o = 0, sig_w = 0;  // this thread's highest group with a non-zero value

  // HBM zeros of the columns the decoder cannot reach
  {
    uint2*         h8 = reinterpret_cast<uint2*>(harq);
    const uint32_t n8 = G.N / 8;
    for (uint32_t i = (cproc - 2) * Z / 8 + threadIdx.x; i < n8; i += blockDim.x) {
      h8[i] = make_uint2(0, 0);
    }
  }

  auto group_value = [&](uint32_t o0) -> uint32_t {
    if (o0 + 4 <= run0_hi) {
      return *reinterpret_cast<const uint32_t*>(S.base + o0);
    }
    if (o0 >= G.sys && o0 + 4 <= run1_hi) {
      const uint32_t x = o0 - G.fillers, k8 = (x & 3u) * 8u; // k8 is the same for every group of the codeblock
      const uint32_t lo = *reinterpret_cast<const uint32_t*>(S.base + (x & ~3u));
      if (k8 == 0) {
        return lo;
      }
      return __funnelshift_r(lo, *reinterpret_cast<const uint32_t*>(S.base + (x & ~3u) + 4), k8);
    }
    if (o0 >= G.zero_from) {
      return 0u;
    }
    const uint32_t old = G.hzero ? 0u : harq4[o0 / 4];
    int            val[4];
#pragma unroll
    for (int b = 0; b < 4; ++b) {
      val[b] = fec_dematch_byte(G, S, o0 + b, (int)(int8_t)(old >> (8 * b)));
    }
    return fec_pack4(val);
  };

  for (uint32_t c1 = cproc; c1 > 2;) {
    uint32_t c0 = ((c1 - 2) * Z + 8 + 2 * Z - 1) / (2 * Z);
    c0          = max(c0, 2u);
    const uint32_t nu = (c1 - c0) * upc;
    for (uint32_t u = threadIdx.x; u < nu; u += blockDim.x) {
      const uint32_t cc  = (upc == 1) ? u : __umulhi(u, inv_upc);
      const uint32_t k   = u - cc * upc;
      const uint32_t col = c0 + cc;
      const uint32_t olo = (col - 2) * Z + 4 * k, ohi = olo + H;
      const uint32_t wlo = group_value(olo), whi = group_value(ohi);
      harq4[olo / 4] = wlo;
      harq4[ohi / 4] = whi;
      if (whi != 0 && ohi >= sig_o) {
        sig_o = ohi, sig_w = whi;
      } else if (wlo != 0 && olo >= sig_o) {
        sig_o = olo, sig_w = wlo;
      }
      uint32_t l01, l23, h01, h23;
      fec_cvt4_sm16(wlo, l01, l23);
      fec_cvt4_sm16(whi, h01, h23);
      FEC_CHECK(col * upc + k < 68u * upc, "direct dematch: pair word index", col * upc + k);
      soft128[col * upc + k] = make_uint4(prmt(l01, h01, 0x5410), prmt(l01, h01, 0x7632), prmt(l23, h23, 0x5410),
                                          prmt(l23, h23, 0x7632));
    }
    __syncthreads();
    c1 = c0;
  }
  return fec_last_sig(sig_o, sig_w);
}

// Returns true if the decoder's soft bits (columns 2 .. of the pair-word array) have been written as well (direct path).
__device__ __forceinline__ bool fec_dematch_packed(const fec_cb_desc& D,
                                                   const int8_t* __restrict__ in,
                                                   int8_t* __restrict__ harq,
                                                   int8_t*   stage,
                                                   uint32_t  stage_cap,
                                                   uint32_t* n_sig_smem,
                                                   uint32_t  K
#ifdef FEC_PROFILE
                                                   ,
                                                   long long& fec_prof_t
#endif
)
{
  const fec_dematch_geom G = fec_dematch_geometry(D);

  fec_dematch_src S;
  S.base   = in;
  S.staged = false;
  S.split  = (G.k0c < G.info) ? min(G.E, G.info - G.k0c) : 0u;
  S.d1     = G.k0c & 3u;
  S.d2     = S.d1 + (G.fillers & 3u);
  const bool can_transpose = (G.E + 8 <= stage_cap) && (G.Kp % 4 == 0) && ((reinterpret_cast<uintptr_t>(in) & 15u) == 0) &&
                             (G.qm == 2 || G.qm == 4 || G.qm == 6 || G.qm == 8);
  // direct path (fec_dematch_direct): also fills the decoder's soft bits. It reads the run behind the filler gap with
  // unaligned word loads when the number of fillers is not a multiple of four, so it never needs the staging shifts.
  const bool direct = can_transpose && G.newd && !G.wrapped && (D.k0 == 0) && (G.Z % 8 == 0) && (D.mode == FEC_MODE_CHAIN);
  if (can_transpose) {
    S.d1 = S.d2 = 0; // no staging shifts: the readers put unaligned words together themselves
  }
  const bool transposed = can_transpose;
  if (transposed) {
    switch (G.qm) {
      case 2:
        fec_stage_transposed<2>(in, stage, G.Kp);
        break;
      case 4:
        fec_stage_transposed<4>(in, stage, G.Kp);
        break;
      case 6:
        fec_stage_transposed<6>(in, stage, G.Kp);
        break;
      default:
        fec_stage_transposed<8>(in, stage, G.Kp);
        break;
    }
    __syncthreads();
    S.base   = stage;
    S.staged = true;
  } else if (G.E + 8 <= stage_cap) {
    // (1) stage and de-interleave: input byte s = i * Qm + j is rate-matched LLR p = j * (E / Qm) + i.
    const uint32_t inv_qm = 0xffffffffu / G.qm + 1;
    auto           put    = [&](uint32_t s, uint32_t byte) {
      const uint32_t i = (G.qm == 1) ? s : __umulhi(s, inv_qm);
      const uint32_t p = (s - i * G.qm) * G.Kp + i;
      FEC_CHECK(fec_dematch_stage_pos(S, p) < stage_cap, "dematch: staging position", p);
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
  FEC_PROF_MARK(1);

  const uint32_t my_sig = direct ? fec_dematch_direct(G, S, harq, reinterpret_cast<uint16_t*>(stage), K)
                                 : fec_dematch_groups(G, S, harq);
  if (my_sig != 0) {
    atomicMax(n_sig_smem, my_sig);
  }
  return direct;
}


__device__ __forceinline__ bool fec_packed_cb_supported(const fec_cb_desc& D)
{
  // Z must split into 2 interleaved lanes: every even lifting size from 4, all three modes. Odd Z (<= 15) and Z = 2 are left
  // to the generic kernel, launched beside this one when a batch can hold such codeblocks (fec_packed_launch).
  return fec_packed_covers(D.lifting_size) && (D.lifting_size / 2 <= blockDim.x);
}

// The result store at the end of a codeblock. In the persistent kernel the codeblock index is re-read from shared memory
// (s_cur) and the pointer is made opaque here, so that neither occupies a register through the row bodies, which are at the
// register limit (a D = 19 row spilt 28 registers with them live).
#define FEC_LATE_RESULT_STORE(results_, cb_, res_)                                                                             \
  do {                                                                                                                         \
    fec_cb_result* r_ = (results_);                                                                                            \
    asm volatile("" : "+l"(r_));                                                                                               \
    r_[*reinterpret_cast<volatile uint32_t*>(fec_s_cur)] = (res_);                                                             \
    (void)(cb_);                                                                                                               \
  } while (0)

__device__ __forceinline__ void fec_packed_cb(const fec_cb_desc& D,
                                           uint32_t           cb,
                                           uint8_t*           smem_raw,
                                           const int8_t* __restrict__ llr_base,
                                           uint8_t* __restrict__ out_base,
                                           fec_cb_result* __restrict__ results,
                                           int8_t* __restrict__ harq_base,
                                           const fec_dev_tables* __restrict__ tab,
                                           const fec_packed_tables* __restrict__ ptab,
                                           uint32_t tmem_a,
                                           uint32_t tmem_b,
                                           uint32_t* fec_s_cur,
                                           bool      same_graph)
{
  __shared__ uint32_t s_nsig_arr[2];
#define s_nsig s_nsig_arr[0]

  const uint32_t Z = D.lifting_size, H = Z / 2;
  const int      bg1 = (D.base_graph == 1);
  const uint32_t Kb = bg1 ? 22u : 10u, nfull = bg1 ? 68u : 52u;
  const uint32_t K = Kb * Z, N = (nfull - 2) * Z;
  const uint32_t ne = bg1 ? 316u : 197u;
  const uint32_t invZ = 0xffffffffu / Z + 1;
  uint32_t tid = threadIdx.x;
#ifndef FEC_EXP_NOOPAQUE
  asm volatile("" : "+r"(tid)); // opaque: nothing derived from it is hoisted out of the persistent loop and kept in registers
#endif

#ifdef FEC_PROFILE
  long long tp0 = clock64(), tp1 = 0, tp2 = 0, tp3 = 0;
  long long fec_prof_t = tp0;
#endif
  const fec_packed_smem_layout L = fec_packed_layout(Z, bg1);
  uint32_t*        state  = reinterpret_cast<uint32_t*>(smem_raw + L.state);
  fec_packed_edge* edges  = reinterpret_cast<fec_packed_edge*>(smem_raw + L.edges);
  uint32_t*        bitsw  = reinterpret_cast<uint32_t*>(smem_raw + L.bitsw);
  uint32_t*        red    = reinterpret_cast<uint32_t*>(smem_raw + L.red);
  uint16_t*        soft16 = reinterpret_cast<uint16_t*>(smem_raw + L.soft);

  // Stage the graph of this codeblock, unless the codeblock this CTA decoded before had the same (BG, Z): the edge records
  // are still in place (nothing else writes that part of shared memory, and the layout is a function of (BG, Z)). The 152
  // codeblocks of a transport block share one graph.
  if (!same_graph) {
    const uint4*   src     = reinterpret_cast<const uint4*>(ptab->edges[D.graph_index]);
    uint4*         dst     = reinterpret_cast<uint4*>(edges);
    const uint32_t soft_sw = (uint32_t)__cvta_generic_to_shared(soft16);
    for (uint32_t i = tid; i < ne; i += blockDim.x) {
      uint4 E = src[i];
      E.z += soft_sw; // column base as a shared-window address
      dst[i] = E;
    }
    // The row program (what fec_packed_iteration reads, one 16-byte record per base row): the row schedule word of constant
    // memory decoded once per graph instead of once per row and iteration.
    const uint32_t nrows = bg1 ? 46u : 42u, row_a0 = fec_packed_row_a(bg1), row_b0 = fec_packed_row_b(bg1);
    const uint32_t edges_sw = (uint32_t)__cvta_generic_to_shared(edges);
    const uint32_t state_sw = (uint32_t)__cvta_generic_to_shared(state);
    for (uint32_t l = tid; l < nrows; l += blockDim.x) {
      const uint32_t info = fec_c_rowinfo[bg1 ? 0 : 1][l];
      const uint32_t so   = (info >> 9) & 0xffu;
      uint4          P;
      P.x = edges_sw + 16u * (info & 0x1ffu);
      P.y = (l >= row_b0) ? so : (state_sw + so * H * 4u);
      P.z = (info >> 17) & 15u;
      P.w = (l >= row_a0) ? 1u : 0u;
      reinterpret_cast<uint4*>(smem_raw + L.prog)[l] = P;
    }
  }
  if (tid == 0) {
    s_nsig = 0;
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
  FEC_PROF_MARK(0);
  // ---- stage 1: the soft buffer (rate dematching + HARQ combining), or the decoder input as given --------------------------
  const fec_soft_store_packed store{soft16, Z, H, invZ};
  const int8_t*               harq = nullptr;
  bool                        direct_fill = false;
  if (D.mode == FEC_MODE_DECODE) {
    const uint32_t n_in   = D.E;
    const uint32_t n_node = (n_in / Z) * Z; // full nodes are clamped to +-64, a partial last node is copied as it is
    uint32_t       my_sig = 0;
    for (uint32_t i = tid; i < Z; i += blockDim.x) {
      reinterpret_cast<uint32_t*>(soft16)[i] = 0; // the two punctured columns
    }
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
    int8_t* hq = harq_base + (size_t)D.harq_slot * FEC_HARQ_SLOT_BYTES;
    harq       = hq;
    // the soft-bit area is the staging area of the dematcher: the decoder's copy is filled afterwards
    direct_fill = fec_dematch_packed(D, in, hq, reinterpret_cast<int8_t*>(soft16), L.total - L.soft, &s_nsig, K
#ifdef FEC_PROFILE
                       ,
                       fec_prof_t
#endif
    );
    if (D.mode == FEC_MODE_DEMATCH) {
      if (tid == 0) {
        res.status  = FEC_RES_DONE;
        results[cb] = res;
      }
      return;
    }
  }
  __syncthreads();
  FEC_PROF_MARK(2);

  // ---- stage 2: layered scaled min-sum -------------------------------------------------------------------------------------
  const uint32_t n_sig      = s_nsig;
  const bool     early_stop = (D.flags & FEC_FLAG_EARLY_STOP) != 0;
  uint8_t*       out        = out_base + (((size_t)D.reserved1 << 32) | D.out_offset); // reserved1: bits 32.. of the offset
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
  // Tensor-memory home of this warp's check state: lane quadrant of the warp, one column slot per group of four warps.
  const uint32_t row_b = fec_packed_row_b(bg1), row_a = fec_packed_row_a(bg1);
  const uint32_t tm_q  = (((tid >> 5) & 3u) * 32u) << 16;
  const uint32_t tm_a  = tmem_a + tm_q + (tid >> 7) * FEC_PACKED_TMEM_A_COLS;
  const uint32_t tm_b  = tmem_b + tm_q + (tid >> 7) * FEC_PACKED_TMEM_B_COLS;
  if (direct_fill) {
    for (uint32_t i = tid; i < Z; i += blockDim.x) {
      reinterpret_cast<uint32_t*>(soft16)[i] = 0; // the two punctured columns (the dematcher has written the others)
    }
  } else if (harq != nullptr) {
    // every staged LLR has been consumed (barrier above): the decoder's soft bits, columns of the layers in use only
    fec_soft_fill_from_harq(harq, soft16, cbl - 2 * Z, Z, H, invZ);
  }
  FEC_PROF_MARK(3);
  {
    // Clear the check state of the rows in use ("no c2v messages yet"): shared memory, region B, region A.
    uint4*         s4 = reinterpret_cast<uint4*>(state);
    const uint32_t nw = (nlayers >= row_b) ? fec_packed_state_words(bg1) : ((fec_c_rowinfo[bg1 ? 0 : 1][nlayers] >> 9) & 0xffu);
    const uint32_t n4 = (nw * H * 4 + 15) / 16;
    for (uint32_t i = tid; i < n4; i += blockDim.x) {
      s4[i] = make_uint4(0, 0, 0, 0);
    }
    if (nlayers > row_b) { // every thread clears its own lane, eight columns per store
      tmem_st8_zero(tm_b);
      tmem_st8_zero(tm_b + 8);
      for (uint32_t c = 0; nlayers > row_a && c < 2u * (nlayers - row_a); c += 8) {
        tmem_st8_zero(tm_a + c);
      }
      tmem_wait_st();
    }
  }
  // (Fetching these three where they are used instead, so that nothing of the CRC lives through the row bodies, shrinks the
  // stack frame from 56 to 40 bytes and was measured at -2.4 % for the headline, +1.4 % for the config-3 shape: ptxas
  // allocates and schedules the row bodies differently with every change around them -- see DESIGN.md 3.3.)
  const uint32_t* x8       = tab->crc_x8[D.crc_type];
  const uint32_t  crc_poly = tab->crc_poly[D.crc_type], crc_top = 1u << tab->crc_len[D.crc_type];
  __syncthreads();
  const uint32_t smul = ptab->scale_mul, sadd = ptab->scale_add;
  const uint32_t prog_s  = (uint32_t)__cvta_generic_to_shared(smem_raw + L.prog);

#ifdef FEC_PROFILE
  tp2 = clock64();
#endif
  FEC_PROF_MARK(4);
  bool     pass = false;
  uint32_t it   = 0;
  for (; it < D.max_iter; ++it) {
    fec_packed_iteration(prog_s, H, nlayers, smul, sadd, tm_b, tm_a);
    __syncthreads(); // the last row's stores, before the hard decision or the next iteration
    FEC_PROF_MARK(5);
    if (early_stop) {
      const bool clean = fec_hard_bits_packed(soft16, bitsw, K, Z, H, invZ);
      FEC_PROF_MARK(6);
      const bool crc   = fec_crc_ok_packed(bitsw, T, x8, crc_poly, crc_top, red);
      FEC_PROF_MARK(7);
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
    fec_hard_bits_packed(soft16, bitsw, K, Z, H, invZ);
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
    FEC_PROF_MARK(8);
    atomicAdd(&fec_prof_acc[15], 1ull);
    res.reserved = (uint32_t)min(255ll, (tp1 - tp0) >> 8) | ((uint32_t)min(255ll, (tp2 - tp1) >> 10) << 8) |
                   ((uint32_t)min(255ll, (tp3 - tp2) >> 12) << 16) | ((uint32_t)min(255ll, (tp4 - tp3) >> 8) << 24);
#endif
    FEC_LATE_RESULT_STORE(results, cb, res);
  }
}

#undef s_nsig
// Dispatcher: every codeblock the packed path covers goes there, the rest (odd Z, dematch-only
// operations, unaligned decoder inputs) runs the generic per-check code in the same launch.
// Persistent: the grid is no larger than what the device holds at once (fec_packed_launch) and every CTA takes codeblocks
// from a counter until the batch is empty. A wave of 444 CTAs ending together and 444 new ones starting costs about 18 us
// on a B200 (measured with %globaltimer, profiles/r02_phase_clocks.txt): 7 % of the headline launch and 27 % of a launch of
// high-rate codeblocks that take 28 us each. Tensor memory is allocated once per CTA.
// work[0]: next codeblock - gridDim.x (the first one of a CTA is blockIdx.x); work[1]: CTAs that are done. The last CTA
// to finish clears both, so a slot is ready for the next launch without a memset.
__global__ void __launch_bounds__(FEC_PACKED_MAX_THREADS, FEC_PACKED_CTAS_PER_SM)
fec_cb_packed_kernel(const fec_cb_desc* __restrict__ descs,
                     uint32_t n,
                     const int8_t* __restrict__ llr_base,
                     uint8_t* __restrict__ out_base,
                     fec_cb_result* __restrict__ results,
                     int8_t* __restrict__ harq_base,
                     const fec_dev_tables* __restrict__ tab,
                     const fec_packed_tables* __restrict__ ptab,
                     uint32_t* __restrict__ work)
{
  extern __shared__ __align__(16) uint8_t smem_raw[];
  __shared__ uint32_t s_tmem_base[2];
  __shared__ uint32_t s_next, s_cur, s_graph;
#ifdef FEC_PROFILE
  unsigned long long prof_g0;
  asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(prof_g0));
  const long long prof_c0 = clock64();
#endif
  // Tensor memory for the check state (see the layout notes at the top): region A takes 64 columns per group of four warps,
  // region B 16 (allocations are powers of two >= 32). Three resident CTAs of six warps take 3 x (128 + 32) of the SM's 512
  // columns. One warp allocates and later frees.
  const uint32_t slots  = (blockDim.x > 128) ? 2u : 1u;
  const uint32_t cols_a = slots * FEC_PACKED_TMEM_A_COLS, cols_b = 32u;
  if (threadIdx.x < 32) {
    tmem_alloc((uint32_t)__cvta_generic_to_shared(&s_tmem_base[0]), cols_a);
    tmem_alloc((uint32_t)__cvta_generic_to_shared(&s_tmem_base[1]), cols_b);
    tmem_relinquish();
  }
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
#ifdef FEC_PROFILE
  const long long prof_c1 = clock64();
#endif
  if (threadIdx.x == 0) {
    s_next  = blockIdx.x;
    s_graph = 0xffffffffu; // no graph staged yet
  }
  __syncthreads();
  for (;;) {
    // Nothing is kept in registers from one codeblock to the next, and the codeblock index lives in shared memory (s_cur):
    // the row bodies are at the register limit.
    const fec_cb_desc* p_descs = descs;
    uint32_t*          p_work  = work;
    asm volatile("" : "+l"(p_descs), "+l"(p_work));
    const uint32_t cb         = *reinterpret_cast<volatile uint32_t*>(&s_next);
    const uint32_t last_graph = *reinterpret_cast<volatile uint32_t*>(&s_graph);
    if (cb >= n) {
      break;
    }
    __syncthreads(); // s_next / s_graph have been read by every thread before thread 0 writes them again
    const fec_cb_desc D = p_descs[cb];
    if (threadIdx.x == 0) {
      s_cur  = cb;
      // in flight while this codeblock is decoded (work == nullptr: fixed stride, for measurements)
      s_next = p_work ? gridDim.x + atomicAdd(&p_work[0], 1u) : cb + gridDim.x;
    }
    // (No call of the generic per-check code from here: a function call inside this loop makes the compiler save and restore
    // 28 registers around the degree-19 row bodies. Codeblocks this kernel does not cover are skipped; the generic kernel,
    // launched beside this one, decodes them.)
    if (fec_packed_cb_supported(D)) {
      fec_packed_cb(D, cb, smem_raw, llr_base, out_base, results, harq_base, tab, ptab,
                    *reinterpret_cast<volatile uint32_t*>(&s_tmem_base[0]), *reinterpret_cast<volatile uint32_t*>(&s_tmem_base[1]),
                    &s_cur, last_graph == D.graph_index);
      if (threadIdx.x == 0) {
        s_graph = D.graph_index;
      }
    }
    __syncthreads(); // every warp is done with this codeblock's shared memory; s_next is visible
  }
#ifdef FEC_PROFILE
  const long long prof_c2 = clock64();
#endif
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  if (threadIdx.x < 32) {
    tmem_dealloc(s_tmem_base[0], cols_a);
    tmem_dealloc(s_tmem_base[1], cols_b);
  }
  if (threadIdx.x == 0 && work != nullptr) {
    if (atomicAdd(&work[1], 1u) == gridDim.x - 1) {
      // every CTA has drawn its last (out of range) number: nobody touches the counters of this launch any more
      work[0] = 0;
      work[1] = 0;
    }
  }
#ifdef FEC_PROFILE
  if (threadIdx.x == 0) {
    unsigned long long prof_g1;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(prof_g1));
    atomicAdd(&fec_prof_acc[9], (unsigned long long)(prof_c1 - prof_c0));
    atomicAdd(&fec_prof_acc[10], (unsigned long long)(clock64() - prof_c2));
    atomicMin(&fec_prof_acc[11], prof_g0);
    atomicMax(&fec_prof_acc[12], prof_g1);
    atomicAdd(&fec_prof_acc[13], prof_g1 - prof_g0);
    atomicAdd(&fec_prof_acc[14], 1ull);
  }
#endif
}

// ---- host side ---------------------------------------------------------------------------------------------------------
// Fills the packed tables from the generic ones. Returns false if the scaling LUT cannot be written as
// (x * mul + add) >> 8 on 0..120 (then the engine keeps to the generic kernel).
// Row schedule words for fec_c_rowinfo (see there). Returns false if a row has no instantiation of the row body.
static inline bool fec_packed_build_rowinfo(uint32_t (*info)[48], const fec_dev_tables& t, const fec_packed_tables& pt)
{
  for (int bg = 0; bg < 2; ++bg) {
    const uint32_t M = bg ? 42 : 46, row_b = fec_packed_row_b(bg == 0), row_a = fec_packed_row_a(bg == 0);
    for (uint32_t m = 0; m < 48; ++m) {
      info[bg][m] = 0;
    }
    for (uint32_t m = 0; m < M; ++m) {
      const uint32_t d = t.deg[bg][m];
      uint32_t       code = FEC_ROW_NONE, so = pt.state_off[bg][m];
      if (m >= row_b) {
        code = (d >= 3 && d <= 10) ? FEC_ROW_T3 + (d - 3) : FEC_ROW_NONE;
        so   = (m >= row_a) ? 2 * (m - row_a) : so;
      } else {
        code = (d == 3) ? FEC_ROW_S3 : (d == 8) ? FEC_ROW_S8 : (d == 9) ? FEC_ROW_S9 : (d == 10) ? FEC_ROW_S10 :
               (d == 19) ? FEC_ROW_S19 : FEC_ROW_NONE;
      }
      if (code == FEC_ROW_NONE || so > 0xff || t.row_start[bg][m] > 0x1ff) {
        return false;
      }
      info[bg][m] = t.row_start[bg][m] | (so << 9) | (code << 17);
    }
  }
  return true;
}

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
    // Rows in shared memory: word offset (above). Region B rows: column offset inside the region. Region A rows: two
    // columns each (the kernel computes 2 (row - row_a) itself).
    const uint32_t row_b = fec_packed_row_b(bg == 0), row_a = fec_packed_row_a(bg == 0);
    if (pt.state_off[bg][row_b] != fec_packed_state_words(bg == 0)) {
      return false;
    }

    uint32_t boff = 0;
    for (uint32_t m = row_b; m < row_a; ++m) {
      pt.state_off[bg][m] = (uint16_t)boff;
      boff += fec_packed_row_words(t.deg[bg][m]);
    }
    if (boff > FEC_PACKED_TMEM_B_COLS) {
      return false;
    }
    for (uint32_t m = row_a; m < (uint32_t)M; ++m) {
      if (t.deg[bg][m] > 8) {
        return false; // region A rows hold two state words
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
        e.c0   = 4u * r - 4u * H + 2u * ((q + 1) & 1); // modulo 2^32
        e.c1   = 4u * H + 2u * (q & 1) - 2u * ((q + 1) & 1);
        e.base = 4u * col * H;
        e.pad  = 0;
      }
    }
  }
  return found;
}

// CTAs of `threads` threads and `smem` bytes one SM holds at once (registers, shared memory, tensor memory columns).
static inline uint32_t fec_packed_ctas_per_sm(uint32_t threads, uint32_t smem)
{
  static std::mutex                    mtx;
  static std::map<uint64_t, uint32_t>  cache; // per (threads, smem): the occupancy query costs a few microseconds
  const uint64_t                       key = ((uint64_t)threads << 32) | smem;
  std::lock_guard<std::mutex>          lk(mtx);
  auto                                 it = cache.find(key);
  if (it != cache.end()) {
    return it->second;
  }
  // Shared memory and tensor memory bound the residency (registers: 96 x 192 threads x 3 fits). Computed from the device
  // attributes rather than cudaOccupancyMaxActiveBlocksPerMultiprocessor, which answered 1 for the 73 KB layout although
  // three CTAs are resident (it does not assume the carve-out the launch picks). An overestimate is harmless: a CTA that is
  // not resident at first starts when another one exits, finds the batch empty and exits too.
  int dev = 0, smem_sm = 0, smem_rsv = 0;
  cudaGetDevice(&dev);
  cudaDeviceGetAttribute(&smem_sm, cudaDevAttrMaxSharedMemoryPerMultiprocessor, dev);
  cudaDeviceGetAttribute(&smem_rsv, cudaDevAttrReservedSharedMemoryPerBlock, dev);
  cudaFuncAttributes fa{};
  cudaFuncGetAttributes(&fa, fec_cb_packed_kernel);
  const uint32_t per_cta   = smem + (uint32_t)fa.sharedSizeBytes + (uint32_t)smem_rsv;
  const uint32_t by_smem   = per_cta ? (uint32_t)smem_sm / per_cta : 1u;
  const uint32_t by_warps  = 64u / (threads / 32u);
  const uint32_t tmem_cols = ((threads > 128) ? 2u : 1u) * FEC_PACKED_TMEM_A_COLS + 32u;
  const uint32_t r         = std::max(1u, std::min(std::min(by_smem, by_warps), std::min(512u / tmem_cols, 32u)));
  cache[key]               = r;
  return r;
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
                                    int                      any_bg1,
                                    uint32_t                 sm_count,
                                    uint32_t*                work,
                                    bool                     leftovers,
                                    uint32_t                 grid_div = 1)
{
  uint32_t threads = ((max_Z / 2 + 31) / 32) * 32;
  threads          = threads < 32 ? 32 : (threads > FEC_PACKED_MAX_THREADS ? FEC_PACKED_MAX_THREADS : threads);
  const uint32_t smem = (max_Z >= 4) ? fec_packed_layout(max_Z & ~1u, any_bg1).total : 0;
  // cudaFuncAttributeMaxDynamicSharedMemorySize is set ONCE, to the largest layout, in fec_engine_create: the attribute is
  // per-function global state and launches from several queues / threads would otherwise race on it.
  if (max_Z >= 4) {
    // FEC_B200_ONE_CTA_PER_CB=1 (measurements only): one CTA per codeblock, as before the kernel became persistent
    static const bool one_per_cb = getenv("FEC_B200_ONE_CTA_PER_CB") != nullptr;
    const uint32_t    resident   = sm_count * fec_packed_ctas_per_sm(threads, smem);
    const uint32_t    cap        = std::max(1u, resident / std::max(1u, grid_div));
    const uint32_t    grid       = one_per_cb ? n : std::min(n, cap);
    static const bool fixed_stride = getenv("FEC_B200_FIXED_STRIDE") != nullptr; // measurements only
    fec_cb_packed_kernel<<<grid, threads, smem, st>>>(descs, n, llr, out, res, harq, tab, ptab, fixed_stride ? nullptr : work);
    if (cudaGetLastError() != cudaSuccess) {
      return FEC_ERR_CUDA;
    }
  }
  if (leftovers) {
    // Odd lifting sizes and Z = 2 (all <= 15): one small CTA per codeblock of the batch; those the packed kernel covers
    // return at once. The two kernels touch disjoint codeblocks, HARQ slots and outputs: no ordering between them matters.
    const uint32_t zl = max_Z < 15 ? max_Z : 15;
    fec_cb_generic_kernel<<<n, 32, fec_generic_layout(zl, any_bg1).total, st>>>(descs, llr, out, res, harq, tab, 1);
    if (cudaGetLastError() != cudaSuccess) {
      return FEC_ERR_CUDA;
    }
  }
  return FEC_OK;
}
