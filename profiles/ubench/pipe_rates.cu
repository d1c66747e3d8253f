// instruction throughput microbenchmarks (per SMSP, warp-instructions per clock)
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>
#define ITERS 512
#define NACC 8
template <int OP>
__device__ __forceinline__ uint32_t op(uint32_t a, uint32_t b, uint32_t c)
{
  uint32_t r;
  if (OP == 0) asm volatile("fma.rn.f16x2 %0, %1, %2, %3;" : "=r"(r) : "r"(a), "r"(b), "r"(c));
  if (OP == 1) asm volatile("add.rn.f16x2 %0, %1, %2;" : "=r"(r) : "r"(a), "r"(b));
  if (OP == 2) asm volatile("min.xorsign.abs.f16x2 %0, %1, %2;" : "=r"(r) : "r"(a), "r"(b));
  if (OP == 3) asm volatile("set.eq.f16x2.f16x2 %0, %1, %2;" : "=r"(r) : "r"(a), "r"(b));
  if (OP == 4) asm volatile("lop3.b32 %0, %1, %2, %3, 0x96;" : "=r"(r) : "r"(a), "r"(b), "r"(c));
  if (OP == 5) asm volatile("mad.lo.u32 %0, %1, %2, %3;" : "=r"(r) : "r"(a), "r"(b), "r"(c));
  if (OP == 6) asm volatile("mul.hi.u32 %0, %1, %2;" : "=r"(r) : "r"(a), "r"(b));
  if (OP == 7) asm volatile("prmt.b32 %0, %1, %2, %3;" : "=r"(r) : "r"(a), "r"(b), "r"(c));
  if (OP == 8) asm volatile("min.f16x2 %0, %1, %2;" : "=r"(r) : "r"(a), "r"(b));
  if (OP == 9) asm volatile("fma.rn.f32 %0, %1, %2, %3;" : "=r"(r) : "r"(a), "r"(b), "r"(c));
  if (OP == 10) asm volatile("add.u32 %0, %1, %2;" : "=r"(r) : "r"(a), "r"(b));
  if (OP == 11) asm volatile("shf.l.wrap.b32 %0, %1, %2, %3;" : "=r"(r) : "r"(a), "r"(b), "r"(c));
  if (OP == 12) asm volatile("{.reg .pred p; setp.ge.u32 p, %1, %2; selp.b32 %0, %1, %3, p;}" : "=r"(r) : "r"(a), "r"(b), "r"(c));
  if (OP == 13) asm volatile("vabsdiff4.u32.u32.u32.add %0, %1, %2, %3;" : "=r"(r) : "r"(a), "r"(b), "r"(c));
  if (OP == 14) asm volatile("fma.rn.bf16x2 %0, %1, %2, %3;" : "=r"(r) : "r"(a), "r"(b), "r"(c));
  if (OP == 15) asm volatile("mul.rn.f16x2 %0, %1, %2;" : "=r"(r) : "r"(a), "r"(b));
  if (OP == 16) asm volatile("max.f16x2 %0, %1, %2;" : "=r"(r) : "r"(a), "r"(b));
  if (OP == 17) asm volatile("set.eq.u32.f16x2 %0, %1, %2;" : "=r"(r) : "r"(a), "r"(b));
  if (OP == 18) asm volatile("min.u32 %0, %1, %2;" : "=r"(r) : "r"(a), "r"(b));
  if (OP == 19) asm volatile("min.f32 %0, %1, %2;" : "=r"(r) : "r"(a), "r"(b));
  if (OP == 20) asm volatile("{.reg .b32 t; abs.f16x2 t, %1; add.rn.f16x2 %0, t, %2;}" : "=r"(r) : "r"(a), "r"(b));
  if (OP == 21) asm volatile("{.reg .b32 t; abs.f16x2 t, %1; set.eq.f16x2.f16x2 %0, t, %2;}" : "=r"(r) : "r"(a), "r"(b));
  if (OP == 22) asm volatile("shl.b32 %0, %1, 3;" : "=r"(r) : "r"(a));
  if (OP == 23) asm volatile("mad.lo.u32 %0, %1, 8, %2;" : "=r"(r) : "r"(a), "r"(b));
  if (OP == 24) asm volatile("{.reg .b32 t; add.u32 t, %1, %2; add.u32 %0, t, %3;}" : "=r"(r) : "r"(a), "r"(b), "r"(c));
  if (OP == 25) asm volatile("min.s16x2 %0, %1, %2;" : "=r"(r) : "r"(a), "r"(b));
  if (OP == 26) asm volatile("{.reg .b32 t; neg.f16x2 t, %1; add.rn.f16x2 %0, t, %2;}" : "=r"(r) : "r"(a), "r"(b));
  if (OP == 27) asm volatile("{.reg .b32 t; abs.f16x2 t, %1; min.f16x2 %0, t, %2;}" : "=r"(r) : "r"(a), "r"(b));
  if (OP == 28) asm volatile("{.reg .b16 lo, hi; mov.b32 {lo, hi}, %1; add.rn.f16 lo, hi, lo; mov.b32 %0, {lo, hi};}" : "=r"(r) : "r"(a));
  if (OP == 29) asm volatile("and.b32 %0, %1, %2;" : "=r"(r) : "r"(a), "r"(b));
  if (OP == 30) asm volatile("xor.b32 %0, %1, 0x80008000;" : "=r"(r) : "r"(a));
  return r;
}
// pairs: OP A and OP B interleaved
template <int A, int B>
__global__ void k(uint32_t* out, uint32_t x, uint32_t y, long long* clk)
{
  uint32_t acc[NACC];
#pragma unroll
  for (int i = 0; i < NACC; ++i) acc[i] = threadIdx.x + i;
  long long t0 = clock64();
  for (int it = 0; it < ITERS; ++it) {
#pragma unroll
    for (int i = 0; i < NACC; ++i) {
      acc[i] = op<A>(acc[i], x, y);
      if (B >= 0) acc[(i + NACC / 2) % NACC] = op<(B < 0 ? 0 : B)>(acc[(i + NACC / 2) % NACC], y, x);
    }
  }
  long long t1 = clock64();
  uint32_t s = 0;
#pragma unroll
  for (int i = 0; i < NACC; ++i) s ^= acc[i];
  out[blockIdx.x * blockDim.x + threadIdx.x] = s;
  if (threadIdx.x == 0 && blockIdx.x == 0) *clk = t1 - t0;
}
// shared memory byte load/store throughput
template <int MODE>
__global__ void ks(uint32_t* out, uint32_t x, long long* clk)
{
  __shared__ uint8_t sm[16384];
  for (int i = threadIdx.x; i < 16384; i += blockDim.x) sm[i] = i;
  __syncthreads();
  uint32_t base = (uint32_t)__cvta_generic_to_shared(sm) + threadIdx.x * 2;
  uint32_t acc = 0;
  long long t0 = clock64();
  for (int it = 0; it < ITERS; ++it) {
#pragma unroll
    for (int i = 0; i < NACC; ++i) {
      uint32_t r;
      uint32_t a = base + ((it * NACC + i) * 2 * 32 & 8191) + x;
      if (MODE == 0) { asm volatile("ld.shared.u8 %0, [%1];" : "=r"(r) : "r"(a)); acc ^= r; }
      if (MODE == 1) { asm volatile("ld.shared.u16 %0, [%1];" : "=r"(r) : "r"(a)); acc ^= r; }
      if (MODE == 2) { asm volatile("st.shared.u8 [%0], %1;" :: "r"(a), "r"(acc + i) : "memory"); }
      if (MODE == 3) { asm volatile("st.shared.u16 [%0], %1;" :: "r"(a), "r"(acc + i) : "memory"); }
      if (MODE == 4) { asm volatile("ld.shared.u32 %0, [%1];" : "=r"(r) : "r"(a * 2)); acc ^= r; }
    }
  }
  long long t1 = clock64();
  out[blockIdx.x * blockDim.x + threadIdx.x] = acc + sm[threadIdx.x];
  if (threadIdx.x == 0 && blockIdx.x == 0) *clk = t1 - t0;
}
template <int A, int B>
void run(const char* name, int warps_per_sm, uint32_t* out, long long* clk)
{
  k<A, B><<<148, warps_per_sm * 32>>>(out, 0x3c003c00u, 0x00010001u, clk);
  cudaDeviceSynchronize();
  k<A, B><<<148, warps_per_sm * 32>>>(out, 0x3c003c00u, 0x00010001u, clk);
  cudaDeviceSynchronize();
  long long c;
  cudaMemcpy(&c, clk, 8, cudaMemcpyDeviceToHost);
  double n = (double)ITERS * NACC * (B >= 0 ? 2 : 1) * (warps_per_sm / 4.0);
  printf("%-22s warps/SM %2d : %.3f warp-instr/clk/SMSP\n", name, warps_per_sm, n / c);
}
template <int M>
void runs(const char* name, int warps_per_sm, uint32_t* out, long long* clk)
{
  ks<M><<<148, warps_per_sm * 32>>>(out, 0, clk);
  cudaDeviceSynchronize();
  ks<M><<<148, warps_per_sm * 32>>>(out, 0, clk);
  cudaDeviceSynchronize();
  long long c;
  cudaMemcpy(&c, clk, 8, cudaMemcpyDeviceToHost);
  double n = (double)ITERS * NACC * warps_per_sm;
  printf("%-22s warps/SM %2d : %.3f warp-instr/clk/SM\n", name, warps_per_sm, n / c);
}
#define R1(A, name) run<A, -1>(name, 4, out, clk); run<A, -1>(name, 12, out, clk);
#define R2(A, B, name) run<A, B>(name, 4, out, clk); run<A, B>(name, 12, out, clk);
int main()
{
  uint32_t* out; long long* clk;
  cudaMalloc(&out, 148 * 1024 * 4); cudaMalloc(&clk, 8);
  R1(8, "HMNMX2 plain") R1(16, "HMNMX2 max") R1(17, "HSET2 u32mask") R1(18, "VIMNMX u32") R1(19, "FMNMX") R1(20, "HADD2 |a|") R1(21, "HSET2 |a|") R1(22, "SHL") R1(23, "IMAD x8") R1(24, "IADD3") R1(25, "VIMNMX s16x2") R1(26, "HADD2 -a") R1(27, "HMNMX2 |a| plain") R1(28, "HADD scalar hi") R1(29, "AND") R1(30, "XOR imm")
  R2(8, 4, "HMNMX2p+LOP3") R2(8, 0, "HMNMX2p+HFMA2") R2(8, 2, "HMNMX2p+HMNMX2xs") R2(24, 4, "IADD3+LOP3") R2(24, 0, "IADD3+HFMA2") R2(8, 24, "HMNMX2p+IADD3") R2(0, 4, "HFMA2+LOP3") R2(17, 0, "HSET2u32+HFMA2") R2(7, 0, "PRMT+HFMA2") R2(9, 4, "FFMA+LOP3") R2(19, 4, "FMNMX+LOP3") R2(19, 0, "FMNMX+HFMA2")
  runs<0>("LDS.U8", 12, out, clk); runs<1>("LDS.U16", 12, out, clk); runs<2>("STS.U8", 12, out, clk); runs<3>("STS.U16", 12, out, clk); runs<4>("LDS.32", 12, out, clk);
  return 0;
}
