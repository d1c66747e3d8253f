// Which SM sub-partition (scheduler) do the warps of three resident 6-warp CTAs land on?
// %warpid is the physical warp slot; slot % 4 is the sub-partition on sm_90 / sm_100.
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o warp_slots warp_slots.cu ; run on the GPU box.
#include <cstdio>
#include <cuda_runtime.h>
#include <vector>

__global__ void __launch_bounds__(192, 3) probe(unsigned* out, int spin)
{
  extern __shared__ unsigned char smem[];
  unsigned warpid, smid;
  asm volatile("mov.u32 %0, %%warpid;" : "=r"(warpid));
  asm volatile("mov.u32 %0, %%smid;" : "=r"(smid));
  if ((threadIdx.x & 31) == 0) {
    out[(blockIdx.x * 8 + (threadIdx.x >> 5)) * 2]     = smid;
    out[(blockIdx.x * 8 + (threadIdx.x >> 5)) * 2 + 1] = warpid;
  }
  // stay resident so that all three CTAs of an SM coexist
  long long t0 = clock64();
  while (clock64() - t0 < spin) {
  }
  if (smem[threadIdx.x] == 77 && spin < 0) {
    out[0] = 1;
  }
}

int main()
{
  int nsm = 0;
  cudaDeviceGetAttribute(&nsm, cudaDevAttrMultiProcessorCount, 0);
  const int threads = 192, smem = 73200;
  cudaFuncSetAttribute(probe, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
  for (int nblk : {nsm * 3, nsm * 6}) {
    unsigned* d;
    cudaMalloc(&d, nblk * 16 * sizeof(unsigned));
    cudaMemset(d, 0xff, nblk * 16 * sizeof(unsigned));
    probe<<<nblk, threads, smem>>>(d, 2000000);
    cudaError_t err = cudaDeviceSynchronize();
    std::vector<unsigned> h(nblk * 16);
    cudaMemcpy(h.data(), d, h.size() * 4, cudaMemcpyDeviceToHost);
    // per SM: warps per sub-partition (all CTAs that ran there), and the pattern of the first CTAs
    std::vector<int> cnt(nsm * 4, 0);
    int pattern[8][4] = {};
    for (int b = 0; b < nblk; ++b) {
      for (int w = 0; w < 6; ++w) {
        unsigned smid = h[(b * 8 + w) * 2], wid = h[(b * 8 + w) * 2 + 1];
        if (smid < (unsigned)nsm) {
          cnt[smid * 4 + wid % 4]++;
          pattern[w][wid % 4]++;
        }
      }
    }
    printf("%d CTAs of %d threads, %d B shared memory (%s)\n", nblk, threads, smem, cudaGetErrorString(err));
    printf("local warp -> sub-partition histogram over all CTAs:\n");
    for (int w = 0; w < 6; ++w) {
      printf("  warp %d: %5d %5d %5d %5d\n", w, pattern[w][0], pattern[w][1], pattern[w][2], pattern[w][3]);
    }
    int hist[64] = {};
    for (int s = 0; s < nsm; ++s) {
      int mx = 0;
      for (int k = 0; k < 4; ++k) {
        mx = cnt[s * 4 + k] > mx ? cnt[s * 4 + k] : mx;
      }
      hist[mx < 63 ? mx : 63]++;
    }
    printf("SMs by the warp count of their fullest sub-partition (total over the launch):");
    for (int k = 0; k < 64; ++k) {
      if (hist[k]) {
        printf("  %d warps: %d SMs", k, hist[k]);
      }
    }
    printf("\nfirst SMs: ");
    for (int s = 0; s < 6; ++s) {
      printf("[%d %d %d %d] ", cnt[s * 4], cnt[s * 4 + 1], cnt[s * 4 + 2], cnt[s * 4 + 3]);
    }
    printf("\nCTA 0 warp slots:");
    for (int w = 0; w < 6; ++w) {
      printf(" %u", h[w * 2 + 1]);
    }
    printf("   CTA 1:");
    for (int w = 0; w < 6; ++w) {
      printf(" %u", h[(8 + w) * 2 + 1]);
    }
    printf("\n\n");
    cudaFree(d);
  }
  return 0;
}
