// profiles/ubench/h2d_bw.cu -- what the host can feed the GPUs: pinned host memory -> device (and back) with cudaMemcpyAsync,
// one GPU at a time and all visible GPUs at once (one host thread and one pinned buffer per GPU). The end-to-end numbers of
// bench.py are bounded by this (8.5 bits of soft-bit input per decoded information bit at the config-3 rate).
//   nvcc -O2 -o h2d_bw h2d_bw.cu && ./h2d_bw [MiB per copy] [copies]
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <cuda_runtime.h>
#include <thread>
#include <vector>

static double run(int dev, size_t bytes, int reps, bool h2d, void* h, void* d, cudaStream_t st)
{
  cudaSetDevice(dev);
  cudaMemcpyAsync(h2d ? d : h, h2d ? h : d, bytes, h2d ? cudaMemcpyHostToDevice : cudaMemcpyDeviceToHost, st);
  cudaStreamSynchronize(st);
  const auto t0 = std::chrono::steady_clock::now();
  for (int i = 0; i < reps; ++i) {
    cudaMemcpyAsync(h2d ? d : h, h2d ? h : d, bytes, h2d ? cudaMemcpyHostToDevice : cudaMemcpyDeviceToHost, st);
  }
  cudaStreamSynchronize(st);
  const double s = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
  return (double)bytes * reps / s / 1e9;
}

int main(int argc, char** argv)
{
  const size_t bytes = (size_t)((argc > 1) ? atoi(argv[1]) : 64) << 20;
  const int    reps  = (argc > 2) ? atoi(argv[2]) : 40;
  int          n     = 0;
  cudaGetDeviceCount(&n);
  std::vector<void*>        h(n), d(n);
  std::vector<cudaStream_t> st(n);
  for (int i = 0; i < n; ++i) {
    cudaSetDevice(i);
    cudaHostAlloc(&h[i], bytes, cudaHostAllocDefault);
    cudaMalloc(&d[i], bytes);
    cudaStreamCreate(&st[i]);
  }
  printf("# %d GPU(s), %zu MiB per copy, %d copies, host threads %u\n", n, bytes >> 20, reps, std::thread::hardware_concurrency());
  for (int dir = 0; dir < 2; ++dir) {
    for (int i = 0; i < n; ++i) {
      printf("%s gpu %d alone            %7.2f GB/s\n", dir ? "d2h" : "h2d", i, run(i, bytes, reps, !dir, h[i], d[i], st[i]));
    }
    if (n > 1) {
      std::vector<double>      bw(n);
      std::vector<std::thread> th;
      for (int i = 0; i < n; ++i) {
        th.emplace_back([&, i] { bw[i] = run(i, bytes, reps, !dir, h[i], d[i], st[i]); });
      }
      double sum = 0;
      for (int i = 0; i < n; ++i) {
        th[i].join();
        sum += bw[i];
      }
      for (int i = 0; i < n; ++i) {
        printf("%s gpu %d with all %d at once %7.2f GB/s\n", dir ? "d2h" : "h2d", i, n, bw[i]);
      }
      printf("%s aggregate               %7.2f GB/s\n", dir ? "d2h" : "h2d", sum);
    }
  }
  // both directions at once on GPU 0 (the engine's pipelines overlap input and output copies)
  if (n > 0) {
    cudaSetDevice(0);
    void *h2, *d2;
    cudaStream_t s2;
    cudaHostAlloc(&h2, bytes, cudaHostAllocDefault);
    cudaMalloc(&d2, bytes);
    cudaStreamCreate(&s2);
    double a = 0, b = 0;
    std::thread t1([&] { a = run(0, bytes, reps, true, h[0], d[0], st[0]); });
    std::thread t2([&] { b = run(0, bytes, reps, false, h2, d2, s2); });
    t1.join();
    t2.join();
    printf("duplex gpu 0: h2d %7.2f GB/s + d2h %7.2f GB/s\n", a, b);
  }
  return 0;
}
