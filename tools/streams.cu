// micro-benchmark: how fast can one kernel stream NR read planes + NW written planes of 1024x1024 doubles (separate 8.4 MB planes)?
#include <cstdio>
#include <cuda_runtime.h>
#include <vector>
struct P { const double* r[32]; double* w[12]; };
template <int NR, int NW>
__global__ void __launch_bounds__(256) k(const __grid_constant__ P p, int SX, int Lm, int Mm) {
  const int i = blockIdx.x * 32 + threadIdx.x, j = blockIdx.y * 8 + threadIdx.y;
  if (i >= Lm || j >= Mm) return;
  const long c = 16 + i + (long)(j + 2) * SX;
  double s = 0;
#pragma unroll
  for (int q = 0; q < NR; ++q) s += p.r[q][c];
#pragma unroll
  for (int q = 0; q < NW; ++q) p.w[q][c] = s + q;
}
template <int NR, int NW>
void run(P p, int SX, int Lm, int Mm, cudaStream_t st) {
  dim3 b(32, 8), g((Lm + 31) / 32, (Mm + 7) / 8);
  cudaEvent_t a, e; cudaEventCreate(&a); cudaEventCreate(&e);
  for (int w = 0; w < 5; ++w) k<NR, NW><<<g, b, 0, st>>>(p, SX, Lm, Mm);
  cudaEventRecord(a, st);
  const int reps = 50;
  for (int w = 0; w < reps; ++w) k<NR, NW><<<g, b, 0, st>>>(p, SX, Lm, Mm);
  cudaEventRecord(e, st); cudaEventSynchronize(e);
  float ms; cudaEventElapsedTime(&ms, a, e);
  const double us = ms * 1e3 / reps, gb = (double)(NR + NW) * 8 * Lm * Mm / 1e9;
  printf("NR=%2d NW=%2d  %.1f us  %.0f GB/s\n", NR, NW, us, gb / (us * 1e-6));
}
int main() {
  const int Lm = 1024, Mm = 1024, SX = 1056; const size_t n = (size_t)SX * (Mm + 4);
  P p; std::vector<double*> all;
  // 60 planes spread in one arena like the library's (others in between to defeat L2: total 60*8.9MB = 534 MB)
  double* arena; cudaMalloc(&arena, 64 * n * 8); cudaMemset(arena, 0, 64 * n * 8);
  for (int q = 0; q < 32; ++q) p.r[q] = arena + (size_t)q * n;
  for (int q = 0; q < 12; ++q) p.w[q] = arena + (size_t)(40 + q) * n;
  cudaStream_t st; cudaStreamCreate(&st);
  run<2, 1>(p, SX, Lm, Mm, st); run<4, 2>(p, SX, Lm, Mm, st); run<8, 2>(p, SX, Lm, Mm, st); run<12, 4>(p, SX, Lm, Mm, st);
  run<16, 4>(p, SX, Lm, Mm, st); run<20, 6>(p, SX, Lm, Mm, st); run<27, 8>(p, SX, Lm, Mm, st); run<32, 8>(p, SX, Lm, Mm, st); run<32, 12>(p, SX, Lm, Mm, st);
  return 0;
}
