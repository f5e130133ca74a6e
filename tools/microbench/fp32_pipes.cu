// Throughput of the fp32 instruction forms the stencil uses, per SM sub-partition.
// Each test: 8 independent register chains per thread, 256 iterations x 8 ops, `W` warps per SMSP.
#include <cstdio>
#include <cuda_runtime.h>

#define ITERS 512
template <int MODE>
__global__ void k(float* out, float a, float b, long long* cyc) {
  // cyc[2*w], cyc[2*w+1] = start / end clock of warp w of block 0
  float2 r[8];
  for (int i = 0; i < 8; ++i) r[i] = make_float2(threadIdx.x * 0.001f + i, i * 0.5f + a);
  float2 ca = make_float2(a, a), cb = make_float2(b, b);
  long long t0 = clock64();
#pragma unroll 1
  for (int it = 0; it < ITERS; ++it) {
#pragma unroll
    for (int i = 0; i < 8; ++i) {
      if (MODE == 0) r[i] = __ffma2_rn(r[i], ca, cb);                                  // FFMA2 reg
      if (MODE == 1) r[i] = __ffma2_rn(r[i], make_float2(0.999f, 0.999f), r[(i + 1) & 7]);  // FFMA2 imm, 2 reg pairs
      if (MODE == 2) r[i] = __fadd2_rn(r[i], r[(i + 3) & 7]);                          // FADD2
      if (MODE == 3) r[i] = __fmul2_rn(r[i], make_float2(0.999f, 0.999f));             // FMUL2 imm
      if (MODE == 4) { r[i].x = __fmaf_rn(r[i].x, a, b); r[i].y = __fmaf_rn(r[i].y, a, b); }   // 2x FFMA reg
      if (MODE == 5) { r[i].x = __fmaf_rn(r[i].x, 0.999f, r[(i+1)&7].x); r[i].y = __fmaf_rn(r[i].y, 0.999f, r[(i+1)&7].y); } // 2x FFMA imm
      if (MODE == 6) { r[i].x = __fadd_rn(r[i].x, r[(i+3)&7].y); r[i].y = __fadd_rn(r[i].y, r[(i+5)&7].x); }  // 2x FADD
      if (MODE == 7) { r[i].x = __shfl_up_sync(0xffffffffu, r[i].x, 1); r[i].y = __fadd_rn(r[i].y, r[(i+5)&7].x); } // SHFL + FADD
      if (MODE == 8) { r[i] = __ffma2_rn(r[i], make_float2(0.999f, 0.999f), r[(i + 1) & 7]); r[(i+4)&7].x = __fadd_rn(r[(i+4)&7].x, r[(i+5)&7].y); } // FFMA2imm + FADD
      if (MODE == 9) { r[i] = __ffma2_rn(r[i], make_float2(0.999f, 0.999f), r[(i + 1) & 7]); r[(i+4)&7].x = __shfl_up_sync(0xffffffffu, r[(i+4)&7].x, 1); } // FFMA2imm + SHFL
    }
  }
  long long t1 = clock64();
  float s = 0;
  for (int i = 0; i < 8; ++i) s += r[i].x + r[i].y;
  out[blockIdx.x * blockDim.x + threadIdx.x] = s;
  if ((threadIdx.x & 31) == 0 && blockIdx.x == 0) { cyc[2 * (threadIdx.x >> 5)] = t0; cyc[2 * (threadIdx.x >> 5) + 1] = t1; }
}

template <int MODE>
void run(const char* name, int ops_per_inner, int warps_per_smsp) {
  float* out; long long* cyc;
  int threads = warps_per_smsp * 4 * 32;
  cudaMalloc(&out, 148 * threads * sizeof(float));
  cudaMalloc(&cyc, 8 * 2 * 64);
  k<MODE><<<148, threads>>>(out, 0.999f, 0.001f, cyc);
  k<MODE><<<148, threads>>>(out, 0.999f, 0.001f, cyc);
  cudaDeviceSynchronize();
  long long hh[128]; cudaMemcpy(hh, cyc, 8 * 2 * 64, cudaMemcpyDeviceToHost);
  long long lo = hh[0], hi = hh[1];
  for (int i = 0; i < warps_per_smsp * 4; ++i) { if (hh[2*i] < lo) lo = hh[2*i]; if (hh[2*i+1] > hi) hi = hh[2*i+1]; }
  long long h = hi - lo;
  double inst = (double)ITERS * 8 * ops_per_inner * warps_per_smsp;  // warp-instructions per SMSP
  printf("%-28s warps/SMSP=%d  cycles=%lld  warp-instr/cycle/SMSP=%.3f  cycles/instr=%.3f\n", name, warps_per_smsp, h,
         inst / h, h / inst);
  cudaFree(out); cudaFree(cyc);
}

// Whole-chip fp32 rate with the packed immediate form (the kernels' hot instruction), event-timed.
static void peak() {
  float* out; long long* cyc;
  const int threads = 512, blocks = 148 * 4;
  cudaMalloc(&out, (size_t)blocks * threads * sizeof(float));
  cudaMalloc(&cyc, 8 * 2 * 64);
  cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
  for (int rep = 0; rep < 3; ++rep) {
    cudaEventRecord(e0);
    for (int i = 0; i < 20; ++i) k<1><<<blocks, threads>>>(out, 0.999f, 0.001f, cyc);
    cudaEventRecord(e1); cudaEventSynchronize(e1);
  }
  float ms; cudaEventElapsedTime(&ms, e0, e1);
  const double fma = 20.0 * blocks * threads * (double)ITERS * 8 * 2;  // FFMA2 = 2 FMA per lane
  printf("PEAK fp32 (FFMA2 imm, whole chip, event-timed): %.1f TFLOP/s  (%.3f ms for %.3e FMA)\n",
         2.0 * fma / (ms * 1e-3) / 1e12, ms, fma);
}

int main() {
  peak();
  for (int w : {1, 3, 4, 8}) {
    if (w == 1) printf("--- latency-ish (1 warp per SMSP, 8 chains)\n"); else printf("--- %d warps per SMSP\n", w);
    switch (w) { default: break; }
    run<0>("FFMA2 reg,reg,reg", 1, w);
    run<1>("FFMA2 reg,imm,reg", 1, w);
    run<2>("FADD2", 1, w);
    run<3>("FMUL2 imm", 1, w);
    run<4>("FFMA reg x2", 2, w);
    run<5>("FFMA imm x2", 2, w);
    run<6>("FADD x2", 2, w);
    run<7>("SHFL + FADD", 2, w);
    run<8>("FFMA2imm + FADD", 2, w);
    run<9>("FFMA2imm + SHFL", 2, w);
  }
  return 0;
}
