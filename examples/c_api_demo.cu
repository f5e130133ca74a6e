/* Minimal C-ABI user of libheat2d_cuda.so (no Python, no torch):
 *   nvcc -Iinclude examples/c_api_demo.cu -Lheat2d_b200/_lib -lheat2d_cuda -o c_api_demo
 *   LD_LIBRARY_PATH=heat2d_b200/_lib ./c_api_demo 2560 2048 1000
 * Solves the reference problem (grad1612_cuda_heat.cu:64-94) with k = 8 temporal blocking
 * and prints the reference's two lines plus the throughput. */
#include <cuda_runtime.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include "heat2d.h"

#define CHECK(x) do { int rc__ = (x); if (rc__ != 0) { fprintf(stderr, "%s failed: %s\n", #x, h2d_last_error()); return 1; } } while (0)
#define CU(x) do { cudaError_t e__ = (x); if (e__ != cudaSuccess) { fprintf(stderr, "%s: %s\n", #x, cudaGetErrorString(e__)); return 1; } } while (0)

int main(int argc, char** argv) {
  const int nx = argc > 1 ? atoi(argv[1]) : 2560, ny = argc > 2 ? atoi(argv[2]) : 2048;
  const int steps = argc > 3 ? atoi(argv[3]) : 1000, k = 8;
  h2d_tile_t t;
  t.nx = nx; t.ny = ny; t.x0 = 0; t.y0 = 0; t.rows = nx; t.cols = ny; t.halo_r = 8; t.halo_c = 32;
  t.pitch = ((int64_t)ny + 64 + 31) / 32 * 32;
  const size_t bytes = (size_t)(nx + 16) * t.pitch * sizeof(float);
  float *a = NULL, *b = NULL, *norm = NULL;
  CU(cudaMalloc(&a, bytes)); CU(cudaMalloc(&b, bytes)); CU(cudaMalloc(&norm, sizeof(float)));
  CU(cudaMemset(a, 0, bytes)); CU(cudaMemset(b, 0, bytes));
  printf("Problem size: %dx%d\nAmount of iterations: %d\n", nx, ny, steps);
  CHECK(h2d_cuda_init(a, &t, H2D_INIT_EXACT, 1, 0));

  h2d_plan_desc_t d;
  memset(&d, 0, sizeof d);
  d.tile = t; d.phys.cx = 0.1; d.phys.cy = 0.1; d.phys.boundary = H2D_BOUNDARY_FIXED; d.phys.arith = H2D_ARITH_FP32;
  d.buf[0] = a; d.buf[1] = b; d.kmax = k;
  h2d_plan_t* plan = h2d_plan_create(&d);
  if (!plan) { fprintf(stderr, "plan: %s\n", h2d_last_error()); return 1; }

  cudaEvent_t e0, e1; CU(cudaEventCreate(&e0)); CU(cudaEventCreate(&e1));
  CU(cudaEventRecord(e0, 0));
  int par = 0, done = 0; uint32_t epoch = 0;
  h2d_plan_t* plans[1] = {plan};
  const int full = steps / k;
  CHECK(h2d_plan_run_many(plans, 1, par, k, full, epoch + 1, NULL, 0));
  epoch += full; par ^= full & 1; done += full * k;
  if (done < steps) { CHECK(h2d_plan_launch(plan, par, steps - done, ++epoch, H2D_NORM_NONE, 0)); par ^= 1; done = steps; }
  CU(cudaEventRecord(e1, 0)); CU(cudaEventSynchronize(e1));
  float ms = 0; CU(cudaEventElapsedTime(&ms, e0, e1));
  printf("Elapsed time: %e sec\n", ms / 1e3);
  printf("%.1f GLUP/s, %lld kernel launches\n", (double)nx * ny * steps / (ms * 1e-3) / 1e9,
         (long long)h2d_plan_launch_count(plan));

  /* how far is the field from the previous buffer?  (allocation-free norm entry point) */
  float* cur = par == 0 ? a : b; float* prev = par == 0 ? b : a;
  CHECK(h2d_cuda_norm_diff_static(cur, prev, &t, H2D_NORM_LINF, norm, 0));
  float hn = 0; CU(cudaMemcpy(&hn, norm, sizeof hn, cudaMemcpyDeviceToHost));
  float centre = 0;
  CU(cudaMemcpy(&centre, cur + (size_t)(nx / 2 + t.halo_r) * t.pitch + t.halo_c + ny / 2, sizeof centre, cudaMemcpyDeviceToHost));
  printf("max |u(t) - u(t-%d)| = %g, u(centre) = %g\n", steps % k ? steps % k : k, hn, centre);
  h2d_plan_destroy(plan);
  cudaFree(a); cudaFree(b); cudaFree(norm);
  return 0;
}
