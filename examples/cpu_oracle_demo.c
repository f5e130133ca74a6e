/* heat2d-b200 from plain C, CPU only: the serial / OpenMP oracle and the dump writers.
 *
 *   cc -O2 -Iinclude examples/cpu_oracle_demo.c -Lheat2d_b200/_lib -lheat2d_cpu \
 *      -Wl,-rpath,$PWD/heat2d_b200/_lib -o cpu_oracle_demo
 *   ./cpu_oracle_demo [nx ny steps out_dir]
 *
 * Does what mpi_heat2Dn.c does on one process (same initial condition, same float
 * coefficients, fixed boundary, same initial.dat / final.dat format, :84-131), then
 * repeats the run with a convergence test in the style of grad1612_mpi_heat.c:262-271.
 * No reference code is used: every number comes from libheat2d_cpu.so.             */
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include "heat2d.h"

#define CHECK(call)                                                              \
  do {                                                                           \
    if ((call) != H2D_OK) {                                                      \
      fprintf(stderr, "%s failed: %s\n", #call, h2d_last_error());               \
      return 1;                                                                  \
    }                                                                            \
  } while (0)

int main(int argc, char** argv) {
  const int nx = argc > 1 ? atoi(argv[1]) : 10;
  const int ny = argc > 2 ? atoi(argv[2]) : 10;
  const long steps = argc > 3 ? atol(argv[3]) : 100;
  const char* dir = argc > 4 ? argv[4] : ".";
  char path[4096];
  if (nx < 1 || ny < 1 || steps < 0) {
    fprintf(stderr, "usage: %s [nx ny steps out_dir]\n", argv[0]);
    return 2;
  }
  printf("%s\nGrid size: X= %d  Y= %d  Time steps= %ld\n", h2d_version(), nx, ny, steps);

  float* u = (float*)malloc(sizeof(float) * (size_t)nx * ny);
  if (!u) return 1;
  /* a dense grid is a tile that covers everything, without margins */
  h2d_tile_t whole = {nx, ny, 0, 0, nx, ny, 0, 0, ny};
  CHECK(h2d_cpu_init(u, &whole, H2D_INIT_REF_INT32, 0));
  snprintf(path, sizeof path, "%s/initial.dat", dir);
  CHECK(h2d_dump_prtdat(path, u, nx, ny));

  /* the given program: float coefficients, fp64 intermediates, unfused (gcc) */
  h2d_phys_t ph;
  memset(&ph, 0, sizeof ph);
  ph.cx = (double)0.1f;
  ph.cy = (double)0.1f;
  ph.boundary = H2D_BOUNDARY_FIXED;
  ph.arith = H2D_ARITH_STRICT;
  long done = (long)h2d_cpu_run(u, nx, ny, steps, &ph, H2D_NORM_NONE, 1, 0.0, 1, NULL);
  if (done != steps) {
    fprintf(stderr, "h2d_cpu_run: %s\n", h2d_last_error());
    return 1;
  }
  snprintf(path, sizeof path, "%s/final.dat", dir);
  CHECK(h2d_dump_prtdat(path, u, nx, ny));
  printf("u(centre) after %ld steps = %.6e\n", done, (double)u[(size_t)(nx / 2) * ny + ny / 2]);

  /* again, stopping when the sum of squared changes of a checked step drops below eps */
  CHECK(h2d_cpu_init(u, &whole, H2D_INIT_EXACT, 0));
  ph.cx = ph.cy = 0.1;
  ph.boundary = H2D_BOUNDARY_GHOST_RING;
  double last = -1.0;
  done = (long)h2d_cpu_run(u, nx, ny, 1000000, &ph, H2D_NORM_L2SQ, 20, 0.1, 2, &last);
  printf("Exiting after %ld iterations (last norm %.3e)\n", done, last);
  /* the grad1612 dump pair: raw fp32 file written tile by tile, then the text form */
  snprintf(path, sizeof path, "%s/final_binary.dat", dir);
  h2d_tile_t top = {nx, ny, 0, 0, nx / 2 > 0 ? nx / 2 : nx, ny, 0, 0, ny};
  CHECK(h2d_dump_binary_tile(path, u, ny, &top, 1));
  if (top.rows < nx) {
    h2d_tile_t bottom = {nx, ny, top.rows, 0, nx - top.rows, ny, 0, 0, ny};
    CHECK(h2d_dump_binary_tile(path, u + (size_t)top.rows * ny, ny, &bottom, 0));
  }
  char text[4096];
  snprintf(text, sizeof text, "%s/final_rowmajor.dat", dir);
  CHECK(h2d_binary_to_text(path, text, nx, ny, 0));
  free(u);
  return 0;
}
