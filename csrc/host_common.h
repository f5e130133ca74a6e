// Host-side helpers shared by the CPU and CUDA libraries (header-only).
#pragma once
#include <cstdarg>
#include <cstdio>
#include <cstdint>
#include "heat2d.h"

#define H2D_VERSION_STRING "heat2d-b200 0.1"

inline char* h2d_err_buf() {
  static thread_local char buf[512] = "";
  return buf;
}

inline int h2d_fail(int code, const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(h2d_err_buf(), 512, fmt, ap);
  va_end(ap);
  return code;
}

inline int h2d_check_tile(const h2d_tile_t* t) {
  if (t->nx < 1 || t->ny < 1) return h2d_fail(H2D_EINVAL, "tile: grid %dx%d is empty", t->nx, t->ny);
  if (t->rows < 1 || t->cols < 1)
    return h2d_fail(H2D_EINVAL, "tile: owned extent %dx%d is empty", t->rows, t->cols);
  if (t->x0 < 0 || t->y0 < 0 || t->x0 + t->rows > t->nx || t->y0 + t->cols > t->ny)
    return h2d_fail(H2D_EINVAL, "tile: [%d+%d, %d+%d] outside the %dx%d grid", t->x0, t->rows,
                    t->y0, t->cols, t->nx, t->ny);
  if (t->halo_r < 0 || t->halo_c < 0) return h2d_fail(H2D_EINVAL, "tile: negative margin");
  if (t->pitch < (int64_t)t->cols + 2 * (int64_t)t->halo_c)
    return h2d_fail(H2D_EINVAL, "tile: pitch %lld < cols + 2*halo_c", (long long)t->pitch);
  return H2D_OK;
}

// x(nx-x-1) y(ny-y-1) peaks at (nx/2)^2 (ny/2)^2: does it fit a signed 64-bit integer?
#ifdef __CUDACC__
__host__ __device__
#endif
inline bool h2d_exact_fits_int64(int32_t nx, int32_t ny) {
  const double peak = 0.25 * (double)nx * (double)nx * 0.25 * (double)ny * (double)ny;
  return peak < 4.0e18;
}

// Value of the initial condition at global (x, y); defined in oracle.cpp and
// duplicated (device) in init.cu.
float h2d_init_value(int32_t nx, int32_t ny, int32_t x, int32_t y, int mode);
