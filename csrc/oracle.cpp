// CPU oracle for heat2d-b200: the spec the GPU kernels are tested against.
//
// Implements every numerical mode of SURVEY.md App. A:
//   arithmetic  fp32 | strict (fp64 intermediates, unfused) | strict-fma
//   boundary    fixed edge cells | ghost ring
//   init        int32-wrapping signed / unsigned | exact
// Parity notes: update() mpi_heat2Dn.c:225-237, the sweeps of
// grad1612_mpi_heat.c:239-259, update<<<>>> grad1612_cuda_heat.cu:55-62,
// inidat() mpi_heat2Dn.c:242-248, convergence grad1612_mpi_heat.c:262-271.
//
// Must be compiled with -ffp-contract=off: the "strict" variant depends on the
// compiler NOT fusing a*b+c, the "strict-fma" variant calls fma() explicitly.
#include <cmath>
#include <cstdint>
#include <cstdio>
#include <cstring>
#include <vector>
#include <algorithm>
#ifdef _OPENMP
#include <omp.h>
#endif
#include "heat2d.h"
#include "host_common.h"

namespace {

struct Coef {
  double cx, cy;      // strict modes
  float cxf, cyf, a0; // fp32 mode
};

inline Coef make_coef(const h2d_phys_t* p) {
  Coef c;
  c.cx = p->cx;
  c.cy = p->cy;
  c.cxf = (float)p->cx;
  c.cyf = (float)p->cy;
  c.a0 = (float)(1.0 - 2.0 * p->cx - 2.0 * p->cy);
  return c;
}

// One grid point.  xm/xp: neighbours in the slow (row) index, ym/yp: in the
// contiguous index.
// always_inline: the row kernels below are compiled once per instruction set, and the
// fma / conversion code must be generated for the set of the kernel it ends up in.
template <int ARITH>
__attribute__((always_inline)) inline float point(float c, float xm, float xp, float ym, float yp,
                                                  const Coef& k) {
  float ew = xp + xm;  // fl32, as the C expression `u[i+1][j] + u[i-1][j]`
  float ns = yp + ym;
  if (ARITH == H2D_ARITH_FP32) {
    // Same operation order as point<H2D_ARITH_FP32>() in device_common.cuh.
    float t = k.a0 * c;
    t = fmaf(k.cyf, ns, t);
    return fmaf(k.cxf, ew, t);
  } else {
    double dc = (double)c;
    double t1 = (double)ew - 2.0 * dc;  // 2*c is exact, so this is one rounding
    double t2 = (double)ns - 2.0 * dc;
    double r;
    if (ARITH == H2D_ARITH_STRICT_FMA) {
      r = fma(k.cy, t2, fma(k.cx, t1, dc));
    } else {
      double a = k.cx * t1;
      double b = dc + a;
      double d = k.cy * t2;
      r = b + d;
    }
    return (float)r;
  }
}

struct Active {
  int32_t x_lo, x_hi, y_lo, y_hi;  // inclusive global range of updated cells
};

inline Active active_range(int32_t nx, int32_t ny, int boundary) {
  if (boundary == H2D_BOUNDARY_FIXED) return {1, nx - 2, 1, ny - 2};
  return {0, nx - 1, 0, ny - 1};
}

// ---------------------------------------------------------------- row kernels
// One output row, columns [y0, y1]: every operand is a plain array read (rows outside
// the grid are a row of zeros, columns outside it are handled by the caller), so the loop
// has no branches and vectorises.  The arithmetic is point<ARITH>() element by element —
// vector instructions round each lane exactly like the scalar ones — and -ffp-contract=off
// still forbids fusing in the strict variant.  Each kernel is compiled three times
// (baseline x86-64, x86-64-v3 = AVX2 + FMA, x86-64-v4 = AVX-512) and the loader picks one at run time
// (target_clones): the library is built on one box and runs on another.
//
// The reference's sweeps are the scalar loops of mpi_heat2Dn.c:225-237 and
// grad1612_mpi_heat.c:239-259; gcc -O3 vectorises those too, which is why its single
// rank (0.50 GLUP/s on this box's CPU) used to beat this oracle (0.19) before this rewrite.
#if defined(__x86_64__) && defined(__GNUC__) && !defined(__clang__) && !defined(H2D_NO_CLONES)
#define H2D_CLONES __attribute__((target_clones("default", "arch=x86-64-v3", "arch=x86-64-v4")))
#else
#define H2D_CLONES
#endif

#define H2D_ROW_KERNEL(NAME, ARITH)                                                             \
  H2D_CLONES void NAME(const float* __restrict__ row, const float* __restrict__ up,            \
                       const float* __restrict__ dn, float* __restrict__ out, int32_t y0,       \
                       int32_t y1, const Coef* kp) {                                            \
    const Coef k = *kp;                                                                         \
    _Pragma("omp simd")                                                                         \
    for (int32_t y = y0; y <= y1; ++y)                                                          \
      out[y] = point<ARITH>(row[y], up[y], dn[y], row[y - 1], row[y + 1], k);                   \
  }
H2D_ROW_KERNEL(row_fp32, H2D_ARITH_FP32)
H2D_ROW_KERNEL(row_strict, H2D_ARITH_STRICT)
H2D_ROW_KERNEL(row_strict_fma, H2D_ARITH_STRICT_FMA)
#undef H2D_ROW_KERNEL

typedef void (*row_fn_t)(const float*, const float*, const float*, float*, int32_t, int32_t, const Coef*);

template <int ARITH>
void sweep_dense(const float* __restrict__ s, float* __restrict__ d, int32_t nx, int32_t ny,
                 const Active& a, const Coef& k, int threads) {
  // Cells outside the active range keep their value.
  (void)threads;
  const row_fn_t fn = ARITH == H2D_ARITH_FP32 ? row_fp32 : ARITH == H2D_ARITH_STRICT ? row_strict : row_strict_fma;
  const std::vector<float> zeros((size_t)ny, 0.0f);   // the row above row 0 / below row nx-1
  const float* zrow = zeros.data();
  // dynamic: on a shared box one preempted thread must not hold up a whole static sweep
#pragma omp parallel for schedule(dynamic, 16) num_threads(threads) if (threads > 1)
  for (int32_t x = 0; x < nx; ++x) {
    const float* row = s + (int64_t)x * ny;
    float* out = d + (int64_t)x * ny;
    if (x < a.x_lo || x > a.x_hi || a.y_lo > a.y_hi) {
      std::memcpy(out, row, sizeof(float) * (size_t)ny);
      continue;
    }
    const float* up = (x > 0) ? row - ny : zrow;
    const float* dn = (x < nx - 1) ? row + ny : zrow;
    // inactive columns at both ends copy through
    for (int32_t y = 0; y < a.y_lo; ++y) out[y] = row[y];
    for (int32_t y = a.y_hi + 1; y < ny; ++y) out[y] = row[y];
    // active columns that have both horizontal neighbours inside the grid
    const int32_t y0 = a.y_lo > 1 ? a.y_lo : 1;
    const int32_t y1 = a.y_hi < ny - 2 ? a.y_hi : ny - 2;
    if (y0 <= y1) fn(row, up, dn, out, y0, y1, &k);
    // active grid-edge columns (ghost-ring mode): the missing neighbour reads 0
    if (a.y_lo == 0) out[0] = point<ARITH>(row[0], up[0], dn[0], 0.0f, ny > 1 ? row[1] : 0.0f, k);
    if (a.y_hi == ny - 1 && ny > 1)
      out[ny - 1] = point<ARITH>(row[ny - 1], up[ny - 1], dn[ny - 1], row[ny - 2], 0.0f, k);
  }
}

void sweep_dispatch(const float* s, float* d, int32_t nx, int32_t ny, const Active& a,
                    const Coef& k, int arith, int threads) {
  switch (arith) {
    case H2D_ARITH_FP32: sweep_dense<H2D_ARITH_FP32>(s, d, nx, ny, a, k, threads); break;
    case H2D_ARITH_STRICT: sweep_dense<H2D_ARITH_STRICT>(s, d, nx, ny, a, k, threads); break;
    default: sweep_dense<H2D_ARITH_STRICT_FMA>(s, d, nx, ny, a, k, threads); break;
  }
}

double norm_dense(const float* a, const float* b, int64_t n, int norm, int threads) {
  double acc = 0.0;
  (void)threads;
  if (norm == H2D_NORM_L2SQ) {
#pragma omp parallel for reduction(+ : acc) num_threads(threads) if (threads > 1)
    for (int64_t i = 0; i < n; ++i) {
      double d = (double)a[i] - (double)b[i];
      acc += d * d;
    }
  } else {
    // NaN propagates (as in the CUDA kernels): a diverged field never looks converged
    int has_nan = 0;
#pragma omp parallel for reduction(max : acc) reduction(| : has_nan) num_threads(threads) if (threads > 1)
    for (int64_t i = 0; i < n; ++i) {
      double d = std::fabs((double)a[i] - (double)b[i]);
      if (d != d) has_nan = 1;
      if (d > acc) acc = d;
    }
    if (has_nan) acc = std::nan("");
  }
  return acc;
}

}  // namespace

float h2d_init_value(int32_t nx, int32_t ny, int32_t x, int32_t y, int mode) {
  if (mode == H2D_INIT_ZERO) return 0.0f;
  if (x < 0 || x >= nx || y < 0 || y >= ny) return 0.0f;
  if (mode == H2D_INIT_EXACT) {
    // int64 is enough up to 65536^2 (peak 1.15e18 < 2^63), SURVEY 7.3.  Beyond that
    // (weak scaling to 131072 x 262144) the two factors, each exact below 2^53, are
    // multiplied in double: one extra rounding, identical on CPU and GPU.
    const int64_t a = (int64_t)x * (int64_t)(nx - x - 1);
    const int64_t b = (int64_t)y * (int64_t)(ny - y - 1);
    if (h2d_exact_fits_int64(nx, ny)) return (float)(a * b);
    return (float)((double)a * (double)b);
  }
  // ((x*(nx-x-1))*y)*(ny-y-1), left to right in 32 bits (mpi_heat2Dn.c:247).
  uint32_t p = (uint32_t)x * (uint32_t)(nx - x - 1);
  p *= (uint32_t)y;
  p *= (uint32_t)(ny - y - 1);
  if (mode == H2D_INIT_REF_UINT32) return (float)p;
  int32_t sp;
  std::memcpy(&sp, &p, sizeof sp);
  return (float)sp;
}

extern "C" {

int h2d_cpu_init(float* base, const h2d_tile_t* t, int init_mode, int with_margins) {
  if (!base || !t) return h2d_fail(H2D_EINVAL, "h2d_cpu_init: null argument");
  if (int rc = h2d_check_tile(t)) return rc;
  const int32_t r_lo = with_margins ? -t->halo_r : 0;
  const int32_t r_hi = with_margins ? t->rows + t->halo_r : t->rows;
  const int32_t c_lo = with_margins ? -t->halo_c : 0;
  const int32_t c_hi = with_margins ? t->cols + t->halo_c : t->cols;
  for (int32_t i = r_lo; i < r_hi; ++i) {
    float* row = base + (int64_t)(i + t->halo_r) * t->pitch + t->halo_c;
    for (int32_t j = c_lo; j < c_hi; ++j)
      row[j] = h2d_init_value(t->nx, t->ny, t->x0 + i, t->y0 + j, init_mode);
  }
  return H2D_OK;
}

int64_t h2d_cpu_run(float* u, int32_t nx, int32_t ny, int64_t steps, const h2d_phys_t* phys,
                    int norm, int64_t interval, double eps, int threads, double* last_norm) {
  if (!u || !phys || nx < 1 || ny < 1 || steps < 0)
    return h2d_fail(H2D_EINVAL, "h2d_cpu_run: bad argument");
  if (norm != H2D_NORM_NONE && interval < 1)
    return h2d_fail(H2D_EINVAL, "h2d_cpu_run: interval must be >= 1");
  const int64_t n = (int64_t)nx * ny;
  std::vector<float> other((size_t)n);
  float* buf[2] = {u, other.data()};
  const Coef k = make_coef(phys);
  const Active a = active_range(nx, ny, phys->boundary);
  int iz = 0;
  int64_t done = steps;
  if (threads < 1) threads = 1;
  for (int64_t s = 0; s < steps; ++s) {
    sweep_dispatch(buf[iz], buf[1 - iz], nx, ny, a, k, phys->arith, threads);
    if (norm != H2D_NORM_NONE && s % interval == 0) {
      double v = norm_dense(buf[1 - iz], buf[iz], n, norm, threads);
      if (last_norm) *last_norm = v;
      if (v < eps) {  // break before the flip: the result is the pre-step field
        done = s;
        break;
      }
    }
    iz = 1 - iz;
  }
  if (iz != 0) std::memcpy(u, buf[iz], sizeof(float) * (size_t)n);
  return done;
}

int h2d_cpu_step_tile(const float* src, float* dst, const h2d_tile_t* t, const h2d_phys_t* phys) {
  if (!src || !dst || !t || !phys) return h2d_fail(H2D_EINVAL, "h2d_cpu_step_tile: null argument");
  if (int rc = h2d_check_tile(t)) return rc;
  if (t->halo_r < 1 || t->halo_c < 1)
    return h2d_fail(H2D_EINVAL, "h2d_cpu_step_tile: needs a margin of at least 1");
  const Coef k = make_coef(phys);
  const Active a = active_range(t->nx, t->ny, phys->boundary);
  const row_fn_t fn = phys->arith == H2D_ARITH_FP32     ? row_fp32
                      : phys->arith == H2D_ARITH_STRICT ? row_strict
                                                        : row_strict_fma;
  // Rows above / below the GRID read as zeros whatever the margin holds; they are indexed
  // with tile-local columns -1 .. cols, hence the offset of one.
  const std::vector<float> zeros((size_t)t->cols + 2, 0.0f);
  const float* zrow = zeros.data() + 1;
  // tile-local range of active columns
  const int32_t j_lo = a.y_lo - t->y0 > 0 ? a.y_lo - t->y0 : 0;
  const int32_t j_hi = a.y_hi - t->y0 < t->cols - 1 ? a.y_hi - t->y0 : t->cols - 1;
  for (int32_t i = 0; i < t->rows; ++i) {
    const float* row = src + (int64_t)(i + t->halo_r) * t->pitch + t->halo_c;
    float* out = dst + (int64_t)(i + t->halo_r) * t->pitch + t->halo_c;
    const int32_t gx = t->x0 + i;
    if (gx < a.x_lo || gx > a.x_hi || j_lo > j_hi) {
      std::memcpy(out, row, sizeof(float) * (size_t)t->cols);
      continue;
    }
    const float* up = (gx > 0) ? row - t->pitch : zrow;
    const float* dn = (gx < t->nx - 1) ? row + t->pitch : zrow;
    for (int32_t j = 0; j < j_lo; ++j) out[j] = row[j];
    for (int32_t j = j_hi + 1; j < t->cols; ++j) out[j] = row[j];
    // Columns whose horizontal neighbours are inside the grid: the margin supplies them
    // at tile edges.  Grid-edge columns (ghost-ring mode) read a zero instead.
    int32_t j0 = j_lo, j1 = j_hi;
    if (t->y0 + j0 == 0) {
      const float yp = t->ny > 1 ? row[j0 + 1] : 0.0f;
      switch (phys->arith) {
        case H2D_ARITH_FP32: out[j0] = point<H2D_ARITH_FP32>(row[j0], up[j0], dn[j0], 0.0f, yp, k); break;
        case H2D_ARITH_STRICT: out[j0] = point<H2D_ARITH_STRICT>(row[j0], up[j0], dn[j0], 0.0f, yp, k); break;
        default: out[j0] = point<H2D_ARITH_STRICT_FMA>(row[j0], up[j0], dn[j0], 0.0f, yp, k); break;
      }
      ++j0;
    }
    if (j1 >= j0 && t->y0 + j1 == t->ny - 1) {
      switch (phys->arith) {
        case H2D_ARITH_FP32: out[j1] = point<H2D_ARITH_FP32>(row[j1], up[j1], dn[j1], row[j1 - 1], 0.0f, k); break;
        case H2D_ARITH_STRICT: out[j1] = point<H2D_ARITH_STRICT>(row[j1], up[j1], dn[j1], row[j1 - 1], 0.0f, k); break;
        default: out[j1] = point<H2D_ARITH_STRICT_FMA>(row[j1], up[j1], dn[j1], row[j1 - 1], 0.0f, k); break;
      }
      --j1;
    }
    if (j0 <= j1) fn(row, up, dn, out, j0, j1, &k);
  }
  return H2D_OK;
}

}  // extern "C"
