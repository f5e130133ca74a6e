// Exercises the CPU library (oracle + dump writers) under AddressSanitizer /
// UndefinedBehaviorSanitizer.  Built and run by tests/test_sanitizers.py; exits
// non-zero on a wrong result, the sanitizers abort on a bad access.
//
// The reference has no sanitizer builds at all ("race
// detection / sanitizers: none"); its init overflows a signed int
// (mpi_heat2Dn.c:247), which is exactly what UBSan would flag there.  Here the
// wrapping modes are computed in unsigned arithmetic, so this driver must stay
// silent even at 640x640 where the reference product overflows.
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <vector>
#include "heat2d.h"

static int failures = 0;
#define CHECK(cond)                                                        \
  do {                                                                     \
    if (!(cond)) {                                                         \
      std::fprintf(stderr, "%s:%d: CHECK failed: %s\n", __FILE__, __LINE__, #cond); \
      ++failures;                                                          \
    }                                                                      \
  } while (0)

static h2d_phys_t phys(int boundary, int arith) {
  h2d_phys_t p;
  std::memset(&p, 0, sizeof p);
  p.cx = 0.1;
  p.cy = 0.1;
  p.boundary = boundary;
  p.arith = arith;
  return p;
}

static std::vector<float> dense_init(int nx, int ny, int mode) {
  std::vector<float> u((size_t)nx * ny);
  h2d_tile_t t = {nx, ny, 0, 0, nx, ny, 0, 0, ny};
  CHECK(h2d_cpu_init(u.data(), &t, mode, 0) == H2D_OK);
  return u;
}

// Whole-grid run == tile-by-tile stepping on padded arrays with hand-copied ghosts.
static void tiles_match_dense(int nx, int ny, int px, int py, int boundary, int arith) {
  const int steps = 5;
  h2d_phys_t p = phys(boundary, arith);
  std::vector<float> ref = dense_init(nx, ny, H2D_INIT_EXACT);
  CHECK(h2d_cpu_run(ref.data(), nx, ny, steps, &p, H2D_NORM_NONE, 1, 0.0, 1, nullptr) == steps);

  std::vector<float> cur = dense_init(nx, ny, H2D_INIT_EXACT), nxt(cur.size());
  const int hr = 1, hc = 1;
  for (int s = 0; s < steps; ++s) {
    for (int bx = 0; bx < px; ++bx)
      for (int by = 0; by < py; ++by) {
        const int x0 = bx * nx / px, x1 = (bx + 1) * nx / px;
        const int y0 = by * ny / py, y1 = (by + 1) * ny / py;
        const int rows = x1 - x0, cols = y1 - y0;
        const int64_t pitch = cols + 2 * hc;
        h2d_tile_t t = {nx, ny, x0, y0, rows, cols, hr, hc, pitch};
        std::vector<float> a((size_t)(rows + 2 * hr) * pitch, 0.f), b(a.size(), -777.f);
        for (int i = -hr; i < rows + hr; ++i)
          for (int j = -hc; j < cols + hc; ++j) {
            const int gx = x0 + i, gy = y0 + j;
            if (gx < 0 || gx >= nx || gy < 0 || gy >= ny) continue;
            a[(size_t)(i + hr) * pitch + hc + j] = cur[(size_t)gx * ny + gy];
          }
        CHECK(h2d_cpu_step_tile(a.data(), b.data(), &t, &p) == H2D_OK);
        for (int i = 0; i < rows; ++i)
          for (int j = 0; j < cols; ++j)
            nxt[(size_t)(x0 + i) * ny + y0 + j] = b[(size_t)(i + hr) * pitch + hc + j];
      }
    cur.swap(nxt);
  }
  CHECK(std::memcmp(cur.data(), ref.data(), ref.size() * sizeof(float)) == 0);
}

static std::string slurp(const std::string& path) {
  std::string out;
  if (FILE* f = std::fopen(path.c_str(), "rb")) {
    char buf[4096];
    size_t n;
    while ((n = std::fread(buf, 1, sizeof buf, f)) > 0) out.append(buf, n);
    std::fclose(f);
  }
  return out;
}

int main(int argc, char** argv) {
  const std::string dir = argc > 1 ? argv[1] : ".";

  // Every init mode, including the sizes where the 32-bit product wraps.
  for (int mode : {H2D_INIT_REF_INT32, H2D_INIT_REF_UINT32, H2D_INIT_EXACT, H2D_INIT_ZERO})
    for (int n : {1, 2, 10, 433, 640}) {
      std::vector<float> u = dense_init(n, n + 3, mode);
      CHECK(u[0] == 0.f);
      if (mode == H2D_INIT_EXACT && n >= 10) CHECK(u[(size_t)(n / 2) * (n + 3) + n / 2] > 0.f);
    }
  // Margins that stick out of the grid are zero, the ones inside hold the neighbour's value.
  {
    h2d_tile_t t = {12, 9, 4, 0, 4, 9, 2, 3, 9 + 6};
    std::vector<float> a((size_t)(4 + 4) * 15, -1.f);
    CHECK(h2d_cpu_init(a.data(), &t, H2D_INIT_EXACT, 1) == H2D_OK);
    CHECK(a[(size_t)2 * 15 + 0] == 0.f);                    // column -3: outside
    CHECK(a[(size_t)0 * 15 + 3 + 4] == 2.f * 9 * 4 * 4);    // row 2 of the grid, column 4
  }
  // Bad arguments are refused, not dereferenced.
  {
    h2d_tile_t bad = {8, 8, 6, 0, 4, 8, 1, 1, 10};
    float x = 0.f;
    CHECK(h2d_cpu_init(&x, &bad, H2D_INIT_EXACT, 0) == H2D_EINVAL);
    CHECK(h2d_cpu_init(nullptr, &bad, H2D_INIT_EXACT, 0) == H2D_EINVAL);
    CHECK(std::strlen(h2d_last_error()) > 0);
  }

  for (int boundary : {H2D_BOUNDARY_FIXED, H2D_BOUNDARY_GHOST_RING})
    for (int arith : {H2D_ARITH_FP32, H2D_ARITH_STRICT, H2D_ARITH_STRICT_FMA}) {
      tiles_match_dense(13, 11, 1, 1, boundary, arith);
      tiles_match_dense(13, 11, 3, 2, boundary, arith);
      tiles_match_dense(5, 4, 5, 4, boundary, arith);  // one cell per tile
    }

  // Degenerate grids: nothing is active with fixed boundaries.
  for (int nx : {1, 2})
    for (int ny : {1, 2, 7}) {
      h2d_phys_t p = phys(H2D_BOUNDARY_FIXED, H2D_ARITH_STRICT);
      std::vector<float> u((size_t)nx * ny, 3.f), v = u;
      CHECK(h2d_cpu_run(u.data(), nx, ny, 3, &p, H2D_NORM_NONE, 1, 0.0, 1, nullptr) == 3);
      CHECK(u == v);
    }

  // Convergence: pre-step field, both norms, threaded == serial.
  for (int norm : {H2D_NORM_L2SQ, H2D_NORM_LINF}) {
    h2d_phys_t p = phys(H2D_BOUNDARY_FIXED, H2D_ARITH_STRICT);
    std::vector<float> a = dense_init(24, 20, H2D_INIT_EXACT), b = a;
    double na = -1, nb = -1;
    const int64_t sa = h2d_cpu_run(a.data(), 24, 20, 4000, &p, norm, 20, 0.5, 1, &na);
    const int64_t sb = h2d_cpu_run(b.data(), 24, 20, 4000, &p, norm, 20, 0.5, 3, &nb);
    CHECK(sa == sb && sa % 20 == 0 && sa < 4000);
    CHECK(a == b);
    CHECK(na < 0.5 && (norm == H2D_NORM_LINF ? na == nb : std::fabs(na - nb) <= 1e-9 * na + 1e-30));
  }

  // Dump writers: three formats + tile-wise binary + conversion back to text.
  {
    const int nx = 7, ny = 5;
    std::vector<float> u = dense_init(nx, ny, H2D_INIT_EXACT);
    const std::string a = dir + "/a.dat", b = dir + "/b.dat", bin = dir + "/u.bin", c = dir + "/c.dat",
                      d = dir + "/d.dat";
    CHECK(h2d_dump_prtdat(a.c_str(), u.data(), nx, ny) == H2D_OK);
    CHECK(h2d_dump_rowmajor(b.c_str(), u.data(), nx, ny) == H2D_OK);
    // two tiles (rows 0-3, rows 4-6), second written first
    h2d_tile_t t1 = {nx, ny, 4, 0, 3, ny, 0, 0, ny}, t0 = {nx, ny, 0, 0, 4, ny, 0, 0, ny};
    CHECK(h2d_dump_binary_tile(bin.c_str(), u.data() + 4 * ny, ny, &t1, 1) == H2D_OK);
    CHECK(h2d_dump_binary_tile(bin.c_str(), u.data(), ny, &t0, 0) == H2D_OK);
    const std::string raw = slurp(bin);
    CHECK(raw.size() == u.size() * sizeof(float));
    CHECK(std::memcmp(raw.data(), u.data(), raw.size()) == 0);
    CHECK(h2d_binary_to_text(bin.c_str(), c.c_str(), nx, ny, 0) == H2D_OK);
    CHECK(h2d_binary_to_text(bin.c_str(), d.c_str(), nx, ny, 1) == H2D_OK);
    CHECK(slurp(c) == slurp(b) && !slurp(b).empty());
    CHECK(slurp(d) == slurp(a) && !slurp(a).empty());
    CHECK(h2d_dump_prtdat((dir + "/no/such/dir/x").c_str(), u.data(), nx, ny) != H2D_OK);
    CHECK(h2d_binary_to_text((dir + "/missing.bin").c_str(), c.c_str(), nx, ny, 0) != H2D_OK);
  }

  if (failures) {
    std::fprintf(stderr, "%d check(s) failed\n", failures);
    return 1;
  }
  std::puts("sanitize_driver: ok");
  return 0;
}
