// Dump writers: the two text formats and the raw binary format of the
// reference programs.  Byte-for-byte compatible with
//   prtdat()                          mpi_heat2Dn.c:253-268
//   the rank-0 text conversion        grad1612_mpi_heat.c:191-203, 286-298
//   initial_binary.dat / final_binary.dat (MPI-IO subarray write)
//                                     grad1612_mpi_heat.c:178-190, 283-285
// The MPI-IO collective is replaced by one pwrite() per owned row at the
// row's offset in the shared file; ranks never overlap, so no locking.
#include <cerrno>
#include <cmath>
#include <cstdint>
#include <cstdio>
#include <cstring>
#include <string>
#include <vector>
#include <fcntl.h>
#include <sys/stat.h>
#include <unistd.h>
#include "heat2d.h"
#include "host_common.h"

namespace {

// "%6.1f" of a float (promoted to double, as printf does), without printf.
// v * 10 is exact in double (24-bit significand times 10), so rounding it to an integer
// with ties-to-even (nearbyint in the default rounding mode) is the same decision printf
// takes on the exact decimal expansion.  The sign is printed whenever the sign bit is set,
// also for results that round to zero ("-0.0"), as printf does.  Huge values, infinities
// and NaNs take the snprintf path.  Returns the number of characters written (< 64).
// tests/test_dump.py compares it with snprintf on tens of millions of bit patterns.
inline int fmt61(char* dst, float v) {
  const double t = (double)v * 10.0;
  if (!(std::fabs(t) < 9.0e18)) return snprintf(dst, 64, "%6.1f", (double)v);
  uint64_t n = (uint64_t)std::fabs(std::nearbyint(t));
  const bool neg = std::signbit(v);
  char tmp[24];
  int len = 0;
  const int frac = (int)(n % 10u);
  n /= 10u;
  do {
    tmp[len++] = (char)('0' + (int)(n % 10u));
    n /= 10u;
  } while (n);
  const int total = (neg ? 1 : 0) + len + 2;
  char* p = dst;
  for (int pad = total; pad < 6; ++pad) *p++ = ' ';
  if (neg) *p++ = '-';
  while (len) *p++ = tmp[--len];
  *p++ = '.';
  *p++ = (char)('0' + frac);
  return (int)(p - dst);
}

// Text files are written line by line; the lines of a batch are formatted in parallel
// (OpenMP, when the library was built with it) and written in order.  `line(i, out)`
// appends line i to `out`.  The reference formats and writes value by value with fprintf
// (mpi_heat2Dn.c:259-266, grad1612_mpi_heat.c:195-200).
template <class LineFn>
int write_lines(const char* path, int64_t n_lines, size_t line_bytes_hint, LineFn line) {
  FILE* fp = fopen(path, "w");
  if (!fp) return h2d_fail(H2D_EIO, "cannot open %s: %s", path, strerror(errno));
  int64_t batch = (int64_t)((64u << 20) / (line_bytes_hint ? line_bytes_hint : 1));   // ~64 MiB of text
  if (batch < 1) batch = 1;
  if (batch > 4096) batch = 4096;
  std::vector<std::string> text((size_t)batch);
  bool ok = true;
  for (int64_t base = 0; base < n_lines && ok; base += batch) {
    const int64_t cnt = n_lines - base < batch ? n_lines - base : batch;
#pragma omp parallel for schedule(dynamic, 1) if (cnt > 1)
    for (int64_t i = 0; i < cnt; ++i) {
      std::string& out = text[(size_t)i];
      out.clear();
      if (out.capacity() < line_bytes_hint) out.reserve(line_bytes_hint);
      line(base + i, out);
    }
    for (int64_t i = 0; i < cnt && ok; ++i)
      ok = fwrite(text[(size_t)i].data(), 1, text[(size_t)i].size(), fp) == text[(size_t)i].size();
  }
  if (fclose(fp) != 0) ok = false;
  if (!ok) return h2d_fail(H2D_EIO, "write to %s failed", path);
  return H2D_OK;
}

}  // namespace

extern "C" {

const char* h2d_last_error(void) { return h2d_err_buf(); }
const char* h2d_version(void) { return H2D_VERSION_STRING; }

int h2d_dump_prtdat(const char* path, const float* u, int32_t nx, int32_t ny) {
  if (!path || !u || nx < 1 || ny < 1) return h2d_fail(H2D_EINVAL, "h2d_dump_prtdat: bad argument");
  // line i holds y = ny-1-i, x ascending
  return write_lines(path, ny, (size_t)nx * 8, [=](int64_t i, std::string& out) {
    const int64_t y = (int64_t)ny - 1 - i;
    char tmp[64];
    for (int32_t x = 0; x < nx; ++x) {
      int n = fmt61(tmp, u[(int64_t)x * ny + y]);
      tmp[n++] = (x != nx - 1) ? ' ' : '\n';
      out.append(tmp, (size_t)n);
    }
  });
}

int h2d_dump_rowmajor(const char* path, const float* u, int32_t nx, int32_t ny) {
  if (!path || !u || nx < 1 || ny < 1)
    return h2d_fail(H2D_EINVAL, "h2d_dump_rowmajor: bad argument");
  return write_lines(path, nx, (size_t)ny * 8 + 1, [=](int64_t x, std::string& out) {
    const float* row = u + x * ny;
    char tmp[64];
    for (int32_t y = 0; y < ny; ++y) {
      int n = fmt61(tmp, row[y]);
      tmp[n++] = ' ';
      out.append(tmp, (size_t)n);
    }
    out.push_back('\n');
  });
}

/* Test hooks: "%6.1f" as the dump writers format it (returns the length), and a bulk
 * comparison with snprintf over the float bit patterns first, first + stride, ...
 * (count of them): returns the number of mismatches, *first_bad = the first one.      */
int h2d_format_6_1f(float v, char* out64) { return out64 ? fmt61(out64, v) : 0; }

int64_t h2d_format_6_1f_check(uint32_t first, uint64_t count, uint32_t stride, uint32_t* first_bad) {
  int64_t bad = 0;
  uint64_t lowest = UINT64_MAX;
#pragma omp parallel for schedule(static) reduction(+ : bad) reduction(min : lowest)
  for (int64_t i = 0; i < (int64_t)count; ++i) {
    const uint32_t bits = first + (uint32_t)((uint64_t)i * stride);
    float v;
    std::memcpy(&v, &bits, sizeof v);
    char a[64], b[64];
    const int na = fmt61(a, v);
    const int nb = snprintf(b, sizeof b, "%6.1f", (double)v);
    if (na != nb || std::memcmp(a, b, (size_t)na) != 0) {
      ++bad;
      if ((uint64_t)i < lowest) lowest = (uint64_t)i;
    }
  }
  if (first_bad && bad) *first_bad = first + (uint32_t)(lowest * stride);
  return bad;
}

int h2d_dump_binary_tile(const char* path, const float* owned, int64_t owned_pitch,
                         const h2d_tile_t* t, int truncate) {
  if (!path || !owned || !t) return h2d_fail(H2D_EINVAL, "h2d_dump_binary_tile: null argument");
  if (t->rows < 1 || t->cols < 1 || owned_pitch < t->cols)
    return h2d_fail(H2D_EINVAL, "h2d_dump_binary_tile: bad extent");
  // No O_TRUNC: like MPI_MODE_WRONLY|MPI_MODE_CREATE the file keeps stale tail
  // bytes unless the caller asks for the exact length.
  int fd = open(path, O_WRONLY | O_CREAT, 0644);
  if (fd < 0) return h2d_fail(H2D_EIO, "cannot open %s: %s", path, strerror(errno));
  if (truncate) {
    if (ftruncate(fd, (off_t)t->nx * t->ny * 4) != 0) {
      close(fd);
      return h2d_fail(H2D_EIO, "ftruncate %s: %s", path, strerror(errno));
    }
  }
  for (int32_t i = 0; i < t->rows; ++i) {
    const char* src = (const char*)(owned + (int64_t)i * owned_pitch);
    size_t left = (size_t)t->cols * 4;
    off_t off = ((off_t)(t->x0 + i) * t->ny + t->y0) * 4;
    while (left) {
      ssize_t w = pwrite(fd, src, left, off);
      if (w < 0) {
        if (errno == EINTR) continue;
        close(fd);
        return h2d_fail(H2D_EIO, "pwrite %s: %s", path, strerror(errno));
      }
      src += w;
      off += w;
      left -= (size_t)w;
    }
  }
  if (close(fd) != 0) return h2d_fail(H2D_EIO, "close %s: %s", path, strerror(errno));
  return H2D_OK;
}

int h2d_binary_to_text(const char* bin_path, const char* txt_path, int32_t nx, int32_t ny,
                       int format) {
  if (!bin_path || !txt_path || nx < 1 || ny < 1)
    return h2d_fail(H2D_EINVAL, "h2d_binary_to_text: bad argument");
  FILE* fs = fopen(bin_path, "rb");
  if (!fs) return h2d_fail(H2D_EIO, "cannot open %s: %s", bin_path, strerror(errno));
  std::vector<float> u((size_t)nx * ny);
  size_t got = fread(u.data(), sizeof(float), u.size(), fs);
  fclose(fs);
  if (got != u.size())
    return h2d_fail(H2D_EIO, "%s: expected %zu floats, read %zu", bin_path, u.size(), got);
  return format == 1 ? h2d_dump_prtdat(txt_path, u.data(), nx, ny)
                     : h2d_dump_rowmajor(txt_path, u.data(), nx, ny);
}

}  // extern "C"
