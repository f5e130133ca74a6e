// Dump writers: the two text formats and the raw binary format of the
// reference programs.  Byte-for-byte compatible with
//   prtdat()                          mpi_heat2Dn.c:253-268
//   the rank-0 text conversion        grad1612_mpi_heat.c:191-203, 286-298
//   initial_binary.dat / final_binary.dat (MPI-IO subarray write)
//                                     grad1612_mpi_heat.c:178-190, 283-285
// The MPI-IO collective is replaced by one pwrite() per owned row at the
// row's offset in the shared file; ranks never overlap, so no locking.
#include <cerrno>
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

// "%6.1f" of a float (promoted to double, as printf does).  Values are at
// most ~1e38, so 64 bytes is plenty.
inline int fmt61(char* dst, float v) { return snprintf(dst, 64, "%6.1f", (double)v); }

struct BufferedFile {
  FILE* fp = nullptr;
  std::vector<char> buf;
  size_t used = 0;
  explicit BufferedFile(const char* path) : buf(1 << 20) { fp = fopen(path, "w"); }
  ~BufferedFile() { close(); }
  bool ok() const { return fp != nullptr; }
  char* reserve(size_t n) {
    if (used + n > buf.size()) flush();
    return buf.data() + used;
  }
  void commit(size_t n) { used += n; }
  void flush() {
    if (fp && used) fwrite(buf.data(), 1, used, fp);
    used = 0;
  }
  int close() {
    int rc = 0;
    if (fp) {
      flush();
      rc = fclose(fp);
      fp = nullptr;
    }
    return rc;
  }
};

}  // namespace

extern "C" {

const char* h2d_last_error(void) { return h2d_err_buf(); }
const char* h2d_version(void) { return H2D_VERSION_STRING; }

int h2d_dump_prtdat(const char* path, const float* u, int32_t nx, int32_t ny) {
  if (!path || !u || nx < 1 || ny < 1) return h2d_fail(H2D_EINVAL, "h2d_dump_prtdat: bad argument");
  BufferedFile f(path);
  if (!f.ok()) return h2d_fail(H2D_EIO, "cannot open %s: %s", path, strerror(errno));
  for (int32_t y = ny - 1; y >= 0; --y) {
    for (int32_t x = 0; x < nx; ++x) {
      char* p = f.reserve(80);
      int n = fmt61(p, u[(int64_t)x * ny + y]);
      p[n++] = (x != nx - 1) ? ' ' : '\n';
      f.commit((size_t)n);
    }
  }
  if (f.close() != 0) return h2d_fail(H2D_EIO, "write to %s failed", path);
  return H2D_OK;
}

int h2d_dump_rowmajor(const char* path, const float* u, int32_t nx, int32_t ny) {
  if (!path || !u || nx < 1 || ny < 1)
    return h2d_fail(H2D_EINVAL, "h2d_dump_rowmajor: bad argument");
  BufferedFile f(path);
  if (!f.ok()) return h2d_fail(H2D_EIO, "cannot open %s: %s", path, strerror(errno));
  for (int32_t x = 0; x < nx; ++x) {
    for (int32_t y = 0; y < ny; ++y) {
      char* p = f.reserve(80);
      int n = fmt61(p, u[(int64_t)x * ny + y]);
      p[n++] = ' ';
      f.commit((size_t)n);
    }
    char* p = f.reserve(1);
    p[0] = '\n';
    f.commit(1);
  }
  if (f.close() != 0) return h2d_fail(H2D_EIO, "write to %s failed", path);
  return H2D_OK;
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
