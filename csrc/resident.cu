// resident solver — small grids: the whole field lives in the shared memory of
// ONE thread-block cluster for the entire run.  Every CTA owns a band of rows
// (double-buffered in its own shared memory); after each step the first and
// last row of the band are stored straight into the neighbouring CTAs' ghost
// rows through distributed shared memory, followed by one cluster barrier.
// All `steps` steps run inside a single launch: no per-step launch latency.
//
// This is the regime where the reference is launch-bound: 9.15 us per step at
// 80x64 with one update<<<>>> launch per step (grad1612_cuda_heat.cu:82-85,
// Report.pdf p26); and it is the in-GPU analogue of the MPI ghost-row exchange
// of mpi_heat2Dn.c:179-192 (rows only: bands span the full width).
#include <cooperative_groups.h>
#include "device_common.cuh"

namespace cg = cooperative_groups;

namespace h2d {

constexpr int RES_THREADS = 1024;
constexpr int RES_MAX_SMEM = 227 * 1024;

// Shared-memory layout of one CTA: two buffers of (band + 2) rows x nyp floats;
// owned column j lives at float RES_LPAD + j of a row, so 4-column groups are
// 16-byte aligned; RES_LPAD floats of zeros on the left, >= 4 on the right.
constexpr int RES_LPAD = 4;
__host__ __device__ inline int res_pitch(int ny) { return ((ny + 3) & ~3) + 2 * RES_LPAD; }

template <int ARITH>
__global__ void __launch_bounds__(RES_THREADS, 1)
resident_kernel(const float* __restrict__ src, float* __restrict__ dst, int64_t pitch,
                int64_t steps, const __grid_constant__ Geo g, int band) {
  cg::cluster_group cluster = cg::this_cluster();
  const int rank = (int)cluster.block_rank();
  const int nblk = (int)cluster.num_blocks();
  extern __shared__ __align__(16) float sm[];
  const int nx = g.nx, ny = g.ny;
  const int nyp = res_pitch(ny);
  const int bstride = (band + 2) * nyp;
  float* buf0 = sm;
  float* buf1 = sm + bstride;
  const int x_lo = rank * band;
  const int x_hi = min(x_lo + band, nx);
  const int my_rows = x_hi - x_lo;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  constexpr int NW = RES_THREADS / 32;

  for (int i = threadIdx.x; i < 2 * bstride; i += RES_THREADS) sm[i] = 0.f;
  __syncthreads();
  // Owned rows plus the two ghost rows (when they are inside the grid).
  for (int r = warp; r < my_rows + 2; r += NW) {
    const int gx = x_lo + r - 1;
    if (gx < 0 || gx >= nx) continue;
    const float* s = src + (int64_t)gx * pitch;
    for (int j = lane; j < ny; j += 32) buf0[r * nyp + RES_LPAD + j] = s[j];
  }
  __syncthreads();

  float* A = buf0;
  float* B = buf1;
  float* up_ghost[2] = {nullptr, nullptr};  // neighbour's bottom ghost row, per buffer
  float* dn_ghost[2] = {nullptr, nullptr};  // neighbour's top ghost row, per buffer
  if (rank > 0) {
    up_ghost[0] = cluster.map_shared_rank(buf0, rank - 1) + (band + 1) * nyp;
    up_ghost[1] = cluster.map_shared_rank(buf1, rank - 1) + (band + 1) * nyp;
  }
  if (rank + 1 < nblk) {
    dn_ghost[0] = cluster.map_shared_rank(buf0, rank + 1);
    dn_ghost[1] = cluster.map_shared_rank(buf1, rank + 1);
  }
  cluster.sync();

  // Work items: (owned row, group of 4 columns), strided over the 1024 threads;
  // one LDS.128 each for the rows above / at / below, two scalar loads for the
  // horizontal neighbours outside the group, packed fp32 math, one STS.128.
  const int groups = (ny + 3) >> 2;
  const int n_items = my_rows * groups;
  int cur = 0;
  for (int64_t s = 0; s < steps; ++s) {
    for (int it = threadIdx.x; it < n_items; it += RES_THREADS) {
      const int r = it / groups;            // 0-based owned row
      const int gq = it - r * groups;       // group index
      const int j = gq << 2;                // first column of the group
      const int gx = x_lo + r;
      const float* c = A + (r + 1) * nyp + RES_LPAD + j;
      const float4 cv = *reinterpret_cast<const float4*>(c);
      const float4 uv = *reinterpret_cast<const float4*>(c - nyp);
      const float4 dv = *reinterpret_cast<const float4*>(c + nyp);
      const float L = c[-1], R = c[4];
      float4 o = point4<ARITH>(uv, cv, dv, L, R, g);
      const bool ract = gx >= g.ax_lo && gx <= g.ax_hi;
      o.x = (ract && j + 0 >= g.ay_lo && j + 0 <= g.ay_hi) ? o.x : cv.x;
      o.y = (ract && j + 1 >= g.ay_lo && j + 1 <= g.ay_hi) ? o.y : cv.y;
      o.z = (ract && j + 2 >= g.ay_lo && j + 2 <= g.ay_hi) ? o.z : cv.z;
      o.w = (ract && j + 3 >= g.ay_lo && j + 3 <= g.ay_hi) ? o.w : cv.w;
      // columns >= ny (padding of a partial last group) are inactive: they copy zeros
      *reinterpret_cast<float4*>(B + (r + 1) * nyp + RES_LPAD + j) = o;
    }
    __syncthreads();
    // Push my edge rows of the new field into the neighbours' ghost rows (DSMEM).
    const int nb = cur ^ 1;
    const int n4 = nyp >> 2;
    if (up_ghost[nb] != nullptr) {
      const float4* row = reinterpret_cast<const float4*>(B + 1 * nyp);
      float4* dstp = reinterpret_cast<float4*>(up_ghost[nb]);
      for (int j = threadIdx.x; j < n4; j += RES_THREADS) dstp[j] = row[j];
    }
    if (dn_ghost[nb] != nullptr) {
      const float4* row = reinterpret_cast<const float4*>(B + my_rows * nyp);
      float4* dstp = reinterpret_cast<float4*>(dn_ghost[nb]);
      for (int j = threadIdx.x; j < n4; j += RES_THREADS) dstp[j] = row[j];
    }
    cluster.sync();
    float* t = A; A = B; B = t;
    cur ^= 1;
  }

  for (int r = 1 + warp; r <= my_rows; r += NW) {
    const int gx = x_lo + r - 1;
    float* d = dst + (int64_t)gx * pitch;
    const float* c = A + r * nyp + RES_LPAD;
    for (int j = lane; j < ny; j += 32) d[j] = c[j];
  }
}

struct ResidentShape {
  int ncta = 0, band = 0;
  size_t smem = 0;
};

// Pick the cluster size: as many CTAs as keeps >= 8 rows per band (more SMs =
// more lanes per step), while every CTA owns at least one row and the two
// buffers fit in shared memory.
static bool resident_shape(int nx, int ny, int max_cluster, ResidentShape* out) {
  for (int pass = 0; pass < 2; ++pass) {
    for (int n = max_cluster; n >= 1; n >>= 1) {
      const int band = (nx + n - 1) / n;
      if ((n - 1) * band >= nx) continue;               // last CTA would own nothing
      if (pass == 0 && n > 1 && band < 8) continue;      // preference only
      const size_t smem = (size_t)2 * (band + 2) * res_pitch(ny) * sizeof(float);
      if (smem > (size_t)RES_MAX_SMEM) break;            // fewer CTAs only need more
      out->ncta = n; out->band = band; out->smem = smem;
      return true;
    }
  }
  return false;
}

}  // namespace h2d

using namespace h2d;

extern "C" {

int h2d_resident_fits(int32_t nx, int32_t ny) {
  if (nx < 1 || ny < 1) return 0;
  ResidentShape s;
  return resident_shape(nx, ny, 16, &s) ? s.ncta : 0;
}

int h2d_resident_run(const float* src, float* dst, int32_t nx, int32_t ny, int64_t pitch,
                     int64_t steps, const h2d_phys_t* phys, void* stream) {
  if (!src || !dst || !phys || nx < 1 || ny < 1 || pitch < ny || steps < 0)
    return h2d_fail(H2D_EINVAL, "h2d_resident_run: bad argument");
  h2d_tile_t t = {nx, ny, 0, 0, nx, ny, 0, 0, pitch};
  Geo g = make_geo(t, *phys);
  void (*fn)(const float*, float*, int64_t, int64_t, const Geo, int) = nullptr;
  switch (phys->arith) {
    case H2D_ARITH_FP32: fn = resident_kernel<H2D_ARITH_FP32>; break;
    case H2D_ARITH_STRICT: fn = resident_kernel<H2D_ARITH_STRICT>; break;
    case H2D_ARITH_STRICT_FMA: fn = resident_kernel<H2D_ARITH_STRICT_FMA>; break;
    default: return h2d_fail(H2D_EINVAL, "h2d_resident_run: unknown arith %d", phys->arith);
  }
  H2D_CUDA_TRY(cudaFuncSetAttribute((const void*)fn, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                    RES_MAX_SMEM));
  H2D_CUDA_TRY(cudaFuncSetAttribute((const void*)fn, cudaFuncAttributeNonPortableClusterSizeAllowed, 1));
  // Try the 16-CTA (non-portable) cluster first, fall back to the portable 8.
  for (int max_cluster = 16; max_cluster >= 8; max_cluster >>= 1) {
    ResidentShape sh;
    if (!resident_shape(nx, ny, max_cluster, &sh)) {
      if (max_cluster == 8)
        return h2d_fail(H2D_EINVAL, "h2d_resident_run: %dx%d does not fit in one cluster", nx, ny);
      continue;
    }
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3(sh.ncta);
    cfg.blockDim = dim3(RES_THREADS);
    cfg.dynamicSmemBytes = sh.smem;
    cfg.stream = (cudaStream_t)stream;
    cudaLaunchAttribute at[1];
    at[0].id = cudaLaunchAttributeClusterDimension;
    at[0].val.clusterDim.x = sh.ncta;
    at[0].val.clusterDim.y = 1;
    at[0].val.clusterDim.z = 1;
    cfg.attrs = at;
    cfg.numAttrs = 1;
    int nclusters = 0;
    cudaError_t oe = cudaOccupancyMaxActiveClusters(&nclusters, (const void*)fn, &cfg);
    if (oe != cudaSuccess || nclusters < 1) {
      cudaGetLastError();
      if (sh.ncta > 8) continue;  // retry with the portable size
      return h2d_fail(H2D_ECUDA, "h2d_resident_run: a cluster of %d CTAs with %zu B smem cannot be scheduled",
                      sh.ncta, sh.smem);
    }
    cudaError_t e = cudaLaunchKernelEx(&cfg, fn, src, dst, pitch, steps, g, sh.band);
    if (e != cudaSuccess)
      return h2d_fail(H2D_ECUDA, "h2d_resident_run: launch failed: %s", cudaGetErrorString(e));
    return H2D_OK;
  }
  return h2d_fail(H2D_EINVAL, "h2d_resident_run: no cluster shape found");
}

}  // extern "C"
