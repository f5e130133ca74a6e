// Plain (non-temporal-blocked) kernels: init, one-step stencil, norm of a
// difference, halo pack/unpack for the NCCL baseline, device info.
//
// Reference counterparts: inidat<<<>>> grad1612_cuda_heat.cu:48-53,
// update<<<>>> :55-62, the convergence loop grad1612_mpi_heat.c:264-267, the
// row/column derived datatypes :140-144, detailsGPU() cuda:24-37.
#include <cstdio>
#include <cstring>
#include <mutex>
#include "device_common.cuh"

namespace h2d {

// ------------------------------------------------------------------ init (K1)
__device__ __forceinline__ float init_value_dev(int32_t nx, int32_t ny, int32_t x, int32_t y,
                                                int mode) {
  if (mode == H2D_INIT_ZERO) return 0.0f;
  if (x < 0 || x >= nx || y < 0 || y >= ny) return 0.0f;
  if (mode == H2D_INIT_EXACT) {
    const long long a = (long long)x * (long long)(nx - x - 1);
    const long long b = (long long)y * (long long)(ny - y - 1);
    if (h2d_exact_fits_int64(nx, ny)) return __ll2float_rn(a * b);
    return __double2float_rn(__dmul_rn((double)a, (double)b));  // same rounding as the oracle
  }
  unsigned p = (unsigned)x * (unsigned)(nx - x - 1);
  p *= (unsigned)y;
  p *= (unsigned)(ny - y - 1);
  return mode == H2D_INIT_REF_UINT32 ? __uint2float_rn(p) : __int2float_rn((int)p);
}

__global__ void __launch_bounds__(256)
init_kernel(float* __restrict__ base, Geo g, int mode, int r_lo, int c_lo, int c_hi) {
  const int i = r_lo + (int)blockIdx.x;
  float* row = base + (int64_t)(i + g.halo_r) * g.pitch + g.halo_c;
  for (int j = c_lo + (int)threadIdx.x; j < c_hi; j += (int)blockDim.x)
    row[j] = init_value_dev(g.nx, g.ny, g.x0 + i, g.y0 + j, mode);
}

// ------------------------------------------------------- one-step stencil (K2/K3)
// One warp = one task = a 128-column strip x RB rows.  Each lane owns 4
// consecutive columns (one 16-byte load per row); the two horizontal
// neighbours that live in other lanes come from warp shuffles, the two that
// live outside the strip from one scalar load each (lanes 0 and 31).  A 3-row
// register window means every element is read once from L2/HBM.
constexpr int K1_RB = 32;

template <int ARITH>
__global__ void __launch_bounds__(128)
step_k1_kernel(const float* __restrict__ src, float* __restrict__ dst,
               const __grid_constant__ Geo g, int n_strips, int n_tasks) {
  const int lane = threadIdx.x & 31;
  const int task = (int)blockIdx.x * 4 + (threadIdx.x >> 5);
  if (task >= n_tasks) return;
  const int strip = task % n_strips;
  const int rblk = task / n_strips;
  const int j0 = strip * 128 + lane * 4;
  // Lanes up to and including the one that holds column `cols` load: that
  // column is the east neighbour's ghost (halo_c >= 4 keeps the float4 in bounds).
  const bool lane_on = j0 <= g.cols;
  const int r_lo = rblk * K1_RB;
  const int r_hi = min(r_lo + K1_RB, g.rows);

  // Per-column state that does not change with the row.
  bool act[4], own[4];
#pragma unroll
  for (int e = 0; e < 4; ++e) {
    const int gy = g.y0 + j0 + e;
    own[e] = (j0 + e) < g.cols;
    act[e] = own[e] && gy >= g.ay_lo && gy <= g.ay_hi;
  }
  const bool full = (j0 + 3) < g.cols;

  const float* p = src + (int64_t)(r_lo - 1 + g.halo_r) * g.pitch + g.halo_c + j0;
  float* q = dst + (int64_t)(r_lo + g.halo_r) * g.pitch + g.halo_c + j0;
  const float4 zero4 = make_float4(0.f, 0.f, 0.f, 0.f);

  auto ld4 = [&](const float* a) -> float4 {
    return lane_on ? __ldg(reinterpret_cast<const float4*>(a)) : zero4;
  };
  float4 up = ld4(p);
  p += g.pitch;
  float4 c = ld4(p);
  // Out-of-strip neighbours of the centre row.
  float eL = (lane == 0 && lane_on) ? __ldg(p - 1) : 0.f;
  float eR = (lane == 31 && j0 < g.cols) ? __ldg(p + 4) : 0.f;

  for (int i = r_lo; i < r_hi; ++i) {
    p += g.pitch;
    const float4 dn = ld4(p);
    const float nL = (lane == 0 && lane_on) ? __ldg(p - 1) : 0.f;
    const float nR = (lane == 31 && j0 < g.cols) ? __ldg(p + 4) : 0.f;
    float L = __shfl_up_sync(0xffffffffu, c.w, 1);
    float R = __shfl_down_sync(0xffffffffu, c.x, 1);
    if (lane == 0) L = eL;
    if (lane == 31) R = eR;
    float4 o = point4<ARITH>(up, c, dn, L, R, g);
    const int gx = g.x0 + i;
    const bool ract = gx >= g.ax_lo && gx <= g.ax_hi;
    o.x = (ract && act[0]) ? o.x : c.x;
    o.y = (ract && act[1]) ? o.y : c.y;
    o.z = (ract && act[2]) ? o.z : c.z;
    o.w = (ract && act[3]) ? o.w : c.w;
    if (full) {
      *reinterpret_cast<float4*>(q) = o;
    } else {
      if (own[0]) q[0] = o.x;
      if (own[1]) q[1] = o.y;
      if (own[2]) q[2] = o.z;
    }
    q += g.pitch;
    up = c; c = dn; eL = nL; eR = nR;
  }
}

// ------------------------------------------------------------ norm of (a - b)
// Two-level deterministic reduction: per-block partials, last block finishes
// in a fixed order.  Scratch layout: [0] = uint32 counter (self-resetting),
// [64 bytes..] = float partials[grid].
constexpr int NORM_THREADS = 256;
constexpr int NORM_ROWS_PER_BLOCK = 8;

__global__ void __launch_bounds__(NORM_THREADS)
norm_diff_kernel(const float* __restrict__ a, const float* __restrict__ b, Geo g, int norm,
                 unsigned* counter, float* partials, float* out) {
  __shared__ float red[NORM_THREADS / 32];
  __shared__ bool is_last;
  const int r0 = (int)blockIdx.x * NORM_ROWS_PER_BLOCK;
  const int r1 = min(r0 + NORM_ROWS_PER_BLOCK, g.rows);
  float acc = 0.f;
  for (int i = r0; i < r1; ++i) {
    const int64_t off = (int64_t)(i + g.halo_r) * g.pitch + g.halo_c;
    for (int j = threadIdx.x; j < g.cols; j += NORM_THREADS) {
      const float d = a[off + j] - b[off + j];
      acc = (norm == H2D_NORM_L2SQ) ? fmaf(d, d, acc) : fmax_nan(acc, fabsf(d));
    }
  }
  acc = (norm == H2D_NORM_L2SQ) ? warp_sum(acc) : warp_max(acc);
  if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = acc;
  __syncthreads();
  if (threadIdx.x == 0) {
    float v = red[0];
    for (int w = 1; w < NORM_THREADS / 32; ++w)
      v = (norm == H2D_NORM_L2SQ) ? v + red[w] : fmax_nan(v, red[w]);
    partials[blockIdx.x] = v;
    __threadfence();
    const unsigned prev = atomicAdd(counter, 1u);
    is_last = (prev == gridDim.x - 1);
  }
  __syncthreads();
  if (!is_last) return;
  __threadfence();
  float v = 0.f;
  for (int i = threadIdx.x; i < (int)gridDim.x; i += NORM_THREADS) {
    const float x = __ldcg(partials + i);
    v = (norm == H2D_NORM_L2SQ) ? v + x : fmax_nan(v, x);
  }
  v = (norm == H2D_NORM_L2SQ) ? warp_sum(v) : warp_max(v);
  if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = v;
  __syncthreads();
  if (threadIdx.x == 0) {
    float t = red[0];
    for (int w = 1; w < NORM_THREADS / 32; ++w)
      t = (norm == H2D_NORM_L2SQ) ? t + red[w] : fmax_nan(t, red[w]);
    *out = t;
    *counter = 0u;  // ready for the next call on this stream
  }
}

// ---- scratch of the no-allocation norm entry point (h2d_cuda_norm_diff_static) ----
// Lives in the library image: every GPU that loads libheat2d_cuda.so carries
// these 64 KiB + 256 B until the process exits.  Same layout as the caller-
// provided scratch of h2d_cuda_norm_diff: a 64-byte header whose first word is
// the block counter, then one float per block.
constexpr int NORM_STATIC_MAX_BLOCKS = H2D_NORM_STATIC_MAX_ROWS / NORM_ROWS_PER_BLOCK;
constexpr size_t NORM_STATIC_BYTES = 256 + (size_t)NORM_STATIC_MAX_BLOCKS * sizeof(float);
static __device__ __align__(256) unsigned char g_norm_static_scratch[NORM_STATIC_BYTES];

// ------------------------------------------------ halo pack / unpack (baseline)
// dir N/S: w rows x cols, row-major.  dir W/E: rows x w, row-major.
__global__ void pack_edge_kernel(const float* __restrict__ base, Geo g, int dir, int w,
                                 float* __restrict__ out, int ghost) {
  // ghost == 0: read the owned edge; ghost == 1: address the margin instead.
  int nr, nc, r0, c0;
  if (dir == H2D_N || dir == H2D_S) {
    nr = w; nc = g.cols; c0 = 0;
    r0 = (dir == H2D_N) ? (ghost ? -w : 0) : (ghost ? g.rows : g.rows - w);
  } else {
    nr = g.rows; nc = w; r0 = 0;
    c0 = (dir == H2D_W) ? (ghost ? -w : 0) : (ghost ? g.cols : g.cols - w);
  }
  const int64_t n = (int64_t)nr * nc;
  for (int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; t < n;
       t += (int64_t)gridDim.x * blockDim.x) {
    const int i = (int)(t / nc), j = (int)(t % nc);
    out[t] = base[(int64_t)(r0 + i + g.halo_r) * g.pitch + g.halo_c + c0 + j];
  }
}

__global__ void unpack_ghost_kernel(float* __restrict__ base, Geo g, int dir, int w,
                                    const float* __restrict__ in) {
  int nr, nc, r0, c0;
  if (dir == H2D_N || dir == H2D_S) {
    nr = w; nc = g.cols; c0 = 0;
    r0 = (dir == H2D_N) ? -w : g.rows;
  } else {
    nr = g.rows; nc = w; r0 = 0;
    c0 = (dir == H2D_W) ? -w : g.cols;
  }
  const int64_t n = (int64_t)nr * nc;
  for (int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; t < n;
       t += (int64_t)gridDim.x * blockDim.x) {
    const int i = (int)(t / nc), j = (int)(t % nc);
    base[(int64_t)(r0 + i + g.halo_r) * g.pitch + g.halo_c + c0 + j] = in[t];
  }
}

}  // namespace h2d

using namespace h2d;

static int check_common(const void* a, const h2d_tile_t* t, const char* who) {
  if (!a || !t) return h2d_fail(H2D_EINVAL, "%s: null argument", who);
  if (int rc = h2d_check_tile(t)) return rc;
  if (t->pitch % 4 != 0 || t->halo_c % 4 != 0)
    return h2d_fail(H2D_EINVAL, "%s: pitch and halo_c must be multiples of 4 floats", who);
  if (((uintptr_t)a & 15) != 0) return h2d_fail(H2D_EINVAL, "%s: base must be 16-byte aligned", who);
  return H2D_OK;
}

extern "C" {

const char* h2d_last_error(void) { return h2d_err_buf(); }
const char* h2d_version(void) { return H2D_VERSION_STRING; }

int h2d_cuda_device_info(int device, char* out, size_t cap) {
  if (!out || cap == 0) return h2d_fail(H2D_EINVAL, "h2d_cuda_device_info: no buffer");
  cudaDeviceProp p;
  H2D_CUDA_TRY(cudaGetDeviceProperties(&p, device));
  int clusters = 0;
  cudaDeviceGetAttribute(&clusters, cudaDevAttrClusterLaunch, device);
  snprintf(out, cap,
           "%s: %d.%d\nGlobal memory: %zu MB\nShared memory per block: %zu KB (opt-in %zu KB)\n"
           "Constant memory: %zu KB\nRegisters per block: %d\nWarp size: %d\nMax threads per block: %d\n"
           "Max block dimensions: [%d, %d, %d]\nMax grid dimensions: [%d, %d, %d]\n"
           "SMs: %d\nL2: %d MB\nCluster launch: %d\n",
           p.name, p.major, p.minor, p.totalGlobalMem >> 20, p.sharedMemPerBlock >> 10,
           p.sharedMemPerBlockOptin >> 10, p.totalConstMem >> 10, p.regsPerBlock, p.warpSize,
           p.maxThreadsPerBlock, p.maxThreadsDim[0], p.maxThreadsDim[1], p.maxThreadsDim[2],
           p.maxGridSize[0], p.maxGridSize[1], p.maxGridSize[2], p.multiProcessorCount,
           p.l2CacheSize >> 20, clusters);
  return H2D_OK;
}

int h2d_cuda_init(float* base, const h2d_tile_t* t, int init_mode, int with_margins, void* stream) {
  if (int rc = check_common(base, t, "h2d_cuda_init")) return rc;
  h2d_phys_t ph = {0.1, 0.1, H2D_BOUNDARY_FIXED, H2D_ARITH_FP32};
  Geo g = make_geo(*t, ph);
  const int r_lo = with_margins ? -t->halo_r : 0;
  const int r_hi = with_margins ? t->rows + t->halo_r : t->rows;
  const int c_lo = with_margins ? -t->halo_c : 0;
  const int c_hi = with_margins ? t->cols + t->halo_c : t->cols;
  init_kernel<<<r_hi - r_lo, 256, 0, (cudaStream_t)stream>>>(base, g, init_mode, r_lo, c_lo, c_hi);
  H2D_CHECK_LAUNCH("init_kernel");
  return H2D_OK;
}

int h2d_cuda_step_k1(const float* src, float* dst, const h2d_tile_t* t, const h2d_phys_t* phys,
                     void* stream) {
  if (int rc = check_common(src, t, "h2d_cuda_step_k1")) return rc;
  if (int rc = check_common(dst, t, "h2d_cuda_step_k1")) return rc;
  if (!phys) return h2d_fail(H2D_EINVAL, "h2d_cuda_step_k1: null phys");
  if (t->halo_r < 1 || t->halo_c < 4)
    return h2d_fail(H2D_EINVAL, "h2d_cuda_step_k1: needs halo_r >= 1 and halo_c >= 4");
  Geo g = make_geo(*t, *phys);
  const int n_strips = (t->cols + 127) / 128;
  const int n_rblk = (t->rows + K1_RB - 1) / K1_RB;
  const int64_t n_tasks64 = (int64_t)n_strips * n_rblk;
  if (n_tasks64 > (int64_t)1 << 30) return h2d_fail(H2D_EINVAL, "h2d_cuda_step_k1: tile too large");
  const int n_tasks = (int)n_tasks64;
  const int grid = (n_tasks + 3) / 4;
  cudaStream_t s = (cudaStream_t)stream;
  switch (phys->arith) {
    case H2D_ARITH_FP32:
      step_k1_kernel<H2D_ARITH_FP32><<<grid, 128, 0, s>>>(src, dst, g, n_strips, n_tasks);
      break;
    case H2D_ARITH_STRICT:
      step_k1_kernel<H2D_ARITH_STRICT><<<grid, 128, 0, s>>>(src, dst, g, n_strips, n_tasks);
      break;
    case H2D_ARITH_STRICT_FMA:
      step_k1_kernel<H2D_ARITH_STRICT_FMA><<<grid, 128, 0, s>>>(src, dst, g, n_strips, n_tasks);
      break;
    default:
      return h2d_fail(H2D_EINVAL, "h2d_cuda_step_k1: unknown arith %d", phys->arith);
  }
  H2D_CHECK_LAUNCH("step_k1_kernel");
  return H2D_OK;
}

size_t h2d_norm_scratch_bytes(const h2d_tile_t* t) {
  if (!t || t->rows < 1) return 0;
  const size_t blocks = ((size_t)t->rows + NORM_ROWS_PER_BLOCK - 1) / NORM_ROWS_PER_BLOCK;
  return 64 + blocks * sizeof(float);
}

int h2d_cuda_norm_diff(const float* a, const float* b, const h2d_tile_t* t, int norm,
                       void* scratch, float* out_dev, void* stream) {
  if (int rc = check_common(a, t, "h2d_cuda_norm_diff")) return rc;
  if (!b || !scratch || !out_dev) return h2d_fail(H2D_EINVAL, "h2d_cuda_norm_diff: null argument");
  if (norm != H2D_NORM_L2SQ && norm != H2D_NORM_LINF)
    return h2d_fail(H2D_EINVAL, "h2d_cuda_norm_diff: unknown norm %d", norm);
  h2d_phys_t ph = {0.1, 0.1, H2D_BOUNDARY_FIXED, H2D_ARITH_FP32};
  Geo g = make_geo(*t, ph);
  const int blocks = (t->rows + NORM_ROWS_PER_BLOCK - 1) / NORM_ROWS_PER_BLOCK;
  unsigned* counter = (unsigned*)scratch;
  float* partials = (float*)((char*)scratch + 64);
  norm_diff_kernel<<<blocks, NORM_THREADS, 0, (cudaStream_t)stream>>>(a, b, g, norm, counter,
                                                                      partials, out_dev);
  H2D_CHECK_LAUNCH("norm_diff_kernel");
  return H2D_OK;
}

// The name `g_norm_static_scratch` is only a host-side stand-in in host code;
// the kernel needs the array's device address, which differs per GPU.  Looked up
// once per device (with that device current) and cached.
static int norm_static_scratch_addr(void** out) {
  constexpr int MAX_DEVICES = 64;
  static std::mutex mu;
  static void* cache[MAX_DEVICES] = {nullptr};
  int dev = -1;
  H2D_CUDA_TRY(cudaGetDevice(&dev));
  if (dev < 0 || dev >= MAX_DEVICES)
    return h2d_fail(H2D_EINVAL, "h2d_cuda_norm_diff_static: device ordinal %d not supported", dev);
  std::lock_guard<std::mutex> lock(mu);
  if (cache[dev] == nullptr) {
    void* p = nullptr;
    cudaError_t e = cudaGetSymbolAddress((void**)&p, g_norm_static_scratch);
    if (e != cudaSuccess || p == nullptr)
      return h2d_fail(H2D_ECUDA, "h2d_cuda_norm_diff_static: cudaGetSymbolAddress failed on device %d: %s",
                      dev, cudaGetErrorString(e));
    cache[dev] = p;
  }
  *out = cache[dev];
  return H2D_OK;
}

// norm of (a - b) without any allocation, for callers that cannot allocate
// (their own pool owns the GPU, or they must not hit cudaMalloc's device-wide
// synchronisation).  Same kernel and same summation order as
// h2d_cuda_norm_diff; only the scratch differs: the fixed array above, shared by
// every call on this GPU.  Rows beyond H2D_NORM_STATIC_MAX_ROWS are refused.
// The first call on a device performs the address look-up (do it before any
// stream capture); calls that may overlap on the GPU race on the shared array.
int h2d_cuda_norm_diff_static(const float* a, const float* b, const h2d_tile_t* t, int norm,
                              float* out_dev, void* stream) {
  if (int rc = check_common(a, t, "h2d_cuda_norm_diff_static")) return rc;
  if (!b || !out_dev) return h2d_fail(H2D_EINVAL, "h2d_cuda_norm_diff_static: null argument");
  if (norm != H2D_NORM_L2SQ && norm != H2D_NORM_LINF)
    return h2d_fail(H2D_EINVAL, "h2d_cuda_norm_diff_static: unknown norm %d", norm);
  if (t->rows > H2D_NORM_STATIC_MAX_ROWS)
    return h2d_fail(H2D_EINVAL,
                    "h2d_cuda_norm_diff_static: tile has %d rows, the built-in scratch covers at most %d; "
                    "use h2d_cuda_norm_diff with caller-provided scratch",
                    t->rows, H2D_NORM_STATIC_MAX_ROWS);
  void* scratch = nullptr;
  if (int rc = norm_static_scratch_addr(&scratch)) return rc;
  h2d_phys_t ph = {0.1, 0.1, H2D_BOUNDARY_FIXED, H2D_ARITH_FP32};
  Geo g = make_geo(*t, ph);
  const int blocks = (t->rows + NORM_ROWS_PER_BLOCK - 1) / NORM_ROWS_PER_BLOCK;
  cudaStream_t s = (cudaStream_t)stream;
  // The array keeps whatever the previous call (or an aborted one) left behind:
  // the block counter must start at zero, the partials are fully overwritten.
  H2D_CUDA_TRY(cudaMemsetAsync(scratch, 0, 64, s));
  norm_diff_kernel<<<blocks, NORM_THREADS, 0, s>>>(a, b, g, norm, (unsigned*)scratch,
                                                   (float*)((char*)scratch + 64), out_dev);
  H2D_CHECK_LAUNCH("norm_diff_kernel (static scratch)");
  return H2D_OK;
}

static int edge_args_ok(const h2d_tile_t* t, int dir, int w, const char* who) {
  if (dir < H2D_N || dir > H2D_E) return h2d_fail(H2D_EINVAL, "%s: dir must be N,S,W or E", who);
  const int lim = (dir == H2D_N || dir == H2D_S) ? t->halo_r : t->halo_c;
  const int ext = (dir == H2D_N || dir == H2D_S) ? t->rows : t->cols;
  if (w < 1 || w > lim || w > ext)
    return h2d_fail(H2D_EINVAL, "%s: width %d outside [1, min(halo=%d, extent=%d)]", who, w, lim, ext);
  return H2D_OK;
}

int h2d_cuda_pack_edge(const float* base, const h2d_tile_t* t, int dir, int w, float* out,
                       void* stream) {
  if (int rc = check_common(base, t, "h2d_cuda_pack_edge")) return rc;
  if (!out) return h2d_fail(H2D_EINVAL, "h2d_cuda_pack_edge: null out");
  if (int rc = edge_args_ok(t, dir, w, "h2d_cuda_pack_edge")) return rc;
  h2d_phys_t ph = {0.1, 0.1, H2D_BOUNDARY_FIXED, H2D_ARITH_FP32};
  Geo g = make_geo(*t, ph);
  const int64_t n = (dir == H2D_N || dir == H2D_S) ? (int64_t)w * t->cols : (int64_t)w * t->rows;
  const int blocks = (int)((n + 255) / 256 > 1184 ? 1184 : (n + 255) / 256);
  pack_edge_kernel<<<blocks, 256, 0, (cudaStream_t)stream>>>(base, g, dir, w, out, 0);
  H2D_CHECK_LAUNCH("pack_edge_kernel");
  return H2D_OK;
}

int h2d_cuda_unpack_ghost(float* base, const h2d_tile_t* t, int dir, int w, const float* in,
                          void* stream) {
  if (int rc = check_common(base, t, "h2d_cuda_unpack_ghost")) return rc;
  if (!in) return h2d_fail(H2D_EINVAL, "h2d_cuda_unpack_ghost: null in");
  if (int rc = edge_args_ok(t, dir, w, "h2d_cuda_unpack_ghost")) return rc;
  h2d_phys_t ph = {0.1, 0.1, H2D_BOUNDARY_FIXED, H2D_ARITH_FP32};
  Geo g = make_geo(*t, ph);
  const int64_t n = (dir == H2D_N || dir == H2D_S) ? (int64_t)w * t->cols : (int64_t)w * t->rows;
  const int blocks = (int)((n + 255) / 256 > 1184 ? 1184 : (n + 255) / 256);
  unpack_ghost_kernel<<<blocks, 256, 0, (cudaStream_t)stream>>>(base, g, dir, w, in);
  H2D_CHECK_LAUNCH("unpack_ghost_kernel");
  return H2D_OK;
}

}  // extern "C"
