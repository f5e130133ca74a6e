// stencil_tk — the flagship kernel: K time steps per launch (temporal
// blocking), input rows streamed by TMA into per-warp shared-memory rings,
// K levels of the stencil pipelined in registers, output written once; the
// same kernel stores the tile's edge bands straight into the neighbour GPUs'
// ghost margins (peer memory over NVLink) and publishes one epoch flag per
// launch, and optionally reduces the last sub-step's (new - old) for the
// convergence test.
//
// What it replaces in the reference: the interior and edge sweeps
// (grad1612_mpi_heat.c:239-259), the 16 persistent halo requests and their
// Startall/Waitall (:210-235,244,274), the strided `column` datatype (:143),
// the convergence loop and MPI_Allreduce (:262-271), and update<<<>>> of
// grad1612_cuda_heat.cu:55-62.  None of that code is reused: the reference
// exchanges one ghost cell per step, this kernel exchanges kmax-wide bands
// once per K steps and never packs anything.
//
// This file is the HOST side: the plan (tensor maps, task lists, scratch), the launch
// entry points and the pipelined upload / solve / download.  The device code lives in
// tk_task.cuh (task bodies) and tk_kernel.cuh (kernel wrapper); the kernels are
// instantiated in tk_inst.cu, one object per (arithmetic, lane layout, extras) combination.
#include <cuda.h>
#include <cstdio>
#include <cstdlib>
#include <cmath>
#include <cstring>
#include <vector>
#include "tk_plan.h"

namespace h2d {

template <bool XTR>
static tk_fn_t tk_pick_x(int k, int arith, int lay, bool imm) {
  switch (arith) {
    case H2D_ARITH_FP32:
      switch (lay) {
        case LAY_V4: return tk_get<H2D_ARITH_FP32, LAY_V4, XTR>(k, imm);
        case LAY_AB2: return tk_get<H2D_ARITH_FP32, LAY_AB2, XTR>(k, imm);
        case LAY_AB4: return tk_get<H2D_ARITH_FP32, LAY_AB4, XTR>(k, imm);
        default: return nullptr;
      }
    case H2D_ARITH_STRICT:
      return lay == LAY_V4 ? tk_get<H2D_ARITH_STRICT, LAY_V4, XTR>(k, false)
           : lay == LAY_V8 ? tk_get<H2D_ARITH_STRICT, LAY_V8, XTR>(k, false) : nullptr;
    case H2D_ARITH_STRICT_FMA:
      return lay == LAY_V4 ? tk_get<H2D_ARITH_STRICT_FMA, LAY_V4, XTR>(k, false)
           : lay == LAY_V8 ? tk_get<H2D_ARITH_STRICT_FMA, LAY_V8, XTR>(k, false) : nullptr;
    default: return nullptr;
  }
}
// `imm`: fp32 mode with exactly the reference coefficients -> immediate-operand kernels.
// `xtr`: the launch forwards halos and/or reduces a norm.
tk_fn_t tk_pick(int k, int arith, int lay, bool imm, bool xtr) {
  return xtr ? tk_pick_x<true>(k, arith, lay, imm) : tk_pick_x<false>(k, arith, lay, imm);
}

bool coef_is_reference(const Geo& g, int arith) {
  return arith == H2D_ARITH_FP32 && g.cx == kRefCx && g.cy == kRefCy && g.a0 == kRefA0;
}

// H2D_DEBUG_NOFWD=1: skip the peer stores (timing experiments); H2D_DEBUG_DELAY=1: random
// per-task delays around the flag protocol (stress test); H2D_STRICT_FAST=0: strict modes
// always use the plain F2F evaluation.
int tk_debug_flags() {
  // read on every launch: sweeps switch these between runs of one process
  const char* e = getenv("H2D_DEBUG_NOFWD");
  const char* d = getenv("H2D_DEBUG_DELAY");
  const char* f = getenv("H2D_STRICT_FAST");
  const char* b = getenv("H2D_FWD_BULK");    // row bands through the bulk-copy engine (default on)
  const char* l = getenv("H2D_FWD_LINES");   // column bands as whole 128-byte lines
  return ((e && e[0] == '1') ? 1 : 0) | ((d && d[0] == '1') ? 2 : 0) | ((f && f[0] == '0') ? 4 : 0) |
         ((b && b[0] == '0') ? 0 : 8) | ((l && l[0] == '1') ? 16 : 0);
}

// margins of a freshly filled padded array -> this rank's packed column-ghost buffers
__global__ void seed_colbuf_kernel(const float* __restrict__ base, Geo g, int fw, float* w, float* e) {
  const int64_t n = (int64_t)g.rows * TK_CW;
  for (int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; t < n; t += (int64_t)gridDim.x * blockDim.x) {
    const int r = (int)(t / TK_CW), j = (int)(t % TK_CW);
    if (j >= fw) continue;
    const float* row = base + (int64_t)(g.halo_r + r) * g.pitch + g.halo_c;
    if (w) w[t] = row[j - fw];
    if (e) e[t] = row[g.cols + j];
  }
}

}  // namespace h2d

using namespace h2d;

// ------------------------------------------------------------------ the plan
static int g_sm_count = 0;
static int sm_count() {
  if (g_sm_count <= 0) {
    int dev = 0, n = 0;
    if (cudaGetDevice(&dev) == cudaSuccess &&
        cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev) == cudaSuccess && n > 0)
      g_sm_count = n;
    else
      g_sm_count = 148;
  }
  return g_sm_count;
}

// Resident warps per SM of the kernels of one plan (TkBounds in tk_kernel.cuh).
static int warps_per_sm(int kmax, int lay) { return tk_min_blocks(kmax, lay) * TK_WARPS; }

// Lane layout of the fp32 kernels.  Default: H2D_FP32_LAYOUT=v4|ab2|ab4, else the measured
// best.  Strict modes always use V4 (their arithmetic is scalar).
static int default_fp32_layout() {
  static int lay = -1;
  if (lay < 0) {
    lay = H2D_DEFAULT_FP32_LAYOUT;
    const char* e = getenv("H2D_FP32_LAYOUT");
    if (e && !strcmp(e, "v4")) lay = LAY_V4;
    if (e && !strcmp(e, "ab2")) lay = LAY_AB2;
    if (e && !strcmp(e, "ab4")) lay = LAY_AB4;
  }
  return lay;
}
// Strict modes: their arithmetic is scalar, so V4 — or, for plans whose sweeps are at most 4
// levels deep, V8 (8 columns per lane).  Default for those: H2D_STRICT_LAYOUT=v4|v8, else the
// measured best (build.py passes V8: strict-fma, k = 4, 32768^2 1438 vs 1366 GLUP/s, 8192 x 16384
// 1280 vs 1219, 16384^2 1322 vs 1320).
#ifndef H2D_DEFAULT_STRICT_SHALLOW_LAYOUT
#define H2D_DEFAULT_STRICT_SHALLOW_LAYOUT LAY_V4
#endif
static int default_strict_shallow_layout() {
  static int lay = -1;
  if (lay < 0) {
    lay = H2D_DEFAULT_STRICT_SHALLOW_LAYOUT;
    const char* e = getenv("H2D_STRICT_LAYOUT");
    if (e && !strcmp(e, "v4")) lay = LAY_V4;
    if (e && !strcmp(e, "v8")) lay = LAY_V8;
  }
  return lay;
}
static int pick_layout(int arith, int requested, int kmax) {
  if (arith != H2D_ARITH_FP32) {
    if (kmax > TK_V8_KMAX || requested == LAY_V4 + 1) return LAY_V4;
    return requested == LAY_V8 + 1 ? LAY_V8 : default_strict_shallow_layout();
  }
  if (requested >= 1 && requested <= LAY_AB4 + 1) return requested - 1;
  return default_fp32_layout();
}

static int choose_chunk_rows(int rows, int cols, int kmax, int lay, int min_waves = 1) {
  // Tasks cost about the same and the block scheduler hands them out in order,
  // so a launch takes about ceil(tasks / slots) "waves" of one task time, and a
  // task time is proportional to the rows it streams, R + 2K (K rows of pipeline
  // fill at both ends) plus ~8 rows of start-up.  A quarter wave is added for the
  // slower edge / boundary tasks.  Minimise  (ceil(waves) + 0.25) * (R + 2K + 8)
  // over the chunk count (constants fitted to sweeps on B200: 32768^2 -> R = 683,
  // 16384^2 -> 342, 8192x32768 -> 342, 4096x32768 -> 228, all within 1 % of the best).
  // slots = SMs x resident warps.  Small grids land in the latency regime (one wave):
  // short chunks win there.
  const int wv = tk_strip_width(lay) - 2 * round_up4(kmax);
  const int n_strips = (cols + wv - 1) / wv;
  const double slots = (double)sm_count() * warps_per_sm(kmax, lay);
  const int r_min = kmax > 8 ? kmax : 8;
  int best_R = rows;
  double best = 1e300;
  for (int c = 1; c <= rows; ++c) {
    const int R = (rows + c - 1) / c;   // equal chunks; the last one may be shorter
    if (R < r_min && c > 1) break;
    const double waves = ceil((double)((rows + R - 1) / R) * n_strips / slots);
    // With neighbours, the edge tasks (first wave) gate the neighbours' NEXT launch: keep
    // them a small part of the launch.
    if (waves < min_waves && R > r_min) continue;
    const double t = (waves + 0.25) * (double)(R + 2 * kmax + 8);
    if (t < best * (1.0 - 1e-9)) { best = t; best_R = R; }
  }
  return best_R;
}

// Waves of tasks a launch of a plan WITH neighbours keeps at least (see choose_chunk_rows).
// Tuning hook for sweeps on several GPUs without rebuilding: H2D_MIN_WAVES=1..16.
static int peer_min_waves() {
  const char* e = getenv("H2D_MIN_WAVES");
  if (e && e[0]) {
    const int v = atoi(e);
    if (v >= 1 && v <= 16) return v;
  }
  return H2D_DEFAULT_MIN_WAVES;
}

static void plan_counts(const h2d_tile_t* t, int kmax, int chunk_rows, int lay, int* n_chunks,
                        int* max_tasks, int min_waves = 1) {
  int R = chunk_rows > 0 ? chunk_rows : choose_chunk_rows(t->rows, t->cols, kmax, lay, min_waves);
  int nc = (t->rows + R - 1) / R;
  if (nc < 1) nc = 1;
  *n_chunks = nc;
  int mt = 0;
  for (int H = 4; H <= round_up4(kmax); H += 4) {
    const int wv = tk_strip_width(lay) - 2 * H;
    const int ns = (t->cols + wv - 1) / wv;
    if (ns * nc > mt) mt = ns * nc;
  }
  *max_tasks = mt;
}

static size_t scratch_layout_bytes(int max_tasks) {
  const size_t ctas = ((size_t)max_tasks + TK_WARPS - 1) / TK_WARPS;
  return 256 + ctas * sizeof(float);
}

typedef CUresult (*encode_tiled_fn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*,
                                    const cuuint64_t*, const cuuint64_t*, const cuuint32_t*,
                                    const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                    CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

// The map covers this GPU's own padded array and nothing else.
static int make_tmap(CUtensorMap* out, float* base, const h2d_tile_t& t, int lay) {
  static encode_tiled_fn fn = nullptr;
  if (!fn) {
    void* p = nullptr;
    cudaDriverEntryPointQueryResult qr;
    H2D_CUDA_TRY(cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &qr));
    if (!p || qr != cudaDriverEntryPointSuccess)
      return h2d_fail(H2D_ECUDA, "cuTensorMapEncodeTiled is not available from this driver");
    fn = (encode_tiled_fn)p;
  }
  const cuuint64_t gdim[2] = {(cuuint64_t)t.pitch, (cuuint64_t)(t.rows + 2 * t.halo_r)};
  const cuuint64_t gstride[1] = {(cuuint64_t)t.pitch * 4};
  const cuuint32_t box[2] = {(cuuint32_t)tk_strip_width(lay), TK_BR};
  const cuuint32_t estr[2] = {1, 1};
  CUresult r = fn(out, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 2, base, gdim, gstride, box, estr,
                  CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE,
                  CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) return h2d_fail(H2D_ECUDA, "cuTensorMapEncodeTiled failed with %d", (int)r);
  return H2D_OK;
}

// Chunk size of the strips that forward column bands: a tuning hook (H2D_EDGE_CHUNK=rows).
// Shorter chunks spread the column-band copies over more, shorter tasks; measured on two B200
// (1 x 2, 32768^2) that made things WORSE (6008 vs 6492 GLUP/s): the cost of a column band is
// not in the issuing warp (whole 128-byte lines and exact 32-byte rows cost the same, and
// without the copies 1 x 2 runs at the 2 x 1 speed, 8625), it is the row-strided remote
// writes themselves.  Default: the same chunks as every other strip.
static int edge_chunk_rows(const h2d_plan* p, int R, int kmax) {
  const bool we = p->desc.peer[H2D_W].buf[0] || p->desc.peer[H2D_E].buf[0];
  const char* e = getenv("H2D_EDGE_CHUNK");
  if (!we || !e || atoi(e) < 8) return R;
  int Re = atoi(e);
  if (Re < kmax) Re = kmax;
  return Re < R ? Re : R;
}

static int build_task_list(h2d_plan* p, int H, TaskList* out) {
  const h2d_tile_t& t = p->desc.tile;
  const int wv = tk_strip_width(p->lay) - 2 * H;
  const int ns = (t.cols + wv - 1) / wv;
  const int nc = p->n_chunks;
  const int R = p->chunk_rows, Re = p->chunk_rows_edge;
  const int nce = (t.rows + Re - 1) / Re;
  const int fw = p->fw;
  std::vector<int2> edge, inner;
  edge.reserve((size_t)2 * (ns + nc + nce));
  inner.reserve((size_t)ns * nc);
  for (int s = 0; s < ns; ++s) {
    const int j_lo = s * wv, j_hi = (j_lo + wv < t.cols) ? j_lo + wv : t.cols;
    const bool col_edge = j_lo < fw || j_hi > t.cols - fw;
    const bool shorter = col_edge && Re < R;    // this strip forwards (or waits for) column bands
    const int rr = shorter ? Re : R, n = shorter ? nce : nc;
    for (int c = 0; c < n; ++c) {
      const int r_lo = c * rr, r_hi = (c == n - 1) ? t.rows : r_lo + rr;
      const bool row_edge = r_lo < fw || r_hi > t.rows - fw;
      if (row_edge || col_edge)
        edge.push_back(make_int2((int)((unsigned)s | 0x80000000u | (shorter ? 0x40000000u : 0u)), c));
      else
        inner.push_back(make_int2(s, c));
    }
  }
  out->n_strips = ns;
  out->n_comm_edge = (int)edge.size();
  out->n_tasks = (int)(edge.size() + inner.size());
  edge.insert(edge.end(), inner.begin(), inner.end());
  H2D_CUDA_TRY(cudaMalloc(&out->dev, edge.size() * sizeof(int2)));
  H2D_CUDA_TRY(cudaMemcpy(out->dev, edge.data(), edge.size() * sizeof(int2), cudaMemcpyHostToDevice));
  return H2D_OK;
}

// Arguments every launch of a plan shares.
static void fill_common(h2d_plan* p, TkArgs& a, int src_parity, int k) {
  memset(&a, 0, sizeof a);
  const int H = round_up4(k);
  a.g = p->geo;
  a.dst = p->desc.buf[1 - src_parity];
  a.src = p->desc.buf[src_parity];
  a.colbuf_in[0] = p->desc.colbuf_in[0][src_parity];
  a.colbuf_in[1] = p->desc.colbuf_in[1][src_parity];
  a.H = H;
  a.wv = tk_strip_width(p->lay) - 2 * H;
  a.fw = p->fw;
  tk_fill_scratch_ptrs(a, p->scratch);
  a.timeout_ns = p->desc.spin_timeout_ns > 0 ? p->desc.spin_timeout_ns : 10000000000LL;
  a.seq = p->seq;
  a.debug = tk_debug_flags();
}
static void fill_peers(h2d_plan* p, TkArgs& a, int src_parity, uint32_t epoch) {
  a.has_peers = p->has_peers ? 1 : 0;
  for (int d = 0; d < H2D_NDIR; ++d) {
    const h2d_peer_t& pr = p->desc.peer[d];
    a.peer[d].buf = pr.buf[1 - src_parity];
    a.peer[d].colbuf = pr.colbuf[1 - src_parity];
    a.peer[d].flag = pr.flag;
    a.peer[d].x0 = pr.x0; a.peer[d].y0 = pr.y0;
    a.peer[d].halo_r = pr.halo_r; a.peer[d].halo_c = pr.halo_c;
    a.peer[d].pitch = pr.pitch;
  }
  a.flags_in = p->desc.flags_in;
  a.epoch = epoch;
}

extern "C" {

int32_t h2d_plan_chunk_rows_ex(const h2d_tile_t* tile, int32_t kmax, int32_t with_peers, int32_t arith,
                               int32_t layout) {
  if (!tile || kmax < 1 || kmax > TK_KMAX || tile->rows < 1 || tile->cols < 1) return 0;
  return choose_chunk_rows(tile->rows, tile->cols, kmax, pick_layout(arith, layout, kmax), with_peers ? peer_min_waves() : 1);
}

int32_t h2d_plan_chunk_rows(const h2d_tile_t* tile, int32_t kmax, int32_t with_peers) {
  return h2d_plan_chunk_rows_ex(tile, kmax, with_peers, H2D_ARITH_FP32, 0);
}

int32_t h2d_plan_strip_width(int32_t arith, int32_t layout) {
  return tk_strip_width(pick_layout(arith, layout, layout == LAY_V8 + 1 ? TK_V8_KMAX : TK_KMAX));
}

size_t h2d_plan_scratch_bytes(const h2d_tile_t* tile, int32_t kmax, int32_t chunk_rows) {
  if (!tile || kmax < 1 || kmax > TK_KMAX || tile->rows < 1 || tile->cols < 1) return 0;
  int mt = 0;
  for (int lay = 0; lay < LAY_COUNT; ++lay) {   // whichever layout the plan ends up with
    int nc, m1, m2;
    plan_counts(tile, kmax, chunk_rows, lay, &nc, &m1);
    plan_counts(tile, kmax, chunk_rows, lay, &nc, &m2, peer_min_waves());   // plans with peers use more, shorter chunks
    if (m1 > mt) mt = m1;
    if (m2 > mt) mt = m2;
  }
  mt += 3 * ((tile->rows + 31) / 32);   // column-band strips in chunks of >= 32 rows
  return scratch_layout_bytes(mt);
}

void h2d_plan_destroy(h2d_plan_t* p) {
  if (!p) return;
  for (auto& l : p->lists)
    if (l.dev) cudaFree(l.dev);
  if (p->own_scratch && p->scratch) cudaFree(p->scratch);
  if (p->flow) cudaFree(p->flow);
  if (p->ev_fork) cudaEventDestroy(p->ev_fork);
  if (p->ev_join) cudaEventDestroy(p->ev_join);
  if (p->s_side) cudaStreamDestroy(p->s_side);
  for (cudaEvent_t e : p->ev_pool) cudaEventDestroy(e);
  if (p->s_h2d) cudaStreamDestroy(p->s_h2d);
  if (p->s_d2h) cudaStreamDestroy(p->s_d2h);
  delete p;
}

h2d_plan_t* h2d_plan_create(const h2d_plan_desc_t* d) {
  if (!d) { h2d_fail(H2D_EINVAL, "h2d_plan_create: null desc"); return nullptr; }
  const h2d_tile_t& t = d->tile;
  if (h2d_check_tile(&t)) return nullptr;
  if (!d->buf[0] || !d->buf[1]) { h2d_fail(H2D_EINVAL, "h2d_plan_create: null buffer"); return nullptr; }
  if (d->kmax < 1 || d->kmax > TK_KMAX) {
    h2d_fail(H2D_EINVAL, "h2d_plan_create: kmax %d outside [1, %d]", d->kmax, TK_KMAX);
    return nullptr;
  }
  const int Hmax = round_up4(d->kmax);
  if (t.halo_r < d->kmax || t.halo_c < Hmax || t.halo_c % 4 != 0 || t.pitch % 4 != 0) {
    h2d_fail(H2D_EINVAL,
             "h2d_plan_create: need halo_r >= kmax (%d), halo_c >= %d and a multiple of 4, pitch %% 4 == 0",
             d->kmax, Hmax);
    return nullptr;
  }
  if ((((uintptr_t)d->buf[0]) | ((uintptr_t)d->buf[1])) & 127) {
    h2d_fail(H2D_EINVAL, "h2d_plan_create: buffers must be 128-byte aligned");
    return nullptr;
  }
  if (d->phys.arith < H2D_ARITH_FP32 || d->phys.arith > H2D_ARITH_STRICT_FMA) {
    h2d_fail(H2D_EINVAL, "h2d_plan_create: unknown arith %d", d->phys.arith);
    return nullptr;
  }
  if (d->layout < 0 || d->layout > LAY_COUNT) {
    h2d_fail(H2D_EINVAL, "h2d_plan_create: unknown layout %d", d->layout);
    return nullptr;
  }
  bool has_peers = false;
  for (int k = 0; k < H2D_NDIR; ++k)
    if (d->peer[k].buf[0] || d->peer[k].buf[1]) {
      if (!d->peer[k].buf[0] || !d->peer[k].buf[1] || !d->peer[k].flag) {
        h2d_fail(H2D_EINVAL, "h2d_plan_create: peer %d needs both buffers and a flag", k);
        return nullptr;
      }
      has_peers = true;
    }
  for (int side = 0; side < 2; ++side)
    if ((d->colbuf_in[side][0] == nullptr) != (d->colbuf_in[side][1] == nullptr)) {
      h2d_fail(H2D_EINVAL, "h2d_plan_create: colbuf_in[%d] needs both parities", side);
      return nullptr;
    }
  if (has_peers && !d->flags_in) {
    h2d_fail(H2D_EINVAL, "h2d_plan_create: peers given but flags_in is null");
    return nullptr;
  }
  if (has_peers && (d->kmax > t.rows || d->kmax > t.cols)) {
    h2d_fail(H2D_EINVAL, "h2d_plan_create: kmax %d exceeds the tile extent %dx%d", d->kmax, t.rows,
             t.cols);
    return nullptr;
  }

  h2d_plan* p = new h2d_plan();
  p->desc = *d;
  p->geo = make_geo(t, d->phys);
  p->has_peers = has_peers;
  p->lay = pick_layout(d->phys.arith, d->layout, d->kmax);
  cudaGetDevice(&p->device);
  int max_tasks = 0;
  const int min_waves = has_peers ? peer_min_waves() : 1;
  plan_counts(&t, d->kmax, d->chunk_rows, p->lay, &p->n_chunks, &max_tasks, min_waves);
  p->chunk_rows = d->chunk_rows > 0 ? d->chunk_rows : choose_chunk_rows(t.rows, t.cols, d->kmax, p->lay, min_waves);
  if (p->n_chunks == 1) p->chunk_rows = t.rows;
  p->chunk_rows_edge = edge_chunk_rows(p, p->chunk_rows, d->kmax);
  // (the column-band strips may hold more, shorter tasks than plan_counts() assumed)
  max_tasks += 3 * ((t.rows + p->chunk_rows_edge - 1) / p->chunk_rows_edge);
  p->fw = d->kmax;
  if (p->fw > t.rows) p->fw = t.rows;
  if (p->fw > t.cols) p->fw = t.cols;
  p->max_ctas = (max_tasks + TK_WARPS - 1) / TK_WARPS;

  if (make_tmap(&p->tmap[0], d->buf[0], t, p->lay) || make_tmap(&p->tmap[1], d->buf[1], t, p->lay)) {
    delete p;
    return nullptr;
  }
  for (int H = 4; H <= Hmax; H += 4)
    if (build_task_list(p, H, &p->lists[H / 4])) {
      h2d_plan_destroy(p);
      return nullptr;
    }
  const size_t sb = scratch_layout_bytes(max_tasks);
  if (d->scratch) {
    p->scratch = d->scratch;
  } else {
    if (cudaMalloc(&p->scratch, sb) != cudaSuccess) {
      h2d_fail(H2D_ENOMEM, "h2d_plan_create: cannot allocate %zu bytes of scratch", sb);
      h2d_plan_destroy(p);
      return nullptr;
    }
    p->own_scratch = true;
  }
  if (cudaMemset(p->scratch, 0, 256) != cudaSuccess) {
    h2d_fail(H2D_ECUDA, "h2d_plan_create: cannot clear the scratch header");
    h2d_plan_destroy(p);
    return nullptr;
  }
  // Opt in to the dynamic shared memory size once per kernel variant.
  const bool imm = coef_is_reference(p->geo, d->phys.arith);
  for (int k = 1; k <= d->kmax; ++k) {
    for (int x = 0; x < 2; ++x) {
      tk_fn_t fn = tk_pick(k, d->phys.arith, p->lay, imm, x != 0);
      if (!fn) continue;   // an experiment build without the shallow kernels: the launch reports it
      cudaFuncSetAttribute((const void*)fn, cudaFuncAttributeMaxDynamicSharedMemorySize,
                           tk_smem_bytes(p->lay, x != 0));
    }
  }
  if (d->team != 0) h2d_plan_team_prepare(p);   // same for the team kernels it may launch
  return p;
}

int32_t h2d_plan_layout(const h2d_plan_t* p) { return p ? p->lay + 1 : 0; }

int h2d_plan_seed_colbuf(h2d_plan_t* p, int parity, void* stream) {
  if (!p || (parity != 0 && parity != 1)) return h2d_fail(H2D_EINVAL, "h2d_plan_seed_colbuf: bad argument");
  const h2d_tile_t& t = p->desc.tile;
  float* w = p->desc.colbuf_in[0][parity];
  float* e = p->desc.colbuf_in[1][parity];
  if (!w && !e) return H2D_OK;
  const int64_t n = (int64_t)t.rows * TK_CW;
  const int blocks = (int)((n + 255) / 256 > 1184 ? 1184 : (n + 255) / 256);
  seed_colbuf_kernel<<<blocks, 256, 0, (cudaStream_t)stream>>>(p->desc.buf[parity], p->geo, p->fw, w, e);
  H2D_CHECK_LAUNCH("seed_colbuf_kernel");
  return H2D_OK;
}

int h2d_plan_launch(h2d_plan_t* p, int src_parity, int k, uint32_t epoch, int norm, void* stream) {
  return h2d_plan_launch_ex(p, src_parity, k, epoch, norm, nullptr, 0u, stream);
}

int h2d_plan_launch_ex(h2d_plan_t* p, int src_parity, int k, uint32_t epoch, int norm,
                       const uint32_t* ctl, uint32_t run_if, void* stream) {
  if (!p) return h2d_fail(H2D_EINVAL, "h2d_plan_launch: null plan");
  if (src_parity != 0 && src_parity != 1) return h2d_fail(H2D_EINVAL, "h2d_plan_launch: bad parity");
  if (k < 1 || k > p->desc.kmax)
    return h2d_fail(H2D_EINVAL, "h2d_plan_launch: k = %d outside [1, kmax = %d]", k, p->desc.kmax);
  if (norm < H2D_NORM_NONE || norm > H2D_NORM_LINF)
    return h2d_fail(H2D_EINVAL, "h2d_plan_launch: unknown norm %d", norm);
#ifdef H2D_TK_FORCE_XTR
  const bool xtr = true;
#else
  const bool xtr = p->has_peers || norm != H2D_NORM_NONE;
#endif
  // h2d_plan_desc_t::team: sweeps without neighbours and without a norm go out as cluster
  // launches whose strips share their edge columns (stencil_team.cu) where the tile allows it
  if (p->desc.team != 0 && !xtr && h2d_plan_team_eligible(p, k, 1))
    return h2d_plan_launch_team_ex(p, src_parity, k, 1, 0, ctl, run_if, stream);
  const bool imm = coef_is_reference(p->geo, p->desc.phys.arith);
  tk_fn_t fn = tk_pick(k, p->desc.phys.arith, p->lay, imm, xtr);
  if (!fn) return h2d_fail(H2D_EINVAL, "h2d_plan_launch: no kernel for k = %d", k);
  const TaskList& tl = p->lists[round_up4(k) / 4];
  cudaStream_t st = (cudaStream_t)stream;

  TkArgs a;
  fill_common(p, a, src_parity, k);
  a.tasks = tl.dev;
  a.n_tasks = tl.n_tasks;
  a.n_tasks_all = tl.n_tasks;
  a.n_chunks = p->n_chunks;
  a.chunk_rows = p->chunk_rows;
  a.chunk_rows_edge = p->chunk_rows_edge;
  a.row_base = 0;
  a.row_end = p->desc.tile.rows;
  a.n_strips = tl.n_strips;
  a.n_comm_edge = tl.n_comm_edge;
  fill_peers(p, a, src_parity, epoch);
  a.norm = norm;
  a.ctl = ctl;
  a.run_if = run_if;

  // Optional (H2D_SPLIT_LAUNCH=1): a launch with neighbours and no norm as TWO kernels side by
  // side — the edge tasks in the forwarding kernel on the caller's stream, all others in the
  // plain kernel on a side stream forked off and joined back around the pair.  It was built
  // when forwarding still lived inside the row loop and edge tasks ran several times longer
  // than interior ones; with the post-pass forwarding one kernel is as fast and costs half
  // the launches, so the split is off by default.
  static const bool split_ok = getenv("H2D_SPLIT_LAUNCH") && getenv("H2D_SPLIT_LAUNCH")[0] == '1';
  const int n_edge = tl.n_comm_edge, n_inner = tl.n_tasks - tl.n_comm_edge;
  tk_fn_t fn_plain = tk_pick(k, p->desc.phys.arith, p->lay, imm, false);
  if (split_ok && p->has_peers && norm == H2D_NORM_NONE && n_edge > 0 && n_inner > 0 && fn_plain) {
    if (!p->s_side) {
      H2D_CUDA_TRY(cudaStreamCreateWithFlags(&p->s_side, cudaStreamNonBlocking));
      H2D_CUDA_TRY(cudaEventCreateWithFlags(&p->ev_fork, cudaEventDisableTiming));
      H2D_CUDA_TRY(cudaEventCreateWithFlags(&p->ev_join, cudaEventDisableTiming));
    }
    H2D_CUDA_TRY(cudaEventRecord(p->ev_fork, st));
    H2D_CUDA_TRY(cudaStreamWaitEvent(p->s_side, p->ev_fork, 0));
    // edge tasks first: their flags gate the neighbours' next launch
    TkArgs ae = a;
    ae.n_tasks = n_edge;
    fn<<<(n_edge + TK_WARPS - 1) / TK_WARPS, TK_WARPS * 32, tk_smem_bytes(p->lay, true), st>>>(p->tmap[src_parity], ae);
    H2D_CHECK_LAUNCH("tk_kernel (edge tasks)");
    TkArgs ai = a;
    ai.tasks = tl.dev + n_edge;
    ai.n_tasks = n_inner;
    ai.has_peers = 0;
    ai.n_comm_edge = 0;
    fn_plain<<<(n_inner + TK_WARPS - 1) / TK_WARPS, TK_WARPS * 32, tk_smem_bytes(p->lay, false), p->s_side>>>(
        p->tmap[src_parity], ai);
    H2D_CHECK_LAUNCH("tk_kernel (interior tasks)");
    H2D_CUDA_TRY(cudaEventRecord(p->ev_join, p->s_side));
    H2D_CUDA_TRY(cudaStreamWaitEvent(st, p->ev_join, 0));
    p->launches += 2;
    p->seq++;
    return H2D_OK;
  }

  const int grid = (tl.n_tasks + TK_WARPS - 1) / TK_WARPS;
  fn<<<grid, TK_WARPS * 32, tk_smem_bytes(p->lay, xtr), st>>>(p->tmap[src_parity], a);
  H2D_CHECK_LAUNCH("tk_kernel");
  p->launches++;
  p->seq++;
  return H2D_OK;
}

int h2d_plan_run_many(h2d_plan_t* const* plans, int n_plans, int src_parity, int k, int n_launches,
                      uint32_t epoch0, const uint32_t* const* ctls, void* stream) {
  if (!plans || n_plans < 1 || n_launches < 0) return h2d_fail(H2D_EINVAL, "h2d_plan_run_many: bad argument");
  int par = src_parity;
  for (int l = 0; l < n_launches; ++l) {
    for (int i = 0; i < n_plans; ++i)
      if (int rc = h2d_plan_launch_ex(plans[i], par, k, epoch0 + (uint32_t)l, H2D_NORM_NONE,
                                      ctls ? ctls[i] : nullptr, 0u, stream))
        return rc;
    par ^= 1;
  }
  return H2D_OK;
}

// One k-step launch restricted to owned rows [row_base, row_end) of a single tile
// (no norm): tasks are enumerated arithmetically, chunk c starts at row_base + c * R.
static int launch_window(h2d_plan* p, int src_parity, int k, int row_base, int row_end, int chunk_rows,
                         cudaStream_t st, uint32_t epoch = 0u, bool unpack = true) {
  if (row_end <= row_base) return H2D_OK;
  const bool peers = p->has_peers;
  tk_fn_t fn = tk_pick(k, p->desc.phys.arith, p->lay, coef_is_reference(p->geo, p->desc.phys.arith), peers);
  if (!fn) return h2d_fail(H2D_EINVAL, "launch_window: no kernel for k = %d", k);
  TkArgs a;
  fill_common(p, a, src_parity, k);
  a.tasks = nullptr;
  if (!unpack) a.colbuf_in[0] = a.colbuf_in[1] = nullptr;   // the ghost columns of this launch came from the host
  a.chunk_rows = chunk_rows;
  a.row_base = row_base;
  a.row_end = row_end;
  a.n_strips = (p->desc.tile.cols + a.wv - 1) / a.wv;
  a.n_chunks = (row_end - row_base + a.chunk_rows - 1) / a.chunk_rows;
  a.chunk_rows_edge = a.chunk_rows;
  a.n_chunks_edge = a.n_chunks;
  a.n_tasks = a.n_strips * a.n_chunks;
  if (peers) {
    // W / E neighbours only (checked by the caller).  In the flag protocol: the first and the
    // last strip of every chunk (they forward the column bands) and every strip whose stored
    // cells depend on a ghost column — with the last strip shifted to end at the tile edge that
    // can only be the one before it, when the last strip owns fewer than fw columns.  These
    // strips run in short chunks (see edge_chunk_rows).
    const int ns = a.n_strips;
    int ne = ns >= 2 ? 2 : 1;
    if (ns >= 3 && p->desc.tile.cols - (ns - 1) * a.wv < a.fw) ne = 3;
    a.n_edge_strips = ne;
    a.chunk_rows_edge = p->chunk_rows_edge < a.chunk_rows ? p->chunk_rows_edge : a.chunk_rows;
    a.n_chunks_edge = (row_end - row_base + a.chunk_rows_edge - 1) / a.chunk_rows_edge;
    a.n_comm_edge = ne * a.n_chunks_edge;
    a.n_tasks = a.n_comm_edge + (ns - ne) * a.n_chunks;
    fill_peers(p, a, src_parity, epoch);
  }
  a.n_tasks_all = a.n_tasks;
  const int grid = (a.n_tasks + TK_WARPS - 1) / TK_WARPS;
  fn<<<grid, TK_WARPS * 32, tk_smem_bytes(p->lay, peers), st>>>(p->tmap[src_parity], a);
  H2D_CHECK_LAUNCH("tk_kernel (row window)");
  p->launches++;
  p->seq++;
  return H2D_OK;
}

// Rows of one wave of window tasks for this plan (see h2d_plan_solve_pipelined): ranks of a
// column-strip run agree on the minimum of this value so that they cut the same blocks.
// Tuning hooks of the pipelined solve (sweeps without rebuilding; defaults unchanged):
// H2D_PIPE_CHUNK_MAX = largest chunk (default 384 rows), H2D_PIPE_BLOCK_WAVES = waves of
// tasks per plateau block (default 4).
static int env_int(const char* name, int lo, int hi, int dflt) {
  const char* e = getenv(name);
  if (e && e[0]) {
    const int v = atoi(e);
    if (v >= lo && v <= hi) return v;
  }
  return dflt;
}
static int pipelined_chunk_rows(const h2d_plan* p) {
  const int cap = env_int("H2D_PIPE_CHUNK_MAX", 16, 1 << 20, 384);
  int R = p->chunk_rows;
  while (R > cap) R = (R + 1) / 2;
  return R;
}
static int pipelined_wave_rows(const h2d_plan* p) {
  const int wv_ = tk_strip_width(p->lay) - 2 * round_up4(p->desc.kmax);
  const int ns_ = (p->desc.tile.cols + wv_ - 1) / wv_;
  const double slots_ = (double)sm_count() * warps_per_sm(p->desc.kmax, p->lay);
  int m1 = (int)(slots_ / ns_);
  if (m1 < 1) m1 = 1;
  return m1 * pipelined_chunk_rows(p);
}

int32_t h2d_plan_wave_rows(const h2d_plan_t* p) { return p ? pipelined_wave_rows(p) : 0; }

int h2d_plan_solve_pipelined(h2d_plan_t* p, const float* host_in, int64_t in_pitch, float* host_out,
                             int64_t out_pitch, int64_t steps, int k, int block_rows, int skew_rows,
                             void* stream, int* final_parity) {
  if (p && p->has_peers) return h2d_fail(H2D_EINVAL, "h2d_plan_solve_pipelined: single-tile plans only (see _ex)");
  return h2d_plan_solve_pipelined_ex(p, host_in, in_pitch, host_out, out_pitch, steps, k, block_rows, skew_rows,
                                     0, 0u, stream, final_parity, nullptr);
}

// One plan's share of a pipelined solve: validation, geometry, copy streams and events.
struct PipeRank {
  h2d_plan* p = nullptr;
  const float* host_in = nullptr;
  float* host_out = nullptr;
  int64_t in_pitch = 0, out_pitch = 0;
  int up_c0 = 0, up_w = 0, R = 0;
  size_t ev_used = 0;
  std::vector<cudaEvent_t> up;   // one event per upload unit
  int waited = -1;               // last upload unit the compute stream waits for
  int d2h_done_rows = 0;         // rows [0, d2h_done_rows) already queued for download

  int new_event(cudaEvent_t* out) {
    if (ev_used == p->ev_pool.size()) {
      cudaEvent_t e;
      H2D_CUDA_TRY(cudaEventCreateWithFlags(&e, cudaEventDisableTiming));
      p->ev_pool.push_back(e);
    }
    *out = p->ev_pool[ev_used++];
    return H2D_OK;
  }
  float* dev_row(int parity, int r) const {
    const h2d_tile_t& t = p->desc.tile;
    return p->desc.buf[parity] + (int64_t)(r + t.halo_r) * t.pitch + t.halo_c;
  }
};

static int pipe_check(PipeRank& q, int k) {
  h2d_plan* p = q.p;
  if (!p || !q.host_in || !q.host_out) return h2d_fail(H2D_EINVAL, "h2d_plan_solve_pipelined: bad argument");
  if (k < 1 || k > p->desc.kmax) return h2d_fail(H2D_EINVAL, "h2d_plan_solve_pipelined: k = %d outside [1, %d]", k, p->desc.kmax);
  const h2d_tile_t& t = p->desc.tile;
  if (q.in_pitch < t.cols || q.out_pitch < t.cols) return h2d_fail(H2D_EINVAL, "h2d_plan_solve_pipelined: host pitch < cols");
  // Neighbours: column strips only.  Every rank then owns ALL rows, cuts the same row
  // blocks and issues the same sequence of window launches, one epoch each; the W / E
  // bands travel per window exactly as they do per launch.  (Row strips cannot be
  // pipelined this way: a strip's top halo is produced by the LAST block of the strip
  // above it.)  The level-0 ghost columns come from the host: host_in must hold
  // TK_WE_COLS valid columns to the left / right of the owned ones where a W / E
  // neighbour exists.
  const bool has_w = p->desc.peer[H2D_W].buf[0] != nullptr, has_e = p->desc.peer[H2D_E].buf[0] != nullptr;
  if (p->has_peers) {
    for (int d = 0; d < H2D_NDIR; ++d)
      if (d != H2D_W && d != H2D_E && p->desc.peer[d].buf[0] != nullptr)
        return h2d_fail(H2D_EINVAL, "h2d_plan_solve_pipelined: neighbours other than W / E (use a 1 x P process grid)");
    if ((t.cols & 3) != 0 || t.cols < 2 * TK_WE_COLS || t.halo_c < TK_WE_COLS ||
        (has_w && p->desc.peer[H2D_W].halo_c < TK_WE_COLS) || (has_e && p->desc.peer[H2D_E].halo_c < TK_WE_COLS))
      return h2d_fail(H2D_EINVAL, "h2d_plan_solve_pipelined: column strips need cols %% 4 == 0, cols >= %d and %d-column margins",
                      2 * TK_WE_COLS, TK_WE_COLS);
    if (q.in_pitch < t.cols + (has_w ? TK_WE_COLS : 0) + (has_e ? TK_WE_COLS : 0))
      return h2d_fail(H2D_EINVAL, "h2d_plan_solve_pipelined: host_in pitch leaves no room for the ghost columns");
  }
  q.up_c0 = has_w ? -TK_WE_COLS : 0;                                   // first uploaded column (tile-local)
  q.up_w = t.cols + (has_w ? TK_WE_COLS : 0) + (has_e ? TK_WE_COLS : 0);
  q.R = pipelined_chunk_rows(p);
  q.ev_used = 0;
  q.waited = -1;
  q.d2h_done_rows = 0;
  return H2D_OK;
}

// The pipelined solve for one or several plans that share one compute stream.  Several plans
// = the tiles of a 1 x P grid emulated in one process: every window is issued for every plan
// before the next one, so a plan's flags are always published by launches that precede its
// neighbours' waits in stream order.  With one plan per process (real ranks) the same code
// runs with n = 1 on every rank.
static int pipelined_solve(PipeRank* qs, int n, int64_t steps, int k, int block_rows, int skew_rows,
                           int wave_rows, uint32_t epoch0, cudaStream_t cs, int* final_parity,
                           uint32_t* epochs_used) {
  if (steps < 0 || n < 1) return h2d_fail(H2D_EINVAL, "h2d_plan_solve_pipelined: bad argument");
  for (int i = 0; i < n; ++i)
    if (int rc = pipe_check(qs[i], k)) return rc;
  const int rows = qs[0].p->desc.tile.rows;
  for (int i = 1; i < n; ++i)
    if (qs[i].p->desc.tile.rows != rows)
      return h2d_fail(H2D_EINVAL, "h2d_plan_solve_pipelined: all tiles must own the same rows");
  uint32_t epoch = epoch0;
  // windows are a few waves of tasks only: shorter chunks than whole-grid launches use
  // (measured on B200, 32768^2: chunks of ~320 rows and blocks of ~7600 rows are best)
  const int S = skew_rows > 0 ? skew_rows : k + 4;       // rows the processed window moves up per launch
  if (S < k + 1) return h2d_fail(H2D_EINVAL, "h2d_plan_solve_pipelined: skew %d must exceed k = %d", S, k);
  const int L = (int)((steps + k - 1) / k);              // launches ("levels") per block
  // Block sizes.  Large blocks compute efficiently (several waves of tasks per launch) but
  // the first block delays the start by its upload and the last one exposes its download, so
  // the sizes ramp up from one wave of tasks to `block_rows` (0 = four waves) and down again.
  int M1 = wave_rows;                                     // rows of one wave of tasks
  if (M1 <= 0) {
    M1 = pipelined_wave_rows(qs[0].p);
    for (int i = 1; i < n; ++i) {
      const int w = pipelined_wave_rows(qs[i].p);
      if (w < M1) M1 = w;
    }
  }
  // How big a block should be depends on what bounds the solve.  With few launches per block
  // (few steps) the copies dominate and only the fill / drain of the pipeline can be saved:
  // small blocks, even below one wave of tasks (32768^2, 20 steps on one B200: 108 ms with
  // 4-wave blocks, 92 ms with half-wave blocks; the PCIe floor is ~87 ms).  With many launches
  // per block the kernels dominate and want whole waves (1000 steps: 4 waves are best).
  // Narrow tiles (column strips on many GPUs) reach one wave only with thousands of rows:
  // the row-count caps keep enough blocks for upload, compute and download to overlap.
  int M;
  if (block_rows > 0) {
    M = block_rows;
  } else if (const int w = env_int("H2D_PIPE_BLOCK_WAVES", 0, 64, 0)) {
    M = w * M1;
  } else {
    const int half = M1 / 2 > 0 ? M1 / 2 : 1;
    M = L <= 8 ? half : L <= 32 ? M1 : L <= 64 ? 2 * M1 : 4 * M1;
    const int cap = L <= 8 ? rows / 8 : L <= 32 ? rows / 6 : rows / 4;
    if (M > cap) M = cap;
    if (M < 64) M = 64 < rows ? 64 : rows;
  }
  if (M < S + 1) M = S + 1;

  // upload in units of U rows, one event per unit; the copy streams start after whatever
  // the caller queued on `stream`
  int U = M >= 2048 ? 1024 : (M >= 512 ? 256 : M);
  if (U > rows) U = rows;
  const int n_units = (rows + U - 1) / U;
  for (int i = 0; i < n; ++i) {
    PipeRank& q = qs[i];
    h2d_plan* p = q.p;
    const h2d_tile_t& t = p->desc.tile;
    if (!p->s_h2d) H2D_CUDA_TRY(cudaStreamCreateWithFlags(&p->s_h2d, cudaStreamNonBlocking));
    if (!p->s_d2h) H2D_CUDA_TRY(cudaStreamCreateWithFlags(&p->s_d2h, cudaStreamNonBlocking));
    cudaEvent_t e_start;
    if (int rc = q.new_event(&e_start)) return rc;
    H2D_CUDA_TRY(cudaEventRecord(e_start, cs));
    H2D_CUDA_TRY(cudaStreamWaitEvent(p->s_h2d, e_start, 0));
    H2D_CUDA_TRY(cudaStreamWaitEvent(p->s_d2h, e_start, 0));
    q.up.assign((size_t)n_units, nullptr);
  }
  // unit-major, so that with several tiles in one process every tile's first rows arrive first
  for (int u = 0; u < n_units; ++u) {
    const int r0 = u * U, r1 = r0 + U < rows ? r0 + U : rows;
    for (int i = 0; i < n; ++i) {
      PipeRank& q = qs[i];
      const h2d_tile_t& t = q.p->desc.tile;
      H2D_CUDA_TRY(cudaMemcpy2DAsync(q.dev_row(0, r0) + q.up_c0, (size_t)t.pitch * 4,
                                     q.host_in + (int64_t)r0 * q.in_pitch + q.up_c0, (size_t)q.in_pitch * 4,
                                     (size_t)q.up_w * 4, (size_t)(r1 - r0), cudaMemcpyHostToDevice, q.p->s_h2d));
      if (int rc = q.new_event(&q.up[(size_t)u])) return rc;
      H2D_CUDA_TRY(cudaEventRecord(q.up[(size_t)u], q.p->s_h2d));
    }
  }

  const int out_par = L & 1;
  const int64_t span = (int64_t)rows + (int64_t)S * (L > 0 ? L - 1 : 0);
  // block boundaries over [0, span): ramp up, plateau of M, ramp down
  std::vector<int64_t> bounds;
  if (L > 0) {
    std::vector<int64_t> ramp;
    for (int64_t sz = M1; sz < M; sz *= 2) ramp.push_back(sz);
    int64_t ramp_sum = 0;
    for (int64_t v : ramp) ramp_sum += v;
    bounds.push_back(0);
    if (block_rows > 0 && block_rows <= M1) ramp.clear(), ramp_sum = 0;   // caller asked for small uniform blocks
    if (2 * ramp_sum + M <= span) {
      int64_t pos = 0;
      for (int64_t v : ramp) { pos += v; bounds.push_back(pos); }
      const int64_t mid_end = span - ramp_sum;
      while (pos + M < mid_end) { pos += M; bounds.push_back(pos); }
      if (mid_end > pos) { pos = mid_end; bounds.push_back(pos); }
      for (size_t i = ramp.size(); i-- > 0;) { pos += ramp[i]; bounds.push_back(pos); }
    } else {
      int64_t pos = 0;
      while (pos < span) { pos += M; bounds.push_back(pos < span ? pos : span); }
    }
    if (bounds.back() < span) bounds.push_back(span);
  }
  const int n_blocks = bounds.empty() ? 0 : (int)bounds.size() - 1;
  for (int b = 0; b < n_blocks; ++b) {
    const int64_t b_lo = bounds[b], b_hi = bounds[b + 1];
    // level 0 of this block reads rows below b_hi + k
    int64_t need_row = b_hi + k;
    if (need_row > rows) need_row = rows;
    const int need_unit = (int)((need_row - 1) / U);
    for (int i = 0; i < n; ++i)
      if (need_unit > qs[i].waited) {
        H2D_CUDA_TRY(cudaStreamWaitEvent(cs, qs[i].up[(size_t)need_unit], 0));
        qs[i].waited = need_unit;
      }
    for (int e = 0; e < L; ++e) {
      int64_t lo = b_lo - (int64_t)S * e, hi = b_hi - (int64_t)S * e;
      if (hi <= 0 || lo >= rows) continue;
      if (lo < 0) lo = 0;
      if (hi > rows) hi = rows;
      // equal depths (20 steps at k = 8 -> 7 + 7 + 6): a trailing shallow launch would be HBM-bound
      const int kk = (int)(steps / L) + (e < (int)(steps % L) ? 1 : 0);
      for (int i = 0; i < n; ++i)
        if (int rc = launch_window(qs[i].p, e & 1, kk, (int)lo, (int)hi, qs[i].R, cs, epoch, e > 0)) return rc;
      ++epoch;   // one epoch per issued window, on every rank alike
    }
    // rows below b_hi - S*(L-1) now hold the final field
    int64_t fin = b_hi - (int64_t)S * (L - 1);
    if (b == n_blocks - 1 || fin > rows) fin = rows;
    for (int i = 0; i < n; ++i) {
      PipeRank& q = qs[i];
      if (fin <= q.d2h_done_rows) continue;
      const h2d_tile_t& t = q.p->desc.tile;
      cudaEvent_t ev;
      if (int rc = q.new_event(&ev)) return rc;
      H2D_CUDA_TRY(cudaEventRecord(ev, cs));
      H2D_CUDA_TRY(cudaStreamWaitEvent(q.p->s_d2h, ev, 0));
      H2D_CUDA_TRY(cudaMemcpy2DAsync(q.host_out + (int64_t)q.d2h_done_rows * q.out_pitch, (size_t)q.out_pitch * 4,
                                     q.dev_row(out_par, q.d2h_done_rows), (size_t)t.pitch * 4, (size_t)t.cols * 4,
                                     (size_t)(fin - q.d2h_done_rows), cudaMemcpyDeviceToHost, q.p->s_d2h));
      q.d2h_done_rows = (int)fin;
    }
  }
  for (int i = 0; i < n; ++i) {
    PipeRank& q = qs[i];
    const h2d_tile_t& t = q.p->desc.tile;
    if (L == 0) {  // nothing to compute: upload, then download the same field
      H2D_CUDA_TRY(cudaStreamWaitEvent(q.p->s_d2h, q.up[(size_t)n_units - 1], 0));
      H2D_CUDA_TRY(cudaMemcpy2DAsync(q.host_out, (size_t)q.out_pitch * 4, q.dev_row(0, 0), (size_t)t.pitch * 4,
                                     (size_t)t.cols * 4, (size_t)rows, cudaMemcpyDeviceToHost, q.p->s_d2h));
    }
    // the caller's stream completes when the last download (and upload) has
    cudaEvent_t e_end;
    if (int rc = q.new_event(&e_end)) return rc;
    H2D_CUDA_TRY(cudaEventRecord(e_end, q.p->s_d2h));
    H2D_CUDA_TRY(cudaStreamWaitEvent(cs, e_end, 0));
    cudaEvent_t e_end2;
    if (int rc = q.new_event(&e_end2)) return rc;
    H2D_CUDA_TRY(cudaEventRecord(e_end2, q.p->s_h2d));
    H2D_CUDA_TRY(cudaStreamWaitEvent(cs, e_end2, 0));
  }
  if (final_parity) *final_parity = out_par;
  if (epochs_used) *epochs_used = epoch - epoch0;
  return H2D_OK;
}

int h2d_plan_solve_pipelined_ex(h2d_plan_t* p, const float* host_in, int64_t in_pitch, float* host_out,
                                int64_t out_pitch, int64_t steps, int k, int block_rows, int skew_rows,
                                int wave_rows, uint32_t epoch0, void* stream, int* final_parity,
                                uint32_t* epochs_used) {
  PipeRank q;
  q.p = p; q.host_in = host_in; q.host_out = host_out; q.in_pitch = in_pitch; q.out_pitch = out_pitch;
  return pipelined_solve(&q, 1, steps, k, block_rows, skew_rows, wave_rows, epoch0, (cudaStream_t)stream,
                         final_parity, epochs_used);
}

int h2d_plan_solve_pipelined_many(h2d_plan_t* const* plans, int n_plans, const float* const* host_in,
                                  const int64_t* in_pitch, float* const* host_out, const int64_t* out_pitch,
                                  int64_t steps, int k, int block_rows, int skew_rows, uint32_t epoch0,
                                  void* stream, int* final_parity, uint32_t* epochs_used) {
  if (!plans || n_plans < 1 || !host_in || !in_pitch || !host_out || !out_pitch)
    return h2d_fail(H2D_EINVAL, "h2d_plan_solve_pipelined_many: bad argument");
  std::vector<PipeRank> qs((size_t)n_plans);
  for (int i = 0; i < n_plans; ++i) {
    qs[(size_t)i].p = plans[i];
    qs[(size_t)i].host_in = host_in[i];
    qs[(size_t)i].host_out = host_out[i];
    qs[(size_t)i].in_pitch = in_pitch[i];
    qs[(size_t)i].out_pitch = out_pitch[i];
  }
  return pipelined_solve(qs.data(), n_plans, steps, k, block_rows, skew_rows, 0, epoch0, (cudaStream_t)stream,
                         final_parity, epochs_used);
}

int64_t h2d_plan_launch_count(const h2d_plan_t* p) { return p ? p->launches : 0; }
uint32_t* h2d_plan_strict_fallback_ptr(h2d_plan_t* p) { return p ? (uint32_t*)p->scratch + 12 : nullptr; }
uint32_t* h2d_plan_team_fallback_ptr(h2d_plan_t* p) { return p ? (uint32_t*)p->scratch + 15 : nullptr; }
float* h2d_plan_norm_ptr(h2d_plan_t* p) { return p ? (float*)((unsigned*)p->scratch + 2) : nullptr; }
uint32_t* h2d_plan_fault_ptr(h2d_plan_t* p) { return p ? (uint32_t*)p->scratch + 3 : nullptr; }
uint64_t* h2d_plan_stats_ptr(h2d_plan_t* p) { return p ? (uint64_t*)((uint32_t*)p->scratch + 8) : nullptr; }

}  // extern "C"
