// tk_task.cuh — device side of the flagship kernel (stencil_tk): arguments, PTX
// wrappers, halo forwarding, and the task bodies.
//
// One warp = one task = (strip of columns, chunk of R rows).  The warp streams
// R + 2K input rows top to bottom through its own 4-stage TMA ring; K levels of
// the stencil are pipelined in registers (three row slots per level, rotating
// roles, so the unrolled loop body has no register moves) and the K-th level is
// stored.  See stencil_tk.cu for what this replaces in the reference.
//
// Two families of task body:
//
//  * run_task_v4 (LAY_V4): 128-column strips, each lane owns 4 adjacent columns.
//    Used by the strict (fp64-intermediate) modes, whose arithmetic is scalar.
//  * run_task_ab (LAY_AB2 / LAY_AB4): fp32 mode.  A strip is cut into a left
//    half A and a right half B; a lane owns NC columns of A and the NC columns
//    of B at the same offset, and keeps them PACKED as (A_j, B_j) in one 64-bit
//    register pair.  Every operation of the update — the horizontal neighbour
//    sum included — is then one aligned FADD2 / FMUL2 / FFMA2 on two points:
//    5 packed instructions per 2 points, the minimum.  (With 4 adjacent
//    columns per lane the horizontal sums pair (c0,c1) with (c2,c3)-shifted-by-
//    one operands, which are not aligned register pairs: they had to be 4
//    scalar FADDs per 4 points, each costing a full slot of the packed pipe.)
//    The seam between the halves is closed by wrapping the lane shuffles
//    around: lane 31's right neighbour in A is lane 0's first column of B.
//    LAY_AB4 (256-column strips, 8 points per lane) also halves the share of
//    recomputed overlap columns (2H of 256 instead of 2H of 128).
#pragma once
#include <cuda.h>
#include <type_traits>
#include "device_common.cuh"

namespace h2d {

constexpr int TK_BR = 6;         // rows per TMA box (a multiple of the 3-row register rotation)
constexpr int TK_STAGES = 4;     // boxes in flight per warp
constexpr int TK_WARPS = 4;      // warps (tasks) per CTA
constexpr int TK_WE_ROWS = 16;   // rows of column-halo data staged per remote store
constexpr int TK_WE_COLS = 32;   // columns forwarded per row and side: one full 128-byte line
constexpr int TK_WE_Q = TK_WE_COLS / 4;                        // float4 pieces per row and side
constexpr int TK_WE_BYTES = 2 * TK_WE_ROWS * TK_WE_COLS * 4;   // per warp: W and E side
constexpr int TK_KMAX = 8;

enum { LAY_V4 = 0, LAY_AB2 = 1, LAY_AB4 = 2, LAY_COUNT = 3 };
template <int LAY> struct Lay;
template <> struct Lay<LAY_V4>  { static constexpr int W = 128, NC = 4, HALF = 0; };
template <> struct Lay<LAY_AB2> { static constexpr int W = 128, NC = 2, HALF = 64; };
template <> struct Lay<LAY_AB4> { static constexpr int W = 256, NC = 4, HALF = 128; };

__host__ __device__ constexpr int tk_strip_width(int lay) { return lay == LAY_AB4 ? 256 : 128; }
__host__ __device__ constexpr int tk_stage_bytes(int lay) { return TK_BR * tk_strip_width(lay) * 4; }
__host__ __device__ constexpr int tk_smem_ring(int lay) { return TK_WARPS * TK_STAGES * tk_stage_bytes(lay); }
constexpr int TK_SMEM_BARS = TK_WARPS * TK_STAGES * 8;
// XTR kernels also carry the per-warp staging block of the batched column halos
__host__ __device__ constexpr int tk_smem_bytes(int lay, bool xtr) {
  return tk_smem_ring(lay) + TK_SMEM_BARS + (xtr ? TK_WARPS * TK_WE_BYTES : 0);
}

struct PeerDev {
  float* buf;       // neighbour's padded array that holds the SAME parity as our dst
  uint32_t* flag;   // where we publish our epoch for that neighbour
  int32_t x0, y0, halo_r, halo_c;
  int64_t pitch;
};

struct TkArgs {
  Geo g;
  float* dst;
  const int2* tasks;     // .x = strip | (comm_edge << 31), .y = chunk
  int32_t n_tasks, n_chunks, chunk_rows, n_comm_edge;
  int32_t n_tasks_all;   // tasks of the whole logical launch (this kernel may run a part of them)
  int32_t row_base, row_end;   // rows [row_base, row_end) are processed; chunk c starts at row_base + c*R
  int32_t n_strips;            // tasks == nullptr: task -> (task % n_strips, task / n_strips)
  int32_t n_edge_strips;       // tasks == nullptr with peers: strips 0, n_strips-1 (, n_strips-2) come first
  int32_t H, wv, fw;     // column overlap, stored columns per strip, forwarded band width
  int32_t has_peers;
  PeerDev peer[H2D_NDIR];
  const uint32_t* flags_in;
  uint32_t epoch;
  unsigned* edge_counter;
  int32_t norm;
  float* cta_partials;
  unsigned* done_counter;
  float* norm_out;
  int64_t timeout_ns;
  uint32_t* fault;
  // Device-side run control: the launch does nothing unless *ctl == run_if
  // (ctl == nullptr: always run).  0 = still iterating; q > 0 = the q-th
  // convergence check stopped the run (set by norm_finish_kernel).
  const uint32_t* ctl;
  uint32_t run_if;
  unsigned long long* stats;   // [0] ns spent waiting for neighbour flags, [1] number of waits that blocked
  // Strict modes: tasks whose data left the fast path's domain in launch seq-1 / seq
  // (three rotating counters: read [(seq+2)%3], add to [seq%3], clear [(seq+1)%3]).
  unsigned* strict_fb;
  uint32_t seq;
  int32_t debug;   // bit 0: skip the peer stores (timing experiments only; H2D_DEBUG_NOFWD=1)
                   // bit 1: random per-task delays around the flag protocol (H2D_DEBUG_DELAY=1)
                   // bit 2: strict modes never take the fast path (H2D_STRICT_FAST=0)
};

// ------------------------------------------------------------ PTX wrappers
__device__ __forceinline__ uint32_t smem_u32(const void* p) {
  return (uint32_t)__cvta_generic_to_shared(p);
}
__device__ __forceinline__ void mbar_init(uint32_t bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_inval(uint32_t bar) {
  asm volatile("mbarrier.inval.shared::cta.b64 [%0];" ::"r"(bar) : "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint32_t bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes)
               : "memory");
}
__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity) {
  asm volatile(
      "{\n\t"
      ".reg .pred p;\n\t"
      "WAIT_%=:\n\t"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n\t"
      "@p bra DONE_%=;\n\t"
      "bra WAIT_%=;\n\t"
      "DONE_%=:\n\t"
      "}" ::"r"(bar), "r"(parity)
      : "memory");
}
// The tensor map always describes THIS GPU's own padded array (TMA reads local memory only).
__device__ __forceinline__ void tma_load_2d(uint32_t dst, const CUtensorMap* map, int c0, int c1,
                                            uint32_t bar) {
  asm volatile(
      "cp.async.bulk.tensor.2d.shared::cluster.global.tile.mbarrier::complete_tx::bytes"
      " [%0], [%1, {%2, %3}], [%4];" ::"r"(dst), "l"(map), "r"(c0), "r"(c1), "r"(bar)
      : "memory");
}
__device__ __forceinline__ float4 lds128(uint32_t addr) {
  float4 v;
  asm volatile("ld.shared.v4.f32 {%0, %1, %2, %3}, [%4];"
               : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w)
               : "r"(addr));
  return v;
}
__device__ __forceinline__ float2 lds64(uint32_t addr) {
  float2 v;
  asm volatile("ld.shared.v2.f32 {%0, %1}, [%2];" : "=f"(v.x), "=f"(v.y) : "r"(addr));
  return v;
}

template <int NV> struct Vals { float v[NV]; };

// --------------------------------------------------- peer forwarding (cold)
// Store the NV just-computed values of owned row `rho`, columns j0..j0+NV-1, into
// every neighbour whose ghost margin contains them.  Called only by tasks that
// touch the tile edge.  `smask` = which of the elements are owned outputs.
template <int NV>
__device__ __noinline__ void forward_row(const TkArgs& a, int rho, int j0, Vals<NV> o, int smask,
                                         bool skip_we) {
  const int fw = a.fw;
#pragma unroll 1
  for (int d = 0; d < H2D_NDIR; ++d) {
    const PeerDev& p = a.peer[d];
    if (p.buf == nullptr) continue;
    const int sx = (d == H2D_N || d == H2D_NW || d == H2D_NE) ? -1
                 : (d == H2D_S || d == H2D_SW || d == H2D_SE) ? 1 : 0;
    const int sy = (d == H2D_W || d == H2D_NW || d == H2D_SW) ? -1
                 : (d == H2D_E || d == H2D_NE || d == H2D_SE) ? 1 : 0;
    if (sx == 0 && skip_we) continue;   // pure W / E: batched by flush_we()
    if (sx < 0 && rho >= fw) continue;
    if (sx > 0 && rho < a.g.rows - fw) continue;
    int m = 0;
#pragma unroll
    for (int e = 0; e < NV; ++e) {
      const int j = j0 + e;
      const bool in = (sy == 0) || (sy < 0 ? j < fw : j >= a.g.cols - fw);
      if (in && ((smask >> e) & 1)) m |= 1 << e;
    }
    if (m == 0 || (a.debug & 1)) continue;
    float* q = p.buf + (int64_t)(a.g.x0 + rho - p.x0 + p.halo_r) * p.pitch +
               (a.g.y0 + j0 - p.y0 + p.halo_c);
    if (NV == 4 && m == 15 && (((uintptr_t)q) & 15) == 0) {
      *reinterpret_cast<float4*>(q) = make_float4(o.v[0], o.v[1], o.v[2], o.v[3]);
    } else if (NV == 2 && m == 3 && (((uintptr_t)q) & 7) == 0) {
      *reinterpret_cast<float2*>(q) = make_float2(o.v[0], o.v[1]);
    } else {
#pragma unroll
      for (int e = 0; e < NV; ++e)
        if ((m >> e) & 1) q[e] = o.v[e];
    }
  }
}

// Column halos, batched and line-sized.  A column halo is only k floats per row,
// and 32-byte remote stores are expensive twice over: they stall the issuing warp
// for about an NVLink round trip each, and a partial-line write costs the
// destination a read-modify-write (1x2 split: 6770 GLUP/s row by row, 7060 with 16
// rows per store instruction, vs 8052 for the 2x1 split on two B200).  So each
// row forwards a whole 128-byte line — the 32 edge columns, which is exactly the
// neighbour's 32-column margin — parked in shared memory and sent every
// TK_WE_ROWS rows: lane l sends pieces l, l+32, ... of the 16 x 8 float4 block.
// `sides` bit 0: this warp holds the W band, bit 1: the E band.
static __device__ __noinline__ void flush_we(const TkArgs& a, uint32_t stage, int rho0, int count, int sides,
                                      int lane) {
  __syncwarp();
#pragma unroll
  for (int side = 0; side < 2; ++side) {
    const PeerDev& p = a.peer[side == 0 ? H2D_W : H2D_E];
    if (p.buf == nullptr || !((sides >> side) & 1)) continue;
#pragma unroll
    for (int it = 0; it < (TK_WE_ROWS * TK_WE_Q) / 32; ++it) {
      const int idx = it * 32 + lane;
      const int row = idx / TK_WE_Q, piece = idx % TK_WE_Q;
      if (row >= count) continue;
      const float4 v = lds128(stage + (uint32_t)(((side * TK_WE_ROWS + row) * TK_WE_Q + piece) * 16));
      const int jcol = (side == 0 ? 0 : a.g.cols - TK_WE_COLS) + piece * 4;   // tile-local column
      float* q = p.buf + (int64_t)(a.g.x0 + rho0 + row - p.x0 + p.halo_r) * p.pitch +
                 (a.g.y0 + jcol - p.y0 + p.halo_c);
      *reinterpret_cast<float4*>(q) = v;
    }
  }
  __syncwarp();
}

// XM: 0 = plain task; 1 = the task feeds the convergence norm; 2 = it also (or
// only) forwards rows to neighbours.  Warp-uniform and resolved at compile time
// so the common case carries no overhead in its inner loop.
constexpr int XM_NONE = 0, XM_NORM = 1, XM_COMM = 2;

// -------------------------------------------------------------------------
// Per-segment bookkeeping shared by both task bodies.  A segment = NV adjacent
// columns a lane owns (V4: one segment of 4; AB: one in each half).
struct SegInfo {
  int j0;        // tile-local column of element 0
  int smask;     // which elements are stored (inside the strip's valid window AND owned)
  int we_off;    // byte offset of this segment inside a staged W/E row block, or -1
};

// Everything a row of output needs besides the values: built once per task.
template <int NSEG>
struct OutCtx {
  SegInfo seg[NSEG];
  bool we_batched, we_perrow, we_skip;
  int we_sides, we_count, we_rho0;
};

template <int NSEG, int NV, bool COMM>
__device__ __forceinline__ void setup_out(OutCtx<NSEG>& oc, const TkArgs& a, bool comm_edge, int strip,
                                          const bool (&seg_valid)[NSEG]) {
  const Geo& g = a.g;
  oc.we_batched = false; oc.we_perrow = false; oc.we_skip = false;
  oc.we_sides = 0; oc.we_count = 0; oc.we_rho0 = 0;
#pragma unroll
  for (int s = 0; s < NSEG; ++s) oc.seg[s].we_off = -1;
  if (COMM) {
    const bool w_peer = a.peer[H2D_W].buf != nullptr, e_peer = a.peer[H2D_E].buf != nullptr;
    const bool w_ok = w_peer && a.peer[H2D_W].halo_c >= TK_WE_COLS;
    const bool e_ok = e_peer && a.peer[H2D_E].halo_c >= TK_WE_COLS;
    // Batched mode is usable when the 32-column bands are whole segments of single lanes
    // and fit the neighbour's margin; it then covers EVERY W/E band of the tile (a strip
    // without band columns simply stages nothing).
    const bool batched = comm_edge && a.fw <= 8 && (g.cols & 3) == 0 && g.cols >= 2 * TK_WE_COLS &&
                         (w_ok || e_ok) && (!w_peer || w_ok) && (!e_peer || e_ok);
    oc.we_skip = batched;
    oc.we_perrow = comm_edge && !batched && (w_peer || e_peer);
    if (batched) {
      // the last strip is shifted to end at the tile edge, so the E band sits in ONE warp
      if (w_ok && strip == 0) oc.we_sides |= 1;
      if (e_ok && strip == a.n_strips - 1) oc.we_sides |= 2;
#pragma unroll
      for (int s = 0; s < NSEG; ++s) {
        const int j0 = oc.seg[s].j0;
        if (!seg_valid[s] || j0 < 0 || j0 + NV > g.cols) continue;   // computed here, owned or not
        if ((oc.we_sides & 1) && j0 < TK_WE_COLS) oc.seg[s].we_off = j0 * 4;
        if ((oc.we_sides & 2) && j0 >= g.cols - TK_WE_COLS)
          oc.seg[s].we_off = TK_WE_ROWS * TK_WE_COLS * 4 + (j0 - (g.cols - TK_WE_COLS)) * 4;
      }
      oc.we_batched = oc.we_sides != 0;
    }
  }
}

// ============================================================================
//                     strict arithmetic: the fast evaluation
// ============================================================================
// The reference evaluates a point as  r = fma(cy, t2, fma(cx, t1, (double)c)),
// t1 = (double)fl32(xp+xm) - 2(double)c, t2 likewise, out = (float)r  — four
// conversions per point, and F2F runs at a quarter of the fp64 rate: the plain
// strict kernel is conversion-bound.  For NON-NEGATIVE FINITE floats the three
// up-conversions can be done with two integer instructions each and no
// special cases:  the bit pattern b of a float, placed as  hi = b >> 3,
// lo = b << 29,  is the double  f * 2^-896  — exactly, for normal floats
// (exponent field e_f lands in an 11-bit field), for zero (all bits 0), and
// for subnormal floats (exponent field 0 -> a subnormal double with the same
// fraction: (m/2^23) 2^-1022 = (m/2^23) 2^-126 * 2^-896).  In that scaled
// domain  t' = ew' - 2c'  is the exactly scaled t (differences of scaled floats
// are multiples of 2^-1045: exact when subnormal, identically rounded when
// normal), and the two FMAs take the scale back through their constants:
//   r = fma(cy 2^896, t2', fma(cx 2^896, t1', c' 2^896))
// has, operation by operation, the same exact real argument and the same single
// rounding as the reference expression.  Cost per point: 3x(shift, shift) +
// 1 DMUL + 4 DFMA + 1 F2F instead of 4 F2F + 4 DFMA.
//
// Outside the domain (a negative value incl. -0, Inf/NaN, or a magnitude
// >= 2^127 whose neighbour sum could overflow) the task is RE-RUN with the
// plain evaluation; `track` (unsigned max of every input and output bit
// pattern) detects it: all such patterns are >= 0x7f000000.
constexpr uint32_t STRICT_FAST_LIMIT = 0x7f000000u;

__device__ __forceinline__ double up_scaled(float f) {
  // (hi, lo) = (b >> 3, b << 29) is the 64-bit product b * 2^29: ONE IMAD.WIDE instead of two
  // shifts — the kernel is bound by instruction issue, not by a pipe
#ifndef H2D_STRICT_SHIFT_PAIR
  unsigned long long v;
  asm("mul.wide.u32 %0, %1, 0x20000000;" : "=l"(v) : "r"(__float_as_uint(f)));
  return __longlong_as_double((long long)v);
#else
  const uint32_t b = __float_as_uint(f);
  return __hiloint2double((int)(b >> 3), (int)(b << 29));
#endif
}

template <int ARITH>
__device__ __forceinline__ float point_fast(float c, float xm, float xp, float ym, float yp,
                                            const Geo& g) {
  const float ew = __fadd_rn(xp, xm);
  const float ns = __fadd_rn(yp, ym);
  const double cs = up_scaled(c);
  const double t1 = __fma_rn(-2.0, cs, up_scaled(ew));
  const double t2 = __fma_rn(-2.0, cs, up_scaled(ns));
  const double dc = __dmul_rn(cs, 0x1p896);
  double r;
  if (ARITH == H2D_ARITH_STRICT_FMA) {
    r = __fma_rn(g.dcy_s, t2, __fma_rn(g.dcx_s, t1, dc));
  } else {
    r = __dadd_rn(__dadd_rn(dc, __dmul_rn(g.dcx_s, t1)), __dmul_rn(g.dcy_s, t2));
  }
  return __double2float_rn(r);
}

template <int ARITH, bool IMM, bool FAST>
__device__ __forceinline__ float4 point4x(const float4& u, const float4& c, const float4& d,
                                          float L, float R, const Geo& g) {
  if (ARITH != H2D_ARITH_FP32 && FAST) {
    float4 o;
    o.x = point_fast<ARITH>(c.x, u.x, d.x, L, c.y, g);
    o.y = point_fast<ARITH>(c.y, u.y, d.y, c.x, c.z, g);
    o.z = point_fast<ARITH>(c.z, u.z, d.z, c.y, c.w, g);
    o.w = point_fast<ARITH>(c.w, u.w, d.w, c.z, R, g);
    return o;
  }
  return point4<ARITH, IMM>(u, c, d, L, R, g);
}

__device__ __forceinline__ uint32_t umax4(uint32_t t, const float4& v) {
  return max(max(t, __float_as_uint(v.x)), max(max(__float_as_uint(v.y), __float_as_uint(v.z)),
                                               __float_as_uint(v.w)));
}

// ============================================================================
//                        task body, 4 adjacent columns per lane
// ============================================================================
// FROZEN: the footprint contains cells that keep their value (global boundary
// or outside the grid) -> per-element select.  FAST (strict modes): see above;
// `track` receives the running maximum of the bit patterns.
template <int K, int ARITH, bool IMM, bool FROZEN, int XM, bool FAST>
__device__ __forceinline__ float run_task_v4(const CUtensorMap* tmap, const TkArgs& a, float* dst_base,
                                             int strip, int chunk, bool comm_edge, uint32_t ring,
                                             uint32_t bars, uint32_t we_stage, int lane, uint32_t& track,
                                             int box0 = 0) {
  // box0: boxes this warp's ring has already carried in this kernel (a task that is run a
  // second time continues the stage / phase sequence of its mbarriers instead of
  // re-initialising them)
  constexpr bool EXTRAS = XM != XM_NONE;   // accumulates the norm (when the launch has one)
  constexpr bool COMM = XM == XM_COMM;     // forwards to neighbours
  constexpr bool TRACK = ARITH != H2D_ARITH_FP32;
  constexpr int W = Lay<LAY_V4>::W;
  constexpr int STAGE_BYTES = TK_BR * W * 4;
  const Geo& g = a.g;
  const int R = a.chunk_rows;
  const int r_lo = a.row_base + chunk * R;
  const int r_hi = min(r_lo + R, a.row_end);
  const int n_out = r_hi - r_lo;
  // The last strip is shifted left so that its valid window ends at the tile edge: its
  // first columns are then computed twice (by it and by its left neighbour) but it OWNS
  // (stores / reduces) only its own range [own_lo, cols).  No strip is narrow and, what
  // matters for column halos, the last 32 columns always sit in ONE warp, so the east
  // band leaves as whole 128-byte lines.
  const int own_lo = strip * a.wv;
  // (rounded up to a multiple of 4: float4 stores and TMA box origins must stay 16-byte aligned)
  const int vbase = (strip == a.n_strips - 1 && g.cols >= a.wv) ? ((g.cols - a.wv + 3) & ~3) : own_lo;
  const int jbase = vbase - a.H;         // tile-local column of lane 0, element 0
  const int j0 = jbase + lane * 4;
  const int row0 = r_lo - K;             // first input row (tile-local)
  constexpr int LAG = 2 * K;  // row steps between reading a row and storing the row K below it
  const int n_in = n_out + LAG;
  const int n_blk = (n_in + TK_BR - 1) / TK_BR;
  const int tc0 = g.halo_c + jbase;      // TMA coordinates of the first box
  const int tc1 = g.halo_r + row0;

  // Which of my 4 columns are stored: inside the strip's valid window and owned.
  const bool lane_in = (lane * 4 >= a.H) && (lane * 4 < W - a.H);
  OutCtx<1> oc;
  oc.seg[0].j0 = j0;
  oc.seg[0].smask = 0;
#pragma unroll
  for (int e = 0; e < 4; ++e)
    if (lane_in && (j0 + e) >= own_lo && (j0 + e) < g.cols) oc.seg[0].smask |= 1 << e;
  const int smask = oc.seg[0].smask;
  const bool st_full = smask == 15;
  const bool st_part = smask != 0 && smask != 15;
  {
    const bool sv[1] = {lane_in};
    setup_out<1, 4, COMM>(oc, a, comm_edge, strip, sv);
  }

  bool act[4] = {true, true, true, true};
  if (FROZEN) {
#pragma unroll
    for (int e = 0; e < 4; ++e) {
      const int gy = g.y0 + j0 + e;
      act[e] = gy >= g.ay_lo && gy <= g.ay_hi;
    }
  }

  auto issue = [&](int b) {
    const int s = (box0 + b) % TK_STAGES;
    const uint32_t bar = bars + s * 8;
    mbar_expect_tx(bar, STAGE_BYTES);
    tma_load_2d(ring + s * STAGE_BYTES, tmap, tc0, tc1 + b * TK_BR, bar);
  };
  if (lane == 0) {
    const int pre = n_blk < TK_STAGES ? n_blk : TK_STAGES;
    for (int b = 0; b < pre; ++b) issue(b);
  }

  // Three row slots per level, rotating roles (up, centre, incoming) with period
  // 3: every producer writes straight into the slot its consumer will read, so a
  // loop body unrolled over 3 rows needs no register-to-register moves.
  float4 st[K][3];
#pragma unroll
  for (int t = 0; t < K; ++t) st[t][0] = st[t][1] = st[t][2] = make_float4(0.f, 0.f, 0.f, 0.f);

  float acc = 0.f;
  uint32_t trk = 0u;
  // Level-K output of iteration i is tile-local row  row0 - K + i;  `ro` counts
  // rows relative to r_lo so that "is an owned output row" is one unsigned compare.
  int ro = -LAG;
  float* q = dst_base + (int64_t)(g.halo_r + r_lo - LAG) * g.pitch + g.halo_c + j0;

  // One row step: the input row is already in st[0][(PH+2)%3]; levels run
  // 0..K-1, each writing its result straight into the next level's incoming slot.
  // FWD: this stretch of rows may forward to neighbours (compile-time, so that the
  // rows of an edge task that forward nothing run the same body as an interior task).
  auto row_step = [&](auto Ptag, auto Ftag) {
    constexpr int PH = decltype(Ptag)::value;
    constexpr bool FWD = COMM && decltype(Ftag)::value;
    constexpr int iu = PH, ic = (PH + 1) % 3, id = (PH + 2) % 3;
    float4 fin;
    float4 cprev = make_float4(0.f, 0.f, 0.f, 0.f);
    if (TRACK) trk = umax4(trk, st[0][id]);
#pragma unroll
    for (int t = 0; t < K; ++t) {
      const float4 up = st[t][iu];
      const float4 c = st[t][ic];
      const float4 down = st[t][id];
      const float L = __shfl_up_sync(0xffffffffu, c.w, 1);
      const float Rn = __shfl_down_sync(0xffffffffu, c.x, 1);
      float4 o = point4x<ARITH, IMM, FAST>(up, c, down, L, Rn, g);
      if (FROZEN) {
        // tile-local row of this level's output
        const int gx = g.x0 + r_lo + ro + (K - 1 - t);
        const bool ract = gx >= g.ax_lo && gx <= g.ax_hi;
        o.x = (ract && act[0]) ? o.x : c.x;
        o.y = (ract && act[1]) ? o.y : c.y;
        o.z = (ract && act[2]) ? o.z : c.z;
        o.w = (ract && act[3]) ? o.w : c.w;
      }
      if (TRACK) trk = umax4(trk, o);
      if (t + 1 < K) {
        st[t + 1][id] = o;
      } else {
        fin = o;
        if (EXTRAS) cprev = c;
      }
    }
    if ((unsigned)ro < (unsigned)n_out) {
      if (st_full) {
        *reinterpret_cast<float4*>(q) = fin;
      } else if (st_part) {
        if (smask & 1) q[0] = fin.x;
        if (smask & 2) q[1] = fin.y;
        if (smask & 4) q[2] = fin.z;
        if (smask & 8) q[3] = fin.w;
      }
      if (EXTRAS && smask != 0) {
        // rows outside the N/S bands have nothing to forward once W/E is batched
        if (FWD && (oc.we_perrow || r_lo + ro < a.fw || r_lo + ro >= g.rows - a.fw)) {
          Vals<4> v; v.v[0] = fin.x; v.v[1] = fin.y; v.v[2] = fin.z; v.v[3] = fin.w;
          forward_row<4>(a, r_lo + ro, j0, v, smask, oc.we_skip);
        }
        if (a.norm != H2D_NORM_NONE) {
          const float d0 = (smask & 1) ? fin.x - cprev.x : 0.f;
          const float d1 = (smask & 2) ? fin.y - cprev.y : 0.f;
          const float d2 = (smask & 4) ? fin.z - cprev.z : 0.f;
          const float d3 = (smask & 8) ? fin.w - cprev.w : 0.f;
          if (a.norm == H2D_NORM_L2SQ) {
            acc = fmaf(d0, d0, acc); acc = fmaf(d1, d1, acc);
            acc = fmaf(d2, d2, acc); acc = fmaf(d3, d3, acc);
          } else {
            acc = fmax_nan(acc, fmax_nan(fmax_nan(fabsf(d0), fabsf(d1)), fmax_nan(fabsf(d2), fabsf(d3))));
          }
        }
      }
    }
    if (FWD && oc.we_batched && (unsigned)ro < (unsigned)n_out) {
      if (oc.we_count == 0) oc.we_rho0 = r_lo + ro;
      if (oc.seg[0].we_off >= 0) {
        const uint32_t slot = we_stage + (uint32_t)(oc.seg[0].we_off + oc.we_count * (TK_WE_COLS * 4));
        asm volatile("st.shared.v4.f32 [%0], {%1, %2, %3, %4};" ::"r"(slot), "f"(fin.x), "f"(fin.y),
                     "f"(fin.z), "f"(fin.w) : "memory");
      }
      ++oc.we_count;
      if (oc.we_count == TK_WE_ROWS || ro == n_out - 1) {
        flush_we(a, we_stage, oc.we_rho0, oc.we_count, oc.we_sides, lane);
        oc.we_count = 0;
      }
    }
    q += g.pitch;
    ++ro;
  };

  auto boxes = [&](auto Ftag, int b_lo, int b_hi) {
#pragma unroll 1
    for (int b = b_lo; b < b_hi; ++b) {
      const int s = (box0 + b) % TK_STAGES;
      mbar_wait(bars + s * 8, (uint32_t)(((box0 + b) / TK_STAGES) & 1));
      const uint32_t sbase = ring + s * STAGE_BYTES + lane * 16;
#pragma unroll 1
      for (int rr = 0; rr < TK_BR; rr += 3) {
        st[0][2] = lds128(sbase + rr * (W * 4));
        row_step(std::integral_constant<int, 0>{}, Ftag);
        st[0][0] = lds128(sbase + (rr + 1) * (W * 4));
        row_step(std::integral_constant<int, 1>{}, Ftag);
        st[0][1] = lds128(sbase + (rr + 2) * (W * 4));
        row_step(std::integral_constant<int, 2>{}, Ftag);
      }
      // Everyone has read stage s: refill it with box b + STAGES.
      __syncwarp();
      if (lane == 0 && b + TK_STAGES < n_blk) issue(b + TK_STAGES);
    }
  };

  if (COMM && comm_edge) {
    // Only the boxes whose output rows lie in the N / S band (or every box, when column
    // bands are forwarded) run the forwarding body; the stretch between them is plain.
    int b1 = n_blk, b2 = n_blk;
    if (!oc.we_batched && !oc.we_perrow) {
      b1 = r_lo < a.fw ? min(n_blk, (a.fw - r_lo + LAG + TK_BR - 1) / TK_BR) : 0;
      b2 = r_hi > g.rows - a.fw ? max(b1, min(n_blk, (g.rows - a.fw - r_lo + LAG) / TK_BR)) : n_blk;
    }
    boxes(std::true_type{}, 0, b1);
    boxes(std::false_type{}, b1, b2);
    boxes(std::true_type{}, b2, n_blk);
  } else {
    boxes(std::false_type{}, 0, n_blk);
  }
  track = trk;
  return acc;
}

// ============================================================================
//               task body, fp32, two half-strips packed in f32x2
// ============================================================================
__device__ __forceinline__ float2 shfl2(float2 v, int src) {
  return make_float2(__shfl_sync(0xffffffffu, v.x, src), __shfl_sync(0xffffffffu, v.y, src));
}

template <int K, int LAY, bool IMM, bool FROZEN, int XM>
__device__ __forceinline__ float run_task_ab(const CUtensorMap* tmap, const TkArgs& a, float* dst_base,
                                             int strip, int chunk, bool comm_edge, uint32_t ring,
                                             uint32_t bars, uint32_t we_stage, int lane) {
  constexpr bool EXTRAS = XM != XM_NONE;
  constexpr bool COMM = XM == XM_COMM;
  constexpr int NC = Lay<LAY>::NC, W = Lay<LAY>::W, HALF = Lay<LAY>::HALF;
  constexpr int STAGE_BYTES = TK_BR * W * 4;
  const Geo& g = a.g;
  const int R = a.chunk_rows;
  const int r_lo = a.row_base + chunk * R;
  const int r_hi = min(r_lo + R, a.row_end);
  const int n_out = r_hi - r_lo;
  const int own_lo = strip * a.wv;
  const int vbase = (strip == a.n_strips - 1 && g.cols >= a.wv) ? ((g.cols - a.wv + 3) & ~3) : own_lo;
  const int jbase = vbase - a.H;
  const int jA = jbase + lane * NC, jB = jA + HALF;
  const int row0 = r_lo - K;
  constexpr int LAG = 2 * K;
  const int n_in = n_out + LAG;
  const int n_blk = (n_in + TK_BR - 1) / TK_BR;
  const int tc0 = g.halo_c + jbase;
  const int tc1 = g.halo_r + row0;

  // valid window of the strip = [H, W - H): only the outer side of each half is cut
  const bool validA = lane * NC >= a.H;
  const bool validB = HALF + lane * NC < W - a.H;
  OutCtx<2> oc;
  oc.seg[0].j0 = jA; oc.seg[1].j0 = jB;
  oc.seg[0].smask = 0; oc.seg[1].smask = 0;
#pragma unroll
  for (int e = 0; e < NC; ++e) {
    if (validA && (jA + e) >= own_lo && (jA + e) < g.cols) oc.seg[0].smask |= 1 << e;
    if (validB && (jB + e) >= own_lo && (jB + e) < g.cols) oc.seg[1].smask |= 1 << e;
  }
  {
    const bool sv[2] = {validA, validB};
    setup_out<2, NC, COMM>(oc, a, comm_edge, strip, sv);
  }
  constexpr int FULL = (1 << NC) - 1;

  bool actA[NC], actB[NC];
#pragma unroll
  for (int e = 0; e < NC; ++e) {
    actA[e] = actB[e] = true;
    if (FROZEN) {
      const int ya = g.y0 + jA + e, yb = g.y0 + jB + e;
      actA[e] = ya >= g.ay_lo && ya <= g.ay_hi;
      actB[e] = yb >= g.ay_lo && yb <= g.ay_hi;
    }
  }

  auto issue = [&](int b) {
    const int s = b % TK_STAGES;
    const uint32_t bar = bars + s * 8;
    mbar_expect_tx(bar, STAGE_BYTES);
    tma_load_2d(ring + s * STAGE_BYTES, tmap, tc0, tc1 + b * TK_BR, bar);
  };
  if (lane == 0) {
    const int pre = n_blk < TK_STAGES ? n_blk : TK_STAGES;
    for (int b = 0; b < pre; ++b) issue(b);
  }

  // st[level][slot][j] = (A_j, B_j)
  float2 st[K][3][NC];
#pragma unroll
  for (int t = 0; t < K; ++t)
#pragma unroll
    for (int s = 0; s < 3; ++s)
#pragma unroll
      for (int j = 0; j < NC; ++j) st[t][s][j] = make_float2(0.f, 0.f);

  const float2 a0 = IMM ? make_float2(kRefA0, kRefA0) : make_float2(g.a0, g.a0);
  const float2 cx = IMM ? make_float2(kRefCx, kRefCx) : make_float2(g.cx, g.cx);
  const float2 cy = IMM ? make_float2(kRefCy, kRefCy) : make_float2(g.cy, g.cy);
  const int lane_l = (lane + 31) & 31, lane_r = (lane + 1) & 31;
  const bool first = lane == 0, last = lane == 31;

  float acc = 0.f;
  int ro = -LAG;
  float* q = dst_base + (int64_t)(g.halo_r + r_lo - LAG) * g.pitch + g.halo_c + jA;

  auto load_row = [&](uint32_t addr, float2 (&dst)[NC]) {
    if constexpr (NC == 4) {
      const float4 va = lds128(addr), vb = lds128(addr + HALF * 4);
      dst[0] = make_float2(va.x, vb.x); dst[1] = make_float2(va.y, vb.y);
      dst[2] = make_float2(va.z, vb.z); dst[3] = make_float2(va.w, vb.w);
    } else {
      const float2 va = lds64(addr), vb = lds64(addr + HALF * 4);
      dst[0] = make_float2(va.x, vb.x); dst[1] = make_float2(va.y, vb.y);
    }
  };

  auto store_seg = [&](float* p, const float (&v)[NC], int smask) {
    if (smask == FULL) {
      if constexpr (NC == 4) *reinterpret_cast<float4*>(p) = make_float4(v[0], v[1], v[2], v[3]);
      else *reinterpret_cast<float2*>(p) = make_float2(v[0], v[1]);
    } else if (smask != 0) {
#pragma unroll
      for (int e = 0; e < NC; ++e)
        if ((smask >> e) & 1) p[e] = v[e];
    }
  };

  auto row_step = [&](auto Ptag, auto Ftag) {
    constexpr int PH = decltype(Ptag)::value;
    constexpr bool FWD = COMM && decltype(Ftag)::value;
    constexpr int iu = PH, ic = (PH + 1) % 3, id = (PH + 2) % 3;
    float2 fin[NC], cprev[NC];
#pragma unroll
    for (int t = 0; t < K; ++t) {
      // neighbours across lanes; the wrap closes the seam between the halves:
      // lane 0's left neighbour in B is lane 31's last column of A, lane 31's right
      // neighbour in A is lane 0's first column of B (the other two wrapped values fall
      // into the strip's overlap zone, whose results are never stored).
      // (the select sits on the SENDING side, so that each shuffle lands directly in one half
      // of an aligned register pair: receiver-side selects cost a register move per pair)
      const float2 cl = st[t][ic][NC - 1], cr = st[t][ic][0];
      const float2 Lp = shfl2(make_float2(cl.x, last ? cl.x : cl.y), lane_l);
      const float2 Rp = shfl2(make_float2(first ? cr.y : cr.x, cr.y), lane_r);
      bool ract = true;
      if (FROZEN) {
        const int gx = g.x0 + r_lo + ro + (K - 1 - t);
        ract = gx >= g.ax_lo && gx <= g.ax_hi;
      }
#pragma unroll
      for (int j = 0; j < NC; ++j) {
        const float2 c = st[t][ic][j];
        const float2 left = j == 0 ? Lp : st[t][ic][(j + NC - 1) % NC];
        const float2 right = j == NC - 1 ? Rp : st[t][ic][(j + 1) % NC];
        const float2 ns = __fadd2_rn(right, left);
        const float2 gg = __ffma2_rn(cy, ns, __fmul2_rn(a0, c));
        const float2 ew = __fadd2_rn(st[t][id][j], st[t][iu][j]);
        float2 o = __ffma2_rn(cx, ew, gg);
        if (FROZEN) {
          o.x = (ract && actA[j]) ? o.x : c.x;
          o.y = (ract && actB[j]) ? o.y : c.y;
        }
        if (t + 1 < K) {
          st[t + 1][id][j] = o;
        } else {
          fin[j] = o;
          if (EXTRAS) cprev[j] = c;
        }
      }
    }
    if ((unsigned)ro < (unsigned)n_out) {
      float fa[NC], fb[NC];
#pragma unroll
      for (int j = 0; j < NC; ++j) { fa[j] = fin[j].x; fb[j] = fin[j].y; }
      store_seg(q, fa, oc.seg[0].smask);
      store_seg(q + HALF, fb, oc.seg[1].smask);
      if (EXTRAS) {
        if (FWD && (oc.we_perrow || r_lo + ro < a.fw || r_lo + ro >= g.rows - a.fw)) {
          Vals<NC> va, vb;
#pragma unroll
          for (int j = 0; j < NC; ++j) { va.v[j] = fa[j]; vb.v[j] = fb[j]; }
          if (oc.seg[0].smask) forward_row<NC>(a, r_lo + ro, jA, va, oc.seg[0].smask, oc.we_skip);
          if (oc.seg[1].smask) forward_row<NC>(a, r_lo + ro, jB, vb, oc.seg[1].smask, oc.we_skip);
        }
        if (a.norm != H2D_NORM_NONE) {
#pragma unroll
          for (int j = 0; j < NC; ++j) {
            const float da = ((oc.seg[0].smask >> j) & 1) ? fin[j].x - cprev[j].x : 0.f;
            const float db = ((oc.seg[1].smask >> j) & 1) ? fin[j].y - cprev[j].y : 0.f;
            if (a.norm == H2D_NORM_L2SQ) {
              acc = fmaf(da, da, acc); acc = fmaf(db, db, acc);
            } else {
              acc = fmax_nan(acc, fmax_nan(fabsf(da), fabsf(db)));
            }
          }
        }
      }
      if (FWD && oc.we_batched) {
        if (oc.we_count == 0) oc.we_rho0 = r_lo + ro;
#pragma unroll
        for (int s = 0; s < 2; ++s) {
          if (oc.seg[s].we_off < 0) continue;
          const uint32_t slot = we_stage + (uint32_t)(oc.seg[s].we_off + oc.we_count * (TK_WE_COLS * 4));
          const float* v = s == 0 ? fa : fb;
          if constexpr (NC == 4)
            asm volatile("st.shared.v4.f32 [%0], {%1, %2, %3, %4};" ::"r"(slot), "f"(v[0]), "f"(v[1]),
                         "f"(v[2]), "f"(v[3]) : "memory");
          else
            asm volatile("st.shared.v2.f32 [%0], {%1, %2};" ::"r"(slot), "f"(v[0]), "f"(v[1]) : "memory");
        }
        ++oc.we_count;
        if (oc.we_count == TK_WE_ROWS || ro == n_out - 1) {
          flush_we(a, we_stage, oc.we_rho0, oc.we_count, oc.we_sides, lane);
          oc.we_count = 0;
        }
      }
    }
    q += g.pitch;
    ++ro;
  };

  auto boxes = [&](auto Ftag, int b_lo, int b_hi) {
#pragma unroll 1
    for (int b = b_lo; b < b_hi; ++b) {
      const int s = b % TK_STAGES;
      mbar_wait(bars + s * 8, (uint32_t)((b / TK_STAGES) & 1));
      const uint32_t sbase = ring + s * STAGE_BYTES + lane * (NC * 4);
#pragma unroll 1
      for (int rr = 0; rr < TK_BR; rr += 3) {
        load_row(sbase + rr * (W * 4), st[0][2]);
        row_step(std::integral_constant<int, 0>{}, Ftag);
        load_row(sbase + (rr + 1) * (W * 4), st[0][0]);
        row_step(std::integral_constant<int, 1>{}, Ftag);
        load_row(sbase + (rr + 2) * (W * 4), st[0][1]);
        row_step(std::integral_constant<int, 2>{}, Ftag);
      }
      __syncwarp();
      if (lane == 0 && b + TK_STAGES < n_blk) issue(b + TK_STAGES);
    }
  };

  if (COMM && comm_edge) {
    int b1 = n_blk, b2 = n_blk;
    if (!oc.we_batched && !oc.we_perrow) {
      b1 = r_lo < a.fw ? min(n_blk, (a.fw - r_lo + LAG + TK_BR - 1) / TK_BR) : 0;
      b2 = r_hi > g.rows - a.fw ? max(b1, min(n_blk, (g.rows - a.fw - r_lo + LAG) / TK_BR)) : n_blk;
    }
    boxes(std::true_type{}, 0, b1);
    boxes(std::false_type{}, b1, b2);
    boxes(std::true_type{}, b2, n_blk);
  } else {
    boxes(std::false_type{}, 0, n_blk);
  }
  return acc;
}

// One entry point for both families.  `track` is meaningful for strict modes only.
template <int K, int ARITH, int LAY, bool IMM, bool FROZEN, int XM, bool FAST>
__device__ __forceinline__ float run_task(const CUtensorMap* tmap, const TkArgs& a, float* dst_base,
                                          int strip, int chunk, bool comm_edge, uint32_t ring,
                                          uint32_t bars, uint32_t we_stage, int lane, uint32_t& track) {
  if constexpr (LAY == LAY_V4) {
    return run_task_v4<K, ARITH, IMM, FROZEN, XM, FAST>(tmap, a, dst_base, strip, chunk, comm_edge, ring, bars,
                                                        we_stage, lane, track);
  } else {
    static_assert(ARITH == H2D_ARITH_FP32 || LAY == LAY_V4, "packed layouts are fp32 only");
    track = 0u;
    return run_task_ab<K, LAY, IMM, FROZEN, XM>(tmap, a, dst_base, strip, chunk, comm_edge, ring, bars,
                                                we_stage, lane);
  }
}

}  // namespace h2d
