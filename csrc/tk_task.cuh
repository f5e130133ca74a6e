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
//  * run_task_v8 (LAY_V8): 256-column strips, each lane owns 8 adjacent columns; sweeps of
//    at most 4 levels (the register state of 8 levels x 4 columns).  For the strict modes at
//    k <= 4: 8 of 256 columns recomputed instead of 8 of 128, half the shuffles per point.
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
constexpr int TK_STAGES = 4;     // boxes in flight per warp (see tk_stages)
constexpr int TK_WARPS = 4;      // warps (tasks) per CTA
constexpr int TK_WE_COLS = 32;   // columns forwarded per row and side: one full 128-byte line
constexpr int TK_CW = 8;         // floats per row of a packed column-ghost buffer (= TK_KMAX)
constexpr int TK_KMAX = 8;

enum { LAY_V4 = 0, LAY_AB2 = 1, LAY_AB4 = 2, LAY_V8 = 3, LAY_COUNT = 4 };
template <int LAY> struct Lay;
template <> struct Lay<LAY_V4>  { static constexpr int W = 128, NC = 4, HALF = 0; };
template <> struct Lay<LAY_AB2> { static constexpr int W = 128, NC = 2, HALF = 64; };
template <> struct Lay<LAY_AB4> { static constexpr int W = 256, NC = 4, HALF = 128; };
template <> struct Lay<LAY_V8>  { static constexpr int W = 256, NC = 8, HALF = 0; };   // K <= 4 only

__host__ __device__ constexpr int tk_strip_width(int lay) { return (lay == LAY_AB4 || lay == LAY_V8) ? 256 : 128; }
__host__ __device__ constexpr int tk_stage_bytes(int lay) { return TK_BR * tk_strip_width(lay) * 4; }
// boxes in flight per warp: V8's boxes are twice as wide and its sweeps at most 4 levels deep,
// two of them keep the ring at the 12 KB per warp of the other 128-column layouts
__host__ __device__ constexpr int tk_stages(int lay) { return lay == LAY_V8 ? 2 : TK_STAGES; }
constexpr int TK_V8_KMAX = 4;
__host__ __device__ constexpr int tk_smem_ring(int lay) { return TK_WARPS * tk_stages(lay) * tk_stage_bytes(lay); }
constexpr int TK_SMEM_BARS = TK_WARPS * TK_STAGES * 8;
__host__ __device__ constexpr int tk_smem_bytes(int lay, bool /*xtr*/ = false) {
  return tk_smem_ring(lay) + TK_SMEM_BARS;
}

struct PeerDev {
  float* buf;       // neighbour's padded array that holds the SAME parity as our dst
  float* colbuf;    // W / E only: the neighbour's packed column-ghost buffer of that parity (or null)
  uint32_t* flag;   // where we publish our epoch for that neighbour
  int32_t x0, y0, halo_r, halo_c;
  int64_t pitch;
};

struct TkArgs {
  Geo g;
  float* dst;
  // Packed column ghosts (tiles with W / E neighbours).  A column band written in place is one
  // short piece per 64-KiB-pitched row of the neighbour's array, and that row-strided remote
  // write stream costs 25-50 % of a launch on long tiles; so the band goes, as ONE contiguous
  // stream, into a packed buffer of the neighbour (TK_CW floats per tile row), and the
  // neighbour's edge tasks copy the rows they are about to read into the margins of their
  // source array first.  colbuf_in[side]: MY packed buffers of the source parity (0 = W
  // ghosts, 1 = E ghosts; null = ghosts arrive in place).  src: the source array (its margins
  // are written by the unpack; the cells are read through the tensor map).
  float* src;
  const float* colbuf_in[2];
  const int2* tasks;     // .x = strip | (comm_edge << 31) | (short chunks << 30), .y = chunk
  int32_t n_tasks, n_chunks, chunk_rows, n_comm_edge;
  int32_t n_tasks_all;   // tasks of the whole logical launch (this kernel may run a part of them)
  // Tasks of the strips that forward column bands (bit 30 of tasks[].x, or the edge strips of
  // a row window) use shorter chunks: chunk c of such a strip starts at row_base + c * chunk_rows_edge
  int32_t chunk_rows_edge, n_chunks_edge;
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
                   // bit 3: row bands go through the bulk-copy engine (H2D_FWD_BULK, default 1)
                   // bit 4: column bands as whole 128-byte lines (H2D_FWD_LINES=1)
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

// ------------------------------------------------- halo forwarding (post-pass)
// A task that touches the tile edge first computes and stores its chunk like any other
// task; THEN the warp copies the parts of what it has just stored that lie in a
// neighbour's ghost margin from its own array into the neighbour's, over NVLink:
//   N / S (and the corner directions): the rows within `fw` of the tile edge, the columns
//       this task owns;
//   W / E: every row of the chunk, the edge columns — as whole 128-byte lines (the 32
//       outermost columns, exactly the neighbour's margin) whenever the geometry allows,
//       because partial-line remote writes cost the destination a read-modify-write.
// Keeping the copies out of the row loop matters more than their own cost: with the
// forwarding inside the loop (round 1) an edge task carried twice the instructions, a
// stack frame and spills, ran several times longer than an interior task and set the
// length of the whole launch (ncu: 21 % of the warp samples on 5 % of the instructions).
// The reference packs halos with MPI datatypes and sends them with persistent requests
// (grad1612_mpi_heat.c:140-144,210-227).
// A remote store instruction completes one NVLink round trip at a time for the warp that
// issued it (measured on B200: ~2.9 us per instruction, whatever its size; shared -> global
// bulk stores of the copy engine serialise the same way, ~2.3 us per 128-byte row, and were
// slower).  So a store moves as much as a warp can — 512 bytes, 4 rows of a column band —
// and the tasks that forward column bands are kept SHORT (TkArgs::chunk_rows_edge): many
// short edge tasks copy side by side in the first wave of a launch, instead of a few long
// ones serialising hundreds of round trips at the end of theirs.
__device__ __forceinline__ void copy_block(const float* src, int64_t spitch, float* dst, int64_t dpitch,
                                           int nrows, int ncols, int lane) {
  if (nrows <= 0 || ncols <= 0) return;
  const bool vec = ((((uintptr_t)src) | ((uintptr_t)dst)) & 15) == 0 && ((spitch | dpitch | ncols) & 3) == 0;
  if (vec) {
    const int q = ncols >> 2, total = nrows * q;
#pragma unroll 4
    for (int i = lane; i < total; i += 32) {
      const int r = i / q, c = (i - r * q) << 2;
      const float4 v = __ldcg(reinterpret_cast<const float4*>(src + (int64_t)r * spitch + c));
      *reinterpret_cast<float4*>(dst + (int64_t)r * dpitch + c) = v;
    }
  } else {
    const int total = nrows * ncols;
    for (int i = lane; i < total; i += 32) {
      const int r = i / ncols, c = i - r * ncols;
      dst[(int64_t)r * dpitch + c] = __ldcg(src + (int64_t)r * spitch + c);
    }
  }
}

// The same copy through the bulk-copy engine: the warp gathers the rows into its (now idle)
// TMA ring in shared memory and one lane issues one asynchronous shared -> global bulk store
// per row (cp.async.bulk with plain addresses: no tensor map, no multicast, and only STORES
// ever target a neighbour's memory).  The engine also takes ~2.3 us per remote row, one after
// the other, but it does so on its own: worth it for long rows (N / S bands), hopeless for
// 32-byte column-band rows.
__device__ __forceinline__ void copy_block_bulk(const float* src, int64_t spitch, float* dst, int64_t dpitch,
                                                int nrows, int ncols, int lane, uint32_t stage, int stage_bytes) {
  const int q = ncols >> 2;                       // float4 pieces per row
  const int rows_per_round = stage_bytes / (ncols * 4);
  for (int r0 = 0; r0 < nrows; r0 += rows_per_round) {
    const int n = min(rows_per_round, nrows - r0);
    // the previous round's bulk stores must have finished reading the staging area
    if (lane == 0) asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory");
    __syncwarp();
    const int total = n * q;
#pragma unroll 4
    for (int i = lane; i < total; i += 32) {
      const int r = i / q, c = (i - r * q) << 2;
      const float4 v = __ldcg(reinterpret_cast<const float4*>(src + (int64_t)(r0 + r) * spitch + c));
      asm volatile("st.shared.v4.f32 [%0], {%1, %2, %3, %4};" ::"r"(stage + (uint32_t)(i * 16)), "f"(v.x),
                   "f"(v.y), "f"(v.z), "f"(v.w) : "memory");
    }
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");   // generic-proxy writes -> async-proxy reads
    __syncwarp();
    if (lane == 0) {
      for (int r = 0; r < n; ++r)
        asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(dst + (int64_t)(r0 + r) * dpitch),
                     "r"(stage + (uint32_t)(r * ncols * 4)), "r"((uint32_t)(ncols * 4)) : "memory");
      asm volatile("cp.async.bulk.commit_group;" ::: "memory");
    }
  }
}

// Columns [c_lo, c_hi) are the ones the task STORED (the shifted last strip stores its whole
// valid window), [own_lo, c_hi) the ones it owns.
static __device__ __noinline__ void forward_task(const TkArgs& a, int strip, int r_lo, int r_hi, int lane,
                                                 uint32_t stage, int stage_bytes) {
  const Geo& g = a.g;
  const int fw = a.fw;
  const int own_lo = strip * a.wv;
  const bool first = strip == 0, last = strip == a.n_strips - 1;
  const int c_hi = last ? g.cols : own_lo + a.wv;
  // whole-line column bands (debug bit 4): 4x the store instructions of exact-width bands
  const bool lines = (a.debug & 16) && (g.cols & 3) == 0 && g.cols >= 2 * TK_WE_COLS && fw <= TK_WE_COLS;
  bool used_bulk = false;
#pragma unroll 1
  for (int d = 0; d < H2D_NDIR; ++d) {
    const PeerDev& p = a.peer[d];
    if (p.buf == nullptr) continue;
    const int sx = (d == H2D_N || d == H2D_NW || d == H2D_NE) ? -1
                 : (d == H2D_S || d == H2D_SW || d == H2D_SE) ? 1 : 0;
    const int sy = (d == H2D_W || d == H2D_NW || d == H2D_SW) ? -1
                 : (d == H2D_E || d == H2D_NE || d == H2D_SE) ? 1 : 0;
    int ra = r_lo, rb = r_hi;
    if (sx < 0) rb = min(rb, fw);
    if (sx > 0) ra = max(ra, g.rows - fw);
    int ca, cb;
    if (sy == 0) {
      ca = own_lo; cb = c_hi;
    } else {
      // the W band belongs to the first strip, the E band to the last one (which has stored
      // all of it); pure W / E bands travel as whole lines where possible
      if (sy < 0 ? !first : !last) continue;
      const int bw = (sx == 0 && lines && p.halo_c >= TK_WE_COLS) ? TK_WE_COLS : fw;
      ca = sy < 0 ? 0 : g.cols - bw;
      cb = sy < 0 ? bw : g.cols;
    }
    if (rb <= ra || cb <= ca) continue;
    const float* src = a.dst + (int64_t)(g.halo_r + ra) * g.pitch + g.halo_c + ca;
    const int nr = rb - ra, ncl = cb - ca;
    if (sx == 0 && sy != 0 && p.colbuf != nullptr && ncl <= TK_CW) {
      // packed: row r of the neighbour's tile (same row block) <- my fw edge columns
      copy_block(src, g.pitch, p.colbuf + (int64_t)(g.x0 + ra - p.x0) * TK_CW, TK_CW, nr, ncl, lane);
      continue;
    }
    float* dst = p.buf + (int64_t)(g.x0 + ra - p.x0 + p.halo_r) * p.pitch + (g.y0 + ca - p.y0 + p.halo_c);
    const bool vec = ((((uintptr_t)src) | ((uintptr_t)dst)) & 15) == 0 && ((g.pitch | p.pitch | ncl) & 3) == 0;
    if ((a.debug & 8) && vec && ncl >= 64 && ncl * 4 <= stage_bytes) {
      copy_block_bulk(src, g.pitch, dst, p.pitch, nr, ncl, lane, stage, stage_bytes);
      used_bulk = true;
    } else {
      copy_block(src, g.pitch, dst, p.pitch, nr, ncl, lane);
    }
  }
  if (used_bulk) {
    // all bulk stores of this task have been performed before the task is counted in
    if (lane == 0) {
      asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");
      asm volatile("fence.proxy.async;" ::: "memory");
    }
    __syncwarp();
  }
}

// Receiver side of the packed column ghosts: copy rows [ra, rb) of my W / E packed buffers
// into the ghost columns of the source array, then make the writes visible to the TMA
// (async proxy) reads that follow.  Adjacent chunks unpack overlapping rows: same values.
static __device__ __noinline__ void unpack_colbuf(const TkArgs& a, bool west, bool east, int ra, int rb, int lane) {
  const Geo& g = a.g;
  ra = max(ra, 0);
  rb = min(rb, g.rows);
  if (rb > ra) {
    if (west && a.colbuf_in[0] != nullptr)
      copy_block(a.colbuf_in[0] + (int64_t)ra * TK_CW, TK_CW,
                 a.src + (int64_t)(g.halo_r + ra) * g.pitch + g.halo_c - a.fw, g.pitch, rb - ra, a.fw, lane);
    if (east && a.colbuf_in[1] != nullptr)
      copy_block(a.colbuf_in[1] + (int64_t)ra * TK_CW, TK_CW,
                 a.src + (int64_t)(g.halo_r + ra) * g.pitch + g.halo_c + g.cols, g.pitch, rb - ra, a.fw, lane);
  }
  asm volatile("fence.proxy.async;" ::: "memory");
  __syncwarp();
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
  // (c * 2^896 by DMUL; an F2F.F64.F32 of c instead was measured slower: 1253 vs 1318 GLUP/s)
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
// or outside the grid) -> per-element select.  NORM: the last level also
// accumulates (new - old) of the cells the task OWNS (a.norm says which norm).
// FAST (strict modes): see above; `track` receives the running maximum of the
// bit patterns.  box0: boxes this warp's ring has already carried in this
// kernel (a task that is run a second time continues the stage / phase sequence
// of its mbarriers instead of re-initialising them).
template <int K, int ARITH, bool IMM, bool FROZEN, bool NORM, bool FAST>
__device__ __forceinline__ float run_task_v4(const CUtensorMap* tmap, const TkArgs& a, float* dst_base,
                                             int strip, int r_lo, int r_hi, uint32_t ring, uint32_t bars, int lane,
                                             uint32_t& track, int box0 = 0) {
  constexpr bool TRACK = ARITH != H2D_ARITH_FP32;
  constexpr int W = Lay<LAY_V4>::W;
  constexpr int STAGE_BYTES = TK_BR * W * 4;
  const Geo& g = a.g;
  const int n_out = r_hi - r_lo;
  // The last strip is shifted left so that its valid window ends at the tile edge: its
  // first columns are then computed twice (by it and by its left neighbour).  It STORES its
  // whole valid window (the doubly computed cells get the same bits from both warps) but
  // OWNS — reduces, forwards as a row band — only [own_lo, cols).  No strip is narrow and,
  // what matters for column halos, the last 32 columns always come from ONE warp, so the
  // east band can leave as whole 128-byte lines.
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

  // Which of my 4 columns are stored / owned: inside the strip's valid window and the tile.
  const bool lane_in = (lane * 4 >= a.H) && (lane * 4 < W - a.H);
  int smask = 0, nmask = 0;
#pragma unroll
  for (int e = 0; e < 4; ++e) {
    if (lane_in && (j0 + e) >= vbase && (j0 + e) < g.cols) smask |= 1 << e;
    if (lane_in && (j0 + e) >= own_lo && (j0 + e) < g.cols) nmask |= 1 << e;
  }
  const bool st_full = smask == 15;
  const bool st_part = smask != 0 && smask != 15;

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
  // (A "skewed" variant in which level t consumes what level t-1 produced in the
  // previous row step — K independent levels per step instead of a K-long chain —
  // was measured 11 % SLOWER on B200: the chain is not the limiter.)
  auto row_step = [&](auto Ptag) {
    constexpr int PH = decltype(Ptag)::value;
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
        if (NORM) cprev = c;
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
      if (NORM && nmask != 0) {
        const float d0 = (nmask & 1) ? fin.x - cprev.x : 0.f;
        const float d1 = (nmask & 2) ? fin.y - cprev.y : 0.f;
        const float d2 = (nmask & 4) ? fin.z - cprev.z : 0.f;
        const float d3 = (nmask & 8) ? fin.w - cprev.w : 0.f;
        if (a.norm == H2D_NORM_L2SQ) {
          acc = fmaf(d0, d0, acc); acc = fmaf(d1, d1, acc);
          acc = fmaf(d2, d2, acc); acc = fmaf(d3, d3, acc);
        } else {
          acc = fmax_nan(acc, fmax_nan(fmax_nan(fabsf(d0), fabsf(d1)), fmax_nan(fabsf(d2), fabsf(d3))));
        }
      }
    }
    q += g.pitch;
    ++ro;
  };

#pragma unroll 1
  for (int b = 0; b < n_blk; ++b) {
    const int s = (box0 + b) % TK_STAGES;
    mbar_wait(bars + s * 8, (uint32_t)(((box0 + b) / TK_STAGES) & 1));
    const uint32_t sbase = ring + s * STAGE_BYTES + lane * 16;
#pragma unroll 1
    for (int rr = 0; rr < TK_BR; rr += 3) {
      st[0][2] = lds128(sbase + rr * (W * 4));
      row_step(std::integral_constant<int, 0>{});
      st[0][0] = lds128(sbase + (rr + 1) * (W * 4));
      row_step(std::integral_constant<int, 1>{});
      st[0][1] = lds128(sbase + (rr + 2) * (W * 4));
      row_step(std::integral_constant<int, 2>{});
    }
    // Everyone has read stage s: refill it with box b + STAGES.
    __syncwarp();
    if (lane == 0 && b + TK_STAGES < n_blk) issue(b + TK_STAGES);
  }
  track = trk;
  return acc;
}

// ============================================================================
//                  task body, 8 adjacent columns per lane (K <= 4)
// ============================================================================
// The same pipeline as run_task_v4 on 256-column strips: a lane owns columns 8 l .. 8 l + 7 as
// two float4 (lo, hi).  A sweep of K <= 4 levels needs a column margin of 4, so a strip stores
// 248 of its 256 columns (120 of 128 in run_task_v4) and a lane shuffles twice per 8 points
// instead of twice per 4; the register state — K levels x 3 rows x 8 columns — is what
// run_task_v4 holds at K = 8.  Two 6-row boxes of 6 KB in flight per warp (tk_stages).
struct Row8 { float4 lo, hi; };

template <int K, int ARITH, bool IMM, bool FROZEN, bool NORM, bool FAST>
__device__ __forceinline__ float run_task_v8(const CUtensorMap* tmap, const TkArgs& a, float* dst_base,
                                             int strip, int r_lo, int r_hi, uint32_t ring, uint32_t bars, int lane,
                                             uint32_t& track, int box0 = 0) {
  static_assert(K <= TK_V8_KMAX, "the 8-column layout holds at most 4 levels");
  constexpr bool TRACK = ARITH != H2D_ARITH_FP32;
  constexpr int W = Lay<LAY_V8>::W;
  constexpr int STAGES = tk_stages(LAY_V8);
  constexpr int STAGE_BYTES = TK_BR * W * 4;
  const Geo& g = a.g;
  const int n_out = r_hi - r_lo;
  const int own_lo = strip * a.wv;
  const int vbase = (strip == a.n_strips - 1 && g.cols >= a.wv) ? ((g.cols - a.wv + 3) & ~3) : own_lo;
  const int jbase = vbase - a.H;
  const int j0 = jbase + lane * 8;
  const int row0 = r_lo - K;
  constexpr int LAG = 2 * K;
  const int n_in = n_out + LAG;
  const int n_blk = (n_in + TK_BR - 1) / TK_BR;
  const int tc0 = g.halo_c + jbase;
  const int tc1 = g.halo_r + row0;

  // stored / owned elements of my 8 columns: inside the strip's valid window and the tile
  int smask = 0, nmask = 0;
#pragma unroll
  for (int e = 0; e < 8; ++e) {
    const bool in = (lane * 8 + e >= a.H) && (lane * 8 + e < W - a.H);
    if (in && (j0 + e) >= vbase && (j0 + e) < g.cols) smask |= 1 << e;
    if (in && (j0 + e) >= own_lo && (j0 + e) < g.cols) nmask |= 1 << e;
  }

  bool act[8] = {true, true, true, true, true, true, true, true};
  if (FROZEN) {
#pragma unroll
    for (int e = 0; e < 8; ++e) {
      const int gy = g.y0 + j0 + e;
      act[e] = gy >= g.ay_lo && gy <= g.ay_hi;
    }
  }

  auto issue = [&](int b) {
    const int s = (box0 + b) % STAGES;
    const uint32_t bar = bars + s * 8;
    mbar_expect_tx(bar, STAGE_BYTES);
    tma_load_2d(ring + s * STAGE_BYTES, tmap, tc0, tc1 + b * TK_BR, bar);
  };
  if (lane == 0) {
    const int pre = n_blk < STAGES ? n_blk : STAGES;
    for (int b = 0; b < pre; ++b) issue(b);
  }

  Row8 st[K][3];
#pragma unroll
  for (int t = 0; t < K; ++t)
#pragma unroll
    for (int q3 = 0; q3 < 3; ++q3) st[t][q3].lo = st[t][q3].hi = make_float4(0.f, 0.f, 0.f, 0.f);

  float acc = 0.f;
  uint32_t trk = 0u;
  int ro = -LAG;
  float* q = dst_base + (int64_t)(g.halo_r + r_lo - LAG) * g.pitch + g.halo_c + j0;

  auto row_step = [&](auto Ptag) {
    constexpr int PH = decltype(Ptag)::value;
    constexpr int iu = PH, ic = (PH + 1) % 3, id = (PH + 2) % 3;
    Row8 fin, cprev;
    cprev.lo = cprev.hi = make_float4(0.f, 0.f, 0.f, 0.f);
    if (TRACK) trk = umax4(umax4(trk, st[0][id].lo), st[0][id].hi);
#pragma unroll
    for (int t = 0; t < K; ++t) {
      const Row8 up = st[t][iu];
      const Row8 c = st[t][ic];
      const Row8 down = st[t][id];
      const float L = __shfl_up_sync(0xffffffffu, c.hi.w, 1);
      const float Rn = __shfl_down_sync(0xffffffffu, c.lo.x, 1);
      Row8 o;
      o.lo = point4x<ARITH, IMM, FAST>(up.lo, c.lo, down.lo, L, c.hi.x, g);
      o.hi = point4x<ARITH, IMM, FAST>(up.hi, c.hi, down.hi, c.lo.w, Rn, g);
      if (FROZEN) {
        const int gx = g.x0 + r_lo + ro + (K - 1 - t);
        const bool ract = gx >= g.ax_lo && gx <= g.ax_hi;
        o.lo.x = (ract && act[0]) ? o.lo.x : c.lo.x;
        o.lo.y = (ract && act[1]) ? o.lo.y : c.lo.y;
        o.lo.z = (ract && act[2]) ? o.lo.z : c.lo.z;
        o.lo.w = (ract && act[3]) ? o.lo.w : c.lo.w;
        o.hi.x = (ract && act[4]) ? o.hi.x : c.hi.x;
        o.hi.y = (ract && act[5]) ? o.hi.y : c.hi.y;
        o.hi.z = (ract && act[6]) ? o.hi.z : c.hi.z;
        o.hi.w = (ract && act[7]) ? o.hi.w : c.hi.w;
      }
      if (TRACK) trk = umax4(umax4(trk, o.lo), o.hi);
      if (t + 1 < K) {
        st[t + 1][id] = o;
      } else {
        fin = o;
        if (NORM) cprev = c;
      }
    }
    if ((unsigned)ro < (unsigned)n_out) {
      if ((smask & 0x0f) == 0x0f) {
        *reinterpret_cast<float4*>(q) = fin.lo;
      } else if (smask & 0x0f) {
        if (smask & 1) q[0] = fin.lo.x;
        if (smask & 2) q[1] = fin.lo.y;
        if (smask & 4) q[2] = fin.lo.z;
        if (smask & 8) q[3] = fin.lo.w;
      }
      if ((smask & 0xf0) == 0xf0) {
        *reinterpret_cast<float4*>(q + 4) = fin.hi;
      } else if (smask & 0xf0) {
        if (smask & 16) q[4] = fin.hi.x;
        if (smask & 32) q[5] = fin.hi.y;
        if (smask & 64) q[6] = fin.hi.z;
        if (smask & 128) q[7] = fin.hi.w;
      }
      if (NORM && nmask != 0) {
        const float d[8] = {fin.lo.x - cprev.lo.x, fin.lo.y - cprev.lo.y, fin.lo.z - cprev.lo.z, fin.lo.w - cprev.lo.w,
                            fin.hi.x - cprev.hi.x, fin.hi.y - cprev.hi.y, fin.hi.z - cprev.hi.z, fin.hi.w - cprev.hi.w};
#pragma unroll
        for (int e = 0; e < 8; ++e) {
          const float de = ((nmask >> e) & 1) ? d[e] : 0.f;
          // the same order of accumulation as a lane of run_task_v4 would not give the same
          // partial sums anyway (other lanes own these columns there); the norm is reduced
          // deterministically for a given layout, as everywhere
          if (a.norm == H2D_NORM_L2SQ) acc = fmaf(de, de, acc);
          else acc = fmax_nan(acc, fabsf(de));
        }
      }
    }
    q += g.pitch;
    ++ro;
  };

#pragma unroll 1
  for (int b = 0; b < n_blk; ++b) {
    const int s = (box0 + b) % STAGES;
    mbar_wait(bars + s * 8, (uint32_t)(((box0 + b) / STAGES) & 1));
    const uint32_t sbase = ring + s * STAGE_BYTES + lane * 32;
#pragma unroll 1
    for (int rr = 0; rr < TK_BR; rr += 3) {
      st[0][2].lo = lds128(sbase + rr * (W * 4));
      st[0][2].hi = lds128(sbase + rr * (W * 4) + 16);
      row_step(std::integral_constant<int, 0>{});
      st[0][0].lo = lds128(sbase + (rr + 1) * (W * 4));
      st[0][0].hi = lds128(sbase + (rr + 1) * (W * 4) + 16);
      row_step(std::integral_constant<int, 1>{});
      st[0][1].lo = lds128(sbase + (rr + 2) * (W * 4));
      st[0][1].hi = lds128(sbase + (rr + 2) * (W * 4) + 16);
      row_step(std::integral_constant<int, 2>{});
    }
    __syncwarp();
    if (lane == 0 && b + STAGES < n_blk) issue(b + STAGES);
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

template <int K, int LAY, bool IMM, bool FROZEN, bool NORM>
__device__ __forceinline__ float run_task_ab(const CUtensorMap* tmap, const TkArgs& a, float* dst_base,
                                             int strip, int r_lo, int r_hi, uint32_t ring, uint32_t bars, int lane) {
  constexpr int NC = Lay<LAY>::NC, W = Lay<LAY>::W, HALF = Lay<LAY>::HALF;
  constexpr int STAGE_BYTES = TK_BR * W * 4;
  const Geo& g = a.g;
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
  int smA = 0, smB = 0, nmA = 0, nmB = 0;
#pragma unroll
  for (int e = 0; e < NC; ++e) {
    if (validA && (jA + e) >= vbase && (jA + e) < g.cols) smA |= 1 << e;
    if (validB && (jB + e) >= vbase && (jB + e) < g.cols) smB |= 1 << e;
    if (validA && (jA + e) >= own_lo && (jA + e) < g.cols) nmA |= 1 << e;
    if (validB && (jB + e) >= own_lo && (jB + e) < g.cols) nmB |= 1 << e;
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

  auto row_step = [&](auto Ptag) {
    constexpr int PH = decltype(Ptag)::value;
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
          if (NORM) cprev[j] = c;
        }
      }
    }
    if ((unsigned)ro < (unsigned)n_out) {
      float fa[NC], fb[NC];
#pragma unroll
      for (int j = 0; j < NC; ++j) { fa[j] = fin[j].x; fb[j] = fin[j].y; }
      store_seg(q, fa, smA);
      store_seg(q + HALF, fb, smB);
      if (NORM) {
#pragma unroll
        for (int j = 0; j < NC; ++j) {
          const float da = ((nmA >> j) & 1) ? fin[j].x - cprev[j].x : 0.f;
          const float db = ((nmB >> j) & 1) ? fin[j].y - cprev[j].y : 0.f;
          if (a.norm == H2D_NORM_L2SQ) {
            acc = fmaf(da, da, acc); acc = fmaf(db, db, acc);
          } else {
            acc = fmax_nan(acc, fmax_nan(fabsf(da), fabsf(db)));
          }
        }
      }
    }
    q += g.pitch;
    ++ro;
  };

#pragma unroll 1
  for (int b = 0; b < n_blk; ++b) {
    const int s = b % TK_STAGES;
    mbar_wait(bars + s * 8, (uint32_t)((b / TK_STAGES) & 1));
    const uint32_t sbase = ring + s * STAGE_BYTES + lane * (NC * 4);
#pragma unroll 1
    for (int rr = 0; rr < TK_BR; rr += 3) {
      load_row(sbase + rr * (W * 4), st[0][2]);
      row_step(std::integral_constant<int, 0>{});
      load_row(sbase + (rr + 1) * (W * 4), st[0][0]);
      row_step(std::integral_constant<int, 1>{});
      load_row(sbase + (rr + 2) * (W * 4), st[0][1]);
      row_step(std::integral_constant<int, 2>{});
    }
    __syncwarp();
    if (lane == 0 && b + TK_STAGES < n_blk) issue(b + TK_STAGES);
  }
  return acc;
}

// One entry point for both families.  `track` is meaningful for strict modes only.
template <int K, int ARITH, int LAY, bool IMM, bool FROZEN, bool NORM, bool FAST>
__device__ __forceinline__ float run_task(const CUtensorMap* tmap, const TkArgs& a, float* dst_base,
                                          int strip, int r_lo, int r_hi, uint32_t ring, uint32_t bars, int lane,
                                          uint32_t& track) {
  if constexpr (LAY == LAY_V4) {
    return run_task_v4<K, ARITH, IMM, FROZEN, NORM, FAST>(tmap, a, dst_base, strip, r_lo, r_hi, ring, bars, lane, track);
  } else if constexpr (LAY == LAY_V8) {
    return run_task_v8<K, ARITH, IMM, FROZEN, NORM, FAST>(tmap, a, dst_base, strip, r_lo, r_hi, ring, bars, lane, track);
  } else {
    static_assert(ARITH == H2D_ARITH_FP32, "packed layouts are fp32 only");
    track = 0u;
    return run_task_ab<K, LAY, IMM, FROZEN, NORM>(tmap, a, dst_base, strip, r_lo, r_hi, ring, bars, lane);
  }
}

}  // namespace h2d
