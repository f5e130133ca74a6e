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
// Work decomposition.  One warp = one task = (strip, chunk):
//   strip  128 consecutive columns (each lane owns 4 = one float4); the outer
//          H = roundup4(K) columns on each side are the overlap whose results
//          go stale as the K levels advance, the middle 128-2H are stored;
//   chunk  R output rows; the warp streams R + 2K input rows top to bottom.
// Each warp runs its own 4-stage TMA ring of 6-row boxes (no block-wide barrier in the loop).
// Per input row the K levels advance by one row each; level t keeps two rows
// (up, centre) in registers, so the state is 2*4*K floats per lane.
#include <cuda.h>
#include <cstdio>
#include <cstdlib>
#include <cmath>
#include <cstring>
#include <type_traits>
#include <vector>
#include "device_common.cuh"

namespace h2d {

constexpr int TK_W = 128;      // floats per strip row
constexpr int TK_BR = 6;       // rows per TMA box (a multiple of the 3-row register rotation)
constexpr int TK_STAGES = 4;   // boxes in flight per warp
constexpr int TK_WARPS = 4;    // warps (tasks) per CTA
constexpr int TK_STAGE_BYTES = TK_BR * TK_W * 4;
constexpr int TK_WE_ROWS = 16;   // rows of column-halo data staged per remote store
constexpr int TK_WE_COLS = 32;   // columns forwarded per row and side: one full 128-byte line
constexpr int TK_WE_Q = TK_WE_COLS / 4;               // float4 pieces per row and side
constexpr int TK_WE_BYTES = 2 * TK_WE_ROWS * TK_WE_COLS * 4;   // per warp: W and E side
constexpr int TK_SMEM_RING = TK_WARPS * TK_STAGES * TK_STAGE_BYTES;
constexpr int TK_SMEM_BARS = TK_WARPS * TK_STAGES * 8;
constexpr int TK_SMEM_BYTES = TK_SMEM_RING + TK_SMEM_BARS + TK_WARPS * TK_WE_BYTES;
constexpr int TK_KMAX = 8;

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
  int32_t row_base, row_end;   // rows [row_base, row_end) are processed; chunk c starts at row_base + c*R
  int32_t n_strips;            // tasks == nullptr: task -> (task % n_strips, task / n_strips)
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
  int32_t debug;   // bit 0: skip the peer stores (timing experiments only; H2D_DEBUG_NOFWD=1)
                   // bit 1: random per-task delays around the flag protocol (H2D_DEBUG_DELAY=1)
};

// ------------------------------------------------------------ PTX wrappers
__device__ __forceinline__ uint32_t smem_u32(const void* p) {
  return (uint32_t)__cvta_generic_to_shared(p);
}
__device__ __forceinline__ void mbar_init(uint32_t bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count) : "memory");
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

// --------------------------------------------------- peer forwarding (cold)
// Store the 4 just-computed values of owned row `rho`, columns j0..j0+3, into
// every neighbour whose ghost margin contains them.  Called only by tasks that
// touch the tile edge.  `smask` = which of the 4 elements are owned outputs.
__device__ __noinline__ void forward_row(const TkArgs& a, int rho, int j0, float4 o, int smask,
                                         bool skip_we) {
  const int fw = a.fw;
  const float v[4] = {o.x, o.y, o.z, o.w};
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
    for (int e = 0; e < 4; ++e) {
      const int j = j0 + e;
      const bool in = (sy == 0) || (sy < 0 ? j < fw : j >= a.g.cols - fw);
      if (in && ((smask >> e) & 1)) m |= 1 << e;
    }
    if (m == 0 || (a.debug & 1)) continue;
    float* q = p.buf + (int64_t)(a.g.x0 + rho - p.x0 + p.halo_r) * p.pitch +
               (a.g.y0 + j0 - p.y0 + p.halo_c);
    if (m == 15 && (((uintptr_t)q) & 15) == 0) {
      *reinterpret_cast<float4*>(q) = o;
    } else {
#pragma unroll
      for (int e = 0; e < 4; ++e)
        if ((m >> e) & 1) q[e] = v[e];
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
// TK_WE_ROWS rows: lane l sends pieces l and l+32 of the 16 x 8 float4 block.
// `own` = which of the 8 pieces per side (bits 0-7: W, 8-15: E) this warp holds.
__device__ __noinline__ void flush_we(const TkArgs& a, uint32_t stage, int rho0, int count, int own,
                                      int lane) {
  __syncwarp();
#pragma unroll
  for (int side = 0; side < 2; ++side) {
    const PeerDev& p = a.peer[side == 0 ? H2D_W : H2D_E];
    if (p.buf == nullptr || ((own >> (8 * side)) & 0xff) == 0) continue;
#pragma unroll
    for (int it = 0; it < (TK_WE_ROWS * TK_WE_Q) / 32; ++it) {
      const int idx = it * 32 + lane;
      const int row = idx / TK_WE_Q, piece = idx % TK_WE_Q;
      if (row >= count || !((own >> (8 * side + piece)) & 1)) continue;
      const float4 v = lds128(stage + (uint32_t)(((side * TK_WE_ROWS + row) * TK_WE_Q + piece) * 16));
      const int jcol = (side == 0 ? 0 : a.g.cols - TK_WE_COLS) + piece * 4;   // tile-local column
      float* q = p.buf + (int64_t)(a.g.x0 + rho0 + row - p.x0 + p.halo_r) * p.pitch +
                 (a.g.y0 + jcol - p.y0 + p.halo_c);
      *reinterpret_cast<float4*>(q) = v;
    }
  }
  __syncwarp();
}

// ------------------------------------------------------------- one task
// FROZEN: the footprint contains cells that keep their value (global boundary
// or outside the grid) -> per-element select.  XM: 0 = plain task; 1 = the task
// feeds the convergence norm; 2 = it also (or only) forwards rows to neighbours.
// Both are warp-uniform and resolved at compile time so the common case
// (interior, no extras) carries no overhead in its inner loop, and a norm-only
// task (every task of a convergence-check launch on one GPU, every interior task
// on several) does not carry the forwarding calls: with them the row loop is 712
// instructions per 3 rows instead of ~440 (397 for a plain task).
constexpr int XM_NONE = 0, XM_NORM = 1, XM_COMM = 2;
template <int K, int ARITH, bool IMM, bool FROZEN, int XM>
__device__ __forceinline__ float run_task(const CUtensorMap* tmap, const TkArgs& a, float* dst_base,
                                          int strip, int chunk, bool comm_edge, uint32_t ring,
                                          uint32_t bars, uint32_t we_stage, int lane) {
  constexpr bool EXTRAS = XM != XM_NONE;   // accumulates the norm (when the launch has one)
  constexpr bool COMM = XM == XM_COMM;     // forwards to neighbours
  const Geo& g = a.g;
  const int R = a.chunk_rows;
  const int r_lo = a.row_base + chunk * R;
  const int r_hi = min(r_lo + R, a.row_end);
  const int n_out = r_hi - r_lo;
  // The last strip is shifted left so that its valid window ends at the tile edge: its
  // first columns are then computed twice (by it and by its left neighbour) but it OWNS
  // (stores / reduces) only its own range [own_lo, cols).  No strip is narrow any more and,
  // what matters for column halos, the last 32 columns always sit in ONE warp, so the east
  // band leaves as whole 128-byte lines (half-line remote writes made every rank with a
  // west neighbour ~2x slower: 1x4 strips of 8192 columns, last strip 16 wide).
  const int own_lo = strip * a.wv;
  // (rounded up to a multiple of 4: float4 stores and TMA box origins must stay 16-byte aligned)
  const int vbase = (strip == a.n_strips - 1 && g.cols >= a.wv) ? ((g.cols - a.wv + 3) & ~3) : own_lo;
  const int jbase = vbase - a.H;         // tile-local column of lane 0, element 0
  const int j0 = jbase + lane * 4;
  const int row0 = r_lo - K;             // first input row (tile-local)
  // (A "skewed" variant in which level t consumes what level t-1 produced in the
  // previous row step — K independent levels per step instead of a K-long chain —
  // was measured 11 % SLOWER on B200: the chain is not the limiter, see DESIGN.md.)
  constexpr int LAG = 2 * K;  // row steps between reading a row and storing the row K below it
  const int n_in = n_out + LAG;
  const int n_blk = (n_in + TK_BR - 1) / TK_BR;
  const int tc0 = g.halo_c + jbase;      // TMA coordinates of the first box
  const int tc1 = g.halo_r + row0;

  // Which of my 4 columns are stored: inside the strip's valid window and owned.
  const bool lane_in = (lane * 4 >= a.H) && (lane * 4 < TK_W - a.H);
  int smask = 0;
#pragma unroll
  for (int e = 0; e < 4; ++e)
    if (lane_in && (j0 + e) >= own_lo && (j0 + e) < g.cols) smask |= 1 << e;
  const bool st_full = smask == 15;
  const bool st_part = smask != 0 && smask != 15;

  // Batched column halos (EXTRAS only): usable when the 32-column bands are whole
  // float4s of single lanes and fit the neighbour's margin.  we_piece: 0..7 = my
  // float4 is that piece of the W band, 8..15 = of the E band, -1 = not in a band.
  int we_piece = -1, we_own = 0, we_count = 0, we_rho0 = 0;
  bool we_batched = false;
  bool we_perrow = false;   // W/E peers exist but their columns must go row by row
  if (COMM) {
    const bool w_ok = a.peer[H2D_W].buf != nullptr && a.peer[H2D_W].halo_c >= TK_WE_COLS;
    const bool e_ok = a.peer[H2D_E].buf != nullptr && a.peer[H2D_E].halo_c >= TK_WE_COLS;
    we_batched = comm_edge && a.fw <= 8 && (g.cols & 3) == 0 && g.cols >= 2 * TK_WE_COLS && (w_ok || e_ok) &&
                 (a.peer[H2D_W].buf == nullptr || w_ok) && (a.peer[H2D_E].buf == nullptr || e_ok);
    if (we_batched && lane_in && j0 >= 0 && j0 + 3 < g.cols) {   // computed here, owned or not
      if (w_ok && j0 >= 0 && j0 < TK_WE_COLS) we_piece = j0 >> 2;
      if (e_ok && j0 >= g.cols - TK_WE_COLS && j0 < g.cols) we_piece = 8 + ((j0 - (g.cols - TK_WE_COLS)) >> 2);
    }
    if (we_batched) {
#pragma unroll
      for (int h = 0; h < 16; ++h)
        if (__ballot_sync(0xffffffffu, we_piece == h) != 0u) we_own |= 1 << h;
    }
    // batched mode covers every W/E column band of the tile; a strip without band
    // columns simply stages nothing
    we_perrow = comm_edge && !we_batched &&
                (a.peer[H2D_W].buf != nullptr || a.peer[H2D_E].buf != nullptr);
  }
  const bool we_skip = we_batched;            // forward_row leaves pure W / E to flush_we
  if (we_own == 0) we_batched = false;        // nothing to stage in this strip

  bool act[4] = {true, true, true, true};
  if (FROZEN) {
#pragma unroll
    for (int e = 0; e < 4; ++e) {
      const int gy = g.y0 + j0 + e;
      act[e] = gy >= g.ay_lo && gy <= g.ay_hi;
    }
  }

  auto issue = [&](int b) {
    const int s = b % TK_STAGES;
    const uint32_t bar = bars + s * 8;
    mbar_expect_tx(bar, TK_STAGE_BYTES);
    tma_load_2d(ring + s * TK_STAGE_BYTES, tmap, tc0, tc1 + b * TK_BR, bar);
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
  // Level-K output of iteration i is tile-local row  row0 - K + i;  `ro` counts
  // rows relative to r_lo so that "is an owned output row" is one unsigned compare.
  int ro = -LAG;
  float* q = dst_base + (int64_t)(g.halo_r + r_lo - LAG) * g.pitch + g.halo_c + j0;

  // One row step: the input row is already in st[0][(PH+2)%3]; levels run
  // 0..K-1, each writing its result straight into the next level's incoming slot.
  auto row_step = [&](auto Ptag) {
    constexpr int PH = decltype(Ptag)::value;
    constexpr int iu = PH, ic = (PH + 1) % 3, id = (PH + 2) % 3;
    float4 fin;
    float4 cprev = make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
    for (int t = 0; t < K; ++t) {
      const float4 up = st[t][iu];
      const float4 c = st[t][ic];
      const float4 down = st[t][id];
      const float L = __shfl_up_sync(0xffffffffu, c.w, 1);
      const float Rn = __shfl_down_sync(0xffffffffu, c.x, 1);
      float4 o = point4<ARITH, IMM>(up, c, down, L, Rn, g);
      if (FROZEN) {
        // tile-local row of this level's output
        const int gx = g.x0 + r_lo + ro + (K - 1 - t);
        const bool ract = gx >= g.ax_lo && gx <= g.ax_hi;
        o.x = (ract && act[0]) ? o.x : c.x;
        o.y = (ract && act[1]) ? o.y : c.y;
        o.z = (ract && act[2]) ? o.z : c.z;
        o.w = (ract && act[3]) ? o.w : c.w;
      }
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
        if (COMM && comm_edge && (we_perrow || r_lo + ro < a.fw || r_lo + ro >= g.rows - a.fw))
          forward_row(a, r_lo + ro, j0, fin, smask, we_skip);
        if (a.norm != H2D_NORM_NONE) {
          const float d0 = (smask & 1) ? fin.x - cprev.x : 0.f;
          const float d1 = (smask & 2) ? fin.y - cprev.y : 0.f;
          const float d2 = (smask & 4) ? fin.z - cprev.z : 0.f;
          const float d3 = (smask & 8) ? fin.w - cprev.w : 0.f;
          if (a.norm == H2D_NORM_L2SQ) {
            acc = fmaf(d0, d0, acc); acc = fmaf(d1, d1, acc);
            acc = fmaf(d2, d2, acc); acc = fmaf(d3, d3, acc);
          } else {
            acc = fmaxf(acc, fmaxf(fmaxf(fabsf(d0), fabsf(d1)), fmaxf(fabsf(d2), fabsf(d3))));
          }
        }
      }
    }
    if (COMM && we_batched && (unsigned)ro < (unsigned)n_out) {
      if (we_count == 0) we_rho0 = r_lo + ro;
      if (we_piece >= 0) {
        const uint32_t slot =
            we_stage + (uint32_t)((((we_piece >> 3) * TK_WE_ROWS + we_count) * TK_WE_Q + (we_piece & 7)) * 16);
        asm volatile("st.shared.v4.f32 [%0], {%1, %2, %3, %4};" ::"r"(slot), "f"(fin.x), "f"(fin.y),
                     "f"(fin.z), "f"(fin.w) : "memory");
      }
      ++we_count;
      if (we_count == TK_WE_ROWS || ro == n_out - 1) {
        flush_we(a, we_stage, we_rho0, we_count, we_own, lane);
        we_count = 0;
      }
    }
    q += g.pitch;
    ++ro;
  };

#pragma unroll 1
  for (int b = 0; b < n_blk; ++b) {
    const int s = b % TK_STAGES;
    mbar_wait(bars + s * 8, (uint32_t)((b / TK_STAGES) & 1));
    const uint32_t sbase = ring + s * TK_STAGE_BYTES + lane * 16;
#pragma unroll 1
    for (int rr = 0; rr < TK_BR; rr += 3) {
      st[0][2] = lds128(sbase + rr * (TK_W * 4));
      row_step(std::integral_constant<int, 0>{});
      st[0][0] = lds128(sbase + (rr + 1) * (TK_W * 4));
      row_step(std::integral_constant<int, 1>{});
      st[0][1] = lds128(sbase + (rr + 2) * (TK_W * 4));
      row_step(std::integral_constant<int, 2>{});
    }
    // Everyone has read stage s: refill it with box b + STAGES.
    __syncwarp();
    if (lane == 0 && b + TK_STAGES < n_blk) issue(b + TK_STAGES);
  }
  return acc;
}

template <int K>
struct TkBounds {
  // K <= 5 fits 128 registers (4 CTAs = 16 warps per SM); deeper pipelines get
  // 168 registers (3 CTAs = 12 warps) instead of spilling.  Measured on B200 at
  // K = 8: 12 warps / 166 regs / no spills 4.43 TLUP/s; 13-14 warps with
  // __maxnreg__ 152 / 144 (32-72 B of spills) 4.31 TLUP/s; 16 warps at 128
  // registers (244 B of spills) 3.9 TLUP/s.
  static constexpr int min_blocks = (K <= 5) ? 4 : 3;
};

// XTR = false: a launch with neither peers nor a convergence norm (the common single-GPU
// launch): the kernel contains no forwarding / reduction code at all, so that code cannot
// cost the hot loop registers.
template <int K, int ARITH, bool IMM, bool XTR>
__global__ void __launch_bounds__(TK_WARPS * 32, TkBounds<K>::min_blocks)
tk_kernel(const __grid_constant__ CUtensorMap tmap, const __grid_constant__ TkArgs a) {
  extern __shared__ __align__(128) unsigned char smem[];
  __shared__ float cta_red[TK_WARPS];
  __shared__ bool cta_last;
  // Converged earlier (or a conditional launch whose condition did not occur):
  // every CTA of every rank reads the same word and leaves; no flag is
  // published and none is awaited, so the ranks stay consistent.
  if (a.ctl != nullptr && *reinterpret_cast<const volatile uint32_t*>(a.ctl) != a.run_if) return;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int task = (int)blockIdx.x * TK_WARPS + warp;
  const uint32_t ring = smem_u32(smem) + warp * (TK_STAGES * TK_STAGE_BYTES);
  const uint32_t bars = smem_u32(smem) + TK_SMEM_RING + warp * TK_STAGES * 8;
  const uint32_t we_stage = smem_u32(smem) + TK_SMEM_RING + TK_SMEM_BARS + warp * TK_WE_BYTES;

  float acc = 0.f;
  if (task < a.n_tasks) {
    if (lane == 0) {
#pragma unroll
      for (int s = 0; s < TK_STAGES; ++s) mbar_init(bars + s * 8, 1);
      asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncwarp();
    int2 tk;
    if (a.tasks != nullptr) {
      tk = a.tasks[task];
    } else if (XTR && a.has_peers) {
      // Row window of a tile with W / E neighbours (pipelined solve): no task list; the
      // tasks of the first and the last strip — the ones that hold the 32-column bands —
      // come first, chunk by chunk, then the interior strips.
      const int ns = a.n_strips, ne = ns >= 2 ? 2 : 1, n_edge = ne * a.n_chunks;
      if (task < n_edge) {
        tk = make_int2((int)(((unsigned)((task % ne) ? ns - 1 : 0)) | 0x80000000u), task / ne);
      } else {
        const int t2 = task - n_edge;
        tk = make_int2(1 + t2 % (ns - ne), t2 / (ns - ne));
      }
    } else {
      tk = make_int2(task % a.n_strips, task / a.n_strips);
    }
    const int strip = tk.x & 0x7fffffff;
    const bool comm_edge = XTR && (tk.x < 0) && a.has_peers;
    const int chunk = tk.y;

    if (comm_edge && (a.debug & 2)) {
      // Stress mode (race detection): shake the order in which ranks reach
      // the flag protocol with a pseudo-random delay of up to ~64 us per edge task.
      const uint32_t h = ((uint32_t)task * 2654435761u) ^ (a.epoch * 40503u) ^ ((uint32_t)a.g.x0 * 97u + (uint32_t)a.g.y0);
      const uint64_t t0 = globaltimer_ns();
      while (globaltimer_ns() - t0 < (uint64_t)((h >> 7) & 0xffffu)) __nanosleep(256);
    }
    if (comm_edge) {
      // Ghost cells of src were written by the neighbours' launch `epoch-1`;
      // their ghost cells of OUR dst parity were last read by that launch too.
      if (lane == 0) {
        for (int d = 0; d < H2D_NDIR; ++d) {
          if (a.peer[d].buf == nullptr) continue;
          // fast path: the neighbour published long ago (its edge tasks run first)
          if ((int32_t)(ld_acquire_sys(a.flags_in + d) - (a.epoch - 1u)) >= 0) continue;
          const uint64_t t0 = globaltimer_ns();
          if (!wait_flag_ge(a.flags_in + d, a.epoch - 1u, a.timeout_ns)) atomicExch(a.fault, 1u + d);
          atomicAdd(a.stats, (unsigned long long)(globaltimer_ns() - t0));   // the mpiP "Waitall" row
          atomicAdd(a.stats + 1, 1ull);
        }
        asm volatile("fence.proxy.async;" ::: "memory");
      }
      __syncwarp();
    }

    // Does the footprint (all K levels) stay inside the updated region?
    const Geo& g = a.g;
    const int R = a.chunk_rows;
    const int r_lo = a.row_base + chunk * R;
    const int r_hi = min(r_lo + R, a.row_end);
    const int jb = ((strip == a.n_strips - 1 && g.cols >= a.wv) ? ((g.cols - a.wv + 3) & ~3) : strip * a.wv) - a.H;
    const bool frozen = (g.x0 + r_lo - K < g.ax_lo) || (g.x0 + r_hi - 1 + K > g.ax_hi) ||
                        (g.y0 + jb < g.ay_lo) || (g.y0 + jb + TK_W - 1 > g.ay_hi);
    constexpr int XC = XTR ? XM_COMM : XM_NONE, XN = XTR ? XM_NORM : XM_NONE;
    if (XTR && comm_edge && frozen) {
      acc = run_task<K, ARITH, IMM, true, XC>(&tmap, a, a.dst, strip, chunk, true, ring, bars, we_stage, lane);
    } else if (XTR && comm_edge) {
      acc = run_task<K, ARITH, IMM, false, XC>(&tmap, a, a.dst, strip, chunk, true, ring, bars, we_stage, lane);
    } else if (XTR && a.norm != H2D_NORM_NONE && frozen) {
      acc = run_task<K, ARITH, IMM, true, XN>(&tmap, a, a.dst, strip, chunk, false, ring, bars, we_stage, lane);
    } else if (XTR && a.norm != H2D_NORM_NONE) {
      acc = run_task<K, ARITH, IMM, false, XN>(&tmap, a, a.dst, strip, chunk, false, ring, bars, we_stage, lane);
    } else if (frozen) {
      acc = run_task<K, ARITH, IMM, true, XM_NONE>(&tmap, a, a.dst, strip, chunk, false, ring, bars, we_stage, lane);
    } else {
      acc = run_task<K, ARITH, IMM, false, XM_NONE>(&tmap, a, a.dst, strip, chunk, false, ring, bars, we_stage, lane);
    }

    if (comm_edge) {
      // Make this warp's peer stores visible system-wide, then count it in; the
      // last edge task of the launch publishes the epoch to every neighbour.
      __threadfence_system();
      __syncwarp();
      if (lane == 0) {
        const unsigned prev = atomicAdd(a.edge_counter, 1u);
        if (prev == (unsigned)a.n_comm_edge - 1u) {
          *a.edge_counter = 0u;
          __threadfence_system();
          for (int d = 0; d < H2D_NDIR; ++d)
            if (a.peer[d].buf != nullptr) st_release_sys(a.peer[d].flag, a.epoch);
        }
      }
    }
  }

  if (!XTR || a.norm == H2D_NORM_NONE) return;
  // Deterministic two-level reduction: warp -> CTA -> last CTA.
  const bool l2 = a.norm == H2D_NORM_L2SQ;
  acc = l2 ? warp_sum(acc) : warp_max(acc);
  if (lane == 0) cta_red[warp] = acc;
  __syncthreads();
  if (threadIdx.x == 0) {
    float v = cta_red[0];
    for (int w = 1; w < TK_WARPS; ++w) v = l2 ? v + cta_red[w] : fmaxf(v, cta_red[w]);
    a.cta_partials[blockIdx.x] = v;
    __threadfence();
    cta_last = atomicAdd(a.done_counter, 1u) == gridDim.x - 1;
  }
  __syncthreads();
  if (!cta_last) return;
  __threadfence();
  float v = 0.f;
  for (int i = threadIdx.x; i < (int)gridDim.x; i += TK_WARPS * 32) {
    const float x = __ldcg(a.cta_partials + i);
    v = l2 ? v + x : fmaxf(v, x);
  }
  v = l2 ? warp_sum(v) : warp_max(v);
  __syncthreads();
  if (lane == 0) cta_red[warp] = v;
  __syncthreads();
  if (threadIdx.x == 0) {
    float t = cta_red[0];
    for (int w = 1; w < TK_WARPS; ++w) t = l2 ? t + cta_red[w] : fmaxf(t, cta_red[w]);
    *a.norm_out = t;
    *a.done_counter = 0u;
  }
}

// ----------------------------------------------------------- dataflow variant
// The same tasks, but ONE launch runs `n_epochs` k-step sweeps: the grid holds one warp
// per (epoch, task) pair.  Every warp draws a ticket from a global counter — tickets,
// not blockIdx, define the order, so nothing depends on how the hardware hands out
// CTAs — and ticket -> (epoch e, task t), epoch-major.  Task (e, t) may start as soon as
// the tasks of epoch e-1 whose stored cells it reads — and whose read windows cover the
// cells it is about to overwrite, the same set — have finished: chunks c-1..c+1 x strips
// s-2..s+2 (a read window reaches K rows and H columns into the adjacent chunks /
// strips; the shifted last strip reaches two strips back).  Each task publishes "epochs
// completed" in done[t] with release semantics.  A task only ever waits on LOWER
// tickets, which are held by warps that are already running, so progress never depends
// on how many CTAs are co-resident.
//
// What this removes is the drain at every launch boundary (the last wave of a launch is
// partly empty and the next launch ramps up again: 125 launches per 1000 steps).  The
// reference has one launch per step and the same drain 1000 times
// (grad1612_cuda_heat.cu:82-85).  Single tile, no norm: the forwarding / reduction code
// stays in tk_kernel<.., XTR = true>.
//
// (A first version looped over tickets inside a persistent CTA; ptxas then put 18
// register moves per 3 rows into the steady-state row loop — 418 instead of 397
// instructions — and it ran 4-7 % slower than tk_kernel on B200.  One task per warp per
// CTA, as in tk_kernel, keeps the row loop's register rotation intact.)
struct FlowArgs {
  float* dst[2];       // destination of even / odd epochs
  unsigned* ticket;    // zeroed by the host before the launch
  uint32_t* done;      // per task (chunk * n_strips + strip): base + epochs completed
  uint32_t base;       // value of "0 epochs completed" for this launch
  int32_t n_epochs;
};

__device__ __forceinline__ uint32_t ld_acquire_gpu(const uint32_t* p) {
  uint32_t v;
  asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ void st_release_gpu(uint32_t* p, uint32_t v) {
  asm volatile("st.release.gpu.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}

constexpr uint32_t TK_FAULT_DEP = 50u;   // fault code: a task dependency never arrived

// Draw the next ticket and wait for its dependencies.  Returns the ticket, or -1 when
// there is nothing (left) to do: all tickets drawn, or a dependency timed out somewhere.
// Out of line: the register allocation of the row loop is sensitive to what is live
// around it.
__device__ __forceinline__ int flow_next(const TkArgs& a, const FlowArgs& pa, int lane) {
  unsigned tk = 0;
  if (lane == 0) tk = atomicAdd(pa.ticket, 1u);
  tk = __shfl_sync(0xffffffffu, tk, 0);
  const int n_tasks = a.n_tasks, ns = a.n_strips, nc = a.n_chunks;
  const int e = (int)(tk / (unsigned)n_tasks);
  if (e >= pa.n_epochs) return -1;
  if (e == 0) return (int)tk;
  const int task = (int)(tk - (unsigned)e * (unsigned)n_tasks);
  const int chunk = task / ns, strip = task - chunk * ns;
  bool ok = true;
  if (lane < 15) {
    const int c2 = chunk + lane / 5 - 1, s2 = strip + lane % 5 - 2;
    if (c2 >= 0 && c2 < nc && s2 >= 0 && s2 < ns) {
      const uint32_t* f = pa.done + (c2 * ns + s2);
      const uint32_t want = pa.base + (uint32_t)e;
      if ((int32_t)(ld_acquire_gpu(f) - want) < 0) {
        const uint64_t t0 = globaltimer_ns();
        while ((int32_t)(ld_acquire_gpu(f) - want) < 0) {
          __nanosleep(100);
          if (*reinterpret_cast<volatile uint32_t*>(a.fault) != 0u) { ok = false; break; }
          if ((int64_t)(globaltimer_ns() - t0) > a.timeout_ns) {
            atomicExch(a.fault, TK_FAULT_DEP);
            ok = false;
            break;
          }
        }
      }
    }
  }
  // somebody gave up (or an earlier fault was seen): leave without computing on stale
  // data; every other warp sees the fault word at its next wait
  if (!__all_sync(0xffffffffu, ok)) return -1;
  // the cells were written with ordinary stores by other SMs; TMA reads them through
  // the async proxy
  asm volatile("fence.proxy.async;" ::: "memory");
  return (int)tk;
}

// Every lane makes its stores visible device-wide, then one release for the task.
__device__ __forceinline__ void flow_publish(const FlowArgs& pa, int task, uint32_t value, int lane) {
  __threadfence();
  asm volatile("fence.proxy.async;" ::: "memory");
  __syncwarp();
  if (lane == 0) st_release_gpu(pa.done + task, value);
}

template <int K, int ARITH, bool IMM>
__global__ void __launch_bounds__(TK_WARPS * 32, TkBounds<K>::min_blocks)
tk_flow_kernel(const __grid_constant__ CUtensorMap tmap0, const __grid_constant__ CUtensorMap tmap1,
               const __grid_constant__ TkArgs a, const __grid_constant__ FlowArgs pa) {
  extern __shared__ __align__(128) unsigned char smem[];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const uint32_t ring = smem_u32(smem) + warp * (TK_STAGES * TK_STAGE_BYTES);
  const uint32_t bars = smem_u32(smem) + TK_SMEM_RING + warp * TK_STAGES * 8;
  const int tk = flow_next(a, pa, lane);
  if (tk < 0) return;
  const int e = tk / a.n_tasks;
  const int task = tk - e * a.n_tasks;
  const int chunk = task / a.n_strips, strip = task - chunk * a.n_strips;
  if (lane == 0) {
#pragma unroll
    for (int s = 0; s < TK_STAGES; ++s) mbar_init(bars + s * 8, 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  __syncwarp();

  const Geo& g = a.g;
  const int R = a.chunk_rows;
  const int r_lo = a.row_base + chunk * R;
  const int r_hi = min(r_lo + R, a.row_end);
  const int jb = ((strip == a.n_strips - 1 && g.cols >= a.wv) ? ((g.cols - a.wv + 3) & ~3) : strip * a.wv) - a.H;
  const bool frozen = (g.x0 + r_lo - K < g.ax_lo) || (g.x0 + r_hi - 1 + K > g.ax_hi) ||
                      (g.y0 + jb < g.ay_lo) || (g.y0 + jb + TK_W - 1 > g.ay_hi);
  const CUtensorMap* tm = (e & 1) ? &tmap1 : &tmap0;
  float* dst = pa.dst[e & 1];
  if (frozen) {
    run_task<K, ARITH, IMM, true, XM_NONE>(tm, a, dst, strip, chunk, false, ring, bars, 0u, lane);
  } else {
    run_task<K, ARITH, IMM, false, XM_NONE>(tm, a, dst, strip, chunk, false, ring, bars, 0u, lane);
  }
  flow_publish(pa, task, pa.base + (uint32_t)e + 1u, lane);
}

typedef void (*tk_flow_fn_t)(const CUtensorMap, const CUtensorMap, const TkArgs, const FlowArgs);

template <int ARITH, bool IMM>
static tk_flow_fn_t tk_flow_pick_k(int k) {
  switch (k) {
    case 1: return tk_flow_kernel<1, ARITH, IMM>;
    case 2: return tk_flow_kernel<2, ARITH, IMM>;
    case 3: return tk_flow_kernel<3, ARITH, IMM>;
    case 4: return tk_flow_kernel<4, ARITH, IMM>;
    case 5: return tk_flow_kernel<5, ARITH, IMM>;
    case 6: return tk_flow_kernel<6, ARITH, IMM>;
    case 7: return tk_flow_kernel<7, ARITH, IMM>;
    case 8: return tk_flow_kernel<8, ARITH, IMM>;
    default: return nullptr;
  }
}
static tk_flow_fn_t tk_flow_pick(int k, int arith, bool imm) {
  switch (arith) {
    case H2D_ARITH_FP32:
      return imm ? tk_flow_pick_k<H2D_ARITH_FP32, true>(k) : tk_flow_pick_k<H2D_ARITH_FP32, false>(k);
    case H2D_ARITH_STRICT: return tk_flow_pick_k<H2D_ARITH_STRICT, false>(k);
    case H2D_ARITH_STRICT_FMA: return tk_flow_pick_k<H2D_ARITH_STRICT_FMA, false>(k);
    default: return nullptr;
  }
}

typedef void (*tk_fn_t)(const CUtensorMap, const TkArgs);

template <int ARITH, bool IMM, bool XTR>
static tk_fn_t tk_pick_k(int k) {
  switch (k) {
    case 1: return tk_kernel<1, ARITH, IMM, XTR>;
    case 2: return tk_kernel<2, ARITH, IMM, XTR>;
    case 3: return tk_kernel<3, ARITH, IMM, XTR>;
    case 4: return tk_kernel<4, ARITH, IMM, XTR>;
    case 5: return tk_kernel<5, ARITH, IMM, XTR>;
    case 6: return tk_kernel<6, ARITH, IMM, XTR>;
    case 7: return tk_kernel<7, ARITH, IMM, XTR>;
    case 8: return tk_kernel<8, ARITH, IMM, XTR>;
    default: return nullptr;
  }
}
template <bool XTR>
static tk_fn_t tk_pick_x(int k, int arith, bool imm) {
  switch (arith) {
    case H2D_ARITH_FP32:
      return imm ? tk_pick_k<H2D_ARITH_FP32, true, XTR>(k) : tk_pick_k<H2D_ARITH_FP32, false, XTR>(k);
    case H2D_ARITH_STRICT: return tk_pick_k<H2D_ARITH_STRICT, false, XTR>(k);
    case H2D_ARITH_STRICT_FMA: return tk_pick_k<H2D_ARITH_STRICT_FMA, false, XTR>(k);
    default: return nullptr;
  }
}
// `imm`: fp32 mode with exactly the reference coefficients -> immediate-operand kernels.
// `xtr`: the launch forwards halos and/or reduces a norm.
static tk_fn_t tk_pick(int k, int arith, bool imm, bool xtr) {
  return xtr ? tk_pick_x<true>(k, arith, imm) : tk_pick_x<false>(k, arith, imm);
}

static bool coef_is_reference(const Geo& g, int arith) {
  return arith == H2D_ARITH_FP32 && g.cx == kRefCx && g.cy == kRefCy && g.a0 == kRefA0;
}

}  // namespace h2d

using namespace h2d;

// ------------------------------------------------------------------ the plan
struct TaskList {
  int2* dev = nullptr;
  int n_tasks = 0, n_strips = 0, n_comm_edge = 0;
};

struct h2d_plan {
  h2d_plan_desc_t desc;
  Geo geo;
  CUtensorMap tmap[2];
  int n_chunks = 0, chunk_rows = 0, fw = 0;
  TaskList lists[TK_KMAX / 4 + 1];  // index H/4 (H = 4, 8)
  void* scratch = nullptr;          // counters + partials
  bool own_scratch = false;
  int max_ctas = 0;
  int64_t launches = 0;
  int device = 0;
  bool has_peers = false;
  // lazily created by h2d_plan_run_dataflow: [0] ticket, [64...] done[] per task
  uint32_t* flow = nullptr;
  size_t flow_tasks = 0;
  uint32_t flow_base = 0;
  // lazily created by h2d_plan_solve_pipelined
  cudaStream_t s_h2d = nullptr, s_d2h = nullptr;
  std::vector<cudaEvent_t> ev_pool;
};

static inline int round_up4(int k) { return (k + 3) & ~3; }

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

static int choose_chunk_rows(int rows, int cols, int kmax, int min_waves = 1) {
  // Tasks cost about the same and the block scheduler hands them out in order,
  // so a launch takes about ceil(tasks / slots) "waves" of one task time, and a
  // task time is proportional to the rows it streams, R + 2K (K rows of pipeline
  // fill at both ends) plus ~8 rows of start-up.  A quarter wave is added for the
  // slower edge / boundary tasks.  Minimise  (ceil(waves) + 0.25) * (R + 2K + 8)
  // over the chunk count (constants fitted to sweeps on B200: 32768^2 -> R = 683,
  // 16384^2 -> 342, 8192x32768 -> 342, 4096x32768 -> 228, all within 1 % of the best).
  // slots = SMs x resident warps (16 up to K = 5, 12 beyond: register-limited).
  // Small grids land in the latency regime (one wave): short chunks win there.
  const int wv = TK_W - 2 * round_up4(kmax);
  const int n_strips = (cols + wv - 1) / wv;
  const double slots = (double)sm_count() * (kmax <= 5 ? 16 : 12);
  const int r_min = kmax > 8 ? kmax : 8;
  int best_R = rows;
  double best = 1e300;
  for (int c = 1; c <= rows; ++c) {
    const int R = (rows + c - 1) / c;   // equal chunks; the last one may be shorter
    if (R < r_min && c > 1) break;
    const double waves = ceil((double)((rows + R - 1) / R) * n_strips / slots);
    // With neighbours, the edge tasks (first wave) gate the neighbours' NEXT launch: keep
    // them a small part of the launch.  (1x4 column strips at 2 waves per launch: every
    // rank waited for its neighbours' edge tasks, 6798 instead of ~15000 GLUP/s on 4 B200.)
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
  return 4;
}

static void plan_counts(const h2d_tile_t* t, int kmax, int chunk_rows, int* n_chunks,
                        int* max_tasks, int min_waves = 1) {
  int R = chunk_rows > 0 ? chunk_rows : choose_chunk_rows(t->rows, t->cols, kmax, min_waves);
  int nc = (t->rows + R - 1) / R;
  if (nc < 1) nc = 1;
  *n_chunks = nc;
  int mt = 0;
  for (int H = 4; H <= round_up4(kmax); H += 4) {
    const int wv = TK_W - 2 * H;
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

static int make_tmap(CUtensorMap* out, float* base, const h2d_tile_t& t) {
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
  const cuuint32_t box[2] = {TK_W, TK_BR};
  const cuuint32_t estr[2] = {1, 1};
  CUresult r = fn(out, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 2, base, gdim, gstride, box, estr,
                  CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE,
                  CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) return h2d_fail(H2D_ECUDA, "cuTensorMapEncodeTiled failed with %d", (int)r);
  return H2D_OK;
}

static int build_task_list(h2d_plan* p, int H, TaskList* out) {
  const h2d_tile_t& t = p->desc.tile;
  const int wv = TK_W - 2 * H;
  const int ns = (t.cols + wv - 1) / wv;
  const int nc = p->n_chunks;
  const int R = p->chunk_rows;
  const int fw = p->fw;
  std::vector<int2> edge, inner;
  edge.reserve((size_t)2 * (ns + nc));
  inner.reserve((size_t)ns * nc);
  for (int c = 0; c < nc; ++c) {
    const int r_lo = c * R, r_hi = (c == nc - 1) ? t.rows : r_lo + R;
    const bool row_edge = r_lo < fw || r_hi > t.rows - fw;
    for (int s = 0; s < ns; ++s) {
      const int j_lo = s * wv, j_hi = (j_lo + wv < t.cols) ? j_lo + wv : t.cols;
      const bool col_edge = j_lo < fw || j_hi > t.cols - fw;
      if (row_edge || col_edge)
        edge.push_back(make_int2((int)((unsigned)s | 0x80000000u), c));
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

extern "C" {

int32_t h2d_plan_chunk_rows(const h2d_tile_t* tile, int32_t kmax, int32_t with_peers) {
  if (!tile || kmax < 1 || kmax > TK_KMAX || tile->rows < 1 || tile->cols < 1) return 0;
  return choose_chunk_rows(tile->rows, tile->cols, kmax, with_peers ? peer_min_waves() : 1);
}

size_t h2d_plan_scratch_bytes(const h2d_tile_t* tile, int32_t kmax, int32_t chunk_rows) {
  if (!tile || kmax < 1 || kmax > TK_KMAX || tile->rows < 1 || tile->cols < 1) return 0;
  int nc, mt, nc2, mt2;
  plan_counts(tile, kmax, chunk_rows, &nc, &mt);
  plan_counts(tile, kmax, chunk_rows, &nc2, &mt2, peer_min_waves());   // plans with peers use more, shorter chunks
  if (mt2 > mt) mt = mt2;
  return scratch_layout_bytes(mt);
}

void h2d_plan_destroy(h2d_plan_t* p) {
  if (!p) return;
  for (auto& l : p->lists)
    if (l.dev) cudaFree(l.dev);
  if (p->own_scratch && p->scratch) cudaFree(p->scratch);
  if (p->flow) cudaFree(p->flow);
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
  bool has_peers = false;
  for (int k = 0; k < H2D_NDIR; ++k)
    if (d->peer[k].buf[0] || d->peer[k].buf[1]) {
      if (!d->peer[k].buf[0] || !d->peer[k].buf[1] || !d->peer[k].flag) {
        h2d_fail(H2D_EINVAL, "h2d_plan_create: peer %d needs both buffers and a flag", k);
        return nullptr;
      }
      has_peers = true;
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
  cudaGetDevice(&p->device);
  int max_tasks = 0;
  const int min_waves = has_peers ? peer_min_waves() : 1;
  plan_counts(&t, d->kmax, d->chunk_rows, &p->n_chunks, &max_tasks, min_waves);
  p->chunk_rows = d->chunk_rows > 0 ? d->chunk_rows : choose_chunk_rows(t.rows, t.cols, d->kmax, min_waves);
  if (p->n_chunks == 1) p->chunk_rows = t.rows;
  p->fw = d->kmax;
  if (p->fw > t.rows) p->fw = t.rows;
  if (p->fw > t.cols) p->fw = t.cols;
  p->max_ctas = (max_tasks + TK_WARPS - 1) / TK_WARPS;

  if (make_tmap(&p->tmap[0], d->buf[0], t) || make_tmap(&p->tmap[1], d->buf[1], t)) {
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
  for (int k = 1; k <= d->kmax; ++k) {
    for (int x = 0; x < 2; ++x) {
      tk_fn_t fn = tk_pick(k, d->phys.arith, coef_is_reference(p->geo, d->phys.arith), x != 0);
      cudaFuncSetAttribute((const void*)fn, cudaFuncAttributeMaxDynamicSharedMemorySize, TK_SMEM_BYTES);
    }
  }
  return p;
}

// H2D_DEBUG_NOFWD=1: skip the peer stores (timing experiments); H2D_DEBUG_DELAY=1: random
// per-task delays around the flag protocol (stress test).
static int debug_flags() {
  static int dbg = -1;
  if (dbg < 0) {
    const char* e = getenv("H2D_DEBUG_NOFWD");
    const char* d = getenv("H2D_DEBUG_DELAY");
    dbg = ((e && e[0] == '1') ? 1 : 0) | ((d && d[0] == '1') ? 2 : 0);
  }
  return dbg;
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
  tk_fn_t fn = tk_pick(k, p->desc.phys.arith, coef_is_reference(p->geo, p->desc.phys.arith), xtr);
  if (!fn) return h2d_fail(H2D_EINVAL, "h2d_plan_launch: no kernel for k = %d", k);
  const int H = round_up4(k);
  const TaskList& tl = p->lists[H / 4];

  TkArgs a;
  memset(&a, 0, sizeof a);
  a.g = p->geo;
  a.dst = p->desc.buf[1 - src_parity];
  a.tasks = tl.dev;
  a.n_tasks = tl.n_tasks;
  a.n_chunks = p->n_chunks;
  a.chunk_rows = p->chunk_rows;
  a.row_base = 0;
  a.row_end = p->desc.tile.rows;
  a.n_strips = tl.n_strips;
  a.n_comm_edge = tl.n_comm_edge;
  a.H = H;
  a.wv = TK_W - 2 * H;
  a.fw = p->fw;
  a.has_peers = p->has_peers ? 1 : 0;
  for (int d = 0; d < H2D_NDIR; ++d) {
    const h2d_peer_t& pr = p->desc.peer[d];
    a.peer[d].buf = pr.buf[1 - src_parity];
    a.peer[d].flag = pr.flag;
    a.peer[d].x0 = pr.x0; a.peer[d].y0 = pr.y0;
    a.peer[d].halo_r = pr.halo_r; a.peer[d].halo_c = pr.halo_c;
    a.peer[d].pitch = pr.pitch;
  }
  a.flags_in = p->desc.flags_in;
  a.epoch = epoch;
  unsigned* hdr = (unsigned*)p->scratch;
  a.edge_counter = hdr + 0;
  a.done_counter = hdr + 1;
  a.norm_out = (float*)(hdr + 2);
  a.fault = hdr + 3;
  a.stats = (unsigned long long*)(hdr + 8);
  a.cta_partials = (float*)((char*)p->scratch + 256);
  a.norm = norm;
  a.timeout_ns = p->desc.spin_timeout_ns > 0 ? p->desc.spin_timeout_ns : 10000000000LL;
  a.ctl = ctl;
  a.run_if = run_if;
  a.debug = debug_flags();

  const int grid = (tl.n_tasks + TK_WARPS - 1) / TK_WARPS;
  fn<<<grid, TK_WARPS * 32, TK_SMEM_BYTES, (cudaStream_t)stream>>>(p->tmap[src_parity], a);
  H2D_CHECK_LAUNCH("tk_kernel");
  p->launches++;
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
// (no peers, no norm): tasks are enumerated arithmetically, chunk c starts at
// row_base + c * R.
static int launch_window(h2d_plan* p, int src_parity, int k, int row_base, int row_end, int chunk_rows,
                         cudaStream_t st, uint32_t epoch = 0u) {
  if (row_end <= row_base) return H2D_OK;
  const bool peers = p->has_peers;
  tk_fn_t fn = tk_pick(k, p->desc.phys.arith, coef_is_reference(p->geo, p->desc.phys.arith), peers);
  if (!fn) return h2d_fail(H2D_EINVAL, "launch_window: no kernel for k = %d", k);
  const int H = round_up4(k);
  TkArgs a;
  memset(&a, 0, sizeof a);
  a.g = p->geo;
  a.dst = p->desc.buf[1 - src_parity];
  a.tasks = nullptr;
  a.chunk_rows = chunk_rows;
  a.row_base = row_base;
  a.row_end = row_end;
  a.H = H;
  a.wv = TK_W - 2 * H;
  a.n_strips = (p->desc.tile.cols + a.wv - 1) / a.wv;
  a.n_chunks = (row_end - row_base + a.chunk_rows - 1) / a.chunk_rows;
  a.n_tasks = a.n_strips * a.n_chunks;
  a.fw = p->fw;
  unsigned* hdr = (unsigned*)p->scratch;
  a.edge_counter = hdr + 0;
  a.done_counter = hdr + 1;
  a.norm_out = (float*)(hdr + 2);
  a.fault = hdr + 3;
  a.stats = (unsigned long long*)(hdr + 8);
  a.cta_partials = (float*)((char*)p->scratch + 256);
  a.timeout_ns = p->desc.spin_timeout_ns > 0 ? p->desc.spin_timeout_ns : 10000000000LL;
  if (peers) {
    // W / E neighbours only (checked by the caller): the first and last strip of every
    // chunk forward the column bands and take part in the epoch protocol.
    a.has_peers = 1;
    a.n_comm_edge = (a.n_strips >= 2 ? 2 : 1) * a.n_chunks;
    for (int d = 0; d < H2D_NDIR; ++d) {
      const h2d_peer_t& pr = p->desc.peer[d];
      a.peer[d].buf = pr.buf[1 - src_parity];
      a.peer[d].flag = pr.flag;
      a.peer[d].x0 = pr.x0; a.peer[d].y0 = pr.y0;
      a.peer[d].halo_r = pr.halo_r; a.peer[d].halo_c = pr.halo_c;
      a.peer[d].pitch = pr.pitch;
    }
    a.flags_in = p->desc.flags_in;
    a.epoch = epoch;
    a.debug = debug_flags();
  }
  const int grid = (a.n_tasks + TK_WARPS - 1) / TK_WARPS;
  fn<<<grid, TK_WARPS * 32, TK_SMEM_BYTES, st>>>(p->tmap[src_parity], a);
  H2D_CHECK_LAUNCH("tk_kernel (row window)");
  p->launches++;
  return H2D_OK;
}

int h2d_plan_run_dataflow(h2d_plan_t* p, int src_parity, int k, int n_epochs, void* stream) {
  if (!p) return h2d_fail(H2D_EINVAL, "h2d_plan_run_dataflow: null plan");
  if (src_parity != 0 && src_parity != 1) return h2d_fail(H2D_EINVAL, "h2d_plan_run_dataflow: bad parity");
  if (k < 1 || k > p->desc.kmax)
    return h2d_fail(H2D_EINVAL, "h2d_plan_run_dataflow: k = %d outside [1, kmax = %d]", k, p->desc.kmax);
  if (n_epochs < 0) return h2d_fail(H2D_EINVAL, "h2d_plan_run_dataflow: negative epoch count");
  if (p->has_peers) return h2d_fail(H2D_EINVAL, "h2d_plan_run_dataflow: single-tile plans only");
  if (n_epochs == 0) return H2D_OK;
  const h2d_tile_t& t = p->desc.tile;
  const int H = round_up4(k);
  const int R = p->chunk_rows;
  // a read window must stay inside the adjacent chunks
  if (p->n_chunks > 1 && R < k)
    return h2d_fail(H2D_EINVAL, "h2d_plan_run_dataflow: chunk_rows %d < k = %d", R, k);
  const bool imm = coef_is_reference(p->geo, p->desc.phys.arith);
  tk_flow_fn_t fn = tk_flow_pick(k, p->desc.phys.arith, imm);
  if (!fn) return h2d_fail(H2D_EINVAL, "h2d_plan_run_dataflow: no kernel for k = %d", k);
  cudaStream_t cs = (cudaStream_t)stream;

  TkArgs a;
  memset(&a, 0, sizeof a);
  a.g = p->geo;
  a.dst = nullptr;   // per epoch: FlowArgs::dst
  a.tasks = nullptr;
  a.chunk_rows = R;
  a.row_base = 0;
  a.row_end = t.rows;
  a.H = H;
  a.wv = TK_W - 2 * H;
  a.n_strips = (t.cols + a.wv - 1) / a.wv;
  a.n_chunks = (t.rows + R - 1) / R;
  a.n_tasks = a.n_strips * a.n_chunks;
  a.fw = p->fw;
  unsigned* hdr = (unsigned*)p->scratch;
  a.edge_counter = hdr + 0;
  a.done_counter = hdr + 1;
  a.norm_out = (float*)(hdr + 2);
  a.fault = hdr + 3;
  a.stats = (unsigned long long*)(hdr + 8);
  a.cta_partials = (float*)((char*)p->scratch + 256);
  a.timeout_ns = p->desc.spin_timeout_ns > 0 ? p->desc.spin_timeout_ns : 10000000000LL;
  if ((int64_t)a.n_tasks * (int64_t)n_epochs >= (int64_t)0x7fffffff)
    return h2d_fail(H2D_EINVAL, "h2d_plan_run_dataflow: %d tasks x %d epochs overflow the ticket counter",
                    a.n_tasks, n_epochs);

  if (p->flow == nullptr || p->flow_tasks < (size_t)a.n_tasks) {
    // first use (or a deeper k with more strips): counters for every task
    if (p->flow) { H2D_CUDA_TRY(cudaStreamSynchronize(cs)); cudaFree(p->flow); p->flow = nullptr; }
    size_t mt = (size_t)a.n_tasks;
    if (mt < (size_t)p->max_ctas * TK_WARPS) mt = (size_t)p->max_ctas * TK_WARPS;
    if (cudaMalloc(&p->flow, (64 + mt) * sizeof(uint32_t)) != cudaSuccess)
      return h2d_fail(H2D_ENOMEM, "h2d_plan_run_dataflow: cannot allocate %zu task counters", mt);
    p->flow_tasks = mt;
    H2D_CUDA_TRY(cudaMemsetAsync(p->flow, 0, (64 + mt) * sizeof(uint32_t), cs));
    p->flow_base = 0;
  }
  if (p->flow_base > 0x70000000u) {   // keep the wrap-safe compare far from the wrap
    H2D_CUDA_TRY(cudaMemsetAsync(p->flow, 0, (64 + p->flow_tasks) * sizeof(uint32_t), cs));
    p->flow_base = 0;
  }
  H2D_CUDA_TRY(cudaFuncSetAttribute((const void*)fn, cudaFuncAttributeMaxDynamicSharedMemorySize, TK_SMEM_BYTES));
  const int64_t total = (int64_t)a.n_tasks * n_epochs;   // one warp per (epoch, task)
  const int grid = (int)((total + TK_WARPS - 1) / TK_WARPS);

  FlowArgs pa;
  pa.dst[0] = p->desc.buf[1 - src_parity];
  pa.dst[1] = p->desc.buf[src_parity];
  pa.ticket = p->flow;
  pa.done = p->flow + 64;
  pa.base = p->flow_base;
  pa.n_epochs = n_epochs;
  H2D_CUDA_TRY(cudaMemsetAsync(p->flow, 0, sizeof(uint32_t), cs));
  fn<<<grid, TK_WARPS * 32, TK_SMEM_BYTES, cs>>>(p->tmap[src_parity], p->tmap[1 - src_parity], a, pa);
  H2D_CHECK_LAUNCH("tk_flow_kernel");
  p->flow_base += (uint32_t)n_epochs;
  p->launches++;
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
  const int wv_ = TK_W - 2 * round_up4(p->desc.kmax);
  const int ns_ = (p->desc.tile.cols + wv_ - 1) / wv_;
  const double slots_ = (double)sm_count() * (p->desc.kmax <= 5 ? 16 : 12);
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
  int M = block_rows > 0 ? block_rows : env_int("H2D_PIPE_BLOCK_WAVES", 1, 64, 4) * M1;
  // Narrow tiles (column strips on many GPUs: 37 task strips at 4096 columns) reach one wave
  // of tasks only with thousands of rows; keep at least ~4 blocks so that upload, compute and
  // download still overlap, but never less than one full wave per window.
  if (block_rows <= 0 && qs[0].p->has_peers && M > rows / 4) M = M1 > rows / 4 ? M1 : rows / 4;
  if (M < S + 1) M = S + 1;

  // upload in units of U rows, one event per unit; the copy streams start after whatever
  // the caller queued on `stream`
  const int U = M1 < 1024 ? (M1 < rows ? M1 : rows) : 1024;
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
      const int kk = (e < L - 1) ? k : (int)(steps - (int64_t)(L - 1) * k);
      for (int i = 0; i < n; ++i)
        if (int rc = launch_window(qs[i].p, e & 1, kk, (int)lo, (int)hi, qs[i].R, cs, epoch)) return rc;
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
float* h2d_plan_norm_ptr(h2d_plan_t* p) { return p ? (float*)((unsigned*)p->scratch + 2) : nullptr; }
uint32_t* h2d_plan_fault_ptr(h2d_plan_t* p) { return p ? (uint32_t*)p->scratch + 3 : nullptr; }
uint64_t* h2d_plan_stats_ptr(h2d_plan_t* p) { return p ? (uint64_t*)((uint32_t*)p->scratch + 8) : nullptr; }

}  // extern "C"
