// tk_team.cuh — the temporal-blocked task body for a TEAM of adjacent strips: the warps of a
// thread-block cluster (4 per CTA) work on neighbouring 128-column strips of the same rows in
// row lock-step and hand each other their edge columns half-way through the K levels, through
// shared memory inside a CTA and through DISTRIBUTED shared memory between the CTAs of the
// cluster.  What it buys: an independent strip (run_task_v4) recomputes H = 8 columns on each
// side at K = 8, 16 of every 128; inside a team the invalid region is refreshed after K/2
// levels, when it is exactly one lane (4 columns) wide, so neighbouring strips overlap by 8
// columns instead of 16 and a strip stores 120 columns instead of 112.  Only the two outer
// edges of the team keep the full margin.
//
// The reference's kernel has no reuse at all (one thread per point, five global loads,
// grad1612_cuda_heat.cu:55-62) and its report gave up on shared-memory tiling (Report.pdf p20).
//
// Pipeline.  run_task_v4 chains the K levels inside one row step: level t+1 consumes, as its
// incoming ("down") row, what level t produced a few instructions earlier.  Here the chain is
// cut between levels M-1 and M (M = K/2): the lower half publishes its row in the exchange
// area, and the upper half picks up the row published in the PREVIOUS row step — lane 0 from
// the left neighbour's lane 30, lane 31 from the right neighbour's lane 1, the other lanes
// their own — so the upper half lags one row step more (LAG = 2K + 1).  One split cluster
// barrier per row step orders the exchange; row step i does
//     wait (phase i-1) -> read row i-1 -> levels 0..M-1 -> write row i -> arrive (phase i) -> levels M..K-1
// with the two parities of the exchange area alternating.  A warp writes only its own CTA's
// shared memory; the two edge lanes of a CTA READ the neighbour CTA's (ld.shared::cluster).
// The barrier is used in its relaxed form: with release / acquire semantics nvcc brackets it
// with MEMBAR.ALL.GPU + CCTL.IVALL, which would drain the kernel's global stores on every row
// step.  Ordering instead rests on (a) the arrive being issued after the first instruction that
// consumes the row just read, i.e. after the read has returned — nobody overwrites a row that
// is still being read —
// and (b) a CTA-scope fence between a warp's shared-memory store and its arrive.
#pragma once
#include "tk_task.cuh"

namespace h2d {

constexpr int TEAM_STRIDE = Lay<LAY_V4>::W - 8;          // column distance of adjacent strips of a team
constexpr int TEAM_SLOTS = TK_WARPS + 2;                 // exchange rows of a CTA: left ghost, its warps, right ghost
constexpr int TEAM_XROW = 32 * 16;                       // one float4 per lane
constexpr int TEAM_XPAR = 4096;                          // one parity (>= TEAM_SLOTS rows): a power of two, so that
constexpr int TEAM_XBYTES = 2 * TEAM_XPAR;               // flipping one address bit switches parity (area aligned to 8 KB)
static_assert(TK_BR % 2 == 0, "a run ends on exchange parity 0");
static_assert(TEAM_SLOTS * TEAM_XROW <= TEAM_XPAR, "exchange rows of one parity");

constexpr int TEAM_MAX_CLUSTER = 4;

// Columns a team of `warps` strips stores (H = column margin of the OUTER edges).
__host__ __device__ constexpr int team_cols(int warps, int H) { return TEAM_STRIDE * warps + 8 - 2 * H; }
// shared memory of a CTA: TMA ring | mbarriers | slack | exchange area (aligned to its size by the kernel)
__host__ __device__ constexpr int tk_team_smem_bytes() { return tk_smem_ring(LAY_V4) + TK_SMEM_BARS + 2 * TEAM_XBYTES; }

struct TeamArgs {
  int32_t n_teams;      // teams side by side across the tile (the last one is shifted left to end at the edge)
  int32_t chunk_rows, n_chunks;
};

struct TeamCtx {
  int jbase;            // tile-local column of lane 0, element 0
  int lo_lane, hi_lane; // lanes whose 4 columns are stored
  uint32_t xrd;         // exchange address this lane reads (parity 0): shared::cta, or shared::cluster if `remote`
  uint32_t xwr;         // exchange address this lane writes (parity 0)
  bool remote;          // this lane reads the neighbour CTA's shared memory
};

// CS = CTAs per cluster, a compile-time constant (__cluster_dims__): without it nvcc emits both
// a cluster and a CTA barrier behind a run-time test of the cluster size.  CS == 1: a team is
// one CTA and the hand-over needs a single bar.sync where the others wait.
template <int CS>
__device__ __forceinline__ void team_arrive() {
  // membar.cta: this thread's shared-memory store has been performed before the arrive can be
  // seen (without it a 4-CTA cluster read a stale row once in a 32768^2 sweep); the .release
  // form of the arrive would add MEMBAR.ALL.GPU, i.e. drain the global stores of every row step
  if (CS > 1) asm volatile("fence.acq_rel.cta;\n\tbarrier.cluster.arrive.relaxed.aligned;" ::: "memory");
}
template <int CS>
__device__ __forceinline__ void team_wait() {
  if (CS > 1) {
    asm volatile("barrier.cluster.wait.aligned;" ::: "memory");
  } else {
    asm volatile("bar.sync 0;" ::: "memory");
  }
}
__device__ __forceinline__ uint32_t cluster_ctarank() {
  uint32_t r;
  asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
  return r;
}
__device__ __forceinline__ uint32_t cluster_nctarank() {
  uint32_t r;
  asm volatile("mov.u32 %0, %%cluster_nctarank;" : "=r"(r));
  return r;
}
__device__ __forceinline__ uint32_t cluster_map(uint32_t local_addr, uint32_t rank) {
  uint32_t r;
  asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(r) : "r"(local_addr), "r"(rank));
  return r;
}
__device__ __forceinline__ void sts128(uint32_t addr, const float4& v) {
  asm volatile("st.shared.v4.f32 [%0], {%1, %2, %3, %4};" ::"r"(addr), "f"(v.x), "f"(v.y), "f"(v.z), "f"(v.w)
               : "memory");
}
__device__ __forceinline__ float4 ld_cluster128(uint32_t addr) {
  float4 v;
  asm volatile("ld.shared::cluster.v4.f32 {%0, %1, %2, %3}, [%4];"
               : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w)
               : "r"(addr)
               : "memory");
  return v;
}

// One strip of a team.  The caller has zeroed the exchange area and ARRIVED once at the
// cluster barrier (so that the first row step finds a phase to wait for); on return every
// arrive has been waited for.  All warps of the cluster run the same number of row steps.
template <int K, int ARITH, bool IMM, bool FROZEN, bool FAST, int CS>
__device__ __forceinline__ void run_task_team(const CUtensorMap* tmap, const TkArgs& a, float* dst_base,
                                              const TeamCtx& tc, int r_lo, int r_hi, uint32_t ring, uint32_t bars,
                                              int lane, uint32_t& track, int box0 = 0) {
  static_assert(K >= 2, "a team needs two levels to split");
  constexpr bool TRACK = ARITH != H2D_ARITH_FP32;
  constexpr int M = K / 2;                 // levels before the hand-over
  constexpr int W = Lay<LAY_V4>::W;
  constexpr int STAGE_BYTES = TK_BR * W * 4;
  const Geo& g = a.g;
  const int n_out = r_hi - r_lo;
  const int jbase = tc.jbase;
  const int j0 = jbase + lane * 4;
  const int row0 = r_lo - K;
  constexpr int LAG = 2 * K + 1;           // the upper half runs one row step behind
  const int n_in = n_out + LAG;
  const int n_blk = (n_in + TK_BR - 1) / TK_BR;
  const int tc0 = g.halo_c + jbase;
  const int tc1 = g.halo_r + row0;

  const bool lane_in = lane >= tc.lo_lane && lane <= tc.hi_lane;
  int smask = 0;
#pragma unroll
  for (int e = 0; e < 4; ++e)
    if (lane_in && (j0 + e) >= 0 && (j0 + e) < g.cols) smask |= 1 << e;
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

  // box0: boxes this warp's ring has already carried (a task that is run a second time continues
  // the stage / phase sequence of its mbarriers; the number of row steps of a run is a multiple
  // of TK_BR = 6, so the exchange parity starts over at 0 as well)
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

  float4 st[K][3];
#pragma unroll
  for (int t = 0; t < K; ++t) st[t][0] = st[t][1] = st[t][2] = make_float4(0.f, 0.f, 0.f, 0.f);

  uint32_t trk = 0u;
  int ro = -LAG;
  float* q = dst_base + (int64_t)(g.halo_r + r_lo - LAG) * g.pitch + g.halo_c + j0;

  // Row step i writes parity i & 1 of the exchange area and reads the other one.  The parity is
  // a run-time address bit, not a template argument: unrolling over 3 rotation phases x 2
  // parities doubles the loop body, and the strict body then (2 450 instructions, 39 KB) no
  // longer fits the instruction cache (ncu: no_instruction 0.04 -> 0.88 warps per issue).
  uint32_t xw = tc.xwr, xr = tc.xrd ^ (uint32_t)TEAM_XPAR;

  // One row step; PH = its phase in the 3-slot rotation.
  auto row_step = [&](auto Ptag) {
    constexpr int PH = decltype(Ptag)::value;
    constexpr int iu = PH, ic = (PH + 1) % 3, id = (PH + 2) % 3;
    float4 fin;
    if (TRACK) trk = umax4(trk, st[0][id]);
    // everybody's previous row is in the other parity
    team_wait<CS>();
    const float4 nx = (CS > 1 && tc.remote) ? ld_cluster128(xr) : lds128(xr);
    xr ^= (uint32_t)TEAM_XPAR;
#pragma unroll
    for (int t = 0; t < K; ++t) {
      const float4 up = st[t][iu];
      const float4 c = st[t][ic];
      const float4 down = st[t][id];
      const float L = __shfl_up_sync(0xffffffffu, c.w, 1);
      const float Rn = __shfl_down_sync(0xffffffffu, c.x, 1);
      float4 o = point4x<ARITH, IMM, FAST>(up, c, down, L, Rn, g);
      if (FROZEN) {
        // tile-local row of this level's output (the lower half is one row step ahead)
        const int gx = g.x0 + r_lo + ro + (K - 1 - t) + (t < M ? 1 : 0);
        const bool ract = gx >= g.ax_lo && gx <= g.ax_hi;
        o.x = (ract && act[0]) ? o.x : c.x;
        o.y = (ract && act[1]) ? o.y : c.y;
        o.z = (ract && act[2]) ? o.z : c.z;
        o.w = (ract && act[3]) ? o.w : c.w;
      }
      if (TRACK) trk = umax4(trk, o);
      // level M has just consumed the row read at the top of this step: in-order issue puts the
      // arrive behind the return of that read (checked in the SASS: UCGABAR_ARV follows the
      // first FADD2 on the LDS / LD destination registers)
      if (t == M) team_arrive<CS>();
      if (t == M - 1) {
        // hand-over
        sts128(xw, o);
        xw ^= (uint32_t)TEAM_XPAR;
        st[M][id] = nx;
      } else if (t + 1 < K) {
        st[t + 1][id] = o;
      } else {
        fin = o;
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
    __syncwarp();
    if (lane == 0 && b + TK_STAGES < n_blk) issue(b + TK_STAGES);
  }
  if (CS > 1) team_wait<CS>();   // the last arrive: nobody leaves while a neighbour may still read its shared memory
  track = trk;
}

}  // namespace h2d
