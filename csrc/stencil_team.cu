// stencil_team — the CLUSTER launch of the temporal-blocked stencil: the tile is cut into
// teams of adjacent strips, one team per thread-block cluster; the strips of a team share
// their edge columns through (distributed) shared memory instead of recomputing them
// (csrc/tk_team.cuh).  Single tile, no convergence norm; h2d_plan_launch_team() is the host
// entry.  Its own translation unit, like the dataflow launch: the flagship kernels are not
// recompiled when this changes.
#include <cuda.h>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include "tk_plan.h"
#include "tk_team.cuh"

namespace h2d {

constexpr int TEAM_SMEM_RING = tk_smem_ring(LAY_V4);
constexpr int TEAM_STAGE_BYTES = tk_stage_bytes(LAY_V4);

// One cluster = one (team, chunk) task; warp `w` of CTA `cr` is strip  cr * 4 + w  of the team.
template <int K, int ARITH, bool IMM, int CS>
__global__ void __cluster_dims__(CS, 1, 1) __launch_bounds__(TK_WARPS * 32, TkBounds<K, LAY_V4>::min_blocks)
tk_team_kernel(const __grid_constant__ CUtensorMap tmap, const __grid_constant__ TkArgs a,
               const __grid_constant__ TeamArgs ta) {
  extern __shared__ __align__(128) unsigned char smem[];
  constexpr bool STRICT = ARITH != H2D_ARITH_FP32;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const uint32_t ring = smem_u32(smem) + warp * (TK_STAGES * TEAM_STAGE_BYTES);
  // ring | mbarriers | exchange area, the latter aligned to its size (the launch reserves the slack)
  const uint32_t bars = smem_u32(smem) + TEAM_SMEM_RING + warp * TK_STAGES * 8;
  const uint32_t xch = (smem_u32(smem) + TEAM_SMEM_RING + TK_SMEM_BARS + TEAM_XBYTES - 1) & ~(uint32_t)(TEAM_XBYTES - 1);
  constexpr int cs = CS;
  // device-side run control and the fault word, as in tk_kernel: the whole grid takes the same exit
  if (a.ctl != nullptr && *reinterpret_cast<const volatile uint32_t*>(a.ctl) != a.run_if) return;
  if (*reinterpret_cast<const volatile uint32_t*>(a.fault) != 0u) return;
  const int cr = CS > 1 ? (int)cluster_ctarank() : 0;

  // exchange area: zero (the first row step reads it), then the arrive the first wait pairs with
  for (int i = threadIdx.x; i < TEAM_XBYTES / 16; i += TK_WARPS * 32)
    sts128(xch + i * 16, make_float4(0.f, 0.f, 0.f, 0.f));
  if (lane == 0) {
#pragma unroll
    for (int s = 0; s < TK_STAGES; ++s) mbar_init(bars + s * 8, 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  __syncwarp();
  team_arrive<CS>();

  const Geo& g = a.g;
  const int task = (int)blockIdx.x / cs;
  const int chunk = task / ta.n_teams, team = task - chunk * ta.n_teams;
  const int tw = cs * TK_WARPS, wpos = cr * TK_WARPS + warp;
  const int tcols = team_cols(tw, a.H);
  const int col0 = (team == ta.n_teams - 1) ? ((g.cols - tcols + 3) & ~3) : team * tcols;
  const int r_lo = chunk * ta.chunk_rows;
  const int r_hi = min(r_lo + ta.chunk_rows, g.rows);

  TeamCtx tc;
  tc.jbase = col0 - a.H + TEAM_STRIDE * wpos;
  tc.lo_lane = wpos == 0 ? a.H / 4 : 1;
  tc.hi_lane = wpos == tw - 1 ? 31 - a.H / 4 : 30;
  // slot of warp w is w + 1 (slots 0 and TK_WARPS + 1 stay zero: the outer edges of the team).
  // Lane 0 reads the left neighbour's lane 30, lane 31 the right neighbour's lane 1; for the
  // first / last warp of a CTA that neighbour lives in the next CTA of the cluster.
  int rd_cta = cr, rd_warp = warp, rd_lane = lane;
  if (lane == 0) { rd_warp = warp - 1; rd_lane = 30; }
  if (lane == 31) { rd_warp = warp + 1; rd_lane = 1; }
  if (rd_warp < 0 && cr > 0) { rd_cta = cr - 1; rd_warp = TK_WARPS - 1; }
  if (rd_warp >= TK_WARPS && cr + 1 < cs) { rd_cta = cr + 1; rd_warp = 0; }
  tc.xrd = xch + (rd_warp + 1) * TEAM_XROW + rd_lane * 16;
  tc.remote = rd_cta != cr;
  if (tc.remote) tc.xrd = cluster_map(tc.xrd, (uint32_t)rd_cta);
  tc.xwr = xch + (warp + 1) * TEAM_XROW + lane * 16;

  // the footprint reaches one row further down than an independent strip's (LAG = 2K + 1)
  const bool frozen = (g.x0 + r_lo - K < g.ax_lo) || (g.x0 + r_hi + K > g.ax_hi) ||
                      (g.y0 + tc.jbase < g.ay_lo) || (g.y0 + tc.jbase + Lay<LAY_V4>::W - 1 > g.ay_hi);
  float* dst = a.dst;
  uint32_t track = 0u;
  if (frozen) {
    run_task_team<K, ARITH, IMM, true, STRICT, CS>(&tmap, a, dst, tc, r_lo, r_hi, ring, bars, lane, track);
  } else {
    run_task_team<K, ARITH, IMM, false, STRICT, CS>(&tmap, a, dst, tc, r_lo, r_hi, ring, bars, lane, track);
  }
  if constexpr (STRICT) {
    // Data outside the domain of the fast evaluation (negative, non-finite, huge) anywhere in
    // the team: a one-CTA team runs the task again with the plain evaluation — all four warps,
    // they work in lock-step — and overwrites every store of the first run.  Larger clusters
    // only count the strips (scratch header word 15, h2d_plan_team_fallback_ptr): their caller
    // repeats the sweep with h2d_plan_launch, whose kernels fall back task by task.
    const bool bad = __any_sync(0xffffffffu, track >= STRICT_FAST_LIMIT);
    if (bad && lane == 0) atomicAdd(a.strict_fb + 3, 1u);   // header word 15
    if (CS == 1) {
      if (__syncthreads_or(bad ? 1 : 0)) {
        const int box0 = (r_hi - r_lo + 2 * K + 1 + TK_BR - 1) / TK_BR;
        run_task_team<K, ARITH, false, true, false, CS>(&tmap, a, dst, tc, r_lo, r_hi, ring, bars, lane, track, box0);
      }
    }
  }
}

typedef void (*tk_team_fn_t)(const CUtensorMap, const TkArgs, const TeamArgs);

template <int ARITH, bool IMM, int CS>
static tk_team_fn_t team_pick_k(int k) {
  switch (k) {
    case 5: return tk_team_kernel<5, ARITH, IMM, CS>;
    case 6: return tk_team_kernel<6, ARITH, IMM, CS>;
    case 7: return tk_team_kernel<7, ARITH, IMM, CS>;
    case 8: return tk_team_kernel<8, ARITH, IMM, CS>;
    default: return nullptr;
  }
}
template <int CS>
static tk_team_fn_t team_pick_cs(int k, int arith, bool imm) {
  switch (arith) {
    case H2D_ARITH_FP32:
      return imm ? team_pick_k<H2D_ARITH_FP32, true, CS>(k) : team_pick_k<H2D_ARITH_FP32, false, CS>(k);
    case H2D_ARITH_STRICT_FMA: return team_pick_k<H2D_ARITH_STRICT_FMA, false, CS>(k);
    case H2D_ARITH_STRICT: return CS == 1 ? team_pick_k<H2D_ARITH_STRICT, false, 1>(k) : nullptr;
    default: return nullptr;
  }
}
static tk_team_fn_t team_pick(int k, int arith, bool imm, int cluster) {
  switch (cluster) {
    case 1: return team_pick_cs<1>(k, arith, imm);
    case 2: return team_pick_cs<2>(k, arith, imm);
    case 4: return team_pick_cs<4>(k, arith, imm);
    default: return nullptr;
  }
}

}  // namespace h2d

using namespace h2d;

extern "C" {

int h2d_plan_launch_team_ex(h2d_plan_t* p, int src_parity, int k, int cluster, int chunk_rows,
                            const uint32_t* ctl, uint32_t run_if, void* stream);

int h2d_plan_launch_team(h2d_plan_t* p, int src_parity, int k, int cluster, int chunk_rows, void* stream) {
  return h2d_plan_launch_team_ex(p, src_parity, k, cluster, chunk_rows, nullptr, 0u, stream);
}

// Opt in to the dynamic shared memory size of the one-CTA team kernels a plan with
// h2d_plan_desc_t::team may launch — which also loads them, so that the first timed sweep
// does not pay for it (lazy module loading cost 1.8 ms of a 16 ms run).
void h2d_plan_team_prepare(const h2d_plan_t* p) {
  if (!p) return;
  const bool imm = coef_is_reference(p->geo, p->desc.phys.arith);
  for (int k = 5; k <= p->desc.kmax && k <= TK_KMAX; ++k) {
    tk_team_fn_t fn = team_pick(k, p->desc.phys.arith, imm, 1);
    if (fn) cudaFuncSetAttribute((const void*)fn, cudaFuncAttributeMaxDynamicSharedMemorySize, tk_team_smem_bytes());
  }
}

int h2d_plan_team_eligible(const h2d_plan_t* p, int k, int cluster) {
  if (!p || p->has_peers || p->lay != LAY_V4 || k < 5 || k > p->desc.kmax || k > TK_KMAX) return 0;
  if (p->desc.phys.arith != H2D_ARITH_FP32 && p->desc.phys.arith != H2D_ARITH_STRICT_FMA &&
      p->desc.phys.arith != H2D_ARITH_STRICT) return 0;
  return p->desc.tile.cols >= team_cols(cluster * TK_WARPS, round_up4(k));
}

int h2d_plan_launch_team_ex(h2d_plan_t* p, int src_parity, int k, int cluster, int chunk_rows,
                            const uint32_t* ctl, uint32_t run_if, void* stream) {
  if (!p) return h2d_fail(H2D_EINVAL, "h2d_plan_launch_team: null plan");
  if (src_parity != 0 && src_parity != 1) return h2d_fail(H2D_EINVAL, "h2d_plan_launch_team: bad parity");
  if (k < 5 || k > p->desc.kmax || k > TK_KMAX)
    return h2d_fail(H2D_EINVAL, "h2d_plan_launch_team: k = %d outside [5, %d] (shallower sweeps recompute one lane "
                    "per side already)", k, p->desc.kmax < TK_KMAX ? p->desc.kmax : TK_KMAX);
  if (cluster != 1 && cluster != 2 && cluster != 4)
    return h2d_fail(H2D_EINVAL, "h2d_plan_launch_team: cluster size %d (1, 2 or 4)", cluster);
  if (p->has_peers) return h2d_fail(H2D_EINVAL, "h2d_plan_launch_team: single-tile plans only");
  if (p->lay != LAY_V4) return h2d_fail(H2D_EINVAL, "h2d_plan_launch_team: the plan must use the v4 lane layout");
  const h2d_tile_t& t = p->desc.tile;
  const int H = round_up4(k);
  const int tw = cluster * TK_WARPS;
  const int tcols = team_cols(tw, H);
  if (t.cols < tcols)
    return h2d_fail(H2D_EINVAL, "h2d_plan_launch_team: tile of %d columns is narrower than one team (%d)", t.cols, tcols);
  const bool imm = coef_is_reference(p->geo, p->desc.phys.arith);
  tk_team_fn_t fn = team_pick(k, p->desc.phys.arith, imm, cluster);
  if (!fn) return h2d_fail(H2D_EINVAL, "h2d_plan_launch_team: no kernel for k = %d, arith %d", k, p->desc.phys.arith);
  cudaStream_t cs = (cudaStream_t)stream;

  TeamArgs ta;
  ta.n_teams = (t.cols + tcols - 1) / tcols;
  // chunk height: whole waves of clusters, the pipeline fill (2k + 9 row steps) weighed against
  // the last, partly empty wave — the rule of choose_chunk_rows with teams for strips
  int R = chunk_rows;
  int& cached = p->team_rows[k - 5][cluster == 1 ? 0 : (cluster == 2 ? 1 : 2)];
  if (R <= 0) R = cached;
  if (R <= 0) {
    int dev = 0, sms = 148;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    const double slots = (double)sms * tk_min_blocks(k, LAY_V4) / cluster;   // clusters in flight
    double best = 1e300;
    R = t.rows;
    for (int c = 1; c <= t.rows; ++c) {
      const int r = (t.rows + c - 1) / c;
      if (r < 8 && c > 1) break;
      const double waves = ceil((double)((t.rows + r - 1) / r) * ta.n_teams / slots);
      // (+0.5, not the +0.25 of independent strips: a team is as slow as its slowest strip, and
      // the strips on the global boundary run the select-heavy body; 16384^2 strict-fma:
      // 432-row tasks in 3 waves 1209 GLUP/s, 261-row tasks in 5 waves 1296)
      const double cost = (waves + 0.5) * (double)(r + 2 * k + 9);
      if (cost < best * (1.0 - 1e-9)) { best = cost; R = r; }
    }
    cached = R;   // the search walks up to `rows` candidates: once per plan, k and cluster size
  }
  ta.chunk_rows = R;
  ta.n_chunks = (t.rows + R - 1) / R;

  TkArgs a;
  memset(&a, 0, sizeof a);
  a.g = p->geo;
  a.dst = p->desc.buf[1 - src_parity];
  a.src = p->desc.buf[src_parity];
  a.row_base = 0;
  a.row_end = t.rows;
  a.H = H;
  a.wv = TEAM_STRIDE;
  a.fw = p->fw;
  tk_fill_scratch_ptrs(a, p->scratch);
  a.timeout_ns = p->desc.spin_timeout_ns > 0 ? p->desc.spin_timeout_ns : 10000000000LL;
  a.ctl = ctl;
  a.run_if = run_if;

  const int smem = tk_team_smem_bytes();
  H2D_CUDA_TRY(cudaFuncSetAttribute((const void*)fn, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
  fn<<<(unsigned)(ta.n_teams * ta.n_chunks * cluster), TK_WARPS * 32, (size_t)smem, cs>>>(p->tmap[src_parity], a, ta);
  H2D_CHECK_LAUNCH("tk_team_kernel");
  p->launches++;
  p->seq++;
  return H2D_OK;
}

}  // extern "C"
