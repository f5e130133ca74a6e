// stencil_flow — the opt-in DATAFLOW launch of the temporal-blocked stencil: the same tasks
// as stencil_tk.cu, but ONE launch runs `n_epochs` k-step sweeps.  Measured 5 % slower than
// one launch per sweep on B200 (profiles/dataflow_probe_b200.json), so it lives in its own
// translation unit, away from the flagship kernels; kept because it is the only way to
// avoid a drained machine between sweeps on grids of a few waves.
#include <cuda.h>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include "tk_plan.h"

namespace h2d {

constexpr int FLOW_W = Lay<LAY_V4>::W;
constexpr int FLOW_STAGE_BYTES = TK_BR * FLOW_W * 4;
constexpr int FLOW_SMEM_RING = TK_WARPS * TK_STAGES * FLOW_STAGE_BYTES;
constexpr int FLOW_SMEM_BYTES = FLOW_SMEM_RING + TK_SMEM_BARS;

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
__global__ void __launch_bounds__(TK_WARPS * 32, TkBounds<K, LAY_V4>::min_blocks)
tk_flow_kernel(const __grid_constant__ CUtensorMap tmap0, const __grid_constant__ CUtensorMap tmap1,
               const __grid_constant__ TkArgs a, const __grid_constant__ FlowArgs pa) {
  extern __shared__ __align__(128) unsigned char smem[];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const uint32_t ring = smem_u32(smem) + warp * (TK_STAGES * FLOW_STAGE_BYTES);
  const uint32_t bars = smem_u32(smem) + FLOW_SMEM_RING + warp * TK_STAGES * 8;
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
                      (g.y0 + jb < g.ay_lo) || (g.y0 + jb + FLOW_W - 1 > g.ay_hi);
  const CUtensorMap* tm = (e & 1) ? &tmap1 : &tmap0;
  float* dst = pa.dst[e & 1];
  uint32_t track = 0u;
  if (frozen) {
    run_task_v4<K, ARITH, IMM, true, false, false>(tm, a, dst, strip, r_lo, r_hi, ring, bars, lane, track);
  } else {
    run_task_v4<K, ARITH, IMM, false, false, false>(tm, a, dst, strip, r_lo, r_hi, ring, bars, lane, track);
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

}  // namespace h2d

using namespace h2d;

extern "C" {

int h2d_plan_run_dataflow(h2d_plan_t* p, int src_parity, int k, int n_epochs, void* stream) {
  if (!p) return h2d_fail(H2D_EINVAL, "h2d_plan_run_dataflow: null plan");
  if (src_parity != 0 && src_parity != 1) return h2d_fail(H2D_EINVAL, "h2d_plan_run_dataflow: bad parity");
  if (k < 1 || k > p->desc.kmax)
    return h2d_fail(H2D_EINVAL, "h2d_plan_run_dataflow: k = %d outside [1, kmax = %d]", k, p->desc.kmax);
  if (n_epochs < 0) return h2d_fail(H2D_EINVAL, "h2d_plan_run_dataflow: negative epoch count");
  if (p->has_peers) return h2d_fail(H2D_EINVAL, "h2d_plan_run_dataflow: single-tile plans only");
  if (p->lay != LAY_V4)
    return h2d_fail(H2D_EINVAL, "h2d_plan_run_dataflow: the plan must use the v4 lane layout (h2d_plan_desc_t::layout = 1)");
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
  a.wv = FLOW_W - 2 * H;
  a.n_strips = (t.cols + a.wv - 1) / a.wv;
  a.n_chunks = (t.rows + R - 1) / R;
  a.n_tasks = a.n_strips * a.n_chunks;
  a.n_tasks_all = a.n_tasks;
  a.fw = p->fw;
  tk_fill_scratch_ptrs(a, p->scratch);
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
  H2D_CUDA_TRY(cudaFuncSetAttribute((const void*)fn, cudaFuncAttributeMaxDynamicSharedMemorySize, FLOW_SMEM_BYTES));
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
  fn<<<grid, TK_WARPS * 32, FLOW_SMEM_BYTES, cs>>>(p->tmap[src_parity], p->tmap[1 - src_parity], a, pa);
  H2D_CHECK_LAUNCH("tk_flow_kernel");
  p->flow_base += (uint32_t)n_epochs;
  p->launches++;
  return H2D_OK;
}

}  // extern "C"
