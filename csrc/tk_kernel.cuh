// tk_kernel.cuh — the __global__ wrapper around the task bodies of tk_task.cuh: run
// control, task enumeration, the epoch-flag protocol of the fused halo exchange, the
// strict-mode fast path / fallback, and the deterministic norm reduction.
//
// The kernel templates are instantiated in tk_inst.cu, compiled once per
// (arithmetic, lane layout, XTR) combination so that the instantiations build in parallel.
#pragma once
#include "tk_task.cuh"

namespace h2d {

// Resident CTAs per SM the kernels are compiled for (__launch_bounds__), also used by the
// host to size tasks (stencil_tk.cu: warps_per_sm).
// V4 / AB2 (12 state registers per level): K <= 5 fits 128 registers (4 CTAs = 16 warps
// per SM); deeper pipelines get 168 registers (3 CTAs = 12 warps) instead of spilling.
// Measured on B200 at K = 8, V4: 12 warps / 166 regs / no spills 4.43 TLUP/s; 13-14 warps
// with __maxnreg__ 152 / 144 (32-72 B of spills) 4.31 TLUP/s; 16 warps at 128 registers
// (244 B of spills) 3.9 TLUP/s.
// AB4 (24 state registers per level): 2 CTAs = 8 warps per SM at K >= 5 (shared memory
// allows no more either: 112 KB per CTA), 3 below.  V8 (24 per level, K <= 4): 3 CTAs.
__host__ __device__ constexpr int tk_min_blocks(int k, int lay) {
#ifdef H2D_TK_MINBLOCKS
  return H2D_TK_MINBLOCKS;   // experiments only
#else
  return lay == LAY_AB4 ? ((k <= 4) ? 3 : 2) : (lay == LAY_V8 ? 3 : ((k <= 5) ? 4 : 3));
#endif
}
template <int K, int LAY>
struct TkBounds {
  static constexpr int min_blocks = tk_min_blocks(K, LAY);
};

// XTR = false: a launch with neither peers nor a convergence norm (the common single-GPU
// launch): the kernel contains no forwarding / reduction code at all, so that code cannot
// cost the hot loop registers.
template <int K, int ARITH, int LAY, bool IMM, bool XTR>
__global__ void __launch_bounds__(TK_WARPS * 32, TkBounds<K, LAY>::min_blocks)
tk_kernel(const __grid_constant__ CUtensorMap tmap, const __grid_constant__ TkArgs a) {
  extern __shared__ __align__(128) unsigned char smem[];
  __shared__ float cta_red[TK_WARPS];
  __shared__ bool cta_last;
  constexpr int W = Lay<LAY>::W;
  constexpr int STAGE_BYTES = TK_BR * W * 4;
  constexpr int STAGES = tk_stages(LAY);
  constexpr int SMEM_RING = TK_WARPS * STAGES * STAGE_BYTES;
  constexpr bool STRICT = ARITH != H2D_ARITH_FP32;
  // Converged earlier (or a conditional launch whose condition did not occur):
  // every CTA of every rank reads the same word and leaves; no flag is
  // published and none is awaited, so the ranks stay consistent.
  if (a.ctl != nullptr && *reinterpret_cast<const volatile uint32_t*>(a.ctl) != a.run_if) return;
  // A flag wait timed out in an earlier launch (dead or stalled peer): the queued launches
  // drain at once instead of each spinning through the watchdog again.
  if (*reinterpret_cast<const volatile uint32_t*>(a.fault) != 0u) return;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int task = (int)blockIdx.x * TK_WARPS + warp;
  const uint32_t ring = smem_u32(smem) + warp * (STAGES * STAGE_BYTES);
  const uint32_t bars = smem_u32(smem) + SMEM_RING + warp * TK_STAGES * 8;
  if (STRICT && blockIdx.x == 0 && threadIdx.x == 0) a.strict_fb[(a.seq + 1u) % 3u] = 0u;

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
      // tasks of the strips that take part in the flag protocol — the first and the last
      // (they forward the 32-column bands) and, when the last strip owns fewer than k columns,
      // the one before it (its results depend on the E ghost columns) — come first, in short
      // chunks, then the interior strips.
      const int ns = a.n_strips, ne = a.n_edge_strips, n_edge = ne * a.n_chunks_edge;
      if (task < n_edge) {
        const int i = task % ne;
        const int s = i == 0 ? 0 : (i == 1 ? ns - 1 : ns - 2);
        tk = make_int2((int)(((unsigned)s) | 0xc0000000u), task / ne);
      } else {
        const int t2 = task - n_edge;
        tk = make_int2(1 + t2 % (ns - ne), t2 / (ns - ne));
      }
    } else {
      tk = make_int2(task % a.n_strips, task / a.n_strips);
    }
    const int strip = tk.x & 0x3fffffff;
    const bool comm_edge = XTR && (tk.x < 0) && a.has_peers;
    const int R = (tk.x & 0x40000000) ? a.chunk_rows_edge : a.chunk_rows;
    const int r_lo = a.row_base + tk.y * R;
    const int r_hi = min(r_lo + R, a.row_end);

    if (comm_edge && (a.debug & 2)) {
      // Stress mode (race detection): shake the order in which ranks reach
      // the flag protocol with a pseudo-random delay of up to ~64 us per edge task.
      const uint32_t h = ((uint32_t)task * 2654435761u) ^ (a.epoch * 40503u) ^ ((uint32_t)a.g.x0 * 97u + (uint32_t)a.g.y0);
      const uint64_t t0 = globaltimer_ns();
      while (globaltimer_ns() - t0 < (uint64_t)((h >> 7) & 0xffffu)) __nanosleep(256);
    }
    bool alive = true;
    if (comm_edge) {
      // Ghost cells of src were written by the neighbours' launch `epoch-1`;
      // their ghost cells of OUR dst parity were last read by that launch too.
      int ok = 1;
      if (lane == 0) {
        for (int d = 0; d < H2D_NDIR; ++d) {
          if (a.peer[d].buf == nullptr) continue;
          // fast path: the neighbour published long ago (its edge tasks run first)
          if ((int32_t)(ld_acquire_sys(a.flags_in + d) - (a.epoch - 1u)) >= 0) continue;
          const uint64_t t0 = globaltimer_ns();
          if (!wait_flag_ge(a.flags_in + d, a.epoch - 1u, a.timeout_ns, a.fault)) {
            atomicCAS(a.fault, 0u, 1u + d);   // the first cause stays
            ok = 0;
          }
          atomicAdd(a.stats, (unsigned long long)(globaltimer_ns() - t0));   // the mpiP "Waitall" row
          atomicAdd(a.stats + 1, 1ull);
          if (!ok) break;
        }
        asm volatile("fence.proxy.async;" ::: "memory");
      }
      // After a timeout this task neither computes on stale ghosts nor counts itself in:
      // the epoch of this launch is never published, so the neighbours stop as well.
      alive = __shfl_sync(0xffffffffu, ok, 0) != 0;
    }

    if (alive) {
      // Does the footprint (all K levels) stay inside the updated region?
      const Geo& g = a.g;
      const int jb = ((strip == a.n_strips - 1 && g.cols >= a.wv) ? ((g.cols - a.wv + 3) & ~3) : strip * a.wv) - a.H;
      const bool frozen = (g.x0 + r_lo - K < g.ax_lo) || (g.x0 + r_hi - 1 + K > g.ax_hi) ||
                          (g.y0 + jb < g.ay_lo) || (g.y0 + jb + W - 1 > g.ay_hi);
      if (XTR && comm_edge && (a.colbuf_in[0] != nullptr || a.colbuf_in[1] != nullptr)) {
        // ghost columns arrive packed: put the rows this task reads into the source array's margins
        const bool west = jb < 0, east = jb + W > g.cols;
        if (west || east) unpack_colbuf(a, west, east, r_lo - K, r_hi + K, lane);
      }
      uint32_t track = 0u;
      bool fast = STRICT && !(a.debug & 4);
      if (STRICT && fast) {
        // most tasks of the previous launch were outside the fast path's domain: skip the try
        const unsigned prev = *reinterpret_cast<const volatile unsigned*>(a.strict_fb + (a.seq + 2u) % 3u);
        if ((int64_t)prev * 4 > (int64_t)a.n_tasks_all) fast = false;
      }
      // Strict modes: try the fast evaluation unless the previous launch said most tasks are
      // outside its domain; a task whose data is (negative, non-finite or huge anywhere in its
      // pipeline) is REDONE with the plain evaluation below — it overwrites every store before
      // anything is forwarded or published.
      bool need_slow = STRICT && !fast;
      const bool with_norm = XTR && a.norm != H2D_NORM_NONE;
      if (need_slow) {
      } else if (with_norm && frozen) {
        acc = run_task<K, ARITH, LAY, IMM, true, XTR, STRICT>(&tmap, a, a.dst, strip, r_lo, r_hi, ring, bars, lane, track);
      } else if (with_norm) {
        acc = run_task<K, ARITH, LAY, IMM, false, XTR, STRICT>(&tmap, a, a.dst, strip, r_lo, r_hi, ring, bars, lane, track);
      } else if (frozen) {
        acc = run_task<K, ARITH, LAY, IMM, true, false, STRICT>(&tmap, a, a.dst, strip, r_lo, r_hi, ring, bars, lane, track);
      } else {
        acc = run_task<K, ARITH, LAY, IMM, false, false, STRICT>(&tmap, a, a.dst, strip, r_lo, r_hi, ring, bars, lane, track);
      }
      if constexpr (STRICT) {
        int box0 = 0;
        if (!need_slow && __any_sync(0xffffffffu, track >= STRICT_FAST_LIMIT)) {
          need_slow = true;
          // the ring's mbarriers have completed one phase per box: the second run continues there
          box0 = (r_hi - r_lo + 2 * K + TK_BR - 1) / TK_BR;
          __syncwarp();
        }
        if (need_slow) {
          // one generic instantiation (boundary selects and norm always compiled in): rare path
          if constexpr (LAY == LAY_V8) {
            acc = run_task_v8<K, ARITH, false, true, XTR, false>(&tmap, a, a.dst, strip, r_lo, r_hi, ring, bars, lane,
                                                                 track, box0);
          } else {
            acc = run_task_v4<K, ARITH, false, true, XTR, false>(&tmap, a, a.dst, strip, r_lo, r_hi, ring, bars, lane,
                                                                 track, box0);
          }
          // feeds the next launch's decision, whichever way this task got here
          const bool bad = __any_sync(0xffffffffu, track >= STRICT_FAST_LIMIT);
          if (bad && lane == 0) atomicAdd(a.strict_fb + a.seq % 3u, 1u);
        }
      }

      if (comm_edge && !(a.debug & 1)) {
        // every lane's stores of this chunk are ordered before the copies that read them back
        __syncwarp();
        forward_task(a, strip, r_lo, r_hi, lane, ring, STAGES * STAGE_BYTES);
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
  }

  if (!XTR || a.norm == H2D_NORM_NONE) return;
  // Deterministic two-level reduction: warp -> CTA -> last CTA.
  const bool l2 = a.norm == H2D_NORM_L2SQ;
  acc = l2 ? warp_sum(acc) : warp_max(acc);
  if (lane == 0) cta_red[warp] = acc;
  __syncthreads();
  if (threadIdx.x == 0) {
    float v = cta_red[0];
    for (int w = 1; w < TK_WARPS; ++w) v = l2 ? v + cta_red[w] : fmax_nan(v, cta_red[w]);
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
    v = l2 ? v + x : fmax_nan(v, x);
  }
  v = l2 ? warp_sum(v) : warp_max(v);
  __syncthreads();
  if (lane == 0) cta_red[warp] = v;
  __syncthreads();
  if (threadIdx.x == 0) {
    float t = cta_red[0];
    for (int w = 1; w < TK_WARPS; ++w) t = l2 ? t + cta_red[w] : fmax_nan(t, cta_red[w]);
    *a.norm_out = t;
    *a.done_counter = 0u;
  }
}

typedef void (*tk_fn_t)(const CUtensorMap, const TkArgs);

// Defined by the instantiation units (tk_inst.cu, one object per combination); nullptr
// for a combination that is not built.
template <int ARITH, int LAY, bool XTR> tk_fn_t tk_get(int k, bool imm);

}  // namespace h2d
