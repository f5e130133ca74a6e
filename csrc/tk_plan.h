// tk_plan.h — host-side plan object shared by stencil_tk.cu (launches, pipelined solve)
// and stencil_flow.cu (the opt-in dataflow launch).
#pragma once
#include <vector>
#include "tk_kernel.cuh"

struct TaskList {
  int2* dev = nullptr;
  int n_tasks = 0, n_strips = 0, n_comm_edge = 0;
};

struct h2d_plan {
  h2d_plan_desc_t desc;
  h2d::Geo geo;
  int lay = h2d::LAY_V4;            // lane layout of this plan's kernels (tk_task.cuh)
  CUtensorMap tmap[2];
  int n_chunks = 0, chunk_rows = 0, fw = 0;
  int chunk_rows_edge = 0;          // chunk size of the strips that forward column bands (== chunk_rows without W/E peers)
  TaskList lists[h2d::TK_KMAX / 4 + 1];  // index H/4 (H = 4, 8)
  void* scratch = nullptr;          // counters + partials
  bool own_scratch = false;
  int max_ctas = 0;
  int64_t launches = 0;             // kernels launched (bench bookkeeping)
  uint32_t seq = 0;                 // logical launches (one k-step advance of the tile = one)
  // fork / join of a launch with neighbours: interior tasks run the plain kernel on `s_side`
  cudaStream_t s_side = nullptr;
  cudaEvent_t ev_fork = nullptr, ev_join = nullptr;
  int device = 0;
  bool has_peers = false;
  // task height of team launches, [k - 5][cluster 1 / 2 / 4]; 0 = not chosen yet (stencil_team.cu)
  int team_rows[4][3] = {{0, 0, 0}, {0, 0, 0}, {0, 0, 0}, {0, 0, 0}};
  // lazily created by h2d_plan_run_dataflow: [0] ticket, [64...] done[] per task
  uint32_t* flow = nullptr;
  size_t flow_tasks = 0;
  uint32_t flow_base = 0;
  // lazily created by h2d_plan_solve_pipelined
  cudaStream_t s_h2d = nullptr, s_d2h = nullptr;
  std::vector<cudaEvent_t> ev_pool;
};

namespace h2d {

inline int round_up4(int k) { return (k + 3) & ~3; }

// Scratch header (256 bytes, 32-bit words): 0 edge counter, 1 done counter, 2 norm value,
// 3 fault word, 8-11 two 64-bit wait statistics, 12-14 strict-mode fallback counters,
// 15 strips of cluster launches (stencil_team.cu) whose data left the fast strict evaluation.
inline void tk_fill_scratch_ptrs(TkArgs& a, void* scratch) {
  unsigned* hdr = (unsigned*)scratch;
  a.edge_counter = hdr + 0;
  a.done_counter = hdr + 1;
  a.norm_out = (float*)(hdr + 2);
  a.fault = hdr + 3;
  a.stats = (unsigned long long*)(hdr + 8);
  a.strict_fb = hdr + 12;
  a.cta_partials = (float*)((char*)scratch + 256);
}

// Kernel for (k, arithmetic, layout, imm, xtr); nullptr if that combination is not built.
tk_fn_t tk_pick(int k, int arith, int lay, bool imm, bool xtr);
bool coef_is_reference(const Geo& g, int arith);
int tk_debug_flags();

}  // namespace h2d

// stencil_team.cu
extern "C" void h2d_plan_team_prepare(const h2d_plan_t* plan);
extern "C" int h2d_plan_team_eligible(const h2d_plan_t* plan, int k, int cluster);
extern "C" int h2d_plan_launch_team_ex(h2d_plan_t* plan, int src_parity, int k, int cluster, int chunk_rows,
                                       const uint32_t* ctl, uint32_t run_if, void* stream);
