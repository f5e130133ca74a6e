/*
 * heat2d.h — C ABI of the heat2d-b200 native libraries.
 *
 *   libheat2d_cpu.so   CPU oracle, dump writers (g++ -fopenmp)
 *   libheat2d_cuda.so  sm_100a kernels + peer-memory runtime (nvcc, static cudart)
 *
 * Every function is plain `extern "C"` so it can be bound from Python (ctypes),
 * C, or anything else without a C++ or torch dependency.  Device pointers are
 * raw addresses (`tensor.data_ptr()` from PyTorch); `stream` is a cudaStream_t
 * passed as `void*` (0 = legacy default stream).
 *
 * Storage convention (same as every reference program, e.g.
 * grad1612_cuda_heat.cu:52,59 and mpi_heat2Dn.c:230): row-major `[x][y]`, `y`
 * contiguous, NX rows by NY columns.
 *
 * A *tile* is the part of the global grid one GPU owns, stored inside a padded
 * array: `halo_r` margin rows above/below and `halo_c` margin columns left and
 * right, `pitch` floats per padded row.  Owned cell (i, j) (0-based inside the
 * tile) lives at `base[(i + halo_r) * pitch + (j + halo_c)]`.  The margins are
 * the ghost region (they replace the MPI ghost ring of
 * grad1612_mpi_heat.c:47-52 and are `halo` cells wide so that k time steps can
 * be taken per exchange).
 *
 * All functions returning int return 0 on success and a negative H2D_E* code
 * on failure; h2d_last_error() gives a human-readable message for the calling
 * thread.
 */
#ifndef HEAT2D_H_
#define HEAT2D_H_

#include <stdint.h>
#include <stddef.h>

#ifdef __cplusplus
extern "C" {
#endif

/* ------------------------------------------------------------------ enums */

/* Which cells are updated.  Reference: SURVEY App. A "Domain of the update". */
enum {
  H2D_BOUNDARY_FIXED = 0,      /* mpi_heat2Dn.c:162-169, grad1612_cuda_heat.cu:58: 1..NX-2 x 1..NY-2 */
  H2D_BOUNDARY_GHOST_RING = 1  /* grad1612_mpi_heat.c:239-259: all cells, zeros outside the grid  */
};

/* How one point is evaluated. */
enum {
  H2D_ARITH_FP32 = 0,        /* throughput mode: fma(cy, ns, fma(cx, ew, a0*c)), all fp32       */
  H2D_ARITH_STRICT = 1,      /* fp64 intermediates, no contraction  (gcc, readme.md:10 flags)    */
  H2D_ARITH_STRICT_FMA = 2   /* fp64 intermediates, 2x DFMA         (nvcc default -fmad=true)    */
};

/* Initial condition u(x,y) = x(NX-x-1) y(NY-y-1)  (mpi_heat2Dn.c:247). */
enum {
  H2D_INIT_REF_INT32 = 0,    /* 32-bit wrapping product, converted as signed   (gcc)            */
  H2D_INIT_REF_UINT32 = 1,   /* 32-bit wrapping product, converted as unsigned (nvcc sm_100a)   */
  H2D_INIT_EXACT = 2,        /* int64 product rounded once to fp32                              */
  H2D_INIT_ZERO = 3
};

/* Convergence norms.  grad1612_mpi_heat.c:264-269 uses the first one. */
enum {
  H2D_NORM_NONE = 0,
  H2D_NORM_L2SQ = 1,         /* sum over cells of (new-old)^2 */
  H2D_NORM_LINF = 2          /* max over cells of |new-old|   */
};

enum {
  H2D_OK = 0,
  H2D_EINVAL = -1,
  H2D_ECUDA = -2,
  H2D_EIO = -3,
  H2D_ETIMEOUT = -4,
  H2D_ENOMEM = -5
};

/* The 8 neighbour directions of a tile (N = smaller x / row index). */
enum { H2D_N = 0, H2D_S, H2D_W, H2D_E, H2D_NW, H2D_NE, H2D_SW, H2D_SE, H2D_NDIR };

/* ---------------------------------------------------------------- structs */

typedef struct h2d_tile {
  int32_t nx, ny;          /* global grid: rows, columns                          */
  int32_t x0, y0;          /* global coordinates of owned cell (0,0)              */
  int32_t rows, cols;      /* owned extent                                        */
  int32_t halo_r, halo_c;  /* margin rows / columns of the padded array           */
  int64_t pitch;           /* floats per padded row; multiple of 4                */
} h2d_tile_t;

typedef struct h2d_phys {
  double cx, cy;           /* coefficients as doubles (0.1, or (double)0.1f for mpi_heat2Dn.c:41-44) */
  int32_t boundary;        /* H2D_BOUNDARY_*                                      */
  int32_t arith;           /* H2D_ARITH_*                                         */
} h2d_phys_t;

/* ------------------------------------------------------------------ common */

const char* h2d_last_error(void);
const char* h2d_version(void);

/* =================================================================== CPU == */

/* Fill the owned cells of a tile (and, when `with_margins` != 0, the margin
 * cells that fall inside the global grid; margin cells outside it become 0).
 * Reference: inidat(), mpi_heat2Dn.c:242-248; grad1612_mpi_heat.c:150-168. */
int h2d_cpu_init(float* base, const h2d_tile_t* tile, int init_mode, int with_margins);

/* Serial / OpenMP oracle: advance a full NX x NY grid (dense, pitch == ny) by
 * up to `steps` steps.  `u` holds the initial field and receives the result.
 * Convergence (norm != NONE): every `interval` steps, on the 0-based step index
 * (the intent of grad1612_mpi_heat.c:262-271), compute the norm of (new - old);
 * stop when it is < eps; the returned field is then the PRE-step one and the
 * return value is that step index (the reference's "Exiting after %d
 * iterations").  Without convergence the return value is `steps`.
 * `threads` <= 1 runs serially; > 1 uses a correct `omp parallel for`
 * (the reference's nested pragmas, grad1612_hybrid_heat.c:243-259, are not
 * reproduced).  `last_norm` (nullable) receives the last evaluated norm.     */
int64_t h2d_cpu_run(float* u, int32_t nx, int32_t ny, int64_t steps,
                    const h2d_phys_t* phys, int norm, int64_t interval, double eps,
                    int threads, double* last_norm);

/* One step on a padded tile array (reads margins of src as ghosts); used to
 * cross-check tile kernels.  Writes owned cells of dst only. */
int h2d_cpu_step_tile(const float* src, float* dst, const h2d_tile_t* tile,
                      const h2d_phys_t* phys);

/* Text dump, format of prtdat() (mpi_heat2Dn.c:253-268): y descending, x
 * across, "%6.1f" separated by one blank, newline after the last value.      */
int h2d_dump_prtdat(const char* path, const float* u, int32_t nx, int32_t ny);
/* Text dump, format of grad1612_mpi_heat.c:195-200: row-major, "%6.1f " and a
 * newline per row (trailing blank kept).                                      */
int h2d_dump_rowmajor(const char* path, const float* u, int32_t nx, int32_t ny);
/* Raw fp32 row-major file (initial_binary.dat / final_binary.dat).  Each rank
 * writes its own rows with pwrite at ((x0+i)*ny + y0)*4 — the replacement for
 * the MPI-IO subarray view of grad1612_mpi_heat.c:178-190.  `truncate` != 0
 * sets the file length to nx*ny*4 first (the reference never truncates).      */
int h2d_dump_binary_tile(const char* path, const float* owned, int64_t owned_pitch,
                         const h2d_tile_t* tile, int truncate);
/* Test hooks for the text writers' own "%6.1f" formatter: format one value into out64
 * (at least 64 bytes; returns the length), and compare it with snprintf over `count` float
 * bit patterns first, first + stride, ... (returns the number of mismatches).           */
int h2d_format_6_1f(float v, char* out64);
int64_t h2d_format_6_1f_check(uint32_t first, uint64_t count, uint32_t stride, uint32_t* first_bad);
/* rank 0: convert the binary file to text (grad1612_mpi_heat.c:191-203). */
int h2d_binary_to_text(const char* bin_path, const char* txt_path, int32_t nx, int32_t ny,
                       int format /*0 = rowmajor, 1 = prtdat*/);

/* ================================================================== CUDA == */

/* Device properties dump, the DEBUG block of grad1612_cuda_heat.cu:24-37.
 * Writes at most `cap` bytes of text into `out`. */
int h2d_cuda_device_info(int device, char* out, size_t cap);

/* K1 — initial condition on the device; same semantics as h2d_cpu_init. */
int h2d_cuda_init(float* base, const h2d_tile_t* tile, int init_mode, int with_margins,
                  void* stream);

/* K2/K3 — one time step, plain kernel (no temporal blocking, no TMA): each
 * warp streams a 128-column strip down a block of rows with a 3-row register
 * window.  Reads ghost cells from the margins of `src`; writes owned cells of
 * `dst`.  Replaces update(), grad1612_cuda_heat.cu:55-62.                     */
int h2d_cuda_step_k1(const float* src, float* dst, const h2d_tile_t* tile,
                     const h2d_phys_t* phys, void* stream);

/* ---- plan: the temporal-blocked, TMA-fed, halo-fusing stencil (K4-K7) ---- */

typedef struct h2d_plan h2d_plan_t;

typedef struct h2d_peer {
  /* One entry per direction; buf[p] == NULL means "no neighbour" (MPI_PROC_NULL,
   * grad1612_mpi_heat.c:79-81).  Pointers are peer-mapped device addresses.    */
  float* buf[2];            /* neighbour's padded arrays, parity 0 / 1            */
  uint32_t* flag;           /* where THIS rank publishes its epoch for that peer  */
  int32_t x0, y0;           /* neighbour tile origin (global)                     */
  int32_t halo_r, halo_c;
  int64_t pitch;
  /* W / E entries only, optional: the neighbour's PACKED column-ghost buffers (parity 0 / 1)
   * that this rank's edge columns go into — the neighbour's h2d_plan_desc_t::colbuf_in of
   * the facing side.  NULL: the columns are stored in place into buf[] (slow on long tiles:
   * one short remote write per row).                                                      */
  float* colbuf[2];
} h2d_peer_t;

typedef struct h2d_plan_desc {
  h2d_tile_t tile;
  h2d_phys_t phys;
  float* buf[2];            /* this rank's padded arrays                          */
  int32_t kmax;             /* largest k that will be launched (<= halo)          */
  int32_t chunk_rows;       /* rows per streaming task; 0 = choose                */
  /* multi-GPU (all optional; zero = single tile) */
  h2d_peer_t peer[H2D_NDIR];
  uint32_t* flags_in;       /* local array [H2D_NDIR]: epoch published BY dir d   */
  /* convergence scratch (optional): >= h2d_plan_scratch_bytes() bytes          */
  void* scratch;
  int64_t spin_timeout_ns;  /* watchdog for flag waits; 0 = 10 s                  */
  /* Lane layout of the fp32 kernels (csrc/tk_task.cuh): 0 = library default
   * (H2D_FP32_LAYOUT=v4|ab2|ab4 overrides it), 1 = v4 (4 adjacent columns per lane,
   * 128-column strips), 2 = ab2 / 3 = ab4 (two half-strips packed in f32x2, 128- /
   * 256-column strips), 4 = v8 (8 adjacent columns per lane, 256-column strips; strict
   * arithmetic, kmax <= 4 only).  Strict arithmetic uses v4 or v8.  A layout that does not exist
   * for the arithmetic / depth falls back to the default.  Results are bit-identical.   */
  int32_t layout;
  /* != 0: k-step sweeps (k >= 5) of a plan without neighbours, launched without a norm, go out as
   * cluster launches whose strips exchange their edge columns through shared memory instead of
   * recomputing them (h2d_plan_launch_team with one CTA per team; tiles narrower than 472
   * columns and all other launches are unaffected).  Same bits.  Measured on B200 in strict-fma:
   * 3-7 % faster at 32768^2, 2 % at 16384^2 with short tasks, slower below; fp32: 3-4 % slower. */
  int32_t team;
  /* Packed column-ghost buffers of THIS rank, [side][parity] (side 0 = ghosts west of the
   * tile, 1 = east), each rows x 8 floats, in memory the W / E neighbour can reach (it gets
   * the pointers as its h2d_peer_t::colbuf).  The neighbour writes its k edge columns of tile
   * row r to floats [8r, 8r + k); this rank's edge tasks copy the rows they read into the
   * margins of the source array first.  After filling a padded array from the host
   * (initial condition, restart, upload) call h2d_plan_seed_colbuf for that parity.  All NULL:
   * ghost columns arrive in place.                                                        */
  float* colbuf_in[2][2];
} h2d_plan_desc_t;

size_t h2d_plan_scratch_bytes(const h2d_tile_t* tile, int32_t kmax, int32_t chunk_rows);
/* Rows per streaming task the plan would choose for this tile (pure host computation:
 * the makespan model of DESIGN.md section 3; with_peers != 0 asks for >= 4 waves). */
int32_t h2d_plan_chunk_rows(const h2d_tile_t* tile, int32_t kmax, int32_t with_peers);
/* The same for a given arithmetic (H2D_ARITH_*) and lane layout (0 = default), and the
 * strip width in columns that combination works with (128 or 256). */
int32_t h2d_plan_chunk_rows_ex(const h2d_tile_t* tile, int32_t kmax, int32_t with_peers, int32_t arith,
                               int32_t layout);
int32_t h2d_plan_strip_width(int32_t arith, int32_t layout);
/* Copy the ghost columns found in the margins of buf[parity] (rows 0 .. rows-1) into this
 * rank's packed column-ghost buffers of that parity: to be called after the padded array was
 * filled by anything but the neighbours' launches.  No-op without colbuf_in.               */
int h2d_plan_seed_colbuf(h2d_plan_t* plan, int parity, void* stream);
/* Lane layout a plan was built with (1 = v4, 2 = ab2, 3 = ab4). */
int32_t h2d_plan_layout(const h2d_plan_t* plan);
h2d_plan_t* h2d_plan_create(const h2d_plan_desc_t* desc);
void h2d_plan_destroy(h2d_plan_t* plan);

/* Advance k steps: reads buf[src_parity], writes buf[1 - src_parity] (owned
 * cells, plus the neighbours' ghost cells when peers are set).  `epoch` is the
 * 1-based launch counter shared by all ranks (flag protocol).
 * norm != NONE additionally reduces the LAST sub-step's (new - old) over the
 * owned cells into result slot `norm_slot` (see h2d_plan_norm_*).            */
int h2d_plan_launch(h2d_plan_t* plan, int src_parity, int k, uint32_t epoch, int norm,
                    void* stream);

/* Same, under device-side run control: the launch is a no-op unless the device
 * word *ctl equals run_if (ctl == NULL: unconditional).  With ctl = the word
 * h2d_cuda_norm_finish sets on convergence and run_if = 0, everything queued
 * after a converged check drains without doing work and WITHOUT the host ever
 * synchronising inside the step loop (the reference stops with a host-side
 * `break` after a blocking MPI_Allreduce, grad1612_mpi_heat.c:268-269).
 * run_if = q (> 0) makes a launch that runs ONLY if check number q converged:
 * used to regenerate the pre-step field the reference returns.               */
int h2d_plan_launch_ex(h2d_plan_t* plan, int src_parity, int k, uint32_t epoch, int norm,
                       const uint32_t* ctl, uint32_t run_if, void* stream);

/* `n_launches` consecutive k-step launches without going back to the caller:
 * launch l uses epoch0 + l and alternates the source parity starting from
 * `src_parity`; with several plans (tiles emulated in one process) each launch
 * index is issued for every plan before the next one starts.  Small grids are
 * bound by the per-launch host cost, which this keeps at the bare
 * cudaLaunchKernel (the reference pays one host launch per time step,
 * grad1612_cuda_heat.cu:82-85).                                               */
int h2d_plan_run_many(h2d_plan_t* const* plans, int n_plans, int src_parity, int k, int n_launches,
                      uint32_t epoch0, const uint32_t* const* ctls /* per plan, or NULL */, void* stream);

/* `n_epochs` k-step sweeps in ONE launch (single-tile plans, no norm): a grid with one
 * warp per (epoch, task); warps draw tickets in that order and start a task as soon as the tasks
 * of the previous epoch that it reads from / writes over have finished, so the machine
 * never drains between sweeps.  Result parity as with h2d_plan_run_many:
 * src_parity ^ (n_epochs & 1).  The first call allocates one counter per task (later
 * calls with the same or a smaller k reuse them).  A dependency that does not arrive
 * within the plan's spin timeout sets the fault word to 50 and the kernel winds down.
 * The reference launches once per time step (grad1612_cuda_heat.cu:82-85).          */
int h2d_plan_run_dataflow(h2d_plan_t* plan, int src_parity, int k, int n_epochs, void* stream);

/* One k-step sweep (5 <= k <= 8) as a CLUSTER launch: the tile is cut into teams of
 * 4 * cluster adjacent strips (cluster = 1, 2 or 4 CTAs), one team per thread-block cluster;
 * the strips of a team advance in row lock-step and hand each other their edge columns
 * half-way through the k levels — through shared memory inside a CTA, through distributed
 * shared memory between the CTAs — so that adjacent strips overlap by 8 columns instead of
 * 16 (csrc/tk_team.cuh).  Single-tile plans, v4 lane layout, fp32 / strict-fma (strict: cluster = 1),
 * no norm.  chunk_rows = 0 picks the task height.  Same result, bit for bit, as
 * h2d_plan_launch.  Strict arithmetic: strips whose data left the domain of the fast evaluation
 * are counted (h2d_plan_team_fallback_ptr); a one-CTA team (cluster = 1) then runs its task
 * again with the plain evaluation, larger clusters leave that to the caller (repeat the sweep
 * with h2d_plan_launch if the counter moved).  The reference has one thread per point and no reuse
 * (grad1612_cuda_heat.cu:55-62).                                                     */
int h2d_plan_launch_team(h2d_plan_t* plan, int src_parity, int k, int cluster, int chunk_rows, void* stream);
uint32_t* h2d_plan_team_fallback_ptr(h2d_plan_t* plan);

/* Upload, solve, download — overlapped.  For a single-tile plan: copies the nx x ny
 * field `host_in` (pinned host memory, `in_pitch` floats per row) into buf[0], advances
 * it `steps` steps in launches of k, and copies the result to `host_out`, with the
 * three phases pipelined: the grid is cut into blocks of `block_rows` rows
 * (0 = about two waves of tasks) and block b runs ALL its launches before block b+1
 * starts, the window of launch e being shifted up by e * skew_rows rows
 * (0 = k + 4; must exceed k) so that every launch finds its k halo rows already
 * computed by the previous launch.  Block b can start as soon as its rows (and k
 * more) have arrived, and its final rows leave while later blocks still compute.
 * Everything is ordered on `stream` for the caller: when work queued on `stream`
 * after this call runs, host_out is complete.  *final_parity = which buffer holds
 * the result.  The reference has no counterpart: its CUDA program never moves the
 * field across PCIe at all (grad1612_cuda_heat.cu:75-92).                        */
int h2d_plan_solve_pipelined(h2d_plan_t* plan, const float* host_in, int64_t in_pitch, float* host_out,
                             int64_t out_pitch, int64_t steps, int k, int block_rows, int skew_rows,
                             void* stream, int* final_parity);

/* The same for a rank of a 1 x P column-strip run (plans whose only neighbours are W and E):
 * every rank owns all rows, cuts the same row blocks and issues the same sequence of window
 * launches — one epoch each, starting at epoch0; *epochs_used tells the caller how far the
 * shared epoch counter advanced — so the column bands and flags travel per window exactly as
 * they do per whole-tile launch.  `wave_rows` (> 0) fixes the block granularity; ranks must
 * pass the same value (the minimum of h2d_plan_wave_rows over the ranks).  host_in points
 * at owned cell (0,0) of this rank's strip inside a host array that also holds 32 valid
 * columns to the left / right of it wherever a W / E neighbour exists (the level-0 ghosts).
 * Row strips cannot be pipelined this way (a strip's top halo is produced by the last block
 * of the strip above), so the call refuses N / S neighbours.                                */
int h2d_plan_solve_pipelined_ex(h2d_plan_t* plan, const float* host_in, int64_t in_pitch, float* host_out,
                                int64_t out_pitch, int64_t steps, int k, int block_rows, int skew_rows,
                                int wave_rows, uint32_t epoch0, void* stream, int* final_parity,
                                uint32_t* epochs_used);
/* The tiles of a 1 x P grid emulated in ONE process (several plans, one compute stream):
 * every window launch is issued for every plan before the next one.  Arrays have one entry
 * per plan; the block granularity is the minimum over the plans.                          */
int h2d_plan_solve_pipelined_many(h2d_plan_t* const* plans, int n_plans, const float* const* host_in,
                                  const int64_t* in_pitch, float* const* host_out, const int64_t* out_pitch,
                                  int64_t steps, int k, int block_rows, int skew_rows, uint32_t epoch0,
                                  void* stream, int* final_parity, uint32_t* epochs_used);
/* Rows that one wave of window tasks covers for this plan (block granularity above). */
int32_t h2d_plan_wave_rows(const h2d_plan_t* plan);

/* Number of kernels launched by this plan since creation (bench bookkeeping). */
int64_t h2d_plan_launch_count(const h2d_plan_t* plan);
/* Device address of this rank's partial norm (1 float, valid after a launch
 * with norm != NONE has completed). */
float* h2d_plan_norm_ptr(h2d_plan_t* plan);
/* Device address of the watchdog word: non-zero after a flag wait timed out
 * (value = 1 + direction that stalled; 50 = a task dependency inside
 * h2d_plan_run_dataflow; 100 + rank = a norm slot). */
uint32_t* h2d_plan_fault_ptr(h2d_plan_t* plan);
/* Device address of two 64-bit counters: [0] nanoseconds edge tasks spent blocked on a
 * neighbour's epoch flag, [1] how many such waits blocked at all — the in-kernel
 * counterpart of the MPI_Waitall row of the reference's mpiP profile (Report.pdf p35-37). */
uint64_t* h2d_plan_stats_ptr(h2d_plan_t* plan);

/* Device address of three rotating 32-bit counters (strict arithmetic only): tasks of the
 * last launches whose data was outside the domain of the fast evaluation and were computed
 * with the plain one (slot = logical launch number % 3). */
uint32_t* h2d_plan_strict_fallback_ptr(h2d_plan_t* plan);

/* ---- resident solver: whole grid in the shared memory of one cluster ------ */
/* All `steps` steps in ONE launch; halo rows travel between the CTAs of the
 * cluster through distributed shared memory.  Returns H2D_EINVAL if the grid
 * does not fit (see h2d_resident_fits).  */
int h2d_resident_fits(int32_t nx, int32_t ny);
int h2d_resident_run(const float* src, float* dst, int32_t nx, int32_t ny, int64_t pitch,
                     int64_t steps, const h2d_phys_t* phys, void* stream);

/* ---- standalone reduction (baseline path and tests) ----------------------- */
/* norm of (a - b) over the owned cells; result written to *out_dev (device).
 * `scratch` >= h2d_norm_scratch_bytes(tile) bytes of device memory.           */
size_t h2d_norm_scratch_bytes(const h2d_tile_t* tile);
int h2d_cuda_norm_diff(const float* a, const float* b, const h2d_tile_t* tile, int norm,
                       void* scratch, float* out_dev, void* stream);

/* Same result as h2d_cuda_norm_diff, for callers that cannot provide scratch
 * memory because they cannot allocate any: programs whose own pool already
 * owns the whole GPU (a cudaMalloc from inside this library would fail or
 * fight that pool), and steady-state loops that must not pay the device-wide
 * synchronisation hidden in cudaMalloc / cudaFree.  The scratch is a fixed
 * array inside the library image instead: 64 KiB + 256 B per GPU that loads
 * the library, held until the process exits.
 *  - Limit: tile->rows <= H2D_NORM_STATIC_MAX_ROWS (any number of columns);
 *    larger tiles are rejected with H2D_EINVAL and a message.
 *  - Allocates nothing: no cudaMalloc / cudaFree, no allocating library call.
 *    The caller still passes the inputs (a, b) and the output (out_dev).
 *  - The first call on each device looks up the array's device address
 *    (cudaGetSymbolAddress, with that device current) and caches it; make that
 *    first call before starting any stream capture.
 *  - All calls on one GPU share the one array: each call clears the words that
 *    must start at zero on `stream` before its kernel, but two calls that can
 *    overlap on the GPU (different streams, no ordering) race and give wrong
 *    sums.  Serialise them, or use h2d_cuda_norm_diff with separate scratch.
 * No counterpart in the reference (its only reduction runs on the CPU,
 * grad1612_mpi_heat.c:264-267). */
#define H2D_NORM_STATIC_MAX_ROWS 131072
int h2d_cuda_norm_diff_static(const float* a, const float* b, const h2d_tile_t* tile, int norm,
                              float* out_dev, void* stream);

/* ---- cross-GPU finish of the convergence value ---------------------------- */
/* Writes `*local_val` into slot `rank` of every peer's value table (peer
 * pointers in vals[r], one table of 2*world floats per rank: [parity][rank]),
 * publishes `seq` in the matching flag table, waits until all `world` slots of
 * the local table carry `seq`, combines them in rank order (sum or max) and
 * stores the total at *out_dev and (total < eps) at *stop_dev.
 * inline_wait == 0 skips the publish/wait and only combines (used when ranks
 * are emulated as consecutive launches on one GPU).                           */
int h2d_cuda_norm_finish(const float* local_val, float* const* vals, uint32_t* const* flags,
                         int rank, int world, uint32_t seq, int norm, float eps,
                         float* out_dev, uint32_t* stop_dev, int phase /*0 publish,1 combine,2 both*/,
                         int64_t spin_timeout_ns, uint32_t* fault_dev, void* stream);
/* Same with run control: does nothing if *ctl != 0 (an earlier check already
 * stopped the run); on convergence stores `seq` into *ctl.                    */
int h2d_cuda_norm_finish_ctl(const float* local_val, float* const* vals, uint32_t* const* flags,
                             int rank, int world, uint32_t seq, int norm, float eps,
                             float* out_dev, uint32_t* ctl, int phase, int64_t spin_timeout_ns,
                             uint32_t* fault_dev, void* stream);

/* ---- halo pack / unpack for the NCCL-halo BASELINE backend (K8) ----------- */
/* Pack the owned edge (width w) facing direction dir (N,S,W,E) into `out`
 * (contiguous), or unpack `in` into the ghost margin facing dir.              */
int h2d_cuda_pack_edge(const float* base, const h2d_tile_t* tile, int dir, int w, float* out,
                       void* stream);
int h2d_cuda_unpack_ghost(float* base, const h2d_tile_t* tile, int dir, int w, const float* in,
                          void* stream);

/* ---- peer-memory arena (CUDA IPC) ----------------------------------------- */
typedef struct h2d_arena h2d_arena_t;
h2d_arena_t* h2d_arena_create(size_t bytes);            /* cudaMalloc + zero          */
void* h2d_arena_ptr(h2d_arena_t* a);
size_t h2d_arena_bytes(h2d_arena_t* a);
int h2d_arena_export(h2d_arena_t* a, uint8_t handle[64]);
void* h2d_arena_import(const uint8_t handle[64]);       /* maps a peer arena          */
int h2d_arena_unmap(void* peer_ptr);
void h2d_arena_destroy(h2d_arena_t* a);

#ifdef __cplusplus
}
#endif
#endif /* HEAT2D_H_ */
