// Peer-memory runtime: a CUDA-IPC arena per rank (tile buffers, flag words and
// reduction tables live in it so every neighbour can map them), and the
// cross-GPU finish of the convergence value.
//
// Replaces, in the reference: MPI_Cart_create / persistent Send_init/Recv_init
// plumbing (grad1612_mpi_heat.c:74-81, 210-227) — peers are plain mapped
// pointers here — and MPI_Allreduce(1 float, SUM) (:268).  torch.distributed
// is used only to ship the 64-byte handles between ranks at start-up.
#include <cstring>
#include "device_common.cuh"

struct h2d_arena {
  void* ptr = nullptr;
  size_t bytes = 0;
  int device = 0;
};

namespace h2d {

// One thread block.  phase bit 0: publish my value to every rank's table;
// phase bit 1: wait for all ranks' values of this sequence number, combine in
// rank order (so every rank gets the same bits) and decide.
__global__ void norm_finish_kernel(const float* local_val, float* const* vals,
                                   uint32_t* const* flags, int rank, int world, uint32_t seq,
                                   int norm, float eps, float* out, uint32_t* stop, int phase,
                                   int64_t timeout_ns, uint32_t* fault, uint32_t* ctl) {
  // An earlier check already stopped the run: every rank sees the same word.
  // (In the split-phase emulation the publish phase of check `seq` runs before any
  // combine phase of the same check, so ctl == seq can only be seen by a combine.)
  if (ctl != nullptr) {
    const uint32_t c = *reinterpret_cast<volatile uint32_t*>(ctl);
    if (c != 0u && c != seq) return;
  }
  const int par = (int)(seq & 1u);
  if (phase & 1) {
    const float v = *local_val;
    for (int r = threadIdx.x; r < world; r += blockDim.x) {
      st_relaxed_sys_f32(vals[r] + par * world + rank, v);
      __threadfence_system();
      st_release_sys(flags[r] + par * world + rank, seq);
    }
  }
  if (!(phase & 2)) return;
  __shared__ int failed;
  if (threadIdx.x == 0) failed = 0;
  __syncthreads();
  for (int r = threadIdx.x; r < world; r += blockDim.x) {
    if (!wait_flag_ge(flags[rank] + par * world + r, seq, timeout_ns, fault)) {
      failed = 1;
      if (fault) atomicCAS(fault, 0u, 100u + r);   // keep the first cause
    }
  }
  __syncthreads();
  if (threadIdx.x == 0) {
    const float* mine = vals[rank] + par * world;
    float t = 0.f;
    for (int r = 0; r < world; ++r) {
      const float x = *((volatile const float*)(mine + r));
      t = (norm == H2D_NORM_L2SQ) ? t + x : fmax_nan(t, x);
    }
    *out = t;
    if (!failed && t < eps) {
      if (stop != nullptr) *stop = 1u;
      if (ctl != nullptr) *ctl = seq;
    }
  }
}

}  // namespace h2d

extern "C" {

h2d_arena_t* h2d_arena_create(size_t bytes) {
  if (bytes == 0) {
    h2d_fail(H2D_EINVAL, "h2d_arena_create: zero bytes");
    return nullptr;
  }
  h2d_arena* a = new h2d_arena();
  cudaGetDevice(&a->device);
  // Plain cudaMalloc (not a pooled / VMM allocation): legacy IPC handles need it.
  cudaError_t e = cudaMalloc(&a->ptr, bytes);
  if (e != cudaSuccess) {
    h2d_fail(H2D_ENOMEM, "h2d_arena_create: cudaMalloc(%zu) failed: %s", bytes, cudaGetErrorString(e));
    delete a;
    return nullptr;
  }
  e = cudaMemset(a->ptr, 0, bytes);
  if (e == cudaSuccess) e = cudaDeviceSynchronize();
  if (e != cudaSuccess) {
    h2d_fail(H2D_ECUDA, "h2d_arena_create: clearing failed: %s", cudaGetErrorString(e));
    cudaFree(a->ptr);
    delete a;
    return nullptr;
  }
  a->bytes = bytes;
  return a;
}

void* h2d_arena_ptr(h2d_arena_t* a) { return a ? a->ptr : nullptr; }
size_t h2d_arena_bytes(h2d_arena_t* a) { return a ? a->bytes : 0; }

int h2d_arena_export(h2d_arena_t* a, uint8_t handle[64]) {
  if (!a || !handle) return h2d_fail(H2D_EINVAL, "h2d_arena_export: null argument");
  static_assert(sizeof(cudaIpcMemHandle_t) == 64, "IPC handle size");
  cudaIpcMemHandle_t h;
  H2D_CUDA_TRY(cudaIpcGetMemHandle(&h, a->ptr));
  memcpy(handle, &h, 64);
  return H2D_OK;
}

void* h2d_arena_import(const uint8_t handle[64]) {
  if (!handle) {
    h2d_fail(H2D_EINVAL, "h2d_arena_import: null handle");
    return nullptr;
  }
  cudaIpcMemHandle_t h;
  memcpy(&h, handle, 64);
  void* p = nullptr;
  cudaError_t e = cudaIpcOpenMemHandle(&p, h, cudaIpcMemLazyEnablePeerAccess);
  if (e != cudaSuccess) {
    h2d_fail(H2D_ECUDA, "cudaIpcOpenMemHandle failed: %s", cudaGetErrorString(e));
    return nullptr;
  }
  return p;
}

int h2d_arena_unmap(void* peer_ptr) {
  if (!peer_ptr) return H2D_OK;
  H2D_CUDA_TRY(cudaIpcCloseMemHandle(peer_ptr));
  return H2D_OK;
}

void h2d_arena_destroy(h2d_arena_t* a) {
  if (!a) return;
  if (a->ptr) cudaFree(a->ptr);
  delete a;
}

static int norm_finish_impl(const float* local_val, float* const* vals, uint32_t* const* flags,
                            int rank, int world, uint32_t seq, int norm, float eps, float* out_dev,
                            uint32_t* stop_dev, uint32_t* ctl, int phase, int64_t spin_timeout_ns,
                            uint32_t* fault_dev, void* stream);

int h2d_cuda_norm_finish(const float* local_val, float* const* vals, uint32_t* const* flags,
                         int rank, int world, uint32_t seq, int norm, float eps, float* out_dev,
                         uint32_t* stop_dev, int phase, int64_t spin_timeout_ns,
                         uint32_t* fault_dev, void* stream) {
  if (!stop_dev) return h2d_fail(H2D_EINVAL, "h2d_cuda_norm_finish: null argument");
  return norm_finish_impl(local_val, vals, flags, rank, world, seq, norm, eps, out_dev, stop_dev, nullptr,
                          phase, spin_timeout_ns, fault_dev, stream);
}

int h2d_cuda_norm_finish_ctl(const float* local_val, float* const* vals, uint32_t* const* flags,
                             int rank, int world, uint32_t seq, int norm, float eps, float* out_dev,
                             uint32_t* ctl, int phase, int64_t spin_timeout_ns, uint32_t* fault_dev,
                             void* stream) {
  if (!ctl) return h2d_fail(H2D_EINVAL, "h2d_cuda_norm_finish_ctl: null ctl");
  if (seq == 0) return h2d_fail(H2D_EINVAL, "h2d_cuda_norm_finish_ctl: seq must be >= 1");
  return norm_finish_impl(local_val, vals, flags, rank, world, seq, norm, eps, out_dev, nullptr, ctl, phase,
                          spin_timeout_ns, fault_dev, stream);
}

static int norm_finish_impl(const float* local_val, float* const* vals, uint32_t* const* flags,
                            int rank, int world, uint32_t seq, int norm, float eps, float* out_dev,
                            uint32_t* stop_dev, uint32_t* ctl, int phase, int64_t spin_timeout_ns,
                            uint32_t* fault_dev, void* stream) {
  if (!local_val || !vals || !flags || !out_dev)
    return h2d_fail(H2D_EINVAL, "h2d_cuda_norm_finish: null argument");
  if (world < 1 || rank < 0 || rank >= world)
    return h2d_fail(H2D_EINVAL, "h2d_cuda_norm_finish: rank %d / world %d", rank, world);
  if (norm != H2D_NORM_L2SQ && norm != H2D_NORM_LINF)
    return h2d_fail(H2D_EINVAL, "h2d_cuda_norm_finish: unknown norm %d", norm);
  const int ph = phase == 0 ? 1 : phase == 1 ? 2 : 3;
  const int64_t to = spin_timeout_ns > 0 ? spin_timeout_ns : 10000000000LL;
  h2d::norm_finish_kernel<<<1, 32, 0, (cudaStream_t)stream>>>(
      local_val, vals, flags, rank, world, seq, norm, eps, out_dev, stop_dev, ph, to, fault_dev, ctl);
  H2D_CHECK_LAUNCH("norm_finish_kernel");
  return H2D_OK;
}

}  // extern "C"
