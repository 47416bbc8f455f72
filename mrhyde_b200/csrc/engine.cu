// Device side of the engine: sm_100a kernels wrapping the phase functions of assemble.h / assemble_fast.h,
// slot-pattern dictionary construction, DBC fix-up kernels, and the C-ABI entry points that need the GPU.
// There is NO CPU fallback: every entry point here returns MRH_ECUDA when the device is unavailable.
#include <cuda_runtime.h>
#include <thrust/device_ptr.h>
#include <thrust/execution_policy.h>
#include <thrust/sequence.h>
#include <thrust/sort.h>
#include <thrust/unique.h>

#include "capi_host.inl"
#include "assemble_fast.h"
#include "assemble_fast2d.h"
#include "assemble_batched.h"
#include "dispatch.h"
#include "generic_kernels.h"
#include "jit.hpp"

using namespace mrh;

#define CUDA_CK(call)                                                                                   \
  do { cudaError_t e_ = (call); if (e_ != cudaSuccess)                                                  \
         throw std::runtime_error(std::string("CUDA error: ") + cudaGetErrorString(e_) + " at " #call); } while (0)

namespace {

template <class T>
struct DevBuf {
  T* p = nullptr; size_t n = 0;
  DevBuf() {}
  DevBuf(const DevBuf&) = delete;
  DevBuf& operator=(const DevBuf&) = delete;
  ~DevBuf() { release(); }
  void release() { if (p) cudaFree(p); p = nullptr; n = 0; }
  void alloc(size_t count) { release(); n = count; if (count) CUDA_CK(cudaMalloc((void**)&p, count * sizeof(T))); }
  void upload(const T* h, size_t count) { alloc(count); if (count) CUDA_CK(cudaMemcpy(p, h, count * sizeof(T), cudaMemcpyHostToDevice)); }
  void upload(const std::vector<T>& v) { upload(v.data(), v.size()); }
};

struct TabDev { DevBuf<double> B, G, VB, VG, W, FR; };
struct ExchangeDev;                    // exchange.inl

struct DeviceImpl {
  int device = 0, num_sms = 148, max_smem_optin = 48 * 1024;
  DevBuf<double> nodes; DevBuf<int> e2n, lids, pat_id, dbc, dbc_diag_dummy;
  DevBuf<uint8_t> fixed; DevBuf<int64_t> rowptr, dbc_slots; DevBuf<int> colind; DevBuf<uint16_t> pat;
  DevBuf<Instr> code; DevBuf<double> cpool;
  TabDev vol_tab, side_tab[6];
  std::vector<std::unique_ptr<DevBuf<int>>> side_elems;
  BasePtrs bp;
  int64_t npatterns = 0, launches = 0;
  bool ready = false;
  cudaEvent_t ev0 = nullptr, ev1 = nullptr;
  double vol_ms = 0.0;
  // staging for the *_host entry point
  DevBuf<double> h_u, h_res, h_jac;
  JitCache jit;
  struct TileDev* tiles = nullptr;     // engine_tiles.inl
  bool jit_used = false;
  // chunked copy-out of the host-buffer entry point (engine_launch.inl build_chunk_plan)
  bool chunk_ready = false;
  std::vector<int64_t> chunk_tile_end, chunk_val_end;
  std::vector<cudaEvent_t> chunk_ev;
  cudaStream_t copy_stream = nullptr;
  const char* last_path = "none";
  std::shared_ptr<void> solver;        // SolverWork (solver.inl), created at the first solve
  ExchangeDev* xch = nullptr;          // NCCL communicator + interface exchange tables (exchange.inl), mrh_comm_init
  int64_t iface_tiles = 0;             // leading tiles of the tile plan that touch DOFs owned by another rank (-1: no split)
  double* static_scratch = nullptr;    // device address of g_static_scratch on this handle's GPU (set at finalize)
};

// ------------------------------------------------------------------------------------------------ kernels
template <bool TRANSIENT, bool ADV, bool JAC>
__global__ void __launch_bounds__(128) thermal_hex_q1_kernel(const __grid_constant__ KParams kp) {
  __shared__ double stage[ThermalHexQ1<TRANSIENT, ADV, JAC>::NSTAGE * 8 * 128];
  for (int64_t e = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; e < kp.nwork; e += (int64_t)gridDim.x * blockDim.x)
    ThermalHexQ1<TRANSIENT, ADV, JAC>::element(kp, e, stage + threadIdx.x, 128);
}

template <bool TRANSIENT, bool ADV, bool JAC>
__global__ void __launch_bounds__(128) thermal_quad_q1_kernel(const __grid_constant__ KParams kp) {
  for (int64_t e = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; e < kp.nwork; e += (int64_t)gridDim.x * blockDim.x)
    ThermalQuadQ1<TRANSIENT, ADV, JAC>::element(kp, e);
}

template <bool TRANSIENT, bool ADV>
__global__ void __launch_bounds__(128) thermal_quad_q1_tiled_kernel(const __grid_constant__ KParams kp) {
  extern __shared__ double dsm[];
  ThermalQuadQ1<TRANSIENT, ADV, true>::tiled(kp, dsm);
}

template <bool TRANSIENT, bool ADV>
__global__ void __launch_bounds__(128) thermal_hex_q1_tiled_kernel(const __grid_constant__ KParams kp) {
  __shared__ double stage[ThermalHexQ1<TRANSIENT, ADV, true>::NSTAGE * 8 * 128];
  extern __shared__ double dsm[];
  ThermalHexQ1<TRANSIENT, ADV, true>::tiled(kp, stage + threadIdx.x, 128, dsm);
}

// row-relative CRS position of column `col` in row `row` (columns ascending)
__device__ __forceinline__ int rel_slot(const int64_t* rowptr, const int* colind, int row, int col) {
  const int64_t b = rowptr[row];
  int lo = 0, hi = (int)(rowptr[row + 1] - b);
  while (lo < hi) { const int mid = (lo + hi) >> 1; if (colind[b + mid] < col) lo = mid + 1; else hi = mid; }
  return (lo < (int)(rowptr[row + 1] - b) && colind[b + lo] == col) ? lo : -1;
}
__device__ __forceinline__ uint64_t mix64(uint64_t h, uint64_t v) {
  h ^= v + 0x9E3779B97F4A7C15ull + (h << 6) + (h >> 2);
  h *= 0xBF58476D1CE4E5B9ull; h ^= h >> 29;
  return h;
}
// pass 1: hash of each element's N x N slot pattern
__global__ void pattern_hash_kernel(const int* lids, const int64_t* rowptr, const int* colind, int64_t nelem, int N,
                                    uint64_t* hash, int* bad) {
  const int64_t e = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (e >= nelem) return;
  const int* l = lids + e * N;
  uint64_t h = 0x1234567ull;
  for (int r = 0; r < N; ++r)
    for (int c = 0; c < N; ++c) {
      const int s = rel_slot(rowptr, colind, l[r], l[c]);
      if (s < 0) atomicExch(bad, 1);
      h = mix64(h, (uint64_t)(s + 1));
    }
  hash[e] = h;
}
// pass 2: materialise the pattern of each representative element
__global__ void pattern_fill_kernel(const int* lids, const int64_t* rowptr, const int* colind, const int* rep, int64_t npat,
                                    int N, uint16_t* pat) {
  const int64_t k = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (k >= npat * N) return;
  const int64_t p = k / N; const int r = (int)(k % N);
  const int* l = lids + (int64_t)rep[p] * N;
  for (int c = 0; c < N; ++c) pat[(p * N + r) * N + c] = (uint16_t)rel_slot(rowptr, colind, l[r], l[c]);
}
// pass 3: pattern id of every element (binary search on the sorted unique hashes) and full verification
__global__ void pattern_assign_kernel(const int* lids, const int64_t* rowptr, const int* colind, int64_t nelem, int N,
                                      const uint64_t* hash, const uint64_t* uhash, int64_t npat, const uint16_t* pat,
                                      int* pat_id, int* bad) {
  const int64_t e = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (e >= nelem) return;
  const uint64_t h = hash[e];
  int64_t lo = 0, hi = npat;
  while (lo < hi) { const int64_t mid = (lo + hi) >> 1; if (uhash[mid] < h) lo = mid + 1; else hi = mid; }
  pat_id[e] = (int)lo;
  const int* l = lids + e * N;
  const uint16_t* P = pat + lo * N * N;
  for (int r = 0; r < N; ++r)
    for (int c = 0; c < N; ++c)
      if ((int)P[r * N + c] != rel_slot(rowptr, colind, l[r], l[c])) atomicExch(bad, 2);
}

// updateJacDBC (assemblyManager.cpp:1167-1180): replaceLocalValues(dof, 1, 1.0, dof)
__global__ void dbc_diag_kernel(const int64_t* slots, int64_t n, double* jac) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n && slots[i] >= 0) jac[slots[i]] = 1.0;
}
__global__ void dbc_slot_kernel(const int* dbc, int64_t n, const int64_t* rowptr, const int* colind, int64_t* slots) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const int s = rel_slot(rowptr, colind, dbc[i], dbc[i]);
  slots[i] = s < 0 ? -1 : rowptr[dbc[i]] + s;
}
// fix zero rows (assemblyManager.cpp:2644-2663): rows whose |.|-sum is below 1e-14 get a unit diagonal
__global__ void fix_zero_rows_kernel(const int64_t* rowptr, const int* colind, int64_t nrows, double* jac) {
  const int64_t r = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (r >= nrows) return;
  double s = 0.0;
  for (int64_t p = rowptr[r]; p < rowptr[r + 1]; ++p) s += fabs(jac[p]);
  if (s < 1.0e-14) { const int d = rel_slot(rowptr, colind, (int)r, (int)r); if (d >= 0) jac[rowptr[r] + d] = 1.0; }
}

// identity rows of the Dirichlet projection system (setDirichlet, assemblyManager.cpp:1926-1943): non-fixed rows get a unit diagonal
__global__ void projection_identity_kernel(const int64_t* rowptr, const int* colind, const uint8_t* fixed, int64_t nrows, double* mass) {
  const int64_t r = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (r >= nrows || fixed[r]) return;
  const int d = rel_slot(rowptr, colind, (int)r, (int)r);
  if (d >= 0) mass[rowptr[r] + d] = 1.0;
}
// u[i] = x[i] on fixed DOFs (SolverManager::setDirichlet, solverManager.cpp:1919-1932)
__global__ void copy_fixed_kernel(const uint8_t* fixed, const double* x, double* u, int64_t n) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n && fixed[i]) u[i] = x[i];
}

inline unsigned blocks_for(int64_t n, int t) { return (unsigned)((n + t - 1) / t); }

// ---- static scratch of the allocation-free entry point (mrh_assemble_res_host_static) --------------------------------
// File-scope device array: every GPU that loads the library holds these 16 MiB until exit.  In HOST code the name is only
// a stand-in for the symbol, not a device address: the address is fetched with cudaGetSymbolAddress, once per device with
// that device current (static_scratch_of, called from device_finalize), and only that pointer is handed to kernels.
// Layout: [ u : MRH_STATIC_SCRATCH_MAX_DOFS | res : MRH_STATIC_SCRATCH_MAX_DOFS ] doubles.
static __device__ __align__(256) double g_static_scratch[2 * (size_t)MRH_STATIC_SCRATCH_MAX_DOFS];
constexpr int MRH_MAX_DEVICES = 64;
double* g_static_scratch_addr[MRH_MAX_DEVICES] = {};

double* static_scratch_of(int device) {
  if (device < 0 || device >= MRH_MAX_DEVICES) throw std::runtime_error("CUDA error: device ordinal beyond the static scratch table");
  if (!g_static_scratch_addr[device]) {
    CUDA_CK(cudaSetDevice(device));
    void* ptr = nullptr;
    const cudaError_t rc = cudaGetSymbolAddress(&ptr, g_static_scratch);
    if (rc != cudaSuccess || !ptr) throw std::runtime_error(std::string("CUDA error: cudaGetSymbolAddress(static scratch) failed: ") + cudaGetErrorString(rc));
    g_static_scratch_addr[device] = static_cast<double*>(ptr);
  }
  return g_static_scratch_addr[device];
}

}  // namespace

#include "engine_tiles.inl"
#include "exchange.inl"
#include "engine_launch.inl"
#include "solver.inl"
