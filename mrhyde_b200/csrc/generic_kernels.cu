// Generic fused assembly kernel (any of the four physics, state-dependent coefficients, volume or side launch):
// the CUDA wrapper around the phase functions of assemble.h.  Own translation unit: its template instantiations
// dominate compile time and rarely change.
#include <cuda_runtime.h>

#include <algorithm>
#include <stdexcept>
#include <string>
#include <type_traits>

#include "dispatch.h"
#include "generic_kernels.h"

using namespace mrh;

#define CUDA_CK(call)                                                                                   \
  do { cudaError_t e_ = (call); if (e_ != cudaSuccess)                                                  \
         throw std::runtime_error(std::string("CUDA error: ") + cudaGetErrorString(e_) + " at " #call); } while (0)

namespace {

template <class C, bool JAC>
__global__ void __launch_bounds__(128) generic_assembly_kernel(const __grid_constant__ KParams kp) {
  extern __shared__ double sm[];
  using G = GenPhases<C>;
  using S = typename std::conditional<JAC, PAD<C::P>, double>::type;
  const int EB = kp.elems_per_block, tid = threadIdx.x, nt = blockDim.x;
  const GenSmem<C> L(EB, kp.tab.nq, JAC, kp.tp.transient != 0);
  for (int64_t blk0 = (int64_t)blockIdx.x * EB; blk0 < kp.nwork; blk0 += (int64_t)gridDim.x * EB) {
    for (int it = tid; it < EB * C::N; it += nt) G::gather_dof(kp, L, sm, blk0, it);
    for (int it = tid; it < EB * C::NV; it += nt) G::gather_vertex(kp, L, sm, blk0, it);
    __syncthreads();
    for (int it = tid; it < EB * L.nq; it += nt) G::geometry(kp, L, sm, blk0, it);
    __syncthreads();
    for (int it = tid; it < EB * L.nq; it += nt) G::template point<S>(kp, L, sm, blk0, it);
    __syncthreads();
    for (int it = tid; it < EB * C::NB * C::NVAR; it += nt) G::contract(kp, L, sm, blk0, it);
    __syncthreads();
  }
}

// L2-projection launch (GenPhases::projection): gather of the element's LIDs and vertices, geometry / basis at the points, then
// one work item per (element, variable, basis function)
template <class C>
__global__ void __launch_bounds__(128) generic_projection_kernel(const __grid_constant__ KParams kp) {
  extern __shared__ double sm[];
  using G = GenPhases<C>;
  const int EB = kp.elems_per_block, tid = threadIdx.x, nt = blockDim.x;
  const GenSmem<C> L(EB, kp.tab.nq, false, false);
  for (int64_t blk0 = (int64_t)blockIdx.x * EB; blk0 < kp.nwork; blk0 += (int64_t)gridDim.x * EB) {
    for (int it = tid; it < EB * C::N; it += nt) {
      const int le = it / C::N, k = it % C::N;
      if (blk0 + le < kp.nwork) reinterpret_cast<int*>(sm + L.oLid())[le * C::N + k] = kp.lids[G::elem_of(kp, blk0 + le) * C::N + k];
    }
    for (int it = tid; it < EB * C::NV; it += nt) G::gather_vertex(kp, L, sm, blk0, it);
    __syncthreads();
    for (int it = tid; it < EB * L.nq; it += nt) G::geometry(kp, L, sm, blk0, it);
    __syncthreads();
    for (int it = tid; it < EB * C::NB * C::NVAR; it += nt) G::projection(kp, L, sm, blk0, it);
    __syncthreads();
  }
}

template <class C>
void launch_projection(int num_sms, KParams& kp, cudaStream_t stream) {
  const size_t per_elem = GenSmem<C>(1, kp.tab.nq, false, false).total_doubles() * sizeof(double);
  int EB = std::max(1, std::min(16, (int)((100 * 1024) / per_elem)));
  const size_t smem = GenSmem<C>(EB, kp.tab.nq, false, false).total_doubles() * sizeof(double);
  if (smem > 227 * 1024) throw std::runtime_error("element block does not fit in shared memory (quadrature too large)");
  auto kern = generic_projection_kernel<C>;
  CUDA_CK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  kp.elems_per_block = EB;
  const int64_t nblk = (kp.nwork + EB - 1) / EB;
  const unsigned grid = (unsigned)std::min<int64_t>(nblk, (int64_t)num_sms * 4);
  kern<<<grid, 128, smem, stream>>>(kp);
  CUDA_CK(cudaGetLastError());
}

template <class C, bool JAC>
void launch_generic(int num_sms, KParams& kp, cudaStream_t stream) {
  const size_t per_elem = GenSmem<C>(1, kp.tab.nq, JAC, kp.tp.transient != 0).total_doubles() * sizeof(double);
  int EB = (int)((100 * 1024) / per_elem);
  if (EB < 1) EB = 1;
  if (EB > 16) EB = 16;
  const size_t smem = GenSmem<C>(EB, kp.tab.nq, JAC, kp.tp.transient != 0).total_doubles() * sizeof(double);
  if (smem > 227 * 1024) throw std::runtime_error("element block does not fit in shared memory (quadrature too large)");
  auto kern = generic_assembly_kernel<C, JAC>;
  CUDA_CK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  int per_sm = 1;
  CUDA_CK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kern, 128, smem));
  if (per_sm < 1) per_sm = 1;
  kp.elems_per_block = EB;
  const int64_t nblk = (kp.nwork + EB - 1) / EB;
  const unsigned grid = (unsigned)std::min<int64_t>(nblk, (int64_t)num_sms * per_sm);
  kern<<<grid, 128, smem, stream>>>(kp);
  CUDA_CK(cudaGetLastError());
}

}  // namespace

namespace mrh {
bool launch_generic_assembly(int phys, int dim, int order, bool jac, int num_sms, KParams& kp, void* stream) {
  return dispatch_cfg(phys, dim, order, [&](auto c) {
    using C = decltype(c);
    if (jac) launch_generic<C, true>(num_sms, kp, (cudaStream_t)stream); else launch_generic<C, false>(num_sms, kp, (cudaStream_t)stream);
  });
}
bool launch_generic_projection(int phys, int dim, int order, int num_sms, KParams& kp, void* stream) {
  return dispatch_cfg(phys, dim, order, [&](auto c) { launch_projection<decltype(c)>(num_sms, kp, (cudaStream_t)stream); });
}
}  // namespace mrh
