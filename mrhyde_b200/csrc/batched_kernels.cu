// Batched fused assembly kernel (volume terms, residual + Jacobian): the CUDA wrapper around the phase functions of
// assemble_batched.h.  Own translation unit like generic_kernels.cu.
#include <cuda_runtime.h>

#include <algorithm>
#include <stdexcept>
#include <string>

#include "assemble_batched.h"
#include "dispatch.h"
#include "generic_kernels.h"

using namespace mrh;

#define CUDA_CK(call)                                                                                   \
  do { cudaError_t e_ = (call); if (e_ != cudaSuccess)                                                  \
         throw std::runtime_error(std::string("CUDA error: ") + cudaGetErrorString(e_) + " at " #call); } while (0)

namespace {

template <class C>
__global__ void __launch_bounds__(32 * MRH_BATCHED_MAXW) batched_assembly_kernel(const __grid_constant__ KParams kp) {
  extern __shared__ __align__(16) double smem_all[];
  batched_body<C, InterpProgs>(kp, smem_all);
}

template <class C>
bool launch_batched(int num_sms, KParams& kp, cudaStream_t stream) {
  int nw; size_t smem;
  if (!batched_shape<C>(kp.tab.nq, kp.tp.transient != 0, nw, smem)) return false;      // caller falls back to the generic kernel
  auto kern = batched_assembly_kernel<C>;
  CUDA_CK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  int per_sm = 1;
  CUDA_CK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kern, 32 * nw, smem));
  if (per_sm < 1) per_sm = 1;
  kp.elems_per_block = nw;
  const int64_t nblk = (kp.nwork + nw - 1) / nw;
  const unsigned grid = (unsigned)std::min<int64_t>(nblk, (int64_t)num_sms * per_sm);
  kern<<<grid, 32 * nw, smem, stream>>>(kp);
  CUDA_CK(cudaGetLastError());
  return true;
}

}  // namespace

namespace mrh {
bool launch_batched_assembly(int phys, int dim, int order, int num_sms, KParams& kp, void* stream) {
  bool launched = false;
  dispatch_cfg(phys, dim, order, [&](auto c) {
    using C = decltype(c);
    launched = launch_batched<C>(num_sms, kp, (cudaStream_t)stream);
  });
  return launched;
}
}  // namespace mrh
