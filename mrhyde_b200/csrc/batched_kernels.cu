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

constexpr int MAXW = 4;       // warps per CTA (fewer when one element's shared-memory block is large)

template <class C>
__global__ void __launch_bounds__(32 * MAXW) batched_assembly_kernel(const __grid_constant__ KParams kp) {
  extern __shared__ __align__(16) double smem_all[];
  using G = GenPhases<C>;
  using B = BatPhases<C>;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nwarp = blockDim.x >> 5;
  const WarpSmem<C> L(kp.tab.nq, kp.tp.transient != 0);
  int* offs = reinterpret_cast<int*>(smem_all);                        // offsets(var, basis) -> element slot, CTA-shared
  for (int k = threadIdx.x; k < C::N; k += blockDim.x) offs[k] = kp.offsets[k];
  __syncthreads();
  double* sm = smem_all + even_up((C::N + 1) / 2) + (size_t)warp * L.total_doubles();
  const int nq = L.nq, nparts = B::geom_parts(nq), n1 = B::num_kind1(kp);
  for (int64_t w = (int64_t)blockIdx.x * nwarp + warp; w < kp.nwork; w += (int64_t)gridDim.x * nwarp) {
    for (int it = lane; it < C::N; it += 32) G::gather_dof(kp, L, sm, w, it);
    for (int it = lane; it < C::NV; it += 32) G::gather_vertex(kp, L, sm, w, it);
    __syncwarp();
    for (int it = lane; it < nq * nparts; it += 32) B::geometry(kp, L, sm, it, nparts);
    __syncwarp();
    for (int it = lane; it < nq * n1; it += 32) B::prep(kp, L, sm, it, n1);
    if (lane == 31) B::element_size(L, sm);
    __syncwarp();
    for (int it = lane; it < nq * B::NCH; it += 32) B::point_chunk(kp, L, sm, it);
    __syncwarp();
    typename B::Acc acc;
    B::contract(L, sm, acc);
    __syncwarp();                                          // every lane is done with D: K may take its place
    B::store_acc(L, sm, acc);
    __syncwarp();
    B::row_setup(kp, L, sm, offs, lane);
    __syncwarp();
    B::scatter(kp, L, sm, offs, G::elem_of(kp, w), lane);
    __syncwarp();                                          // the next element reuses the block
  }
}

template <class C>
bool launch_batched(int num_sms, KParams& kp, cudaStream_t stream) {
  const size_t per = WarpSmem<C>(kp.tab.nq, kp.tp.transient != 0).total_doubles() * sizeof(double);
  const size_t head = even_up((C::N + 1) / 2) * sizeof(double);
  int nw = 0; size_t best = 0;                                         // warps per CTA that put the most warps on one SM
  for (int c = MAXW; c >= 1; --c) {
    const size_t total = ((size_t)227 * 1024 / (head + c * per + 1024)) * c;
    if (total > best) { best = total; nw = c; }
  }
  if (nw == 0) return false;                                           // caller falls back to the generic kernel
  const size_t smem = head + nw * per;
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
