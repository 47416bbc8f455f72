// Batched fused assembly kernel (volume terms, residual + Jacobian): the CUDA wrapper around the phase functions of
// assemble_batched.h.  Own translation unit like generic_kernels.cu.
#include <cuda_runtime.h>

#include <algorithm>
#include <stdexcept>
#include <string>

#if defined(MRH_BAT_TU_PHYS)
#define MRH_DISPATCH_ONLY_PHYS MRH_BAT_TU_PHYS
#endif
#include "assemble_batched.h"
#include "dispatch.h"
#include "generic_kernels.h"

using namespace mrh;

#define CUDA_CK(call)                                                                                   \
  do { cudaError_t e_ = (call); if (e_ != cudaSuccess)                                                  \
         throw std::runtime_error(std::string("CUDA error: ") + cudaGetErrorString(e_) + " at " #call); } while (0)

namespace {

// register budget per configuration (measured: capping the 3-D Navier-Stokes point phase below its 255 registers to fit
// 10 warps per SM instead of 8 spills 0.5 KB per thread and is slower: 29.2 ms instead of 23.3-25.7 ms at 128^3)
// Default: two CTAs of up to 192 threads per SM, i.e. at most 168 registers (12 warps per SM); the Navier-Stokes point
// phase is left its full register budget.
template <class C> struct BatBounds { static constexpr int THREADS = 32 * MRH_BATCHED_MAXW, MINB = (C::PHYS == PHYS_NS) ? 1 : 2; };
#if MRH_NS_HALF_CHUNKS
template <> struct BatBounds<GenCfg<PHYS_NS, 3, 1>> { static constexpr int THREADS = 160, MINB = 2; };   // <= 204 registers: 10 warps per SM
#endif

// SIDES: the side-launch branches (side measure, normals, boundary() and Flux terms) exist only in the instantiation that
// serves the side launches; the volume instantiation compiles them away
template <class C, bool SIDES>
__global__ void __launch_bounds__(BatBounds<C>::THREADS, BatBounds<C>::MINB) batched_assembly_kernel(const __grid_constant__ KParams kp) {
  extern __shared__ __align__(16) double smem_all[];
  batched_body<C, InterpProgs, SIDES>(kp, smem_all);
}

template <class C, bool SIDES>
bool launch_batched(int num_sms, KParams& kp, cudaStream_t stream) {
  auto kern = batched_assembly_kernel<C, SIDES>;
  const bool tr = kp.tp.transient != 0;
  static int set_smem = 0;                                   // per instantiation: largest opt-in size requested so far
  const int want = (int)std::min<size_t>(batched_smem<C>(kp.tab.nq, tr, BatBounds<C>::THREADS / 32), (size_t)227 * 1024);
  if (want > set_smem) { CUDA_CK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, want)); set_smem = want; }
  int per_sm = 0;
  const int nw = batched_pick_warps<C>(kp.tab.nq, tr, [&](int threads, size_t smem) {
    if (threads > BatBounds<C>::THREADS) return 0;
    int b = 0; return cudaOccupancyMaxActiveBlocksPerMultiprocessor(&b, kern, threads, smem) == cudaSuccess ? b : 0; }, per_sm);
  if (nw == 0) return false;                                 // caller falls back to the generic kernel
  kp.elems_per_block = nw;
  const int64_t nblk = (kp.nwork + nw - 1) / nw;
  const unsigned grid = (unsigned)std::min<int64_t>(nblk, (int64_t)num_sms * per_sm);
  kern<<<grid, 32 * nw, batched_smem<C>(kp.tab.nq, tr, nw), stream>>>(kp);
  CUDA_CK(cudaGetLastError());
  return true;
}

}  // namespace

// The file is compiled once per physics module for the volume instantiations (MRH_BAT_TU_PHYS = 0..6) and once more for the
// side-launch instantiations and the entry point (build.py; all in parallel).  Not only for build time: what the compiler makes of one
// kernel depended on what else was in its module - with all 24 configurations in one translation unit the 3-D Navier-Stokes volume
// kernel lost its per-lane-chunk specialisation of the point phase (30 600 -> 16 500 SASS lines, 24.2 -> 28.5 ms at 128^3, same source).
#define MRH_CAT2(a, b) a##b
#define MRH_CAT(a, b) MRH_CAT2(a, b)
namespace mrh {
#if defined(MRH_BAT_TU_PHYS)
bool MRH_CAT(launch_batched_volume_p, MRH_BAT_TU_PHYS)(int dim, int order, int num_sms, KParams& kp, void* stream) {
  bool launched = false;
  dispatch_cfg(MRH_BAT_TU_PHYS, dim, order, [&](auto c) {
    using C = decltype(c);
    launched = launch_batched<C, false>(num_sms, kp, (cudaStream_t)stream);
  });
  return launched;
}
#else
bool launch_batched_volume_p0(int, int, int, KParams&, void*); bool launch_batched_volume_p1(int, int, int, KParams&, void*);
bool launch_batched_volume_p2(int, int, int, KParams&, void*); bool launch_batched_volume_p3(int, int, int, KParams&, void*);
bool launch_batched_volume_p4(int, int, int, KParams&, void*); bool launch_batched_volume_p5(int, int, int, KParams&, void*);
bool launch_batched_volume_p6(int, int, int, KParams&, void*);
bool launch_batched_assembly(int phys, int dim, int order, int num_sms, KParams& kp, void* stream) {
  if (kp.is_side) {
    bool launched = false;
    dispatch_cfg(phys, dim, order, [&](auto c) {
      using C = decltype(c);
      launched = launch_batched<C, true>(num_sms, kp, (cudaStream_t)stream);
    });
    return launched;
  }
  switch (phys) {
    case PHYS_THERMAL: return launch_batched_volume_p0(dim, order, num_sms, kp, stream);
    case PHYS_CDR: return launch_batched_volume_p1(dim, order, num_sms, kp, stream);
    case PHYS_LE: return launch_batched_volume_p2(dim, order, num_sms, kp, stream);
    case PHYS_NS: return launch_batched_volume_p3(dim, order, num_sms, kp, stream);
    case PHYS_STOKES: return launch_batched_volume_p4(dim, order, num_sms, kp, stream);
    case PHYS_HELMHOLTZ: return launch_batched_volume_p5(dim, order, num_sms, kp, stream);
    case PHYS_BURGERS: return launch_batched_volume_p6(dim, order, num_sms, kp, stream);
    default: return false;
  }
}
#endif
}  // namespace mrh
