// Run-time (physics, dim, order) -> compile-time kernel configuration.
#pragma once
#include "assemble.h"

namespace mrh {

template <class F>
inline bool dispatch_cfg(int phys, int dim, int order, F&& f) {
  // MRH_DISPATCH_ONLY_PHYS (a translation unit per physics module, see batched_kernels.cu): only that module's cases are instantiated
#if defined(MRH_DISPATCH_ONLY_PHYS)
#define MRH_CASE(PH, D, O) if constexpr (PH == MRH_DISPATCH_ONLY_PHYS) { if (phys == PH && dim == D && order == O) { f(GenCfg<PH, D, O>{}); return true; } }
#else
#define MRH_CASE(PH, D, O) if (phys == PH && dim == D && order == O) { f(GenCfg<PH, D, O>{}); return true; }
#endif
  MRH_CASE(PHYS_THERMAL, 2, 1) MRH_CASE(PHYS_THERMAL, 2, 2) MRH_CASE(PHYS_THERMAL, 3, 1) MRH_CASE(PHYS_THERMAL, 3, 2)
  MRH_CASE(PHYS_CDR, 2, 1) MRH_CASE(PHYS_CDR, 2, 2) MRH_CASE(PHYS_CDR, 3, 1) MRH_CASE(PHYS_CDR, 3, 2)
  MRH_CASE(PHYS_LE, 2, 1) MRH_CASE(PHYS_LE, 2, 2) MRH_CASE(PHYS_LE, 3, 1)
  MRH_CASE(PHYS_NS, 2, 1) MRH_CASE(PHYS_NS, 2, 2) MRH_CASE(PHYS_NS, 3, 1)
  MRH_CASE(PHYS_STOKES, 2, 1) MRH_CASE(PHYS_STOKES, 2, 2) MRH_CASE(PHYS_STOKES, 3, 1)
  MRH_CASE(PHYS_HELMHOLTZ, 2, 1) MRH_CASE(PHYS_HELMHOLTZ, 2, 2) MRH_CASE(PHYS_HELMHOLTZ, 3, 1)
  MRH_CASE(PHYS_BURGERS, 2, 1) MRH_CASE(PHYS_BURGERS, 2, 2) MRH_CASE(PHYS_BURGERS, 3, 1) MRH_CASE(PHYS_BURGERS, 3, 2)
#undef MRH_CASE
  return false;
}

}  // namespace mrh
