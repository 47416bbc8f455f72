// compile-only probe: register / spill / code size of the specialised kernel without rebuilding the whole library
#include "../mrhyde_b200/csrc/assemble_fast.h"
using namespace mrh;
template <bool T, bool A, bool J>
__global__ void __launch_bounds__(128) probe(const __grid_constant__ KParams kp) {
  __shared__ double stage[ThermalHexQ1<T, A, J>::NSTAGE * 8 * 128];
  for (int64_t e = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; e < kp.nwork; e += (int64_t)gridDim.x * blockDim.x)
    ThermalHexQ1<T, A, J>::element(kp, e, stage + threadIdx.x, 128);
}
template __global__ void probe<false, false, true>(const __grid_constant__ KParams);
