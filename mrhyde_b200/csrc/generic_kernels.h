#pragma once
#include "params.h"
namespace mrh {
// launches the generic fused kernel for (phys, dim, order); false when no such configuration is compiled in.
// Throws std::runtime_error("CUDA error: ...") on launch failure.
bool launch_generic_assembly(int phys, int dim, int order, bool jac, int num_sms, KParams& kp, void* stream);
}
