#pragma once
#include "params.h"
namespace mrh {
// launches the generic fused kernel for (phys, dim, order); false when no such configuration is compiled in.
// Throws std::runtime_error("CUDA error: ...") on launch failure.
bool launch_generic_assembly(int phys, int dim, int order, bool jac, int num_sms, KParams& kp, void* stream);
// launches the L2-projection kernel (GenPhases::projection; kp.ph.proj selects Dirichlet-side or initial-volume mode)
bool launch_generic_projection(int phys, int dim, int order, int num_sms, KParams& kp, void* stream);
// launches the batched volume kernel (residual + Jacobian, assemble_batched.h); false when the configuration is not
// compiled in or one element does not fit the shared-memory layout (the caller then uses the generic kernel).
bool launch_batched_assembly(int phys, int dim, int order, int num_sms, KParams& kp, void* stream);
}
