// mg_kernels_fast.cuh -- specialised stage kernels for the hot configurations (scalar real
// (2d+1)-point star stencils: 2D/3D Poisson).  Each fast_* hook returns true when it issued the
// stage; false falls through to the generic kernel.  Results are bit-identical to the generic
// kernels (same operand order, no FMA contraction).
#pragma once
#include "mg_kernels_generic.cuh"
#include "mg_workspace.cuh"

namespace mg {

template <typename T, int D, int F>
inline bool fast_smooth(mg_workspace*, Level&, Smoother&) { return false; }
template <typename T, int D, int F>
inline bool fast_residual(mg_workspace*, Level&) { return false; }
template <typename T, int D, int F>
inline bool fast_residual_restrict(mg_workspace*, Level&, Level&) { return false; }
template <typename T, int D, int F>
inline bool fast_prolong_add(mg_workspace*, Level&, Level&, const FieldSet<T>&, double) { return false; }
template <typename T, int D, int F>
inline bool fast_norm(mg_workspace*, Level&) { return false; }

}  // namespace mg
