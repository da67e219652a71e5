// sddmm_staged.cuh -- placeholder until the staged SDDMM kernel lands.
#pragma once
#include "sddmm_gather.cuh"
namespace tsp {
inline bool sddmm_staged_applicable(const SddmmArgs&, int) { return false; }
inline int launch_sddmm_staged(const SddmmArgs&, int, cudaStream_t) { return -1; }
}  // namespace tsp
