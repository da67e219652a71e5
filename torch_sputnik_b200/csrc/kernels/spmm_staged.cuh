// spmm_staged.cuh -- placeholder until the staged SpMM kernel lands.
#pragma once
#include "spmm_gather.cuh"
namespace tsp {
inline bool spmm_staged_applicable(const SpmmArgs&, int) { return false; }
inline int launch_spmm_staged(const SpmmArgs&, int, cudaStream_t) { return -1; }
}  // namespace tsp
