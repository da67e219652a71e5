// sddmm_staged.cuh -- entry points of the non-gather SDDMM variants (see sddmm_tiled.cuh).
#pragma once
#include "sddmm_tiled.cuh"
