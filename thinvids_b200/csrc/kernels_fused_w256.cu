// Fused tile kernel, 256-column tiles (see kernels_fused_impl.cuh).
#define TVS_HALVES 2
#define TVS_FUSED_SUFFIX _w256
#include "kernels_fused_impl.cuh"
