// Fused tile kernel, 128-column tiles (see kernels_fused_impl.cuh).
#define TVS_HALVES 1
#define TVS_FUSED_SUFFIX _w128
#include "kernels_fused_impl.cuh"
