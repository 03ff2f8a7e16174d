// Placeholder until the fused tile kernel lands: every shape uses the two-pass path.
#include "tvs_internal.h"
namespace tvs {
struct FusedPlan {};
FusedPlan *fused_plan_create(const Plan &, const AxisTable[4], int, std::string &err) { err = "not built"; return nullptr; }
void fused_plan_destroy(FusedPlan *p) { delete p; }
size_t fused_plan_device_bytes(const FusedPlan *) { return 0; }
int fused_launch(const FusedPlan *, const Plan &, const SampleView[], const SampleView[], int, cudaStream_t) { return -1; }
}  // namespace tvs
