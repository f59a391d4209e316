// pmps_ppe.cu — placeholder until the PPE kernels land (next commit)
#include "pmps_device.cuh"
int pmps_ppe_solve(pmps_handle* h) { return pmps_fail(h, PMPS_E_UNSUPPORTED, "pressure Poisson solve not built yet"); }
