// Temporary stubs (replaced as the kernels land).
#include "large.hpp"
namespace aslam {
aslam_status large_ukf_create(aslam_large*) { set_error("ukf large: not built yet"); return ASLAM_ERR_UNSUPPORTED; }
void large_ukf_destroy(aslam_large*) {}
aslam_status large_ukf_step(aslam_large*, double, double, float, double, double, float, uint32_t) { return ASLAM_ERR_UNSUPPORTED; }
aslam_status large_ukf_load_z(aslam_large*, const double*) { return ASLAM_ERR_UNSUPPORTED; }
aslam_status large_ukf_status(aslam_large*, uint32_t*, int32_t*) { return ASLAM_ERR_UNSUPPORTED; }
}  // namespace aslam
