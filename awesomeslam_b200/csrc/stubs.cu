// Temporary stubs (replaced as the kernels land).
#include "internal.hpp"
namespace aslam {
aslam_status launch_ukf_batched(const aslam_config&, const BatchState&, const BatchInput&, cudaStream_t) {
    set_error("ukf batched: not built yet");
    return ASLAM_ERR_UNSUPPORTED;
}
}  // namespace aslam
