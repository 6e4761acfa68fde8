// Temporary stubs (replaced as the kernels land).
#include "internal.hpp"
namespace aslam {
aslam_status launch_ukf_batched(const aslam_config&, const BatchState&, const BatchInput&, cudaStream_t) {
    set_error("ukf batched: not built yet");
    return ASLAM_ERR_UNSUPPORTED;
}
}  // namespace aslam
#define STUB { aslam::set_error("large-map: not built yet"); return ASLAM_ERR_UNSUPPORTED; }
extern "C" {
aslam_status aslam_large_ekf_create(const aslam_config*, uint32_t, uint32_t, uint32_t, int32_t, aslam_large**) STUB
aslam_status aslam_large_ukf_create(const aslam_config*, uint32_t, int32_t, aslam_large**) STUB
aslam_status aslam_large_destroy(aslam_large*) STUB
aslam_status aslam_large_sync(aslam_large*) STUB
void* aslam_large_stream(aslam_large*) { return nullptr; }
aslam_status aslam_large_load_state(aslam_large*, const double*, const double*, const double*, size_t) STUB
aslam_status aslam_large_get_state(aslam_large*, double*, double*, size_t) STUB
aslam_status aslam_large_step(aslam_large*, double, double, float, double, double, float, const float*, const float*,
                              uint32_t, int32_t*) STUB
aslam_status aslam_large_step_phase1(aslam_large*, double, double, float, double, double, float, const float*,
                                     const float*, uint32_t, uint32_t*) STUB
aslam_status aslam_large_exchange_buffer(aslam_large*, double**, size_t*) STUB
aslam_status aslam_large_step_phase2(aslam_large*, int32_t*) STUB
uint64_t aslam_large_launch_count(aslam_large*) { return 0; }
aslam_status aslam_large_last_kernel_ms(aslam_large*, float*) STUB
aslam_status aslam_large_status(aslam_large*, uint32_t*, int32_t*) STUB
}
