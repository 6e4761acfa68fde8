// NVTX ranges around the public step calls ("tracing"): a timeline tool (Nsight Systems) shows
// one named range per filter step / replay launch / laser batch on the calling thread. nvtx3 is header-only; without a
// tool attached a push/pop pair is a couple of indirect calls through a null-initialised table.
#pragma once
#include <nvtx3/nvToolsExt.h>

namespace aslam {
struct NvtxRange {
    explicit NvtxRange(const char* name) { nvtxRangePushA(name); }
    ~NvtxRange() { nvtxRangePop(); }
    NvtxRange(const NvtxRange&) = delete;
    NvtxRange& operator=(const NvtxRange&) = delete;
};
}  // namespace aslam
#define ASLAM_NVTX(name) ::aslam::NvtxRange aslam_nvtx_range_(name)
