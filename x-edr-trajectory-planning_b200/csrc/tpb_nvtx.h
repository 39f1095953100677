// NVTX ranges around the stages of the hot path (the reference has no
// tracing; a timeline tool shows these as "tpb:<stage>" on the launching thread).
// nvtx3 is header-only: without an attached tool a range is two no-op calls.
#pragma once

#include <nvtx3/nvToolsExt.h>

namespace tpb {
struct NvtxRange {
  explicit NvtxRange(const char* name) { nvtxRangePushA(name); }
  ~NvtxRange() { nvtxRangePop(); }
  NvtxRange(const NvtxRange&) = delete;
  NvtxRange& operator=(const NvtxRange&) = delete;
};
}  // namespace tpb

#define TPB_NVTX_CAT2(a, b) a##b
#define TPB_NVTX_CAT(a, b) TPB_NVTX_CAT2(a, b)
#define TPB_NVTX(name) ::tpb::NvtxRange TPB_NVTX_CAT(tpb_nvtx_range_, __LINE__)(name)
