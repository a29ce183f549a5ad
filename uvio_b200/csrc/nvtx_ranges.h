// NVTX ranges around the host-side phases of a feed (visible in Nsight Systems timelines; no cost when no tool is attached:
// nvtx3 is header-only and resolves its injection library lazily).  The host simulation (tests/hostsim) compiles them away.
#pragma once
#if !defined(UVIO_NO_GRAPHS) && __has_include(<nvtx3/nvToolsExt.h>)
#include <nvtx3/nvToolsExt.h>
namespace uvio {
struct NvtxRange {
  explicit NvtxRange(const char *name) { nvtxRangePushA(name); }
  ~NvtxRange() { nvtxRangePop(); }
  NvtxRange(const NvtxRange &) = delete;
  NvtxRange &operator=(const NvtxRange &) = delete;
};
}  // namespace uvio
#define UVIO_NVTX(name) ::uvio::NvtxRange uvio_nvtx_range_##__LINE__(name)
#else
#define UVIO_NVTX(name) do { } while (0)
#endif
