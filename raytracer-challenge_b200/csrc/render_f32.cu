// F32 FAST kernel — placeholder translation unit until the fast path lands (next commit).
#include "rtc_common.h"
namespace rtc {
cudaError_t launch_render_f32(const DevScene32&, const CamParams&, void*, uint8_t*, Counters*, cudaStream_t) {
    return cudaErrorNotSupported;
}
}  // namespace rtc
