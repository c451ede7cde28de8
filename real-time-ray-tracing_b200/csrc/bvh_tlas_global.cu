// TLAS build for scenes with more than 1024 batches (generalisation rule B7) — placeholder until the
// multi-CTA global path lands; reports a capacity error instead of producing a wrong hierarchy.
#include "context.h"
namespace rtrt {
int launch_build_tlas_global(Context& c) {
    return c.fail(RTRT_ERR_CAPACITY, "rtrt_build_bvh: more than 1024 batches not supported yet");
}
}  // namespace rtrt
