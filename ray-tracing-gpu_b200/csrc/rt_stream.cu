// rt_stream.cu -- streamed (tiled brute-force) path and LBVH entry points.
#include "rt_context.h"

namespace rtb {
int render_streamed(rtb200_ctx* c, const rtb200_frame*, const FrameParams&, int, bool, rtb200_stats*, uint32_t*)
{
    c->last_error = "streamed path not available in this build";
    return RTB200_E_STATE;
}
}

extern "C" int rtb200_build_bvh(rtb200_ctx* ctx)
{
    if (!ctx) return RTB200_E_ARG;
    ctx->last_error = "LBVH not available in this build";
    return RTB200_E_STATE;
}
