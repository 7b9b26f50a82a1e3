/* spk_render.cu — placeholder until the path tracer lands (next commit). */
#include "spk_context.h"
using namespace spk;
extern "C"
{
int spk_cuda_render(SpkContext *, const SpkRenderParams *)
{
    set_error("spk_cuda_render: not built yet");
    return SPK_ERR_UNSUPPORTED;
}
int spk_cuda_resolve(SpkContext *, uint32_t, uint32_t, void *, size_t)
{
    set_error("spk_cuda_resolve: not built yet");
    return SPK_ERR_UNSUPPORTED;
}
}
