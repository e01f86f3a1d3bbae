// temporary: K10/K12/K13/ReSTIR entry points until their kernels land
#include "rt2_internal.cuh"
namespace rt2 {
int lights_prepare(rt2_ctx* ctx) { return fail(ctx, RT2_E_STATE, "prepare_lights not built yet"); }
int mips_generate(rt2_ctx* ctx, int) { return fail(ctx, RT2_E_STATE, "mip_pass not built yet"); }
int post_process(rt2_ctx* ctx, uint32_t) { return fail(ctx, RT2_E_STATE, "postprocess not built yet"); }
void passes_release(rt2_ctx*) {}
int render_restir(rt2_ctx* ctx, uint32_t, const rt2_launch_params&) { return fail(ctx, RT2_E_STATE, "ReSTIR passes not built yet (set refrence_mode = 1)"); }
bool pass_buffer(rt2_ctx*, int, void**, size_t*) { return false; }
}
