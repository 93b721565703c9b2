// bsp_numeric_f32_tc.cu -- float32 tile products on the 5th-generation tensor cores (tcgen05).
// Placeholder until the kernel lands: reports "no kernel for this k" so the caller falls back
// to the fp32 SIMT kernel.
#include "bsp_internal.h"

int bsp_numeric_f32_tc(bsp_ctx *ctx, const bsp_mat *A, const bsp_mat *, const bsp_symbolic_out *, float *, int32_t *)
{
    snprintf(ctx->err, sizeof(ctx->err), "no float32 tensor-core kernel for k = %d", A->k);
    return BSP_ERR_LIMIT;
}
