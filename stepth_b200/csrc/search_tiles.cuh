// search_tiles.cuh -- K1 production search (placeholder wiring: dispatches to the in-order kernel until the
// tiled kernel lands in the next commit).
#pragma once
#include "ctx.cuh"
#include "search_literal.cuh"

namespace stepth {

inline int32_t launch_search_tiles_dense(stepth_ctx *ctx, cudaStream_t st, const SearchArgs &a, uint32_t n_pairs) {
    dim3 grid((a.n_queries + 255) / 256, n_pairs);
    search_literal_kernel<<<grid, 256, 0, st>>>(a);
    ctx->launches++;
    STEPTH_CUDA_TRY(ctx, cudaGetLastError());
    return STEPTH_OK;
}

}  // namespace stepth
