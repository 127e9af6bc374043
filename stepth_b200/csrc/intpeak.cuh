// intpeak.cuh -- measured integer-ALU peak of this device for the search's instruction class (the
// `P_INT` of t_roof = max(B_alg / BW_HBM, O_alg / P_INT)).  The ring search is 32-bit logic / add / shift work
// (LOP3, IADD3, SHF -- the alu pipe; tensor cores and FP pipes do nothing for it), so the roofline denominator is the
// rate at which the SMs retire such instructions: eight independent LOP3 -> IADD3 chains per thread, every SM full.
// Measurement aid for bench.py; it is not on the data path.
#pragma once
#include "ctx.cuh"

namespace stepth {

constexpr int IP_CHAINS = 8, IP_UNROLL = 16;

// LOGIC_ONLY = false: LOP3 + add (the compiler issues the add as IMAD.IADD on the fma pipe: both integer pipes busy, the
// highest rate any 32-bit integer mix can reach); true: LOP3 + SHF, alu pipe only (the search's own instruction class)
template <bool LOGIC_ONLY>
__global__ void __launch_bounds__(256) int_peak_kernel(uint32_t *sink, uint32_t b, uint32_t c, int iters) {
    uint32_t a[IP_CHAINS];
#pragma unroll
    for (int k = 0; k < IP_CHAINS; ++k) a[k] = threadIdx.x * 2654435761u + k + blockIdx.x;
#pragma unroll 1
    for (int i = 0; i < iters; ++i) {
#pragma unroll
        for (int u = 0; u < IP_UNROLL; ++u) {
#pragma unroll
            for (int k = 0; k < IP_CHAINS; ++k) {
                const uint32_t t = (a[k] ^ b) & (a[k] | c);                       // LOP3
                a[k] = LOGIC_ONLY ? __funnelshift_l(t, t, 5) : t + c;            // SHF | IMAD.IADD
            }
        }
    }
    uint32_t x = 0;
#pragma unroll
    for (int k = 0; k < IP_CHAINS; ++k) x ^= a[k];
    if (x == 0x12345u) sink[0] = x;  // keeps the chains alive; practically never taken
}

// -> thread-level 32-bit integer ops per second (LOP3/IADD3 class), best of `reps`
inline int32_t measure_int_peak(stepth_ctx *ctx, bool logic_only, double *ops_per_s) {
    cudaStream_t st = ctx->stream;
    int32_t rc = ensure(ctx, ctx->worklist, 256);
    if (rc != STEPTH_OK) return rc;
    const int iters = 2048, grid = ctx->sm_count * 16;
    cudaEvent_t e0, e1;
    STEPTH_CUDA_TRY(ctx, cudaEventCreate(&e0));
    STEPTH_CUDA_TRY(ctx, cudaEventCreate(&e1));
    double best = 0.0;
    for (int rep = 0; rep < 6; ++rep) {
        STEPTH_CUDA_TRY(ctx, cudaEventRecord(e0, st));
        if (logic_only) int_peak_kernel<true><<<grid, 256, 0, st>>>((uint32_t *)ctx->worklist.p, 0x9e3779b9u + rep, 0x7f4a7c15u, iters);
        else int_peak_kernel<false><<<grid, 256, 0, st>>>((uint32_t *)ctx->worklist.p, 0x9e3779b9u + rep, 0x7f4a7c15u, iters);
        ctx->launches++;
        STEPTH_CUDA_TRY(ctx, cudaEventRecord(e1, st));
        STEPTH_CUDA_TRY(ctx, cudaEventSynchronize(e1));
        float ms = 0.f;
        STEPTH_CUDA_TRY(ctx, cudaEventElapsedTime(&ms, e0, e1));
        const double ops = 2.0 * IP_CHAINS * IP_UNROLL * (double)iters * 256.0 * grid;
        if (rep > 0 && ms > 0.f) best = std::max(best, ops / (ms * 1e-3));
    }
    cudaEventDestroy(e0);
    cudaEventDestroy(e1);
    *ops_per_s = best;
    return STEPTH_OK;
}

}  // namespace stepth
