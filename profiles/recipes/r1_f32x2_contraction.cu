// ptxas 12.9, sm_100a: mul.rn.f32x2 followed by add.rn.f32x2 comes out as ONE FFMA2 (fused), with or without -fmad=false:
//   nvcc -gencode arch=compute_100a,code=sm_100a -fmad=false -c f32x2_contraction.cu && cuobjdump -sass f32x2_contraction.o | grep FFMA2
// which is why K2 (finish.cuh) keeps scalar __fmul_rn / __fadd_rn: the reference rounds the product and the sum separately.
#include <cuda_runtime.h>
#include <cstdint>
__device__ __forceinline__ uint64_t mul2(uint64_t a, uint64_t b){ uint64_t d; asm("mul.rn.f32x2 %0, %1, %2;" : "=l"(d) : "l"(a), "l"(b)); return d; }
__device__ __forceinline__ uint64_t add2(uint64_t a, uint64_t b){ uint64_t d; asm("add.rn.f32x2 %0, %1, %2;" : "=l"(d) : "l"(a), "l"(b)); return d; }
__global__ void k(const uint64_t* a, const uint64_t* b, uint64_t* c) {
    int i = threadIdx.x;
    uint64_t x = a[i], y = b[i];
    c[i] = add2(mul2(x, y), x);
}
