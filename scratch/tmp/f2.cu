#include <cuda_runtime.h>
#include <cstdint>
__device__ __forceinline__ uint64_t mul2(uint64_t a, uint64_t b){ uint64_t d; asm("mul.rn.f32x2 %0, %1, %2;" : "=l"(d) : "l"(a), "l"(b)); return d; }
__device__ __forceinline__ uint64_t add2(uint64_t a, uint64_t b){ uint64_t d; asm("add.rn.f32x2 %0, %1, %2;" : "=l"(d) : "l"(a), "l"(b)); return d; }
__global__ void k(const uint64_t* a, const uint64_t* b, uint64_t* c) {
    int i = threadIdx.x;
    uint64_t x = a[i], y = b[i];
    c[i] = add2(mul2(x, y), x);
}
