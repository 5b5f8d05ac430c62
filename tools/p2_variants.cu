// Micro-benchmark of Poseidon2 code-generation variants (experiments only; the product kernel is csrc/poseidon2.cuh).
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 --expt-relaxed-constexpr -I openvm_b200/csrc tools/p2_variants.cu -o tools/bin/p2_variants
#include <cstdio>
#include <cstdlib>
#include <cuda_runtime.h>
#include "field.cuh"
#include "poseidon2_constants.cuh"
using namespace b2;

static __constant__ uint32_t C_EXT[8][16] = B2_P2_EXT_ALL;
static __constant__ uint32_t C_PARTIAL[13] = B2_P2_PARTIAL;
static __constant__ uint32_t C_ZERO = 0u;
static __constant__ uint32_t C_UMAX = 0xFFFFFFFFu;

// ADD mode: 0 plain (ptxas chooses), 1 three-input add with an opaque zero, 2 add fused with min against opaque UINT_MAX
template <int MODE> __device__ __forceinline__ uint32_t madd(uint32_t a, uint32_t b) {
    uint32_t s;
    if (MODE == 1) s = a + b + C_ZERO;
    else if (MODE == 2) s = min(a + b, C_UMAX);
    else s = a + b;
    return min(s - P, s);
}
template <int MODE> __device__ __forceinline__ uint32_t msub(uint32_t a, uint32_t b) {
    uint32_t d;
    if (MODE == 1) d = a - b + C_ZERO;
    else if (MODE == 2) d = min(a - b, C_UMAX);
    else d = a - b;
    return min(d + P, d);
}
static __constant__ uint32_t C_27 = 27u, C_31 = 31u;
// ALUM: m = lo * p^-1 = lo + (lo << 27) + (lo << 31) with shift amounts the compiler cannot see (no IMAD.SHL)
template <bool ALUM = false>
__device__ __forceinline__ int32_t smont(int32_t a, int32_t b) {
    const int64_t t = (int64_t)a * b;
    const uint32_t lo = (uint32_t)t;
    const uint32_t m = ALUM ? lo + (lo << C_27) + (lo << C_31) : lo * P_INV;
    const int32_t u = __mulhi((int32_t)m, (int32_t)P);
    return (int32_t)(t >> 32) - u;
}
template <int SM = 0>
__device__ __forceinline__ uint32_t sbox(uint32_t x, uint32_t rc) {
    const int32_t y = (int32_t)(x + rc - P);
    const int32_t y2 = smont<(SM & 1) != 0>(y, y), y3 = smont<(SM & 2) != 0>(y2, y), y4 = smont<(SM & 4) != 0>(y2, y2);
    const uint32_t y7 = (uint32_t)smont<(SM & 8) != 0>(y3, y4);
    return min(y7 + P, y7);
}
template <int K, bool NEG> __device__ __forceinline__ uint32_t mul2n(uint32_t x) {
    const uint32_t lo = x & ((1u << K) - 1), hi = x >> K;
    const uint32_t t = lo * (15u << (27 - K));
    const uint32_t r = NEG ? t - hi : hi - t;
    return min(r + P, r);
}
__device__ __forceinline__ uint32_t halve(uint32_t a) { return (a >> 1) + (a & 1u) * ((P + 1) >> 1); }

template <int M> __device__ __forceinline__ void external(uint32_t (&s)[16]) {
#pragma unroll
    for (int b = 0; b < 16; b += 4) {
        uint32_t x0 = s[b], x1 = s[b + 1], x2 = s[b + 2], x3 = s[b + 3];
        uint32_t t01 = madd<M>(x0, x1), t23 = madd<M>(x2, x3), t0123 = madd<M>(t01, t23);
        uint32_t t01123 = madd<M>(t0123, x1), t01233 = madd<M>(t0123, x3);
        s[b + 3] = madd<M>(t01233, madd<M>(x0, x0));
        s[b + 1] = madd<M>(t01123, madd<M>(x2, x2));
        s[b] = madd<M>(t01123, t01);
        s[b + 2] = madd<M>(t01233, t23);
    }
    uint32_t sums[4];
#pragma unroll
    for (int k = 0; k < 4; k++) sums[k] = madd<M>(madd<M>(s[k], s[4 + k]), madd<M>(s[8 + k], s[12 + k]));
#pragma unroll
    for (int i = 0; i < 16; i++) s[i] = madd<M>(s[i], sums[i & 3]);
}
template <int M> __device__ __forceinline__ void internal(uint32_t (&s)[16]) {
    uint32_t part = madd<M>(madd<M>(madd<M>(madd<M>(s[1], s[2]), madd<M>(s[3], s[4])), madd<M>(madd<M>(s[5], s[6]), madd<M>(s[7], s[8]))),
                            madd<M>(madd<M>(madd<M>(s[9], s[10]), madd<M>(s[11], s[12])), madd<M>(madd<M>(s[13], s[14]), s[15])));
    uint32_t full = madd<M>(part, s[0]);
    uint32_t h3 = halve(s[3]), h6 = halve(s[6]);
    uint32_t d4 = madd<M>(s[4], s[4]), d5 = madd<M>(s[5], s[5]), d7 = madd<M>(s[7], s[7]), d8 = madd<M>(s[8], s[8]);
    uint32_t y[16];
    y[0] = msub<M>(part, s[0]);
    y[1] = madd<M>(full, s[1]);
    y[2] = madd<M>(full, madd<M>(s[2], s[2]));
    y[3] = madd<M>(full, h3);
    y[4] = madd<M>(full, madd<M>(d4, s[4]));
    y[5] = madd<M>(full, madd<M>(d5, d5));
    y[6] = msub<M>(full, h6);
    y[7] = msub<M>(full, madd<M>(d7, s[7]));
    y[8] = msub<M>(full, madd<M>(d8, d8));
    y[9] = madd<M>(full, mul2n<8, false>(s[9]));
    y[10] = madd<M>(full, mul2n<2, false>(s[10]));
    y[11] = madd<M>(full, mul2n<3, false>(s[11]));
    y[12] = madd<M>(full, mul2n<27, false>(s[12]));
    y[13] = madd<M>(full, mul2n<8, true>(s[13]));
    y[14] = madd<M>(full, mul2n<4, true>(s[14]));
    y[15] = madd<M>(full, mul2n<27, true>(s[15]));
#pragma unroll
    for (int i = 0; i < 16; i++) s[i] = y[i];
}
template <int ME, int MI, int SM = 0> __device__ __forceinline__ void permute(uint32_t (&s)[16]) {
#pragma unroll 1
    for (int r = 0; r < 9; r++) {
        external<ME>(s);
        if (r == 4) {
#pragma unroll 1
            for (int q = 0; q < 13; q++) { s[0] = sbox<SM>(s[0], C_PARTIAL[q]); internal<MI>(s); }
        }
        if (r < 8) {
#pragma unroll
            for (int i = 0; i < 16; i++) s[i] = sbox<SM>(s[i], C_EXT[r][i]);
        }
    }
}
template <int ME, int MI, int MINB, int SM = 0>
__global__ void __launch_bounds__(128, MINB) kern(uint32_t* sink, uint32_t iters) {
    uint32_t s[16];
#pragma unroll
    for (int k = 0; k < 16; k++) s[k] = (threadIdx.x * 2654435761u + k * 40503u + blockIdx.x) % P;
    for (uint32_t i = 0; i < iters; i++) permute<ME, MI, SM>(s);
    uint32_t acc = 0;
#pragma unroll
    for (int k = 0; k < 16; k++) acc ^= s[k];
    sink[blockIdx.x * blockDim.x + threadIdx.x] = acc;
}
template <int ME, int MI, int MINB, int SM = 0> void run(const char* name, uint32_t* sink, uint32_t* ref, size_t n) {
    const uint32_t iters = 64;
    const unsigned blocks = (unsigned)(n / 128);
    kern<ME, MI, MINB, SM><<<blocks, 128>>>(sink, 4);
    cudaEvent_t a, b;
    cudaEventCreate(&a); cudaEventCreate(&b);
    float best = 1e9f;
    for (int rep = 0; rep < 3; rep++) {
        cudaEventRecord(a);
        kern<ME, MI, MINB, SM><<<blocks, 128>>>(sink, iters);
        cudaEventRecord(b);
        cudaEventSynchronize(b);
        float ms; cudaEventElapsedTime(&ms, a, b);
        if (ms < best) best = ms;
    }
    uint32_t h[4];
    cudaMemcpy(h, sink, 16, cudaMemcpyDeviceToHost);
    bool ok = true;
    if (ref[0] == 0 && ref[1] == 0) for (int i = 0; i < 4; i++) ref[i] = h[i];
    else for (int i = 0; i < 4; i++) ok &= ref[i] == h[i];
    cudaFuncAttributes fa;
    cudaFuncGetAttributes(&fa, kern<ME, MI, MINB, SM>);
    printf("%-34s %7.3f G perm/s  regs %3d  %s\n", name, n * (double)iters / best / 1e6, fa.numRegs, ok ? "same result" : "RESULT DIFFERS");
}
int main() {
    const size_t n = 148 * 2048;
    uint32_t* sink;
    cudaMalloc(&sink, n * 4);
    uint32_t ref[4] = {0, 0, 0, 0};
    run<0, 0, 8>("plain adds (product kernel)", sink, ref, n);
    run<1, 0, 8>("E: IADD3+zero", sink, ref, n);
    run<0, 1, 8>("I: IADD3+zero", sink, ref, n);
    run<1, 1, 8>("E+I: IADD3+zero", sink, ref, n);
    run<2, 0, 8>("E: VIADDMNMX+umax", sink, ref, n);
    run<0, 2, 8>("I: VIADDMNMX+umax", sink, ref, n);
    run<0, 0, 6>("plain, 6 blocks/SM (80 regs)", sink, ref, n);
    run<0, 0, 4>("plain, 4 blocks/SM (128 regs)", sink, ref, n);
    run<0, 0, 8, 1>("m on ALU: product 1 of 4", sink, ref, n);
    run<0, 0, 8, 9>("m on ALU: products 1,4", sink, ref, n);
    run<0, 0, 8, 6>("m on ALU: products 2,3", sink, ref, n);
    run<0, 0, 8, 15>("m on ALU: all 4", sink, ref, n);
    cudaError_t e = cudaDeviceSynchronize();
    if (e != cudaSuccess) { printf("cuda error %s\n", cudaGetErrorString(e)); return 1; }
    return 0;
}
