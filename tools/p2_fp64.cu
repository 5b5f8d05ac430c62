// Experiment: Poseidon2-BabyBear on the FP64 pipe, alone and next to the integer permutation of the product
// (warps of one block split between the two code paths, work handed out by an atomic counter).
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 --expt-relaxed-constexpr -I openvm_b200/csrc tools/p2_fp64.cu -o tools/bin/p2_fp64
// FP64 form: values are integers held in doubles, canonical domain; a product is the error-free pair (h, l) = a*b
// followed by q = round(h / p) through the 1.5*2^52 constant and r = (h - q p) + l, exact for |a|,|b| < 2^36
// (tools/fp64_modmul.cu checks 2^26 random pairs). Additions are plain DADD/DFMA with no correction; magnitudes
// are bounded per layer (see the comments) and folded with a 3-operation reduction where they would outgrow 2^36.
#include <cstdio>
#include <cstdlib>
#include <cuda_runtime.h>
#include "poseidon2_f64.cuh"
using namespace b2;

__device__ __forceinline__ void dpermute(double (&s)[16]) { poseidon2_permute_f64(s); }
__device__ __forceinline__ double u2d(uint32_t x) { return __hiloint2double(0x43300000, (int)x) - 4503599627370496.0; }
// canonical double in (-p, p) -> canonical u32 in [0, p)
__device__ __forceinline__ uint32_t d2u(double r) {
    r = dred(r);
    const double t = (r < 0.0 ? r + D_P : r) + 4503599627370496.0;
    return (uint32_t)__double2loint(t);
}

__global__ void check_kernel(uint32_t* bad) {
    const uint32_t tid = blockIdx.x * blockDim.x + threadIdx.x;
    Fp s[16];
    double d[16];
#pragma unroll
    for (int k = 0; k < 16; k++) {
        const uint32_t c = (tid * 2654435761u + k * 40503u + 12345u) % P;
        s[k] = Fp::from_canonical(c);
        d[k] = u2d(c);
    }
    for (int it = 0; it < 3; it++) {
        poseidon2_permute(s);
        dpermute(d);
    }
#pragma unroll
    for (int k = 0; k < 16; k++)
        if (s[k].to_canonical() != d2u(d[k])) atomicAdd(bad, 1u);
}

// NI of the 8 warps of a block run the integer permutation, the rest the FP64 one; a chunk = CH chained permutations
constexpr int CH = 16;
template <int NI>
__global__ void __launch_bounds__(256) mixed_kernel(uint32_t* sink, unsigned* ctr, unsigned chunks, unsigned* done_by) {
    const unsigned warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    uint32_t acc = 0;
    unsigned mine = 0;
    if (warp < NI) {
        for (;;) {
            unsigned c = 0;
            if (lane == 0) c = atomicAdd(ctr, 1u);
            c = __shfl_sync(0xffffffffu, c, 0);
            if (c >= chunks) break;
            Fp s[16];
#pragma unroll
            for (int k = 0; k < 16; k++) s[k] = Fp::raw((c * 2654435761u + lane * 97u + k * 40503u) % P);
            for (int i = 0; i < CH; i++) poseidon2_permute(s);
#pragma unroll
            for (int k = 0; k < 16; k++) acc ^= s[k].v;
            mine++;
        }
        if (lane == 0) atomicAdd(&done_by[0], mine);
    } else {
        for (;;) {
            unsigned c = 0;
            if (lane == 0) c = atomicAdd(ctr, 1u);
            c = __shfl_sync(0xffffffffu, c, 0);
            if (c >= chunks) break;
            double s[16];
#pragma unroll
            for (int k = 0; k < 16; k++) s[k] = u2d((c * 2654435761u + lane * 97u + k * 40503u) % P);
            for (int i = 0; i < CH; i++) dpermute(s);
#pragma unroll
            for (int k = 0; k < 16; k++) acc ^= d2u(s[k]);
            mine++;
        }
        if (lane == 0) atomicAdd(&done_by[1], mine);
    }
    sink[blockIdx.x * blockDim.x + threadIdx.x] = acc;
}

template <int NI> void run(uint32_t* sink, unsigned* ctr, int blocks_per_sm) {
    const unsigned blocks = 148 * blocks_per_sm;
    const unsigned chunks = 148 * 8 * 8 * 48;  // warp-chunks of 32 x CH permutations
    cudaFuncAttributes fa;
    cudaFuncGetAttributes(&fa, mixed_kernel<NI>);
    float best = 1e9f;
    unsigned h[3] = {0, 0, 0};
    for (int rep = 0; rep < 3; rep++) {
        cudaMemset(ctr, 0, 12);
        cudaEvent_t a, b;
        cudaEventCreate(&a); cudaEventCreate(&b);
        cudaEventRecord(a);
        mixed_kernel<NI><<<blocks, 256>>>(sink, ctr, rep == 0 ? chunks / 8 : chunks, ctr + 1);
        cudaEventRecord(b);
        cudaEventSynchronize(b);
        float ms; cudaEventElapsedTime(&ms, a, b);
        if (rep && ms < best) best = ms;
        cudaMemcpy(h, ctr, 12, cudaMemcpyDeviceToHost);
    }
    printf("int warps %d / fp64 warps %d, %d blocks/SM: %7.3f G perm/s  regs %3d  chunks int %u fp64 %u\n", NI, 8 - NI, blocks_per_sm,
           (double)chunks * 32 * CH / best / 1e6, fa.numRegs, h[1], h[2]);
}

int main() {
    uint32_t* sink;
    unsigned* ctr;
    cudaMalloc(&sink, 148 * 8 * 256 * 4);
    cudaMalloc(&ctr, 16);
    cudaMemset(ctr, 0, 16);
    check_kernel<<<64, 128>>>(ctr);
    unsigned bad = 0;
    cudaMemcpy(&bad, ctr, 4, cudaMemcpyDeviceToHost);
    printf("fp64 permutation vs integer permutation: %u mismatching words of %u\n", bad, 64u * 128 * 16);
    for (int bps = 3; bps <= 5; bps++) {
        run<8>(sink, ctr, bps);
        run<6>(sink, ctr, bps);
        run<5>(sink, ctr, bps);
        run<4>(sink, ctr, bps);
        run<3>(sink, ctr, bps);
        run<2>(sink, ctr, bps);
        run<0>(sink, ctr, bps);
    }
    cudaError_t e = cudaDeviceSynchronize();
    printf("%s\n", cudaGetErrorString(e));
    return e != cudaSuccess;
}
