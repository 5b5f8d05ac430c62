// Experiment: does the FP64 pipe of a B200 run BabyBear products concurrently with the integer-multiply pipe?
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 tools/fp64_modmul.cu -o tools/bin/fp64_modmul
// Prints products/s for integer-only (signed Montgomery), FP64-only (error-free product + magic rounding) and mixed chains.
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>

constexpr uint32_t P = 2013265921u, P_INV = 0x88000001u;  // p^-1 mod 2^32

__device__ __forceinline__ int32_t smont(int32_t a, int32_t b) {
    const int64_t t = (int64_t)a * b;
    const uint32_t m = (uint32_t)t * P_INV;
    return (int32_t)(t >> 32) - __mulhi((int32_t)m, (int32_t)P);
}
// a * b mod p in (-p, p) for |a|, |b| < 2^36, all operations exact or error-free
__device__ __forceinline__ double dmod(double a, double b) {
    const double MAGIC = 6755399441055744.0;  // 1.5 * 2^52: rounds to nearest integer
    const double h = a * b;
    const double l = fma(a, b, -h);
    const double q = fma(h, 1.0 / (double)P, MAGIC) - MAGIC;
    return fma(-q, (double)P, h) + l;
}

template <int NI, int ND>
__global__ void chain(uint32_t* out, int iters, uint32_t seed) {
    const uint32_t tid = blockIdx.x * blockDim.x + threadIdx.x;
    int32_t xi[NI > 0 ? NI : 1];
    double xd[ND > 0 ? ND : 1];
#pragma unroll
    for (int k = 0; k < NI; k++) xi[k] = (int32_t)((tid * 2654435761u + seed + k) % P);
#pragma unroll
    for (int k = 0; k < ND; k++) xd[k] = (double)((tid * 2246822519u + seed + 77 * k) % P);
    for (int it = 0; it < iters; it++) {
#pragma unroll
        for (int k = 0; k < NI; k++) xi[k] = smont(xi[k], xi[k]);
#pragma unroll
        for (int k = 0; k < ND; k++) xd[k] = dmod(xd[k], xd[k]);
    }
    uint32_t acc = 0;
#pragma unroll
    for (int k = 0; k < NI; k++) acc += (uint32_t)xi[k];
#pragma unroll
    for (int k = 0; k < ND; k++) acc += (uint32_t)(int32_t)xd[k];
    out[tid] = acc;
}

// correctness of dmod against 64-bit integer arithmetic
__global__ void check(uint32_t* bad, uint32_t seed) {
    const uint32_t tid = blockIdx.x * blockDim.x + threadIdx.x;
    uint64_t s = tid * 0x9E3779B97F4A7C15ull + seed;
    for (int it = 0; it < 256; it++) {
        s = s * 6364136223846793005ull + 1442695040888963407ull;
        const int64_t a = (int64_t)(s >> 28) - ((int64_t)1 << 35), b = (int64_t)((s * 0xD6E8FEB86659FD93ull) >> 28) - ((int64_t)1 << 35);
        const double r = dmod((double)a, (double)b);
        const __int128 prod = (__int128)a * b;
        int64_t want = (int64_t)(prod % (int64_t)P);
        int64_t got = (int64_t)r % (int64_t)P;
        if (want < 0) want += P;
        if (got < 0) got += P;
        if (want != got || r >= (double)P || r <= -(double)P) atomicAdd(bad, 1u);
    }
}

template <int NI, int ND> double run(uint32_t* out, int iters) {
    const int blocks = 148 * 16, threads = 256;
    chain<NI, ND><<<blocks, threads>>>(out, 16, 1);
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0); cudaEventCreate(&e1);
    cudaEventRecord(e0);
    chain<NI, ND><<<blocks, threads>>>(out, iters, 2);
    cudaEventRecord(e1);
    cudaEventSynchronize(e1);
    float ms;
    cudaEventElapsedTime(&ms, e0, e1);
    const double prods = (double)blocks * threads * iters * (NI + ND);
    printf("int chains %d, fp64 chains %d: %.3f ms, %.3f T products/s\n", NI, ND, ms, prods / ms / 1e9);
    return prods / ms / 1e9;
}

int main() {
    uint32_t *out, *bad, hbad = 0;
    cudaMalloc(&out, 148 * 16 * 256 * 4);
    cudaMalloc(&bad, 4);
    cudaMemset(bad, 0, 4);
    check<<<1024, 256>>>(bad, 12345);
    cudaMemcpy(&hbad, bad, 4, cudaMemcpyDeviceToHost);
    printf("dmod mismatches: %u of %u\n", hbad, 1024u * 256 * 256);
    const int iters = 4096;
    run<4, 0>(out, iters);
    run<8, 0>(out, iters);
    run<0, 4>(out, iters);
    run<0, 8>(out, iters);
    run<4, 4>(out, iters);
    run<4, 2>(out, iters);
    run<6, 2>(out, iters);
    run<8, 4>(out, iters);
    run<4, 1>(out, iters);
    printf("%s\n", cudaGetErrorString(cudaDeviceSynchronize()));
    return 0;
}
