// Probe: how do the integer-multiply pipe, the ALU pipe and the FP64 pipe of a B200 share the issue/dispatch port?
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 tools/dispatch_probe.cu -o tools/bin/dispatch_probe
// Each kernel runs independent dependency chains of one or two instruction kinds (8 chains per kind per thread, 16 warps
// per SM sub-partition) and reports warp-instructions issued per cycle and sub-partition. A kind that alone reaches 0.5
// (one per 2 cycles) and loses nothing next to another kind has its own pipe and a free dispatch slot; a pair whose
// total stays at the slower kind's rate shares something.
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>

enum Kind { K_NONE = 0, K_IMAD, K_WIDE, K_HI, K_IADD3 /* LOP3: a plain ALU-pipe instruction ptxas cannot turn into IMAD.IADD */, K_VMNMX, K_DFMA, K_FFMA };

template <int K> struct Chain {
    uint32_t a[8]; uint64_t w[8]; double d[8]; float f[8];
    __device__ void init(uint32_t s) {
#pragma unroll
        for (int i = 0; i < 8; i++) { a[i] = s * 2654435761u + i * 40503u + 1u; w[i] = a[i]; d[i] = (double)(a[i] & 1023u) + 1.0; f[i] = (float)(a[i] & 255u) + 1.0f; }
    }
    // one PTX instruction per chain element, volatile so that nothing is merged or dropped (SASS checked with cuobjdump)
    __device__ void step(uint32_t c, uint32_t c2) {
#pragma unroll
        for (int i = 0; i < 8; i++) {
            if (K == K_IMAD) asm volatile("mad.lo.u32 %0, %0, %1, %2;" : "+r"(a[i]) : "r"(c), "r"(c2));
            // both halves of the product feed the chain (one ALU add joins them: the ALU pipe runs next to the multiplier,
            // see IMAD + LOP3), otherwise ptxas narrows the instruction to a 32-bit IMAD
            if (K == K_WIDE) asm volatile("{ .reg .u64 t; .reg .u32 l, h; mul.wide.u32 t, %0, %1; mov.b64 {l, h}, t; add.u32 %0, l, h; }" : "+r"(a[i]) : "r"(c));
            if (K == K_HI) asm volatile("mul.hi.u32 %0, %0, %1;" : "+r"(a[i]) : "r"(c));
            if (K == K_IADD3) asm volatile("lop3.b32 %0, %0, %1, %2, 0x96;" : "+r"(a[i]) : "r"(c), "r"(c2));
            if (K == K_VMNMX) asm volatile("{ .reg .u32 t; add.u32 t, %0, %1; min.u32 %0, t, %2; }" : "+r"(a[i]) : "r"(c), "r"(c2));
            if (K == K_DFMA) asm volatile("fma.rn.f64 %0, %0, %1, %2;" : "+d"(d[i]) : "d"(1.0000001), "d"(1e-9));
            if (K == K_FFMA) asm volatile("fma.rn.f32 %0, %0, %1, %2;" : "+f"(f[i]) : "f"(1.0000001f), "f"(1e-9f));
        }
    }
    __device__ uint32_t fold() {
        uint32_t r = 0;
#pragma unroll
        for (int i = 0; i < 8; i++) r ^= a[i] ^ (uint32_t)w[i] ^ (uint32_t)(w[i] >> 32) ^ (uint32_t)__double2loint(d[i]) ^ __float_as_uint(f[i]);
        return r;
    }
};

template <int KA, int KB>
__global__ void __launch_bounds__(512) probe(uint32_t* sink, int iters, uint32_t c, uint32_t c2) {
    Chain<KA> x; Chain<KB> y;
    x.init(threadIdx.x + blockIdx.x * blockDim.x);
    y.init(threadIdx.x * 7 + 3);
    for (int it = 0; it < iters; it++) {
        x.step(c, c2);
        if (KB != K_NONE) y.step(c2, c);
    }
    sink[blockIdx.x * blockDim.x + threadIdx.x] = x.fold() ^ y.fold();
}

template <int KA, int KB> void run(const char* name, uint32_t* sink, int sms, double mhz) {
    const int iters = 4096, blocks = sms * 4, threads = 512;   // 64 warps per SM
    probe<KA, KB><<<blocks, threads>>>(sink, 16, 3u, 5u);
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0); cudaEventCreate(&e1);
    cudaEventRecord(e0);
    probe<KA, KB><<<blocks, threads>>>(sink, iters, 3u, 5u);
    cudaEventRecord(e1);
    cudaEventSynchronize(e1);
    float ms;
    cudaEventElapsedTime(&ms, e0, e1);
    const double per_kind = (double)iters * 8 * (threads / 32) * 4 / 4;   // warp instructions per kind per sub-partition (4 blocks / 4 SMSPs)
    const double cycles = ms * 1e-3 * mhz * 1e6;
    const int kinds = KB == K_NONE ? 1 : 2;
    printf("%-28s %8.3f ms   %.3f warp-instr / cycle / SMSP in total (%.3f per kind)\n", name, ms, kinds * per_kind / cycles, per_kind / cycles);
}

int main() {
    cudaDeviceProp p;
    cudaGetDeviceProperties(&p, 0);
    int khz = 0;
    cudaDeviceGetAttribute(&khz, cudaDevAttrClockRate, 0);
    const double mhz = khz / 1000.0;
    printf("%s, %d SMs, %0.f MHz (rates assume this clock)\n", p.name, p.multiProcessorCount, mhz);
    uint32_t* sink;
    cudaMalloc(&sink, (size_t)p.multiProcessorCount * 4 * 512 * 4);
    const int sms = p.multiProcessorCount;
    run<K_IMAD, K_NONE>("IMAD (32-bit)", sink, sms, mhz);
    run<K_WIDE, K_NONE>("IMAD.WIDE", sink, sms, mhz);
    run<K_HI, K_NONE>("IMAD.HI", sink, sms, mhz);
    run<K_IADD3, K_NONE>("LOP3 (ALU)", sink, sms, mhz);
    run<K_VMNMX, K_NONE>("VIADDMNMX", sink, sms, mhz);
    run<K_DFMA, K_NONE>("DFMA", sink, sms, mhz);
    run<K_FFMA, K_NONE>("FFMA", sink, sms, mhz);
    run<K_IMAD, K_IADD3>("IMAD + LOP3", sink, sms, mhz);
    run<K_WIDE, K_IADD3>("IMAD.WIDE + LOP3", sink, sms, mhz);
    run<K_WIDE, K_VMNMX>("IMAD.WIDE + VIADDMNMX", sink, sms, mhz);
    run<K_DFMA, K_IADD3>("DFMA + LOP3", sink, sms, mhz);
    run<K_DFMA, K_IMAD>("DFMA + IMAD", sink, sms, mhz);
    run<K_DFMA, K_WIDE>("DFMA + IMAD.WIDE", sink, sms, mhz);
    run<K_DFMA, K_FFMA>("DFMA + FFMA", sink, sms, mhz);
    run<K_IMAD, K_FFMA>("IMAD + FFMA", sink, sms, mhz);
    run<K_WIDE, K_FFMA>("IMAD.WIDE + FFMA", sink, sms, mhz);
    printf("%s\n", cudaGetErrorString(cudaDeviceSynchronize()));
    return 0;
}
