// b2s_merkle_root_static: the Poseidon2 Merkle root of ONE already-extended device matrix, computed without any
// allocation. No counterpart in the reference; added for callers that cannot call
// cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid it).
//
// The digest layers of the tree -- the scratch that mmcs_build (merkle.cu) takes from the engine's stream-ordered pool --
// live here in one file-scope __device__ array of B2S_STATIC_SCRATCH_BYTES (4 MiB) that every GPU loading the library
// holds until the process exits. The kernels are a one-matrix variant of tree_level_kernel (merkle.cu): one node per
// thread, 8 word-planes per layer, same sponge (rate 8, overwrite mode, no padding) and same 2-to-1 compression, so the
// root equals b2s_mmcs_commit's on the same matrix bit for bit.
//
// The array's NAME is a host-side stand-in in host code, not a device address: its device address is taken once per
// device with cudaGetSymbolAddress (device current), and that pointer is what the kernels receive.
//
// Every call on a device shares the array; it keeps the last call's contents. Nothing in it has to start at zero: each
// level reads only digests the previous launch of the same call wrote (layers are laid out for the call's own height),
// so stale words are never read. Two calls that overlap on the GPU (different streams, or different host threads on
// different streams) race on the array: the caller must order them.
#include "common.cuh"
#include "poseidon2.cuh"

#include <mutex>

namespace b2 {
void set_global_error(const std::string& msg);   // api.cu

constexpr uint64_t STATIC_SCRATCH_WORDS = B2S_STATIC_SCRATCH_BYTES / 4;
static_assert(STATIC_SCRATCH_WORDS >= 2 * (uint64_t)B2S_STATIC_MAX_ROWS * 8, "layers of the tallest tree must fit");

static __device__ __align__(256) uint32_t g_static_scratch[STATIC_SCRATCH_WORDS];

constexpr int NA_THREADS = 128;

// mode 0: out[i] = sponge(row i of the matrix)      mode 2: out[i] = compress(prev[2i], prev[2i+1])
// root8 != nullptr (n == 1): the single node is also written there as 8 canonical words
__global__ void __launch_bounds__(NA_THREADS, 8)
    static_tree_level_kernel(const uint32_t* __restrict__ base, uint64_t stride, uint32_t ncols, uint64_t n,
                             const uint32_t* prev, uint32_t* out, int mode, uint32_t* root8) {
    const uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    Fp s[16];
#pragma unroll
    for (int k = 0; k < 16; k++) s[k] = Fp::zero();
    if (mode == 0) {
        const uint32_t chunks = (ncols + 7) >> 3;
#pragma unroll 1
        for (uint32_t it = 0; it < chunks; it++) {
            const uint32_t c = it << 3;
#pragma unroll
            for (int k = 0; k < 8; k++)
                if (c + k < ncols) s[k] = Fp::raw(__ldg(base + (uint64_t)(c + k) * stride + i));
            poseidon2_permute(s);
        }
    } else {
        const uint64_t np = 2 * n;
#pragma unroll
        for (int w = 0; w < 8; w++) {
            const uint2 lr = __ldcg(reinterpret_cast<const uint2*>(prev + (uint64_t)w * np + 2 * i));
            s[w] = Fp::raw(lr.x);
            s[8 + w] = Fp::raw(lr.y);
        }
        poseidon2_permute(s);
    }
#pragma unroll
    for (int w = 0; w < 8; w++) out[(uint64_t)w * n + i] = s[w].v;
    if (root8) {
#pragma unroll
        for (int w = 0; w < 8; w++) root8[w] = s[w].to_canonical();
    }
}

// device address of the scratch array on the current device, looked up once per device
static uint32_t* static_scratch_ptr(int device) {
    constexpr int MAX_DEVICES = 64;
    static std::mutex mu;
    static uint32_t* ptrs[MAX_DEVICES] = {nullptr};
    B2_REQUIRE(device >= 0 && device < MAX_DEVICES, B2S_ERR_BAD_ARG, "merkle_root_static: device ordinal out of range");
    std::lock_guard<std::mutex> lk(mu);
    if (!ptrs[device]) {
        uint32_t* p = nullptr;
        B2_CUDA(cudaGetSymbolAddress((void**)&p, g_static_scratch));
        B2_REQUIRE(p != nullptr, B2S_ERR_INTERNAL, "merkle_root_static: null scratch address");
        ptrs[device] = p;
    }
    return ptrs[device];
}

static void merkle_root_static(const uint32_t* d_matrix, uint64_t height, uint64_t width, uint32_t* d_root8,
                               cudaStream_t stream) {
    B2_REQUIRE(d_matrix && d_root8, B2S_ERR_BAD_ARG, "merkle_root_static: null pointer");
    B2_REQUIRE(height >= 1 && is_pow2(height), B2S_ERR_SHAPE, "merkle_root_static: height must be a power of two");
    B2_REQUIRE(height <= B2S_STATIC_MAX_ROWS, B2S_ERR_SHAPE,
               "merkle_root_static: height " + std::to_string(height) + " exceeds B2S_STATIC_MAX_ROWS (" +
                   std::to_string(B2S_STATIC_MAX_ROWS) + " rows, the size of the static scratch array)");
    B2_REQUIRE(width >= 1 && width <= B2S_STATIC_MAX_WIDTH, B2S_ERR_SHAPE,
               "merkle_root_static: width " + std::to_string(width) + " outside [1, B2S_STATIC_MAX_WIDTH = " +
                   std::to_string(B2S_STATIC_MAX_WIDTH) + "]");
    int device = 0;
    B2_CUDA(cudaGetDevice(&device));
    uint32_t* scratch = static_scratch_ptr(device);
    // layers back to back: leaves (height nodes) at word 0, then height/2 nodes, ... each 8 planes
    uint32_t* layer = scratch;
    uint64_t n = height;
    static_tree_level_kernel<<<(unsigned)((n + NA_THREADS - 1) / NA_THREADS), NA_THREADS, 0, stream>>>(
        d_matrix, height, (uint32_t)width, n, nullptr, layer, 0, n == 1 ? d_root8 : nullptr);
    B2_CHECK_LAUNCH();
    while (n > 1) {
        uint32_t* next = layer + n * 8;
        n >>= 1;
        static_tree_level_kernel<<<(unsigned)((n + NA_THREADS - 1) / NA_THREADS), NA_THREADS, 0, stream>>>(
            nullptr, 0, 0, n, layer, next, 2, n == 1 ? d_root8 : nullptr);
        B2_CHECK_LAUNCH();
        layer = next;
    }
}

}  // namespace b2

extern "C" int32_t b2s_merkle_root_static(const uint32_t* d_matrix, uint64_t height, uint64_t width, uint32_t* d_root8,
                                          void* cuda_stream) {
    try {
        b2::merkle_root_static(d_matrix, height, width, d_root8, (cudaStream_t)cuda_stream);
        return B2S_OK;
    } catch (const b2::StatusError& se) {
        b2::set_global_error(se.msg);
        return se.code;
    } catch (...) {
        b2::set_global_error("merkle_root_static: unknown failure");
        return B2S_ERR_INTERNAL;
    }
}
