// C ABI of libb2stark.so (include/b2stark.h). Every entry point converts internal failures into status codes;
// there is no CPU fallback behind any of them.
#include "common.cuh"
#include "bn254.cuh"

#include <mutex>

using namespace b2;

namespace b2 {
void free_scales(b2s_engine* e);
// Message of the last failure on THIS host thread (b2s_last_error(NULL)): thread-local, so K proving threads cannot read
// each other's messages or a string another thread is rewriting.
static thread_local std::string g_last_error;
void set_global_error(const std::string& msg) { g_last_error = msg; }
}  // namespace b2

// ---- engine internals ------------------------------------------------------------------------------------
void* b2s_engine::dmalloc(size_t bytes) {
    if (bytes == 0) bytes = 4;
    void* p = nullptr;
    auto it = block_cache.find({stream, bytes});
    if (it != block_cache.end() && !it->second.empty()) {
        p = it->second.back();
        it->second.pop_back();
        cache_bytes -= bytes;
    } else {
        cudaError_t err = cudaMallocAsync(&p, bytes, stream);
        if (err == cudaErrorMemoryAllocation) {   // give the cached blocks back and try once more
            cudaGetLastError();
            flush_cache();
            err = cudaMallocAsync(&p, bytes, stream);
        }
        if (err == cudaErrorMemoryAllocation) {
            cudaGetLastError();
            fail(B2S_ERR_OOM, "device allocation of " + std::to_string(bytes) + " bytes failed");
        }
        B2_CUDA(err);
    }
    mem_in_use += bytes;
    if (mem_in_use > mem_peak) mem_peak = mem_in_use;
    return p;
}
void b2s_engine::dfree(void* p, size_t bytes) {
    if (!p) return;
    if (bytes == 0) bytes = 4;
    mem_in_use -= bytes;
    block_cache[{stream, bytes}].push_back(p);
    cache_bytes += bytes;
    if (cache_bytes > cache_cap) flush_cache();
}
void b2s_engine::flush_cache() {
    for (auto& kv : block_cache)
        for (void* q : kv.second) cudaFreeAsync(q, kv.first.first);
    block_cache.clear();
    cache_bytes = 0;
}
void* b2s_engine::host_pinned(size_t bytes) {
    if (bytes > pinned_bytes) {
        B2_CUDA(cudaStreamSynchronize(stream));   // nothing may still be landing in the old buffer
        if (pinned) B2_CUDA(cudaFreeHost(pinned));
        pinned = nullptr;
        pinned_bytes = 0;
        const size_t want = std::max<size_t>(bytes + bytes / 2, (size_t)1 << 20);
        B2_CUDA(cudaMallocHost(&pinned, want));
        pinned_bytes = want;
    }
    return pinned;
}
cudaEvent_t b2s_engine::get_event() {
    if (!event_pool.empty()) {
        cudaEvent_t ev = event_pool.back();
        event_pool.pop_back();
        return ev;
    }
    cudaEvent_t ev;
    B2_CUDA(cudaEventCreate(&ev));
    return ev;
}
void b2s_engine::stage_begin(Stage s) {
    if (!metrics_enabled) return;
    cudaEvent_t a = get_event();
    B2_CUDA(cudaEventRecord(a, stream));
    timers[s].pending.push_back({a, nullptr});
}
void b2s_engine::stage_end(Stage s) {
    if (!metrics_enabled) return;
    // close the innermost open interval of this stage
    for (auto it = timers[s].pending.rbegin(); it != timers[s].pending.rend(); ++it) {
        if (it->second == nullptr) {
            cudaEvent_t b = get_event();
            if (cudaEventRecord(b, stream) == cudaSuccess) it->second = b;
            return;
        }
    }
}
void b2s_engine::collect_timers() {
    B2_CUDA(cudaStreamSynchronize(stream));
    for (int s = 0; s < ST_COUNT; s++) {
        for (auto& pr : timers[s].pending) {
            if (pr.first && pr.second) {
                float ms = 0;
                if (cudaEventElapsedTime(&ms, pr.first, pr.second) == cudaSuccess) timers[s].ms += ms;
            }
            if (pr.first) event_pool.push_back(pr.first);
            if (pr.second) event_pool.push_back(pr.second);
        }
        timers[s].pending.clear();
    }
}

namespace {

template <typename F>
int32_t guarded(const b2s_engine* e_const, F&& fn) {
    b2s_engine* e = const_cast<b2s_engine*>(e_const);
    try {
        if (e) B2_CUDA(cudaSetDevice(e->device));
        fn();
        return B2S_OK;
    } catch (const StatusError& se) {
        if (e) e->last_error = se.msg;
        set_global_error(se.msg);
        return se.code;
    } catch (const std::bad_alloc&) {
        if (e) e->last_error = "host allocation failed";
        set_global_error("host allocation failed");
        return B2S_ERR_OOM;
    } catch (const std::exception& ex) {
        if (e) e->last_error = ex.what();
        set_global_error(ex.what());
        return B2S_ERR_INTERNAL;
    } catch (...) {
        if (e) e->last_error = "unknown failure";
        set_global_error("unknown failure");
        return B2S_ERR_INTERNAL;
    }
}

}  // namespace
namespace b2 {
b2s_matrix* new_matrix(b2s_engine* e, uint64_t height, uint64_t width) {
    b2s_matrix* m = new b2s_matrix();
    m->eng = e;
    m->height = height;
    m->width = width;
    m->owned = true;
    m->d = (uint32_t*)e->dmalloc(m->bytes());
    return m;
}

void free_pcs(b2s_pcs_data* d) {
    if (!d) return;
    b2s_engine* e = d->eng;
    for (size_t i = 0; i < d->layers.size(); i++) e->dfree(d->layers[i], d->layer_nodes[i] * 32);
    for (size_t i = 0; i < d->mats.size(); i++)
        if (d->owns[i]) b2s_matrix_free(d->mats[i]);
    delete d;
}

// One tall host-resident trace -> (LDE, Merkle tree), with the upload hidden behind the arithmetic: column groups are
// copied on a second stream into a double buffer while the previous group is extended and absorbed into the per-row
// sponge states. LDE and hashing are multiply-bound, so PCIe (~55 GB/s) is the faster side: for
// 2^24 x 256 a group is 39 ms of copy against 56 ms of arithmetic. The trace itself is not kept on the device.
b2s_pcs_data* commit_host_streamed(b2s_engine* e, const uint32_t* host, uint64_t N, uint64_t W, int layout, int repr) {
    StageScope total(e, ST_COMMIT);
    B2_REQUIRE(is_pow2(N) && W > 0, B2S_ERR_SHAPE, "commit: height must be a power of two and width > 0");
    const uint32_t lb = e->params.log_blowup;
    const unsigned logn = log2_exact(N);
    B2_REQUIRE(logn + lb <= TWO_ADICITY, B2S_ERR_SHAPE, "commit: trace too tall for this blowup");
    const uint64_t M = N << lb;
    const uint64_t gq = e->hash_kind ? 16 : 8;          // columns per sponge block: a group ends on a block boundary
    const size_t state_bytes = e->hash_kind ? 96 : 64;  // sponge state of one row between groups
    uint64_t G = (((uint64_t)1 << 27) / N) & ~(gq - 1);
    // B2S_COMMIT_HOST_GROUP_COLS (tuning / test aid): columns per upload group, rounded down to whole sponge blocks
    if (const char* env = getenv("B2S_COMMIT_HOST_GROUP_COLS")) G = strtoull(env, nullptr, 10) & ~(gq - 1);
    if (G < gq) G = gq;
    if (G > W) G = W;
    const uint64_t ngroups = (W + G - 1) / G;
    const bool row_major = layout == B2S_LAYOUT_ROW_MAJOR;
    if (!e->copy_stream) B2_CUDA(cudaStreamCreateWithFlags(&e->copy_stream, cudaStreamNonBlocking));
    b2s_pcs_data* d = new b2s_pcs_data();
    d->eng = e;
    uint32_t *stage[2] = {nullptr, nullptr}, *raw[2] = {nullptr, nullptr}, *state = nullptr, *leaf = nullptr;
    cudaEvent_t copied[2] = {e->get_event(), e->get_event()}, consumed[2] = {e->get_event(), e->get_event()};
    auto cleanup = [&]() {
        for (int k = 0; k < 2; k++) {
            e->dfree(stage[k], G * N * 4);
            if (raw[k]) e->dfree(raw[k], G * N * 4);
            e->event_pool.push_back(copied[k]);
            e->event_pool.push_back(consumed[k]);
        }
        if (state) e->dfree(state, M * state_bytes);
    };
    try {
        b2s_matrix* lde = new_matrix(e, M, W);
        d->mats.push_back(lde);
        d->owns.push_back(true);
        for (int k = 0; k < 2; k++) {
            stage[k] = (uint32_t*)e->dmalloc(G * N * 4);
            if (row_major) raw[k] = (uint32_t*)e->dmalloc(G * N * 4);
        }
        if (ngroups > 1) state = (uint32_t*)e->dmalloc(M * state_bytes);
        leaf = (uint32_t*)e->dmalloc(M * 32);
        // the buffers come from the stream-ordered allocator of e->stream: order the copy stream behind them
        B2_CUDA(cudaEventRecord(consumed[0], e->stream));
        B2_CUDA(cudaStreamWaitEvent(e->copy_stream, consumed[0], 0));
        for (uint64_t g = 0; g < ngroups; g++) {
            const uint64_t c0 = g * G, gw = std::min(G, W - c0);
            const int k = (int)(g & 1);
            if (g >= 2) B2_CUDA(cudaStreamWaitEvent(e->copy_stream, consumed[k], 0));
            if (row_major)
                B2_CUDA(cudaMemcpy2DAsync(raw[k], gw * 4, host + c0, W * 4, gw * 4, N, cudaMemcpyHostToDevice, e->copy_stream));
            else
                B2_CUDA(cudaMemcpyAsync(stage[k], host + c0 * N, gw * N * 4, cudaMemcpyHostToDevice, e->copy_stream));
            B2_CUDA(cudaEventRecord(copied[k], e->copy_stream));
            B2_CUDA(cudaStreamWaitEvent(e->stream, copied[k], 0));
            if (row_major) transpose_to_colmajor(e, raw[k], stage[k], N, gw, repr == B2S_REPR_CANONICAL);
            else if (repr == B2S_REPR_CANONICAL) convert_repr(e, stage[k], gw * N, true);
            {
                StageScope sc(e, ST_LDE);
                lde_columns(e, stage[k], N, logn, gw, lb, Fp::one(), lde->d + c0 * M, M);
            }
            B2_CUDA(cudaEventRecord(consumed[k], e->stream));
            leaf_absorb(e, lde->d + c0 * M, (uint32_t)gw, M, state, leaf, g == 0, g + 1 == ngroups);
        }
        B2_CUDA(cudaStreamSynchronize(e->copy_stream));
        mmcs_build(e, d, leaf);
        leaf = nullptr;
    } catch (...) {
        cudaStreamSynchronize(e->copy_stream);
        cudaStreamSynchronize(e->stream);
        if (leaf) e->dfree(leaf, M * 32);
        cleanup();
        free_pcs(d);
        throw;
    }
    cleanup();
    return d;
}

b2s_pcs_data* commit_matrices(b2s_engine* e, const b2s_matrix* const* mats, const Fp* shifts, size_t n) {
    StageScope total(e, ST_COMMIT);
    b2s_pcs_data* d = new b2s_pcs_data();
    d->eng = e;
    try {
        const uint32_t lb = e->params.log_blowup;
        bool any_side = false;
        for (size_t i = 0; i < n; i++) {
            const b2s_matrix* in = mats[i];
            B2_REQUIRE(in != nullptr, B2S_ERR_BAD_ARG, "null matrix");
            B2_REQUIRE(is_pow2(in->height), B2S_ERR_SHAPE, "commit: trace height must be a power of two");
            B2_REQUIRE(in->width > 0, B2S_ERR_SHAPE, "commit: zero-width trace");
            b2s_matrix* lde = new_matrix(e, in->height << lb, in->width);
            d->mats.push_back(lde);
            d->owns.push_back(true);
            // Short matrices are a handful of latency-bound launches each (a segment commits 34 to 68 of them): when there
            // are several, each goes to a side stream and they overlap; tall ones fill the chip alone and stay in order.
            const bool side = n > 1 && (in->height << lb) * in->width <= ((uint64_t)1 << 24);
            struct Restore { b2s_engine* e; cudaStream_t s; ~Restore() { e->stream = s; } } restore{e, e->stream};
            if (side) { e->stream = aux_fork(e, (unsigned)i); any_side = true; }
            StageScope sc(e, ST_LDE);
            lde_columns(e, in->d, in->height, log2_exact(in->height), in->width, lb, shifts ? shifts[i] : Fp::one(), lde->d,
                        lde->height);
        }
        if (any_side) aux_join(e);
        mmcs_build(e, d);
    } catch (...) {
        // nothing on a side stream may still be writing into the matrices that are about to be released
        for (auto st : e->aux_streams) cudaStreamSynchronize(st);
        if (e->hash_stream) cudaStreamSynchronize(e->hash_stream);
        free_pcs(d);
        throw;
    }
    return d;
}
}  // namespace b2

extern "C" {

int32_t b2s_engine_new(const b2s_fri_params* params, int32_t cuda_device, b2s_engine** out) {
    if (!params || !out) return B2S_ERR_BAD_ARG;
    *out = nullptr;
    return guarded(nullptr, [&] {
        int count = 0;
        cudaError_t err = cudaGetDeviceCount(&count);
        if (err != cudaSuccess || count == 0) {
            cudaGetLastError();
            fail(B2S_ERR_NO_DEVICE, "no CUDA device available: b2stark has no CPU fallback");
        }
        B2_REQUIRE(cuda_device >= 0 && cuda_device < count, B2S_ERR_BAD_ARG, "cuda_device out of range");
        B2_REQUIRE(params->log_blowup >= 1 && params->log_blowup <= 4, B2S_ERR_BAD_ARG, "log_blowup must be in [1,4]");
        B2_CUDA(cudaSetDevice(cuda_device));
        cudaDeviceProp prop;
        B2_CUDA(cudaGetDeviceProperties(&prop, cuda_device));
        B2_REQUIRE(prop.major >= 10, B2S_ERR_NO_DEVICE,
                   std::string("device ") + prop.name + " is not sm_100-class; this library is built for sm_100a only");
        B2_REQUIRE(params->log_final_poly_len == 0, B2S_ERR_BAD_ARG,
                   "log_final_poly_len must be 0 (the prover emits a constant final polynomial)");
        struct Holder {   // the engine is released if stream or pool set-up throws
            b2s_engine* e = new b2s_engine();
            ~Holder() {
                if (!e) return;
                if (e->own_stream && e->stream) cudaStreamDestroy(e->stream);
                delete e;
            }
        } hold;
        b2s_engine* e = hold.e;
        e->device = cuda_device;
        e->num_sms = prop.multiProcessorCount;
        e->params = *params;
        B2_CUDA(cudaStreamCreateWithFlags(&e->stream, cudaStreamNonBlocking));
        e->own_stream = true;
        cudaMemPool_t pool;
        B2_CUDA(cudaDeviceGetDefaultMemPool(&pool, cuda_device));
        uint64_t thresh = UINT64_MAX;
        B2_CUDA(cudaMemPoolSetAttribute(pool, cudaMemPoolAttrReleaseThreshold, &thresh));
        *out = e;
        hold.e = nullptr;
    });
}

void b2s_engine_free(b2s_engine* e) {
    if (!e) return;
    cudaSetDevice(e->device);
    cudaStreamSynchronize(e->stream);
    try {
        e->collect_timers();
        free_twiddles(e);
        free_scales(e);
        if (e->tile_ctr) e->dfree(e->tile_ctr, 256);
        if (e->pinned) cudaFreeHost(e->pinned);
        free_selectors(e);
    } catch (...) {
    }
    e->flush_cache();
    for (auto ev : e->event_pool) cudaEventDestroy(ev);
    cudaStreamSynchronize(e->stream);
    if (e->copy_stream) cudaStreamDestroy(e->copy_stream);
    for (auto st : e->aux_streams) cudaStreamDestroy(st);
    if (e->hash_stream) cudaStreamDestroy(e->hash_stream);
    if (e->own_stream) cudaStreamDestroy(e->stream);
    delete e;
}

int32_t b2s_engine_fri_params(const b2s_engine* e, b2s_fri_params* out) {
    if (!e || !out) return B2S_ERR_BAD_ARG;
    *out = e->params;
    return B2S_OK;
}

const char* b2s_last_error(const b2s_engine* e) {
    if (e) return e->last_error.c_str();
    return g_last_error.c_str();   // this thread's own message
}

void* b2s_engine_stream(const b2s_engine* e) { return e ? (void*)e->stream : nullptr; }

int32_t b2s_engine_set_stream(b2s_engine* e, void* cuda_stream) {
    if (!e) return B2S_ERR_BAD_ARG;
    return guarded(e, [&] {
        B2_CUDA(cudaStreamSynchronize(e->stream));
        e->flush_cache();   // cached blocks are keyed by the stream they were freed on
        if (e->own_stream) B2_CUDA(cudaStreamDestroy(e->stream));
        e->stream = (cudaStream_t)cuda_stream;
        e->own_stream = false;
    });
}

int32_t b2s_engine_synchronize(b2s_engine* e) {
    if (!e) return B2S_ERR_BAD_ARG;
    return guarded(e, [&] { B2_CUDA(cudaStreamSynchronize(e->stream)); });
}

int32_t b2s_reset_metrics(b2s_engine* e) {
    if (!e) return B2S_ERR_BAD_ARG;
    return guarded(e, [&] {
        e->collect_timers();
        for (int s = 0; s < ST_COUNT; s++) e->timers[s].ms = 0;
        for (auto& l : e->launches) l = 0;
        for (auto& l : e->quotient_launches) l = 0;
        e->permutations = 0;
    });
}

int32_t b2s_get_metrics(b2s_engine* e, b2s_metrics* out) {
    if (!e || !out) return B2S_ERR_BAD_ARG;
    return guarded(e, [&] {
        e->collect_timers();
        memset(out, 0, sizeof(*out));
        out->lde_ms = e->timers[ST_LDE].ms;
        out->leaf_hash_ms = e->timers[ST_LEAF].ms;
        out->compress_ms = e->timers[ST_COMPRESS].ms;
        out->commit_total_ms = e->timers[ST_COMMIT].ms;
        out->main_trace_commit_time_ms = e->timers[ST_MAIN_COMMIT].ms;
        out->generate_perm_trace_time_ms = e->timers[ST_PERM_GEN].ms;
        out->perm_trace_commit_time_ms = e->timers[ST_PERM_COMMIT].ms;
        out->quotient_poly_compute_time_ms = e->timers[ST_QUOTIENT].ms;
        out->quotient_poly_commit_time_ms = e->timers[ST_QUOTIENT_COMMIT].ms;
        out->pcs_opening_time_ms = e->timers[ST_OPENING].ms;
        out->ood_eval_ms = e->timers[ST_OOD].ms;
        out->reduced_opening_ms = e->timers[ST_REDUCED].ms;
        out->fri_commit_phase_ms = e->timers[ST_FRI_COMMIT].ms;
        out->stark_prove_excluding_trace_time_ms = e->timers[ST_PROVE].ms;
        out->quotient_jit_launches = e->quotient_launches[0];
        out->quotient_interpreter_launches = e->quotient_launches[1];
        out->lde_launches = e->launches[0];
        out->leaf_hash_launches = e->launches[1];
        out->compress_launches = e->launches[2];
        out->other_launches = e->launches[3];
        out->poseidon2_permutations = e->permutations;
    });
}

int32_t b2s_device_memory(const b2s_engine* e, uint64_t* in_use, uint64_t* peak) {
    if (!e) return B2S_ERR_BAD_ARG;
    if (in_use) *in_use = e->mem_in_use;
    if (peak) *peak = e->mem_peak;
    return B2S_OK;
}

// ---- matrices ------------------------------------------------------------------------------------------------
int32_t b2s_matrix_upload(b2s_engine* e, const uint32_t* host, uint64_t height, uint64_t width, int32_t layout,
                          int32_t repr, b2s_matrix** out) {
    if (!e || !out || (!host && height * width)) return B2S_ERR_BAD_ARG;
    *out = nullptr;
    return guarded(e, [&] {
        B2_REQUIRE(layout == B2S_LAYOUT_ROW_MAJOR || layout == B2S_LAYOUT_COL_MAJOR, B2S_ERR_BAD_ARG, "bad layout");
        B2_REQUIRE(repr == B2S_REPR_CANONICAL || repr == B2S_REPR_MONTY, B2S_ERR_BAD_ARG, "bad repr");
        b2s_matrix* m = new_matrix(e, height, width);
        size_t bytes = m->bytes();
        if (bytes) {
            if (layout == B2S_LAYOUT_COL_MAJOR) {
                B2_CUDA(cudaMemcpyAsync(m->d, host, bytes, cudaMemcpyHostToDevice, e->stream));
                if (repr == B2S_REPR_CANONICAL) convert_repr(e, m->d, (size_t)height * width, true);
            } else {
                uint32_t* tmp = (uint32_t*)e->dmalloc(bytes);
                B2_CUDA(cudaMemcpyAsync(tmp, host, bytes, cudaMemcpyHostToDevice, e->stream));
                transpose_to_colmajor(e, tmp, m->d, height, width, repr == B2S_REPR_CANONICAL);
                e->dfree(tmp, bytes);
            }
        }
        *out = m;
    });
}

int32_t b2s_matrix_download(b2s_engine* e, const b2s_matrix* m, uint32_t* host, int32_t layout, int32_t repr) {
    if (!e || !m || !host) return B2S_ERR_BAD_ARG;
    return guarded(e, [&] {
        size_t bytes = m->bytes();
        if (!bytes) return;
        uint32_t* tmp = (uint32_t*)e->dmalloc(bytes);
        if (layout == B2S_LAYOUT_ROW_MAJOR) {
            transpose_to_rowmajor(e, m->d, tmp, m->height, m->width, repr == B2S_REPR_CANONICAL);
        } else {
            B2_CUDA(cudaMemcpyAsync(tmp, m->d, bytes, cudaMemcpyDeviceToDevice, e->stream));
            if (repr == B2S_REPR_CANONICAL) convert_repr(e, tmp, (size_t)m->height * m->width, false);
        }
        B2_CUDA(cudaMemcpyAsync(host, tmp, bytes, cudaMemcpyDeviceToHost, e->stream));
        B2_CUDA(cudaStreamSynchronize(e->stream));
        e->dfree(tmp, bytes);
    });
}

int32_t b2s_matrix_adopt_device(b2s_engine* e, void* device_ptr, uint64_t height, uint64_t width, b2s_matrix** out) {
    if (!e || !out || !device_ptr) return B2S_ERR_BAD_ARG;
    b2s_matrix* m = new b2s_matrix();
    m->eng = e;
    m->d = (uint32_t*)device_ptr;
    m->height = height;
    m->width = width;
    m->owned = false;
    *out = m;
    return B2S_OK;
}

int32_t b2s_matrix_alloc(b2s_engine* e, uint64_t height, uint64_t width, b2s_matrix** out) {
    if (!e || !out) return B2S_ERR_BAD_ARG;
    *out = nullptr;
    return guarded(e, [&] { *out = new_matrix(e, height, width); });
}

int32_t b2s_matrix_fill_synthetic(b2s_engine* e, b2s_matrix* m, uint64_t seed) {
    if (!e || !m) return B2S_ERR_BAD_ARG;
    return guarded(e, [&] { fill_synthetic(e, m->d, m->height, m->width, seed); });
}

int32_t b2s_matrix_fill_chain(b2s_engine* e, b2s_matrix* m, uint64_t seed, uint32_t pv_out[2]) {
    if (!e || !m || !pv_out) return B2S_ERR_BAD_ARG;
    return guarded(e, [&] { fill_chain(e, m->d, m->height, m->width, seed, pv_out); });
}

uint64_t b2s_matrix_height(const b2s_matrix* m) { return m ? m->height : 0; }
uint64_t b2s_matrix_width(const b2s_matrix* m) { return m ? m->width : 0; }
void* b2s_matrix_device_ptr(const b2s_matrix* m) { return m ? (void*)m->d : nullptr; }

void b2s_matrix_free(b2s_matrix* m) {
    if (!m) return;
    if (m->owned && m->d) {
        cudaSetDevice(m->eng->device);
        m->eng->dfree(m->d, m->bytes());
    }
    delete m;
}

// ---- commitment --------------------------------------------------------------------------------------------
int32_t b2s_coset_lde(b2s_engine* e, const b2s_matrix* in, uint32_t log_blowup, uint32_t shift_in, b2s_matrix** out) {
    if (!e || !in || !out) return B2S_ERR_BAD_ARG;
    *out = nullptr;
    return guarded(e, [&] {
        B2_REQUIRE(is_pow2(in->height), B2S_ERR_SHAPE, "lde: height must be a power of two");
        B2_REQUIRE(shift_in > 0 && shift_in < P, B2S_ERR_BAD_ARG, "lde: shift must be a non-zero canonical element");
        unsigned logn = log2_exact(in->height);
        b2s_matrix* m = new_matrix(e, in->height << log_blowup, in->width);
        try {
            StageScope sc(e, ST_LDE);
            lde_columns(e, in->d, in->height, logn, in->width, log_blowup, Fp::from_canonical(shift_in), m->d, m->height);
        } catch (...) {
            b2s_matrix_free(m);
            throw;
        }
        *out = m;
    });
}

int32_t b2s_shared_alloc(b2s_engine* e, uint64_t bytes, void** device_ptr, uint8_t handle64[64]) {
    if (!e || !device_ptr || !handle64 || !bytes) return B2S_ERR_BAD_ARG;
    *device_ptr = nullptr;
    return guarded(e, [&] {
        static_assert(sizeof(cudaIpcMemHandle_t) == 64, "ipc handle size");
        void* p = nullptr;
        B2_CUDA(cudaMalloc(&p, bytes));   // not from the stream-ordered pool: pool memory cannot be exported
        cudaIpcMemHandle_t h;
        cudaError_t err = cudaIpcGetMemHandle(&h, p);
        if (err != cudaSuccess) {
            cudaFree(p);
            B2_CUDA(err);
        }
        memcpy(handle64, &h, 64);
        *device_ptr = p;
    });
}
int32_t b2s_shared_free(b2s_engine* e, void* device_ptr) {
    if (!e) return B2S_ERR_BAD_ARG;
    return guarded(e, [&] {
        B2_CUDA(cudaStreamSynchronize(e->stream));
        B2_CUDA(cudaFree(device_ptr));
    });
}
int32_t b2s_shared_open(b2s_engine* e, const uint8_t handle64[64], void** device_ptr) {
    if (!e || !device_ptr || !handle64) return B2S_ERR_BAD_ARG;
    *device_ptr = nullptr;
    return guarded(e, [&] {
        cudaIpcMemHandle_t h;
        memcpy(&h, handle64, 64);
        B2_CUDA(cudaIpcOpenMemHandle(device_ptr, h, cudaIpcMemLazyEnablePeerAccess));
    });
}
int32_t b2s_shared_close(b2s_engine* e, void* device_ptr) {
    if (!e) return B2S_ERR_BAD_ARG;
    return guarded(e, [&] {
        B2_CUDA(cudaStreamSynchronize(e->stream));
        B2_CUDA(cudaIpcCloseMemHandle(device_ptr));
    });
}
int32_t b2s_coset_lde_scatter(b2s_engine* e, const b2s_matrix* in, uint32_t log_blowup, uint32_t shift_in,
                              void* const* peer_bufs, uint32_t n_peers, uint64_t total_width, uint64_t col0) {
    if (!e || !in || !peer_bufs) return B2S_ERR_BAD_ARG;
    return guarded(e, [&] {
        B2_REQUIRE(is_pow2(in->height), B2S_ERR_SHAPE, "lde: height must be a power of two");
        B2_REQUIRE(shift_in > 0 && shift_in < P, B2S_ERR_BAD_ARG, "lde: shift must be a non-zero canonical element");
        B2_REQUIRE(n_peers >= 1 && n_peers <= B2_MAX_PEERS && is_pow2(n_peers), B2S_ERR_BAD_ARG,
                   "lde scatter: the number of peers must be a power of two, at most 16");
        B2_REQUIRE(col0 + in->width <= total_width, B2S_ERR_SHAPE, "lde scatter: column shard outside the trace");
        const uint64_t rows = in->height << log_blowup;
        B2_REQUIRE(rows / n_peers >= 4, B2S_ERR_SHAPE, "lde scatter: fewer than 4 rows per peer");
        PeerScatter ps{};
        ps.n_peers = n_peers;
        ps.log_rows = log2_exact(rows / n_peers);
        ps.col0 = col0;
        for (uint32_t i = 0; i < n_peers; i++) {
            B2_REQUIRE(peer_bufs[i] != nullptr, B2S_ERR_BAD_ARG, "lde scatter: null peer buffer");
            ps.peer[i] = (uint32_t*)peer_bufs[i];
        }
        StageScope sc(e, ST_LDE);
        lde_columns(e, in->d, in->height, log2_exact(in->height), in->width, log_blowup, Fp::from_canonical(shift_in), nullptr, 0,
                    &ps);
    });
}

static void root_to_host(b2s_engine* e, b2s_pcs_data* d, uint32_t root8[8]) {
    uint32_t raw[8];
    B2_CUDA(cudaMemcpyAsync(raw, d->layers.back(), 32, cudaMemcpyDeviceToHost, e->stream));
    B2_CUDA(cudaStreamSynchronize(e->stream));
    for (int i = 0; i < 8; i++) root8[i] = e->hash_kind ? raw[i] : Fp::raw(raw[i]).to_canonical();
}

int32_t b2s_mmcs_commit(b2s_engine* e, const b2s_matrix* const* mats, size_t n_mats, uint32_t root8[8],
                        b2s_pcs_data** out) {
    if (!e || !mats || !n_mats || !root8 || !out) return B2S_ERR_BAD_ARG;
    *out = nullptr;
    return guarded(e, [&] {
        b2s_pcs_data* d = new b2s_pcs_data();
        d->eng = e;
        for (size_t i = 0; i < n_mats; i++) {
            B2_REQUIRE(mats[i] != nullptr, B2S_ERR_BAD_ARG, "null matrix");
            b2s_matrix* view = new b2s_matrix(*mats[i]);
            view->owned = false;
            d->mats.push_back(view);
            d->owns.push_back(true);  // frees the view object only (owned == false keeps the buffer)
        }
        try {
            mmcs_build(e, d);
            root_to_host(e, d, root8);
        } catch (...) {
            free_pcs(d);
            throw;
        }
        *out = d;
    });
}

int32_t b2s_commit(b2s_engine* e, const b2s_matrix* const* mats, size_t n_mats, uint32_t root8[8], b2s_pcs_data** out) {
    if (!e || !mats || !n_mats || !root8 || !out) return B2S_ERR_BAD_ARG;
    *out = nullptr;
    return guarded(e, [&] {
        b2s_pcs_data* d = commit_matrices(e, mats, nullptr, n_mats);
        try {
            root_to_host(e, d, root8);
        } catch (...) {
            free_pcs(d);
            throw;
        }
        *out = d;
    });
}

// ---- proving key / prove ---------------------------------------------------------------------------------------
int32_t b2s_pk_load(b2s_engine* e, const uint32_t* key_words, size_t n_words, b2s_pk** out) {
    if (!e || !key_words || !out) return B2S_ERR_BAD_ARG;
    *out = nullptr;
    return guarded(e, [&] { *out = pk_load(e, key_words, n_words); });
}

void b2s_pk_free(b2s_pk* pk) {
    if (!pk) return;
    cudaSetDevice(pk->eng->device);
    pk_free(pk);
}

int32_t b2s_prove(b2s_engine* e, const b2s_pk* pk, const b2s_air_ctx* airs, size_t n_airs, uint8_t** proof_bytes,
                  size_t* proof_len) {
    if (!e || !pk || !airs || !n_airs || !proof_bytes || !proof_len) return B2S_ERR_BAD_ARG;
    *proof_bytes = nullptr;
    *proof_len = 0;
    return guarded(e, [&] {
        std::vector<uint8_t> bytes;
        prove(e, pk, airs, n_airs, bytes);
        uint8_t* buf = (uint8_t*)malloc(bytes.size() ? bytes.size() : 1);
        if (!buf) fail(B2S_ERR_OOM, "proof buffer allocation failed");
        memcpy(buf, bytes.data(), bytes.size());
        *proof_bytes = buf;
        *proof_len = bytes.size();
    });
}

int32_t b2s_commit_host(b2s_engine* e, const uint32_t* host, uint64_t height, uint64_t width, int32_t layout, int32_t repr,
                        uint32_t* root8, b2s_pcs_data** out) {
    if (!e || !host || !out) return B2S_ERR_BAD_ARG;
    *out = nullptr;
    return guarded(e, [&] {
        B2_REQUIRE((layout == B2S_LAYOUT_ROW_MAJOR || layout == B2S_LAYOUT_COL_MAJOR) &&
                       (repr == B2S_REPR_CANONICAL || repr == B2S_REPR_MONTY),
                   B2S_ERR_BAD_ARG, "commit_host: bad layout / representation");
        b2s_pcs_data* d = commit_host_streamed(e, host, height, width, layout, repr);
        if (root8) {
            try {
                uint32_t r[8];
                B2_CUDA(cudaMemcpyAsync(r, d->layers.back(), 32, cudaMemcpyDeviceToHost, e->stream));
                B2_CUDA(cudaStreamSynchronize(e->stream));
                for (int k = 0; k < 8; k++) root8[k] = e->hash_kind ? r[k] : Fp::raw(r[k]).to_canonical();
            } catch (...) {
                free_pcs(d);
                throw;
            }
        }
        *out = d;
    });
}

int32_t b2s_engine_set_pow_witnesses(b2s_engine* e, uint32_t logup_witness, uint32_t fri_witness) {
    if (!e) return B2S_ERR_BAD_ARG;
    return guarded(e, [&] {
        B2_REQUIRE((logup_witness == B2S_POW_SEARCH || logup_witness < P) && (fri_witness == B2S_POW_SEARCH || fri_witness < P),
                   B2S_ERR_BAD_ARG, "pow witnesses must be canonical field elements or B2S_POW_SEARCH");
        e->pow_override[0] = logup_witness;
        e->pow_override[1] = fri_witness;
    });
}

int32_t b2s_debug_quotient_ptx(const uint32_t* key_words, size_t n_words, uint32_t log_blowup, uint32_t air_index, char** ptx_out) {
    if (!key_words || !ptx_out) return B2S_ERR_BAD_ARG;
    *ptx_out = nullptr;
    return guarded(nullptr, [&] {
        b2s_pk* pk = pk_load(nullptr, key_words, n_words, true, log_blowup);
        try {
            B2_REQUIRE(air_index < pk->airs.size(), B2S_ERR_BAD_ARG, "air index out of range");
            const std::string ptx = quotient_jit_ptx(pk->airs[air_index]);
            char* buf = (char*)malloc(ptx.size() + 1);
            if (!buf) fail(B2S_ERR_OOM, "ptx buffer allocation failed");
            memcpy(buf, ptx.c_str(), ptx.size() + 1);
            *ptx_out = buf;
        } catch (...) {
            pk_free(pk);
            throw;
        }
        pk_free(pk);
    });
}

int32_t b2s_verify(b2s_engine* e, const b2s_pk* pk, const uint8_t* proof_bytes, size_t proof_len) {
    if (!e || !pk || !proof_bytes) return B2S_ERR_BAD_ARG;
    return guarded(e, [&] {
        B2_REQUIRE(pk->eng == e, B2S_ERR_BAD_ARG, "verify: key was loaded on a different engine");
        verify(e, pk, proof_bytes, proof_len);
    });
}

// The verifier is host code (transcript, Merkle paths, constraint evaluation at zeta): this entry needs no device and no
// engine -- the key blob is parsed on the host (preprocessed traces are skipped, their commitments kept).
int32_t b2s_verify_host(const b2s_fri_params* params, int32_t hash_kind, const uint32_t* key_words, size_t n_words,
                        const uint8_t* proof_bytes, size_t proof_len) {
    if (!params || !key_words || !proof_bytes) return B2S_ERR_BAD_ARG;
    return guarded(nullptr, [&] {
        B2_REQUIRE(hash_kind == B2S_HASH_POSEIDON2_BABYBEAR || hash_kind == B2S_HASH_POSEIDON2_BN254, B2S_ERR_BAD_ARG, "unknown hash kind");
        B2_REQUIRE(params->log_blowup >= 1 && params->log_blowup <= 4 && params->log_final_poly_len == 0, B2S_ERR_BAD_ARG,
                   "verify: unsupported FRI parameters");
        b2s_engine host;          // parameters only: no stream, no device memory, nothing to release
        host.params = *params;
        host.hash_kind = hash_kind;
        b2s_pk* pk = pk_load(&host, key_words, n_words, true, params->log_blowup);
        try {
            verify(&host, pk, proof_bytes, proof_len);
        } catch (...) {
            pk_free(pk);
            throw;
        }
        pk_free(pk);
    });
}

int32_t b2s_proof_hint_stream(b2s_engine* e, const uint8_t* proof_bytes, size_t proof_len, uint32_t** words, size_t* n_words) {
    if (!proof_bytes || !words || !n_words) return B2S_ERR_BAD_ARG;
    *words = nullptr;
    *n_words = 0;
    return guarded(e, [&] {
        // the hint layout is the inner verifier program's (hints.rs); root proofs feed the static halo2 verifier instead
        B2_REQUIRE(!e || e->hash_kind == 0, B2S_ERR_BAD_ARG, "hint stream: defined for BabyBearPoseidon2Config proofs only");
        const std::vector<uint32_t> v = proof_hint_stream(proof_bytes, proof_len);
        uint32_t* out = (uint32_t*)malloc(v.size() * 4);
        B2_REQUIRE(out != nullptr, B2S_ERR_OOM, "hint stream: out of host memory");
        memcpy(out, v.data(), v.size() * 4);
        *words = out;
        *n_words = v.size();
    });
}

int32_t b2s_proof_witness_stream(b2s_engine* e, const uint8_t* proof_bytes, size_t proof_len, uint32_t** words, size_t* n_words) {
    if (!proof_bytes || !words || !n_words) return B2S_ERR_BAD_ARG;
    *words = nullptr;
    *n_words = 0;
    return guarded(e, [&] {
        const std::vector<uint32_t> v = proof_witness_stream(proof_bytes, proof_len);
        uint32_t* out = (uint32_t*)malloc(v.size() * 4);
        B2_REQUIRE(out != nullptr, B2S_ERR_OOM, "witness stream: out of host memory");
        memcpy(out, v.data(), v.size() * 4);
        *words = out;
        *n_words = v.size();
    });
}

size_t b2s_pcs_num_matrices(const b2s_pcs_data* d) { return d ? d->mats.size() : 0; }
uint32_t b2s_pcs_log_max_height(const b2s_pcs_data* d) { return d ? d->log_max_height : 0; }
const b2s_matrix* b2s_pcs_matrix(const b2s_pcs_data* d, size_t i) { return (d && i < d->mats.size()) ? d->mats[i] : nullptr; }

int32_t b2s_pcs_open(b2s_engine* e, const b2s_pcs_data* d, uint64_t index, uint32_t* rows_out, uint32_t* proof_out) {
    if (!e || !d) return B2S_ERR_BAD_ARG;
    return guarded(e, [&] { gather_open(e, d, index, rows_out, proof_out); });
}

int32_t b2s_pcs_layer(b2s_engine* e, const b2s_pcs_data* d, uint32_t level, uint32_t* out) {
    if (!e || !d || !out) return B2S_ERR_BAD_ARG;
    return guarded(e, [&] { download_layer(e, d, level, out); });
}

void b2s_pcs_free(b2s_pcs_data* d) {
    if (!d) return;
    cudaSetDevice(d->eng->device);
    free_pcs(d);
}

// ---- primitives ----------------------------------------------------------------------------------------------
int32_t b2s_poseidon2_permute(b2s_engine* e, uint32_t* host_states, size_t n_states) {
    if (!e || (!host_states && n_states)) return B2S_ERR_BAD_ARG;
    return guarded(e, [&] {
        if (!n_states) return;
        size_t bytes = n_states * 64;
        uint32_t* d = (uint32_t*)e->dmalloc(bytes);
        B2_CUDA(cudaMemcpyAsync(d, host_states, bytes, cudaMemcpyHostToDevice, e->stream));
        poseidon2_permute_states(e, d, n_states);
        B2_CUDA(cudaMemcpyAsync(host_states, d, bytes, cudaMemcpyDeviceToHost, e->stream));
        B2_CUDA(cudaStreamSynchronize(e->stream));
        e->dfree(d, bytes);
    });
}

int32_t b2s_engine_set_hash(b2s_engine* e, int32_t kind) {
    if (!e) return B2S_ERR_BAD_ARG;
    return guarded(e, [&] {
        B2_REQUIRE(kind == B2S_HASH_POSEIDON2_BABYBEAR || kind == B2S_HASH_POSEIDON2_BN254, B2S_ERR_BAD_ARG, "unknown hash kind");
        e->hash_kind = kind;
    });
}
int32_t b2s_engine_hash(const b2s_engine* e) { return e ? e->hash_kind : B2S_ERR_BAD_ARG; }

int32_t b2s_poseidon2_bn254_permute(b2s_engine* e, uint32_t* host_states, size_t n_states) {
    if (!e || (!host_states && n_states)) return B2S_ERR_BAD_ARG;
    return guarded(e, [&] {
        if (!n_states) return;
        size_t bytes = n_states * 96;
        uint32_t* d = (uint32_t*)e->dmalloc(bytes);
        B2_CUDA(cudaMemcpyAsync(d, host_states, bytes, cudaMemcpyHostToDevice, e->stream));
        bn_permute_states(e, d, n_states);
        B2_CUDA(cudaMemcpyAsync(host_states, d, bytes, cudaMemcpyDeviceToHost, e->stream));
        B2_CUDA(cudaStreamSynchronize(e->stream));
        e->dfree(d, bytes);
    });
}

// the same permutation on the host (the verifier and the transcript of the root configuration hash on the CPU)
int32_t b2s_host_poseidon2_bn254_permute(uint32_t* states, size_t n_states) {
    if (!states && n_states) return B2S_ERR_BAD_ARG;
    for (size_t i = 0; i < n_states; i++) {
        Fr s[3];
        for (int k = 0; k < 3; k++) {
            uint32_t w[8];
            for (int j = 0; j < 8; j++) w[j] = states[24 * i + 8 * k + j];
            s[k] = fr_from_words(w);
        }
        poseidon2_bn254_permute(s);
        for (int k = 0; k < 3; k++) {
            uint32_t w[8];
            fr_to_words(s[k], w);
            for (int j = 0; j < 8; j++) states[24 * i + 8 * k + j] = w[j];
        }
    }
    return B2S_OK;
}

int32_t b2s_bench_modmul(b2s_engine* e, uint64_t threads, uint32_t iters, int32_t ilp, double* ms_out) {
    if (!e || !ms_out) return B2S_ERR_BAD_ARG;
    return guarded(e, [&] { *ms_out = bench_modmul(e, threads, iters, ilp); });
}
int32_t b2s_bench_poseidon2(b2s_engine* e, uint64_t threads, uint32_t iters, double* ms_out) {
    if (!e || !ms_out) return B2S_ERR_BAD_ARG;
    return guarded(e, [&] { *ms_out = bench_poseidon2(e, threads, iters); });
}

void b2s_free(void* p) { free(p); }

}  // extern "C"
