// StarkEngine::prove on the device: quotient evaluation, out-of-domain opening, FRI commit phase, proof-of-work
// and query openings, driven by a host-side Fiat-Shamir transcript, emitted in the reference's proof wire format.
//
// Replaces (reference, external): openvm-stark-backend `MultiTraceStarkProver::prove` / `QuotientCommitter` /
// `OpeningProver` and Plonky3 `TwoAdicFriPcs::open`, reached from crates/vm/src/arch/vm.rs:694,1005. The in-tree
// specification followed step by step is the verifier twin:
//   extensions/native/recursion/src/stark/mod.rs:130-834,903-1126   transcript, rounds, constraint folding
//   extensions/native/recursion/src/fri/two_adic_pcs.rs:41-376      opened-value order, reduced openings, queries
//   extensions/native/recursion/src/fri/mod.rs:33-170               fold formula and beta^2 roll-in
//   extensions/native/recursion/src/fri/domain.rs:59-155            selectors, quotient chunk domains
//   crates/sdk/src/codec.rs:92-330                                  wire format
//
// Device work (all integer-pipe bound, one thread per row unless noted):
//   quotient_kernel       interprets the AIR's register-allocated constraint program on every row of the quotient
//                         domain (= the first N*qd rows of the bit-reversed LDE), folds with alpha, divides by Z_H
//   bary_weights_kernel   x_i / (z - x_i) for the N points of LDE block 0, once per (height, point)
//   bary_dot_multi_kernel per column sum_i p(x_i) * weight_i  -> p(z); all columns of one (height, point) per launch
//   reduced_open_height_kernel   ro_h[row] = sum over the matrices of height h of alpha^o * (sum_t alpha^t p_t(z) - sum_t alpha^t p_t(x)) / (z - x)
//   fri_pack / fri_fold   commit-phase leaves (2 EF = 8 words per row) and the arity-2 fold with beta^2 roll-in
//   pow_grind_kernel      parallel search for the minimum proof-of-work witness
//   gather_rows_kernel    all query openings (rows and Merkle siblings) of all rounds in one launch
#include "common.cuh"
#include "host_challenger.hpp"
#include "poseidon2.cuh"

#include <algorithm>
#include <exception>
#include <array>
#include <chrono>
#include <deque>
#include <atomic>
#include <map>
#include <thread>

namespace b2 {

// ---- key blob -> device key --------------------------------------------------------------------------------
enum NodeKind { N_CONST = 0, N_PREP, N_MAIN, N_PERM, N_PUBLIC, N_CHALLENGE, N_EXPOSED, N_FIRST, N_LAST, N_TRANS,
                N_ADD, N_SUB, N_NEG, N_MUL };
enum Op { OP_LOAD = 0, OP_PUBLIC, OP_CONST, OP_FIRST, OP_LAST, OP_TRANS, OP_ADD, OP_SUB, OP_NEG, OP_MUL, OP_ACC,
          // extension-field values (after-challenge columns, challenges, exposed values and what depends on them)
          OP_LOADP, OP_CHAL, OP_EXPOSED, OP_ADD_EE, OP_ADD_EF, OP_SUB_EE, OP_SUB_EF, OP_SUB_FE, OP_NEG_E, OP_MUL_EE,
          OP_MUL_EF, OP_ACC_E,
          // interaction program (LogUp trace generation)
          OP_IFIELD, OP_IEND };

struct HostNode { uint32_t kind, a, b, c; };
struct HostInteraction { uint32_t bus, type, count_node; std::vector<uint32_t> fields; };

// Compile the DAG (topological node list) into a straight-line program over two small slot files (base field and
// extension field): dead nodes are dropped, a slot is recycled after the last use of its value.
//   quotient mode:    roots = constraints; constraint k of C becomes OP_ACC / OP_ACC_E with the index of its alpha
//                     power (alpha^(C-1-k): Horner order of the reference's constraint folder)
//   interaction mode: roots = message fields and counts of the interactions; after the node code, per chunk and per
//                     interaction: OP_IFIELD (slot, j) ... OP_IEND (flags: bit0 receive, bit1 last of its chunk)
struct Compiled {
    std::vector<uint4> program;
    uint32_t slots_f = 1, slots_e = 1;
};
static Compiled compile_dag(const AirKeyDev& air, const std::vector<HostNode>& nodes, const std::vector<uint32_t>& cidx,
                            const std::vector<HostInteraction>* inter, const std::vector<std::vector<uint32_t>>* chunks) {
    const size_t n = nodes.size();
    const uint32_t NEVER = 0xFFFFFFFFu;
    std::vector<uint32_t> last_use(n, 0);
    std::vector<char> live(n, 0), is_e(n, 0);
    auto root = [&](uint32_t c, bool pin) {
        B2_REQUIRE(c < n, B2S_ERR_BAD_ARG, "key: node index out of range");
        live[c] = 1;
        if (pin) last_use[c] = NEVER;
    };
    if (inter) {
        for (auto& it : *inter) { root(it.count_node, true); for (uint32_t f : it.fields) root(f, true); }
    } else {
        for (uint32_t c : cidx) root(c, false);
    }
    for (size_t i = n; i-- > 0;) {
        if (!live[i]) continue;
        const HostNode& nd = nodes[i];
        if (nd.kind >= N_ADD) {
            B2_REQUIRE(nd.a < i && (nd.kind == N_NEG || nd.b < i), B2S_ERR_BAD_ARG, "key: DAG is not topologically ordered");
            live[nd.a] = 1;
            if (nd.kind != N_NEG) live[nd.b] = 1;
        }
    }
    for (size_t i = 0; i < n; i++) {
        if (!live[i]) continue;
        const HostNode& nd = nodes[i];
        if (last_use[i] != NEVER) last_use[i] = std::max<uint32_t>(last_use[i], (uint32_t)i);
        if (nd.kind == N_PERM || nd.kind == N_CHALLENGE || nd.kind == N_EXPOSED) is_e[i] = 1;
        if (nd.kind >= N_ADD) {
            if (last_use[nd.a] != NEVER) last_use[nd.a] = (uint32_t)i;
            if (nd.kind != N_NEG && last_use[nd.b] != NEVER) last_use[nd.b] = (uint32_t)i;
            is_e[i] = is_e[nd.a] || (nd.kind != N_NEG && is_e[nd.b]);
        }
        B2_REQUIRE(!(inter && is_e[i]), B2S_ERR_BAD_ARG, "key: interaction message depends on a challenge-phase value");
    }
    // column index of a variable inside the AIR's pointer table: [preprocessed][part 0][part 1]...[after-challenge]
    std::vector<uint32_t> part_base;
    uint32_t base = air.width_prep;
    for (uint32_t w : air.cached_widths) { part_base.push_back(base); base += w; }
    if (air.width_common) { part_base.push_back(base); base += air.width_common; }
    const uint32_t perm_base = base;
    auto part_width = [&](uint32_t p) { return p < air.cached_widths.size() ? air.cached_widths[p] : air.width_common; };

    Compiled out;
    std::vector<uint32_t> slot_of(n, NEVER), free_f, free_e;
    std::vector<std::vector<uint32_t>> expiring(n);
    for (size_t i = 0; i < n; i++) if (live[i] && last_use[i] != NEVER) expiring[last_use[i]].push_back((uint32_t)i);
    uint32_t nf = 0, ne = 0;
    const uint32_t C = (uint32_t)cidx.size();
    std::vector<std::vector<uint32_t>> acc_of(n);
    if (!inter) for (uint32_t k = 0; k < C; k++) acc_of[cidx[k]].push_back(C - 1 - k);
    auto emit = [&](uint32_t op, uint32_t d, uint32_t a, uint32_t b2) { out.program.push_back(make_uint4(op, d, a, b2)); };
    for (size_t i = 0; i < n; i++) {
        if (!live[i]) continue;
        const HostNode& nd = nodes[i];
        uint32_t sa = 0, sb = 0;
        bool ea = false, eb = false;
        if (nd.kind >= N_ADD) {
            sa = slot_of[nd.a]; ea = is_e[nd.a];
            if (nd.kind != N_NEG) { sb = slot_of[nd.b]; eb = is_e[nd.b]; }
        }
        uint32_t dst;
        if (is_e[i]) { if (!free_e.empty()) { dst = free_e.back(); free_e.pop_back(); } else dst = ne++; }
        else { if (!free_f.empty()) { dst = free_f.back(); free_f.pop_back(); } else dst = nf++; }
        slot_of[i] = dst;
        switch (nd.kind) {
            case N_CONST:
                B2_REQUIRE(nd.a < P, B2S_ERR_BAD_ARG, "key: non-canonical constant");
                emit(OP_CONST, dst, Fp::from_canonical(nd.a).v, 0);
                break;
            case N_PREP:
                B2_REQUIRE(nd.a < air.width_prep && nd.b < 2, B2S_ERR_BAD_ARG, "key: preprocessed variable out of range");
                emit(OP_LOAD, dst, nd.a, nd.b);
                break;
            case N_MAIN:
                B2_REQUIRE(nd.c < part_base.size() && nd.a < part_width(nd.c) && nd.b < 2, B2S_ERR_BAD_ARG,
                           "key: main variable out of range");
                emit(OP_LOAD, dst, part_base[nd.c] + nd.a, nd.b);
                break;
            case N_PERM:
                B2_REQUIRE(nd.a < air.width_after && nd.b < 2, B2S_ERR_BAD_ARG, "key: after-challenge variable out of range");
                emit(OP_LOADP, dst, perm_base + 4 * nd.a, nd.b);
                break;
            case N_PUBLIC:
                B2_REQUIRE(nd.a < air.num_pv, B2S_ERR_BAD_ARG, "key: public value out of range");
                emit(OP_PUBLIC, dst, nd.a, 0);
                break;
            case N_CHALLENGE:
                B2_REQUIRE(nd.a < 2 && air.width_after, B2S_ERR_BAD_ARG, "key: challenge out of range");
                emit(OP_CHAL, dst, nd.a, 0);
                break;
            case N_EXPOSED:
                B2_REQUIRE(nd.a < air.num_exposed, B2S_ERR_BAD_ARG, "key: exposed value out of range");
                emit(OP_EXPOSED, dst, nd.a, 0);
                break;
            case N_FIRST: emit(OP_FIRST, dst, 0, 0); break;
            case N_LAST: emit(OP_LAST, dst, 0, 0); break;
            case N_TRANS: emit(OP_TRANS, dst, 0, 0); break;
            case N_ADD:
                if (!ea && !eb) emit(OP_ADD, dst, sa, sb);
                else if (ea && eb) emit(OP_ADD_EE, dst, sa, sb);
                else if (ea) emit(OP_ADD_EF, dst, sa, sb);
                else emit(OP_ADD_EF, dst, sb, sa);
                break;
            case N_SUB:
                if (!ea && !eb) emit(OP_SUB, dst, sa, sb);
                else if (ea && eb) emit(OP_SUB_EE, dst, sa, sb);
                else if (ea) emit(OP_SUB_EF, dst, sa, sb);
                else emit(OP_SUB_FE, dst, sa, sb);
                break;
            case N_NEG: emit(ea ? OP_NEG_E : OP_NEG, dst, sa, 0); break;
            case N_MUL:
                if (!ea && !eb) emit(OP_MUL, dst, sa, sb);
                else if (ea && eb) emit(OP_MUL_EE, dst, sa, sb);
                else if (ea) emit(OP_MUL_EF, dst, sa, sb);
                else emit(OP_MUL_EF, dst, sb, sa);
                break;
            default: fail(B2S_ERR_BAD_ARG, "key: unknown node kind");
        }
        for (uint32_t pw : acc_of[i]) emit(is_e[i] ? OP_ACC_E : OP_ACC, 0, dst, pw);
        for (uint32_t j : expiring[i]) (is_e[j] ? free_e : free_f).push_back(slot_of[j]);
    }
    if (inter) {
        for (auto& ch : *chunks)
            for (size_t q = 0; q < ch.size(); q++) {
                B2_REQUIRE(ch[q] < inter->size(), B2S_ERR_BAD_ARG, "key: interaction chunk index out of range");
                const HostInteraction& it = (*inter)[ch[q]];
                for (size_t j = 0; j < it.fields.size(); j++) emit(OP_IFIELD, 0, slot_of[it.fields[j]], (uint32_t)j);
                emit(OP_IEND, (it.type ? 1u : 0u) | (q + 1 == ch.size() ? 2u : 0u), slot_of[it.count_node], ch[q]);
            }
    }
    out.slots_f = std::max<uint32_t>(nf, 1);
    out.slots_e = std::max<uint32_t>(ne, 1);
    return out;
}

// Every node of the DAG, live or dead, must be well-formed: the verifier evaluates all of them at zeta
// (verifier.cu fold_constraints), so a malformed dead node would load, prove, and then fail every verification.
static void validate_nodes(const AirKeyDev& air, const std::vector<HostNode>& nodes) {
    const size_t parts = air.cached_widths.size() + (air.width_common ? 1 : 0);
    auto part_width = [&](uint32_t p) { return p < air.cached_widths.size() ? air.cached_widths[p] : air.width_common; };
    for (size_t i = 0; i < nodes.size(); i++) {
        const HostNode& nd = nodes[i];
        bool ok;
        switch (nd.kind) {
            case N_CONST: ok = nd.a < P; break;
            case N_PREP: ok = nd.a < air.width_prep && nd.b < 2; break;
            case N_MAIN: ok = nd.c < parts && nd.a < part_width(nd.c) && nd.b < 2; break;
            case N_PERM: ok = nd.a < air.width_after && nd.b < 2; break;
            case N_PUBLIC: ok = nd.a < air.num_pv; break;
            case N_CHALLENGE: ok = nd.a < 2 && air.width_after; break;
            case N_EXPOSED: ok = nd.a < air.num_exposed; break;
            case N_FIRST: case N_LAST: case N_TRANS: ok = true; break;
            case N_ADD: case N_SUB: case N_MUL: ok = nd.a < i && nd.b < i; break;
            case N_NEG: ok = nd.a < i; break;
            default: ok = false;
        }
        B2_REQUIRE(ok, B2S_ERR_BAD_ARG, "key: malformed DAG node " + std::to_string(i) + " (kind " + std::to_string(nd.kind) + ")");
    }
}

static void root_monty(b2s_engine* e, const b2s_pcs_data* d, uint32_t out[8]) {
    B2_CUDA(cudaMemcpyAsync(out, d->layers.back(), 32, cudaMemcpyDeviceToHost, e->stream));
    B2_CUDA(cudaStreamSynchronize(e->stream));
}

void pk_free(b2s_pk* pk) {
    if (!pk) return;
    b2s_engine* e = pk->eng;
    if (!e) { delete pk; return; }
    for (auto& a : pk->airs) {
        if (a.prep_pcs) free_pcs(a.prep_pcs);
        if (a.prep_trace) b2s_matrix_free(a.prep_trace);
        if (a.d_program) e->dfree(a.d_program, a.program.size() * sizeof(uint4));
        if (a.d_iprogram) e->dfree(a.d_iprogram, a.iprogram.size() * sizeof(uint4));
    }
    delete pk;
}

// host_only: parse and validate without touching a device (e may be null); used to inspect the generated quotient PTX
b2s_pk* pk_load(b2s_engine* e, const uint32_t* w, size_t n, bool host_only, uint32_t log_blowup) {
    if (!host_only) log_blowup = e->params.log_blowup;
    size_t i = 0;
    auto rd = [&]() -> uint32_t {
        if (i >= n) fail(B2S_ERR_BAD_ARG, "key blob truncated");
        return w[i++];
    };
    B2_REQUIRE(rd() == 0x4B503242u, B2S_ERR_BAD_ARG, "key blob: bad magic");
    const uint32_t version = rd();   // 2: no trace-height constraints; 3: they follow the AIRs
    B2_REQUIRE(version == 2 || version == 3, B2S_ERR_BAD_ARG, "key blob: unsupported version");
    b2s_pk* pk = new b2s_pk();
    pk->eng = host_only ? nullptr : e;
    pk->hash_kind = e ? e->hash_kind : 0;
    const bool root = pk->hash_kind != 0;
    try {
        for (int j = 0; j < 8; j++) {
            uint32_t v = rd();
            B2_REQUIRE(root || v < P, B2S_ERR_BAD_ARG, "key blob: non-canonical pre_hash");
            pk->pre_hash[j] = root ? v : Fp::from_canonical(v).v;
        }
        pk->log_up_pow_bits = rd();
        const uint32_t na = rd();
        B2_REQUIRE(na <= (1u << 20), B2S_ERR_BAD_ARG, "key blob: AIR count");
        pk->airs.resize(na);
        for (auto& a : pk->airs) {
            a.width_prep = rd();
            uint32_t nc = rd();
            B2_REQUIRE(nc <= 1024, B2S_ERR_BAD_ARG, "key blob: cached main count");
            a.cached_widths.resize(nc);
            for (auto& c : a.cached_widths) c = rd();
            a.width_common = rd(); a.width_after = rd(); a.log_qd = rd(); a.num_pv = rd(); a.num_exposed = rd();
            B2_REQUIRE(a.num_exposed == (a.width_after ? 1u : 0u) && a.width_after < (1u << 20), B2S_ERR_BAD_ARG,
                       "key: after-challenge shape");
            if (a.log_qd > log_blowup) fail(B2S_ERR_CONSTRAINT_DEGREE, "key: quotient degree exceeds the engine's blowup");
            if (a.width_prep) {
                for (int j = 0; j < 8; j++) {
                    const uint32_t v = rd();
                    B2_REQUIRE(root || v < P, B2S_ERR_BAD_ARG, "key blob: non-canonical preprocessed commitment");
                    a.prep_commit[j] = root ? v : Fp::from_canonical(v).v;
                }
                a.prep_log_height = rd();
                B2_REQUIRE(a.prep_log_height + log_blowup <= TWO_ADICITY, B2S_ERR_SHAPE, "key: preprocessed height");
                const size_t cnt = (size_t)a.width_prep << a.prep_log_height;
                B2_REQUIRE(i + cnt <= n, B2S_ERR_BAD_ARG, "key blob truncated (preprocessed trace)");
                if (host_only) {
                    i += cnt;
                } else {
                B2_REQUIRE(b2s_matrix_upload(e, w + i, (uint64_t)1 << a.prep_log_height, a.width_prep, B2S_LAYOUT_ROW_MAJOR,
                                             B2S_REPR_CANONICAL, &a.prep_trace) == 0,
                           B2S_ERR_INTERNAL, std::string("key: preprocessed upload failed: ") + e->last_error);
                i += cnt;
                const b2s_matrix* m = a.prep_trace;
                a.prep_pcs = commit_matrices(e, &m, nullptr, 1);
                uint32_t root[8];
                root_monty(e, a.prep_pcs, root);
                B2_REQUIRE(memcmp(root, a.prep_commit, 32) == 0, B2S_ERR_BAD_ARG,
                           "key: preprocessed commitment does not match the committed trace");
                }
            }
            const uint32_t nn = rd();
            B2_REQUIRE((size_t)nn * 4 <= n - i, B2S_ERR_BAD_ARG, "key blob truncated (nodes)");
            std::vector<HostNode> nodes(nn);
            for (auto& nd : nodes) { nd.kind = rd(); nd.a = rd(); nd.b = rd(); nd.c = rd(); }
            const uint32_t ncst = rd();
            B2_REQUIRE(ncst <= n - i, B2S_ERR_BAD_ARG, "key blob truncated (constraints)");
            std::vector<uint32_t> cidx(ncst);
            for (auto& c : cidx) c = rd();
            const uint32_t ni = rd();
            B2_REQUIRE(ni <= n - i, B2S_ERR_BAD_ARG, "key blob truncated (interactions)");
            std::vector<HostInteraction> inter(ni);
            for (auto& it : inter) {
                it.bus = rd(); it.type = rd(); it.count_node = rd();
                const uint32_t nf = rd();
                B2_REQUIRE(nf <= n - i, B2S_ERR_BAD_ARG, "key blob truncated (interaction fields)");
                it.fields.resize(nf);
                for (auto& f : it.fields) f = rd();
                a.inter_bus.push_back(it.bus);
                a.inter_len.push_back(nf);
                a.max_fields = std::max(a.max_fields, nf);
            }
            const uint32_t nch = rd();
            B2_REQUIRE(nch <= n - i, B2S_ERR_BAD_ARG, "key blob truncated (chunks)");
            std::vector<std::vector<uint32_t>> chunks(nch);
            for (auto& chv : chunks) {
                const uint32_t len = rd();
                B2_REQUIRE(len <= n - i, B2S_ERR_BAD_ARG, "key blob truncated (chunk)");
                chv.resize(len);
                for (auto& x : chv) x = rd();
            }
            B2_REQUIRE((ni == 0) == (a.width_after == 0) && (a.width_after == 0 || a.width_after == nch + 1), B2S_ERR_BAD_ARG,
                       "key: interaction / after-challenge shape mismatch");
            validate_nodes(a, nodes);
            a.num_chunks = nch;
            for (auto& nd : nodes) a.nodes.push_back(make_uint4(nd.kind, nd.a, nd.b, nd.c));
            a.constraint_idx = cidx;
            auto up = [&](const std::vector<uint4>& prog) -> uint4* {
                if (prog.empty() || host_only) return nullptr;
                uint4* d = (uint4*)e->dmalloc(prog.size() * sizeof(uint4));
                B2_CUDA(cudaMemcpyAsync(d, prog.data(), prog.size() * sizeof(uint4), cudaMemcpyHostToDevice, e->stream));
                return d;
            };
            Compiled q = compile_dag(a, nodes, cidx, nullptr, nullptr);
            a.program = std::move(q.program);
            a.num_slots = q.slots_f; a.num_slots_e = q.slots_e; a.num_constraints = (uint32_t)cidx.size();
            a.d_program = up(a.program);
            if (ni) {
                Compiled ip = compile_dag(a, nodes, {}, &inter, &chunks);
                a.iprogram = std::move(ip.program);
                a.islots = ip.slots_f;
                a.d_iprogram = up(a.iprogram);
            }
        }
        if (version >= 3) {
            const uint32_t nhc = rd();
            B2_REQUIRE((uint64_t)nhc * (pk->airs.size() + 1) <= n - i, B2S_ERR_BAD_ARG, "key blob truncated (trace-height constraints)");
            pk->height_coefficients.resize(nhc);
            pk->height_thresholds.resize(nhc);
            for (uint32_t c = 0; c < nhc; c++) {
                pk->height_coefficients[c].resize(pk->airs.size());
                for (auto& x : pk->height_coefficients[c]) x = rd();
                pk->height_thresholds[c] = rd();
                B2_REQUIRE(pk->height_thresholds[c] <= P, B2S_ERR_BAD_ARG, "key: trace-height threshold above the field order");
            }
        }
        B2_REQUIRE(i == n, B2S_ERR_BAD_ARG, "key blob: trailing words");
        if (host_only) return pk;
        B2_CUDA(cudaStreamSynchronize(e->stream));
        // quotient kernels: straight-line PTX per AIR, compiled by the driver on a few host threads (qjit.cu)
        if (quotient_jit_enabled()) {
            const auto t_jit = std::chrono::steady_clock::now();
            std::vector<std::pair<size_t, size_t>> work;  // (air, module)
            for (size_t k = 0; k < pk->airs.size(); k++) {
                const size_t nm = quotient_jit_prepare(pk->airs[k]);
                for (size_t m = 0; m < nm; m++) work.push_back({k, m});
            }
            const unsigned nthreads = (unsigned)std::min<size_t>(work.size(), std::max(1u, std::min(16u, std::thread::hardware_concurrency())));
            std::vector<std::string> errors(work.size());
            std::atomic<size_t> next{0};
            std::vector<std::thread> pool;
            const int device = e->device;
            for (unsigned t = 0; t < nthreads; t++)
                pool.emplace_back([&, device] {
                    cudaSetDevice(device);
                    for (size_t w; (w = next.fetch_add(1)) < work.size();) {
                        try {
                            quotient_jit_compile(pk->airs[work[w].first], work[w].second);
                        } catch (const StatusError& se) {
                            errors[w] = se.msg;
                        } catch (const std::exception& ex) {
                            errors[w] = ex.what();
                        }
                    }
                });
            for (auto& th : pool) th.join();
            for (size_t w = 0; w < work.size(); w++)
                if (!errors[w].empty()) fail(B2S_ERR_INTERNAL, "air " + std::to_string(work[w].first) + ": " + errors[w]);
            for (auto& a : pk->airs) { a.jit_ptx.clear(); a.jit_ptx.shrink_to_fit(); }
            if (getenv("B2S_PROVE_TRACE"))
                fprintf(stderr, "[b2s trace] quotient JIT: %zu kernels on %u threads in %.1f ms\n", work.size(), nthreads,
                        std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t_jit).count());
        }
    } catch (...) {
        pk_free(pk);
        throw;
    }
    return pk;
}

// ---- quotient ---------------------------------------------------------------------------------------------------
struct QuotientParams {
    const uint4* program;
    uint32_t prog_len;
    const uint32_t* const* cols;   // LDE column base pointers, [preprocessed][part 0][part 1]...
    const uint32_t* publics;       // Montgomery
    const uint32_t* alpha_pows;    // [num_constraints][4], alpha^j
    const uint32_t* ef_consts;     // LogUp challenges [2][4] then the AIR's exposed cumulative sum [4]
    uint32_t* out;                 // qd chunk matrices, each 4 columns x N rows, column-major, natural row order
    uint32_t n, b, lq;
    uint32_t gq;                   // generator of order 2^(n+lq)
    uint32_t ginv;                 // w_n^-1
    uint32_t zh[16], inv_zh[16];   // Z_H(x) = x^N - 1 takes qd values on the quotient domain: index kq mod qd
};

// Thread r owns LDE row r (r < N*qd: the rows of the quotient domain are a prefix of the bit-reversed LDE), so the
// `local` reads are coalesced; the `next` row differs from r by one XOR mask for whole warps except at carries.
// Base-field values live in slot[], extension-field values (4 words) in eslot[]; both are thread-local arrays
// indexed by the program, i.e. local memory whose accesses are coalesced across the warp (same index in every lane).
B2_D Fp4 ld4(const uint32_t* s) { return Fp4(Fp::raw(s[0]), Fp::raw(s[1]), Fp::raw(s[2]), Fp::raw(s[3])); }
B2_D void st4(uint32_t* s, const Fp4& v) { s[0] = v.c[0].v; s[1] = v.c[1].v; s[2] = v.c[2].v; s[3] = v.c[3].v; }

template <int NSLOT, int NE>
__global__ void __launch_bounds__(128) quotient_kernel(QuotientParams p) {
    const uint32_t r = blockIdx.x * blockDim.x + threadIdx.x;
    const uint32_t lm = p.n + p.b;
    if (r >= (1u << (p.n + p.lq))) return;
    const uint32_t K = lm ? (__brev(r) >> (32 - lm)) : 0u;
    const uint32_t kq = K >> (p.b - p.lq);
    const uint32_t Kn = (K + (1u << p.b)) & ((1u << lm) - 1);
    const uint32_t rn = lm ? (__brev(Kn) >> (32 - lm)) : 0u;
    const uint32_t rows[2] = {r, rn};
    const Fp x = Fp::raw(M_GENERATOR) * fp_pow(Fp::raw(p.gq), kq);
    const uint32_t qi = kq & ((1u << p.lq) - 1);
    const Fp zh = Fp::raw(p.zh[qi]);
    const Fp ginv = Fp::raw(p.ginv);
    bool have_first = false, have_last = false;
    Fp is_first, is_last;
    uint32_t slot[NSLOT];
    uint32_t eslot[NE * 4];
    Fp acc[4];
    for (uint32_t pc = 0; pc < p.prog_len; pc++) {
        const uint4 ins = __ldg(p.program + pc);
        switch (ins.x) {
            case OP_LOAD: slot[ins.y] = __ldg(p.cols[ins.z] + rows[ins.w]); break;
            case OP_PUBLIC: slot[ins.y] = __ldg(p.publics + ins.z); break;
            case OP_CONST: slot[ins.y] = ins.z; break;
            case OP_FIRST:
                if (!have_first) { is_first = zh * fp_inv(x - Fp::one()); have_first = true; }
                slot[ins.y] = is_first.v;
                break;
            case OP_LAST:
                if (!have_last) { is_last = zh * fp_inv(x - ginv); have_last = true; }
                slot[ins.y] = is_last.v;
                break;
            case OP_TRANS: slot[ins.y] = (x - ginv).v; break;
            case OP_ADD: slot[ins.y] = (Fp::raw(slot[ins.z]) + Fp::raw(slot[ins.w])).v; break;
            case OP_SUB: slot[ins.y] = (Fp::raw(slot[ins.z]) - Fp::raw(slot[ins.w])).v; break;
            case OP_NEG: slot[ins.y] = (-Fp::raw(slot[ins.z])).v; break;
            case OP_MUL: slot[ins.y] = (Fp::raw(slot[ins.z]) * Fp::raw(slot[ins.w])).v; break;
            case OP_ACC: {
                const uint4 ap = __ldg(reinterpret_cast<const uint4*>(p.alpha_pows) + ins.w);
                const Fp v = Fp::raw(slot[ins.z]);
                acc[0] += v * Fp::raw(ap.x); acc[1] += v * Fp::raw(ap.y);
                acc[2] += v * Fp::raw(ap.z); acc[3] += v * Fp::raw(ap.w);
                break;
            }
            case OP_LOADP: {
                uint32_t* d = eslot + 4 * ins.y;
#pragma unroll
                for (int e = 0; e < 4; e++) d[e] = __ldg(p.cols[ins.z + e] + rows[ins.w]);
                break;
            }
            case OP_CHAL: st4(eslot + 4 * ins.y, ld4(p.ef_consts + 4 * ins.z)); break;
            case OP_EXPOSED: st4(eslot + 4 * ins.y, ld4(p.ef_consts + 8 + 4 * ins.z)); break;
            case OP_ADD_EE: st4(eslot + 4 * ins.y, ld4(eslot + 4 * ins.z) + ld4(eslot + 4 * ins.w)); break;
            case OP_ADD_EF: st4(eslot + 4 * ins.y, ld4(eslot + 4 * ins.z) + Fp::raw(slot[ins.w])); break;
            case OP_SUB_EE: st4(eslot + 4 * ins.y, ld4(eslot + 4 * ins.z) - ld4(eslot + 4 * ins.w)); break;
            case OP_SUB_EF: st4(eslot + 4 * ins.y, ld4(eslot + 4 * ins.z) - Fp::raw(slot[ins.w])); break;
            case OP_SUB_FE: st4(eslot + 4 * ins.y, -(ld4(eslot + 4 * ins.w) - Fp::raw(slot[ins.z]))); break;
            case OP_NEG_E: st4(eslot + 4 * ins.y, -ld4(eslot + 4 * ins.z)); break;
            case OP_MUL_EE: st4(eslot + 4 * ins.y, ld4(eslot + 4 * ins.z) * ld4(eslot + 4 * ins.w)); break;
            case OP_MUL_EF: st4(eslot + 4 * ins.y, ld4(eslot + 4 * ins.z) * Fp::raw(slot[ins.w])); break;
            default: {  // OP_ACC_E
                const Fp4 t = ld4(eslot + 4 * ins.z) * ld4(p.alpha_pows + 4 * ins.w);
#pragma unroll
                for (int c = 0; c < 4; c++) acc[c] += t.c[c];
            }
        }
    }
    const Fp iz = Fp::raw(p.inv_zh[qi]);
    const uint64_t N = (uint64_t)1 << p.n;
    const uint32_t row = kq >> p.lq;
#pragma unroll
    for (int c = 0; c < 4; c++) p.out[((uint64_t)qi * 4 + c) * N + row] = (acc[c] * iz).v;
}

template <int NSLOT, int NE>
static void launch_quotient_t(b2s_engine* e, const QuotientParams& p) {
    const uint32_t rows = 1u << (p.n + p.lq);
    quotient_kernel<NSLOT, NE><<<(rows + 127) / 128, 128, 0, e->stream>>>(p);
    B2_CHECK_LAUNCH();
    e->launches[3]++;
    e->quotient_launches[1]++;
}
template <int NSLOT>
static void launch_quotient_e(b2s_engine* e, uint32_t eslots, const QuotientParams& p) {
    if (eslots <= 1) launch_quotient_t<NSLOT, 1>(e, p);
    else if (eslots <= 64) launch_quotient_t<NSLOT, 64>(e, p);
    else if (eslots <= 1024) launch_quotient_t<NSLOT, 1024>(e, p);
    else fail(B2S_ERR_CONSTRAINT_DEGREE, "quotient: constraint program needs more than 1024 live extension-field values");
}
static void launch_quotient(b2s_engine* e, uint32_t slots, uint32_t eslots, const QuotientParams& p) {
    if (slots <= 32) launch_quotient_e<32>(e, eslots, p);
    else if (slots <= 128) launch_quotient_e<128>(e, eslots, p);
    else if (slots <= 512) launch_quotient_e<512>(e, eslots, p);
    else if (slots <= 2048) launch_quotient_e<2048>(e, eslots, p);
    else if (slots <= 8192) launch_quotient_e<8192>(e, eslots, p);
    else fail(B2S_ERR_CONSTRAINT_DEGREE, "quotient: constraint program needs more than 8192 live values");
}

// ---- FRI-LogUp after-challenge trace (stark/mod.rs:307-358; formula: DESIGN.md 4) ----------------------------------
struct LogupParams {
    const uint4* program;
    uint32_t prog_len;
    const uint32_t* const* cols;   // TRACE column base pointers (height N), [preprocessed][cached parts][common]
    const uint32_t* publics;
    const uint32_t* betas;         // [max_fields + 1][4], beta^j
    const uint32_t* consts;        // [num_interactions][4]: alpha + beta^len * (bus + 1)
    uint32_t* out;                 // after-challenge matrix, 4*(chunks+1) columns x N rows, column-major
    uint32_t* row_sums;            // 4 planes x N: sum over chunks of this row's reciprocal sums
    uint32_t n;
};
// blockDim = (rows per block, S): S lanes share a trace row (natural order). Every lane evaluates the row's message fields and
// counts; chunk c -- one reciprocal-sum column,
//   perm_c[row] = sum_{i in chunk} (+-count_i) / (alpha + sum_j beta^j field_ij + beta^len_i (bus_i + 1))
// -- is computed by lane c mod S. An extension-field inversion per interaction makes a 321-interaction row a ~1.4 ms chain
// for one thread (r2h: logup_kernel<2048> 3 launches, 4.4 ms); short, wide AIRs get S up to 32, tall ones S = 1.
template <int NSLOT>
__global__ void __launch_bounds__(1024) logup_kernel(LogupParams p) {
    extern __shared__ uint32_t red[];   // [S][4][rows per block], only for S > 1
    const uint32_t S = blockDim.y, sy = threadIdx.y, rx = threadIdx.x, R = blockDim.x;
    const uint32_t r = blockIdx.x * R + rx;
    const uint32_t N = 1u << p.n;
    const bool live = r < N;
    Fp4 total = Fp4::zero();
    if (live) {
        const uint32_t rows[2] = {r, (r + 1) & (N - 1)};
        uint32_t slot[NSLOT];
        Fp4 den = Fp4::zero(), chunk = Fp4::zero();
        uint32_t c = 0;
        bool mine = sy == 0;
        for (uint32_t pc = 0; pc < p.prog_len; pc++) {
            const uint4 ins = __ldg(p.program + pc);
            switch (ins.x) {
                case OP_LOAD: slot[ins.y] = __ldg(p.cols[ins.z] + rows[ins.w]); break;
                case OP_PUBLIC: slot[ins.y] = __ldg(p.publics + ins.z); break;
                case OP_CONST: slot[ins.y] = ins.z; break;
                case OP_FIRST: slot[ins.y] = r == 0 ? R_MOD_P : 0u; break;
                case OP_LAST: slot[ins.y] = r == N - 1 ? R_MOD_P : 0u; break;
                case OP_TRANS: slot[ins.y] = r == N - 1 ? 0u : R_MOD_P; break;
                case OP_ADD: slot[ins.y] = (Fp::raw(slot[ins.z]) + Fp::raw(slot[ins.w])).v; break;
                case OP_SUB: slot[ins.y] = (Fp::raw(slot[ins.z]) - Fp::raw(slot[ins.w])).v; break;
                case OP_NEG: slot[ins.y] = (-Fp::raw(slot[ins.z])).v; break;
                case OP_MUL: slot[ins.y] = (Fp::raw(slot[ins.z]) * Fp::raw(slot[ins.w])).v; break;
                case OP_IFIELD:
                    if (mine) den += ld4(p.betas + 4 * ins.w) * Fp::raw(slot[ins.z]);
                    break;
                default: {  // OP_IEND
                    if (mine) {
                        den += ld4(p.consts + 4 * ins.w);
                        Fp cnt = Fp::raw(slot[ins.z]);
                        if (ins.y & 1) cnt = -cnt;
                        chunk += fp4_inv(den) * cnt;
                        den = Fp4::zero();
                    }
                    if (ins.y & 2) {
                        if (mine) {
#pragma unroll
                            for (int e = 0; e < 4; e++) p.out[((uint64_t)4 * c + e) * N + r] = chunk.c[e].v;
                            total += chunk;
                            chunk = Fp4::zero();
                        }
                        c++;
                        mine = (c & (S - 1)) == sy;
                    }
                }
            }
        }
    }
    if (S > 1) {
#pragma unroll
        for (int e = 0; e < 4; e++) red[(sy * 4 + e) * R + rx] = total.c[e].v;
        __syncthreads();
        for (uint32_t s = S >> 1; s > 0; s >>= 1) {
            if (sy < s) {
#pragma unroll
                for (int e = 0; e < 4; e++)
                    red[(sy * 4 + e) * R + rx] = (Fp::raw(red[(sy * 4 + e) * R + rx]) + Fp::raw(red[((sy + s) * 4 + e) * R + rx])).v;
            }
            __syncthreads();
        }
        if (sy == 0)
#pragma unroll
            for (int e = 0; e < 4; e++) total.c[e] = Fp::raw(red[e * R + rx]);
    }
    if (live && sy == 0) {
#pragma unroll
        for (int e = 0; e < 4; e++) p.row_sums[(uint64_t)e * N + r] = total.c[e].v;
    }
}
static void launch_logup(b2s_engine* e, uint32_t slots, uint32_t num_chunks, const LogupParams& p) {
    const uint32_t N = 1u << p.n;
    // lanes per row: a power of two, at most 32, at most the number of chunks, ~2^15 threads per launch for short traces
    uint32_t S = 1;
    while (S < 32 && 2 * S <= num_chunks && ((uint64_t)N * 2 * S) <= (1u << 15)) S *= 2;
    const dim3 block(S == 1 ? 128 : 32, S);
    const dim3 grid((N + block.x - 1) / block.x);
    const size_t smem = S > 1 ? (size_t)S * 4 * block.x * 4 : 0;
    if (slots <= 32) logup_kernel<32><<<grid, block, smem, e->stream>>>(p);
    else if (slots <= 256) logup_kernel<256><<<grid, block, smem, e->stream>>>(p);
    else if (slots <= 2048) logup_kernel<2048><<<grid, block, smem, e->stream>>>(p);
    else fail(B2S_ERR_CONSTRAINT_DEGREE, "logup: interaction program needs more than 2048 live values");
    B2_CHECK_LAUNCH();
    e->launches[3]++;
}

// inclusive prefix sums (mod p) along each of the 4 planes: blocks of 1024 elements, block totals scanned recursively
constexpr int SCAN_THREADS = 256, SCAN_ITEMS = 4, SCAN_BLOCK = SCAN_THREADS * SCAN_ITEMS;
__global__ void __launch_bounds__(SCAN_THREADS) scan_block_kernel(const uint32_t* __restrict__ in, uint32_t* __restrict__ out,
                                                                  uint32_t* __restrict__ totals, uint32_t n, uint32_t nblocks) {
    __shared__ uint32_t sm[SCAN_THREADS];
    const uint32_t plane = blockIdx.y;
    const uint32_t base = blockIdx.x * SCAN_BLOCK + threadIdx.x * SCAN_ITEMS;
    Fp v[SCAN_ITEMS];
    Fp run = Fp::zero();
#pragma unroll
    for (int k = 0; k < SCAN_ITEMS; k++) {
        if (base + k < n) run += Fp::raw(in[(uint64_t)plane * n + base + k]);
        v[k] = run;
    }
    sm[threadIdx.x] = run.v;
    __syncthreads();
    for (int off = 1; off < SCAN_THREADS; off <<= 1) {
        Fp t = Fp::raw(sm[threadIdx.x]);
        if ((int)threadIdx.x >= off) t += Fp::raw(sm[threadIdx.x - off]);
        __syncthreads();
        sm[threadIdx.x] = t.v;
        __syncthreads();
    }
    const Fp prefix = threadIdx.x ? Fp::raw(sm[threadIdx.x - 1]) : Fp::zero();
#pragma unroll
    for (int k = 0; k < SCAN_ITEMS; k++)
        if (base + k < n) out[(uint64_t)plane * n + base + k] = (v[k] + prefix).v;
    if (totals && threadIdx.x == SCAN_THREADS - 1) totals[(uint64_t)plane * nblocks + blockIdx.x] = sm[SCAN_THREADS - 1];
}
__global__ void scan_add_kernel(uint32_t* __restrict__ data, const uint32_t* __restrict__ block_prefix, uint32_t n, uint32_t nblocks) {
    const uint32_t plane = blockIdx.y;
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const uint32_t blk = i / SCAN_BLOCK;
    if (!blk) return;
    uint32_t* d = data + (uint64_t)plane * n + i;
    *d = (Fp::raw(*d) + Fp::raw(block_prefix[(uint64_t)plane * nblocks + blk - 1])).v;
}
struct Scratch;
static void scan_planes(b2s_engine* e, Scratch& scratch, const uint32_t* in, uint32_t* out, uint32_t n);

// ---- out-of-domain evaluation (barycentric over LDE block 0) ---------------------------------------------------------
// Block 0 of an LDE of a height-N trace holds p(x_i), x_i = g * w_N^{bitrev_n(i)}, i < N. For z off the coset:
//   p(z) = ((z/g)^N - 1) / N * sum_i p(x_i) * x_i / (z - x_i)
__global__ void bary_weights_kernel(uint32_t* __restrict__ wts, uint32_t n, Fp4 z, uint32_t wn) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    const uint32_t N = 1u << n;
    if (i >= N) return;
    const uint32_t e = n ? (__brev(i) >> (32 - n)) : 0u;
    const Fp x = Fp::raw(M_GENERATOR) * fp_pow(Fp::raw(wn), e);
    Fp4 d = z - x;
    Fp4 w = fp4_inv(d) * x;
#pragma unroll
    for (int c = 0; c < 4; c++) wts[(uint64_t)c * N + i] = w.c[c].v;
}

constexpr int DOT_THREADS = 256;
// All columns of all matrices of one trace height at one point in one launch (a segment opens ~200 (matrix, point) pairs;
// one launch pair and one read-back each was 400 launches and 200 host waits): column blockIdx.x of the group is described by
// BaryCol {column base (block 0 of its LDE), where its partial sums go}.
struct BaryCol {
    const uint32_t* col;
    uint32_t partial_off;   // word offset of this column's [splits][4] partial sums
    uint32_t splits;
    uint32_t scale_idx;     // which ((z/g)^N - 1) / N factor applies
    uint32_t pad;
};
__global__ void __launch_bounds__(DOT_THREADS) bary_dot_multi_kernel(const BaryCol* __restrict__ cols, const uint32_t* __restrict__ wts,
                                                                     uint32_t n, uint32_t* __restrict__ partial) {
    __shared__ uint32_t sm[4][DOT_THREADS];
    const uint32_t N = 1u << n;
    const BaryCol bc = cols[blockIdx.x];
    const uint32_t split = blockIdx.y, splits = gridDim.y;
    const uint32_t per = (N + splits - 1) / splits;
    const uint32_t lo = split * per, hi = min(N, lo + per);
    const uint32_t* __restrict__ colp = bc.col;
    Fp acc[4];
    for (uint32_t i = lo + threadIdx.x; i < hi; i += DOT_THREADS) {
        const Fp v = Fp::raw(colp[i]);
#pragma unroll
        for (int c = 0; c < 4; c++) acc[c] += v * Fp::raw(wts[(uint64_t)c * N + i]);
    }
#pragma unroll
    for (int c = 0; c < 4; c++) sm[c][threadIdx.x] = acc[c].v;
    __syncthreads();
    for (int s = DOT_THREADS / 2; s > 0; s >>= 1) {
        if (threadIdx.x < s) {
#pragma unroll
            for (int c = 0; c < 4; c++) sm[c][threadIdx.x] = (Fp::raw(sm[c][threadIdx.x]) + Fp::raw(sm[c][threadIdx.x + s])).v;
        }
        __syncthreads();
    }
    if (threadIdx.x < 4) partial[bc.partial_off + split * 4 + threadIdx.x] = sm[threadIdx.x][0];
}
// every opened value of the proof: sums a column's partial dot products and applies its barycentric factor
__global__ void bary_finish_multi_kernel(const BaryCol* __restrict__ cols, uint32_t ncols, const uint32_t* __restrict__ partial,
                                         const uint32_t* __restrict__ scales, uint32_t* __restrict__ vals) {
    const uint32_t c = blockIdx.x * blockDim.x + threadIdx.x;
    if (c >= ncols) return;
    const BaryCol bc = cols[c];
    Fp4 s = Fp4::zero();
    for (uint32_t sp = 0; sp < bc.splits; sp++) {
        const uint4 q = *reinterpret_cast<const uint4*>(partial + bc.partial_off + sp * 4);
        s.c[0] += Fp::raw(q.x); s.c[1] += Fp::raw(q.y); s.c[2] += Fp::raw(q.z); s.c[3] += Fp::raw(q.w);
    }
    s = s * ld4(scales + 4 * bc.scale_idx);
    *reinterpret_cast<uint4*>(vals + (uint64_t)c * 4) = make_uint4(s.c[0].v, s.c[1].v, s.c[2].v, s.c[3].v);
}

// ---- reduced openings ------------------------------------------------------------------------------------------------
// All matrices of one LDE height in one launch. Every matrix of height 2^h is opened at zeta and, for trace matrices, at
// zeta * g_N (quotient chunks: zeta only), so the two inverses 1 / (z_j - x) -- an exponentiation for x and two extension
// inversions, ~250 products, more than the column sums of a narrow matrix -- are paid once per row instead of once per
// row and matrix, and ro is written once instead of read-modify-written per matrix (the segment proof had 40 launches).
// The kernel is bound by its products (leaf-aggregation shape: ~27 G for the per-row part, ~24 G for the column sums, ~10 G
// per (matrix, row)): batching the loads or moving the column sums to FP64 dot products did not change its 20 ms.
// Descriptor of matrix m (RO_DESC_WORDS words): pointer lo/hi, width, npoints, cz[2][4], coef[2][4].
constexpr uint32_t RO_DESC_WORDS = 20;
__global__ void __launch_bounds__(128) reduced_open_height_kernel(const uint32_t* __restrict__ descs, uint32_t nmats, uint32_t h,
                                                                  const uint32_t* __restrict__ alpha_pows,
                                                                  uint32_t* __restrict__ ro, Fp4 z0, Fp4 z1, uint32_t two_points,
                                                                  uint32_t wh) {
    const uint32_t row = blockIdx.x * blockDim.x + threadIdx.x;
    const uint32_t H = 1u << h;
    if (row >= H) return;
    const uint32_t e = h ? (__brev(row) >> (32 - h)) : 0u;
    const Fp x = Fp::raw(M_GENERATOR) * fp_pow(Fp::raw(wh), e);
    Fp4 inv0, inv1 = Fp4::zero();
    if (two_points) fp4_inv_pair(z0 - x, z1 - x, inv0, inv1);
    else inv0 = fp4_inv(z0 - x);
    // per point: sum over the matrices of alpha^o (cz - inner); the division by (z_j - x) is applied once at the end
    Fp4 acc0 = Fp4::zero(), acc1 = Fp4::zero();
#pragma unroll 1
    for (uint32_t m = 0; m < nmats; m++) {
        const uint32_t* __restrict__ d = descs + (uint64_t)m * RO_DESC_WORDS;
        const uint32_t* __restrict__ mat = reinterpret_cast<const uint32_t*>(((uint64_t)__ldg(d + 1) << 32) | __ldg(d));
        const uint32_t width = __ldg(d + 2), npoints = __ldg(d + 3);
        Fp4 inner = Fp4::zero();
        for (uint32_t t = 0; t < width; t++) {
            const Fp v = Fp::raw(__ldg(mat + (uint64_t)t * H + row));
            const uint4 ap = __ldg(reinterpret_cast<const uint4*>(alpha_pows) + t);
            inner.c[0] += v * Fp::raw(ap.x); inner.c[1] += v * Fp::raw(ap.y);
            inner.c[2] += v * Fp::raw(ap.z); inner.c[3] += v * Fp::raw(ap.w);
        }
        const Fp4 cz0(Fp::raw(__ldg(d + 4)), Fp::raw(__ldg(d + 5)), Fp::raw(__ldg(d + 6)), Fp::raw(__ldg(d + 7)));
        const Fp4 cf0(Fp::raw(__ldg(d + 12)), Fp::raw(__ldg(d + 13)), Fp::raw(__ldg(d + 14)), Fp::raw(__ldg(d + 15)));
        acc0 += cf0 * (cz0 - inner);
        if (npoints > 1) {
            const Fp4 cz1(Fp::raw(__ldg(d + 8)), Fp::raw(__ldg(d + 9)), Fp::raw(__ldg(d + 10)), Fp::raw(__ldg(d + 11)));
            const Fp4 cf1(Fp::raw(__ldg(d + 16)), Fp::raw(__ldg(d + 17)), Fp::raw(__ldg(d + 18)), Fp::raw(__ldg(d + 19)));
            acc1 += cf1 * (cz1 - inner);
        }
    }
    Fp4 acc = acc0 * inv0;
    if (two_points) acc += acc1 * inv1;
#pragma unroll
    for (int c = 0; c < 4; c++) ro[(uint64_t)c * H + row] = acc.c[c].v;
}

// The same for short LDE heights (< 2^17 rows), where one thread per row leaves the chip empty and a 3164-column row is a
// 140 us dependent chain: S threads share a row, thread (rx, sy) takes columns sy, sy + S, ... of every matrix (loads stay
// coalesced across the 32 rows of a warp), the S partial sums meet in shared memory and lane sy = 0 applies the inverses.
// Sum over matrices of coef * (cz - inner) is linear in inner, so partial inners give partial sums; cz is added by sy = 0.
constexpr int RO_ROWS = 32;
__global__ void __launch_bounds__(1024) reduced_open_split_kernel(const uint32_t* __restrict__ descs, uint32_t nmats, uint32_t h,
                                                                  const uint32_t* __restrict__ alpha_pows, uint32_t* __restrict__ ro,
                                                                  Fp4 z0, Fp4 z1, uint32_t two_points, uint32_t wh) {
    extern __shared__ uint32_t red[];   // [S][8][RO_ROWS]
    const uint32_t S = blockDim.y, rx = threadIdx.x, sy = threadIdx.y;
    const uint32_t H = 1u << h;
    const uint32_t row = blockIdx.x * RO_ROWS + rx;
    const bool live = row < H;
    Fp4 acc0 = Fp4::zero(), acc1 = Fp4::zero();
    if (live) {
#pragma unroll 1
        for (uint32_t m = 0; m < nmats; m++) {
            const uint32_t* __restrict__ d = descs + (uint64_t)m * RO_DESC_WORDS;
            const uint32_t* __restrict__ mat = reinterpret_cast<const uint32_t*>(((uint64_t)__ldg(d + 1) << 32) | __ldg(d));
            const uint32_t width = __ldg(d + 2), npoints = __ldg(d + 3);
            if (sy >= width && sy != 0) continue;
            Fp4 inner = Fp4::zero();
            for (uint32_t t = sy; t < width; t += S) {
                const Fp v = Fp::raw(__ldg(mat + (uint64_t)t * H + row));
                const uint4 ap = __ldg(reinterpret_cast<const uint4*>(alpha_pows) + t);
                inner.c[0] += v * Fp::raw(ap.x); inner.c[1] += v * Fp::raw(ap.y);
                inner.c[2] += v * Fp::raw(ap.z); inner.c[3] += v * Fp::raw(ap.w);
            }
            const Fp4 cf0(Fp::raw(__ldg(d + 12)), Fp::raw(__ldg(d + 13)), Fp::raw(__ldg(d + 14)), Fp::raw(__ldg(d + 15)));
            Fp4 term0 = -inner;
            if (sy == 0) term0 += Fp4(Fp::raw(__ldg(d + 4)), Fp::raw(__ldg(d + 5)), Fp::raw(__ldg(d + 6)), Fp::raw(__ldg(d + 7)));
            acc0 += cf0 * term0;
            if (npoints > 1) {
                const Fp4 cf1(Fp::raw(__ldg(d + 16)), Fp::raw(__ldg(d + 17)), Fp::raw(__ldg(d + 18)), Fp::raw(__ldg(d + 19)));
                Fp4 term1 = -inner;
                if (sy == 0) term1 += Fp4(Fp::raw(__ldg(d + 8)), Fp::raw(__ldg(d + 9)), Fp::raw(__ldg(d + 10)), Fp::raw(__ldg(d + 11)));
                acc1 += cf1 * term1;
            }
        }
    }
#pragma unroll
    for (int c = 0; c < 4; c++) {
        red[(sy * 8 + c) * RO_ROWS + rx] = acc0.c[c].v;
        red[(sy * 8 + 4 + c) * RO_ROWS + rx] = acc1.c[c].v;
    }
    __syncthreads();
    for (uint32_t s = S >> 1; s > 0; s >>= 1) {
        if (sy < s) {
#pragma unroll
            for (int c = 0; c < 8; c++)
                red[(sy * 8 + c) * RO_ROWS + rx] = (Fp::raw(red[(sy * 8 + c) * RO_ROWS + rx]) + Fp::raw(red[((sy + s) * 8 + c) * RO_ROWS + rx])).v;
        }
        __syncthreads();
    }
    if (sy != 0 || !live) return;
#pragma unroll
    for (int c = 0; c < 4; c++) { acc0.c[c] = Fp::raw(red[c * RO_ROWS + rx]); acc1.c[c] = Fp::raw(red[(4 + c) * RO_ROWS + rx]); }
    const uint32_t e = h ? (__brev(row) >> (32 - h)) : 0u;
    const Fp x = Fp::raw(M_GENERATOR) * fp_pow(Fp::raw(wh), e);
    Fp4 inv0, inv1 = Fp4::zero();
    if (two_points) fp4_inv_pair(z0 - x, z1 - x, inv0, inv1);
    else inv0 = fp4_inv(z0 - x);
    Fp4 acc = acc0 * inv0;
    if (two_points) acc += acc1 * inv1;
#pragma unroll
    for (int c = 0; c < 4; c++) ro[(uint64_t)c * H + row] = acc.c[c].v;
}

// ---- FRI commit phase ------------------------------------------------------------------------------------------------
// planes[4][2R] (bit-reversed evaluations) -> leaf matrix 8 columns x R rows: column 4k+c of row p = coefficient c of
// element 2p+k  (fri/mod.rs:93-116: a commit-phase leaf is the pair (e0, e1) flattened)
__global__ void fri_pack_kernel(const uint32_t* __restrict__ planes, uint32_t* __restrict__ leaf, uint32_t R) {
    const uint32_t p = blockIdx.x * blockDim.x + threadIdx.x;
    if (p >= R) return;
#pragma unroll
    for (int c = 0; c < 4; c++) {
        const uint2 pr = *reinterpret_cast<const uint2*>(planes + (uint64_t)c * 2 * R + 2 * p);
        leaf[(uint64_t)c * R + p] = pr.x;
        leaf[(uint64_t)(4 + c) * R + p] = pr.y;
    }
}
// next[p] = e0 + (beta - x0)(e1 - e0)/(x1 - x0) + beta^2 * roll[p],  x0 = w_lh^{bitrev_lh(2p)}, x1 = -x0
__global__ void fri_fold_kernel(const uint32_t* __restrict__ cur, uint32_t* __restrict__ next,
                                const uint32_t* __restrict__ roll, uint32_t lh, uint32_t w, Fp4 beta, Fp4 beta2) {
    const uint32_t R = 1u << (lh - 1);
    const uint32_t p = blockIdx.x * blockDim.x + threadIdx.x;
    if (p >= R) return;
    Fp4 e0, e1;
#pragma unroll
    for (int c = 0; c < 4; c++) {
        const uint2 pr = *reinterpret_cast<const uint2*>(cur + (uint64_t)c * 2 * R + 2 * p);
        e0.c[c] = Fp::raw(pr.x);
        e1.c[c] = Fp::raw(pr.y);
    }
    const uint32_t ex = lh > 1 ? (__brev(p) >> (32 - (lh - 1))) : 0u;  // bitrev_lh(2p) = bitrev_{lh-1}(p)
    const Fp x0 = fp_pow(Fp::raw(w), ex);
    const Fp inv = fp_inv(-(x0 + x0));
    Fp4 f = e0 + (beta - x0) * ((e1 - e0) * inv);
    if (roll) {
        Fp4 r;
#pragma unroll
        for (int c = 0; c < 4; c++) r.c[c] = Fp::raw(roll[(uint64_t)c * R + p]);
        f += beta2 * r;
    }
#pragma unroll
    for (int c = 0; c < 4; c++) next[(uint64_t)c * R + p] = f.c[c].v;
}

// ---- proof of work ------------------------------------------------------------------------------------------------------
struct PowParams { uint32_t st[16]; uint32_t in_ptr, bits, base; };
__global__ void __launch_bounds__(128) pow_grind_kernel(PowParams p, uint32_t count, uint32_t* result) {
    const uint32_t t = blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= count) return;
    const uint32_t w = p.base + t;
    if (w >= P) return;
    Fp s[16];
#pragma unroll
    for (int i = 0; i < 16; i++) s[i] = Fp::raw(p.st[i]);
    const Fp wm = Fp::from_canonical(w);
#pragma unroll
    for (int i = 0; i < 8; i++) if ((uint32_t)i == p.in_ptr) s[i] = wm;
    poseidon2_permute(s);
    const uint32_t v = s[7].to_canonical();
    if ((v & ((1u << p.bits) - 1)) == 0) atomicMin(result, w);
}

// ---- host-side helpers ----------------------------------------------------------------------------------------------------
struct Scratch {  // device buffers released when the proof is done (or abandoned)
    b2s_engine* e;
    struct Buf { void* p; size_t bytes; cudaStream_t stream; };
    std::vector<Buf> bufs;   // with the stream each was allocated on: it goes back to that stream's cache (side-stream scans)
    std::vector<b2s_pcs_data*> pcs;
    std::deque<std::vector<uint32_t>> host_words;        // sources of asynchronous uploads, alive until the proof is done
    std::deque<std::vector<const uint32_t*>> host_ptrs;
    explicit Scratch(b2s_engine* e_) : e(e_) {}
    // upload without a stream round trip: the host copy moves into the scratch and outlives the copy
    uint32_t* upload(std::vector<uint32_t>&& h) {
        host_words.push_back(std::move(h));
        const std::vector<uint32_t>& v = host_words.back();
        uint32_t* d = words(std::max<size_t>(v.size(), 1));
        if (!v.empty()) B2_CUDA(cudaMemcpyAsync(d, v.data(), v.size() * 4, cudaMemcpyHostToDevice, e->stream));
        return d;
    }
    const uint32_t** upload(std::vector<const uint32_t*>&& h) {
        host_ptrs.push_back(std::move(h));
        const std::vector<const uint32_t*>& v = host_ptrs.back();
        const uint32_t** d = (const uint32_t**)bytes(std::max<size_t>(v.size(), 1) * sizeof(void*));
        if (!v.empty()) B2_CUDA(cudaMemcpyAsync(d, v.data(), v.size() * sizeof(void*), cudaMemcpyHostToDevice, e->stream));
        return d;
    }
    uint32_t* words(size_t n) {
        void* p = e->dmalloc(n * 4);
        bufs.push_back({p, n * 4, e->stream});
        return (uint32_t*)p;
    }
    void* bytes(size_t n) {
        void* p = e->dmalloc(n);
        bufs.push_back({p, n, e->stream});
        return p;
    }
    ~Scratch() {
        if (std::uncaught_exceptions() > 0) {   // a failed proof: nothing may still be running on any stream of the engine
            for (auto st : e->aux_streams) cudaStreamSynchronize(st);
            if (e->hash_stream) cudaStreamSynchronize(e->hash_stream);
            cudaStreamSynchronize(e->stream);
        }
        for (auto* d : pcs) free_pcs(d);
        // by now every side stream has joined the engine's stream and the proof has synchronised (or failed and synchronised
        // below): a buffer returns to the cache of the stream that allocated it, where the next proof asks for it again
        cudaStream_t main_stream = e->stream;
        for (auto& b : bufs) {
            e->stream = b.stream;
            e->dfree(b.p, b.bytes);
        }
        e->stream = main_stream;
    }
};

static void scan_planes(b2s_engine* e, Scratch& scratch, const uint32_t* in, uint32_t* out, uint32_t n) {
    const uint32_t nblocks = (n + SCAN_BLOCK - 1) / SCAN_BLOCK;
    uint32_t* totals = nblocks > 1 ? scratch.words((size_t)4 * nblocks) : nullptr;
    scan_block_kernel<<<dim3(nblocks, 4), SCAN_THREADS, 0, e->stream>>>(in, out, totals, n, nblocks);
    B2_CHECK_LAUNCH();
    e->launches[3]++;
    if (nblocks > 1) {
        uint32_t* prefix = scratch.words((size_t)4 * nblocks);
        scan_planes(e, scratch, totals, prefix, nblocks);
        scan_add_kernel<<<dim3((n + 255) / 256, 4), 256, 0, e->stream>>>(out, prefix, n, nblocks);
        B2_CHECK_LAUNCH();
        e->launches[3]++;
    }
}

struct ByteWriter {
    std::vector<uint8_t>& buf;
    void u32(uint32_t x) {  // little-endian host (x86-64 / aarch64): the wire format is the in-memory word
        const uint8_t* b = reinterpret_cast<const uint8_t*>(&x);
        buf.insert(buf.end(), b, b + 4);
    }
    void f(Fp x) { u32(x.to_canonical()); }
    void ef(const Fp4& e) { for (int i = 0; i < 4; i++) f(e.c[i]); }
    void efs(const std::vector<Fp4>& v) {   // one growth per vector: the opened values are ~10^5 elements of a segment proof
        const size_t at = buf.size();
        buf.resize(at + 4 + 16 * v.size());
        uint32_t tmp = (uint32_t)v.size();
        memcpy(buf.data() + at, &tmp, 4);
        size_t pos = at + 4;
        for (auto& e : v)
            for (int i = 0; i < 4; i++) {
                tmp = e.c[i].to_canonical();
                memcpy(buf.data() + pos, &tmp, 4);
                pos += 4;
            }
    }
    bool raw_digests = false;  // root configuration: a digest is already the canonical limbs of its scalar
    void digest_monty(const uint32_t* d) {
        if (raw_digests) return words(d, 8);
        for (int i = 0; i < 8; i++) f(Fp::raw(d[i]));
    }
    void words(const uint32_t* w, size_t n) {  // already canonical
        const uint8_t* b = reinterpret_cast<const uint8_t*>(w);
        if (n) buf.insert(buf.end(), b, b + 4 * n);
    }
};

struct RoundMat {
    const b2s_matrix* lde;
    unsigned log_n;                          // log2 of the trace height
    std::vector<Fp4> points;
    std::vector<std::vector<Fp4>> values;    // [point][column]
};
struct Round {
    const b2s_pcs_data* pcs;
    std::vector<RoundMat> mats;              // caller order
};

static Fp4 fp4_from(Fp a) { return Fp4(a); }
static Fp4 fp4_pow2k(Fp4 x, unsigned k) { for (unsigned i = 0; i < k; i++) x = x * x; return x; }

static void upload_ef_powers(b2s_engine* e, uint32_t* d, const Fp4& base, size_t count) {
    std::vector<uint32_t> h(count * 4);
    Fp4 p = Fp4::one();
    for (size_t i = 0; i < count; i++) {
        for (int c = 0; c < 4; c++) h[4 * i + c] = p.c[c].v;
        p = p * base;
    }
    B2_CUDA(cudaMemcpyAsync(d, h.data(), h.size() * 4, cudaMemcpyHostToDevice, e->stream));
    B2_CUDA(cudaStreamSynchronize(e->stream));  // h is a stack-lifetime buffer
}

// minimum proof-of-work witness for the transcript state `ch` (duplex.rs:122-130); advances the transcript
static Fp grind_device(b2s_engine* e, Scratch& scratch, HostChallenger& ch, unsigned bits, uint32_t override_witness) {
    B2_REQUIRE(bits < 31, B2S_ERR_BAD_ARG, "prove: proof-of-work bits too large");
    if (override_witness != 0xFFFFFFFFu) {
        const Fp w = Fp::from_canonical(override_witness);
        B2_REQUIRE(ch.check_witness(bits, w), B2S_ERR_BAD_ARG, "prove: the supplied proof-of-work witness does not satisfy the transcript");
        return w;
    }
    uint32_t* d_res = scratch.words(1);
    uint32_t bn_state[24];
    for (int k = 0; k < 3; k++)
        for (int j = 0; j < 8; j++) bn_state[8 * k + j] = ch.bst[k].l[j];
    PowParams pp{};
    for (int i = 0; i < 16; i++) pp.st[i] = ch.st[i].v;
    pp.in_ptr = (uint32_t)ch.in_ptr;
    pp.bits = bits;
    uint32_t found = 0xFFFFFFFFu;
    // first batch: 4 x the expected witness (found with probability 1 - e^-4); later batches double
    uint32_t batch = std::min(1u << 22, std::max(1u << 14, 4u << bits));
    for (uint64_t base = 0; base < P && found == 0xFFFFFFFFu; base += batch, batch = std::min(1u << 22, batch * 2)) {
        pp.base = (uint32_t)base;
        B2_CUDA(cudaMemsetAsync(d_res, 0xFF, 4, e->stream));
        if (ch.root) bn_pow_grind(e, bn_state, ch.bin, (uint32_t)ch.in_ptr, bits, (uint32_t)base, batch, d_res);
        else pow_grind_kernel<<<batch / 128, 128, 0, e->stream>>>(pp, batch, d_res);
        B2_CHECK_LAUNCH();
        e->launches[3]++;
        e->permutations += batch;
        B2_CUDA(cudaMemcpyAsync(&found, d_res, 4, cudaMemcpyDeviceToHost, e->stream));
        B2_CUDA(cudaStreamSynchronize(e->stream));
    }
    B2_REQUIRE(found != 0xFFFFFFFFu, B2S_ERR_INTERNAL, "prove: no proof-of-work witness exists");
    const Fp witness = Fp::from_canonical(found);
    B2_REQUIRE(ch.check_witness(bits, witness), B2S_ERR_INTERNAL, "prove: device proof-of-work witness rejected by the host transcript");
    return witness;
}

// B2S_PROVE_TRACE=1: wall-clock marks of the host thread through one proof (stderr); the stream is drained at each mark
struct HostTrace {
    b2s_engine* e;
    bool on;
    int level;   // B2S_PROVE_TRACE=2 adds marks inside the FRI commit phase
    std::chrono::steady_clock::time_point t0, last;
    explicit HostTrace(b2s_engine* e_) : e(e_), on(getenv("B2S_PROVE_TRACE") != nullptr), level(on ? atoi(getenv("B2S_PROVE_TRACE")) : 0) {
        t0 = last = std::chrono::steady_clock::now();
    }
    void mark(const char* what, int min_level = 1) {
        if (!on || level < min_level) return;
        auto a = std::chrono::steady_clock::now();
        cudaStreamSynchronize(e->stream);
        auto b = std::chrono::steady_clock::now();
        fprintf(stderr, "[b2s trace] %-28s host %8.3f ms  drain %8.3f ms  (t = %8.3f ms)\n", what,
                std::chrono::duration<double, std::milli>(a - last).count(), std::chrono::duration<double, std::milli>(b - a).count(),
                std::chrono::duration<double, std::milli>(b - t0).count());
        last = b;
    }
};

// ---- prove -------------------------------------------------------------------------------------------------------------------
void prove(b2s_engine* e, const b2s_pk* pk, const b2s_air_ctx* airs, size_t n_airs, std::vector<uint8_t>& out) {
    B2_REQUIRE(pk->eng == e, B2S_ERR_BAD_ARG, "prove: key was loaded on a different engine");
    StageScope total_scope(e, ST_PROVE);
    Scratch scratch(e);
    HostTrace trace(e);
    const unsigned b = e->params.log_blowup;
    B2_REQUIRE(pk->hash_kind == e->hash_kind, B2S_ERR_BAD_ARG, "prove: the key was loaded under a different hash configuration");
    HostChallenger ch(e->hash_kind);
    std::vector<unsigned> logn(n_airs);
    std::vector<const AirKeyDev*> keys(n_airs);
    unsigned log_max_height = 0;
    for (size_t i = 0; i < n_airs; i++) {
        const b2s_air_ctx& a = airs[i];
        B2_REQUIRE(a.air_id < pk->airs.size(), B2S_ERR_BAD_ARG, "prove: air_id out of range");
        B2_REQUIRE(i == 0 || a.air_id > airs[i - 1].air_id, B2S_ERR_BAD_ARG, "prove: air ids must be strictly increasing");
        const AirKeyDev& k = pk->airs[a.air_id];
        keys[i] = &k;
        B2_REQUIRE(a.n_public_values == k.num_pv && (k.num_pv == 0 || a.public_values), B2S_ERR_SHAPE, "prove: public value count");
        B2_REQUIRE(a.n_cached == k.cached_widths.size(), B2S_ERR_SHAPE, "prove: cached main count");
        B2_REQUIRE((k.width_common != 0) == (a.common_main != nullptr), B2S_ERR_SHAPE, "prove: common main presence");
        uint64_t h = 0;
        if (a.common_main) {
            B2_REQUIRE(a.common_main->width == k.width_common, B2S_ERR_SHAPE, "prove: common main width");
            h = a.common_main->height;
        }
        for (uint32_t p = 0; p < a.n_cached; p++) {
            const b2s_pcs_data* c = a.cached[p];
            B2_REQUIRE(c && c->mats.size() == 1 && c->mats[0]->width == k.cached_widths[p], B2S_ERR_SHAPE, "prove: cached main shape");
            uint64_t hc = c->mats[0]->height >> b;
            B2_REQUIRE(h == 0 || h == hc, B2S_ERR_SHAPE, "prove: partition heights differ");
            h = hc;
        }
        B2_REQUIRE(h && is_pow2(h), B2S_ERR_SHAPE, "prove: trace height must be a power of two");
        if (k.width_after)
            for (uint32_t p = 0; p < a.n_cached; p++) {
                const b2s_matrix* t = a.cached_traces ? a.cached_traces[p] : nullptr;
                B2_REQUIRE(t && t->width == k.cached_widths[p] && t->height == h, B2S_ERR_SHAPE,
                           "prove: an AIR with interactions needs its cached main traces (cached_traces)");
            }
        logn[i] = log2_exact(h);
        B2_REQUIRE(logn[i] + b <= TWO_ADICITY, B2S_ERR_SHAPE, "prove: trace too tall for this blowup");
        B2_REQUIRE(!k.width_prep || k.prep_log_height == logn[i], B2S_ERR_SHAPE, "prove: preprocessed trace height mismatch");
        log_max_height = std::max(log_max_height, logn[i]);
    }
    {   // crates/vm/src/arch/vm.rs:627-652: a proving context whose heights break a linear constraint of the key is refused
        std::vector<uint32_t> ids(n_airs);
        for (size_t i = 0; i < n_airs; i++) ids[i] = airs[i].air_id;
        const int bad = pk->violated_height_constraint(ids, logn, n_airs);
        B2_REQUIRE(bad < 0, B2S_ERR_SHAPE, "prove: trace heights violate linear constraint " + std::to_string(bad) + " of the key");
    }
    // transcript prefix (stark/mod.rs:140-149, 231-238, 247-254)
    ch.observe_digest(pk->pre_hash);
    ch.observe_canonical((uint32_t)n_airs);
    for (size_t i = 0; i < n_airs; i++) ch.observe_canonical(airs[i].air_id);
    for (size_t i = 0; i < n_airs; i++)
        for (uint32_t j = 0; j < airs[i].n_public_values; j++) {
            B2_REQUIRE(airs[i].public_values[j] < P, B2S_ERR_BAD_ARG, "prove: non-canonical public value");
            ch.observe_canonical(airs[i].public_values[j]);
        }
    for (size_t i = 0; i < n_airs; i++) if (keys[i]->width_prep) ch.observe_digest(keys[i]->prep_commit);
    // main commitments: cached mains in (AIR, partition) order, then one tree over all common mains
    std::vector<std::array<uint32_t, 8>> main_commits;
    for (size_t i = 0; i < n_airs; i++)
        for (uint32_t p = 0; p < airs[i].n_cached; p++) {
            std::array<uint32_t, 8> r;
            root_monty(e, airs[i].cached[p], r.data());
            main_commits.push_back(r);
        }
    std::vector<const b2s_matrix*> commons;
    std::vector<size_t> common_of;
    for (size_t i = 0; i < n_airs; i++) if (airs[i].common_main) { commons.push_back(airs[i].common_main); common_of.push_back(i); }
    B2_REQUIRE(!commons.empty(), B2S_ERR_SHAPE, "prove: no common main trace");
    b2s_pcs_data* common_pcs;
    {
        StageScope ms(e, ST_MAIN_COMMIT);
        common_pcs = commit_matrices(e, commons.data(), nullptr, commons.size());
        scratch.pcs.push_back(common_pcs);
        std::array<uint32_t, 8> r;
        root_monty(e, common_pcs, r.data());
        main_commits.push_back(r);
    }
    for (auto& c : main_commits) ch.observe_digest(c.data());
    for (size_t i = 0; i < n_airs; i++) ch.observe_canonical(logn[i]);

    trace.mark("main commit");
    // ---- FRI-LogUp challenge phase (stark/mod.rs:307-358): PoW, two challenges, one after-challenge matrix per
    // interacting AIR = [reciprocal sum of each chunk | running sum phi], all committed in one tree
    bool phase = false;
    for (auto* k : keys) phase |= k->width_after != 0;
    Fp logup_witness;
    Fp4 challenges[2];
    std::vector<Fp4> cumsum(n_airs, Fp4::zero());
    std::vector<size_t> after_of;
    std::vector<b2s_matrix> after_views;
    b2s_pcs_data* after_pcs = nullptr;
    std::array<uint32_t, 8> after_commit{};
    if (phase) {
        e->stage_begin(ST_PERM_GEN);
        logup_witness = grind_device(e, scratch, ch, pk->log_up_pow_bits, e->pow_override[0]);
        challenges[0] = ch.sample_ext();
        challenges[1] = ch.sample_ext();
        trace.mark("  logup: pow + challenges", 2);
        for (size_t i = 0; i < n_airs; i++) if (keys[i]->width_after) after_of.push_back(i);
        after_views.reserve(after_of.size());
        std::vector<uint32_t*> last_rows;
        unsigned side_next = 0;
        for (size_t i : after_of) {
            const AirKeyDev& k = *keys[i];
            const unsigned n = logn[i];
            const uint64_t N = (uint64_t)1 << n;
            std::vector<const uint32_t*> cols;
            auto push_cols = [&](const b2s_matrix* m) { for (uint64_t c = 0; c < m->width; c++) cols.push_back(m->d + c * m->height); };
            if (k.width_prep) push_cols(k.prep_trace);
            for (uint32_t p = 0; p < airs[i].n_cached; p++) push_cols(airs[i].cached_traces[p]);
            if (k.width_common) push_cols(airs[i].common_main);
            // host constants: beta powers, per-interaction alpha + beta^len (bus + 1), public values
            std::vector<uint32_t> hc;
            Fp4 bp = Fp4::one();
            std::vector<Fp4> betas(k.max_fields + 1);
            for (uint32_t j = 0; j <= k.max_fields; j++) { betas[j] = bp; bp = bp * challenges[1]; }
            for (auto& v : betas) for (int c = 0; c < 4; c++) hc.push_back(v.c[c].v);
            const size_t consts_off = hc.size();
            for (size_t t = 0; t < k.inter_bus.size(); t++) {
                const Fp4 v = challenges[0] + betas[k.inter_len[t]] * Fp::from_canonical((k.inter_bus[t] + 1) % P);
                for (int c = 0; c < 4; c++) hc.push_back(v.c[c].v);
            }
            const size_t pv_off = hc.size();
            for (uint32_t j = 0; j < k.num_pv; j++) hc.push_back(Fp::from_canonical(airs[i].public_values[j]).v);
            hc.push_back(0);
            uint32_t* d_hc = scratch.upload(std::move(hc));
            const uint32_t** d_cols = scratch.upload(std::move(cols));
            LogupParams lp{};
            lp.program = k.d_iprogram; lp.prog_len = (uint32_t)k.iprogram.size();
            lp.cols = d_cols; lp.publics = d_hc + pv_off; lp.betas = d_hc; lp.consts = d_hc + consts_off;
            lp.out = scratch.words((size_t)4 * k.width_after * N);
            lp.row_sums = scratch.words((size_t)4 * N);
            lp.n = n;
            // every AIR's trace kernel and scan on a side stream: they are short, latency-bound and independent
            // (34 AIRs: 4.4 ms one after the other). The uploads above are on the engine's stream: fork after them.
            uint32_t* phi = lp.out + (size_t)4 * k.num_chunks * N;
            {
                cudaStream_t main_stream = e->stream;
                cudaStream_t side = aux_fork(e, side_next++);
                e->stream = side;
                try {
                    launch_logup(e, k.islots, k.num_chunks, lp);
                    scan_planes(e, scratch, lp.row_sums, phi, (uint32_t)N);
                } catch (...) {
                    e->stream = main_stream;
                    throw;
                }
                e->stream = main_stream;
            }
            last_rows.push_back(phi);
            b2s_matrix v;
            v.eng = e; v.d = lp.out; v.height = N; v.width = (uint64_t)4 * k.width_after; v.owned = false;
            after_views.push_back(v);
        }
        aux_join(e);
        trace.mark("  logup: traces enqueued", 2);
        // cumulative sums = last row of each phi column
        {
            std::vector<OpenDesc> descs;
            for (size_t a = 0; a < after_of.size(); a++) {
                const uint64_t N = (uint64_t)1 << logn[after_of[a]];
                descs.push_back(OpenDesc{last_rows[a], N, N - 1, 4, (uint32_t)(4 * a)});
            }
            std::vector<uint32_t> sums(4 * after_of.size());
            gather_rows(e, descs, sums.size(), sums.data());
            for (size_t a = 0; a < after_of.size(); a++) {
                Fp4& cs = cumsum[after_of[a]];
                for (int c = 0; c < 4; c++) cs.c[c] = Fp::from_canonical(sums[4 * a + c]);
                ch.observe_ext(cs);
            }
        }
        trace.mark("  logup: sums on host", 2);
        e->stage_end(ST_PERM_GEN);
        StageScope cs(e, ST_PERM_COMMIT);
        std::vector<const b2s_matrix*> ptrs;
        for (auto& v : after_views) ptrs.push_back(&v);
        after_pcs = commit_matrices(e, ptrs.data(), nullptr, ptrs.size());
        trace.mark("  logup: commit enqueued", 2);
        scratch.pcs.push_back(after_pcs);
        root_monty(e, after_pcs, after_commit.data());
        ch.observe_digest(after_commit.data());
    }
    const Fp4 alpha = ch.sample_ext();

    trace.mark("logup phase");
    // ---- quotient (stark/mod.rs:366-416; fri/domain.rs:59-155)
    std::vector<b2s_matrix> chunk_views;   // one N x 4 matrix per (AIR, chunk)
    std::vector<Fp> chunk_shifts;
    std::vector<unsigned> chunk_logn;
    {
        StageScope qs(e, ST_QUOTIENT);
        uint32_t max_c = 1;
        for (auto* k : keys) max_c = std::max(max_c, k->num_constraints);
        uint32_t* d_apow = scratch.words((size_t)max_c * 4);
        upload_ef_powers(e, d_apow, alpha, max_c);
        size_t total_chunks = 0;
        for (auto* k : keys) total_chunks += (size_t)1 << k->log_qd;
        chunk_views.reserve(total_chunks);
        QuotientBatch qbatch;
        bool any_jit = false;
        for (auto* k : keys) any_jit |= !k->jit_kernels.empty();
        if (any_jit) quotient_batch_begin(e, qbatch);
        for (size_t i = 0; i < n_airs; i++) {
            const AirKeyDev& k = *keys[i];
            const unsigned n = logn[i], lq = k.log_qd;
            const uint64_t N = (uint64_t)1 << n;
            // column pointer table: [preprocessed][cached parts][common]
            std::vector<const uint32_t*> cols;
            auto push_cols = [&](const b2s_matrix* m) { for (uint64_t c = 0; c < m->width; c++) cols.push_back(m->d + c * m->height); };
            if (k.width_prep) push_cols(k.prep_pcs->mats[0]);
            for (uint32_t p = 0; p < airs[i].n_cached; p++) push_cols(airs[i].cached[p]->mats[0]);
            if (k.width_common) {
                size_t pos = std::find(common_of.begin(), common_of.end(), i) - common_of.begin();
                push_cols(common_pcs->mats[pos]);
            }
            std::vector<uint32_t> efc(12, 0);
            if (k.width_after) {
                size_t pos = std::find(after_of.begin(), after_of.end(), i) - after_of.begin();
                push_cols(after_pcs->mats[pos]);
                for (int c = 0; c < 4; c++) { efc[c] = challenges[0].c[c].v; efc[4 + c] = challenges[1].c[c].v; efc[8 + c] = cumsum[i].c[c].v; }
            }
            uint32_t* d_efc = scratch.upload(std::move(efc));
            std::vector<uint32_t> pv(std::max<uint32_t>(k.num_pv, 1), 0);
            for (uint32_t j = 0; j < k.num_pv; j++) pv[j] = Fp::from_canonical(airs[i].public_values[j]).v;
            uint32_t* d_pv = scratch.upload(std::move(pv));
            const uint32_t** d_cols = scratch.upload(std::move(cols));
            QuotientParams qp{};
            qp.program = k.d_program; qp.prog_len = (uint32_t)k.program.size();
            qp.cols = d_cols; qp.publics = d_pv; qp.alpha_pows = d_apow; qp.ef_consts = d_efc;
            qp.out = scratch.words(((size_t)4 << lq) * N);
            qp.n = n; qp.b = b; qp.lq = lq;
            const Fp gq = two_adic_generator(n + lq);
            qp.gq = gq.v;
            qp.ginv = fp_inv(two_adic_generator(n)).v;
            // x^N on the quotient domain: (g * gq^k)^N = g^N * w_qd^(k mod qd)
            const Fp gN = fp_pow(Fp::raw(M_GENERATOR), N);
            const Fp wq = two_adic_generator(lq);
            Fp wpow = Fp::one();
            for (uint32_t j = 0; j < (1u << lq); j++) {
                Fp zh = gN * wpow - Fp::one();
                qp.zh[j] = zh.v;
                qp.inv_zh[j] = fp_inv(zh).v;
                wpow *= wq;
            }
            if (!k.jit_kernels.empty()) {
                // straight-line kernels generated from the DAG at key load (qjit.cu); same inputs, same output layout
                std::vector<const uint32_t*> mats;   // one entry per partition: a key may carry up to 1024 cached mains
                if (k.width_prep) mats.push_back(k.prep_pcs->mats[0]->d);
                for (uint32_t p = 0; p < airs[i].n_cached; p++) mats.push_back(airs[i].cached[p]->mats[0]->d);
                if (k.width_common) {
                    size_t pos = std::find(common_of.begin(), common_of.end(), i) - common_of.begin();
                    mats.push_back(common_pcs->mats[pos]->d);
                }
                if (k.width_after) {
                    size_t pos = std::find(after_of.begin(), after_of.end(), i) - after_of.begin();
                    mats.push_back(after_pcs->mats[pos]->d);
                }
                B2_REQUIRE(mats.size() == k.jit_num_mats, B2S_ERR_INTERNAL, "quotient JIT: matrix count mismatch");
                uint32_t* accbuf = k.jit_kernels.size() > 1 ? scratch.words((k.jit_kernels.size() * 4) << (n + lq)) : nullptr;
                quotient_jit_launch(qbatch, k, mats.data(), d_pv, d_apow, d_efc, accbuf, qp.out, n, lq);
            } else {
                launch_quotient(e, k.num_slots, k.num_slots_e, qp);
            }
            Fp sh = Fp::raw(M_GENERATOR);
            for (uint32_t j = 0; j < (1u << lq); j++) {
                b2s_matrix v;
                v.eng = e; v.d = qp.out + (size_t)j * 4 * N; v.height = N; v.width = 4; v.owned = false;
                chunk_views.push_back(v);
                chunk_shifts.push_back(sh);
                chunk_logn.push_back(n);
                sh *= gq;
            }
        }
        if (any_jit) quotient_batch_end(qbatch);
    }
    trace.mark("quotient eval");
    std::vector<const b2s_matrix*> chunk_ptrs;
    for (auto& v : chunk_views) chunk_ptrs.push_back(&v);
    b2s_pcs_data* quotient_pcs;
    std::array<uint32_t, 8> quotient_commit;
    {
        StageScope cs(e, ST_QUOTIENT_COMMIT);
        quotient_pcs = commit_matrices(e, chunk_ptrs.data(), chunk_shifts.data(), chunk_ptrs.size());
        scratch.pcs.push_back(quotient_pcs);
        root_monty(e, quotient_pcs, quotient_commit.data());
    }
    ch.observe_digest(quotient_commit.data());
    const Fp4 zeta = ch.sample_ext();

    trace.mark("quotient commit");
    // ---- opening rounds: [preprocessed per AIR][cached mains][common main][quotient]  (stark/mod.rs:418-698)
    StageScope open_scope(e, ST_OPENING);
    std::vector<Round> rounds;
    auto trace_mat = [&](const b2s_matrix* lde, size_t i) {
        return RoundMat{lde, logn[i], {zeta, zeta * two_adic_generator(logn[i])}, {}};
    };
    for (size_t i = 0; i < n_airs; i++)
        if (keys[i]->width_prep) rounds.push_back(Round{keys[i]->prep_pcs, {trace_mat(keys[i]->prep_pcs->mats[0], i)}});
    const size_t first_cached = rounds.size();
    for (size_t i = 0; i < n_airs; i++)
        for (uint32_t p = 0; p < airs[i].n_cached; p++)
            rounds.push_back(Round{airs[i].cached[p], {trace_mat(airs[i].cached[p]->mats[0], i)}});
    const size_t common_round = rounds.size();
    {
        Round r{common_pcs, {}};
        for (size_t m = 0; m < common_of.size(); m++) r.mats.push_back(trace_mat(common_pcs->mats[m], common_of[m]));
        rounds.push_back(r);
    }
    const size_t after_round = rounds.size();
    if (phase) {
        Round r{after_pcs, {}};
        for (size_t m = 0; m < after_of.size(); m++) r.mats.push_back(trace_mat(after_pcs->mats[m], after_of[m]));
        rounds.push_back(r);
    }
    const size_t quotient_round = rounds.size();
    {
        Round r{quotient_pcs, {}};
        for (size_t m = 0; m < chunk_views.size(); m++) r.mats.push_back(RoundMat{quotient_pcs->mats[m], chunk_logn[m], {zeta}, {}});
        rounds.push_back(r);
    }
    // (i) out-of-domain values. Weights are shared by all matrices of one trace height: index 0 = zeta, 1 = zeta * w_n.
    {
        StageScope ood_scope(e, ST_OOD);
        std::map<unsigned, std::array<uint32_t*, 2>> weights;
        auto get_weights = [&](unsigned n, int which, const Fp4& z) {
            auto& slot = weights[n];
            if (!slot[which]) {
                slot[which] = scratch.words((size_t)4 << n);
                const uint32_t N = 1u << n;
                bary_weights_kernel<<<(N + 127) / 128, 128, 0, e->stream>>>(slot[which], n, z, two_adic_generator(n).v);
                B2_CHECK_LAUNCH();
                e->launches[3]++;
            }
            return slot[which];
        };
        // Value order = transcript order (round -> matrix -> point -> column, two_adic_pcs.rs:63-95): global column id c has
        // its value at vals[4 c]. Columns are then grouped by (trace height, point) for the launches.
        struct Job { size_t r, m, p; size_t first_col; };
        std::vector<Job> jobs;
        std::vector<BaryCol> cols;
        std::vector<uint32_t> scales;
        std::map<std::pair<unsigned, int>, std::vector<uint32_t>> by_group;   // (log_n, point index) -> column ids
        size_t partial_words = 0;
        const Fp ginv = fp_inv(Fp::raw(M_GENERATOR));
        for (size_t r = 0; r < rounds.size(); r++)
            for (size_t m = 0; m < rounds[r].mats.size(); m++)
                for (size_t p = 0; p < rounds[r].mats[m].points.size(); p++) {
                    const RoundMat& rm = rounds[r].mats[m];
                    const uint32_t splits = (uint32_t)std::min<uint64_t>(64, std::max<uint64_t>(1, ((uint64_t)1 << rm.log_n) / 4096));
                    const Fp ninv = fp_inv(Fp::from_canonical((uint32_t)(((uint64_t)1 << rm.log_n) % P)));
                    const Fp4 scale = (fp4_pow2k(rm.points[p] * ginv, rm.log_n) - Fp::one()) * ninv;
                    const uint32_t scale_idx = (uint32_t)(scales.size() / 4);
                    for (int c = 0; c < 4; c++) scales.push_back(scale.c[c].v);
                    jobs.push_back(Job{r, m, p, cols.size()});
                    auto& grp = by_group[{rm.log_n, (int)p}];
                    for (uint64_t c = 0; c < rm.lde->width; c++) {
                        grp.push_back((uint32_t)cols.size());
                        cols.push_back(BaryCol{rm.lde->d + c * rm.lde->height, (uint32_t)partial_words, splits, scale_idx, 0});
                        partial_words += (size_t)splits * 4;
                    }
                }
        B2_REQUIRE(partial_words < 0xFFFFFFFFull, B2S_ERR_INTERNAL, "prove: too many opened columns");
        const size_t ncols = cols.size(), val_words = ncols * 4;
        // device copies: the column table once in value order (for the finish kernel) and once grouped (for the dot kernels)
        std::vector<uint32_t> table_words((ncols * 2) * (sizeof(BaryCol) / 4));
        {
            BaryCol* t = reinterpret_cast<BaryCol*>(table_words.data());
            size_t at = 0;
            for (auto& c : cols) t[at++] = c;
            for (auto& kv : by_group)
                for (uint32_t id : kv.second) t[at++] = cols[id];
        }
        static_assert(sizeof(BaryCol) == 24, "BaryCol layout");
        const BaryCol* d_cols = reinterpret_cast<const BaryCol*>(scratch.upload(std::move(table_words)));
        const uint32_t* d_scales = scratch.upload(std::move(scales));
        uint32_t* d_partial = scratch.words(std::max<size_t>(partial_words, 4));
        uint32_t* d_vals = scratch.words(std::max<size_t>(val_words, 4));
        uint32_t* h_vals = (uint32_t*)e->host_pinned(val_words * 4);
        {
            size_t at = ncols;
            for (auto& kv : by_group) {
                const unsigned n = kv.first.first;
                const int which = kv.first.second;
                // every matrix of this height shares the point: zeta (0) or zeta * w_n (1)
                const Fp4 z = which == 0 ? zeta : zeta * two_adic_generator(n);
                uint32_t* w = get_weights(n, which, z);
                const uint32_t splits = cols[kv.second[0]].splits;
                bary_dot_multi_kernel<<<dim3((unsigned)kv.second.size(), splits), DOT_THREADS, 0, e->stream>>>(d_cols + at, w, n, d_partial);
                B2_CHECK_LAUNCH();
                e->launches[3]++;
                at += kv.second.size();
            }
        }
        bary_finish_multi_kernel<<<(unsigned)((ncols + 127) / 128), 128, 0, e->stream>>>(d_cols, (uint32_t)ncols, d_partial, d_scales, d_vals);
        B2_CHECK_LAUNCH();
        e->launches[3]++;
        B2_CUDA(cudaMemcpyAsync(h_vals, d_vals, val_words * 4, cudaMemcpyDeviceToHost, e->stream));
        trace.mark("ood values enqueued");
        B2_CUDA(cudaStreamSynchronize(e->stream));
        for (auto& j : jobs) {
            RoundMat& rm = rounds[j.r].mats[j.m];
            std::vector<Fp4> vals(rm.lde->width);
            const uint32_t* src = h_vals + j.first_col * 4;
            for (uint64_t c = 0; c < rm.lde->width; c++) {
                vals[c] = Fp4(Fp::raw(src[4 * c]), Fp::raw(src[4 * c + 1]), Fp::raw(src[4 * c + 2]), Fp::raw(src[4 * c + 3]));
                ch.observe_ext(vals[c]);
            }
            if (rm.values.size() <= j.p) rm.values.resize(j.p + 1);
            rm.values[j.p] = std::move(vals);
        }
    }
    const Fp4 fri_alpha = ch.sample_ext();

    trace.mark("observe opened values");
    // (ii) reduced openings per LDE log height; the alpha power runs on per height (two_adic_pcs.rs:175-196, 313-325)
    const unsigned L = log_max_height + b;
    std::map<unsigned, uint32_t*> ro;
    {
        StageScope ro_scope(e, ST_REDUCED);
        uint64_t max_w = 1;
        for (auto& r : rounds) for (auto& m : r.mats) max_w = std::max<uint64_t>(max_w, m.lde->width);
        // alpha^t for t up to the widest matrix; the host copy lives in the scratch until the proof is done
        std::vector<uint32_t> apw((max_w + 1) * 4);
        std::vector<Fp4> ap_host(max_w + 1);
        {
            Fp4 pw = Fp4::one();
            for (uint64_t t = 0; t <= max_w; t++) {
                ap_host[t] = pw;
                for (int c = 0; c < 4; c++) apw[4 * t + c] = pw.c[c].v;
                pw = pw * fri_alpha;
            }
        }
        const uint32_t* d_apow = scratch.upload(std::move(apw));
        std::map<unsigned, Fp4> alpha_pow;
        struct HeightBatch { std::vector<uint32_t> descs; uint32_t nmats = 0; Fp4 z[2]; uint32_t npoints = 0; };
        std::map<unsigned, HeightBatch> batches;
        for (auto& r : rounds)
            for (auto& m : r.mats) {
                const unsigned h = m.log_n + b;
                if (!ro.count(h)) {
                    ro[h] = scratch.words((size_t)4 << h);
                    alpha_pow[h] = Fp4::one();
                }
                HeightBatch& hb = batches[h];
                B2_REQUIRE(m.points.size() >= 1 && m.points.size() <= 2, B2S_ERR_INTERNAL, "prove: a matrix is opened at 1 or 2 points");
                for (size_t j = 0; j < m.points.size(); j++) {
                    if (hb.npoints <= j) { hb.z[j] = m.points[j]; hb.npoints = (uint32_t)j + 1; }
                    B2_REQUIRE(hb.z[j] == m.points[j], B2S_ERR_INTERNAL, "prove: matrices of one height share their opening points");
                }
                const uint64_t mp = (uint64_t)(uintptr_t)m.lde->d;
                uint32_t d[RO_DESC_WORDS] = {(uint32_t)mp, (uint32_t)(mp >> 32), (uint32_t)m.lde->width, (uint32_t)m.points.size()};
                const Fp4 aw = ap_host[m.lde->width];
                for (size_t j = 0; j < m.points.size(); j++) {
                    Fp4 cz = Fp4::zero();
                    for (size_t t = 0; t < m.values[j].size(); t++) cz += m.values[j][t] * ap_host[t];
                    for (int c = 0; c < 4; c++) { d[4 + 4 * j + c] = cz.c[c].v; d[12 + 4 * j + c] = alpha_pow[h].c[c].v; }
                    alpha_pow[h] = alpha_pow[h] * aw;
                }
                hb.descs.insert(hb.descs.end(), d, d + RO_DESC_WORDS);
                hb.nmats++;
            }
        // uploads first (engine's stream), then one kernel per height, each on a side stream
        std::map<unsigned, const uint32_t*> desc_dev;
        for (auto& kv : batches) desc_dev[kv.first] = scratch.upload(std::move(kv.second.descs));
        unsigned ro_side = 0;
        cudaStream_t ro_main = e->stream;
        for (auto& kv : batches) {
            const unsigned h = kv.first;
            HeightBatch& hb = kv.second;
            const uint32_t nm = hb.nmats;
            const uint32_t* d_descs = desc_dev[h];
            struct Restore { b2s_engine* e; cudaStream_t s; ~Restore() { e->stream = s; } } restore{e, ro_main};
            e->stream = aux_fork(e, ro_side++);
            if (h >= 17) {
                reduced_open_height_kernel<<<((1u << h) + 127) / 128, 128, 0, e->stream>>>(
                    d_descs, nm, h, d_apow, ro[h], hb.z[0], hb.npoints > 1 ? hb.z[1] : Fp4::zero(), hb.npoints > 1 ? 1u : 0u,
                    two_adic_generator(h).v);
            } else {
                // S threads per row, S = 2^(17 - h) capped at 32: the launch has ~2^17 threads whatever the height
                const uint32_t S = std::min(32u, 1u << (17 - h));
                reduced_open_split_kernel<<<((1u << h) + RO_ROWS - 1) / RO_ROWS, dim3(RO_ROWS, S), (size_t)S * 8 * RO_ROWS * 4, e->stream>>>(
                    d_descs, nm, h, d_apow, ro[h], hb.z[0], hb.npoints > 1 ? hb.z[1] : Fp4::zero(), hb.npoints > 1 ? 1u : 0u,
                    two_adic_generator(h).v);
            }
            B2_CHECK_LAUNCH();
            e->launches[3]++;
        }
        aux_join(e);
    }

    trace.mark("reduced openings");
    // (iii) FRI commit phase (two_adic_pcs.rs:111-130; fri/mod.rs:139-165)
    std::vector<b2s_pcs_data*> fri_trees;
    std::vector<b2s_matrix*> fri_leaves;
    std::vector<std::array<uint32_t, 8>> fri_commits;
    uint32_t* folded = ro[L];
    struct LeafGuard {
        std::vector<b2s_matrix*>& v;
        ~LeafGuard() { for (auto* m : v) b2s_matrix_free(m); }
    } leaf_guard{fri_leaves};
    e->stage_begin(ST_FRI_COMMIT);
    for (unsigned lh = L; lh > b; lh--) {
        const uint32_t R = 1u << (lh - 1);
        b2s_matrix* leaf = new_matrix(e, R, 8);
        fri_leaves.push_back(leaf);
        fri_pack_kernel<<<(R + 127) / 128, 128, 0, e->stream>>>(folded, leaf->d, R);
        B2_CHECK_LAUNCH();
        e->launches[3]++;
        b2s_pcs_data* t = new b2s_pcs_data();
        t->eng = e;
        b2s_matrix* view = new b2s_matrix(*leaf);
        view->owned = false;
        t->mats.push_back(view);
        t->owns.push_back(true);
        scratch.pcs.push_back(t);
        mmcs_build(e, t);
        trace.mark("  fri layer: tree enqueued", 2);
        fri_trees.push_back(t);
        std::array<uint32_t, 8> root;
        root_monty(e, t, root.data());
        trace.mark("  fri layer: root on host", 2);
        fri_commits.push_back(root);
        ch.observe_digest(root.data());
        const Fp4 beta = ch.sample_ext();
        uint32_t* next = scratch.words((size_t)4 * R);
        auto it = ro.find(lh - 1);
        fri_fold_kernel<<<(R + 127) / 128, 128, 0, e->stream>>>(folded, next, it != ro.end() ? it->second : nullptr, lh,
                                                               two_adic_generator(lh).v, beta, beta * beta);
        B2_CHECK_LAUNCH();
        e->launches[3]++;
        folded = next;
        trace.mark("  fri layer: fold", 2);
    }
    e->stage_end(ST_FRI_COMMIT);
    trace.mark("fri commit phase");
    // final polynomial: log_final_poly_len = 0, the constant (two_adic_pcs.rs:52-58)
    Fp4 final_poly;
    {
        const uint32_t len = 1u << b;
        std::vector<uint32_t> h((size_t)4 * len);
        B2_CUDA(cudaMemcpyAsync(h.data(), folded, h.size() * 4, cudaMemcpyDeviceToHost, e->stream));
        B2_CUDA(cudaStreamSynchronize(e->stream));
        for (int c = 0; c < 4; c++) final_poly.c[c] = Fp::raw(h[(size_t)c * len]);
        for (uint32_t i = 1; i < len; i++)
            for (int c = 0; c < 4; c++)
                B2_REQUIRE(h[(size_t)c * len + i] == h[(size_t)c * len], B2S_ERR_INTERNAL, "fri: folded polynomial is not constant");
    }
    ch.observe_ext(final_poly);
    // (iv) proof of work: minimum witness (duplex.rs:122-130)
    const Fp pow_witness = grind_device(e, scratch, ch, e->params.proof_of_work_bits, e->pow_override[1]);
    trace.mark("final poly + pow");
    // (v) queries: every opening of every query in one gather (two_adic_pcs.rs:170-372; fri/mod.rs:72-135)
    const unsigned nq = e->params.num_queries;
    std::vector<uint32_t> indices(nq);
    for (auto& ix : indices) ix = ch.sample_bits(L);
    std::vector<OpenDesc> descs;
    size_t off = 0;
    auto push = [&](const uint32_t* base, uint64_t stride, uint64_t row, uint32_t count, bool digest = false) {
        B2_REQUIRE(off + count < 0xFFFFFFFFull, B2S_ERR_INTERNAL, "prove: opening buffer too large");
        descs.push_back(OpenDesc{base, stride, row, digest ? digest_words_desc(e) : count, (uint32_t)off});
        off += count;
    };
    for (uint32_t ix : indices) {
        for (auto& r : rounds) {
            const b2s_pcs_data* d = r.pcs;
            const uint64_t idx = (uint64_t)ix >> (L - d->log_max_height);
            for (auto* m : d->mats) push(m->d, m->height, idx >> (d->log_max_height - log2_exact(m->height)), (uint32_t)m->width);
            for (unsigned l = 0; l < d->log_max_height; l++) push(d->layers[l], d->layer_nodes[l], (idx >> l) ^ 1, 8, true);
        }
        for (size_t i = 0; i < fri_trees.size(); i++) {
            const b2s_pcs_data* d = fri_trees[i];
            const uint64_t pos = (uint64_t)ix >> i, pair = pos >> 1;
            // sibling value: the other element of the leaf pair = 4 columns of the leaf matrix
            const b2s_matrix* leaf = d->mats[0];
            push(leaf->d + (uint64_t)4 * ((pos & 1) ^ 1) * leaf->height, leaf->height, pair, 4);
            for (unsigned l = 0; l < d->log_max_height; l++) push(d->layers[l], d->layer_nodes[l], (pair >> l) ^ 1, 8, true);
        }
    }
    std::vector<uint32_t> opened(off);
    gather_rows(e, descs, off, opened.data());

    trace.mark("query gather");
    // ---- proof bytes (crates/sdk/src/codec.rs:92-330) -------------------------------------------------------------------
    out.clear();
    out.reserve(opened.size() * 4 + (1u << 16));
    ByteWriter w{out};
    w.raw_digests = e->hash_kind != 0;
    w.u32(1);  // CODEC_VERSION
    w.u32((uint32_t)main_commits.size());
    for (auto& c : main_commits) w.digest_monty(c.data());
    w.u32(phase ? 1 : 0);  // after_challenge commitments: one per challenge phase
    if (phase) w.digest_monty(after_commit.data());
    w.digest_monty(quotient_commit.data());
    // FriProof
    w.u32((uint32_t)fri_commits.size());
    for (auto& c : fri_commits) w.digest_monty(c.data());
    w.u32(nq);
    {
        size_t pos = 0;
        for (unsigned q = 0; q < nq; q++) {
            w.u32((uint32_t)rounds.size());
            for (auto& r : rounds) {
                const b2s_pcs_data* d = r.pcs;
                w.u32((uint32_t)d->mats.size());
                for (auto* m : d->mats) {
                    w.u32((uint32_t)m->width);
                    w.words(opened.data() + pos, m->width);
                    pos += m->width;
                }
                w.u32(d->log_max_height);
                w.words(opened.data() + pos, (size_t)8 * d->log_max_height);
                pos += (size_t)8 * d->log_max_height;
            }
            w.u32((uint32_t)fri_trees.size());
            for (auto* d : fri_trees) {
                w.words(opened.data() + pos, 4);
                pos += 4;
                w.u32(d->log_max_height);
                w.words(opened.data() + pos, (size_t)8 * d->log_max_height);
                pos += (size_t)8 * d->log_max_height;
            }
        }
    }
    w.u32(1);
    w.ef(final_poly);
    w.f(pow_witness);
    // OpenedValues
    auto adjacent = [&](const RoundMat& m) { w.efs(m.values[0]); w.efs(m.values[1]); };
    w.u32((uint32_t)first_cached);
    for (size_t r = 0; r < first_cached; r++) adjacent(rounds[r].mats[0]);
    w.u32((uint32_t)(common_round - first_cached + 1));
    for (size_t r = first_cached; r < common_round; r++) { w.u32(1); adjacent(rounds[r].mats[0]); }
    w.u32((uint32_t)rounds[common_round].mats.size());
    for (auto& m : rounds[common_round].mats) adjacent(m);
    w.u32(phase ? 1 : 0);  // after_challenge phases
    if (phase) {
        w.u32((uint32_t)rounds[after_round].mats.size());
        for (auto& m : rounds[after_round].mats) adjacent(m);
    }
    w.u32((uint32_t)n_airs);
    {
        size_t m = 0;
        for (size_t i = 0; i < n_airs; i++) {
            const uint32_t qd = 1u << keys[i]->log_qd;
            w.u32(qd);
            for (uint32_t j = 0; j < qd; j++) w.efs(rounds[quotient_round].mats[m++].values[0]);
        }
    }
    // per_air
    w.u32((uint32_t)n_airs);
    for (size_t i = 0; i < n_airs; i++) {
        w.u32(airs[i].air_id);
        w.u32(1u << logn[i]);
        if (keys[i]->width_after) { w.u32(1); w.u32(1); w.ef(cumsum[i]); }  // exposed_values_after_challenge: [phase][value]
        else w.u32(0);
        w.u32(airs[i].n_public_values);
        w.words(airs[i].public_values, airs[i].n_public_values);
    }
    out.push_back(0);       // RapPhaseSeqKind::FriLogUp
    if (phase) w.f(logup_witness); else w.u32(0xFFFFFFFFu);
    B2_CUDA(cudaStreamSynchronize(e->stream));
    trace.mark("encode");
}

}  // namespace b2
