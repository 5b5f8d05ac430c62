// ORACLE — TEST INFRASTRUCTURE ONLY (see babybear.hpp header).
//
// Proving/verifying key in this repo's flat "key blob" format (the reference's MultiStarkProvingKey /
// MultiStarkVerifyingKey are serde types of the absent openvm-stark-backend; their content as used in-tree:
// extensions/native/recursion/src/types.rs:19-95) and evaluation of a SymbolicExpressionDag
// (extensions/native/recursion/src/stark/mod.rs:1023-1085: nodes in topological order, constraints folded
// Horner-style  acc = acc * alpha + node[constraint_idx[k]]).
//
// Blob = little-endian u32 words:
//   'B2PK', version, pre_hash[8], log_up_pow_bits, num_airs, then per AIR:
//     width_preprocessed, n_cached, cached_widths[n_cached], width_common_main, width_after_challenge (EF columns),
//     log_quotient_degree, num_public_values, num_exposed_values,
//     [if width_preprocessed: commit[8], log_height, row-major canonical data[height*width]]
//     num_nodes, nodes[num_nodes][4] = {kind, a, b, c}, num_constraints, constraint_idx[num_constraints]
//     num_interactions, per interaction {bus, type (0 send, 1 receive), count_node, n_fields, field_nodes[n_fields]},
//     num_chunks, per chunk {n, interaction indices[n]}     (FRI-LogUp: one reciprocal-sum column per chunk)
// Node kinds: 0 Constant(a) | 1 Preprocessed(col a, offset b) | 2 Main(col a, offset b, part c) |
//   3 Permutation(EF col a, offset b) | 4 Public(a) | 5 Challenge(a) | 6 Exposed(a) |
//   7 IsFirstRow | 8 IsLastRow | 9 IsTransition | 10 Add(a,b) | 11 Sub(a,b) | 12 Neg(a) | 13 Mul(a,b)
#pragma once
#include "babybear.hpp"
#include "poseidon2.hpp"
#include <stdexcept>
#include <string>

namespace orc {

enum NodeKind { N_CONST = 0, N_PREP, N_MAIN, N_PERM, N_PUBLIC, N_CHALLENGE, N_EXPOSED, N_FIRST, N_LAST, N_TRANS,
                N_ADD, N_SUB, N_NEG, N_MUL };
struct Node { uint32_t kind, a, b, c; };

struct Interaction { uint32_t bus, type, count_node; std::vector<uint32_t> fields; };

struct AirKey {
    std::vector<Interaction> interactions;
    std::vector<std::vector<uint32_t>> chunks;
    uint32_t width_prep = 0;
    std::vector<uint32_t> cached_widths;
    uint32_t width_common = 0, width_after = 0, log_qd = 0, num_pv = 0, num_exposed = 0;
    Digest prep_commit;
    uint32_t prep_log_height = 0;
    std::vector<F> prep;  // row-major
    std::vector<Node> nodes;
    std::vector<uint32_t> constraint_idx;
    size_t num_parts() const { return cached_widths.size() + (width_common ? 1 : 0); }
    uint32_t part_width(size_t p) const { return p < cached_widths.size() ? cached_widths[p] : width_common; }
};

// LinearConstraint of the verifying key (extensions/native/recursion/src/types.rs:70; checked at
// extensions/native/recursion/src/stark/mod.rs:836-901 and, on the proving side, crates/vm/src/arch/vm.rs:627-652):
// sum_i coefficients[air i] * height_i < threshold over the AIRs present in a proof
struct HeightConstraint {
    std::vector<uint32_t> coefficients;  // one per AIR of the key
    uint32_t threshold = 0;              // <= p (view.rs:64)
};
struct KeyBlob {
    Digest pre_hash;
    uint32_t log_up_pow_bits = 0;
    std::vector<AirKey> airs;
    std::vector<HeightConstraint> height_constraints;
    // index of the first violated constraint for the given (air id, height) pairs, or -1
    int violated_height_constraint(const std::vector<uint32_t>& ids, const std::vector<unsigned>& log_heights) const {
        for (size_t c = 0; c < height_constraints.size(); c++) {
            unsigned __int128 sum = 0;
            for (size_t k = 0; k < ids.size(); k++)
                sum += (unsigned __int128)height_constraints[c].coefficients[ids[k]] << log_heights[k];
            if (sum >= height_constraints[c].threshold) return (int)c;
        }
        return -1;
    }
};

inline KeyBlob parse_key(const uint32_t* w, size_t n) {
    size_t i = 0;
    auto rd = [&]() -> uint32_t { if (i >= n) throw std::runtime_error("key blob truncated"); return w[i++]; };
    if (rd() != 0x4B503242u) throw std::runtime_error("key blob: bad magic");
    const uint32_t version = rd();   // 2: no trace-height constraints; 3: they follow the AIRs
    if (version != 2 && version != 3) throw std::runtime_error("key blob: bad version");
    KeyBlob k;
    for (int j = 0; j < 8; j++) k.pre_hash[j] = F(rd());
    k.log_up_pow_bits = rd();
    uint32_t na = rd();
    k.airs.resize(na);
    for (auto& a : k.airs) {
        a.width_prep = rd();
        a.cached_widths.resize(rd());
        for (auto& c : a.cached_widths) c = rd();
        a.width_common = rd(); a.width_after = rd(); a.log_qd = rd(); a.num_pv = rd(); a.num_exposed = rd();
        if (a.width_prep) {
            for (int j = 0; j < 8; j++) a.prep_commit[j] = F(rd());
            a.prep_log_height = rd();
            size_t cnt = ((size_t)a.width_prep) << a.prep_log_height;
            a.prep.resize(cnt);
            for (size_t j = 0; j < cnt; j++) a.prep[j] = F(rd());
        }
        a.nodes.resize(rd());
        for (auto& nd : a.nodes) { nd.kind = rd(); nd.a = rd(); nd.b = rd(); nd.c = rd(); }
        a.constraint_idx.resize(rd());
        for (auto& c : a.constraint_idx) c = rd();
        a.interactions.resize(rd());
        for (auto& it : a.interactions) {
            it.bus = rd(); it.type = rd(); it.count_node = rd();
            it.fields.resize(rd());
            for (auto& f : it.fields) f = rd();
        }
        a.chunks.resize(rd());
        for (auto& ch : a.chunks) { ch.resize(rd()); for (auto& x : ch) x = rd(); }
        if (a.interactions.empty() != (a.width_after == 0) || (a.width_after && a.width_after != a.chunks.size() + 1) ||
            a.num_exposed != (a.width_after ? 1u : 0u))
            throw std::runtime_error("key blob: interaction / after-challenge shape mismatch");
    }
    if (version >= 3) {
        k.height_constraints.resize(rd());
        for (auto& hc : k.height_constraints) {
            hc.coefficients.resize(k.airs.size());
            for (auto& c : hc.coefficients) c = rd();
            hc.threshold = rd();
            if (hc.threshold > P) throw std::runtime_error("key blob: trace-height threshold above the field order");
        }
    }
    if (i != n) throw std::runtime_error("key blob: trailing words");
    return k;
}

// Row views for one evaluation: T = F (on-domain rows) or EF (out-of-domain opened values).
template <typename T>
struct EvalInputs {
    const T* prep[2] = {nullptr, nullptr};          // [offset][col]
    std::vector<const T*> main[2];                   // [offset][part][col]
    const EF* perm[2] = {nullptr, nullptr};          // [offset][EF col]
    const F* publics = nullptr;
    const EF* challenges = nullptr;
    const EF* exposed = nullptr;
    T is_first, is_last, is_transition;
};

inline EF lift(F x) { return EF(x); }
inline EF lift(const EF& x) { return x; }

// Evaluates every node (as EF, the common super-type). Nodes that need inputs the caller did not provide
// (perm / challenges / exposed == nullptr) evaluate to zero; callers only read nodes that do not depend on them.
template <typename T>
inline std::vector<EF> eval_dag_nodes(const AirKey& air, const EvalInputs<T>& in) {
    std::vector<EF> v(air.nodes.size());
    for (size_t i = 0; i < air.nodes.size(); i++) {
        const Node& nd = air.nodes[i];
        if ((nd.kind == N_PERM && !in.perm[0]) || (nd.kind == N_CHALLENGE && !in.challenges) || (nd.kind == N_EXPOSED && !in.exposed))
            continue;
        switch (nd.kind) {
            case N_CONST: v[i] = EF(F(nd.a)); break;
            case N_PREP: v[i] = lift(in.prep[nd.b][nd.a]); break;
            case N_MAIN: v[i] = lift(in.main[nd.b][nd.c][nd.a]); break;
            case N_PERM: v[i] = in.perm[nd.b][nd.a]; break;
            case N_PUBLIC: v[i] = EF(in.publics[nd.a]); break;
            case N_CHALLENGE: v[i] = in.challenges[nd.a]; break;
            case N_EXPOSED: v[i] = in.exposed[nd.a]; break;
            case N_FIRST: v[i] = lift(in.is_first); break;
            case N_LAST: v[i] = lift(in.is_last); break;
            case N_TRANS: v[i] = lift(in.is_transition); break;
            case N_ADD: v[i] = v[nd.a] + v[nd.b]; break;
            case N_SUB: v[i] = v[nd.a] - v[nd.b]; break;
            case N_NEG: v[i] = -v[nd.a]; break;
            case N_MUL: v[i] = v[nd.a] * v[nd.b]; break;
            default: throw std::runtime_error("dag: bad node kind");
        }
    }
    return v;
}

// alpha-folded constraint value: acc = acc * alpha + node[constraint_idx[k]]
template <typename T>
inline EF eval_dag_fold(const AirKey& air, const EvalInputs<T>& in, const EF& alpha) {
    std::vector<EF> v = eval_dag_nodes(air, in);
    EF acc;
    for (uint32_t c : air.constraint_idx) acc = acc * alpha + v[c];
    return acc;
}

}  // namespace orc
