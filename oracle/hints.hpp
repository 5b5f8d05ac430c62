// ORACLE — TEST INFRASTRUCTURE ONLY (see babybear.hpp header).
//
// Hint stream of a proof for the recursive verifier program: a restatement of `Hintable::write`, one overload per
// `impl Hintable<InnerConfig> for T` of the reference, composed the way the reference composes them:
//   extensions/native/recursion/src/hints.rs:71-81 (usize), :84-94 (Val), :97-112 (Challenge), :134-154 (Vec<I>),
//   :156-170 (Vec<usize>), :188-198 (Vec<Val>), :200-216 (Vec<Challenge>), :218-245 (Vec<Vec<Challenge>>),
//   :272-316 (Proof), :318-342 (AirProofData), :344-366 (OpeningProof), :368-398 (OpenedValues),
//   :400-415 (AdjacentOpenedValues), :417-444 (Commitments);
//   extensions/native/recursion/src/fri/hints.rs:22-36 (digest), :38-60 (commit-phase step), :64-88 (query proof),
//   :92-122 (FriProof), :124-150 (batch opening), :157-162 (opening proof).
// In-tree known answers: hints.rs:455-531 (test_var_array, test_ext_array), reproduced in tests/test_hints.py.
#pragma once
#include "proof.hpp"
#include <algorithm>

namespace orc {

using Stream = std::vector<std::vector<uint32_t>>;
inline void extend(Stream& s, const Stream& t) { s.insert(s.end(), t.begin(), t.end()); }

inline Stream write_usize(size_t x) { return {{(uint32_t)x}}; }
inline Stream write(const F& x) { return {{x.v}}; }
inline Stream write(const EF& x) { return {{x.c[0].v}, {x.c[1].v}, {x.c[2].v}, {x.c[3].v}}; }
inline Stream write(const Digest& d) {
    Stream s;
    for (int i = 0; i < 8; i++) s.push_back({d[i].v});
    return s;
}
inline Stream write_usizes(const std::vector<size_t>& v) {
    std::vector<uint32_t> one;
    for (size_t x : v) one.push_back((uint32_t)x);
    return {one};
}
inline Stream write(const std::vector<F>& v) {
    std::vector<uint32_t> one;
    for (const F& x : v) one.push_back(x.v);
    return {one};
}
inline Stream write(const std::vector<EF>& v) {
    std::vector<uint32_t> flat;
    for (const EF& x : v) for (int c = 0; c < 4; c++) flat.push_back(x.c[c].v);
    return {{(uint32_t)v.size()}, flat};
}
inline Stream write(const AdjacentOpened& a) {
    Stream s = write(a.local);
    extend(s, write(a.next));
    return s;
}
inline Stream write_opening_proof(const std::vector<Digest>& v) {
    std::vector<uint32_t> flat;
    for (const Digest& d : v) for (int i = 0; i < 8; i++) flat.push_back(d[i].v);
    return {{(uint32_t)v.size()}, flat};
}
inline Stream write(const BatchOpening& b) {
    std::vector<uint32_t> flat;
    for (auto& row : b.opened_values) for (const F& x : row) flat.push_back(x.v);
    Stream s = {{(uint32_t)flat.size()}, flat};
    extend(s, write_opening_proof(b.opening_proof));
    return s;
}
inline Stream write(const CommitPhaseStep& c) {
    Stream s = write(c.sibling_value);
    extend(s, write_opening_proof(c.opening_proof));
    return s;
}
// the blanket impl for Vec<I: VecAutoHintable>: [len] followed by every element's stream
template <class T>
inline Stream write_vec(const std::vector<T>& v);
inline Stream write(const std::vector<AdjacentOpened>& v);
inline Stream write(const std::vector<std::vector<EF>>& v);
inline Stream write(const QueryProof& q) {
    Stream s = write_vec(q.input_proof);
    extend(s, write_vec(q.commit_phase_openings));
    return s;
}
inline Stream write(const AirProofData& a) {
    Stream s = write_usize(a.air_id);
    size_t lg = 0;
    while (((size_t)1 << lg) < a.degree) lg++;
    if (((size_t)1 << lg) != a.degree) throw std::runtime_error("hints: degree is not a power of two");
    extend(s, write_usize(lg));
    extend(s, write(a.exposed_values_after_challenge));
    extend(s, write(a.public_values));
    return s;
}
template <class T>
inline Stream write_vec(const std::vector<T>& v) {
    Stream s = write_usize(v.size());
    for (const T& x : v) extend(s, write(x));
    return s;
}
inline Stream write(const std::vector<AdjacentOpened>& v) { return write_vec(v); }
inline Stream write(const std::vector<std::vector<EF>>& v) { return write_vec(v); }
inline Stream write(const std::vector<std::vector<AdjacentOpened>>& v) { return write_vec(v); }
inline Stream write(const std::vector<std::vector<std::vector<EF>>>& v) { return write_vec(v); }

inline Stream write(const FriProof& f) {
    Stream s = write_vec(f.commit_phase_commits);
    extend(s, write_vec(f.query_proofs));
    extend(s, write(f.final_poly));
    s.push_back({f.pow_witness.v});
    return s;
}

inline Stream hint_stream(const Proof& p) {
    Stream s;
    // Commitments
    extend(s, write_vec(p.main_trace_commits));
    extend(s, write_vec(p.after_challenge_commits));
    extend(s, write(p.quotient_commit));
    // OpeningProof { proof, values }
    extend(s, write(p.fri));
    extend(s, write(p.opened_prep));
    extend(s, write(p.opened_main));
    extend(s, write(p.opened_quotient));
    extend(s, write(p.opened_after));
    // per_air, air_perm_by_height, logup witness
    extend(s, write_vec(p.per_air));
    std::vector<size_t> perm(p.per_air.size());
    for (size_t i = 0; i < perm.size(); i++) perm[i] = i;
    std::stable_sort(perm.begin(), perm.end(), [&](size_t a, size_t b) { return p.per_air[a].degree > p.per_air[b].degree; });
    extend(s, write_usizes(perm));
    extend(s, write(p.has_logup_witness ? p.logup_pow_witness : F()));
    return s;
}

}  // namespace orc
