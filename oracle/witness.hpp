// ORACLE — TEST INFRASTRUCTURE ONLY (see babybear.hpp header).
//
// Root proof -> witness of the static (halo2) verifier: Witnessable::write for Proof<BabyBearPoseidon2RootConfig>,
// restated from extensions/native/recursion/src/witness.rs:36-293 and fri/witness.rs:19-141. The witness is three
// append-only streams -- vars (BN254 scalars: every digest), felts (BabyBear: opened rows, public values, the two
// proof-of-work witnesses), exts (extension elements: opened values, FRI sibling values, final polynomial, exposed sums);
// lengths, air ids, degrees and the height permutation are constants of the static circuit and write nothing.
// Both sides of this format are in the tree (write and read), so unlike the hashes it is pinned in-tree.
#pragma once
#include "proof.hpp"

namespace orc {

struct Witness {
    std::vector<Digest> vars;   // one scalar each, as the eight canonical limbs a Digest carries under the root configuration
    std::vector<F> felts;
    std::vector<EF> exts;
};

namespace wit {
inline void write(Witness& w, const AdjacentOpened& a);
inline void write(Witness& w, const BatchOpening& b);
inline void write(Witness& w, const CommitPhaseStep& s);
inline void write(Witness& w, const QueryProof& q);
inline void write(Witness& w, const AirProofData& a);
inline void write(Witness& w, const Digest& d) { w.vars.push_back(d); }                         // OuterCom, witness.rs:73-87
inline void write(Witness& w, F x) { w.felts.push_back(x); }                                     // OuterVal, :49-59
inline void write(Witness& w, const EF& e) { w.exts.push_back(e); }                              // OuterChallenge, :61-71
template <class T>
inline void write(Witness& w, const std::vector<T>& v) { for (const T& x : v) write(w, x); }     // Vec<I>, :113-125
inline void write(Witness& w, const AdjacentOpened& a) { write(w, a.local); write(w, a.next); }  // :276-290
inline void write(Witness& w, const BatchOpening& b) {                                           // fri/witness.rs:85-117
    for (auto& row : b.opened_values) write(w, row);
    write(w, b.opening_proof);
}
inline void write(Witness& w, const CommitPhaseStep& s) { write(w, s.sibling_value); write(w, s.opening_proof); }  // :19-38
inline void write(Witness& w, const QueryProof& q) { write(w, q.input_proof); write(w, q.commit_phase_openings); }  // :40-59
inline void write(Witness& w, const FriProof& f) {                                               // :61-83
    write(w, f.commit_phase_commits);
    write(w, f.query_proofs);
    write(w, f.final_poly);
    write(w, f.pow_witness);
}
inline void write(Witness& w, const AirProofData& a) {                                           // witness.rs:192-216
    write(w, a.exposed_values_after_challenge);
    write(w, a.public_values);
}
}  // namespace wit

inline Witness witness_stream(const Proof& p) {                                                  // witness.rs:147-190
    Witness w;
    // Commitments: after_challenge, main_trace, quotient (witness.rs:218-240)
    wit::write(w, p.after_challenge_commits);
    wit::write(w, p.main_trace_commits);
    wit::write(w, p.quotient_commit);
    // OpeningProof: proof, then values in the order preprocessed, main, quotient, after_challenge (:242-274)
    wit::write(w, p.fri);
    wit::write(w, p.opened_prep);
    wit::write(w, p.opened_main);
    wit::write(w, p.opened_quotient);
    wit::write(w, p.opened_after);
    wit::write(w, p.per_air);
    wit::write(w, p.has_logup_witness ? p.logup_pow_witness : F());
    return w;
}

}  // namespace orc
