// ORACLE — TEST INFRASTRUCTURE ONLY (see babybear.hpp header).
//
// verify_two_adic_pcs, transliterated from extensions/native/recursion/src/fri/two_adic_pcs.rs:41-376 and
// verify_query from extensions/native/recursion/src/fri/mod.rs:33-170 (verify_batch: mmcs.hpp).
#pragma once
#include "challenger.hpp"
#include "mmcs.hpp"
#include "proof.hpp"

namespace orc {

struct VerifyMat {
    unsigned log_n;                             // log2 of the trace domain size
    std::vector<EF> points;
    std::vector<std::vector<EF>> values;        // [point][col]
};
struct VerifyRound {
    Digest commit;
    std::vector<VerifyMat> mats;                // caller order
};

struct VerifyError : std::runtime_error {
    using std::runtime_error::runtime_error;
};

inline void verify_pcs(const std::vector<VerifyRound>& rounds, const FriProof& proof, unsigned log_max_height,
                       unsigned log_blowup, unsigned num_queries, unsigned pow_bits, Challenger& ch) {
    if (proof.final_poly.size() != 1) throw VerifyError("InvalidProofShape: final_poly");
    const EF final_poly_ct = proof.final_poly[0];
    for (auto& r : rounds)
        for (auto& m : r.mats)
            for (auto& pt : m.values)
                for (auto& v : pt) ch.observe_ext(v);
    const EF alpha = ch.sample_ext();
    if (proof.query_proofs.size() != num_queries) throw VerifyError("InvalidProofShape: query count");
    if (proof.commit_phase_commits.size() != log_max_height) throw VerifyError("InvalidProofShape: commit phase");
    std::vector<EF> betas;
    for (auto& c : proof.commit_phase_commits) { ch.observe_digest(c); betas.push_back(ch.sample_ext()); }
    ch.observe_ext(final_poly_ct);
    if (!ch.check_witness(pow_bits, proof.pow_witness)) throw VerifyError("InvalidPowWitness");
    const unsigned L = log_max_height + log_blowup;
    for (auto& qp : proof.query_proofs) {
        const size_t index = ch.sample_bits(L);
        std::vector<EF> ro(TWO_ADICITY + 2), alpha_pow(TWO_ADICITY + 2, EF(F(1)));
        if (qp.input_proof.size() != rounds.size()) throw VerifyError("InvalidProofShape: input_proof");
        for (size_t ri = 0; ri < rounds.size(); ri++) {
            const VerifyRound& r = rounds[ri];
            const BatchOpening& bo = qp.input_proof[ri];
            if (bo.opened_values.size() != r.mats.size()) throw VerifyError("InvalidProofShape: opened_values");
            unsigned log_batch_max = 0;
            std::vector<size_t> heights;
            for (auto& m : r.mats) { log_batch_max = std::max(log_batch_max, m.log_n + log_blowup); heights.push_back((size_t)1 << (m.log_n + log_blowup)); }
            for (size_t mi = 0; mi < r.mats.size(); mi++) {
                const VerifyMat& m = r.mats[mi];
                const unsigned h = m.log_n + log_blowup;
                // x = g * w_h^{bitrev_h(index >> (L - h))}
                const F x = F(GENERATOR) * fpow(two_adic_generator(h), bitrev64(index >> (L - h), h));
                const std::vector<F>& px = bo.opened_values[mi];
                for (size_t pi = 0; pi < m.points.size(); pi++) {
                    const std::vector<EF>& pz = m.values[pi];
                    if (pz.size() != px.size()) throw VerifyError("InvalidProofShape: opened row width");
                    EF n, ap(F(1));
                    for (size_t t = 0; t < pz.size(); t++) { n += (pz[t] - px[t]) * ap; ap *= alpha; }
                    ro[h] += n * einv(m.points[pi] - x) * alpha_pow[h];
                    alpha_pow[h] = alpha_pow[h] * ap;
                }
            }
            if (!mmcs_verify_batch(r.commit, heights, index >> (L - log_batch_max), bo.opened_values, bo.opening_proof))
                throw VerifyError("InvalidOpeningArgument: input batch");
        }
        // verify_query
        EF folded = ro[L];
        F x = fpow(two_adic_generator(L), bitrev64(index, L));
        if (qp.commit_phase_openings.size() != proof.commit_phase_commits.size())
            throw VerifyError("InvalidProofShape: commit_phase_openings");
        for (size_t i = 0; i < proof.commit_phase_commits.size(); i++) {
            const unsigned log_folded = L - (unsigned)i - 1;
            const CommitPhaseStep& st = qp.commit_phase_openings[i];
            const unsigned bit = (index >> i) & 1;
            EF evals[2] = {folded, folded};
            evals[bit ^ 1] = st.sibling_value;
            std::vector<std::vector<F>> row(1);
            for (int k = 0; k < 2; k++) for (int c = 0; c < 4; c++) row[0].push_back(evals[k].c[c]);
            if (!mmcs_verify_batch(proof.commit_phase_commits[i], {(size_t)1 << log_folded}, index >> (i + 1), row, st.opening_proof))
                throw VerifyError("InvalidOpeningArgument: commit phase");
            // xs[bit] = x, xs[bit^1] = -x
            F xs0 = bit ? -x : x, xs1 = bit ? x : -x;
            folded = evals[0] + (betas[i] - xs0) * (evals[1] - evals[0]) * finv(xs1 - xs0);
            x = x * x;
            folded += betas[i] * betas[i] * ro[log_folded];
        }
        if (folded != final_poly_ct) throw VerifyError("FinalPolyMismatch");
    }
}

}  // namespace orc
