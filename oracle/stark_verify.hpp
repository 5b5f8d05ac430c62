// ORACLE — TEST INFRASTRUCTURE ONLY (see babybear.hpp header).
//
// MultiTraceStarkVerifier::verify_raps, transliterated from the in-tree recursive verifier
//   extensions/native/recursion/src/stark/mod.rs:130-834 (transcript, rounds), :903-1126 (constraints, quotient)
//   extensions/native/recursion/src/fri/domain.rs:59-155 (selectors, zerofiers, chunk domains)
// including the single FRI-LogUp challenge phase (stark/mod.rs:307-358, 1129-1160).
#pragma once
#include "pcs_verify.hpp"
#include "stark.hpp"

namespace orc {

inline EF ext_selectors_zh(unsigned log_n, F shift, const EF& point) {  // zp_at_point: (point/shift)^N - 1
    return epow(point * finv(shift), (uint64_t)1 << log_n) - F(1);
}

inline void stark_verify(const KeyBlob& key, const Proof& proof, const FriParams& fp) {
    const unsigned b = fp.log_blowup;
    Challenger ch;
    const size_t na = proof.per_air.size();
    std::vector<uint32_t> ids;
    std::vector<std::vector<F>> pvs;
    std::vector<unsigned> logn;
    if (na == 0) throw VerifyError("InvalidProofShape: no AIRs");
    for (auto& a : proof.per_air) {
        if (a.air_id >= key.airs.size() || (!ids.empty() && a.air_id <= ids.back())) throw VerifyError("InvalidProofShape: air ids");
        const AirKey& k = key.airs[a.air_id];
        if (a.public_values.size() != k.num_pv) throw VerifyError("InvalidProofShape: public values");
        if (!a.degree || (a.degree & (a.degree - 1))) throw VerifyError("InvalidProofShape: degree");
        unsigned l = log2u(a.degree);
        if (l + b > TWO_ADICITY) throw VerifyError("InvalidProofShape: degree too large");
        if (k.num_exposed) {
            if (a.exposed_values_after_challenge.size() != 1 || a.exposed_values_after_challenge[0].size() != 1)
                throw VerifyError("InvalidProofShape: exposed values");
        } else if (!a.exposed_values_after_challenge.empty()) {
            throw VerifyError("InvalidProofShape: exposed values without interactions");
        }
        ids.push_back(a.air_id); pvs.push_back(a.public_values); logn.push_back(l);
    }
    // stark/mod.rs:222-228, 836-901: the heights claimed by the proof satisfy the key's linear constraints
    if (key.violated_height_constraint(ids, logn) >= 0) throw VerifyError("InvalidProofShape: trace heights violate a linear constraint");
    observe_prefix(ch, key, ids, pvs);
    size_t num_cached = 0, num_prep = 0;
    for (uint32_t id : ids) { num_cached += key.airs[id].cached_widths.size(); num_prep += key.airs[id].width_prep ? 1 : 0; }
    if (proof.main_trace_commits.size() != num_cached + 1) throw VerifyError("InvalidProofShape: main commits");
    size_t num_after = 0;
    for (uint32_t id : ids) num_after += key.airs[id].width_after ? 1 : 0;
    const bool phase = num_after != 0;
    if (phase != !proof.after_challenge_commits.empty() || phase != !proof.opened_after.empty() || phase != proof.has_logup_witness ||
        proof.after_challenge_commits.size() > 1 || proof.opened_after.size() > 1)
        throw VerifyError("InvalidProofShape: challenge phase");
    for (auto& c : proof.main_trace_commits) ch.observe_digest(c);
    for (unsigned l : logn) ch.observe(F(l));
    EF challenges[2];
    if (phase) {
        if (!ch.check_witness(key.log_up_pow_bits, proof.logup_pow_witness)) throw VerifyError("InvalidLogUpPowWitness");
        challenges[0] = ch.sample_ext();
        challenges[1] = ch.sample_ext();
        EF total;
        for (size_t i = 0; i < na; i++)
            if (key.airs[ids[i]].num_exposed) {
                const EF& cs = proof.per_air[i].exposed_values_after_challenge[0][0];
                ch.observe_ext(cs);
                total += cs;
            }
        if (!total.is_zero()) throw VerifyError("NonZeroCumulativeSum");
        ch.observe_digest(proof.after_challenge_commits[0]);
    }
    const EF alpha = ch.sample_ext();
    ch.observe_digest(proof.quotient_commit);
    const EF zeta = ch.sample_ext();
    // --- rounds
    if (proof.opened_prep.size() != num_prep || proof.opened_main.size() != num_cached + 1 ||
        proof.opened_quotient.size() != na)
        throw VerifyError("InvalidProofShape: opened values");
    std::vector<VerifyRound> rounds;
    auto trace_mat = [&](size_t i, const AdjacentOpened& ov, size_t width) {
        if (ov.local.size() != width || ov.next.size() != width) throw VerifyError("InvalidProofShape: opened width");
        return VerifyMat{logn[i], {zeta, zeta * two_adic_generator(logn[i])}, {ov.local, ov.next}};
    };
    size_t pi = 0;
    for (size_t i = 0; i < na; i++) {
        const AirKey& k = key.airs[ids[i]];
        if (k.width_prep) { rounds.push_back(VerifyRound{k.prep_commit, {trace_mat(i, proof.opened_prep[pi], k.width_prep)}}); pi++; }
    }
    size_t mc = 0;
    VerifyRound common_round;
    const std::vector<AdjacentOpened>& common_vals = proof.opened_main[num_cached];
    size_t cm = 0;
    unsigned log_max_height = 0;
    for (size_t i = 0; i < na; i++) {
        const AirKey& k = key.airs[ids[i]];
        log_max_height = std::max(log_max_height, logn[i]);
        for (uint32_t w : k.cached_widths) {
            if (proof.opened_main[mc].size() != 1) throw VerifyError("InvalidProofShape: cached main");
            rounds.push_back(VerifyRound{proof.main_trace_commits[mc], {trace_mat(i, proof.opened_main[mc][0], w)}});
            mc++;
        }
        if (k.width_common) {
            if (cm >= common_vals.size()) throw VerifyError("InvalidProofShape: common main");
            common_round.mats.push_back(trace_mat(i, common_vals[cm++], k.width_common));
        }
    }
    if (cm != common_vals.size()) throw VerifyError("InvalidProofShape: common main count");
    common_round.commit = proof.main_trace_commits[num_cached];
    rounds.push_back(common_round);
    if (phase) {
        VerifyRound ar;
        ar.commit = proof.after_challenge_commits[0];
        if (proof.opened_after[0].size() != num_after) throw VerifyError("InvalidProofShape: after-challenge openings");
        size_t am = 0;
        for (size_t i = 0; i < na; i++)
            if (key.airs[ids[i]].width_after) { ar.mats.push_back(trace_mat(i, proof.opened_after[0][am], 4 * key.airs[ids[i]].width_after)); am++; }
        rounds.push_back(ar);
    }
    VerifyRound qround;
    qround.commit = proof.quotient_commit;
    for (size_t i = 0; i < na; i++) {
        const AirKey& k = key.airs[ids[i]];
        if (proof.opened_quotient[i].size() != ((size_t)1 << k.log_qd)) throw VerifyError("InvalidProofShape: quotient chunks");
        for (auto& chunk : proof.opened_quotient[i]) {
            if (chunk.size() != 4) throw VerifyError("InvalidProofShape: quotient chunk width");
            qround.mats.push_back(VerifyMat{logn[i], {zeta}, {chunk}});
        }
    }
    rounds.push_back(qround);
    verify_pcs(rounds, proof.fri, log_max_height, b, fp.num_queries, fp.pow_bits, ch);
    // --- constraints at zeta
    pi = 0; mc = 0; cm = 0;
    size_t am = 0;
    for (size_t i = 0; i < na; i++) {
        const AirKey& k = key.airs[ids[i]];
        const unsigned n = logn[i];
        EvalInputs<EF> in;
        if (k.width_prep) { in.prep[0] = proof.opened_prep[pi].local.data(); in.prep[1] = proof.opened_prep[pi].next.data(); pi++; }
        for (size_t p = 0; p < k.cached_widths.size(); p++) {
            in.main[0].push_back(proof.opened_main[mc][0].local.data());
            in.main[1].push_back(proof.opened_main[mc][0].next.data());
            mc++;
        }
        if (k.width_common) { in.main[0].push_back(common_vals[cm].local.data()); in.main[1].push_back(common_vals[cm].next.data()); cm++; }
        std::vector<EF> perm[2];
        if (k.width_after) {
            // unflatten: EF column c = sum_e opened(base column 4c+e) * x^e   (stark/mod.rs:1035-1047)
            const AdjacentOpened& ov = proof.opened_after[0][am++];
            const std::vector<EF>* src[2] = {&ov.local, &ov.next};
            for (int o = 0; o < 2; o++) {
                perm[o].resize(k.width_after);
                for (size_t c = 0; c < k.width_after; c++)
                    for (int e = 0; e < 4; e++) { EF mono; mono.c[e] = F(1); perm[o][c] += mono * (*src[o])[4 * c + e]; }
                in.perm[o] = perm[o].data();
            }
            in.challenges = challenges;
            in.exposed = proof.per_air[i].exposed_values_after_challenge[0].data();
        }
        in.publics = proof.per_air[i].public_values.data();
        // selectors_at_point on the trace domain (shift 1)
        const F ginv = finv(two_adic_generator(n));
        const EF zh = epow(zeta, (uint64_t)1 << n) - F(1);
        in.is_first = zh * einv(zeta - F(1));
        in.is_last = zh * einv(zeta - ginv);
        in.is_transition = zeta - ginv;
        const EF inv_zeroifier = einv(zh);
        const EF folded = eval_dag_fold(k, in, alpha);
        // recompute_quotient: chunk domains = split of GENERATOR * H_{N*qd}
        const size_t qd = (size_t)1 << k.log_qd;
        const F gq = two_adic_generator(n + k.log_qd);
        std::vector<F> shifts(qd);
        for (size_t j = 0; j < qd; j++) shifts[j] = F(GENERATOR) * fpow(gq, j);
        EF quotient;
        for (size_t ci = 0; ci < qd; ci++) {
            EF zps(F(1));
            for (size_t j = 0; j < qd; j++) {
                if (j == ci) continue;
                zps = zps * ext_selectors_zh(n, shifts[j], zeta) * einv(ext_selectors_zh(n, shifts[j], EF(shifts[ci])));
            }
            EF val;
            for (int e = 0; e < 4; e++) {
                EF mono; mono.c[e] = F(1);
                val += mono * proof.opened_quotient[i][ci][e];
            }
            quotient += zps * val;
        }
        if (folded * inv_zeroifier != quotient) throw VerifyError("OodEvaluationMismatch");
    }
}

}  // namespace orc
