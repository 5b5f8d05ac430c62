// ORACLE — TEST INFRASTRUCTURE ONLY (see babybear.hpp header).
//
// StarkEngine::prove (absent dependency openvm-stark-backend @ 002f52e; PARITY UNPINNED), restated so that the
// transcript matches the in-tree verifier twin extensions/native/recursion/src/stark/mod.rs:130-834 step by step
// (SURVEY.md 3.2 / 8g): any number of cached mains, optional preprocessed trace, quotient degree 2^k <= 2^log_blowup,
// and the single FRI-LogUp challenge phase (stark/mod.rs:307-358): PoW, two EF challenges, one reciprocal-sum column
// per interaction chunk plus the running-sum column, exposed cumulative sum. The LogUp denominators follow
//   alpha + sum_j beta^j field_j + beta^len (bus + 1)        (recalled from the absent dependency; the constraints
// that pin it for a given key are part of that key's DAG, so prover and verifier stay consistent by construction).
#pragma once
#include "air.hpp"
#include "pcs.hpp"

namespace orc {

struct AirInput {
    uint32_t air_id = 0;
    std::vector<ColMatrix> cached;   // cached main partitions (each its own commitment)
    ColMatrix common;                // common main (width may be 0)
    std::vector<F> public_values;
};
struct FriParams { unsigned log_blowup = 1, num_queries = 2, pow_bits = 0; };

inline ColMatrix to_col_major(const std::vector<F>& row_major, size_t h, size_t w) {
    ColMatrix m(h, w);
    for (size_t r = 0; r < h; r++) for (size_t c = 0; c < w; c++) m.at(r, c) = row_major[r * w + c];
    return m;
}

// selectors of the trace domain H_N (shift 1) at a base-field point x off the domain (fri/domain.rs:59-75)
struct Selectors { F is_first, is_last, is_transition, inv_zeroifier; };
inline Selectors selectors_at(unsigned log_n, F x) {
    F g = two_adic_generator(log_n), ginv = finv(g);
    F zh = fpow(x, (uint64_t)1 << log_n) - F(1);
    return Selectors{zh * finv(x - F(1)), zh * finv(x - ginv), x - ginv, finv(zh)};
}

inline void observe_prefix(Challenger& ch, const KeyBlob& key, const std::vector<uint32_t>& air_ids,
                           const std::vector<std::vector<F>>& pvs) {
    ch.observe_digest(key.pre_hash);
    ch.observe(F((uint32_t)air_ids.size()));
    for (uint32_t id : air_ids) ch.observe(F(id));
    for (auto& pv : pvs) for (F v : pv) ch.observe(v);
    for (uint32_t id : air_ids) if (key.airs[id].width_prep) ch.observe_digest(key.airs[id].prep_commit);
}

inline Proof stark_prove(const KeyBlob& key, const std::vector<AirInput>& airs, const FriParams& fp, int threads) {
    const unsigned b = fp.log_blowup;
    Proof proof;
    Challenger ch;
    std::vector<uint32_t> ids;
    std::vector<std::vector<F>> pvs;
    std::vector<unsigned> logn;
    for (auto& a : airs) {
        if (a.air_id >= key.airs.size()) throw std::runtime_error("prove: air_id out of range");
        if (!ids.empty() && a.air_id <= ids.back()) throw std::runtime_error("prove: air ids must be strictly increasing");
        const AirKey& k = key.airs[a.air_id];
        if (a.public_values.size() != k.num_pv) throw std::runtime_error("prove: public value count");
        if (a.cached.size() != k.cached_widths.size()) throw std::runtime_error("prove: cached main count");
        size_t h = k.width_common ? a.common.height : (a.cached.empty() ? 0 : a.cached[0].height);
        if (!h || (h & (h - 1))) throw std::runtime_error("prove: trace height must be a power of two");
        if (k.log_qd > b) throw std::runtime_error("prove: quotient degree exceeds blowup");
        ids.push_back(a.air_id); pvs.push_back(a.public_values); logn.push_back(log2u(h));
    }
    if (key.violated_height_constraint(ids, logn) >= 0)   // vm.rs:627-652 (LinearTraceHeightConstraintExceeded)
        throw std::runtime_error("prove: trace heights violate a linear constraint of the key");
    observe_prefix(ch, key, ids, pvs);
    // --- commitments: preprocessed (from the key), cached mains (one tree each), all common mains in one tree
    std::vector<std::shared_ptr<Committed>> prep(airs.size()), cached;
    std::vector<std::pair<size_t, size_t>> cached_of;  // (air position, partition)
    for (size_t i = 0; i < airs.size(); i++) {
        const AirKey& k = key.airs[ids[i]];
        if (k.width_prep) {
            if (k.prep_log_height != logn[i]) throw std::runtime_error("prove: preprocessed height mismatch");
            prep[i] = pcs_commit({to_col_major(k.prep, (size_t)1 << k.prep_log_height, k.width_prep)}, {F(1)}, b, threads);
            if (prep[i]->root() != k.prep_commit) throw std::runtime_error("prove: preprocessed commitment mismatch");
        }
        for (size_t p = 0; p < airs[i].cached.size(); p++) {
            cached.push_back(pcs_commit({airs[i].cached[p]}, {F(1)}, b, threads));
            cached_of.push_back({i, p});
            proof.main_trace_commits.push_back(cached.back()->root());
        }
    }
    std::vector<ColMatrix> commons;
    std::vector<size_t> common_of;  // air position of each common-main matrix
    for (size_t i = 0; i < airs.size(); i++)
        if (key.airs[ids[i]].width_common) { commons.push_back(airs[i].common); common_of.push_back(i); }
    std::shared_ptr<Committed> common;
    if (!commons.empty()) {
        common = pcs_commit(commons, std::vector<F>(commons.size(), F(1)), b, threads);
        proof.main_trace_commits.push_back(common->root());
    } else {
        throw std::runtime_error("prove: no common main trace");
    }
    for (auto& c : proof.main_trace_commits) ch.observe_digest(c);
    for (unsigned l : logn) ch.observe(F(l));
    // --- FRI-LogUp phase
    bool any_inter = false;
    for (uint32_t id : ids) any_inter |= !key.airs[id].interactions.empty();
    EF challenges[2];
    std::vector<EF> cumsum(airs.size());
    std::shared_ptr<Committed> after;
    std::vector<size_t> after_of;  // air position of each after-challenge matrix
    if (any_inter) {
        proof.has_logup_witness = true;
        proof.logup_pow_witness = ch.grind(key.log_up_pow_bits);
        challenges[0] = ch.sample_ext();
        challenges[1] = ch.sample_ext();
        std::vector<ColMatrix> perms;
        for (size_t i = 0; i < airs.size(); i++) {
            const AirKey& k = key.airs[ids[i]];
            if (k.interactions.empty()) continue;
            const size_t N = (size_t)1 << logn[i], nch = k.chunks.size();
            size_t max_len = 0;
            for (auto& it : k.interactions) max_len = std::max(max_len, it.fields.size());
            std::vector<EF> betas(max_len + 1, EF(F(1)));
            for (size_t j = 1; j <= max_len; j++) betas[j] = betas[j - 1] * challenges[1];
            ColMatrix prepm;
            if (k.width_prep) prepm = to_col_major(k.prep, N, k.width_prep);
            ColMatrix pm(N, 4 * (nch + 1));
            std::vector<EF> row_sums(N);
            parallel_for(N, threads, [&](size_t lo, size_t hi) {
                for (size_t r = lo; r < hi; r++) {
                    const size_t rr[2] = {r, (r + 1) % N};
                    std::vector<std::vector<F>> rows[2];
                    std::vector<F> prow[2];
                    EvalInputs<F> in;
                    for (int o = 0; o < 2; o++) {
                        for (size_t p = 0; p < k.num_parts(); p++) {
                            const ColMatrix& m = p < airs[i].cached.size() ? airs[i].cached[p] : airs[i].common;
                            rows[o].emplace_back(m.width);
                            for (size_t c = 0; c < m.width; c++) rows[o][p][c] = m.at(rr[o], c);
                        }
                        for (auto& v : rows[o]) in.main[o].push_back(v.data());
                        if (k.width_prep) {
                            prow[o].resize(k.width_prep);
                            for (size_t c = 0; c < k.width_prep; c++) prow[o][c] = prepm.at(rr[o], c);
                            in.prep[o] = prow[o].data();
                        }
                    }
                    in.publics = airs[i].public_values.data();
                    in.is_first = F(r == 0 ? 1 : 0); in.is_last = F(r == N - 1 ? 1 : 0); in.is_transition = F(r == N - 1 ? 0 : 1);
                    std::vector<EF> v = eval_dag_nodes(k, in);
                    EF row_sum;
                    for (size_t c = 0; c < nch; c++) {
                        EF s;
                        for (uint32_t ii : k.chunks[c]) {
                            const Interaction& it = k.interactions[ii];
                            EF den = challenges[0];
                            for (size_t j = 0; j < it.fields.size(); j++) den += betas[j] * v[it.fields[j]];
                            den += betas[it.fields.size()] * F(it.bus + 1);
                            EF cnt = v[it.count_node];
                            if (it.type) cnt = -cnt;
                            s += cnt * einv(den);
                        }
                        for (int e = 0; e < 4; e++) pm.at(r, 4 * c + e) = s.c[e];
                        row_sum += s;
                    }
                    row_sums[r] = row_sum;
                }
            });
            EF phi;
            for (size_t r = 0; r < N; r++) {
                phi += row_sums[r];
                for (int e = 0; e < 4; e++) pm.at(r, 4 * nch + e) = phi.c[e];
            }
            cumsum[i] = phi;
            perms.push_back(std::move(pm));
            after_of.push_back(i);
        }
        for (size_t i : after_of) ch.observe_ext(cumsum[i]);
        after = pcs_commit(perms, std::vector<F>(perms.size(), F(1)), b, threads);
        proof.after_challenge_commits.push_back(after->root());
        ch.observe_digest(after->root());
    }
    const EF alpha = ch.sample_ext();
    // --- quotient
    std::vector<ColMatrix> chunks;
    std::vector<F> chunk_shifts;
    for (size_t i = 0; i < airs.size(); i++) {
        const AirKey& k = key.airs[ids[i]];
        const unsigned n = logn[i], lq = k.log_qd;
        const size_t N = (size_t)1 << n, Q = N << lq, M = N << b;
        std::vector<const ColMatrix*> parts;
        for (size_t ci = 0; ci < cached.size(); ci++) if (cached_of[ci].first == i) parts.push_back(&cached[ci]->ldes[0]);
        if (k.width_common) {
            size_t pos = std::find(common_of.begin(), common_of.end(), i) - common_of.begin();
            parts.push_back(&common->ldes[pos]);
        }
        const F gq = two_adic_generator(n + lq);
        std::vector<ColMatrix> cm;
        for (size_t j = 0; j < ((size_t)1 << lq); j++) cm.emplace_back(N, 4);
        const ColMatrix* perm_lde = nullptr;
        if (!k.interactions.empty()) {
            size_t pos = std::find(after_of.begin(), after_of.end(), i) - after_of.begin();
            perm_lde = &after->ldes[pos];
        }
        parallel_for(Q, threads, [&](size_t lo, size_t hi) {
            std::vector<std::vector<F>> rows[2];
            std::vector<F> prow[2];
            std::vector<EF> perm_row[2];
            for (size_t kq = lo; kq < hi; kq++) {
                const size_t K = kq << (b - lq);
                const size_t r[2] = {(size_t)bitrev64(K, n + b), (size_t)bitrev64((K + ((size_t)1 << b)) % M, n + b)};
                EvalInputs<F> in;
                for (int o = 0; o < 2; o++) {
                    rows[o].assign(parts.size(), {});
                    for (size_t p = 0; p < parts.size(); p++) {
                        rows[o][p].resize(parts[p]->width);
                        for (size_t c = 0; c < parts[p]->width; c++) rows[o][p][c] = parts[p]->at(r[o], c);
                        in.main[o].push_back(rows[o][p].data());
                    }
                    if (k.width_prep) {
                        prow[o].resize(k.width_prep);
                        for (size_t c = 0; c < k.width_prep; c++) prow[o][c] = prep[i]->ldes[0].at(r[o], c);
                        in.prep[o] = prow[o].data();
                    }
                }
                if (perm_lde) {
                    for (int o = 0; o < 2; o++) {
                        perm_row[o].resize(k.width_after);
                        for (size_t c = 0; c < k.width_after; c++)
                            for (int e = 0; e < 4; e++) perm_row[o][c].c[e] = perm_lde->at(r[o], 4 * c + e);
                        in.perm[o] = perm_row[o].data();
                    }
                    in.challenges = challenges;
                    in.exposed = &cumsum[i];
                }
                in.publics = airs[i].public_values.data();
                const F x = F(GENERATOR) * fpow(gq, kq);
                Selectors s = selectors_at(n, x);
                in.is_first = s.is_first; in.is_last = s.is_last; in.is_transition = s.is_transition;
                EF q = eval_dag_fold(k, in, alpha) * s.inv_zeroifier;
                const size_t chunk = kq & (((size_t)1 << lq) - 1), row = kq >> lq;
                for (int c = 0; c < 4; c++) cm[chunk].at(row, c) = q.c[c];
            }
        });
        for (size_t j = 0; j < cm.size(); j++) { chunks.push_back(std::move(cm[j])); chunk_shifts.push_back(F(GENERATOR) * fpow(gq, j)); }
    }
    auto quotient = pcs_commit(chunks, chunk_shifts, b, threads);
    proof.quotient_commit = quotient->root();
    ch.observe_digest(proof.quotient_commit);
    const EF zeta = ch.sample_ext();
    // --- opening rounds: [prep per AIR][cached mains][common main][quotient]
    std::vector<OpenRound> rounds;
    auto trace_points = [&](size_t i) { return std::vector<EF>{zeta, zeta * two_adic_generator(logn[i])}; };
    for (size_t i = 0; i < airs.size(); i++) if (prep[i]) rounds.push_back(OpenRound{prep[i], {trace_points(i)}, {}});
    const size_t first_cached = rounds.size();
    for (size_t ci = 0; ci < cached.size(); ci++) rounds.push_back(OpenRound{cached[ci], {trace_points(cached_of[ci].first)}, {}});
    {
        OpenRound r{common, {}, {}};
        for (size_t pos : common_of) r.points.push_back(trace_points(pos));
        rounds.push_back(r);
    }
    if (after) {
        OpenRound r{after, {}, {}};
        for (size_t pos : after_of) r.points.push_back(trace_points(pos));
        rounds.push_back(r);
    }
    {
        OpenRound r{quotient, {}, {}};
        for (size_t j = 0; j < chunks.size(); j++) r.points.push_back({zeta});
        rounds.push_back(r);
    }
    proof.fri = pcs_open(rounds, ch, b, fp.num_queries, fp.pow_bits, threads);
    // --- opened values in proof layout
    size_t ri = 0;
    for (size_t i = 0; i < airs.size(); i++) if (prep[i]) { auto& v = rounds[ri++].values[0]; proof.opened_prep.push_back({v[0], v[1]}); }
    for (size_t ci = 0; ci < cached.size(); ci++) { auto& v = rounds[first_cached + ci].values[0]; proof.opened_main.push_back({AdjacentOpened{v[0], v[1]}}); }
    ri = first_cached + cached.size();
    {
        std::vector<AdjacentOpened> part;
        for (auto& v : rounds[ri].values) part.push_back({v[0], v[1]});
        proof.opened_main.push_back(part);
        ri++;
    }
    if (after) {
        std::vector<AdjacentOpened> part;
        for (auto& v : rounds[ri].values) part.push_back({v[0], v[1]});
        proof.opened_after.push_back(part);
        ri++;
    }
    {
        size_t m = 0;
        for (size_t i = 0; i < airs.size(); i++) {
            std::vector<std::vector<EF>> per_air;
            for (size_t j = 0; j < ((size_t)1 << key.airs[ids[i]].log_qd); j++) per_air.push_back(rounds[ri].values[m++][0]);
            proof.opened_quotient.push_back(per_air);
        }
    }
    for (size_t i = 0; i < airs.size(); i++) {
        AirProofData d;
        d.air_id = ids[i]; d.degree = 1u << logn[i]; d.public_values = airs[i].public_values;
        if (!key.airs[ids[i]].interactions.empty()) d.exposed_values_after_challenge = {{cumsum[i]}};
        proof.per_air.push_back(d);
    }
    return proof;
}

}  // namespace orc
