// b2s_verify: StarkEngine::verify(&vk, &proof) for BabyBearPoseidon2Config.
//
// Replaces (reference, external): openvm-stark-backend `MultiTraceStarkVerifier::verify` called at
// crates/vm/src/arch/vm.rs:1074,1127 and crates/sdk/src/lib.rs:735-737. The reference verifies on the host in a few
// milliseconds and so does this file: it is sequential transcript work (a few thousand Poseidon2 permutations), no
// device kernel is involved and none would help. Follows the in-tree twin step by step:
//   extensions/native/recursion/src/stark/mod.rs:130-834   transcript, opening rounds
//   extensions/native/recursion/src/stark/mod.rs:903-1126  constraint folding, quotient recombination
//   extensions/native/recursion/src/fri/two_adic_pcs.rs:41-376, fri/mod.rs:33-299  PCS / FRI / MMCS verify_batch
//   extensions/native/recursion/src/fri/domain.rs:59-155   selectors and zerofiers
//   crates/sdk/src/codec.rs:332-642                        proof decoding
#include "common.cuh"
#include "host_challenger.hpp"

#include <algorithm>
#include <array>

namespace b2 {
namespace {

[[noreturn]] void reject(const std::string& why) { fail(B2S_ERR_VERIFY_FAILED, why); }

using Digest = std::array<Fp, 8>;

struct AdjacentOpened { std::vector<Fp4> local, next; };
struct BatchOpening { std::vector<std::vector<Fp>> opened_values; std::vector<Digest> opening_proof; };
struct CommitPhaseStep { Fp4 sibling_value; std::vector<Digest> opening_proof; };
struct QueryProof { std::vector<BatchOpening> input_proof; std::vector<CommitPhaseStep> commit_phase_openings; };
struct AirProofData {
    uint32_t air_id = 0, degree = 0;
    std::vector<std::vector<Fp4>> exposed;
    std::vector<Fp> public_values;
};
struct DecodedProof {
    std::vector<Digest> main_commits, after_commits, fri_commits;
    Digest quotient_commit;
    std::vector<QueryProof> queries;
    std::vector<Fp4> final_poly;
    Fp pow_witness;
    std::vector<AdjacentOpened> opened_prep;
    std::vector<std::vector<AdjacentOpened>> opened_main, opened_after;
    std::vector<std::vector<std::vector<Fp4>>> opened_quotient;
    std::vector<AirProofData> per_air;
    bool has_logup_witness = false;
    Fp logup_witness;
};

// crates/sdk/src/codec.rs:272-320: lengths and field elements are u32 LE, EF = 4 words, digest = 8 words
struct ProofReader {
    const uint8_t* p;
    size_t n, i = 0;
    bool root = false;
    uint32_t u32() {
        if (i + 4 > n) reject("InvalidProofShape: proof truncated");
        uint32_t x = (uint32_t)p[i] | ((uint32_t)p[i + 1] << 8) | ((uint32_t)p[i + 2] << 16) | ((uint32_t)p[i + 3] << 24);
        i += 4;
        return x;
    }
    uint32_t len() {
        uint32_t l = u32();
        if ((size_t)l > n - i) reject("InvalidProofShape: length prefix exceeds the proof");
        return l;
    }
    Fp f() {
        uint32_t x = u32();
        if (x >= P) reject("InvalidProofShape: non-canonical field element");
        return Fp::from_canonical(x);
    }
    Fp4 ef() { Fp4 e; for (int k = 0; k < 4; k++) e.c[k] = f(); return e; }
    // root configuration: a digest is one BN254 scalar, eight limbs of its canonical value (kept as raw words)
    Digest digest() {
        Digest d;
        if (!root) { for (int k = 0; k < 8; k++) d[k] = f(); return d; }
        constexpr uint32_t N[8] = B2_BN_MODULUS;
        for (int k = 0; k < 8; k++) d[k] = Fp::raw(u32());
        bool below = false;
        for (int k = 7; k >= 0 && !below; k--) {
            if (d[k].v > N[k]) break;
            if (d[k].v < N[k]) below = true;
        }
        if (!below) reject("InvalidProofShape: non-canonical digest");
        return d;
    }
    std::vector<Fp4> efs() { std::vector<Fp4> v(len()); for (auto& e : v) e = ef(); return v; }
    std::vector<Fp> fs() { std::vector<Fp> v(len()); for (auto& e : v) e = f(); return v; }
    std::vector<Digest> digests() { std::vector<Digest> v(len()); for (auto& d : v) d = digest(); return v; }
    AdjacentOpened adj() { AdjacentOpened a; a.local = efs(); a.next = efs(); return a; }
    std::vector<AdjacentOpened> adjs() { std::vector<AdjacentOpened> v(len()); for (auto& a : v) a = adj(); return v; }
};

DecodedProof decode(const uint8_t* bytes, size_t n, bool root = false) {
    ProofReader r{bytes, n};
    r.root = root;
    if (r.u32() != 1) reject("InvalidProofShape: codec version");
    DecodedProof p;
    p.main_commits = r.digests();
    p.after_commits = r.digests();
    p.quotient_commit = r.digest();
    p.fri_commits = r.digests();
    p.queries.resize(r.len());
    for (auto& q : p.queries) {
        q.input_proof.resize(r.len());
        for (auto& b : q.input_proof) {
            b.opened_values.resize(r.len());
            for (auto& v : b.opened_values) v = r.fs();
            b.opening_proof = r.digests();
        }
        q.commit_phase_openings.resize(r.len());
        for (auto& s : q.commit_phase_openings) { s.sibling_value = r.ef(); s.opening_proof = r.digests(); }
    }
    p.final_poly = r.efs();
    p.pow_witness = r.f();
    p.opened_prep = r.adjs();
    p.opened_main.resize(r.len());
    for (auto& part : p.opened_main) part = r.adjs();
    p.opened_after.resize(r.len());
    for (auto& ph : p.opened_after) ph = r.adjs();
    p.opened_quotient.resize(r.len());
    for (auto& air : p.opened_quotient) {
        air.resize(r.len());
        for (auto& chunk : air) chunk = r.efs();
    }
    p.per_air.resize(r.len());
    for (auto& a : p.per_air) {
        a.air_id = r.u32();
        a.degree = r.u32();
        a.exposed.resize(r.len());
        for (auto& v : a.exposed) v = r.efs();
        a.public_values = r.fs();
    }
    if (r.i >= r.n || r.p[r.i++] != 0) reject("InvalidProofShape: rap phase kind");
    const uint32_t wv = r.u32();
    if (wv != 0xFFFFFFFFu) {
        if (wv >= P) reject("InvalidProofShape: non-canonical logup witness");
        p.has_logup_witness = true;
        p.logup_witness = Fp::from_canonical(wv);
    }
    if (r.i != r.n) reject("InvalidProofShape: trailing bytes");
    return p;
}

// PaddingFreeSponge<16, 8, 8> and TruncatedPermutation (crates/vm/src/arch/hasher/poseidon2.rs:33-43)
struct HostHasher {
    HostPoseidon2 perm;
    bool root = false;  // MultiField32PaddingFreeSponge<_, _, _, 3, 16, 1> / TruncatedPermutation<_, 2, 1, 3> (bn254.cuh)
    Digest hash_rows(const std::vector<const std::vector<Fp>*>& rows) const {
        if (root) {
            Fr st[3] = {fr_zero(), fr_zero(), fr_zero()};
            uint32_t buf[16];
            int k = 0;
            auto flush = [&]() {
                for (int h = 0; 8 * h < k; h++) {
                    uint32_t w[8] = {0, 0, 0, 0, 0, 0, 0, 0};
                    for (int j = 0; j < 8 && 8 * h + j < k; j++) w[j] = buf[8 * h + j];
                    st[h] = bn_reduce_32(w);
                }
                poseidon2_bn254_permute(st);
                k = 0;
            };
            for (auto* row : rows)
                for (Fp x : *row) {
                    buf[k++] = x.to_canonical();
                    if (k == 16) flush();
                }
            if (k) flush();
            uint32_t w[8];
            fr_to_words(st[0], w);
            Digest d;
            for (int j = 0; j < 8; j++) d[j] = Fp::raw(w[j]);
            return d;
        }
        Fp st[16];
        int k = 0;
        for (auto* row : rows)
            for (Fp x : *row) {
                st[k++] = x;
                if (k == 8) { perm.permute(st); k = 0; }
            }
        if (k) perm.permute(st);
        Digest d;
        for (int j = 0; j < 8; j++) d[j] = st[j];
        return d;
    }
    Digest compress(const Digest& l, const Digest& r) const {
        if (root) {
            uint32_t a[8], b[8], o[8];
            for (int j = 0; j < 8; j++) { a[j] = l[j].v; b[j] = r[j].v; }
            bn_compress(a, b, o);
            Digest d;
            for (int j = 0; j < 8; j++) d[j] = Fp::raw(o[j]);
            return d;
        }
        Fp st[16];
        for (int j = 0; j < 8; j++) { st[j] = l[j]; st[8 + j] = r[j]; }
        perm.permute(st);
        Digest d;
        for (int j = 0; j < 8; j++) d[j] = st[j];
        return d;
    }
};

// Mmcs::verify_batch over matrices of mixed power-of-two heights (fri/mod.rs:185-299; poseidon2/README.md:33-41)
bool verify_batch(const HostHasher& hh, const Digest& commit, const std::vector<unsigned>& log_heights, uint64_t index,
                  const std::vector<std::vector<Fp>>& rows, const std::vector<Digest>& proof) {
    const size_t n = log_heights.size();
    if (rows.size() != n || n == 0) return false;
    unsigned lmax = 0;
    for (unsigned l : log_heights) lmax = std::max(lmax, l);
    if (proof.size() != lmax) return false;
    auto rows_at = [&](unsigned l) {
        std::vector<const std::vector<Fp>*> v;
        for (size_t k = 0; k < n; k++) if (log_heights[k] == l) v.push_back(&rows[k]);
        return v;
    };
    Digest node = hh.hash_rows(rows_at(lmax));
    for (unsigned i = 0; i < lmax; i++) {
        node = ((index >> i) & 1) ? hh.compress(proof[i], node) : hh.compress(node, proof[i]);
        auto inject = rows_at(lmax - 1 - i);
        if (!inject.empty()) node = hh.compress(node, hh.hash_rows(inject));
    }
    for (int j = 0; j < 8; j++) if (node[j] != commit[j]) return false;
    return true;
}

struct VerifyMat {
    unsigned log_n;
    std::vector<Fp4> points;
    std::vector<const std::vector<Fp4>*> values;  // [point] -> columns
};
struct VerifyRound {
    Digest commit;
    std::vector<VerifyMat> mats;
};

Fp4 from_base(Fp x) { return Fp4(x); }
Fp4 monomial_sum(const std::vector<Fp4>& v, size_t off) {  // sum_e v[off + e] * x^e  (stark/mod.rs:1035-1047)
    Fp4 r = Fp4::zero();
    for (int e = 0; e < 4; e++) {
        Fp4 mono = Fp4::zero();
        mono.c[e] = Fp::one();
        r += mono * v[off + e];
    }
    return r;
}

// verify_two_adic_pcs + verify_query (two_adic_pcs.rs:41-376; fri/mod.rs:33-170)
void verify_pcs(const HostHasher& hh, const std::vector<VerifyRound>& rounds, const DecodedProof& proof, unsigned log_max_height,
                const b2s_fri_params& fp, HostChallenger& ch) {
    const unsigned b = fp.log_blowup;
    if (proof.final_poly.size() != 1) reject("InvalidProofShape: final_poly");
    const Fp4 final_poly = proof.final_poly[0];
    for (auto& r : rounds)
        for (auto& m : r.mats)
            for (auto* pt : m.values)
                for (auto& v : *pt) ch.observe_ext(v);
    const Fp4 alpha = ch.sample_ext();
    if (proof.queries.size() != fp.num_queries) reject("InvalidProofShape: query count");
    if (proof.fri_commits.size() != log_max_height) reject("InvalidProofShape: commit phase");
    std::vector<Fp4> betas;
    for (auto& c : proof.fri_commits) {
        ch.observe_digest(c);
        betas.push_back(ch.sample_ext());
    }
    ch.observe_ext(final_poly);
    if (!ch.check_witness(fp.proof_of_work_bits, proof.pow_witness)) reject("InvalidPowWitness");
    const unsigned L = log_max_height + b;
    for (auto& qp : proof.queries) {
        const uint64_t index = ch.sample_bits(L);
        std::vector<Fp4> ro(TWO_ADICITY + 2, Fp4::zero()), alpha_pow(TWO_ADICITY + 2, Fp4::one());
        if (qp.input_proof.size() != rounds.size()) reject("InvalidProofShape: input_proof");
        for (size_t ri = 0; ri < rounds.size(); ri++) {
            const VerifyRound& r = rounds[ri];
            const BatchOpening& bo = qp.input_proof[ri];
            if (bo.opened_values.size() != r.mats.size()) reject("InvalidProofShape: opened_values");
            unsigned log_batch_max = 0;
            std::vector<unsigned> log_heights;
            for (auto& m : r.mats) { log_batch_max = std::max(log_batch_max, m.log_n + b); log_heights.push_back(m.log_n + b); }
            for (size_t mi = 0; mi < r.mats.size(); mi++) {
                const VerifyMat& m = r.mats[mi];
                const unsigned h = m.log_n + b;
                const Fp x = Fp::raw(M_GENERATOR) * fp_pow(two_adic_generator(h), bitrev32((uint32_t)(index >> (L - h)), h));
                const std::vector<Fp>& px = bo.opened_values[mi];
                for (size_t pi = 0; pi < m.points.size(); pi++) {
                    const std::vector<Fp4>& pz = *m.values[pi];
                    if (pz.size() != px.size()) reject("InvalidProofShape: opened row width");
                    Fp4 acc = Fp4::zero(), ap = Fp4::one();
                    for (size_t t = 0; t < pz.size(); t++) { acc += (pz[t] - px[t]) * ap; ap = ap * alpha; }
                    ro[h] += acc * fp4_inv(m.points[pi] - x) * alpha_pow[h];
                    alpha_pow[h] = alpha_pow[h] * ap;
                }
            }
            if (!verify_batch(hh, r.commit, log_heights, index >> (L - log_batch_max), bo.opened_values, bo.opening_proof))
                reject("InvalidOpeningArgument: input batch");
        }
        Fp4 folded = ro[L];
        Fp x = fp_pow(two_adic_generator(L), bitrev32((uint32_t)index, L));
        if (qp.commit_phase_openings.size() != proof.fri_commits.size()) reject("InvalidProofShape: commit_phase_openings");
        for (size_t i = 0; i < proof.fri_commits.size(); i++) {
            const unsigned log_folded = L - (unsigned)i - 1;
            const CommitPhaseStep& st = qp.commit_phase_openings[i];
            const unsigned bit = (index >> i) & 1;
            Fp4 evals[2] = {folded, folded};
            evals[bit ^ 1] = st.sibling_value;
            std::vector<std::vector<Fp>> row(1);
            for (int k = 0; k < 2; k++) for (int c = 0; c < 4; c++) row[0].push_back(evals[k].c[c]);
            if (!verify_batch(hh, proof.fri_commits[i], {log_folded}, index >> (i + 1), row, st.opening_proof))
                reject("InvalidOpeningArgument: commit phase");
            const Fp xs0 = bit ? -x : x, xs1 = bit ? x : -x;
            folded = evals[0] + (evals[1] - evals[0]) * (betas[i] - xs0) * fp_inv(xs1 - xs0);
            x = x * x;
            folded += betas[i] * betas[i] * ro[log_folded];
        }
        if (folded != final_poly) reject("FinalPolyMismatch");
    }
}

}  // namespace

// every node of the symbolic DAG at the out-of-domain point, folded Horner-style over the constraint list
// (stark/mod.rs:1023-1085)
struct ZetaInputs {
    const std::vector<Fp4>* prep[2] = {nullptr, nullptr};
    std::vector<const std::vector<Fp4>*> main[2];
    std::vector<Fp4> perm[2];
    const std::vector<Fp>* publics = nullptr;
    Fp4 challenges[2];
    Fp4 exposed;
    bool has_phase = false;
    Fp4 is_first, is_last, is_transition;
};

static Fp4 fold_constraints(const AirKeyDev& k, const ZetaInputs& in, const Fp4& alpha) {
    const size_t n = k.nodes.size();
    std::vector<Fp4> v(n, Fp4::zero());
    auto bad = [&]() { reject("InvalidProofShape: constraint DAG refers to a value the proof does not carry"); };
    for (size_t i = 0; i < n; i++) {
        const uint4 nd = k.nodes[i];
        switch (nd.x) {
            case 0: v[i] = Fp4(Fp::from_canonical(nd.y % P)); break;
            case 1:
                if (nd.z > 1 || !in.prep[nd.z] || nd.y >= in.prep[nd.z]->size()) bad();
                v[i] = (*in.prep[nd.z])[nd.y];
                break;
            case 2:
                if (nd.z > 1 || nd.w >= in.main[nd.z].size() || nd.y >= in.main[nd.z][nd.w]->size()) bad();
                v[i] = (*in.main[nd.z][nd.w])[nd.y];
                break;
            case 3:
                if (nd.z > 1 || nd.y >= in.perm[nd.z].size()) bad();
                v[i] = in.perm[nd.z][nd.y];
                break;
            case 4:
                if (nd.y >= in.publics->size()) bad();
                v[i] = Fp4((*in.publics)[nd.y]);
                break;
            case 5:
                if (nd.y > 1 || !in.has_phase) bad();
                v[i] = in.challenges[nd.y];
                break;
            case 6:
                if (nd.y != 0 || !in.has_phase) bad();
                v[i] = in.exposed;
                break;
            case 7: v[i] = in.is_first; break;
            case 8: v[i] = in.is_last; break;
            case 9: v[i] = in.is_transition; break;
            case 10: case 11: case 13:
                if (nd.y >= i || nd.z >= i) bad();
                v[i] = nd.x == 10 ? v[nd.y] + v[nd.z] : nd.x == 11 ? v[nd.y] - v[nd.z] : v[nd.y] * v[nd.z];
                break;
            case 12:
                if (nd.y >= i) bad();
                v[i] = -v[nd.y];
                break;
            default: bad();
        }
    }
    Fp4 acc = Fp4::zero();
    for (uint32_t c : k.constraint_idx) acc = acc * alpha + v[c];
    return acc;
}

void verify(const b2s_engine* e, const b2s_pk* pk, const uint8_t* bytes, size_t len) {
    const b2s_fri_params& fp = e->params;
    const unsigned b = fp.log_blowup;
    if (pk->hash_kind != e->hash_kind) reject("InvalidProofShape: the key was loaded under a different hash configuration");
    const bool root = e->hash_kind != 0;
    const DecodedProof proof = decode(bytes, len, root);
    HostHasher hh;
    hh.root = root;
    HostChallenger ch(e->hash_kind);
    const size_t na = proof.per_air.size();
    if (na == 0) reject("InvalidProofShape: no AIRs");
    std::vector<const AirKeyDev*> keys(na);
    std::vector<unsigned> logn(na);
    for (size_t i = 0; i < na; i++) {
        const AirProofData& a = proof.per_air[i];
        if (a.air_id >= pk->airs.size() || (i && a.air_id <= proof.per_air[i - 1].air_id)) reject("InvalidProofShape: air ids");
        const AirKeyDev& k = pk->airs[a.air_id];
        keys[i] = &k;
        if (a.public_values.size() != k.num_pv) reject("InvalidProofShape: public values");
        if (!a.degree || (a.degree & (a.degree - 1))) reject("InvalidProofShape: degree");
        logn[i] = log2_exact(a.degree);
        if (logn[i] + b > TWO_ADICITY) reject("InvalidProofShape: degree too large");
        if (k.width_prep && k.prep_log_height != logn[i]) reject("InvalidProofShape: degree differs from the preprocessed trace");
        if (k.width_after) {
            if (a.exposed.size() != 1 || a.exposed[0].size() != 1) reject("InvalidProofShape: exposed values");
        } else if (!a.exposed.empty()) {
            reject("InvalidProofShape: exposed values without interactions");
        }
    }
    {   // stark/mod.rs:222-228, 836-901: the claimed heights satisfy the key's linear constraints
        std::vector<uint32_t> ids(na);
        for (size_t i = 0; i < na; i++) ids[i] = proof.per_air[i].air_id;
        if (pk->violated_height_constraint(ids, logn, na) >= 0) reject("InvalidProofShape: trace heights violate a linear constraint");
    }
    // transcript prefix (stark/mod.rs:140-149, 231-254)
    ch.observe_digest(pk->pre_hash);
    ch.observe_canonical((uint32_t)na);
    for (auto& a : proof.per_air) ch.observe_canonical(a.air_id);
    for (auto& a : proof.per_air) for (Fp v : a.public_values) ch.observe(v);
    size_t num_cached = 0, num_prep = 0, num_after = 0;
    for (auto* k : keys) {
        if (k->width_prep) { ch.observe_digest(k->prep_commit); num_prep++; }
        num_cached += k->cached_widths.size();
        num_after += k->width_after ? 1 : 0;
    }
    if (proof.main_commits.size() != num_cached + 1) reject("InvalidProofShape: main commits");
    const bool phase = num_after != 0;
    if (phase != !proof.after_commits.empty() || phase != !proof.opened_after.empty() || phase != proof.has_logup_witness ||
        proof.after_commits.size() > 1 || proof.opened_after.size() > 1)
        reject("InvalidProofShape: challenge phase");
    for (auto& c : proof.main_commits) ch.observe_digest(c);
    for (unsigned l : logn) ch.observe_canonical(l);
    Fp4 challenges[2];
    if (phase) {  // stark/mod.rs:307-358, 1129-1160
        if (!ch.check_witness(pk->log_up_pow_bits, proof.logup_witness)) reject("InvalidLogUpPowWitness");
        challenges[0] = ch.sample_ext();
        challenges[1] = ch.sample_ext();
        Fp4 total = Fp4::zero();
        for (size_t i = 0; i < na; i++)
            if (keys[i]->width_after) {
                const Fp4& cs = proof.per_air[i].exposed[0][0];
                ch.observe_ext(cs);
                total += cs;
            }
        if (total != Fp4::zero()) reject("NonZeroCumulativeSum");
        ch.observe_digest(proof.after_commits[0]);
    }
    const Fp4 alpha = ch.sample_ext();
    ch.observe_digest(proof.quotient_commit);
    const Fp4 zeta = ch.sample_ext();

    // opening rounds: [preprocessed per AIR][cached mains][common main][after challenge][quotient]
    if (proof.opened_prep.size() != num_prep || proof.opened_main.size() != num_cached + 1 || proof.opened_quotient.size() != na)
        reject("InvalidProofShape: opened values");
    std::vector<VerifyRound> rounds;
    auto trace_mat = [&](size_t i, const AdjacentOpened& ov, size_t width) {
        if (ov.local.size() != width || ov.next.size() != width) reject("InvalidProofShape: opened width");
        return VerifyMat{logn[i], {zeta, zeta * two_adic_generator(logn[i])}, {&ov.local, &ov.next}};
    };
    auto digest_of = [](const uint32_t* monty8) { Digest d; for (int j = 0; j < 8; j++) d[j] = Fp::raw(monty8[j]); return d; };
    size_t pi = 0, mc = 0, cm = 0, am = 0;
    for (size_t i = 0; i < na; i++)
        if (keys[i]->width_prep) {
            rounds.push_back(VerifyRound{digest_of(keys[i]->prep_commit), {trace_mat(i, proof.opened_prep[pi], keys[i]->width_prep)}});
            pi++;
        }
    VerifyRound common_round;
    const std::vector<AdjacentOpened>& common_vals = proof.opened_main[num_cached];
    unsigned log_max_height = 0;
    for (size_t i = 0; i < na; i++) {
        const AirKeyDev& k = *keys[i];
        log_max_height = std::max(log_max_height, logn[i]);
        for (uint32_t w : k.cached_widths) {
            if (proof.opened_main[mc].size() != 1) reject("InvalidProofShape: cached main");
            rounds.push_back(VerifyRound{proof.main_commits[mc], {trace_mat(i, proof.opened_main[mc][0], w)}});
            mc++;
        }
        if (k.width_common) {
            if (cm >= common_vals.size()) reject("InvalidProofShape: common main");
            common_round.mats.push_back(trace_mat(i, common_vals[cm++], k.width_common));
        }
    }
    if (cm != common_vals.size() || cm == 0) reject("InvalidProofShape: common main count");
    common_round.commit = proof.main_commits[num_cached];
    rounds.push_back(common_round);
    if (phase) {
        VerifyRound ar;
        ar.commit = proof.after_commits[0];
        if (proof.opened_after[0].size() != num_after) reject("InvalidProofShape: after-challenge openings");
        for (size_t i = 0; i < na; i++)
            if (keys[i]->width_after) ar.mats.push_back(trace_mat(i, proof.opened_after[0][am++], (size_t)4 * keys[i]->width_after));
        rounds.push_back(ar);
    }
    VerifyRound qround;
    qround.commit = proof.quotient_commit;
    for (size_t i = 0; i < na; i++) {
        if (proof.opened_quotient[i].size() != ((size_t)1 << keys[i]->log_qd)) reject("InvalidProofShape: quotient chunks");
        for (auto& chunk : proof.opened_quotient[i]) {
            if (chunk.size() != 4) reject("InvalidProofShape: quotient chunk width");
            qround.mats.push_back(VerifyMat{logn[i], {zeta}, {&chunk}});
        }
    }
    rounds.push_back(qround);
    verify_pcs(hh, rounds, proof, log_max_height, fp, ch);

    // constraints at zeta against the recombined quotient (stark/mod.rs:903-1126; fri/domain.rs:59-155)
    pi = mc = cm = am = 0;
    for (size_t i = 0; i < na; i++) {
        const AirKeyDev& k = *keys[i];
        const unsigned n = logn[i];
        ZetaInputs in;
        if (k.width_prep) { in.prep[0] = &proof.opened_prep[pi].local; in.prep[1] = &proof.opened_prep[pi].next; pi++; }
        for (size_t p = 0; p < k.cached_widths.size(); p++) {
            in.main[0].push_back(&proof.opened_main[mc][0].local);
            in.main[1].push_back(&proof.opened_main[mc][0].next);
            mc++;
        }
        if (k.width_common) { in.main[0].push_back(&common_vals[cm].local); in.main[1].push_back(&common_vals[cm].next); cm++; }
        if (k.width_after) {
            const AdjacentOpened& ov = proof.opened_after[0][am++];
            const std::vector<Fp4>* src[2] = {&ov.local, &ov.next};
            for (int o = 0; o < 2; o++)
                for (uint32_t c = 0; c < k.width_after; c++) in.perm[o].push_back(monomial_sum(*src[o], (size_t)4 * c));
            in.has_phase = true;
            in.challenges[0] = challenges[0];
            in.challenges[1] = challenges[1];
            in.exposed = proof.per_air[i].exposed[0][0];
        }
        in.publics = &proof.per_air[i].public_values;
        const Fp ginv = fp_inv(two_adic_generator(n));
        const Fp4 zh = fp4_pow(zeta, (uint64_t)1 << n) - Fp::one();
        in.is_first = zh * fp4_inv(zeta - Fp::one());
        in.is_last = zh * fp4_inv(zeta - ginv);
        in.is_transition = zeta - ginv;
        const Fp4 folded = fold_constraints(k, in, alpha);
        // quotient chunk j lives on the coset GENERATOR * w_{N qd}^j * H_N; recombine with the Lagrange-like zerofier ratios
        const size_t qd = (size_t)1 << k.log_qd;
        const Fp gq = two_adic_generator(n + k.log_qd);
        std::vector<Fp> shifts(qd);
        for (size_t j = 0; j < qd; j++) shifts[j] = Fp::raw(M_GENERATOR) * fp_pow(gq, j);
        auto zp_at = [&](Fp shift, const Fp4& point) { return fp4_pow(point * fp_inv(shift), (uint64_t)1 << n) - Fp::one(); };
        Fp4 quotient = Fp4::zero();
        for (size_t ci = 0; ci < qd; ci++) {
            Fp4 zps = Fp4::one();
            for (size_t j = 0; j < qd; j++)
                if (j != ci) zps = zps * zp_at(shifts[j], zeta) * fp4_inv(zp_at(shifts[j], Fp4(shifts[ci])));
            quotient += zps * monomial_sum(proof.opened_quotient[i][ci], 0);
        }
        if (folded * fp4_inv(zh) != quotient) reject("OodEvaluationMismatch");
    }
}

// ---- proof -> hint stream of the recursive verifier -------------------------------------------------------------------
// The next aggregation layer reads a proof through `hint` instructions; the stream is what `Hintable::write` produces
// for Proof<BabyBearPoseidon2Config> (extensions/native/recursion/src/hints.rs:272-316 and the impls it calls:
// hints.rs:71-266,318-444, fri/hints.rs:22-163; used by crates/sdk/src/prover/agg.rs:307-321): a sequence of vectors of
// field elements. Layout rules followed here, in the reference's words:
//   usize / Val            one vector of one word                      (hints.rs:71-94)
//   Challenge              four vectors of one word                    (hints.rs:97-112)
//   Vec<usize>, Vec<Val>   ONE vector with all words, no length        (hints.rs:156-198)
//   Vec<Challenge>         [len], then one vector of 4 len words       (hints.rs:200-216)
//   Vec<T> otherwise       [len], then every element's own stream      (hints.rs:134-154)
//   digest                 eight vectors of one word                   (fri/hints.rs:22-36)
//   opening proof          [number of digests], then one flat vector   (fri/hints.rs:157-162)
//   batch opening          [total words], one flat vector of the opened rows, then its opening proof (fri/hints.rs:124-150)
// Output: n_vectors, then per vector its length and its canonical words.
namespace {
struct HintWriter {
    std::vector<uint32_t> out;
    uint32_t vectors = 0;
    HintWriter() { out.push_back(0); }
    void one(uint32_t x) { out.push_back(1); out.push_back(x); vectors++; }
    void begin(size_t n) { out.push_back((uint32_t)n); vectors++; }
    void f(Fp x) { one(x.to_canonical()); }
    void ef(const Fp4& e) { for (int c = 0; c < 4; c++) f(e.c[c]); }
    void digest(const Digest& d) { for (int i = 0; i < 8; i++) f(d[i]); }
    void digests(const std::vector<Digest>& v) { one((uint32_t)v.size()); for (auto& d : v) digest(d); }
    void efs(const std::vector<Fp4>& v) {
        one((uint32_t)v.size());
        begin(4 * v.size());
        for (auto& e : v) for (int c = 0; c < 4; c++) out.push_back(e.c[c].to_canonical());
    }
    void opening_proof(const std::vector<Digest>& v) {
        one((uint32_t)v.size());
        begin(8 * v.size());
        for (auto& d : v) for (int i = 0; i < 8; i++) out.push_back(d[i].to_canonical());
    }
    void adj(const AdjacentOpened& a) { efs(a.local); efs(a.next); }
    void adjs(const std::vector<AdjacentOpened>& v) { one((uint32_t)v.size()); for (auto& a : v) adj(a); }
};
}  // namespace

std::vector<uint32_t> proof_hint_stream(const uint8_t* bytes, size_t len) {
    const DecodedProof p = decode(bytes, len);
    HintWriter w;
    // Commitments (hints.rs:417-444)
    w.digests(p.main_commits);
    w.digests(p.after_commits);
    w.digest(p.quotient_commit);
    // OpeningProof = FriProof, then OpenedValues (hints.rs:344-366)
    w.digests(p.fri_commits);                                   // fri/hints.rs:101-122
    w.one((uint32_t)p.queries.size());
    for (auto& q : p.queries) {                                  // fri/hints.rs:64-88
        w.one((uint32_t)q.input_proof.size());
        for (auto& b : q.input_proof) {
            size_t total = 0;
            for (auto& row : b.opened_values) total += row.size();
            w.one((uint32_t)total);
            w.begin(total);
            for (auto& row : b.opened_values) for (Fp x : row) w.out.push_back(x.to_canonical());
            w.opening_proof(b.opening_proof);
        }
        w.one((uint32_t)q.commit_phase_openings.size());
        for (auto& s : q.commit_phase_openings) {                // fri/hints.rs:38-60
            w.ef(s.sibling_value);
            w.opening_proof(s.opening_proof);
        }
    }
    w.efs(p.final_poly);
    w.f(p.pow_witness);
    // OpenedValues: preprocessed, main, QUOTIENT, after_challenge (hints.rs:368-398; not the wire order)
    w.adjs(p.opened_prep);
    w.one((uint32_t)p.opened_main.size());
    for (auto& part : p.opened_main) w.adjs(part);
    w.one((uint32_t)p.opened_quotient.size());
    for (auto& air : p.opened_quotient) {
        w.one((uint32_t)air.size());
        for (auto& chunk : air) w.efs(chunk);
    }
    w.one((uint32_t)p.opened_after.size());
    for (auto& ph : p.opened_after) w.adjs(ph);
    // per_air (hints.rs:318-342): air_id, log2(degree), exposed values, public values as one vector
    w.one((uint32_t)p.per_air.size());
    for (auto& a : p.per_air) {
        if (a.degree == 0 || (a.degree & (a.degree - 1))) reject("InvalidProofShape: degree is not a power of two");
        w.one(a.air_id);
        w.one(log2_exact(a.degree));
        w.one((uint32_t)a.exposed.size());
        for (auto& v : a.exposed) w.efs(v);
        w.begin(a.public_values.size());
        for (Fp x : a.public_values) w.out.push_back(x.to_canonical());
    }
    // air_perm_by_height: indices stably sorted by descending degree, one vector (hints.rs:302-305)
    std::vector<uint32_t> perm(p.per_air.size());
    for (size_t i = 0; i < perm.size(); i++) perm[i] = (uint32_t)i;
    std::stable_sort(perm.begin(), perm.end(), [&](uint32_t x, uint32_t y) { return p.per_air[x].degree > p.per_air[y].degree; });
    w.begin(perm.size());
    for (uint32_t i : perm) w.out.push_back(i);
    // logup proof-of-work witness, 0 when the proof has no challenge phase (hints.rs:306-312)
    w.one(p.has_logup_witness ? p.logup_witness.to_canonical() : 0u);
    w.out[0] = w.vectors;
    return w.out;
}

// Root proof -> witness of the static halo2 verifier (b2s_proof_witness_stream): what Witnessable::write produces for
// Proof<BabyBearPoseidon2RootConfig> (extensions/native/recursion/src/witness.rs:36-293, fri/witness.rs:19-141), i.e. the input
// crates/sdk/src/prover/halo2.rs feeds the wrapper circuit. Three append-only streams: vars (every digest, one scalar of
// eight canonical limbs), felts (opened rows, public values, both proof-of-work witnesses), exts (opened values, FRI
// siblings, final polynomial, exposed sums). Lengths, air ids, degrees and the height permutation are constants of the
// static circuit and write nothing. Layout returned: n_vars, n_felts, n_exts, vars, felts, exts.
std::vector<uint32_t> proof_witness_stream(const uint8_t* bytes, size_t len) {
    const DecodedProof p = decode(bytes, len, true);
    std::vector<uint32_t> vars, felts, exts;
    auto var = [&](const Digest& d) { for (int k = 0; k < 8; k++) vars.push_back(d[k].v); };
    auto felt = [&](Fp x) { felts.push_back(x.to_canonical()); };
    auto ext = [&](const Fp4& e) { for (int k = 0; k < 4; k++) exts.push_back(e.c[k].to_canonical()); };
    auto adjs = [&](const std::vector<AdjacentOpened>& v) {
        for (auto& a : v) {
            for (auto& e : a.local) ext(e);
            for (auto& e : a.next) ext(e);
        }
    };
    // Commitments are written after_challenge, main_trace, quotient (witness.rs:218-240)
    for (auto& c : p.after_commits) var(c);
    for (auto& c : p.main_commits) var(c);
    var(p.quotient_commit);
    // FriProof (fri/witness.rs:61-83)
    for (auto& c : p.fri_commits) var(c);
    for (auto& q : p.queries) {
        for (auto& b : q.input_proof) {
            for (auto& row : b.opened_values) for (Fp x : row) felt(x);
            for (auto& d : b.opening_proof) var(d);
        }
        for (auto& st : q.commit_phase_openings) {
            ext(st.sibling_value);
            for (auto& d : st.opening_proof) var(d);
        }
    }
    for (auto& e : p.final_poly) ext(e);
    felt(p.pow_witness);
    // OpenedValues: preprocessed, main, quotient, after_challenge (witness.rs:250-274)
    adjs(p.opened_prep);
    for (auto& part : p.opened_main) adjs(part);
    for (auto& air : p.opened_quotient) for (auto& chunk : air) for (auto& e : chunk) ext(e);
    for (auto& ph : p.opened_after) adjs(ph);
    for (auto& a : p.per_air) {
        for (auto& v : a.exposed) for (auto& e : v) ext(e);
        for (Fp x : a.public_values) felt(x);
    }
    felts.push_back(p.has_logup_witness ? p.logup_witness.to_canonical() : 0u);
    std::vector<uint32_t> out = {(uint32_t)(vars.size() / 8), (uint32_t)felts.size(), (uint32_t)(exts.size() / 4)};
    out.insert(out.end(), vars.begin(), vars.end());
    out.insert(out.end(), felts.begin(), felts.end());
    out.insert(out.end(), exts.begin(), exts.end());
    return out;
}

}  // namespace b2
