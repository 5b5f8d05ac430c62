// ORACLE — TEST INFRASTRUCTURE ONLY (see babybear.hpp header).
//
// Proof<BabyBearPoseidon2Config> and its wire format, following crates/sdk/src/codec.rs:27,92-121 (Proof),
// :130-168 (OpeningProof / OpenedValues), :170-196 (AirProofData), :198-258 (FriProof / QueryProof),
// :261-270 (logup witness or u32::MAX), :272-320 (EF = 4 words, digest = 8 words, lengths and F as u32 LE).
#pragma once
#include "babybear.hpp"
#include "poseidon2.hpp"
#include <stdexcept>

namespace orc {

struct AdjacentOpened { std::vector<EF> local, next; };
struct BatchOpening { std::vector<std::vector<F>> opened_values; std::vector<Digest> opening_proof; };
struct CommitPhaseStep { EF sibling_value; std::vector<Digest> opening_proof; };
struct QueryProof { std::vector<BatchOpening> input_proof; std::vector<CommitPhaseStep> commit_phase_openings; };
struct FriProof {
    std::vector<Digest> commit_phase_commits;
    std::vector<QueryProof> query_proofs;
    std::vector<EF> final_poly;
    F pow_witness;
};
struct AirProofData {
    uint32_t air_id = 0, degree = 0;
    std::vector<std::vector<EF>> exposed_values_after_challenge;
    std::vector<F> public_values;
};
struct Proof {
    std::vector<Digest> main_trace_commits, after_challenge_commits;
    Digest quotient_commit;
    FriProof fri;
    std::vector<AdjacentOpened> opened_prep;                  // per AIR with a preprocessed trace
    std::vector<std::vector<AdjacentOpened>> opened_main;     // per main commitment, per matrix
    std::vector<std::vector<AdjacentOpened>> opened_after;    // per phase, per matrix
    std::vector<std::vector<std::vector<EF>>> opened_quotient;  // per AIR, per chunk, 4 values
    std::vector<AirProofData> per_air;
    bool has_logup_witness = false;
    F logup_pow_witness;
};

struct Writer {
    std::vector<uint8_t> buf;
    void u32(uint32_t x) { for (int i = 0; i < 4; i++) buf.push_back((uint8_t)(x >> (8 * i))); }
    void f(F x) { u32(x.v); }
    void ef(const EF& e) { for (int i = 0; i < 4; i++) f(e.c[i]); }
    void digest(const Digest& d) { for (int i = 0; i < 8; i++) f(d[i]); }
    void efs(const std::vector<EF>& v) { u32((uint32_t)v.size()); for (auto& e : v) ef(e); }
    void fs(const std::vector<F>& v) { u32((uint32_t)v.size()); for (auto& e : v) f(e); }
    void digests(const std::vector<Digest>& v) { u32((uint32_t)v.size()); for (auto& d : v) digest(d); }
    void adj(const AdjacentOpened& a) { efs(a.local); efs(a.next); }
    void adjs(const std::vector<AdjacentOpened>& v) { u32((uint32_t)v.size()); for (auto& a : v) adj(a); }
};

inline std::vector<uint8_t> encode_proof(const Proof& p) {
    Writer w;
    w.u32(1);  // CODEC_VERSION
    w.digests(p.main_trace_commits);
    w.digests(p.after_challenge_commits);
    w.digest(p.quotient_commit);
    // FriProof
    w.digests(p.fri.commit_phase_commits);
    w.u32((uint32_t)p.fri.query_proofs.size());
    for (auto& q : p.fri.query_proofs) {
        w.u32((uint32_t)q.input_proof.size());
        for (auto& b : q.input_proof) {
            w.u32((uint32_t)b.opened_values.size());
            for (auto& v : b.opened_values) w.fs(v);
            w.digests(b.opening_proof);
        }
        w.u32((uint32_t)q.commit_phase_openings.size());
        for (auto& s : q.commit_phase_openings) { w.ef(s.sibling_value); w.digests(s.opening_proof); }
    }
    w.efs(p.fri.final_poly);
    w.f(p.fri.pow_witness);
    // OpenedValues
    w.adjs(p.opened_prep);
    w.u32((uint32_t)p.opened_main.size());
    for (auto& part : p.opened_main) w.adjs(part);
    w.u32((uint32_t)p.opened_after.size());
    for (auto& ph : p.opened_after) w.adjs(ph);
    w.u32((uint32_t)p.opened_quotient.size());
    for (auto& air : p.opened_quotient) {
        w.u32((uint32_t)air.size());
        for (auto& chunk : air) w.efs(chunk);
    }
    // per_air
    w.u32((uint32_t)p.per_air.size());
    for (auto& a : p.per_air) {
        w.u32(a.air_id);
        w.u32(a.degree);
        w.u32((uint32_t)a.exposed_values_after_challenge.size());
        for (auto& v : a.exposed_values_after_challenge) w.efs(v);
        w.fs(a.public_values);
    }
    w.buf.push_back(0);  // RapPhaseSeqKind::FriLogUp
    if (p.has_logup_witness) w.f(p.logup_pow_witness); else w.u32(0xFFFFFFFFu);
    return w.buf;
}

struct Reader {
    const uint8_t* p; size_t n, i = 0;
    uint32_t u32() {
        if (i + 4 > n) throw std::runtime_error("proof truncated");
        uint32_t x = (uint32_t)p[i] | ((uint32_t)p[i + 1] << 8) | ((uint32_t)p[i + 2] << 16) | ((uint32_t)p[i + 3] << 24);
        i += 4;
        return x;
    }
    uint32_t len() { uint32_t l = u32(); if ((size_t)l > n) throw std::runtime_error("proof: bad length"); return l; }
    F f() { uint32_t x = u32(); if (x >= P) throw std::runtime_error("proof: non-canonical field element"); return F(x); }
    EF ef() { EF e; for (int k = 0; k < 4; k++) e.c[k] = f(); return e; }
    // root configuration: a digest is the canonical scalar in eight limbs (there is no reference wire format for it --
    // crates/sdk/src/codec.rs covers BabyBearPoseidon2Config only; the inner layout is kept, limbs in place of words)
    Digest digest() {
        Digest d;
        if (!hash_kind()) { for (int k = 0; k < 8; k++) d[k] = f(); return d; }
        uint64_t w[4];
        for (int k = 0; k < 8; k++) d[k].v = u32();
        for (int k = 0; k < 4; k++) w[k] = (uint64_t)d[2 * k].v | ((uint64_t)d[2 * k + 1].v << 32);
        if (bn::geq_mod(w)) throw std::runtime_error("proof: non-canonical digest");
        return d;
    }
    std::vector<EF> efs() { std::vector<EF> v(len()); for (auto& e : v) e = ef(); return v; }
    std::vector<F> fs() { std::vector<F> v(len()); for (auto& e : v) e = f(); return v; }
    std::vector<Digest> digests() { std::vector<Digest> v(len()); for (auto& d : v) d = digest(); return v; }
    AdjacentOpened adj() { AdjacentOpened a; a.local = efs(); a.next = efs(); return a; }
    std::vector<AdjacentOpened> adjs() { std::vector<AdjacentOpened> v(len()); for (auto& a : v) a = adj(); return v; }
};

inline Proof decode_proof(const uint8_t* bytes, size_t n) {
    Reader r{bytes, n};
    if (r.u32() != 1) throw std::runtime_error("proof: codec version");
    Proof p;
    p.main_trace_commits = r.digests();
    p.after_challenge_commits = r.digests();
    p.quotient_commit = r.digest();
    p.fri.commit_phase_commits = r.digests();
    p.fri.query_proofs.resize(r.len());
    for (auto& q : p.fri.query_proofs) {
        q.input_proof.resize(r.len());
        for (auto& b : q.input_proof) {
            b.opened_values.resize(r.len());
            for (auto& v : b.opened_values) v = r.fs();
            b.opening_proof = r.digests();
        }
        q.commit_phase_openings.resize(r.len());
        for (auto& s : q.commit_phase_openings) { s.sibling_value = r.ef(); s.opening_proof = r.digests(); }
    }
    p.fri.final_poly = r.efs();
    p.fri.pow_witness = r.f();
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
        a.exposed_values_after_challenge.resize(r.len());
        for (auto& v : a.exposed_values_after_challenge) v = r.efs();
        a.public_values = r.fs();
    }
    if (r.i >= r.n || r.p[r.i++] != 0) throw std::runtime_error("proof: rap phase kind");
    uint32_t wv = r.u32();
    if (wv != 0xFFFFFFFFu) { p.has_logup_witness = true; p.logup_pow_witness = F(wv); }
    if (r.i != r.n) throw std::runtime_error("proof: trailing bytes");
    return p;
}

}  // namespace orc
