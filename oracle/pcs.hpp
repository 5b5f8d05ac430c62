// ORACLE — TEST INFRASTRUCTURE ONLY (see babybear.hpp header).
//
// TwoAdicFriPcs (absent dependency p3-fri @ 539bbc84; PARITY UNPINNED): commit / open (prover) restated so that
// the in-tree verifier twin accepts it:
//   extensions/native/recursion/src/fri/two_adic_pcs.rs:41-376  (transcript order, reduced openings, queries)
//   extensions/native/recursion/src/fri/mod.rs:33-170           (fold formula, beta^2 roll-in, commit-phase leaves)
// Matrices are column-major (mmcs.hpp MatView).
#pragma once
#include "challenger.hpp"
#include "dft.hpp"
#include "mmcs.hpp"
#include "proof.hpp"
#include <map>
#include <memory>

namespace orc {

struct ColMatrix {
    std::vector<F> d;
    size_t height = 0, width = 0;
    ColMatrix() {}
    ColMatrix(size_t h, size_t w) : d(h * w), height(h), width(w) {}
    F& at(size_t r, size_t c) { return d[c * height + r]; }
    F at(size_t r, size_t c) const { return d[c * height + r]; }
    MatView view() const { return MatView{d.data(), height, width}; }
};
inline unsigned log2u(size_t x) { unsigned l = 0; while (((size_t)1 << l) < x) l++; return l; }

// PcsData: LDEs (bit-reversed rows, on GENERATOR*H) + the MMCS tree over them, plus what open() needs to
// evaluate the committed polynomials out of domain (the original evaluations and their domain shift).
struct Committed {
    std::vector<ColMatrix> evals;   // as committed: evaluations over shift*H_N, natural order
    std::vector<F> shifts;
    std::vector<ColMatrix> ldes;
    MerkleTree tree;
    Digest root() const { return tree.root(); }
};

inline std::shared_ptr<Committed> pcs_commit(std::vector<ColMatrix> mats, std::vector<F> shifts, unsigned log_blowup,
                                             int threads) {
    auto c = std::make_shared<Committed>();
    c->evals = std::move(mats);
    c->shifts = std::move(shifts);
    std::vector<MatView> views;
    for (size_t i = 0; i < c->evals.size(); i++) {
        const ColMatrix& m = c->evals[i];
        unsigned logn = log2u(m.height);
        ColMatrix l(m.height << log_blowup, m.width);
        parallel_for(m.width, threads, [&](size_t lo, size_t hi) {
            for (size_t col = lo; col < hi; col++)
                coset_lde_column(m.d.data() + col * m.height, logn, log_blowup, c->shifts[i], l.d.data() + col * l.height);
        });
        c->ldes.push_back(std::move(l));
    }
    for (auto& l : c->ldes) views.push_back(l.view());
    c->tree.build(views, threads);
    return c;
}

// evaluate every column of matrix i (a polynomial of degree < N given on shift*H_N) at the EF point z
inline std::vector<EF> eval_columns_at(const Committed& c, size_t i, const EF& z) {
    const ColMatrix& m = c.evals[i];
    unsigned logn = log2u(m.height);
    EF u = z * finv(c.shifts[i]);
    std::vector<EF> out(m.width);
    std::vector<F> coef(m.height);
    for (size_t col = 0; col < m.width; col++) {
        for (size_t r = 0; r < m.height; r++) coef[r] = m.at(r, col);
        idft(coef.data(), logn);
        EF acc;
        for (size_t k = m.height; k-- > 0;) acc = acc * u + coef[k];
        out[col] = acc;
    }
    return out;
}

struct OpenRound {
    std::shared_ptr<Committed> data;
    std::vector<std::vector<EF>> points;                  // per matrix
    std::vector<std::vector<std::vector<EF>>> values;     // [mat][point][col], filled by pcs_open
};

// x of LDE row `row` of an LDE of log height h: GENERATOR * w_h^{bitrev_h(row)}  (two_adic_pcs.rs:178-186)
inline F lde_point(unsigned h, size_t row) { return F(GENERATOR) * fpow(two_adic_generator(h), bitrev64(row, h)); }

inline FriProof pcs_open(std::vector<OpenRound>& rounds, Challenger& ch, unsigned log_blowup, unsigned num_queries,
                         unsigned pow_bits, int threads) {
    // (i) out-of-domain values, observed round -> mat -> point -> column
    for (auto& r : rounds) {
        r.values.resize(r.data->evals.size());
        for (size_t m = 0; m < r.data->evals.size(); m++)
            for (auto& z : r.points[m]) r.values[m].push_back(eval_columns_at(*r.data, m, z));
    }
    for (auto& r : rounds)
        for (auto& mat : r.values)
            for (auto& pt : mat)
                for (auto& v : pt) ch.observe_ext(v);
    const EF alpha = ch.sample_ext();
    // (ii) reduced openings per LDE log height; alpha power runs on per height across rounds/mats/points
    std::map<unsigned, std::vector<EF>> ro;
    std::map<unsigned, EF> alpha_pow;
    unsigned log_max = 0;
    for (auto& r : rounds)
        for (size_t m = 0; m < r.data->ldes.size(); m++) {
            const ColMatrix& l = r.data->ldes[m];
            unsigned h = log2u(l.height);
            log_max = std::max(log_max, h);
            if (!ro.count(h)) { ro[h].assign(l.height, EF()); alpha_pow[h] = EF(F(1)); }
            std::vector<EF>& acc = ro[h];
            for (size_t pi = 0; pi < r.points[m].size(); pi++) {
                const EF z = r.points[m][pi];
                const std::vector<EF>& pz = r.values[m][pi];
                const EF ap0 = alpha_pow[h];
                parallel_for(l.height, threads, [&](size_t lo, size_t hi) {
                    for (size_t row = lo; row < hi; row++) {
                        EF n, ap(F(1));
                        for (size_t c = 0; c < l.width; c++) { n += (pz[c] - l.at(row, c)) * ap; ap *= alpha; }
                        acc[row] += n * einv(z - lde_point(h, row)) * ap0;
                    }
                });
                alpha_pow[h] = ap0 * epow(alpha, l.width);
            }
        }
    // (iii) commit phase
    FriProof proof;
    std::vector<EF> folded = ro[log_max];
    std::vector<std::unique_ptr<MerkleTree>> trees;
    std::vector<ColMatrix> leaves;
    leaves.reserve(log_max);
    const F half = finv(F(2));
    for (unsigned lh = log_max; lh > log_blowup; lh--) {
        size_t rows = folded.size() / 2;
        leaves.emplace_back(rows, 8);
        ColMatrix& lm = leaves.back();
        for (size_t p = 0; p < rows; p++)
            for (int k = 0; k < 2; k++)
                for (int c = 0; c < 4; c++) lm.at(p, 4 * k + c) = folded[2 * p + k].c[c];
        trees.emplace_back(new MerkleTree());
        trees.back()->build({lm.view()}, threads);
        proof.commit_phase_commits.push_back(trees.back()->root());
        ch.observe_digest(trees.back()->root());
        const EF beta = ch.sample_ext();
        std::vector<EF> next(rows);
        const F w = two_adic_generator(lh);
        for (size_t p = 0; p < rows; p++) {
            // x0 = w_lh^{bitrev_lh(2p)}, x1 = -x0:  e0 + (beta - x0)(e1 - e0)/(x1 - x0)
            F x0 = fpow(w, bitrev64(2 * p, lh));
            EF e0 = folded[2 * p], e1 = folded[2 * p + 1];
            next[p] = e0 + (beta - x0) * (e1 - e0) * finv(F(0) - x0 - x0);
        }
        (void)half;
        auto it = ro.find(lh - 1);
        if (it != ro.end()) {
            EF b2 = beta * beta;
            for (size_t p = 0; p < rows; p++) next[p] += b2 * it->second[p];
        }
        folded.swap(next);
    }
    for (size_t i = 1; i < folded.size(); i++)
        if (folded[i] != folded[0]) throw std::runtime_error("fri: final polynomial is not constant");
    proof.final_poly = {folded[0]};
    ch.observe_ext(folded[0]);
    proof.pow_witness = ch.grind(pow_bits);
    // (iv) queries
    for (unsigned q = 0; q < num_queries; q++) {
        size_t index = ch.sample_bits(log_max);
        QueryProof qp;
        for (auto& r : rounds) {
            BatchOpening bo;
            size_t idx = index >> (log_max - r.data->tree.log_max_height);
            r.data->tree.open(idx, bo.opened_values, bo.opening_proof);
            qp.input_proof.push_back(std::move(bo));
        }
        for (size_t i = 0; i < trees.size(); i++) {
            size_t pos = index >> i, pair = pos >> 1;
            CommitPhaseStep st;
            for (int c = 0; c < 4; c++) st.sibling_value.c[c] = leaves[i].at(pair, 4 * ((pos & 1) ^ 1) + c);
            std::vector<std::vector<F>> rows_unused;
            trees[i]->open(pair, rows_unused, st.opening_proof);
            qp.commit_phase_openings.push_back(std::move(st));
        }
        proof.query_proofs.push_back(std::move(qp));
    }
    return proof;
}

}  // namespace orc
