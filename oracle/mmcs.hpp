// ORACLE — TEST INFRASTRUCTURE ONLY (see babybear.hpp header).
//
// MerkleTreeMmcs<.., 8> over matrices of mixed power-of-two heights (hints.rs:39-46).
// In-tree specification followed here:
//   extensions/native/circuit/src/poseidon2/README.md:9-17,33-41  (commit procedure / verify_batch)
//   extensions/native/recursion/src/fri/mod.rs:185-299            (verify_batch in the verifier)
//   extensions/native/recursion/src/fri/two_adic_pcs.rs:24-40     (sorted by height, opened in caller order)
// Matrices are column-major here (F6: GPU DeviceMatrix layout): element (r,c) = data[c*height + r].
#pragma once
#include "poseidon2.hpp"
#include <algorithm>
#include <thread>
#include <functional>

namespace orc {

struct MatView {
    const F* data;   // column-major
    size_t height;   // power of two
    size_t width;
    F at(size_t r, size_t c) const { return data[c * height + r]; }
};

inline void parallel_for(size_t n, int threads, const std::function<void(size_t, size_t)>& fn) {
    if (threads <= 1 || n < 2) { fn(0, n); return; }
    size_t t = std::min<size_t>(threads, n);
    std::vector<std::thread> th;
    size_t chunk = (n + t - 1) / t;
    for (size_t i = 0; i < t; i++) {
        size_t lo = i * chunk, hi = std::min(n, lo + chunk);
        if (lo >= hi) break;
        th.emplace_back(fn, lo, hi);
    }
    for (auto& x : th) x.join();
}

struct MerkleTree {
    std::vector<MatView> mats;            // caller order
    std::vector<size_t> order;            // indices sorted by height descending (stable)
    std::vector<std::vector<Digest>> layers;  // layers[0] = leaves (tallest), last = root (len 1)
    size_t log_max_height = 0;

    Digest root() const { return layers.back()[0]; }

    // digest of row r across all matrices of exactly `height`, in sorted order
    Digest row_hash(size_t height, size_t r) const {
        Sponge s;
        for (size_t k : order) {
            const MatView& m = mats[k];
            if (m.height != height) continue;
            for (size_t c = 0; c < m.width; c++) s.absorb(m.at(r, c));
        }
        return s.finish();
    }
    bool has_height(size_t h) const {
        for (auto& m : mats) if (m.height == h) return true;
        return false;
    }

    void build(const std::vector<MatView>& ms, int threads = 1) {
        mats = ms;
        order.resize(ms.size());
        for (size_t i = 0; i < ms.size(); i++) order[i] = i;
        std::stable_sort(order.begin(), order.end(), [&](size_t a, size_t b) { return ms[a].height > ms[b].height; });
        size_t hmax = ms[order[0]].height;
        log_max_height = 0;
        while (((size_t)1 << log_max_height) < hmax) log_max_height++;
        layers.clear();
        layers.emplace_back(hmax);
        parallel_for(hmax, threads, [&](size_t lo, size_t hi) {
            for (size_t r = lo; r < hi; r++) layers[0][r] = row_hash(hmax, r);
        });
        size_t h = hmax;
        while (h > 1) {
            size_t nh = h / 2;
            std::vector<Digest> next(nh);
            const std::vector<Digest>& prev = layers.back();
            bool inject = has_height(nh);
            parallel_for(nh, threads, [&](size_t lo, size_t hi) {
                for (size_t i = lo; i < hi; i++) {
                    Digest d = compress(prev[2 * i], prev[2 * i + 1]);
                    if (inject) d = compress(d, row_hash(nh, i));
                    next[i] = d;
                }
            });
            layers.push_back(std::move(next));
            h = nh;
        }
    }

    // open_batch(index): rows in caller order, siblings bottom-up
    void open(size_t index, std::vector<std::vector<F>>& rows, std::vector<Digest>& proof) const {
        rows.clear();
        proof.clear();
        for (auto& m : mats) {
            size_t lh = 0;
            while (((size_t)1 << lh) < m.height) lh++;
            size_t r = index >> (log_max_height - lh);
            std::vector<F> row(m.width);
            for (size_t c = 0; c < m.width; c++) row[c] = m.at(r, c);
            rows.push_back(std::move(row));
        }
        for (size_t i = 0; i < log_max_height; i++) proof.push_back(layers[i][(index >> i) ^ 1]);
    }
};

// verify_batch (fri/mod.rs:185-299; poseidon2/README.md:33-41). dims = heights in caller order.
inline bool mmcs_verify_batch(const Digest& commit, const std::vector<size_t>& heights, size_t index,
                              const std::vector<std::vector<F>>& rows, const std::vector<Digest>& proof) {
    size_t n = heights.size();
    if (rows.size() != n || n == 0) return false;
    std::vector<size_t> ord(n);
    for (size_t i = 0; i < n; i++) ord[i] = i;
    std::stable_sort(ord.begin(), ord.end(), [&](size_t a, size_t b) { return heights[a] > heights[b]; });
    size_t hmax = heights[ord[0]];
    size_t logh = 0;
    while (((size_t)1 << logh) < hmax) logh++;
    if (proof.size() != logh) return false;
    auto hash_at = [&](size_t h, bool& any) {
        Sponge s;
        any = false;
        for (size_t k : ord) if (heights[k] == h) { any = true; for (F x : rows[k]) s.absorb(x); }
        return s.finish();
    };
    bool any;
    Digest node = hash_at(hmax, any);
    size_t h = hmax;
    for (size_t i = 0; i < logh; i++) {
        bool bit = (index >> i) & 1;
        node = bit ? compress(proof[i], node) : compress(node, proof[i]);
        h >>= 1;
        Digest inj = hash_at(h, any);
        if (any) node = compress(node, inj);
    }
    return node == commit;
}

}  // namespace orc
