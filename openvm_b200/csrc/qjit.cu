// Quotient evaluation without an interpreter: at key-load time every AIR's constraint DAG is turned into straight-line
// PTX (one thread per row of the quotient domain, every intermediate value a virtual register) and handed to the
// driver's JIT through the runtime's library-loading API (cudaLibraryLoadData accepts PTX text; no NVRTC needed).
//
// Replaces (reference, external): the per-AIR constraint evaluation of openvm-stark-backend's `QuotientCommitter`
// (CPU: a generic `eval` over the symbolic DAG with packed field types; reference GPU backend: an interpreter over a
// rule list). In-tree specification: extensions/native/recursion/src/stark/mod.rs:903-1126 (folding), fri/domain.rs:59-155
// (selectors). The interpreter in prover.cu stays as the path for keys whose JIT is disabled (B2S_QUOTIENT_JIT=0);
// r1e capture: it issues ~45 instructions per DAG operation (dispatch, operand decode, slot traffic), the generated
// code ~7 for a base-field product and 3 for an addition.
//
// Long constraint lists are split into several kernels of bounded size (ptxas time grows faster than linearly in the
// size of a basic block). The kernels of one AIR are INDEPENDENT: kernel k folds its own constraints (with their own powers
// of alpha) into partial plane k; a fixed kernel adds the planes, divides by the zerofier and writes the quotient chunks.
// All quotient kernels of a proof are spread over a few streams between one fork and one join (QuotientBatch): a 128-row
// AIR with 82 kernels is latency-bound per kernel (~30 us each), so they have to overlap (r2h: 252 launches, 8.4 ms serial).
#include "common.cuh"

#include <cstdarg>
#include <functional>
#include <mutex>
#include <unordered_map>

namespace b2 {

namespace {

constexpr uint32_t NONE = 0xFFFFFFFFu;
enum NodeKind { N_CONST = 0, N_PREP, N_MAIN, N_PERM, N_PUBLIC, N_CHALLENGE, N_EXPOSED, N_FIRST, N_LAST, N_TRANS,
                N_ADD, N_SUB, N_NEG, N_MUL };

struct Val {  // a base-field value (one register) or an extension-field value (four)
    uint32_t r[4] = {NONE, NONE, NONE, NONE};
    bool ext = false;
};

struct Emitter {
    std::string body;
    uint32_t nr = 0, nrd = 0;
    size_t lines = 0;
    void line(const char* fmt, ...) {
        char buf[256];
        va_list ap;
        va_start(ap, fmt);
        vsnprintf(buf, sizeof buf, fmt, ap);
        va_end(ap);
        body += buf;
        body += '\n';
        lines++;
    }
    uint32_t r() { return nr++; }
    uint32_t rd() { return nrd++; }
    // ---- base field, Montgomery form, every result in [0, p)
    uint32_t imm(uint32_t v) { uint32_t o = r(); line("mov.u32 %%r%u, %u;", o, v); return o; }
    uint32_t add(uint32_t a, uint32_t b) {
        uint32_t s = r(), t = r(), o = r();
        line("add.u32 %%r%u, %%r%u, %%r%u;", s, a, b);
        line("sub.u32 %%r%u, %%r%u, %u;", t, s, P);
        line("min.u32 %%r%u, %%r%u, %%r%u;", o, s, t);
        return o;
    }
    uint32_t sub(uint32_t a, uint32_t b) {
        uint32_t d = r(), t = r(), o = r();
        line("sub.u32 %%r%u, %%r%u, %%r%u;", d, a, b);
        line("add.u32 %%r%u, %%r%u, %u;", t, d, P);
        line("min.u32 %%r%u, %%r%u, %%r%u;", o, d, t);
        return o;
    }
    uint32_t neg(uint32_t a) {
        uint32_t d = r(), t = r(), o = r();
        line("neg.s32 %%r%u, %%r%u;", d, a);
        line("add.u32 %%r%u, %%r%u, %u;", t, d, P);
        line("min.u32 %%r%u, %%r%u, %%r%u;", o, d, t);
        return o;
    }
    uint32_t mul(uint32_t a, uint32_t b) {
        uint32_t t = rd(), u = rd(), lo = r(), m = r(), hi = r(), s = r(), o = r();
        line("mul.wide.u32 %%rd%u, %%r%u, %%r%u;", t, a, b);
        line("cvt.u32.u64 %%r%u, %%rd%u;", lo, t);
        line("mul.lo.u32 %%r%u, %%r%u, %u;", m, lo, P_NEG_INV);
        line("mad.wide.u32 %%rd%u, %%r%u, %u, %%rd%u;", u, m, P, t);
        line("shr.u64 %%rd%u, %%rd%u, 32;", u, u);
        line("cvt.u32.u64 %%r%u, %%rd%u;", hi, u);
        line("sub.u32 %%r%u, %%r%u, %u;", s, hi, P);
        line("min.u32 %%r%u, %%r%u, %%r%u;", o, hi, s);
        return o;
    }
    // ---- degree-4 extension, x^4 = 11
    Val lift(uint32_t a, uint32_t zero) { Val v; v.ext = true; v.r[0] = a; v.r[1] = v.r[2] = v.r[3] = zero; return v; }
    Val eadd(const Val& a, const Val& b) { Val o; o.ext = true; for (int i = 0; i < 4; i++) o.r[i] = add(a.r[i], b.r[i]); return o; }
    Val esub(const Val& a, const Val& b) { Val o; o.ext = true; for (int i = 0; i < 4; i++) o.r[i] = sub(a.r[i], b.r[i]); return o; }
    Val eneg(const Val& a) { Val o; o.ext = true; for (int i = 0; i < 4; i++) o.r[i] = neg(a.r[i]); return o; }
    Val eadd_f(const Val& a, uint32_t f) { Val o = a; o.r[0] = add(a.r[0], f); return o; }
    Val esub_f(const Val& a, uint32_t f) { Val o = a; o.r[0] = sub(a.r[0], f); return o; }
    Val emul_f(const Val& a, uint32_t f) { Val o; o.ext = true; for (int i = 0; i < 4; i++) o.r[i] = mul(a.r[i], f); return o; }
    Val emul(const Val& a, const Val& b, uint32_t w11) {
        uint32_t pr[4][4];
        for (int i = 0; i < 4; i++) for (int j = 0; j < 4; j++) pr[i][j] = mul(a.r[i], b.r[j]);
        uint32_t t4 = add(add(pr[1][3], pr[2][2]), pr[3][1]);
        uint32_t t5 = add(pr[2][3], pr[3][2]);
        uint32_t t6 = pr[3][3];
        Val o;
        o.ext = true;
        o.r[0] = add(pr[0][0], mul(w11, t4));
        o.r[1] = add(add(pr[0][1], pr[1][0]), mul(w11, t5));
        o.r[2] = add(add(add(pr[0][2], pr[1][1]), pr[2][0]), mul(w11, t6));
        o.r[3] = add(add(add(pr[0][3], pr[1][2]), pr[2][1]), pr[3][0]);
        return o;
    }
};

struct Gen {
    const AirKeyDev& air;
    Emitter em;
    std::vector<Val> memo;        // per node, valid inside the current kernel
    std::vector<char> done;
    std::map<uint64_t, uint32_t> col_memo;  // (matrix, column, offset) -> register
    std::map<uint32_t, uint32_t> pub_memo, const_memo;
    uint32_t zero = NONE, w11 = NONE, sel[3] = {NONE, NONE, NONE};
    Val chal[2], exposed;
    std::vector<uint32_t> part_mat;
    uint32_t after_mat = 0, num_mats = 0;
    uint32_t acc[4];

    explicit Gen(const AirKeyDev& a) : air(a), memo(a.nodes.size()), done(a.nodes.size(), 0) {
        uint32_t m = air.width_prep ? 1u : 0u;
        for (size_t k = 0; k < air.cached_widths.size(); k++) part_mat.push_back(m++);
        if (air.width_common) part_mat.push_back(m++);
        after_mat = m;
        num_mats = m + (air.width_after ? 1u : 0u);
    }
    void reset_kernel() {
        em = Emitter();
        std::fill(done.begin(), done.end(), 0);
        col_memo.clear(); pub_memo.clear(); const_memo.clear();
        zero = w11 = NONE;
        sel[0] = sel[1] = sel[2] = NONE;
        chal[0] = chal[1] = exposed = Val();
    }
    uint32_t get_zero() { if (zero == NONE) zero = em.imm(0); return zero; }
    uint32_t get_w11() { if (w11 == NONE) w11 = em.imm(M_EXT_W); return w11; }
    uint32_t column(uint32_t mat, uint32_t col, uint32_t off) {
        const uint64_t key = ((uint64_t)mat << 40) | ((uint64_t)col << 1) | off;
        auto it = col_memo.find(key);
        if (it != col_memo.end()) return it->second;
        uint32_t a = em.rd(), o = em.r();
        em.line("mad.lo.u64 %%rd%u, %%rdM, %u, %%rdB%u_%u;", a, col, mat, off);
        em.line("ld.global.nc.u32 %%r%u, [%%rd%u];", o, a);
        col_memo[key] = o;
        return o;
    }
    Val ef_load(const char* base, uint32_t word_off) {
        Val v;
        v.ext = true;
        for (int i = 0; i < 4; i++) { v.r[i] = em.r(); em.line("ld.global.nc.u32 %%r%u, [%s+%u];", v.r[i], base, 4 * (word_off + i)); }
        return v;
    }
    Val leaf(uint32_t i) {
        const uint4 nd = air.nodes[i];
        Val v;
        switch (nd.x) {
            case N_CONST: {
                auto it = const_memo.find(nd.y);
                if (it == const_memo.end()) it = const_memo.emplace(nd.y, em.imm(Fp::from_canonical(nd.y).v)).first;
                v.r[0] = it->second;
                break;
            }
            case N_PREP: v.r[0] = column(0, nd.y, nd.z); break;
            case N_MAIN: v.r[0] = column(part_mat[nd.w], nd.y, nd.z); break;
            case N_PERM:
                v.ext = true;
                for (int k = 0; k < 4; k++) v.r[k] = column(after_mat, 4 * nd.y + k, nd.z);
                break;
            case N_PUBLIC: {
                auto it = pub_memo.find(nd.y);
                if (it == pub_memo.end()) {
                    uint32_t o = em.r();
                    em.line("ld.global.nc.u32 %%r%u, [%%rdPub+%u];", o, 4 * nd.y);
                    it = pub_memo.emplace(nd.y, o).first;
                }
                v.r[0] = it->second;
                break;
            }
            case N_CHALLENGE:
                if (chal[nd.y].r[0] == NONE) chal[nd.y] = ef_load("%rdEf", 4 * nd.y);
                v = chal[nd.y];
                break;
            case N_EXPOSED:
                if (exposed.r[0] == NONE) exposed = ef_load("%rdEf", 8);
                v = exposed;
                break;
            default: {  // N_FIRST, N_LAST, N_TRANS: planes 0..2 of the selector table
                const int k = nd.x - N_FIRST;
                if (sel[k] == NONE) { sel[k] = em.r(); em.line("ld.global.nc.u32 %%r%u, [%%rdSel%d];", sel[k], k); }
                v.r[0] = sel[k];
            }
        }
        return v;
    }
    // value of node `root`, computing what the current kernel has not computed yet (explicit stack: DAGs are deep)
    Val eval(uint32_t root) {
        std::vector<std::pair<uint32_t, int>> stack{{root, 0}};
        while (!stack.empty()) {
            const uint32_t i = stack.back().first;
            if (done[i]) { stack.pop_back(); continue; }
            const uint4 nd = air.nodes[i];
            if (nd.x < N_ADD) { memo[i] = leaf(i); done[i] = 1; stack.pop_back(); continue; }
            if (!done[nd.y]) { stack.push_back({nd.y, 0}); continue; }
            if (nd.x != N_NEG && !done[nd.z]) { stack.push_back({nd.z, 0}); continue; }
            const Val a = memo[nd.y], b = nd.x == N_NEG ? Val() : memo[nd.z];
            Val o;
            switch (nd.x) {
                case N_ADD:
                    if (!a.ext && !b.ext) o.r[0] = em.add(a.r[0], b.r[0]);
                    else if (a.ext && b.ext) o = em.eadd(a, b);
                    else o = a.ext ? em.eadd_f(a, b.r[0]) : em.eadd_f(b, a.r[0]);
                    break;
                case N_SUB:
                    if (!a.ext && !b.ext) o.r[0] = em.sub(a.r[0], b.r[0]);
                    else if (a.ext && b.ext) o = em.esub(a, b);
                    else if (a.ext) o = em.esub_f(a, b.r[0]);
                    else o = em.eneg(em.esub_f(b, a.r[0]));
                    break;
                case N_NEG:
                    if (a.ext) o = em.eneg(a); else o.r[0] = em.neg(a.r[0]);
                    break;
                default:
                    if (!a.ext && !b.ext) o.r[0] = em.mul(a.r[0], b.r[0]);
                    else if (a.ext && b.ext) o = em.emul(a, b, get_w11());
                    else o = a.ext ? em.emul_f(a, b.r[0]) : em.emul_f(b, a.r[0]);
            }
            memo[i] = o;
            done[i] = 1;
            stack.pop_back();
        }
        return memo[root];
    }
    // acc += value * alpha^power
    void accumulate(const Val& v, uint32_t power) {
        Val ap = ef_load("%rdAlpha", 4 * power);
        Val t = v.ext ? em.emul(v, ap, get_w11()) : em.emul_f(ap, v.r[0]);
        for (int c = 0; c < 4; c++) acc[c] = em.add(acc[c], t.r[c]);
    }

    std::string prologue(const std::string& name) const {
        std::string s = ".visible .entry " + name + "(\n";
        for (uint32_t m = 0; m < num_mats; m++) s += "  .param .u64 p_mat" + std::to_string(m) + ",\n";
        s += "  .param .u64 p_pub, .param .u64 p_alpha, .param .u64 p_ef, .param .u64 p_sel, .param .u64 p_acc, .param .u64 p_out,\n"
             "  .param .u32 p_n, .param .u32 p_b, .param .u32 p_lq)\n{\n";
        return s;
    }
};

// one kernel covering constraints [k0, k1); first / last decide where the accumulator comes from and goes to
std::string gen_kernel(Gen& g, const std::string& name, uint32_t k0, uint32_t k1, bool first, bool last) {
    g.reset_kernel();
    Emitter& em = g.em;
    const uint32_t C = (uint32_t)g.air.constraint_idx.size();
    // accumulator
    if (first) {
        const uint32_t z = g.get_zero();
        for (int c = 0; c < 4; c++) g.acc[c] = z;
    } else {
        for (int c = 0; c < 4; c++) { g.acc[c] = em.r(); em.line("ld.global.u32 %%r%u, [%%rdAcc%d];", g.acc[c], c); }
    }
    for (uint32_t k = k0; k < k1; k++) g.accumulate(g.eval(g.air.constraint_idx[k]), C - 1 - k);
    if (last) {
        uint32_t iz = em.r();
        em.line("ld.global.nc.u32 %%r%u, [%%rdSel3];", iz);
        for (int c = 0; c < 4; c++) {
            const uint32_t q = em.mul(g.acc[c], iz);
            em.line("st.global.u32 [%%rdOut%d], %%r%u;", c, q);
        }
    } else {
        for (int c = 0; c < 4; c++) em.line("st.global.u32 [%%rdAcc%d], %%r%u;", c, g.acc[c]);
    }
    std::string s = g.prologue(name);
    char buf[512];
    snprintf(buf, sizeof buf, ".reg .b32 %%r<%u>;\n.reg .b64 %%rd<%u>;\n.reg .pred %%p<2>;\n", em.nr + 1, em.nrd + 1);
    s += buf;
    s += ".reg .b32 %n, %b, %lq, %gid, %rows, %lm, %sh, %K, %Kn, %rn, %kq, %qi, %row, %t0, %t1, %t2, %one;\n"
         ".reg .b64 %rdM, %rdPub, %rdAlpha, %rdEf, %rdSel0, %rdSel1, %rdSel2, %rdSel3, %rdT0, %rdT1, %rdRows4;\n"
         ".reg .b64 %rdAcc0, %rdAcc1, %rdAcc2, %rdAcc3, %rdOut0, %rdOut1, %rdOut2, %rdOut3;\n";
    for (uint32_t m = 0; m < g.num_mats; m++) {
        snprintf(buf, sizeof buf, ".reg .b64 %%rdB%u_0, %%rdB%u_1;\n", m, m);
        s += buf;
    }
    s += "ld.param.u32 %n, [p_n];\nld.param.u32 %b, [p_b];\nld.param.u32 %lq, [p_lq];\n"
         "mov.u32 %t0, %ctaid.x;\nmov.u32 %t1, %ntid.x;\nmov.u32 %t2, %tid.x;\nmad.lo.u32 %gid, %t0, %t1, %t2;\n"
         "mov.u32 %one, 1;\nadd.u32 %t0, %n, %lq;\nshl.b32 %rows, %one, %t0;\nsetp.ge.u32 %p0, %gid, %rows;\n@%p0 bra DONE;\n"
         // K = bitrev_lm(r); next row = bitrev_lm(K + 2^b); kq = K >> (b - lq)
         "add.u32 %lm, %n, %b;\nsub.u32 %sh, 32, %lm;\nbrev.b32 %t0, %gid;\nshr.u32 %K, %t0, %sh;\n"
         "shl.b32 %t0, %one, %b;\nadd.u32 %Kn, %K, %t0;\nshl.b32 %t1, %one, %lm;\nsub.u32 %t1, %t1, 1;\nand.b32 %Kn, %Kn, %t1;\n"
         "brev.b32 %t0, %Kn;\nshr.u32 %rn, %t0, %sh;\n"
         "sub.u32 %t0, %b, %lq;\nshr.u32 %kq, %K, %t0;\nshl.b32 %t1, %one, %lq;\nsub.u32 %t1, %t1, 1;\nand.b32 %qi, %kq, %t1;\nshr.u32 %row, %kq, %lq;\n"
         "mov.u64 %rdM, 4;\nshl.b64 %rdM, %rdM, %lm;\n"
         "ld.param.u64 %rdPub, [p_pub];\nld.param.u64 %rdAlpha, [p_alpha];\nld.param.u64 %rdEf, [p_ef];\n"
         // selector planes: is_first, is_last, is_transition, 1/Z_H at [k * rows + r]
         "cvt.u64.u32 %rdRows4, %rows;\nshl.b64 %rdRows4, %rdRows4, 2;\n"
         "ld.param.u64 %rdT0, [p_sel];\nmul.wide.u32 %rdT1, %gid, 4;\nadd.u64 %rdSel0, %rdT0, %rdT1;\n"
         "add.u64 %rdSel1, %rdSel0, %rdRows4;\nadd.u64 %rdSel2, %rdSel1, %rdRows4;\nadd.u64 %rdSel3, %rdSel2, %rdRows4;\n"
         "ld.param.u64 %rdT0, [p_acc];\nadd.u64 %rdAcc0, %rdT0, %rdT1;\n"
         "add.u64 %rdAcc1, %rdAcc0, %rdRows4;\nadd.u64 %rdAcc2, %rdAcc1, %rdRows4;\nadd.u64 %rdAcc3, %rdAcc2, %rdRows4;\n"
         // out[((qi * 4 + c) << n) + row]
         "shl.b32 %t0, %qi, 2;\nshl.b32 %t0, %t0, %n;\nadd.u32 %t0, %t0, %row;\nmul.wide.u32 %rdT1, %t0, 4;\n"
         "ld.param.u64 %rdT0, [p_out];\nadd.u64 %rdOut0, %rdT0, %rdT1;\n"
         "mov.u64 %rdT0, 4;\nshl.b64 %rdT0, %rdT0, %n;\n"
         "add.u64 %rdOut1, %rdOut0, %rdT0;\nadd.u64 %rdOut2, %rdOut1, %rdT0;\nadd.u64 %rdOut3, %rdOut2, %rdT0;\n";
    for (uint32_t m = 0; m < g.num_mats; m++) {
        snprintf(buf, sizeof buf,
                 "ld.param.u64 %%rdT0, [p_mat%u];\nmul.wide.u32 %%rdT1, %%gid, 4;\nadd.u64 %%rdB%u_0, %%rdT0, %%rdT1;\n"
                 "mul.wide.u32 %%rdT1, %%rn, 4;\nadd.u64 %%rdB%u_1, %%rdT0, %%rdT1;\n", m, m, m);
        s += buf;
    }
    s += em.body;
    s += "DONE:\nret;\n}\n";
    return s;
}

std::mutex g_cache_mutex;
struct CachedLib { cudaLibrary_t lib; cudaKernel_t kernel; };
std::unordered_map<std::string, CachedLib> g_cache;  // PTX text -> loaded library (process lifetime)

}  // namespace

bool quotient_jit_enabled() {
    const char* env = getenv("B2S_QUOTIENT_JIT");
    return !(env && env[0] == '0');
}

// PTX modules of the quotient kernels of one AIR, one entry point ("qk") per module so that the modules can be
// compiled on different host threads (host only: no device is touched)
static std::vector<std::string> generate_ptx(const AirKeyDev& a, uint32_t& num_mats) {
    Gen g(a);
    // split the constraint list so that one kernel stays near `limit` PTX lines
    size_t limit = 6000;
    if (const char* env = getenv("B2S_QUOTIENT_JIT_LINES")) limit = strtoul(env, nullptr, 10);
    const uint32_t C = (uint32_t)a.constraint_idx.size();
    // first pass: cut points (generate, measure, cut) -- cheap, it is string building only
    std::vector<uint32_t> cuts{0};
    {
        g.reset_kernel();
        const uint32_t z = g.get_zero();
        for (int c = 0; c < 4; c++) g.acc[c] = z;
        for (uint32_t k = 0; k < C; k++) {
            g.accumulate(g.eval(a.constraint_idx[k]), C - 1 - k);
            if (g.em.lines >= limit && k + 1 < C) {
                cuts.push_back(k + 1);
                g.reset_kernel();
                const uint32_t z2 = g.get_zero();
                for (int c = 0; c < 4; c++) g.acc[c] = z2;
            }
        }
        cuts.push_back(C);
    }
    std::vector<std::string> mods;
    const size_t nk = cuts.size() - 1;
    for (size_t i = 0; i < nk; i++)
        mods.push_back(".version 8.8\n.target sm_100a\n.address_size 64\n\n" +
                       gen_kernel(g, "qk", cuts[i], cuts[i + 1], true, nk == 1));
    num_mats = g.num_mats;
    return mods;
}
std::string quotient_jit_ptx(const AirKeyDev& a) {
    uint32_t nm;
    std::string all;
    if (!a.constraint_idx.empty())
        for (auto& m : generate_ptx(a, nm)) all += m + "\n// ---- next module\n";
    return all;
}

// Generates the PTX modules of one AIR; returns how many there are (each is compiled by quotient_jit_compile).
size_t quotient_jit_prepare(AirKeyDev& a) {
    a.jit_kernels.clear();
    a.jit_ptx.clear();
    if (a.constraint_idx.empty()) return 0;
    a.jit_ptx = generate_ptx(a, a.jit_num_mats);
    a.jit_kernels.assign(a.jit_ptx.size(), nullptr);
    return a.jit_ptx.size();
}

// JIT-compiles module k of `a` (thread-safe: called from several host threads at key load; identical modules of
// different engines / keys share one loaded library for the life of the process)
void quotient_jit_compile(AirKeyDev& a, size_t k) {
    const std::string& ptx = a.jit_ptx[k];
    {
        std::lock_guard<std::mutex> lock(g_cache_mutex);
        auto it = g_cache.find(ptx);
        if (it != g_cache.end()) {
            a.jit_kernels[k] = (void*)it->second.kernel;
            return;
        }
    }
    char log[8192];
    log[0] = 0;
    cudaJitOption opts[3] = {cudaJitErrorLogBuffer, cudaJitErrorLogBufferSizeBytes, cudaJitOptimizationLevel};
    void* vals[3] = {(void*)log, (void*)(uintptr_t)sizeof(log), (void*)(uintptr_t)3};
    cudaLibrary_t lib = nullptr;
    cudaError_t err = cudaLibraryLoadData(&lib, ptx.c_str(), opts, vals, 3, nullptr, nullptr, 0);
    if (err != cudaSuccess) {
        cudaGetLastError();
        fail(B2S_ERR_INTERNAL, std::string("quotient JIT: ") + cudaGetErrorString(err) + ": " + log);
    }
    cudaKernel_t kern = nullptr;
    err = cudaLibraryGetKernel(&kern, lib, "qk");
    if (err != cudaSuccess) {
        cudaGetLastError();
        cudaLibraryUnload(lib);
        fail(B2S_ERR_INTERNAL, std::string("quotient JIT: kernel lookup failed: ") + cudaGetErrorString(err));
    }
    std::lock_guard<std::mutex> lock(g_cache_mutex);
    auto ins = g_cache.emplace(ptx, CachedLib{lib, kern});
    if (!ins.second) cudaLibraryUnload(lib);  // another thread loaded the same text first
    a.jit_kernels[k] = (void*)ins.first->second.kernel;
}

// ---- selector table: is_first, is_last, is_transition, 1/Z_H on the quotient domain (fri/domain.rs:59-75) -------------
__global__ void selectors_kernel(uint32_t* __restrict__ out, uint32_t n, uint32_t b, uint32_t lq, uint32_t gq, uint32_t ginv_m,
                                 const uint32_t* __restrict__ zh, const uint32_t* __restrict__ inv_zh) {
    const uint32_t r = blockIdx.x * blockDim.x + threadIdx.x;
    const uint32_t rows = 1u << (n + lq);
    if (r >= rows) return;
    const uint32_t lm = n + b;
    const uint32_t K = __brev(r) >> (32 - lm);
    const uint32_t kq = K >> (b - lq);
    const Fp x = Fp::raw(M_GENERATOR) * fp_pow(Fp::raw(gq), kq);
    const uint32_t qi = kq & ((1u << lq) - 1);
    const Fp z = Fp::raw(zh[qi]), ginv = Fp::raw(ginv_m);
    out[r] = (z * fp_inv(x - Fp::one())).v;
    out[rows + r] = (z * fp_inv(x - ginv)).v;
    out[2 * (size_t)rows + r] = (x - ginv).v;
    out[3 * (size_t)rows + r] = inv_zh[qi];
}

const uint32_t* quotient_selectors(b2s_engine* e, unsigned n, unsigned lq) {
    const unsigned b = e->params.log_blowup;
    const uint64_t key = ((uint64_t)n << 16) | ((uint64_t)lq << 8) | b;
    auto it = e->selector_cache.find(key);
    if (it != e->selector_cache.end()) return it->second;
    if (e->selector_cache_bytes > ((size_t)1 << 30)) free_selectors(e);
    const uint32_t rows = 1u << (n + lq);
    uint32_t* d = (uint32_t*)e->dmalloc((size_t)rows * 16);
    uint32_t h[32];
    const uint64_t N = (uint64_t)1 << n;
    const Fp gN = fp_pow(Fp::raw(M_GENERATOR), N), wq = two_adic_generator(lq);
    Fp wpow = Fp::one();
    for (uint32_t j = 0; j < (1u << lq); j++) {
        const Fp zh = gN * wpow - Fp::one();
        h[j] = zh.v;
        h[16 + j] = fp_inv(zh).v;
        wpow *= wq;
    }
    uint32_t* dz = (uint32_t*)e->dmalloc(128);
    B2_CUDA(cudaMemcpyAsync(dz, h, 128, cudaMemcpyHostToDevice, e->stream));
    selectors_kernel<<<(rows + 127) / 128, 128, 0, e->stream>>>(d, n, b, lq, two_adic_generator(n + lq).v,
                                                                  fp_inv(two_adic_generator(n)).v, dz, dz + 16);
    B2_CHECK_LAUNCH();
    B2_CUDA(cudaStreamSynchronize(e->stream));  // h is a stack buffer
    e->dfree(dz, 128);
    e->launches[3]++;
    e->selector_cache[key] = d;
    e->selector_cache_bytes += (size_t)rows * 16;
    return d;
}

void free_selectors(b2s_engine* e) {
    for (auto& kv : e->selector_cache) {
        const unsigned n = (unsigned)(kv.first >> 16), lq = (unsigned)((kv.first >> 8) & 0xFF);
        e->dfree(kv.second, ((size_t)16) << (n + lq));
    }
    e->selector_cache.clear();
    e->selector_cache_bytes = 0;
}

// adds the K partial planes of an AIR, divides by the zerofier and scatters into the quotient chunk matrices
// (out[((qi * 4 + c) << n) + row], as the generated single-kernel code does)
__global__ void quotient_finish_kernel(const uint32_t* __restrict__ acc, uint32_t K, const uint32_t* __restrict__ sel,
                                       uint32_t* __restrict__ out, uint32_t n, uint32_t b, uint32_t lq) {
    const uint32_t gid = blockIdx.x * blockDim.x + threadIdx.x;
    const uint32_t rows = 1u << (n + lq);
    if (gid >= rows) return;
    const uint32_t lm = n + b;
    const uint32_t Kb = __brev(gid) >> (32 - lm);
    const uint32_t kq = Kb >> (b - lq);
    const uint32_t qi = kq & ((1u << lq) - 1), row = kq >> lq;
    const Fp iz = Fp::raw(sel[3 * (size_t)rows + gid]);
#pragma unroll
    for (int c = 0; c < 4; c++) {
        Fp sum = Fp::zero();
        for (uint32_t k = 0; k < K; k++) sum += Fp::raw(acc[((size_t)k * 4 + c) * rows + gid]);
        out[(((size_t)qi * 4 + c) << n) + row] = (sum * iz).v;
    }
}

void ensure_aux_streams(b2s_engine* e) {
    if (e->aux_streams.empty()) {
        unsigned ns = 8;
        if (const char* env = getenv("B2S_QUOTIENT_STREAMS")) ns = (unsigned)std::max(1, std::min(32, atoi(env)));
        for (unsigned i = 0; i < ns; i++) {
            cudaStream_t st;
            B2_CUDA(cudaStreamCreateWithFlags(&st, cudaStreamNonBlocking));
            e->aux_streams.push_back(st);
        }
    }
}
// aux stream k starts behind everything queued on the engine's stream so far
cudaStream_t aux_fork(b2s_engine* e, unsigned k) {
    ensure_aux_streams(e);
    cudaStream_t st = e->aux_streams[k % e->aux_streams.size()];
    cudaEvent_t ev = e->get_event();
    B2_CUDA(cudaEventRecord(ev, e->stream));
    B2_CUDA(cudaStreamWaitEvent(st, ev, 0));
    e->event_pool.push_back(ev);
    return st;
}
// the engine's stream continues behind everything queued on the aux streams
void aux_join(b2s_engine* e) {
    for (auto st : e->aux_streams) {
        cudaEvent_t ev = e->get_event();
        B2_CUDA(cudaEventRecord(ev, st));
        B2_CUDA(cudaStreamWaitEvent(e->stream, ev, 0));
        e->event_pool.push_back(ev);
    }
}

void quotient_batch_begin(b2s_engine* e, QuotientBatch& qb) {
    qb.e = e;
    qb.next = 0;
    qb.finishes.clear();
    ensure_aux_streams(e);
}

// Launches the JIT-compiled kernels of `a` on the batch's streams. mats: LDE matrices in the order
// [preprocessed][cached parts][common][after]. acc: K * 4 * 2^(n + lq) words when the AIR has K > 1 kernels.
void quotient_jit_launch(QuotientBatch& qb, const AirKeyDev& a, const uint32_t* const* mats, const uint32_t* publics,
                         const uint32_t* alpha_pows, const uint32_t* ef_consts, uint32_t* acc, uint32_t* out, unsigned n, unsigned lq) {
    b2s_engine* e = qb.e;
    const uint32_t* sel = quotient_selectors(e, n, lq);
    uint32_t un = n, ub = e->params.log_blowup, ulq = lq;
    const uint32_t rows = 1u << (n + lq);
    const size_t K = a.jit_kernels.size();
    std::vector<const uint32_t*> mat_vals(mats, mats + a.jit_num_mats);
    // the side streams this AIR uses start behind everything queued on the engine's stream so far: the committed LDEs, the
    // allocations, and this AIR's own constant uploads (public values, challenges, column tables)
    {
        cudaEvent_t ev = e->get_event();
        B2_CUDA(cudaEventRecord(ev, e->stream));
        const size_t ns = std::min(K, e->aux_streams.size());
        for (size_t j = 0; j < ns; j++) B2_CUDA(cudaStreamWaitEvent(e->aux_streams[(qb.next + j) % e->aux_streams.size()], ev, 0));
        e->event_pool.push_back(ev);
    }
    for (size_t k = 0; k < K; k++) {
        uint32_t* acc_k = acc ? acc + k * 4 * (size_t)rows : nullptr;
        std::vector<void*> args;
        for (auto& m : mat_vals) args.push_back((void*)&m);
        args.push_back((void*)&publics);
        args.push_back((void*)&alpha_pows);
        args.push_back((void*)&ef_consts);
        args.push_back((void*)&sel);
        args.push_back((void*)&acc_k);
        args.push_back((void*)&out);
        args.push_back((void*)&un);
        args.push_back((void*)&ub);
        args.push_back((void*)&ulq);
        cudaStream_t st = e->aux_streams[qb.next++ % e->aux_streams.size()];
        B2_CUDA(cudaLaunchKernel((const void*)a.jit_kernels[k], dim3((rows + 127) / 128), dim3(128), args.data(), 0, st));
        e->launches[3]++;
        e->quotient_launches[0]++;
    }
    if (K > 1) qb.finishes.push_back(QuotientBatch::Finish{acc, (uint32_t)K, sel, out, n, lq});
}

void quotient_batch_end(QuotientBatch& qb) {
    b2s_engine* e = qb.e;
    aux_join(e);
    for (auto& f : qb.finishes) {
        const uint32_t rows = 1u << (f.n + f.lq);
        quotient_finish_kernel<<<(rows + 127) / 128, 128, 0, e->stream>>>(f.acc, f.K, f.sel, f.out, f.n, e->params.log_blowup, f.lq);
        B2_CHECK_LAUNCH();
        e->launches[3]++;
    }
    qb.finishes.clear();
}

}  // namespace b2
