// Level-program graph: the host-side replacement of "issue one tfhe-rs PBS at a time under rayon"
// (reference src/ciphertext/*.rs, every k.k.<op> call site).  String and radix
// operations are written against this builder; every value is a shortint block (plaintext space
// [0,16) + padding bit).  A node is either an INPUT block, or a PBS node
//     out = LUT[ sum_t coef_t * node_t + const ]
// or a LIN node (a linear combination that must exist as a ciphertext because it is an output).
// Linear operations are free: they only build `Lin` expressions that are flattened into the
// prologue of the consuming PBS (K1 gather kernel).  depth(node) = 1 + max depth(sources), so all
// independent LUT evaluations of an operation land in the same level automatically.
#pragma once
#include <array>
#include <cstdio>
#include <cstdint>
#include <algorithm>
#include <map>
#include <stdexcept>
#include <string>
#include <vector>

namespace fhs_host {

using Val = int32_t;
struct Term { Val src; int32_t coef; };

// Linear expression over materialised nodes.  lo/hi bound the plaintext value (message units,
// before reduction mod 32); noise is the variance multiplier in units of one PBS output.
struct Lin {
    std::vector<Term> t;
    int64_t c = 0;
    int32_t lo = 0, hi = 0;
    float noise = 0.f;
    bool is_const() const { return t.empty(); }
};

enum Kind : uint8_t { K_INPUT = 0, K_PBS = 1, K_LIN = 2 };
struct Node {
    Kind kind;
    int32_t lut;       // K_PBS: index into luts
    uint32_t t0, tn;   // range in terms
    int64_t c;         // plaintext constant (message units)
    int32_t level;     // 0 for inputs
    int32_t lo, hi;    // output value bounds
    int32_t in_lo = 0, in_hi = 0;  // K_PBS / K_LIN: bounds the builder claims for the linear combination
    uint8_t signed_in = 0;         // K_PBS: input may be negative (negacyclic LUT use)
};

using Table = std::array<uint64_t, 16>;

class Graph {
public:
    std::vector<Node> nodes;
    std::vector<Term> terms;
    std::vector<Table> luts;
    std::map<Table, int> lut_index;
    std::map<std::vector<int64_t>, Val> cse;   // (kind, table, signedness, constant, sorted terms) -> node
    static std::vector<int64_t> cse_key(Kind kind, int lut_id, bool signed_in, const Lin &e) {
        std::vector<std::pair<Val, int32_t>> ts;
        for (auto &x : e.t) ts.emplace_back(x.src, x.coef);
        std::sort(ts.begin(), ts.end());
        std::vector<int64_t> key{(int64_t)kind, (int64_t)lut_id, signed_in ? 1 : 0, e.c};
        for (auto &x : ts) { key.push_back(x.first); key.push_back(x.second); }
        return key;
    }
    float max_input_noise = 0.f;   // largest variance multiplier fed to any PBS
    int64_t pbs_count = 0;

    Lin input(int hi = 3) {
        nodes.push_back(Node{K_INPUT, -1, 0, 0, 0, 0, 0, hi});
        return leaf((Val)nodes.size() - 1, 0, hi, 1.f);
    }
    static Lin constant(int64_t v) {
        Lin e; e.c = v; e.lo = e.hi = (int32_t)v; return e;
    }
    static Lin leaf(Val v, int lo, int hi, float noise) {
        Lin e; e.t.push_back(Term{v, 1}); e.lo = lo; e.hi = hi; e.noise = noise; return e;
    }
    int lut(const Table &tb) {
        auto it = lut_index.find(tb);
        if (it != lut_index.end()) return it->second;
        luts.push_back(tb);
        lut_index[tb] = (int)luts.size() - 1;
        return (int)luts.size() - 1;
    }
    template <class F> int lut_fn(F f) {
        Table tb;
        for (int x = 0; x < 16; x++) tb[x] = (uint64_t)f(x) & 31u;
        return lut(tb);
    }
    int level_of(const Lin &e) const {
        int l = 0;
        for (auto &x : e.t) l = std::max(l, nodes[x.src].level);
        return l;
    }
    // out = table[e] (e must lie in [0,16), or in (-16,16) when `signed_in`: the padding bit then
    // makes the PBS return -table[e+16] for negative e, tfhe-rs' negacyclic convention).
    template <class F> Lin pbs(const Lin &e, F f, int out_lo, int out_hi, bool signed_in = false) {
        if (!signed_in && (e.lo < 0 || e.hi > 15))
            throw std::logic_error("pbs input out of the 4-bit plaintext space: [" + std::to_string(e.lo) + "," + std::to_string(e.hi) + "]");
        if (signed_in && (e.lo < -15 || e.hi > 15)) throw std::logic_error("signed pbs input out of range");
        if (e.is_const()) {  // plaintext folding
            int64_t x = e.c;
            int64_t v = (x >= 0) ? (int64_t)(f((int)x) & 31) : -(int64_t)(f((int)(x + 16)) & 31);
            return constant(v);
        }
        // a table applied to a*y + c where y is itself the output of a table on some linear combination L is ONE
        // table on L: one level less on that path, and one evaluation less when y has no other reader
        if (!signed_in && e.t.size() == 1 && nodes[e.t[0].src].kind == K_PBS && !nodes[e.t[0].src].signed_in) {
            const Node sn = nodes[e.t[0].src];
            const Table &t1 = luts[sn.lut];
            Table g2;
            for (int v = 0; v < 16; v++) {
                const int64_t u = ((int64_t)e.t[0].coef * (int64_t)t1[v] + e.c) & 31;
                g2[v] = u < 16 ? ((uint64_t)f((int)u) & 31u) : 0u;    // u >= 16: not reachable for an unsigned input
            }
            Lin inner;
            inner.t.assign(terms.begin() + sn.t0, terms.begin() + sn.t0 + sn.tn);
            inner.c = sn.c;
            inner.lo = sn.in_lo;
            inner.hi = sn.in_hi;
            inner.noise = 0.f;                  // already counted when the inner evaluation was made
            return pbs_table(inner, lut(g2), out_lo, out_hi, false);
        }
        return pbs_table(e, lut_fn(f), out_lo, out_hi, signed_in);
    }
    Lin pbs_table(const Lin &e, int id, int out_lo, int out_hi, bool signed_in) {
        // common-subexpression elimination: the same table on the same linear combination is one evaluation
        // (the builders ask again for zero masks, equality bits, ... from several helpers: 35 % of split-256)
        std::vector<int64_t> key = cse_key(K_PBS, id, signed_in, e);
        auto hit = cse.find(key);
        if (hit != cse.end()) {
            Node &old = nodes[hit->second];
            old.lo = std::max(old.lo, (int32_t)out_lo);     // both callers' bounds hold for the one value
            old.hi = std::min(old.hi, (int32_t)out_hi);
            old.in_lo = std::max(old.in_lo, e.lo);
            old.in_hi = std::min(old.in_hi, e.hi);
            return leaf(hit->second, out_lo, out_hi, 1.f);
        }
        cse.emplace(std::move(key), (Val)nodes.size());
        Node n{K_PBS, id, (uint32_t)terms.size(), (uint32_t)e.t.size(), e.c, level_of(e) + 1, out_lo, out_hi};
        n.in_lo = e.lo; n.in_hi = e.hi; n.signed_in = signed_in ? 1 : 0;
        terms.insert(terms.end(), e.t.begin(), e.t.end());
        nodes.push_back(n);
        if (e.noise > max_input_noise) max_input_noise = e.noise;
#ifdef FHS_NOISE_DEBUG
        if (e.noise > 40.f) {
            fprintf(stderr, "noise %.0f terms %zu lvl %d coefs:", e.noise, e.t.size(), level_of(e) + 1);
            for (size_t q = 0; q < e.t.size() && q < 12; q++) fprintf(stderr, " %d", e.t[q].coef);
            fprintf(stderr, " lut:");
            for (int x = 0; x < 16; x++) fprintf(stderr, "%d,", (int)(luts[id][x] & 31));
            fprintf(stderr, "\n");
        }
#endif
        pbs_count++;
        return leaf((Val)nodes.size() - 1, out_lo, out_hi, 1.f);
    }
    // Force `e` to exist as one ciphertext (outputs of an operation).
    Val materialize(const Lin &e) {
        if (e.t.size() == 1 && e.t[0].coef == 1 && e.c == 0) return e.t[0].src;
        std::vector<int64_t> key = cse_key(K_LIN, -1, false, e);
        auto hit = cse.find(key);
        if (hit != cse.end()) return hit->second;
        cse.emplace(std::move(key), (Val)nodes.size());
        Node n{K_LIN, -1, (uint32_t)terms.size(), (uint32_t)e.t.size(), e.c, level_of(e), e.lo, e.hi};
        n.in_lo = e.lo; n.in_hi = e.hi;
        terms.insert(terms.end(), e.t.begin(), e.t.end());
        nodes.push_back(n);
        return (Val)nodes.size() - 1;
    }
    // Drops every node no output depends on (inputs stay: their order is the operand layout) and renumbers;
    // `outputs` holds node ids (>= 0) or encoded constants (< 0) and is rewritten in place.
    size_t prune(std::vector<int64_t> &outputs) {
        const size_t N = nodes.size();
        std::vector<char> live(N, 0);
        for (size_t i = 0; i < N; i++) if (nodes[i].kind == K_INPUT) live[i] = 1;
        for (int64_t o : outputs) if (o >= 0) live[(size_t)o] = 1;
        for (size_t i = N; i-- > 0;) {
            if (!live[i]) continue;
            for (uint32_t t = nodes[i].t0; t < nodes[i].t0 + nodes[i].tn; t++) live[(size_t)terms[t].src] = 1;
        }
        std::vector<Val> renum(N, -1);
        std::vector<Node> nn;
        std::vector<Term> nt;
        size_t dropped = 0;
        for (size_t i = 0; i < N; i++) {
            if (!live[i]) { dropped++; if (nodes[i].kind == K_PBS) pbs_count--; continue; }
            Node n = nodes[i];
            const uint32_t t0 = (uint32_t)nt.size();
            for (uint32_t t = n.t0; t < n.t0 + n.tn; t++) nt.push_back(Term{renum[(size_t)terms[t].src], terms[t].coef});
            n.t0 = t0;
            renum[i] = (Val)nn.size();
            nn.push_back(n);
        }
        nodes.swap(nn);
        terms.swap(nt);
        for (int64_t &o : outputs) if (o >= 0) o = renum[(size_t)o];
        cse.clear();                                   // ids changed; the graph is final at this point
        return dropped;
    }
    int depth() const {
        int d = 0;
        for (auto &n : nodes) d = std::max(d, n.level);
        return d;
    }
};

// ---- linear algebra on Lin (free: no PBS) -------------------------------------------------------
inline Lin scale(const Lin &a, int k) {
    Lin r = a;
    for (auto &x : r.t) x.coef *= k;
    r.c *= k;
    r.lo = std::min(a.lo * k, a.hi * k);
    r.hi = std::max(a.lo * k, a.hi * k);
    r.noise = a.noise * (float)(k * k);
    if (k == 0) { r.t.clear(); r.noise = 0; }
    return r;
}
inline Lin add(const Lin &a, const Lin &b) {
    Lin r = a;
    for (auto &y : b.t) {
        bool merged = false;
        for (auto &x : r.t)
            if (x.src == y.src) { x.coef += y.coef; merged = true; break; }
        if (!merged) r.t.push_back(y);
    }
    size_t w = 0;
    for (size_t i = 0; i < r.t.size(); i++)
        if (r.t[i].coef != 0) r.t[w++] = r.t[i];
    r.t.resize(w);
    r.c += b.c;
    r.lo = a.lo + b.lo;
    r.hi = a.hi + b.hi;
    r.noise = a.noise + b.noise;
    return r;
}
inline Lin sub(const Lin &a, const Lin &b) { return add(a, scale(b, -1)); }
inline Lin addc(const Lin &a, int64_t k) { return add(a, Graph::constant(k)); }
inline Lin bnot(const Lin &a) { Lin r = sub(Graph::constant(1), a); r.lo = 0; r.hi = 1; return r; }
// The caller knows a tighter bound than interval arithmetic gives (e.g. one-hot sums).
inline Lin bounded(Lin a, int lo, int hi) { a.lo = lo; a.hi = hi; return a; }

}  // namespace fhs_host
