// Radix-integer and boolean primitives as level programs (replaces the tfhe::integer::ServerKey
// methods fhe_string calls: SURVEY.md 2.2 -- boolean_bitand/or/not, scalar_eq/ne/gt/ge/le,
// eq/lt/ge, if_then_else, add/sub/scalar_add, sum of selected values, ...).  The algorithms are
// free (only decrypted results must match); these are chosen for low depth and
// few LUT evaluations at message_modulus = carry_modulus = 4:
//   * two clean blocks pack into one 4-bit value 4*hi+lo;
//   * two packed values are compared with ONE signed PBS on their difference (padding bit /
//     negacyclic trick, SURVEY.md 7 "negacyclic semantics are relied upon");
//   * selections use 4*cond + block <= 7 in one PBS per block.
#pragma once
#include <algorithm>
#include <functional>

#include "graph.hpp"

namespace fhs_host {

using Bool = Lin;                // value in {0,1}
using Rad = std::vector<Lin>;    // little-endian base-4 digits, each in [0,3]

// Largest variance multiplier (units of one PBS output) a one-hot digit sum may reach before its groups
// are refreshed (round 1: 160).  64 keeps the per-evaluation failure probability of those PBS at 2^-40.0
// against 2^-40.3 for a single fresh input (DESIGN.md 4 "Noise budget") and still fits the index sum of a
// 65-char string into 10 groups, i.e. no extra level.  The largest multipliers left in any program are the
// fan-in-4 sign reduction (85) and the 8-candidate selections of the chunked match scan (129 .. 162).
constexpr float kMaxLinNoise = 64.f;

struct Ops {
    Graph &g;
    explicit Ops(Graph &g_) : g(g_) {}

    // ---------------- booleans ----------------
    Bool band(const Bool &a, const Bool &b) {
        return g.pbs(add(a, b), [](int x) { return x == 2; }, 0, 1);
    }
    Bool bor(const Bool &a, const Bool &b) {
        return g.pbs(add(a, b), [](int x) { return x >= 1; }, 0, 1);
    }
    // AND / OR of any number of booleans: sums of <= 15 terms, one PBS per chunk, log depth.
    Bool all(std::vector<Bool> v) {
        if (v.empty()) return Graph::constant(1);
        while (v.size() > 1) {
            std::vector<Bool> nx;
            for (size_t i = 0; i < v.size(); i += 15) {
                size_t n = std::min<size_t>(15, v.size() - i);
                Lin s = Graph::constant(0);
                for (size_t j = 0; j < n; j++) s = add(s, v[i + j]);
                int nn = (int)n;
                nx.push_back(n == 1 ? v[i] : g.pbs(s, [nn](int x) { return x == nn; }, 0, 1));
            }
            v.swap(nx);
        }
        return v[0];
    }
    Bool any(std::vector<Bool> v) {
        if (v.empty()) return Graph::constant(0);
        while (v.size() > 1) {
            std::vector<Bool> nx;
            for (size_t i = 0; i < v.size(); i += 15) {
                size_t n = std::min<size_t>(15, v.size() - i);
                Lin s = Graph::constant(0);
                for (size_t j = 0; j < n; j++) s = add(s, v[i + j]);
                nx.push_back(n == 1 ? v[i] : g.pbs(s, [](int x) { return x >= 1; }, 0, 1));
            }
            v.swap(nx);
        }
        return v[0];
    }
    // Inclusive prefix OR: p[i] = v[0] | ... | v[i]; two/three-level blocked scan (chunks of 15).
    std::vector<Bool> prefix_or(const std::vector<Bool> &v) {
        const size_t n = v.size();
        std::vector<Bool> p(n);
        if (n == 0) return p;
        const size_t C = 14;  // 14 bools + 1 carry <= 15
        const size_t nch = (n + C - 1) / C;
        std::vector<Bool> tot(nch);
        std::vector<Lin> run(n);
        for (size_t ch = 0; ch < nch; ch++) {
            Lin s = Graph::constant(0);
            for (size_t i = ch * C; i < std::min(n, (ch + 1) * C); i++) { s = add(s, v[i]); run[i] = s; }
            tot[ch] = s;  // raw sum (<= 15)
        }
        if (nch > 1 && nch <= 16) {
            // two levels: chunk-local prefix bits, then OR with the COUNT of non-empty earlier chunks
            // (the chunk's own last prefix bit is its "non-empty" flag; the count stays below 16)
            std::vector<Bool> local(n);
            for (size_t i = 0; i < n; i++) local[i] = g.pbs(run[i], [](int x) { return x >= 1; }, 0, 1);
            Lin before = Graph::constant(0);
            for (size_t ch = 0; ch < nch; ch++) {
                const size_t lo = ch * C, hi = std::min(n, (ch + 1) * C);
                for (size_t i = lo; i < hi; i++)
                    p[i] = ch == 0 ? local[i] : g.pbs(add(local[i], before), [](int x) { return x >= 1; }, 0, 1);
                before = add(before, local[hi - 1]);
            }
            return p;
        }
        std::vector<Bool> carry;  // OR of all chunks strictly before ch
        if (nch > 1) {
            std::vector<Bool> tb(nch - 1);
            for (size_t ch = 0; ch + 1 < nch; ch++) tb[ch] = g.pbs(tot[ch], [](int x) { return x >= 1; }, 0, 1);
            std::vector<Bool> pc = prefix_or(tb);
            carry.push_back(Graph::constant(0));
            for (size_t ch = 0; ch + 1 < nch; ch++) carry.push_back(pc[ch]);
        } else {
            carry.push_back(Graph::constant(0));
        }
        for (size_t i = 0; i < n; i++) {
            Lin s = add(run[i], carry[i / C]);
            p[i] = g.pbs(s, [](int x) { return x >= 1; }, 0, 1);
        }
        return p;
    }

    // ---------------- packing ----------------
    // pairs of clean blocks -> 4-bit values (last one may be a single block)
    std::vector<Lin> pack(const Rad &a) {
        std::vector<Lin> p;
        for (size_t i = 0; i < a.size(); i += 2)
            p.push_back(i + 1 < a.size() ? add(a[i], scale(a[i + 1], 4)) : a[i]);
        return p;
    }
    static std::vector<int> pack_scalar(uint64_t c, size_t nblocks) {
        std::vector<int> p;
        for (size_t i = 0; i < nblocks; i += 2) {
            int lo = (int)((c >> (2 * i)) & 3), hi = (i + 1 < nblocks) ? (int)((c >> (2 * i + 2)) & 3) : 0;
            p.push_back(lo + 4 * hi);
        }
        return p;
    }
    static bool scalar_fits(uint64_t c, size_t nblocks) { return nblocks >= 32 || (c >> (2 * nblocks)) == 0; }

    // ---------------- equality ----------------
    std::vector<Bool> eq_bits(const Rad &a, const Rad &b) {  // one bool per packed pair
        std::vector<Lin> pa = pack(a), pb = pack(b);
        std::vector<Bool> r;
        for (size_t i = 0; i < pa.size(); i++)
            r.push_back(g.pbs(sub(pa[i], pb[i]), [](int x) { return x == 0; }, 0, 1, true));
        return r;
    }
    Bool eq(const Rad &a, const Rad &b) { return all(eq_bits(a, b)); }
    std::vector<Bool> eq_scalar_bits(const Rad &a, uint64_t c) {
        std::vector<Lin> pa = pack(a);
        std::vector<int> pc = pack_scalar(c, a.size());
        std::vector<Bool> r;
        for (size_t i = 0; i < pa.size(); i++) {
            int ci = pc[i];
            r.push_back(g.pbs(pa[i], [ci](int x) { return x == ci; }, 0, 1));
        }
        return r;
    }
    Bool eq_scalar(const Rad &a, uint64_t c) {
        if (!scalar_fits(c, a.size())) return Graph::constant(0);
        return all(eq_scalar_bits(a, c));
    }
    // a == 0 for a radix of <= 5 clean blocks: one PBS on the digit sum
    Bool is_zero(const Rad &a) {
        std::vector<Bool> parts;
        for (size_t i = 0; i < a.size(); i += 5) {
            Lin s = Graph::constant(0);
            for (size_t j = i; j < std::min(a.size(), i + 5); j++) s = add(s, a[j]);
            parts.push_back(g.pbs(s, [](int x) { return x == 0; }, 0, 1));
        }
        return all(parts);
    }

    // ---------------- ordering ----------------
    // sign in {-1: a<b, 0: a==b, +1: a>b} per packed pair, most significant pair last: one signed PBS on
    // the difference (negative inputs come back as -LUT[x+16], the negacyclic convention)
    std::vector<Lin> sign_pairs(const std::vector<Lin> &pa, const std::vector<Lin> &pb) {
        std::vector<Lin> s;
        for (size_t i = 0; i < pa.size(); i++)
            s.push_back(g.pbs(sub(pa[i], pb[i]), [](int x) { return x == 0 ? 0 : 1; }, -1, 1, true));
        return s;
    }
    // Reduce signs (index 0 least significant) to one and map it through `fin` (argument 0: less,
    // 1: equal, 2: greater).  Four signs at a time: 8 s3 + 4 s2 + 2 s1 + s0 lies in [-15, 15] and has
    // the sign of the most significant non-zero s_k, so the same signed sign-LUT combines them
    // (log4 depth; the pairwise 3*hi+lo scheme needed log2).  The last <= 3 signs are shifted to an
    // unsigned value and go through `fin` directly.
    Bool reduce_sign(std::vector<Lin> s, std::function<int(int)> fin) {
        if (s.empty()) return Graph::constant(fin(1));
        while (s.size() > 3) {
            std::vector<Lin> nx;
            for (size_t i = 0; i < s.size(); i += 4) {
                const size_t n = std::min<size_t>(4, s.size() - i);
                if (n == 1) { nx.push_back(s[i]); continue; }
                Lin x = Graph::constant(0);
                for (size_t k = 0; k < n; k++) x = add(x, scale(s[i + k], 1 << k));
                nx.push_back(g.pbs(x, [](int v) { return v == 0 ? 0 : 1; }, -1, 1, true));
            }
            s.swap(nx);
        }
        Lin x = Graph::constant(0);
        int off = 0;
        for (size_t k = 0; k < s.size(); k++) { x = add(x, scale(s[k], 1 << k)); off += 1 << k; }
        x = bounded(addc(x, off), 0, 2 * off);
        return g.pbs(x, [fin, off](int v) { const int d = v - off; return fin(d < 0 ? 0 : (d == 0 ? 1 : 2)); }, 0, 1);
    }
    // generic comparison with final map fin(sign)
    Bool cmp(const Rad &a, const Rad &b, std::function<int(int)> fin) {
        return reduce_sign(sign_pairs(pack(a), pack(b)), fin);
    }
    Bool cmp_scalar(const Rad &a, uint64_t c, std::function<int(int)> fin) {
        if (!scalar_fits(c, a.size())) return Graph::constant(fin(0));  // a < c always (fin(0) = "less")
        std::vector<int> pc = pack_scalar(c, a.size());
        std::vector<Lin> pb;
        for (int v : pc) pb.push_back(Graph::constant(v));
        return reduce_sign(sign_pairs(pack(a), pb), fin);
    }
    Bool lt(const Rad &a, const Rad &b) { return cmp(a, b, [](int s) { return s == 0; }); }
    Bool le(const Rad &a, const Rad &b) { return cmp(a, b, [](int s) { return s <= 1; }); }
    Bool gt(const Rad &a, const Rad &b) { return cmp(a, b, [](int s) { return s == 2; }); }
    Bool ge(const Rad &a, const Rad &b) { return cmp(a, b, [](int s) { return s >= 1; }); }
    Bool lt_scalar(const Rad &a, uint64_t c) { return cmp_scalar(a, c, [](int s) { return s == 0; }); }
    Bool gt_scalar(const Rad &a, uint64_t c) { return cmp_scalar(a, c, [](int s) { return s == 2; }); }

    // ---------------- selection ----------------
    // c ? a : 0   (logic.rs:13 if_then_else_zero -- one PBS per block instead of a radix multiply)
    Rad select_zero(const Bool &c, const Rad &a) {
        Rad r;
        for (auto &blk : a) {
            if (blk.is_const() && blk.c == 0) { r.push_back(Graph::constant(0)); continue; }
            r.push_back(g.pbs(add(scale(c, 4), blk), [](int x) { return x >= 4 ? (x & 3) : 0; }, 0, 3));
        }
        return r;
    }
    Rad if_then_else(const Bool &c, const Rad &a, const Rad &b) {
        Rad r;
        for (size_t i = 0; i < a.size(); i++) {
            const bool az = a[i].is_const() && a[i].c == 0, bz = b[i].is_const() && b[i].c == 0;
            Lin t = az ? Graph::constant(0) : g.pbs(add(scale(c, 4), a[i]), [](int x) { return x >= 4 ? (x & 3) : 0; }, 0, 3);
            Lin f = bz ? Graph::constant(0) : g.pbs(add(scale(c, 4), b[i]), [](int x) { return x < 4 ? x : 0; }, 0, 3);
            r.push_back(bounded(add(t, f), 0, 3));
        }
        return r;
    }
    // c ? a : b with ONE evaluation per block, for chains of selections on the same operands (barrel shifters).
    // c7 = 7 c (bit7 below makes it with the evaluation that would have made c).  x = c7 + (a - b) + 3 lies in
    // [0, 13] and determines (c, a - b); the table x >= 7 ? x - 10 : 0 returns c (a - b) in [-3, 3]; the result is
    // b + that, linear.  It carries b's noise + 1, so a chain must END with if_then_else (two evaluations per
    // block, fresh result); a link whose input would exceed the linear-combination budget throws.
    Rad mux7(const Bool &c7, const Rad &a, const Rad &b) {
        Rad r;
        for (size_t i = 0; i < a.size(); i++) {
            const bool az = a[i].is_const() && a[i].c == 0, bz = b[i].is_const() && b[i].c == 0;
            if (az && bz) { r.push_back(Graph::constant(0)); continue; }
            if (c7.noise + a[i].noise + b[i].noise > kMaxLinNoise) throw std::logic_error("mux7: chain too long, refresh with if_then_else");
            if (az) { r.push_back(g.pbs(add(c7, b[i]), [](int x) { return x < 7 ? x : 0; }, 0, 3)); continue; }
            if (bz) { r.push_back(g.pbs(add(c7, a[i]), [](int x) { return x >= 7 ? x - 7 : 0; }, 0, 3)); continue; }
            Lin x = bounded(addc(add(c7, sub(a[i], b[i])), 3), 0, 13);
            Lin f = g.pbs(x, [](int v) { return v >= 7 ? v - 10 : 0; }, -3, 3);
            r.push_back(bounded(add(b[i], f), 0, 3));
        }
        return r;
    }
    // would a mux7 link on these operands stay inside the linear-combination budget?  (c7 is a fresh evaluation)
    static bool mux7_fits(const std::vector<Rad> &xs) {
        float worst = 0.f;
        for (auto &r : xs) for (auto &blk : r) worst = std::max(worst, blk.noise);
        return 1.f + 2.f * worst <= kMaxLinNoise;
    }
    Bool if_then_else_bool(const Bool &c, const Bool &a, const Bool &b) {
        // 4c + 2a + b <= 7
        return g.pbs(add(add(scale(c, 4), scale(a, 2)), b), [](int x) { return x >= 4 ? ((x >> 1) & 1) : (x & 1); }, 0, 1);
    }
    // c ? scalar : 0 (logic.rs:19 scalar_if_then_else_zero): linear, no PBS
    Rad bool_times_scalar(const Bool &c, uint64_t v, size_t nblocks) {
        Rad r;
        for (size_t i = 0; i < nblocks; i++) r.push_back(bounded(scale(c, (int)((v >> (2 * i)) & 3)), 0, (int)((v >> (2 * i)) & 3)));
        return r;
    }
    static Rad scalar(uint64_t v, size_t nblocks) {
        Rad r;
        for (size_t i = 0; i < nblocks; i++) r.push_back(Graph::constant((int64_t)((v >> (2 * i)) & 3)));
        return r;
    }
    Rad from_bool(const Bool &b, size_t nblocks) {
        Rad r{b};
        for (size_t i = 1; i < nblocks; i++) r.push_back(Graph::constant(0));
        return r;
    }
    // Sum of radix values of which AT MOST ONE is non-zero (one-hot selections).  Linear; where the
    // variance of a digit's sum would exceed kMaxLinNoise PBS outputs the terms are cut into groups below that
    // bound, every group is refreshed by an identity LUT IN THE SAME LEVEL and the fresh group sums are
    // added (a running sum with refreshes along the way would cost one level per refresh).
    Rad onehot_sum(const std::vector<Rad> &v, size_t nblocks) {
        Rad acc(nblocks);
        for (size_t i = 0; i < nblocks; i++) {
            std::vector<Lin> groups;
            Lin cur = Graph::constant(0);
            for (auto &r : v) {
                if (r[i].is_const() && r[i].c == 0) continue;
                if (cur.noise + r[i].noise > kMaxLinNoise) { groups.push_back(cur); cur = Graph::constant(0); }
                cur = bounded(add(cur, r[i]), 0, 3);
            }
            if (groups.empty()) {
                acc[i] = cur;
            } else {
                groups.push_back(cur);
                Lin sum = Graph::constant(0);
                for (auto &gsum : groups)
                    sum = bounded(add(sum, g.pbs(gsum, [](int x) { return x & 3; }, 0, 3)), 0, 3);
                acc[i] = sum;
            }
            // callers pack / subtract the result (variance x17 .. x34): hand back fresh digits
            if (acc[i].noise > 10.f) acc[i] = g.pbs(acc[i], [](int x) { return x & 3; }, 0, 3);
        }
        return acc;
    }

    // ---------------- addition ----------------
    // blocks s_i in [0,7] (digit sums incl. injected +1) -> clean digits of the sum mod 4^n
    Rad propagate(const std::vector<Lin> &s) {
        const size_t n = s.size();
        if (n == 0) return {};
        // carry state per block: 0 none, 1 propagate (s==3), 2 generate (s>=4)
        std::vector<Lin> st(n);
        for (size_t i = 0; i + 1 < n; i++)
            st[i] = s[i].is_const() ? Graph::constant(s[i].c >= 4 ? 2 : (s[i].c == 3 ? 1 : 0))
                                    : g.pbs(s[i], [](int x) { return x >= 4 ? 2 : (x == 3 ? 1 : 0); }, 0, 2);
        // inclusive prefix of the carry operator (Hillis-Steele): pre[i] = state of blocks 0..i
        std::vector<Lin> pre(st.begin(), st.begin() + (n ? n - 1 : 0));
        for (size_t d = 1; d < pre.size(); d *= 2) {
            std::vector<Lin> nx = pre;
            for (size_t i = d; i < pre.size(); i++)
                nx[i] = g.pbs(add(scale(pre[i], 3), pre[i - d]),
                              [](int v) { int hi = v / 3, lo = v % 3; return hi == 1 ? lo : hi; }, 0, 2);
            pre.swap(nx);
        }
        Rad out(n);
        for (size_t i = 0; i < n; i++) {
            if (i == 0) { out[i] = g.pbs(s[i], [](int x) { return x & 3; }, 0, 3); continue; }
            // carry-in = (pre[i-1] == 2); s_i + 8*... does not fit: fold carry via 2-step value
            Lin cin = g.pbs(pre[i - 1], [](int v) { return v == 2; }, 0, 1);
            out[i] = g.pbs(bounded(add(s[i], cin), 0, 8), [](int x) { return x & 3; }, 0, 3);
        }
        return out;
    }
    Rad add_rad(const Rad &a, const Rad &b) {
        std::vector<Lin> s;
        for (size_t i = 0; i < a.size(); i++) s.push_back(add(a[i], b[i]));
        return propagate(s);
    }
    Rad add_scalar(const Rad &a, uint64_t c) {
        std::vector<Lin> s;
        for (size_t i = 0; i < a.size(); i++) s.push_back(addc(a[i], (int64_t)((c >> (2 * i)) & 3)));
        return propagate(s);
    }
    Rad sub_rad(const Rad &a, const Rad &b) {  // a - b mod 4^n = a + ~b + 1
        std::vector<Lin> s;
        for (size_t i = 0; i < a.size(); i++) {
            Lin x = bounded(add(a[i], sub(Graph::constant(3), b[i])), 0, 6);
            if (i == 0) x = bounded(addc(x, 1), 0, 7);
            s.push_back(x);
        }
        return propagate(s);
    }
    // cnt + b for a boolean b (ripple over the digits; in a chain over positions the ripples pipeline)
    Rad inc_if(const Rad &cnt, const Bool &b) {
        Rad r;
        Lin carry = b;
        for (size_t k = 0; k < cnt.size(); k++) {
            Lin s = bounded(add(cnt[k], carry), 0, 4);
            if (s.is_const()) { r.push_back(Graph::constant(s.c & 3)); carry = Graph::constant(s.c >> 2); continue; }
            r.push_back(g.pbs(s, [](int x) { return x & 3; }, 0, 3));
            if (k + 1 < cnt.size()) carry = (carry.is_const() && carry.c == 0) ? Graph::constant(0) : g.pbs(s, [](int x) { return x >> 2; }, 0, 1);
        }
        return r;
    }
    // low bit / high bit of a clean block
    Bool bit0(const Lin &blk) { return g.pbs(blk, [](int x) { return x & 1; }, 0, 1); }
    Bool bit1(const Lin &blk) { return g.pbs(blk, [](int x) { return (x >> 1) & 1; }, 0, 1); }
    // bit `bit` of a base-4 digit as 0 / 7: the condition operand of mux7
    Bool bit7(const Lin &blk, int bit) { return g.pbs(blk, [bit](int x) { return ((x >> bit) & 1) ? 7 : 0; }, 0, 7); }
};

}  // namespace fhs_host
