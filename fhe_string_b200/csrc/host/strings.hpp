// String operations as level programs.  Mirrors the public methods of the reference's FheString
// (src/ciphertext/{mod,compare,convert,search,trim,split,replace}.rs): same names, same argument
// meaning, same decrypted results (std::str semantics).  The algorithms are
// re-derived for batching and depth: prefix scans instead of serial folds, barrel
// shifts instead of per-index char_at, one-PBS selections instead of radix multiplies.
#pragma once
#include "radix.hpp"

namespace fhs_host {

struct FStr { std::vector<Rad> ch; };           // chars (4 blocks each), zero-terminated, zero-padded
struct FOptUsize { Bool is_some; Rad val; };
struct FOptStr { Bool is_some; FStr val; };
struct FSlice { Bool is_some; Rad start, end; };
struct FSliceVec { FStr s; std::vector<FSlice> v; };

struct StrOps : Ops {
    size_t nb_char, nb_usize;
    StrOps(Graph &g_, size_t nbc, size_t nbu) : Ops(g_), nb_char(nbc), nb_usize(nbu) {}

    Rad zero_char() const { return scalar(0, nb_char); }
    const Rad &at(const FStr &s, size_t i, const Rad &z) const { return i < s.ch.size() ? s.ch[i] : z; }
    Rad usize(uint64_t v) const { return scalar(v, nb_usize); }

    // [s[i] == 0] for every char
    std::vector<Bool> zero_mask(const FStr &s) {
        std::vector<Bool> z;
        for (auto &c : s.ch) z.push_back(is_zero(c));
        return z;
    }

    // compare.rs:15
    Bool is_empty(const FStr &s) { return is_zero(s.ch[0]); }

    // mod.rs:132  len = sum_i i * [s[i-1] != 0 && s[i] == 0]   (one-hot -> linear digits)
    Rad len(const FStr &s) { return len_from_mask(zero_mask(s)); }
    Rad len_from_mask(const std::vector<Bool> &z) {
        std::vector<Rad> terms;
        for (size_t i = 1; i < z.size(); i++) {
            Bool t = g.pbs(add(bnot(z[i - 1]), z[i]), [](int x) { return x == 2; }, 0, 1);
            terms.push_back(bool_times_scalar(t, i, nb_usize));
        }
        return onehot_sum(terms, nb_usize);
    }

    // compare.rs:21  equal over the common chars, the longer operand's overhang must start with 0
    Bool eq(const FStr &a, const FStr &b) {
        const size_t n = std::min(a.ch.size(), b.ch.size());
        std::vector<Bool> bits;
        for (size_t i = 0; i < n; i++) {
            auto e = eq_bits(a.ch[i], b.ch[i]);
            bits.insert(bits.end(), e.begin(), e.end());
        }
        if (a.ch.size() > n) bits.push_back(is_zero(a.ch[n]));
        if (b.ch.size() > n) bits.push_back(is_zero(b.ch[n]));
        return all(bits);
    }
    Bool ne(const FStr &a, const FStr &b) { return bnot(eq(a, b)); }

    // compare.rs:63  lexicographic, 0 < every char, shorter operand padded with zeros
    Bool cmp_str(const FStr &a, const FStr &b, std::function<int(int)> fin) {
        const size_t n = std::max(a.ch.size(), b.ch.size());
        Rad z = zero_char();
        std::vector<Lin> signs;  // least significant first = last char first
        for (size_t k = 0; k < n; k++) {
            size_t i = n - 1 - k;
            auto s = sign_pairs(pack(at(a, i, z)), pack(at(b, i, z)));
            signs.insert(signs.end(), s.begin(), s.end());
        }
        return reduce_sign(signs, fin);
    }
    Bool lt(const FStr &a, const FStr &b) { return cmp_str(a, b, [](int s) { return s == 0; }); }
    Bool le(const FStr &a, const FStr &b) { return cmp_str(a, b, [](int s) { return s <= 1; }); }
    Bool gt(const FStr &a, const FStr &b) { return cmp_str(a, b, [](int s) { return s == 2; }); }
    Bool ge(const FStr &a, const FStr &b) { return cmp_str(a, b, [](int s) { return s >= 1; }); }

    // convert.rs:14-45  chars are (b3 b2 b1 b0) base 4; 'a'..'z' = 97..122 and 'A'..'Z' = 65..90
    // differ only in b2 (+-2 with no carry/borrow), so the result is b2 -+ 2*[in range]: 3 PBS/char.
    Rad case_shift(const Rad &c, bool to_upper) {
        Lin hi = add(c[2], scale(c[3], 4)), lo = add(c[0], scale(c[1], 4));
        const int h0 = to_upper ? 6 : 4;  // high pair of the first letter of the source range
        // source range: hi == h0 && lo >= 1  ||  hi == h0+1 && lo <= 10
        Lin hc = g.pbs(hi, [h0](int x) { return x == h0 ? 1 : (x == h0 + 1 ? 2 : 0); }, 0, 2);
        Lin lc = g.pbs(lo, [](int x) { return (x >= 1 ? 1 : 0) + (x <= 10 ? 2 : 0); }, 1, 3);
        Lin two = g.pbs(add(scale(hc, 4), lc), [](int v) {
            int h = v >> 2, l = v & 3;
            return ((h == 1 && (l & 1)) || (h == 2 && (l & 2))) ? 2 : 0; }, 0, 2);
        Rad r = c;
        r[2] = bounded(to_upper ? sub(c[2], two) : add(c[2], two), 0, 3);
        return r;
    }
    FStr to_uppercase(const FStr &s) { FStr r; for (auto &c : s.ch) r.ch.push_back(case_shift(c, true)); return r; }
    FStr to_lowercase(const FStr &s) { FStr r; for (auto &c : s.ch) r.ch.push_back(case_shift(c, false)); return r; }
    // compare.rs:113
    Bool eq_ignore_ascii_case(const FStr &a, const FStr &b) { return eq(to_lowercase(a), to_lowercase(b)); }

    // convert.rs:14 / :22  FheAsciiChar::is_uppercase (65..90) / is_lowercase (97..122): the same range test
    // case_shift uses, as a boolean
    Bool is_case(const Rad &c, bool upper) {
        Lin hi = add(c[2], scale(c[3], 4)), lo = add(c[0], scale(c[1], 4));
        const int h0 = upper ? 4 : 6;
        Lin hc = g.pbs(hi, [h0](int x) { return x == h0 ? 1 : (x == h0 + 1 ? 2 : 0); }, 0, 2);
        Lin lc = g.pbs(lo, [](int x) { return (x >= 1 ? 1 : 0) + (x <= 10 ? 2 : 0); }, 1, 3);
        return g.pbs(add(scale(hc, 4), lc), [](int v) {
            int h = v >> 2, l = v & 3;
            return ((h == 1 && (l & 1)) || (h == 2 && (l & 2))) ? 1 : 0; }, 0, 1);
    }

    // trim.rs:15  whitespace = 9..13 or 32
    Bool is_whitespace(const Rad &c) {
        Lin hi = add(c[2], scale(c[3], 4)), lo = add(c[0], scale(c[1], 4));
        Lin hc = g.pbs(hi, [](int x) { return x == 0 ? 1 : (x == 2 ? 2 : 0); }, 0, 2);
        Lin lc = g.pbs(lo, [](int x) { return (x >= 9 && x <= 13) ? 1 : (x == 0 ? 2 : 0); }, 0, 2);
        return g.pbs(add(scale(hc, 3), lc), [](int v) { int h = v / 3, l = v % 3; return (h == 1 && l == 1) || (h == 2 && l == 2); }, 0, 1);
    }

    // ---- pattern matching (compare.rs:119 substr_eq at every offset, search.rs:33 find_all) ----
    // m[i] = for all j: a[i+j] == b[j] || b[j] == 0, with a read as 0 past its end;
    // if `limit_to_len`, additionally i <= len(a) (so that "" matches only up to the end).
    std::vector<Bool> find_all(const FStr &a, const FStr &b, const std::vector<Bool> &az, bool limit_to_len) {
        std::vector<Bool> bz = zero_mask(b);
        const size_t L1 = a.ch.size(), M1 = b.ch.size();
        std::vector<Bool> m(L1);
        for (size_t i = 0; i < L1; i++) {
            std::vector<Bool> terms;
            for (size_t j = 0; j < M1; j++) {
                if (i + j >= L1) { terms.push_back(bz[j]); continue; }
                auto e = eq_bits(a.ch[i + j], b.ch[j]);  // 2 bits
                Lin x = scale(bz[j], 2);
                for (auto &bit : e) x = add(x, bit);
                terms.push_back(g.pbs(x, [](int v) { return v >= 2; }, 0, 1));
            }
            if (limit_to_len && i >= 1) terms.push_back(bnot(az[i - 1]));
            m[i] = all(terms);
        }
        return m;
    }
    // search.rs:17
    Bool contains(const FStr &a, const FStr &b) { return any(find_all(a, b, {}, false)); }
    // search.rs:216
    Bool starts_with(const FStr &a, const FStr &b) {
        FStr head; head.ch.assign(a.ch.begin(), a.ch.begin() + 1);
        // only offset 0 is needed
        std::vector<Bool> bz = zero_mask(b);
        std::vector<Bool> terms;
        for (size_t j = 0; j < b.ch.size(); j++) {
            if (j >= a.ch.size()) { terms.push_back(bz[j]); continue; }
            auto e = eq_bits(a.ch[j], b.ch[j]);
            Lin x = scale(bz[j], 2);
            for (auto &bit : e) x = add(x, bit);
            terms.push_back(g.pbs(x, [](int v) { return v >= 2; }, 0, 1));
        }
        return all(terms);
    }
    // first (or last) set position of v as an option: one-hot via prefix OR
    FOptUsize index_of(const std::vector<Bool> &v, bool reverse) {
        std::vector<Bool> w = v;
        if (reverse) std::reverse(w.begin(), w.end());
        std::vector<Bool> p = prefix_or(w);
        std::vector<Rad> terms;
        for (size_t k = 0; k < w.size(); k++) {
            // the prefix OR is monotone, so "first set position" is its discrete derivative: no PBS
            Bool first = (k == 0) ? p[0] : bounded(sub(p[k], p[k - 1]), 0, 1);
            size_t idx = reverse ? w.size() - 1 - k : k;
            terms.push_back(bool_times_scalar(first, idx, nb_usize));
        }
        return FOptUsize{p.back(), onehot_sum(terms, nb_usize)};
    }
    // search.rs:23 / :155
    FOptUsize find(const FStr &a, const FStr &b) { return index_of(find_all(a, b, zero_mask(a), true), false); }
    FOptUsize rfind(const FStr &a, const FStr &b) { return index_of(find_all(a, b, zero_mask(a), true), true); }
    // search.rs:222  some offset where a[i..] == b exactly (zeros included; a is 0 past its end)
    std::vector<Bool> suffix_matches(const FStr &a, const FStr &b) {
        const size_t L1 = a.ch.size(), M1 = b.ch.size();
        Rad z = zero_char();
        std::vector<Bool> bz = zero_mask(b);
        std::vector<Bool> m(L1);
        for (size_t i = 0; i < L1; i++) {
            std::vector<Bool> terms;
            for (size_t j = 0; j < M1; j++) {
                if (i + j >= L1) { terms.push_back(bz[j]); continue; }
                auto e = eq_bits(a.ch[i + j], b.ch[j]);
                terms.insert(terms.end(), e.begin(), e.end());
            }
            m[i] = all(terms);
        }
        return m;
    }
    Bool ends_with(const FStr &a, const FStr &b) { return any(suffix_matches(a, b)); }

    // ---- shifting / truncation ----
    // out[i] = a[i + idx] (0 past the end): log-depth barrel shifter on the bits of idx
    FStr shift_left(const FStr &a, const Rad &idx) {
        FStr cur = a;
        const size_t L1 = a.ch.size();
        Rad z = zero_char();
        // the last in-range stage is a two-evaluation selection (fresh result), the ones before it mux7 links
        int last_blk = -1, last_bit = -1;
        for (size_t blk = 0; blk < idx.size(); blk++)
            for (int bit = 0; bit < 2; bit++)
                if (((size_t)1 << (2 * blk + bit)) < L1 && !(idx[blk].is_const() && ((idx[blk].c >> bit) & 1) == 0)) { last_blk = (int)blk; last_bit = bit; }
        for (size_t blk = 0; blk < idx.size(); blk++) {
            for (int bit = 0; bit < 2; bit++) {
                const size_t sh = (size_t)1 << (2 * blk + bit);
                if (sh >= L1) {
                    // any set higher bit empties the string
                    Bool hb = bit ? bit1(idx[blk]) : bit0(idx[blk]);
                    if (hb.is_const() && hb.c == 0) continue;
                    FStr nx;
                    for (size_t i = 0; i < L1; i++) nx.ch.push_back(select_zero(bnot(hb), cur.ch[i]));
                    cur = nx;
                    continue;
                }
                if (idx[blk].is_const() && ((idx[blk].c >> bit) & 1) == 0) continue;
                // (a chain that has used up the noise budget -- very long strings -- is refreshed in between)
                const bool fresh = ((int)blk == last_blk && bit == last_bit) || idx[blk].is_const() || !mux7_fits(cur.ch);
                Bool sb = fresh ? (bit ? bit1(idx[blk]) : bit0(idx[blk])) : bit7(idx[blk], bit);
                FStr nx;
                for (size_t i = 0; i < L1; i++) {
                    const Rad &sh_c = (i + sh < L1) ? cur.ch[i + sh] : z;
                    nx.ch.push_back(fresh ? if_then_else(sb, sh_c, cur.ch[i]) : mux7(sb, sh_c, cur.ch[i]));
                }
                cur = nx;
            }
        }
        return cur;
    }
    // keep[i] ? a[i] : 0
    FStr mask_chars(const FStr &a, const std::vector<Bool> &keep) {
        FStr r;
        for (size_t i = 0; i < a.ch.size(); i++) r.ch.push_back(select_zero(keep[i], a.ch[i]));
        return r;
    }
    // mod.rs:165 substr_to: a[i] if i < index
    FStr substr_to(const FStr &a, const Rad &index) {
        std::vector<Bool> keep;
        for (size_t i = 0; i < a.ch.size(); i++) keep.push_back(gt_scalar(index, i));
        return mask_chars(a, keep);
    }
    FStr substr_from(const FStr &a, const Rad &index) { return shift_left(a, index); }
    // mod.rs:198 substr_end: a[start..end] = chars i + start < end, shifted left by start
    FStr substr_end(const FStr &a, const Rad &start, const Rad &end) { return shift_left(substr_to(a, end), start); }
    // mod.rs:217 char_at: a[i], 0 if out of bounds -- one-hot over the positions (the reference sums
    // if_then_else_zero(i == j, a[j]) the same way)
    Rad char_at(const FStr &a, const Rad &i) {
        std::vector<Rad> terms;
        for (size_t j = 0; j < a.ch.size(); j++) terms.push_back(select_zero(eq_scalar(i, j), a.ch[j]));
        return onehot_sum(terms, nb_char);
    }
    // compare.rs:119 substr_eq at the CLEAR offset i: for all j: a[i+j] == b[j] || b[j] == 0, a read as 0
    // past its end (the reference pads with substr_clear and demands an empty overhang: the same predicate)
    Bool substr_eq(const FStr &a, size_t i, const FStr &b) {
        std::vector<Bool> bz = zero_mask(b);
        std::vector<Bool> terms;
        for (size_t j = 0; j < b.ch.size(); j++) {
            if (i + j >= a.ch.size()) { terms.push_back(bz[j]); continue; }
            auto e = eq_bits(a.ch[i + j], b.ch[j]);
            Lin x = scale(bz[j], 2);
            for (auto &bit : e) x = add(x, bit);
            terms.push_back(g.pbs(x, [](int v) { return v >= 2; }, 0, 1));
        }
        return all(terms);
    }

    // trim.rs:31
    FStr trim_start(const FStr &a) {
        // amount = index of the first char that is not whitespace (the terminator is not whitespace)
        std::vector<Bool> nws;
        for (auto &c : a.ch) nws.push_back(bnot(is_whitespace(c)));
        FOptUsize first = index_of(nws, false);
        return shift_left(a, first.val);
    }
    // trim.rs:43  zero everything after the last char that is neither whitespace nor 0
    FStr trim_end(const FStr &a) {
        std::vector<Bool> solid;
        std::vector<Bool> az = zero_mask(a);
        for (size_t i = 0; i < a.ch.size(); i++)
            solid.push_back(g.pbs(add(is_whitespace(a.ch[i]), az[i]), [](int x) { return x == 0; }, 0, 1));
        std::reverse(solid.begin(), solid.end());
        std::vector<Bool> keep = prefix_or(solid);  // suffix OR
        std::reverse(keep.begin(), keep.end());
        return mask_chars(a, keep);
    }
    // trim.rs:62
    FStr trim(const FStr &a) { return trim_start(trim_end(a)); }
    // trim.rs:89
    FOptStr strip_prefix(const FStr &a, const FStr &b) {
        return FOptStr{starts_with(a, b), shift_left(a, len(b))};
    }
    // trim.rs:101  keep a[i] while a[i + len(b)] != 0
    FOptStr strip_suffix(const FStr &a, const FStr &b) {
        FStr sh = shift_left(a, len(b));
        std::vector<Bool> keep;
        for (auto &c : sh.ch) keep.push_back(bnot(is_zero(c)));
        return FOptStr{ends_with(a, b), mask_chars(a, keep)};
    }
};

}  // namespace fhs_host
