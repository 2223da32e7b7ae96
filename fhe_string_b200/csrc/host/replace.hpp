// replace / replacen (reference src/ciphertext/replace.rs:17-104) as a level program.
// The reference walks the OUTPUT serially with char_at / element_at per position.  Here the result
// is a concatenation of per-position segments  seg_i = (match starts at i ? s : "") ++ (i covered by
// a match ? "" : a[i])  built by a balanced tree of encrypted-length concatenations:
//   concat(A, B) = A + shift_right(B, len(A))   (blockwise add: the operands never overlap)
// so the depth is O(log^2 L) instead of O(L) and all shifts of one tree level share their levels.
#pragma once
#include "split.hpp"

namespace fhs_host {

struct ReplaceOps : SplitOps {
    using SplitOps::SplitOps;

    // chars beyond len are 0.  `ch` holds min(true_max, l) chars (the output capacity l truncates the
    // BUFFER, replace.rs:45-101 keeps the first l chars of the result); `len` is the untruncated length
    // in as many base-4 digits as `true_max` needs, so it neither wraps at FheUsize::max nor lags behind
    // the buffer when the result is longer than l.
    struct Buf { std::vector<Rad> ch; Rad len; size_t true_max; };

    static size_t digits_for(size_t v) { size_t w = 1; while (w < 32 && (v >> (2 * w)) != 0) w++; return w; }
    // same value in `w` digits: constant zeros appended, provably-zero top digits dropped
    static Rad fit(Rad r, size_t w) {
        while (r.size() < w) r.push_back(Graph::constant(0));
        r.resize(w);
        return r;
    }
    static bool rad_is_zero_const(const Rad &r) {
        for (auto &b : r) if (!(b.is_const() && b.c == 0)) return false;
        return true;
    }
    Rad add_chars(const Rad &x, const Rad &y) {  // disjoint supports: at most one is non-zero
        Rad r;
        for (size_t k = 0; k < x.size(); k++) r.push_back(bounded(add(x[k], y[k]), 0, 3));
        return r;
    }
    // B shifted right by `amount` (<= max_amount) inside a buffer of `cap` chars; a shift of cap or more
    // leaves nothing of B in the buffer
    std::vector<Rad> shift_right(const std::vector<Rad> &b, const Rad &amount, size_t max_amount, size_t cap) {
        Rad z = zero_char();
        std::vector<Rad> cur(cap, z);
        for (size_t j = 0; j < std::min(cap, b.size()); j++) cur[j] = b[j];
        std::vector<Bool> overflow;
        // the last in-range stage is a two-evaluation selection (fresh result: what concat adds to A must not
        // bring a chain's noise along), the ones before it are one-evaluation mux7 links
        auto dead_bit = [&](size_t blk, int bit) { return amount[blk].is_const() && ((amount[blk].c >> bit) & 1) == 0; };
        int last_blk = -1, last_bit = -1;
        for (size_t blk = 0; blk < amount.size(); blk++)
            for (int bit = 0; bit < 2; bit++) {
                const size_t sh = (size_t)1 << (2 * blk + bit);
                if (sh <= max_amount && sh < cap && !dead_bit(blk, bit)) { last_blk = (int)blk; last_bit = bit; }
            }
        for (size_t blk = 0; blk < amount.size(); blk++) {
            for (int bit = 0; bit < 2; bit++) {
                const size_t sh = (size_t)1 << (2 * blk + bit);
                if (sh > max_amount) break;
                if (dead_bit(blk, bit)) continue;
                if (sh >= cap) { overflow.push_back(bit ? bit1(amount[blk]) : bit0(amount[blk])); continue; }
                const bool fresh = ((int)blk == last_blk && bit == last_bit) || amount[blk].is_const() || !mux7_fits(cur);
                Bool sb = fresh ? (bit ? bit1(amount[blk]) : bit0(amount[blk])) : bit7(amount[blk], bit);
                std::vector<Rad> nx(cap);
                for (size_t j = 0; j < cap; j++) {
                    const Rad &from = (j >= sh) ? cur[j - sh] : z;
                    if (rad_is_zero_const(from) && rad_is_zero_const(cur[j])) { nx[j] = z; continue; }
                    nx[j] = fresh ? if_then_else(sb, from, cur[j]) : mux7(sb, from, cur[j]);
                }
                cur.swap(nx);
            }
        }
        if (!overflow.empty()) {
            Bool keep = bnot(any(overflow));
            for (size_t j = 0; j < cap; j++)
                if (!rad_is_zero_const(cur[j])) cur[j] = select_zero(keep, cur[j]);
        }
        return cur;
    }
    Buf concat(const Buf &A, const Buf &B, size_t limit) {
        Buf R;
        R.true_max = A.true_max + B.true_max;
        const size_t cap = std::min(R.true_max, limit);
        std::vector<Rad> bs = shift_right(B.ch, A.len, A.true_max, cap);
        for (size_t j = 0; j < cap; j++) R.ch.push_back(j < A.ch.size() ? add_chars(A.ch[j], bs[j]) : bs[j]);
        const size_t w = digits_for(R.true_max);
        R.len = add_rad(fit(A.len, w), fit(B.len, w));
        return R;
    }

    // replace.rs:37 replace_opt; n == nullptr replaces every match
    FStr replace_impl(const FStr &a, const FStr &p, const FStr &s, const Rad *n, size_t l) {
        const size_t L1 = a.ch.size(), Ms = s.ch.size() - 1, Mp = p.ch.size() - 1;
        std::vector<Bool> az = zero_mask(a), pz = zero_mask(p), sz = zero_mask(s);
        std::vector<Bool> m = find_all(a, p, az, true);
        std::vector<Bool> chosen = non_overlapping(m, pz, false);
        if (n) {
            // the first n matches are used: match i is used iff fewer than n matches precede it.  Counting
            // ALL earlier matches (not the used ones) keeps the counter off the feedback path, so the
            // increments ripple alongside the selection chain instead of adding ~7 levels per position.
            Rad cnt = usize(0);
            std::vector<Bool> below(L1);
            for (size_t i = 0; i < L1; i++) {
                below[i] = Ops::lt(cnt, *n);
                cnt = inc_if(cnt, chosen[i]);
            }
            for (size_t i = 0; i < L1; i++) chosen[i] = band(chosen[i], below[i]);
        }
        Rad slen = len_from_mask(sz);
        std::vector<Buf> nodes;
        for (size_t i = 0; i < L1; i++) {
            // covered_i = sum_{d < Mp} chosen[i-d] & [len(p) > d]   (at most one term)
            Lin cov = Graph::constant(0);
            for (size_t d = 0; d < Mp && d <= i; d++) cov = add(cov, band(chosen[i - d], bnot(pz[d])));
            cov = bounded(cov, 0, 1);
            if (Ms > 0) {
                Buf seg;
                seg.true_max = Ms;
                for (size_t t = 0; t < Ms; t++) seg.ch.push_back(select_zero(chosen[i], s.ch[t]));
                seg.len = fit(select_zero(chosen[i], slen), digits_for(Ms));
                nodes.push_back(seg);
            }
            Buf lit;
            lit.true_max = 1;
            Bool keep = bnot(cov);
            lit.ch.push_back(select_zero(keep, a.ch[i]));
            Bool nonzero = g.pbs(add(keep, bnot(az[i])), [](int x) { return x == 2; }, 0, 1);
            lit.len = from_bool(nonzero, 1);
            nodes.push_back(lit);
        }
        while (nodes.size() > 1) {
            std::vector<Buf> nx;
            for (size_t i = 0; i + 1 < nodes.size(); i += 2) nx.push_back(concat(nodes[i], nodes[i + 1], l));
            if (nodes.size() & 1) nx.push_back(nodes.back());
            nodes.swap(nx);
        }
        FStr out;
        Rad z = zero_char();
        for (size_t j = 0; j < l; j++) out.ch.push_back(j < nodes[0].ch.size() ? nodes[0].ch[j] : z);
        out.ch.push_back(z);  // terminator
        return out;
    }
    FStr replace(const FStr &a, const FStr &p, const FStr &s, size_t l) { return replace_impl(a, p, s, nullptr, l); }
    FStr replacen(const FStr &a, const FStr &p, const FStr &s, const Rad &n, size_t l) { return replace_impl(a, p, s, &n, l); }
};

}  // namespace fhs_host
