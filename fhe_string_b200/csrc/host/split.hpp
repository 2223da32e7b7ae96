// split family (reference src/ciphertext/split.rs:238-515, search.rs:68-152) as level programs.
// Pieces are described by their END: every non-overlapping match at position c closes a piece
// [S, c) where S = (previous match) + len(p) (or 0); a final piece [S, len(a)) follows.  This gives
// Rust's results for every pattern, including the empty one ("abc".split("") = ["","a","b","c",""],
// tests/split.rs:33-47), without special cases.  The greedy non-overlap decision is the only serial
// part: one PBS of depth per position.
#pragma once
#include <cstdlib>
#include <functional>
#include <map>

#include "strings.hpp"

namespace fhs_host {

struct SplitOps : StrOps {
    using StrOps::StrOps;

    // Greedy non-overlapping selection among candidate match starts m[i] for an encrypted pattern
    // b of at most M chars: chosen[i] = m[i] && no chosen match in the previous len(b)-1 positions
    // (forward) / next positions (reverse).  gp[d] = [len(b) > d] = [b[d] != 0].
    //
    // One scan step, with the earlier decisions supplied by `get` (k = index in scan order):
    Bool overlap_step(const Bool &mk, size_t k, size_t M, const std::vector<Bool> &bz,
                      const std::function<Bool(size_t)> &get) {
        // far = sum_{d >= 2} chosen[k-d] & gp[d]   (at most one term is 1)
        Lin far = Graph::constant(0);
        for (size_t d = 2; d < M && d <= k; d++) {
            Bool c = get(k - d);
            if (c.is_const()) { if (c.c) far = add(far, bnot(bz[d])); continue; }
            far = add(far, band(c, bnot(bz[d])));
        }
        far = bounded(far, 0, 1);
        Lin near = Graph::constant(0);  // chosen[k-1] + gp[1]: busy iff == 2
        if (k >= 1 && M >= 2) near = add(get(k - 1), bnot(bz[1]));
        Lin x = add(add(scale(mk, 8), scale(far, 4)), near);
        return g.pbs(x, [](int v) { return (v >> 3) == 1 && ((v >> 2) & 1) == 0 && (v & 3) < 2; }, 0, 1);
    }
    // The scan is a finite automaton whose state is the distance h to the nearest earlier chosen
    // match (h < M matters, anything else = "none").  For M <= 8 the positions are cut into chunks of
    // ~sqrt(L) and every chunk after the first is scanned for ALL M possible entry states at once
    // (the hypothetical history is a plaintext constant); the true entry state of chunk j, a one-hot
    // over h read off the resolved tail of chunk j-1, then selects among the M candidate results.
    // Depth: chunk + 1 per chunk boundary instead of one level per position.
    static constexpr size_t kSpecMaxPattern = 8;
    std::vector<Bool> non_overlapping(const std::vector<Bool> &m, const std::vector<Bool> &bz, bool reverse) {
        const size_t L1 = m.size();
        const size_t M = bz.size() - 1;  // max pattern length
        serial_keep = M >= 2;
        if (!serial_keep) return m;      // patterns of at most one char cannot overlap
        std::vector<Bool> chosen(L1);
        auto at = [&](size_t k) -> size_t { return reverse ? L1 - 1 - k : k; };
        size_t chunk = 1;
        while (chunk * chunk < L1) chunk++;     // depth ~ chunk + L / chunk: one level per chunk boundary
        if (chunk < M) chunk = M;
        if (M > kSpecMaxPattern || L1 < 2 * chunk + 2) {
            for (size_t k = 0; k < L1; k++)
                chosen[at(k)] = overlap_step(m[at(k)], k, M, bz, [&](size_t kk) { return chosen[at(kk)]; });
            return chosen;
        }
        serial_keep = false;
        std::vector<Bool> act(L1);       // resolved decisions, scan order
        for (size_t k = 0; k < std::min(chunk, L1); k++)
            act[k] = overlap_step(m[at(k)], k, M, bz, [&](size_t kk) { return act[kk]; });
        // candidate results of every later chunk under every entry state
        struct Cand { size_t p0, p1; std::vector<std::vector<Bool>> c; };   // c[h][k - p0]
        std::vector<Cand> cands;
        for (size_t p0 = chunk; p0 < L1; p0 += chunk) {
            Cand cd{p0, std::min(L1, p0 + chunk), {}};
            for (size_t h = 0; h < M; h++) {
                std::vector<Bool> c(cd.p1 - p0);
                for (size_t k = p0; k < cd.p1; k++)
                    c[k - p0] = overlap_step(m[at(k)], k, M, bz, [&](size_t kk) -> Bool {
                        if (kk >= p0) return c[kk - p0];
                        return Graph::constant(h > 0 && kk + h == p0 ? 1 : 0);
                    });
                cd.c.push_back(c);
            }
            cands.push_back(cd);
        }
        for (auto &cd : cands) {
            // entry state h is "hot" iff e[h] == 1, with LINEAR e[h] <= 1:
            //   e[h] = act[p0-h] - (act[p0-1] + .. + act[p0-h+1])   (h >= 1: chosen at distance h, nothing nearer)
            //   e[0] = 1 - (act[p0-1] + .. + act[p0-M+1])            (nothing chosen within M-1)
            // so "hot[h] and candidate" is the single signed table [e[h] + c == 2]: one level per chunk boundary
            // (round 1 made hot[h] with one table and the AND with a second one)
            std::vector<Lin> e(M);
            Lin seen = Graph::constant(0);
            for (size_t h = 1; h < M; h++) {
                const Bool &a = act[cd.p0 - h];
                e[h] = bounded(sub(a, seen), -(int)(h - 1), 1);
                seen = add(seen, a);
            }
            e[0] = bounded(sub(Graph::constant(1), seen), 2 - (int)M, 1);
            for (size_t k = cd.p0; k < cd.p1; k++) {
                Lin v = Graph::constant(0);
                for (size_t h = 0; h < M; h++) {
                    const Bool &c = cd.c[h][k - cd.p0];
                    if (c.is_const() && c.c == 0) continue;
                    Lin x = add(e[h], c);
                    v = add(v, g.pbs(bounded(x, e[h].lo, 2), [](int t) { return t == 2; }, 0, 1, true));
                }
                act[k] = bounded(v, 0, 1);
            }
        }
        for (size_t k = 0; k < L1; k++) chosen[at(k)] = act[k];
        return chosen;
    }

    // "value of the last set element" under the associative operator
    //   (f1,v1) o (f2,v2) = (f1|f2, f2 ? v2 : v1),   invariant v = 0 whenever f = 0,
    // so that f2 ? v2 : v1 = v2 + (f2 ? 0 : v1): one PBS per block and one for the flag, all in one level.
    struct Last { Bool f; Rad v; };
    Last last_combine(const Last &lo, const Last &hi) {
        Last r;
        r.f = bor(lo.f, hi.f);
        Rad keep_lo = select_zero(bnot(hi.f), lo.v);
        for (size_t b = 0; b < hi.v.size(); b++) r.v.push_back(bounded(add(hi.v[b], keep_lo[b]), 0, 3));
        return r;
    }
    // inclusive prefix under last_combine: Brent-Kung (2 log2 n levels, < 2n combines) -- replaces the
    // one-level-per-position chain of the reference's reverse loop (split.rs:274-319)
    std::vector<Last> last_set_scan(const std::vector<Last> &x) {
        const size_t n = x.size();
        if (n <= 1) return x;
        std::vector<Last> z(n / 2);
        for (size_t k = 0; k < n / 2; k++) z[k] = last_combine(x[2 * k], x[2 * k + 1]);
        std::vector<Last> w = last_set_scan(z);
        std::vector<Last> y(n);
        y[0] = x[0];
        for (size_t k = 0; k < n / 2; k++) y[2 * k + 1] = w[k];
        for (size_t k = 1; 2 * k < n; k++) y[2 * k] = last_combine(w[k - 1], x[2 * k]);
        return y;
    }

    enum Mode { SPLIT, INCLUSIVE, TERMINATOR };
    bool serial_keep = false;   // set when the candidate selection below is a serial chain

    // Common engine.  `limit` (optional) = encrypted maximum number of pieces (splitn).
    FSliceVec split_engine(const FStr &a, const std::vector<Bool> &chosen, const Rad &plen, Mode mode,
                           bool reverse_output, const Rad *limit, const std::vector<Bool> *piece_ok) {
        const size_t L1 = a.ch.size();
        Rad alen = len(a);
        std::vector<Bool> keep = chosen;
        if (limit) {
            // a match is used iff fewer than limit-1 matches precede it in scan order.  The counter runs
            // over ALL earlier matches, not the used ones (same predicate, no feedback through `keep`).
            Rad cnt = usize(0);
            Rad lim1 = sub_rad(*limit, usize(1));          // limit - 1 (limit == 0 handled below)
            Bool lim_zero = is_zero_rad(*limit);
            for (size_t k = 0; k < L1; k++) {
                size_t i = reverse_output ? L1 - 1 - k : k;
                Bool ok = Ops::lt(cnt, lim1);
                keep[i] = g.pbs(add(add(chosen[i], ok), bnot(lim_zero)), [](int x) { return x == 3; }, 0, 1);
                cnt = inc_if(cnt, chosen[i]);
            }
        }
        // running piece start S_i (before position i) = i' + len(p) of the last kept match i' < i, else 0
        std::vector<Rad> S(L1 + 1);
        S[0] = usize(0);
        if (serial_keep && !reverse_output) {
            // keep[] is itself a forward serial chain: the running start rides along, one level per position
            for (size_t i = 0; i < L1; i++) {
                Rad target = add_rad_scalar_len(plen, i);       // i + len(p)
                S[i + 1] = if_then_else(keep[i], target, S[i]);
            }
        } else {
            std::vector<Last> leaves(L1);
            for (size_t i = 0; i < L1; i++)
                leaves[i] = Last{keep[i], select_zero(keep[i], add_rad_scalar_len(plen, i))};
            std::vector<Last> pre = last_set_scan(leaves);
            for (size_t i = 0; i < L1; i++) {
                S[i + 1] = pre[i].v;
                for (auto &blk : S[i + 1])
                    if (blk.noise > 24.f) blk = g.pbs(blk, [](int x) { return x & 3; }, 0, 3);
            }
        }
        FSliceVec out;
        out.s = a;
        for (size_t i = 0; i < L1; i++) {
            FSlice e;
            e.is_some = keep[i];
            if (piece_ok) e.is_some = band(keep[i], (*piece_ok)[i]);
            e.start = S[i];
            e.end = (mode == INCLUSIVE) ? add_rad_scalar_len(plen, i) : usize(i);
            out.v.push_back(e);
        }
        FSlice fin;
        fin.start = S[L1];
        fin.end = alen;
        if (mode == SPLIT && !piece_ok) fin.is_some = Graph::constant(1);
        else fin.is_some = Ops::lt(S[L1], alen);                 // drop an empty trailing piece
        if (limit) fin.is_some = band(fin.is_some, bnot(is_zero_rad(*limit)));
        out.v.push_back(fin);
        if (reverse_output) std::reverse(out.v.begin(), out.v.end());
        return out;
    }
    Bool is_zero_rad(const Rad &r) { return is_zero(r); }
    // memoised i + len(p)
    std::map<size_t, Rad> plen_cache;
    const Rad *plen_key = nullptr;
    Rad add_rad_scalar_len(const Rad &plen, size_t i) {
        if (plen_key != &plen) { plen_cache.clear(); plen_key = &plen; }
        auto it = plen_cache.find(i);
        if (it != plen_cache.end()) return it->second;
        Rad r = i == 0 ? plen : add_scalar(plen, i);
        plen_cache[i] = r;
        return r;
    }

    FSliceVec split_pattern(const FStr &a, const FStr &b, Mode mode, bool reverse, const Rad *limit) {
        std::vector<Bool> az = zero_mask(a), bz = zero_mask(b);
        std::vector<Bool> m = find_all(a, b, az, true);
        std::vector<Bool> chosen = non_overlapping(m, bz, reverse);
        Rad plen = len_from_mask(bz);
        return split_engine(a, chosen, plen, mode, reverse, limit, nullptr);
    }
    // split.rs:339-413
    FSliceVec split(const FStr &a, const FStr &b) { return split_pattern(a, b, SPLIT, false, nullptr); }
    FSliceVec rsplit(const FStr &a, const FStr &b) { return split_pattern(a, b, SPLIT, true, nullptr); }
    FSliceVec split_inclusive(const FStr &a, const FStr &b) { return split_pattern(a, b, INCLUSIVE, false, nullptr); }
    FSliceVec split_terminator(const FStr &a, const FStr &b) { return split_pattern(a, b, TERMINATOR, false, nullptr); }
    FSliceVec rsplit_terminator(const FStr &a, const FStr &b) { return split_pattern(a, b, TERMINATOR, true, nullptr); }
    FSliceVec splitn(const FStr &a, const Rad &n, const FStr &b) { return split_pattern(a, b, SPLIT, false, &n); }
    FSliceVec rsplitn(const FStr &a, const Rad &n, const FStr &b) { return split_pattern(a, b, SPLIT, true, &n); }
    // split.rs:415  separators are single whitespace chars, empty pieces are dropped
    FSliceVec split_ascii_whitespace(const FStr &a) {
        std::vector<Bool> ws, ok;
        for (auto &c : a.ch) ws.push_back(is_whitespace(c));
        for (size_t i = 0; i < a.ch.size(); i++) ok.push_back(i == 0 ? Graph::constant(0) : bnot(ws[i - 1]));
        Rad one = usize(1);
        serial_keep = false;
        return split_engine(a, ws, one, SPLIT, false, nullptr, &ok);
    }
    // split.rs:505/511
    struct OnceResult { Bool is_some; FStr first, second; };
    OnceResult split_once(const FStr &a, const FStr &b, bool reverse) {
        FOptUsize f = reverse ? rfind(a, b) : find(a, b);
        Rad after = add_rad(f.val, len(b));
        return OnceResult{f.is_some, substr_to(a, f.val), shift_left(a, after)};
    }
};

}  // namespace fhs_host
