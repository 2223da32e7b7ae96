// add / insert / repeat (reference src/ciphertext/insert.rs:14-124) as level programs.
// The reference walks the OUTPUT position by position with char_at on encrypted indices (and a
// radix rem for repeat).  Here all three are encrypted-length concatenations (replace.hpp concat:
// A + shift_right(B, len(A)), a log-depth barrel shifter), so no division and no per-position
// char_at is needed:
//   insert(a, index, b) = a[..index] ++ b ++ a[index..]
//   add(a, b)           = a ++ b                                   (insert at len(a), insert.rs:14-17)
//   repeat(a, n, l)     = ++_{k : bit k of n set} a^(2^k)          (a^(2^k) by repeated doubling),
//                         which has length n * len(a) by construction, truncated to l chars.
#pragma once
#include "replace.hpp"

namespace fhs_host {

struct InsertOps : ReplaceOps {
    using ReplaceOps::ReplaceOps;

    size_t max_str_len() const { return nb_usize >= 16 ? ~(size_t)0 : (((size_t)1 << (2 * nb_usize)) - 1); }

    Buf buf_of(const FStr &s) {
        Buf b;
        b.true_max = s.ch.size() - 1;
        for (size_t i = 0; i + 1 < s.ch.size(); i++) b.ch.push_back(s.ch[i]);
        b.len = fit(len(s), digits_for(b.true_max));
        return b;
    }
    FStr str_of(const Buf &b, size_t l) {
        FStr out;
        Rad z = zero_char();
        for (size_t j = 0; j < l; j++) out.ch.push_back(j < b.ch.size() ? b.ch[j] : z);
        out.ch.push_back(z);  // terminator
        return out;
    }

    // insert.rs:14
    FStr add_str(const FStr &a, const FStr &b) {
        const size_t l = std::min(a.ch.size() - 1 + b.ch.size() - 1, max_str_len());
        return str_of(concat(buf_of(a), buf_of(b), l), l);
    }

    // insert.rs:48; index <= len(a) (the reference panics otherwise; here the tail is dropped)
    FStr insert(const FStr &a, const Rad &index, const FStr &b) {
        const size_t l = std::min(a.ch.size() - 1 + b.ch.size() - 1, max_str_len());
        Buf head;
        FStr h = substr_to(a, index);
        head.true_max = a.ch.size() - 1;
        for (size_t i = 0; i + 1 < h.ch.size(); i++) head.ch.push_back(h.ch[i]);
        head.len = index;
        FStr t = shift_left(a, index);
        Buf tail = buf_of(a);            // same capacity; its length is not needed afterwards
        for (size_t i = 0; i + 1 < t.ch.size(); i++) tail.ch[i] = t.ch[i];
        Buf hb = concat(head, buf_of(b), l);
        return str_of(concat(hb, tail, l), l);
    }

    // insert.rs:20; clear capacity l as in the reference (clamped to the maximum string length)
    FStr repeat(const FStr &a, const Rad &n, size_t l) {
        l = std::min(l, max_str_len());
        // a^(2^k): the buffer is truncated to l chars, the tracked length is not (replace.hpp Buf), so a
        // full buffer shifts everything appended later out of range
        Buf pw = buf_of(a);
        if (pw.ch.size() > l) pw.ch.resize(l);
        Buf acc;
        bool have = false;
        const size_t nbits = 2 * n.size();
        for (size_t k = 0; k < nbits; k++) {
            Bool bit = (k & 1) ? bit1(n[k / 2]) : bit0(n[k / 2]);
            if (!(bit.is_const() && bit.c == 0)) {
                Buf term;
                term.true_max = pw.true_max;
                for (auto &c : pw.ch) term.ch.push_back(select_zero(bit, c));
                term.len = select_zero(bit, pw.len);
                acc = have ? concat(acc, term, l) : term;
                have = true;
            }
            if (k + 1 < nbits) pw = concat(pw, pw, l);
        }
        if (!have) { acc.true_max = 0; acc.len = Rad{Graph::constant(0)}; }
        return str_of(acc, l);
    }

    // ---- FheStringSliceVector::len / get (split.rs:32-83) ----
    // number of present pieces: sum of the is_some flags (chunks of 3 bools fit a digit, then radix adds)
    Rad slices_len(const FSliceVec &sv) {
        Rad acc = usize(0);
        for (auto &e : sv.v) acc = inc_if(acc, e.is_some);
        return acc;
    }
    // split.rs:86 truncate(i): entry j stays present iff fewer than i present entries precede it
    std::vector<Bool> slices_truncate(const FSliceVec &sv, const Rad &i) {
        Rad n = usize(0);
        std::vector<Bool> out;
        for (auto &e : sv.v) {
            out.push_back(band(Ops::lt(n, i), e.is_some));
            n = inc_if(n, e.is_some);
        }
        return out;
    }
    // the i-th present piece: one-hot over the entries by (is_some_j && #present before j == i)
    FOptStr slices_get(const FSliceVec &sv, const Rad &i) {
        Rad n = usize(0);
        std::vector<Bool> hit;
        std::vector<Rad> st, en;
        for (auto &e : sv.v) {
            Bool sel = band(e.is_some, Ops::eq(i, n));
            hit.push_back(sel);
            st.push_back(select_zero(sel, e.start));
            en.push_back(select_zero(sel, e.end));
            n = inc_if(n, e.is_some);
        }
        Rad start = onehot_sum(st, nb_usize), end = onehot_sum(en, nb_usize);
        // mod.rs substr_end(start, end) = chars [start, end): keep i < end, then shift left by start
        FStr piece = shift_left(substr_to(sv.s, end), start);
        return FOptStr{any(hit), piece};
    }
};

}  // namespace fhs_host
