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
        b.max_len = s.ch.size() - 1;
        for (size_t i = 0; i + 1 < s.ch.size(); i++) b.ch.push_back(s.ch[i]);
        b.len = len(s);
        return b;
    }
    FStr str_of(const Buf &b, size_t l) {
        FStr out;
        Rad z = zero_char();
        for (size_t j = 0; j < l; j++) out.ch.push_back(j < b.ch.size() ? b.ch[j] : z);
        out.ch.push_back(z);  // terminator
        return out;
    }

    // min(x + y, cap) without wrapping mod 4^nb_usize: one extra (most significant) block
    Rad add_len_clamped(const Rad &x, const Rad &y, size_t cap) {
        Rad xe = x, ye = y;
        xe.push_back(Graph::constant(0));
        ye.push_back(Graph::constant(0));
        Rad s = add_rad(xe, ye);
        Bool over = gt_scalar(s, cap);
        Rad r = if_then_else(over, scalar(cap, s.size()), s);
        r.pop_back();    // <= cap < 4^nb_usize
        for (auto &blk : r) blk = g.pbs(blk, [](int v) { return v & 3; }, 0, 3);   // fresh digits for the shifter
        return r;
    }
    // replace.hpp concat with the tracked length clamped to the capacity, so that a full buffer
    // shifts everything appended later out of range instead of wrapping around
    Buf concat_c(const Buf &A, const Buf &B, size_t limit) {
        Buf R;
        R.max_len = std::min(A.max_len + B.max_len, limit);
        const size_t cap = R.max_len;
        std::vector<Rad> bs = shift_right(B.ch, A.len, A.max_len, cap);
        for (size_t j = 0; j < cap; j++) R.ch.push_back(j < A.ch.size() ? add_chars(A.ch[j], bs[j]) : bs[j]);
        R.len = add_len_clamped(A.len, B.len, cap);
        return R;
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
        head.max_len = a.ch.size() - 1;
        for (size_t i = 0; i + 1 < h.ch.size(); i++) head.ch.push_back(h.ch[i]);
        head.len = index;
        FStr t = shift_left(a, index);
        Buf tail = buf_of(a);            // same capacity; its length is not needed afterwards
        for (size_t i = 0; i + 1 < t.ch.size(); i++) tail.ch[i] = t.ch[i];
        Buf hb = concat_c(head, buf_of(b), l);
        return str_of(concat(hb, tail, l), l);
    }

    // insert.rs:20; clear capacity l as in the reference (clamped to the maximum string length)
    FStr repeat(const FStr &a, const Rad &n, size_t l) {
        l = std::min(l, max_str_len());
        Buf pw = buf_of(a);              // a^(2^k), truncated to l chars, length clamped to its capacity
        if (pw.max_len > l) {
            pw.ch.resize(l);
            pw.max_len = l;
            pw.len = add_len_clamped(pw.len, usize(0), l);
        }
        Buf acc;
        bool have = false;
        const size_t nbits = 2 * n.size();
        for (size_t k = 0; k < nbits; k++) {
            Bool bit = (k & 1) ? bit1(n[k / 2]) : bit0(n[k / 2]);
            if (!(bit.is_const() && bit.c == 0)) {
                Buf term;
                term.max_len = pw.max_len;
                for (auto &c : pw.ch) term.ch.push_back(select_zero(bit, c));
                term.len = select_zero(bit, pw.len);
                acc = have ? concat_c(acc, term, l) : term;
                have = true;
            }
            if (k + 1 < nbits) pw = concat_c(pw, pw, l);
        }
        if (!have) { acc.max_len = 0; acc.len = usize(0); }
        return str_of(acc, l);
    }

    // ---- FheStringSliceVector::len / get (split.rs:32-83) ----
    // number of present pieces: sum of the is_some flags (chunks of 3 bools fit a digit, then radix adds)
    Rad slices_len(const FSliceVec &sv) {
        Rad acc = usize(0);
        for (auto &e : sv.v) acc = inc_if(acc, e.is_some);
        return acc;
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
