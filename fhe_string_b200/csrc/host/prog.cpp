// fhs_prog: build / export / run level programs (include/fhs_string.h).
#include <algorithm>
#include <atomic>
#include <chrono>
#include <climits>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <memory>
#include <string>
#include <vector>

#include "../../../include/fhs_string.h"
#include "insert.hpp"

using namespace fhs_host;

static std::atomic<uint64_t> g_prog_serial{1};

struct fhs_prog {
    const uint64_t serial = g_prog_serial++;   // owner id of this program's cached executions (fhs_ctx_cache_*)
    Graph g;
    std::vector<int64_t> outputs;  // node id, or -1-c for a clear constant c
    std::string desc;
    size_t n_inputs = 0;
    float max_output_noise = 0.f;  // largest variance multiplier (PBS outputs) an output block carries
};

namespace {

thread_local std::string g_host_err;

struct Builder {
    fhs_prog &P;
    InsertOps ops;
    Builder(fhs_prog &p, size_t nbc, size_t nbu) : P(p), ops(p.g, nbc, nbu) {}

    FStr in_str(size_t n_chars) {
        FStr s;
        for (size_t i = 0; i < n_chars; i++) {
            Rad c;
            for (size_t b = 0; b < ops.nb_char; b++) c.push_back(P.g.input(3));
            s.ch.push_back(c);
        }
        return s;
    }
    Rad in_usize() {
        Rad r;
        for (size_t b = 0; b < ops.nb_usize; b++) r.push_back(P.g.input(3));
        return r;
    }
    void out_lin(const Lin &e) {
        if (e.is_const()) { P.outputs.push_back(-1 - (e.c & 31)); return; }
        if (e.noise > P.max_output_noise) P.max_output_noise = e.noise;
        P.outputs.push_back(P.g.materialize(e));
    }
    void out_rad(const Rad &r) { for (auto &b : r) out_lin(b); }
    void out_str(const FStr &s) { for (auto &c : s.ch) out_rad(c); }
};

int build_op(const std::string &op, Builder &B, const uint32_t *n_chars, size_t n_strings,
             size_t n_usize_args, const uint64_t *scalars, size_t n_scalars) {
    InsertOps &o = B.ops;
    std::vector<FStr> S;
    for (size_t i = 0; i < n_strings; i++) {
        if (n_chars[i] == 0) { g_host_err = "a string has at least its terminator"; return FHS_ERR_INVALID; }
        S.push_back(B.in_str(n_chars[i]));
    }
    std::vector<Rad> U;
    for (size_t i = 0; i < n_usize_args; i++) U.push_back(B.in_usize());
    // slices_len / slices_get take a FheStringSliceVector exactly as split* returns it: the string,
    // then scalars[0] entries of (is_some, start, end); slices_get is followed by the index usize.
    FSliceVec SV;
    if (op == "slices_len" || op == "slices_get" || op == "slices_truncate") {
        if (n_strings != 1 || n_usize_args != 0 || n_scalars < 1) { g_host_err = "wrong operand count for op " + op; return FHS_ERR_INVALID; }
        SV.s = S[0];
        for (uint64_t e = 0; e < scalars[0]; e++) {
            FSlice sl;
            sl.is_some = B.P.g.input(1);
            sl.start = B.in_usize();
            sl.end = B.in_usize();
            SV.v.push_back(sl);
        }
        if (op == "slices_get" || op == "slices_truncate") U.push_back(B.in_usize());
    }
    B.P.n_inputs = B.P.g.nodes.size();
    auto need = [&](size_t ns, size_t nu) { return n_strings == ns && n_usize_args == nu; };
    auto bad = [&]() { g_host_err = "wrong operand count for op " + op; return FHS_ERR_INVALID; };
    auto str_desc = [](const FStr &s) { return "str:" + std::to_string(s.ch.size()); };

    if (op == "is_empty") { if (!need(1, 0)) return bad(); B.out_lin(o.is_empty(S[0])); B.P.desc = "bool"; }
    else if (op == "len") { if (!need(1, 0)) return bad(); B.out_rad(o.len(S[0])); B.P.desc = "usize"; }
    else if (op == "eq" || op == "ne" || op == "lt" || op == "le" || op == "gt" || op == "ge" ||
             op == "eq_ignore_ascii_case" || op == "contains" || op == "starts_with" || op == "ends_with") {
        if (!need(2, 0)) return bad();
        Bool r = op == "eq" ? o.eq(S[0], S[1]) : op == "ne" ? o.ne(S[0], S[1]) : op == "lt" ? o.lt(S[0], S[1])
               : op == "le" ? o.le(S[0], S[1]) : op == "gt" ? o.gt(S[0], S[1]) : op == "ge" ? o.ge(S[0], S[1])
               : op == "eq_ignore_ascii_case" ? o.eq_ignore_ascii_case(S[0], S[1])
               : op == "contains" ? o.contains(S[0], S[1]) : op == "starts_with" ? o.starts_with(S[0], S[1])
               : o.ends_with(S[0], S[1]);
        B.out_lin(r); B.P.desc = "bool";
    }
    else if (op == "find" || op == "rfind") {
        if (!need(2, 0)) return bad();
        FOptUsize r = op == "find" ? o.find(S[0], S[1]) : o.rfind(S[0], S[1]);
        B.out_lin(r.is_some); B.out_rad(r.val); B.P.desc = "opt_usize";
    }
    else if (op == "to_uppercase" || op == "to_lowercase" || op == "trim" || op == "trim_start" || op == "trim_end") {
        if (!need(1, 0)) return bad();
        FStr r = op == "to_uppercase" ? o.to_uppercase(S[0]) : op == "to_lowercase" ? o.to_lowercase(S[0])
               : op == "trim" ? o.trim(S[0]) : op == "trim_start" ? o.trim_start(S[0]) : o.trim_end(S[0]);
        B.out_str(r); B.P.desc = str_desc(r);
    }
    else if (op == "strip_prefix" || op == "strip_suffix") {
        if (!need(2, 0)) return bad();
        FOptStr r = op == "strip_prefix" ? o.strip_prefix(S[0], S[1]) : o.strip_suffix(S[0], S[1]);
        B.out_lin(r.is_some); B.out_str(r.val); B.P.desc = "opt_" + str_desc(r.val);
    }
    else if (op == "char_is_uppercase" || op == "char_is_lowercase" || op == "char_is_whitespace") {
        // FheAsciiChar predicates (convert.rs:14,22, trim.rs:15): operand = one char (a "string" of 1 char)
        if (!need(1, 0) || S[0].ch.size() != 1) return bad();
        B.out_lin(op == "char_is_whitespace" ? o.is_whitespace(S[0].ch[0]) : o.is_case(S[0].ch[0], op == "char_is_uppercase"));
        B.P.desc = "bool";
    }
    else if (op == "char_at") {          // mod.rs:217
        if (!need(1, 1)) return bad();
        B.out_rad(o.char_at(S[0], U[0])); B.P.desc = "char";
    }
    else if (op == "substr_end") {       // mod.rs:198  operands: a; usize start, end
        if (!need(1, 2)) return bad();
        FStr r = o.substr_end(S[0], U[0], U[1]);
        B.out_str(r); B.P.desc = str_desc(r);
    }
    else if (op == "substr_eq") {        // compare.rs:119  operands: a, b; scalars[0] = clear offset i
        if (!need(2, 0) || n_scalars < 1) return bad();
        B.out_lin(o.substr_eq(S[0], (size_t)scalars[0], S[1])); B.P.desc = "bool";
    }
    else if (op == "substr_to" || op == "substr_from") {
        if (!need(1, 1)) return bad();
        FStr r = op == "substr_to" ? o.substr_to(S[0], U[0]) : o.substr_from(S[0], U[0]);
        B.out_str(r); B.P.desc = str_desc(r);
    }
    else if (op == "split" || op == "rsplit" || op == "split_inclusive" || op == "split_terminator" ||
             op == "rsplit_terminator" || op == "splitn" || op == "rsplitn" || op == "split_ascii_whitespace") {
        FSliceVec r;
        if (op == "split_ascii_whitespace") { if (!need(1, 0)) return bad(); r = o.split_ascii_whitespace(S[0]); }
        else if (op == "splitn" || op == "rsplitn") {
            if (!need(2, 1)) return bad();
            r = op == "splitn" ? o.splitn(S[0], U[0], S[1]) : o.rsplitn(S[0], U[0], S[1]);
        } else {
            if (!need(2, 0)) return bad();
            r = op == "split" ? o.split(S[0], S[1]) : op == "rsplit" ? o.rsplit(S[0], S[1])
              : op == "split_inclusive" ? o.split_inclusive(S[0], S[1])
              : op == "split_terminator" ? o.split_terminator(S[0], S[1]) : o.rsplit_terminator(S[0], S[1]);
        }
        B.out_str(r.s);
        for (auto &e : r.v) { B.out_lin(e.is_some); B.out_rad(e.start); B.out_rad(e.end); }
        B.P.desc = "slices:" + std::to_string(r.v.size()) + ":" + str_desc(r.s);
    }
    else if (op == "split_once" || op == "rsplit_once") {
        if (!need(2, 0)) return bad();
        auto r = o.split_once(S[0], S[1], op == "rsplit_once");
        B.out_lin(r.is_some); B.out_str(r.first); B.out_str(r.second);
        B.P.desc = "opt_pair:" + std::to_string(r.first.ch.size()) + ":" + std::to_string(r.second.ch.size());
    }
    else if (op == "replace" || op == "replacen") {
        // operands: a, p, s [, n]; scalars[0] = l, the output capacity (replace.rs:17,23), clamped
        // to the maximum string length like replace.rs:45
        if (!(n_strings == 3 && n_usize_args == (op == "replacen" ? 1u : 0u)) || n_scalars < 1) return bad();
        uint64_t maxlen = (o.nb_usize >= 16) ? ~0ull : ((1ull << (2 * o.nb_usize)) - 1);
        size_t l = (size_t)std::min<uint64_t>(scalars[0], maxlen);
        FStr r = op == "replace" ? o.replace(S[0], S[1], S[2], l) : o.replacen(S[0], S[1], S[2], U[0], l);
        B.out_str(r); B.P.desc = str_desc(r);
    }
    else if (op == "slices_len") { B.out_rad(o.slices_len(SV)); B.P.desc = "usize"; }
    else if (op == "slices_truncate") {   // split.rs:86: the new is_some flags, entry order unchanged
        for (auto &b : o.slices_truncate(SV, U[0])) B.out_lin(b);
        B.P.desc = "bools:" + std::to_string(SV.v.size());
    }
    else if (op == "slices_get") {
        FOptStr r = o.slices_get(SV, U[0]);
        B.out_lin(r.is_some); B.out_str(r.val); B.P.desc = "opt_" + str_desc(r.val);
    }
    else if (op == "add") {           // insert.rs:14  operands: a, b
        if (!need(2, 0)) return bad();
        FStr r = o.add_str(S[0], S[1]);
        B.out_str(r); B.P.desc = str_desc(r);
    }
    else if (op == "insert") {        // insert.rs:48  operands: a, s; usize index
        if (!need(2, 1)) return bad();
        FStr r = o.insert(S[0], U[0], S[1]);
        B.out_str(r); B.P.desc = str_desc(r);
    }
    else if (op == "repeat") {        // insert.rs:20  operands: a; usize n; scalars[0] = l
        if (!need(1, 1) || n_scalars < 1) return bad();
        FStr r = o.repeat(S[0], U[0], (size_t)scalars[0]);
        B.out_str(r); B.P.desc = str_desc(r);
    }
    else { g_host_err = "unknown op " + op; return FHS_ERR_INVALID; }
    return FHS_OK;
}

}  // namespace

extern "C" {

int fhs_prog_build(const char *op, const uint32_t *n_chars, size_t n_strings, size_t n_usize_args,
                   const uint64_t *scalars, size_t n_scalars, uint32_t nbc, uint32_t nbu,
                   fhs_prog **out) {
    if (!op || !out || (n_strings && !n_chars) || nbc != 4 || nbu == 0 || nbu > 16) {
        g_host_err = "fhs_prog_build: bad arguments (num_blocks_char must be 4)";
        return FHS_ERR_INVALID;
    }
    *out = nullptr;
    fhs_prog *p = new fhs_prog();
    try {
        Builder B(*p, nbc, nbu);
        int rc = build_op(op, B, n_chars, n_strings, n_usize_args, scalars, n_scalars);
        if (rc != FHS_OK) { delete p; return rc; }
        p->g.prune(p->outputs);     // evaluations no output depends on (duplicates are merged at creation)
    } catch (const std::exception &e) {
        g_host_err = std::string("building ") + op + ": " + e.what();
        delete p;
        return FHS_ERR_INVALID;
    }
    *out = p;
    return FHS_OK;
}

void fhs_prog_destroy(fhs_prog *p) {
    if (!p) return;
    fhs_cache_drop_owner(p->serial);   // pools / staging parked on any live context
    delete p;
}

const char *fhs_host_last_error(void) { return g_host_err.c_str(); }

int fhs_prog_info(const fhs_prog *p, uint64_t info[8]) {
    if (!p || !info) return FHS_ERR_INVALID;
    info[0] = p->g.nodes.size();
    info[1] = p->n_inputs;
    info[2] = p->outputs.size();
    info[3] = p->g.terms.size();
    info[4] = p->g.luts.size();
    info[5] = (uint64_t)p->g.depth();
    info[6] = (uint64_t)p->g.pbs_count;
    info[7] = (uint64_t)(p->g.max_input_noise + 0.999f);
    return FHS_OK;
}

const char *fhs_prog_output_desc(const fhs_prog *p) { return p ? p->desc.c_str() : ""; }

int fhs_prog_export(const fhs_prog *p, uint8_t *kind, int32_t *lut, uint32_t *t0, uint32_t *tn,
                    int64_t *cst, int32_t *level, int32_t *term_src, int32_t *term_coef,
                    uint64_t *luts, int64_t *outputs) {
    if (!p) return FHS_ERR_INVALID;
    const Graph &g = p->g;
    for (size_t i = 0; i < g.nodes.size(); i++) {
        const Node &n = g.nodes[i];
        kind[i] = n.kind; lut[i] = n.lut; t0[i] = n.t0; tn[i] = n.tn; cst[i] = n.c; level[i] = n.level;
    }
    for (size_t i = 0; i < g.terms.size(); i++) { term_src[i] = g.terms[i].src; term_coef[i] = g.terms[i].coef; }
    for (size_t i = 0; i < g.luts.size(); i++) memcpy(luts + 16 * i, g.luts[i].data(), 16 * sizeof(uint64_t));
    for (size_t i = 0; i < p->outputs.size(); i++) outputs[i] = p->outputs[i];
    return FHS_OK;
}

// Value ranges the builder relied on, per node: in_lo / in_hi bound the linear combination that feeds a
// PBS (or is materialised), lo / hi bound the node's output, signed_in marks PBS nodes whose input may be
// negative.  The test suite's clear-text evaluator asserts them, so that a wrong `bounded()`
// override in the op builders fails at the node where it happens.
double fhs_prog_output_noise(const fhs_prog *p) { return p ? (double)p->max_output_noise : -1.0; }

int fhs_prog_export_ranges(const fhs_prog *p, int32_t *in_lo, int32_t *in_hi, int32_t *out_lo, int32_t *out_hi,
                           uint8_t *signed_in) {
    if (!p || !in_lo || !in_hi || !out_lo || !out_hi || !signed_in) return FHS_ERR_INVALID;
    const Graph &g = p->g;
    for (size_t i = 0; i < g.nodes.size(); i++) {
        const Node &n = g.nodes[i];
        in_lo[i] = n.in_lo; in_hi[i] = n.in_hi; out_lo[i] = n.lo; out_hi[i] = n.hi; signed_in[i] = n.signed_in;
    }
    return FHS_OK;
}

}  // extern "C"

// ------------------------------------------------------------------------------------------------
// Executor: level by level on a device-resident pool (fhs_pool / fhs_level of include/fhs_cuda.h).

// Slice of a level of B evaluations owned by `rank`.  A level of at most shard_min evaluations (74 =
// one wave of the pair blind-rotation kernel, 1.7 ms however few ciphertexts it has) is computed
// redundantly by every rank; anything larger is split, because ceil(B / world) ciphertexts per GPU
// finish sooner than B on one GPU even after the all-gather (B = 148: 2.3 ms on one GPU, 1.7 ms +
// exchange on two; B = 592: 6.2 ms on one GPU, 1.7 ms + exchange on eight).
static void level_slice(size_t B, int rank, int world, size_t shard_min, size_t *per, size_t *lo, size_t *hi) {
    if (world <= 1 || B <= shard_min) { *per = 0; *lo = 0; *hi = B; return; }
    *per = (B + world - 1) / world;
    *lo = std::min(B, *per * (size_t)rank);
    *hi = std::min(B, *per * (size_t)(rank + 1));
}

// ---- level balancing ---------------------------------------------------------------------------------
// A level costs what its WAVES cost, not what its evaluations cost: on a B200 (148 SMs) one blind-rotation
// wave takes 1.65 ms for up to 74 ciphertexts (pair kernel), 2.3 ms up to 148 and 3.6 ms up to 296 (wide
// kernel), 5.28 ms up to 370 (wide + pair) and 5.9 ms per round of 592 (throughput kernel); measured, DESIGN.md
// section 4.  "As soon as possible" levels therefore waste time whenever a level sits just above a wave
// boundary while some of its evaluations are not needed until later.  balance_levels keeps the depth and moves
// evaluations with slack to the next level where that lowers the modelled time of the two levels, a few
// sweeps front to back.  Every rank runs the same deterministic pass on the same graph, so the plans agree.
static double wave_ms(size_t b) {
    if (b == 0) return 0.0;
    const size_t full = b / 592, r = b % 592;
    double t = 5.9 * (double)full;
    if (r == 0) t += 0.0;
    else if (r <= 74) t += 1.65;
    else if (r <= 148) t += 2.3;
    else if (r <= 296) t += 3.6;
    else if (r <= 370) t += 5.28;
    else t += 5.9;
    return t + 0.05;            // gather + digit + keyswitch launches of the level
}
static double level_ms(size_t B, int world, size_t shard_min) {
    if (world <= 1 || B <= shard_min) return wave_ms(B);
    return wave_ms((B + world - 1) / world) + 0.02;   // + exchange
}

// returns the level of every node (inputs 0), starting from the levels the graph builder assigned
static std::vector<int32_t> balance_levels(const Graph &g, int world, size_t shard_min) {
    const size_t N = g.nodes.size();
    std::vector<int32_t> asap(N);
    for (size_t i = 0; i < N; i++) asap[i] = g.nodes[i].level;
    const int D = g.depth();
    const char *knob = getenv("FHS_LEVEL_BALANCE");
    if ((knob && atoi(knob) == 0) || D < 2) return asap;
    // LUT evaluations that read node i, directly or through linear nodes (nodes are in topological order)
    std::vector<std::vector<int32_t>> readers(N);
    for (size_t c = N; c-- > 0;) {
        const Node &n = g.nodes[c];
        for (uint32_t t = n.t0; t < n.t0 + n.tn; t++) {
            std::vector<int32_t> &r = readers[(size_t)g.terms[t].src];
            if (n.kind == K_PBS) r.push_back((int32_t)c);
            else r.insert(r.end(), readers[c].begin(), readers[c].end());
        }
    }
    for (auto &r : readers) { std::sort(r.begin(), r.end()); r.erase(std::unique(r.begin(), r.end()), r.end()); }
    const size_t unit = world > 1 ? (size_t)world : 1;
    // movable evaluations are taken in order of the room they still have (earliest reader furthest away first),
    // so that what moves now can move again later; ties by node id
    bool use_room = false;      // the whole search runs twice, with and without this ordering; the better one is kept
    auto by_room = [&](std::vector<int32_t> &mv, const std::vector<int32_t> &lv) {
        if (!use_room) return;
        std::vector<std::pair<int32_t, int32_t>> key;
        key.reserve(mv.size());
        for (int32_t i : mv) {
            int32_t first = INT32_MAX;
            for (int32_t c : readers[i]) first = std::min(first, lv[c]);
            key.emplace_back(-first, i);
        }
        std::sort(key.begin(), key.end());
        for (size_t q = 0; q < mv.size(); q++) mv[q] = key[q].second;
    };
    auto targets = [&](size_t b) {
        std::vector<size_t> v;
        static const size_t tiers[5] = {74, 148, 296, 370, 592};
        for (size_t th : tiers)
            for (size_t k = 0; k < 8; k++) {
                const size_t x = (th + 592 * k) * unit;
                if (x < b) v.push_back(x);
            }
        if (unit > 1 && shard_min < b) v.push_back(shard_min);
        v.push_back(0);
        return v;
    };
    auto total_ms = [&](const std::vector<int32_t> &lvl) {
        std::vector<size_t> cnt(D + 2, 0);
        for (size_t i = 0; i < N; i++) if (g.nodes[i].kind == K_PBS) cnt[lvl[i]]++;
        double t = 0;
        for (int l = 1; l <= D; l++) t += level_ms(cnt[l], world, shard_min);
        return t;
    };
    // one strategy: sweeps over the levels (front to back, or back to front so that readers make room first);
    // an evaluation may move from level l to l + 1 when every LUT that reads it runs at l + 2 or later
    auto run = [&](bool back_first) {
        std::vector<int32_t> lvl = asap;
        std::vector<std::vector<int32_t>> at(D + 2);
        for (size_t i = 0; i < N; i++)
            if (g.nodes[i].kind == K_PBS) at[lvl[i]].push_back((int32_t)i);
        for (int sweep = 0; sweep < 6; sweep++) {
            bool any = false;
            for (int step = 1; step < D; step++) {
                const int l = back_first ? D - step : step;
                std::vector<int32_t> &cur = at[l], &nxt = at[l + 1];
                const size_t b = cur.size();
                if (b == 0) continue;
                std::vector<int32_t> mv;
                for (int32_t i : cur) {
                    bool free_to_move = true;
                    for (int32_t c : readers[i]) if (lvl[c] < l + 2) { free_to_move = false; break; }
                    if (free_to_move) mv.push_back(i);
                }
                if (mv.empty()) continue;
                const double before = level_ms(b, world, shard_min) + level_ms(nxt.size(), world, shard_min);
                double best = 1e-9;
                size_t best_k = 0;
                for (size_t tgt : targets(b)) {
                    const size_t k = b - tgt;
                    if (k == 0 || k > mv.size()) continue;
                    const double gain = before - level_ms(b - k, world, shard_min) - level_ms(nxt.size() + k, world, shard_min);
                    if (gain > best) { best = gain; best_k = k; }
                }
                if (!best_k) continue;
                by_room(mv, lvl);
                mv.resize(best_k);
                std::sort(mv.begin(), mv.end());        // ascending node ids
                std::vector<int32_t> keep;
                keep.reserve(b - best_k);
                for (int32_t i : cur)
                    if (!std::binary_search(mv.begin(), mv.end(), i)) keep.push_back(i);
                for (int32_t i : mv) { lvl[i] = l + 1; nxt.push_back(i); }
                std::sort(nxt.begin(), nxt.end());      // execution order inside a level stays the node order
                cur.swap(keep);
                any = true;
            }
            if (!any) break;
        }
        return lvl;
    };
    std::vector<int32_t> final_lvl = asap;
    double final_ms = total_ms(asap);
    for (int order = 0; order < 2; order++) {
    use_room = order == 1;
    std::vector<int32_t> lvl = asap;
    double best_ms = total_ms(asap);
    for (int strategy = 0; strategy < 2; strategy++) {
        std::vector<int32_t> cand = run(strategy == 1);
        const double ms = total_ms(cand);
        if (ms < best_ms - 1e-9) { best_ms = ms; lvl.swap(cand); }
    }
    // Second pass on the better schedule: two-level lookahead.  Moving k evaluations from level l to l + 1 may only
    // pay once k2 evaluations of the (grown) level l + 1 move on to l + 2; the pairwise sweeps above never take
    // the first half of such a move.  Same validity rule, gain measured over the three levels.
    {
        std::vector<std::vector<int32_t>> at(D + 3);
        for (size_t i = 0; i < N; i++)
            if (g.nodes[i].kind == K_PBS) at[lvl[i]].push_back((int32_t)i);
        auto free_at = [&](int32_t i, int l) {       // may node i (at level l) move to l + 1 ?
            for (int32_t c : readers[i]) if (lvl[c] < l + 2) return false;
            return true;
        };
        for (int round = 0; round < 4; round++) {
            bool any = false;
            for (int l = 1; l + 2 <= D; l++) {
                std::vector<int32_t> mv;
                for (int32_t i : at[l]) if (free_at(i, l)) mv.push_back(i);
                if (mv.empty()) continue;
                by_room(mv, lvl);
                const size_t b0 = at[l].size(), b1 = at[l + 1].size(), b2 = at[l + 2].size();
                const double before = level_ms(b0, world, shard_min) + level_ms(b1, world, shard_min) + level_ms(b2, world, shard_min);
                double best = 1e-9;
                size_t best_k = 0, best_k2 = 0;
                for (size_t tgt : targets(b0)) {
                    const size_t k = b0 - tgt;
                    if (k == 0 || k > mv.size()) continue;
                    for (size_t q = 0; q < k; q++) lvl[mv[q]] = l + 1;          // tentatively
                    size_t n2 = 0;                                               // how many of level l + 1 could move on
                    for (int32_t i : at[l + 1]) if (free_at(i, l + 1)) n2++;
                    for (size_t q = 0; q < k; q++) if (free_at(mv[q], l + 1)) n2++;
                    const size_t b1k = b1 + k;
                    std::vector<size_t> t2 = targets(b1k);
                    t2.push_back(b1k);                                           // k2 = 0
                    for (size_t tgt2 : t2) {
                        const size_t k2 = b1k - tgt2;
                        if (k2 > n2) continue;
                        const double gain = before - level_ms(b0 - k, world, shard_min) - level_ms(b1k - k2, world, shard_min) -
                                            level_ms(b2 + k2, world, shard_min);
                        if (gain > best) { best = gain; best_k = k; best_k2 = k2; }
                    }
                    for (size_t q = 0; q < k; q++) lvl[mv[q]] = l;               // undo
                }
                if (!best_k) continue;
                std::vector<int32_t> moved(mv.begin(), mv.begin() + best_k), keep;
                std::sort(moved.begin(), moved.end());
                for (int32_t i : at[l]) if (!std::binary_search(moved.begin(), moved.end(), i)) keep.push_back(i);
                at[l].swap(keep);
                for (int32_t i : moved) { lvl[i] = l + 1; at[l + 1].push_back(i); }
                std::sort(at[l + 1].begin(), at[l + 1].end());
                if (best_k2) {
                    std::vector<int32_t> mv2, keep2;
                    for (int32_t i : at[l + 1]) if (free_at(i, l + 1)) mv2.push_back(i);
                    by_room(mv2, lvl);
                    mv2.resize(best_k2);
                    std::sort(mv2.begin(), mv2.end());
                    for (int32_t i : at[l + 1]) if (!std::binary_search(mv2.begin(), mv2.end(), i)) keep2.push_back(i);
                    at[l + 1].swap(keep2);
                    for (int32_t i : mv2) { lvl[i] = l + 2; at[l + 2].push_back(i); }
                    std::sort(at[l + 2].begin(), at[l + 2].end());
                }
                any = true;
            }
            if (!any) break;
        }
        best_ms = total_ms(lvl);
    }
    if (best_ms < final_ms - 1e-9) { final_ms = best_ms; final_lvl.swap(lvl); }
    }
    std::vector<int32_t> lvl = final_lvl;
    const double best_ms = final_ms;
    // linear nodes follow their sources; then check every edge (fall back to the builder's levels on any doubt)
    for (size_t i = 0; i < N; i++) {
        const Node &n = g.nodes[i];
        if (n.kind != K_LIN) continue;
        int32_t l = 0;
        for (uint32_t t = n.t0; t < n.t0 + n.tn; t++) l = std::max(l, lvl[g.terms[t].src]);
        lvl[i] = l;
    }
    bool ok = true;
    for (size_t i = 0; i < N && ok; i++) {
        const Node &n = g.nodes[i];
        for (uint32_t t = n.t0; t < n.t0 + n.tn && ok; t++) {
            const int32_t ls = lvl[g.terms[t].src];
            ok = n.kind == K_PBS ? ls < lvl[i] : ls <= lvl[i];
        }
        ok = ok && lvl[i] <= D;
    }
    if (getenv("FHS_BALANCE_DEBUG"))
        fprintf(stderr, "balance_levels: world %d, %zu nodes, depth %d: %.2f -> %.2f ms (model)%s\n", world, N, D,
                total_ms(asap), best_ms, ok ? "" : "  INVALID, falling back");
    return ok ? lvl : asap;
}

// The level every node runs at on `world` GPUs after balance_levels (what build_exec executes), and the
// modelled time of the as-soon-as-possible and of the balanced schedule: for tests and for tools.
extern "C" int fhs_prog_export_schedule(const fhs_prog *p, int world, size_t shard_min, int32_t *level, double model_ms[2]) {
    if (!p || !level || world < 1) return FHS_ERR_INVALID;
    if (shard_min == 0) shard_min = 74;
    const Graph &g = p->g;
    const std::vector<int32_t> lvl = balance_levels(g, world, shard_min);
    const int D = g.depth();
    std::vector<size_t> c0(D + 2, 0), c1(D + 2, 0);
    for (size_t i = 0; i < g.nodes.size(); i++) {
        level[i] = lvl[i];
        if (g.nodes[i].kind == K_PBS) { c0[g.nodes[i].level]++; c1[lvl[i]]++; }
    }
    if (model_ms) {
        model_ms[0] = model_ms[1] = 0.0;
        for (int l = 1; l <= D; l++) { model_ms[0] += level_ms(c0[l], world, shard_min); model_ms[1] += level_ms(c1[l], world, shard_min); }
    }
    return FHS_OK;
}

extern "C" int fhs_prog_plan(const fhs_prog *p, int rank, int world, size_t shard_min, uint64_t *levels_out,
                             size_t cap_levels, size_t *n_levels) {
    // levels_out[4*l + {0,1,2,3}] = B, blocks per rank (0 = replicated), lo, hi of this rank
    if (!p || !n_levels || world < 1 || rank < 0 || rank >= world) return FHS_ERR_INVALID;
    const Graph &g = p->g;
    const int D = g.depth();
    std::vector<size_t> cnt(D + 1, 0);
    const std::vector<int32_t> lvl = balance_levels(g, world, shard_min ? shard_min : 74);
    for (size_t i = 0; i < g.nodes.size(); i++) if (g.nodes[i].kind == K_PBS) cnt[lvl[i]]++;
    size_t k = 0;
    for (int l = 1; l <= D; l++) {
        size_t per, lo, hi;
        if (cnt[l] == 0) continue;          // a level the balancing pass emptied
        level_slice(cnt[l], rank, world, shard_min, &per, &lo, &hi);
        if (levels_out && k < cap_levels) { levels_out[4 * k] = cnt[l]; levels_out[4 * k + 1] = per; levels_out[4 * k + 2] = lo; levels_out[4 * k + 3] = hi; }
        k++;
    }
    *n_levels = k;
    return FHS_OK;
}

// Execution state of one program on one context (and one rank / world / shard threshold): the level
// plan, the device pool with the descriptors of every level resident, and page-locked staging buffers.
// Built at the first call and parked in the context's attachment cache (fhs_ctx_cache_*), so that every
// later call of the same operation shape is transfers + kernel launches only: no re-planning, no pool
// allocation, no descriptor upload (round 1 redid all three per call).
namespace {

struct Grp { size_t row0, rows, out_base, per, lo, hi; bool is_pbs; };

struct Exec {
    fhs_ctx *ctx = nullptr;
    fhs_pool *pool = nullptr;
    std::vector<Grp> groups;
    Grp OG{};
    size_t n_inputs = 0, n_out = 0, n_pbs = 0, n_levels = 0, n_sharded = 0, width = 0, slots = 0;
    uint64_t *h_in = nullptr, *h_out = nullptr;   // page-locked staging for pageable caller buffers
    ~Exec() {
        if (pool) fhs_pool_destroy(pool);
        if (h_in) fhs_host_free(ctx, h_in);
        if (h_out) fhs_host_free(ctx, h_out);
    }
};

void exec_dtor(void *e) { delete static_cast<Exec *>(e); }

int build_exec(const fhs_prog *p, fhs_ctx *ctx, int rank, int world, size_t shard_min, bool own_comm, Exec **out) {
    const Graph &g = p->g;
    fhs_params prm;
    int rc = fhs_ctx_get_params(ctx, &prm);
    if (rc != FHS_OK) return rc;
    const size_t width = (size_t)prm.glwe_dim * prm.poly_size + 1;
    const uint64_t delta = (1ull << 63) / ((uint64_t)prm.message_modulus * prm.carry_modulus);
    const size_t N = g.nodes.size();
    const int D = g.depth();
    // slots: inputs first, then per level the PBS nodes, then the linear outputs of that level,
    // finally the program outputs gathered into consecutive slots (one read-back at the end)
    std::vector<std::vector<int32_t>> pbs_at(D + 1), lin_at(D + 1);
    const std::vector<int32_t> lvl = balance_levels(g, world, shard_min);
    for (size_t i = 0; i < N; i++) {
        if (g.nodes[i].kind == K_PBS) pbs_at[lvl[i]].push_back((int32_t)i);
        else if (g.nodes[i].kind == K_LIN) lin_at[lvl[i]].push_back((int32_t)i);
    }
    std::vector<uint32_t> slot(N, 0);
    size_t next = 0;
    for (size_t i = 0; i < N; i++)
        if (g.nodes[i].kind == K_INPUT) slot[i] = (uint32_t)next++;
    if (next != p->n_inputs) return FHS_ERR_INVALID;
    std::unique_ptr<Exec> E(new Exec());
    E->ctx = ctx;
    E->n_inputs = p->n_inputs;
    E->width = width;
    std::vector<uint32_t> off{0}, idx, lut;
    std::vector<int32_t> coef;
    std::vector<uint64_t> cst;
    size_t max_rows = 1;
    auto plan_group = [&](const std::vector<int32_t> &ids, bool is_pbs) {
        if (ids.empty()) return;
        Grp G{cst.size(), ids.size(), next, 0, 0, ids.size(), is_pbs};
        for (size_t q = 0; q < ids.size(); q++) {
            const Node &n = g.nodes[ids[q]];
            for (uint32_t t = n.t0; t < n.t0 + n.tn; t++) {
                idx.push_back(slot[g.terms[t].src]);
                coef.push_back(g.terms[t].coef);
            }
            off.push_back((uint32_t)idx.size());
            cst.push_back((uint64_t)n.c * delta);
            lut.push_back(is_pbs ? (uint32_t)n.lut : 0u);
            slot[ids[q]] = (uint32_t)(next + q);
        }
        if (is_pbs) {
            level_slice(ids.size(), rank, world, shard_min, &G.per, &G.lo, &G.hi);
            max_rows = std::max(max_rows, G.hi - G.lo);
            E->n_pbs += ids.size();
            E->n_levels++;
            if (G.per) E->n_sharded++;
        }
        next += G.per ? G.per * (size_t)world : ids.size();   // sharded levels are padded to per * world
        E->groups.push_back(G);
    };
    for (int l = 0; l <= D; l++) { plan_group(pbs_at[l], true); plan_group(lin_at[l], false); }
    // output gather group: identity rows (or clear constants as trivial ciphertexts)
    const size_t n_out = p->outputs.size();
    E->n_out = n_out;
    E->OG = Grp{cst.size(), n_out, next, 0, 0, n_out, false};
    for (size_t k = 0; k < n_out; k++) {
        if (p->outputs[k] >= 0) {
            idx.push_back(slot[p->outputs[k]]);
            coef.push_back(1);
            cst.push_back(0);
        } else {
            cst.push_back((uint64_t)(-1 - p->outputs[k]) * delta);
        }
        off.push_back((uint32_t)idx.size());
        lut.push_back(0u);
    }
    next += n_out;
    E->slots = next ? next : 1;
    std::vector<uint64_t> tables(g.luts.size() * 16);
    for (size_t i = 0; i < g.luts.size(); i++) memcpy(&tables[16 * i], g.luts[i].data(), 16 * sizeof(uint64_t));
    // a program with split levels gets a pool the other ranks can store into (peer window); the call is
    // collective, and every rank builds the execution state of an operation at the same point
    if (world > 1 && E->n_sharded && own_comm) rc = fhs_pool_create_shared(ctx, E->slots, &E->pool);
    else rc = fhs_pool_create(ctx, E->slots, &E->pool);
    if (rc != FHS_OK) return rc;
    if ((rc = fhs_pool_load_program(E->pool, cst.size(), idx.size(), off.data(), idx.data(), coef.data(), cst.data(),
                                    lut.data(), tables.data(), g.luts.size(), max_rows)) != FHS_OK) return rc;
    *out = E.release();
    return FHS_OK;
}

// staging buffers are made on first need (callers that pass page-locked or device memory never need them)
int stage_buf(Exec *E, uint64_t **buf, size_t blocks) {
    if (*buf) return FHS_OK;
    void *q = nullptr;
    int rc = fhs_host_alloc(E->ctx, (blocks ? blocks : 1) * E->width * sizeof(uint64_t), &q);
    if (rc == FHS_OK) *buf = static_cast<uint64_t *>(q);
    return rc;
}

}  // namespace

extern "C" int fhs_prog_run(const fhs_prog *p, fhs_ctx *ctx, const uint64_t *in_blocks,
                            uint64_t *out_blocks, double stats[4]) {
    return fhs_prog_run_sharded(p, ctx, in_blocks, out_blocks, stats, 0, 1, 0, nullptr, nullptr);
}

extern "C" int fhs_prog_run_sharded(const fhs_prog *p, fhs_ctx *ctx, const uint64_t *in_blocks,
                                    uint64_t *out_blocks, double stats[4], int rank, int world,
                                    size_t shard_min, fhs_allgather_fn allgather, void *user) {
    if (!p || !ctx || (!in_blocks && p->n_inputs) || !out_blocks) return FHS_ERR_INVALID;
    fhs_operand op{in_blocks, nullptr, p->n_inputs};
    double st[8];
    int rc = fhs_prog_run_ex(p, ctx, &op, p->n_inputs ? 1 : 0, out_blocks, nullptr, st, rank, world, shard_min, allgather, user);
    if (stats) { stats[0] = st[0]; stats[1] = st[1]; stats[2] = st[2]; stats[3] = st[3]; }
    return rc;
}

extern "C" int fhs_prog_run_ex(const fhs_prog *p, fhs_ctx *ctx, const fhs_operand *ops, size_t n_ops,
                               uint64_t *out_host, uint64_t *out_dev, double stats[8], int rank, int world,
                               size_t shard_min, fhs_allgather_fn allgather, void *user) {
    if (!p || !ctx || (!ops && n_ops)) return FHS_ERR_INVALID;
    if (world < 1 || rank < 0 || rank >= world) return FHS_ERR_INVALID;
    if (world > 1 && !allgather) {
        // no callback: the context's own NCCL communicator (fhs_comm_init) must match rank / world
        int cr = -1, cw = 0;
        if (fhs_comm_info(ctx, &cr, &cw) != FHS_OK || cr != rank || cw != world) return FHS_ERR_COMM;
    }
    if (shard_min == 0) shard_min = 74;
    size_t given = 0;
    for (size_t i = 0; i < n_ops; i++) {
        if (ops[i].n_blocks && !ops[i].host && !ops[i].dev) return FHS_ERR_INVALID;
        given += ops[i].n_blocks;
    }
    if (given != p->n_inputs) { g_host_err = "operand blocks do not add up to the program's inputs"; return FHS_ERR_INVALID; }
    auto t_start = std::chrono::steady_clock::now();
    // ---- cached execution state ----
    const uint64_t key[4] = {p->serial, (uint64_t)rank, (uint64_t)world, (uint64_t)shard_min};
    void *obj = nullptr;
    int rc = fhs_ctx_cache_get(ctx, key, &obj);
    if (rc != FHS_OK) return rc;
    Exec *E = static_cast<Exec *>(obj);
    const bool cached = E != nullptr;
    if (!E) {
        if ((rc = build_exec(p, ctx, rank, world, shard_min, allgather == nullptr, &E)) != FHS_OK) { delete E; return rc; }
        if ((rc = fhs_ctx_cache_put(ctx, key, E, exec_dtor, E->slots * E->width * sizeof(uint64_t))) != FHS_OK) { delete E; return rc; }
    }
    const double plan_ms = std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t_start).count();
    const size_t width = E->width, n_out = E->n_out;
    const uint64_t launches0 = fhs_launch_count(ctx);
    fhs_pool *pool = E->pool;
    const bool peer_window = world > 1 && fhs_pool_is_shared(pool);
    bool timed = fhs_ctx_timer_start(ctx) == FHS_OK;
    do {
        // ---- operands: device arrays are copied on the device; host arrays go by DMA straight from
        // page-locked memory, pageable ones through the page-locked staging buffer ----
        size_t slot = 0;
        for (size_t i = 0; i < n_ops && rc == FHS_OK; i++) {
            const size_t nb = ops[i].n_blocks;
            if (!nb) continue;
            if (ops[i].dev) {
                rc = fhs_pool_copy_in(pool, slot, nb, ops[i].dev);
            } else if (fhs_host_is_pinned(ops[i].host)) {
                rc = fhs_pool_enqueue_write(pool, slot, nb, ops[i].host);
            } else {
                if ((rc = stage_buf(E, &E->h_in, E->n_inputs)) != FHS_OK) break;
                memcpy(E->h_in + slot * width, ops[i].host, nb * width * sizeof(uint64_t));
                rc = fhs_pool_enqueue_write(pool, slot, nb, E->h_in + slot * width);
            }
            slot += nb;
        }
        if (rc != FHS_OK) break;
        for (const Grp &G : E->groups) {
            if (G.hi > G.lo && (rc = fhs_level_run(pool, G.row0 + G.lo, G.hi - G.lo, G.is_pbs ? 1 : 0, G.out_base + G.lo)) != FHS_OK) break;
            if (G.per) {  // exchange result blocks: rank q owns [out_base + q*per, out_base + (q+1)*per)
                uint64_t *base = fhs_pool_device_ptr(pool) + G.out_base * width;
                const size_t bytes = G.per * width * sizeof(uint64_t);
                if (!allgather && peer_window) {
                    // one kernel of our own: stores into the peers' pools over NVLink + sequence numbers
                    if ((rc = fhs_pool_peer_allgather(pool, G.out_base, G.per, G.hi - G.lo)) != FHS_OK) break;
                } else if (!allgather) {
                    // ncclAllGather in place, on the stream the level's kernels run on: no host sync
                    if ((rc = fhs_comm_allgather(ctx, base, bytes)) != FHS_OK) break;
                } else {
                    fhs_ctx_synchronize(ctx);
                    if (allgather(user, base, bytes) != 0) { rc = FHS_ERR_COMM; break; }
                }
            }
        }
        if (rc != FHS_OK) break;
        if (n_out) {
            if ((rc = fhs_level_run(pool, E->OG.row0, n_out, 0, E->OG.out_base)) != FHS_OK) break;
            if (out_dev && (rc = fhs_pool_copy_out(pool, E->OG.out_base, n_out, out_dev)) != FHS_OK) break;
        }
        // nothing below reads blocks another rank wrote: the peers may start overwriting them
        if (peer_window && !allgather && (rc = fhs_pool_peer_run_done(pool)) != FHS_OK) break;
        if (timed) fhs_ctx_timer_stop(ctx);
        if (n_out && out_host) {
            if (fhs_host_is_pinned(out_host)) {
                rc = fhs_pool_enqueue_read(pool, E->OG.out_base, n_out, out_host);
            } else {
                if ((rc = stage_buf(E, &E->h_out, n_out)) != FHS_OK) break;
                if ((rc = fhs_pool_enqueue_read(pool, E->OG.out_base, n_out, E->h_out)) != FHS_OK) break;
                if ((rc = fhs_ctx_synchronize(ctx)) != FHS_OK) break;
                memcpy(out_host, E->h_out, n_out * width * sizeof(uint64_t));
            }
        }
    } while (0);
    const int src = fhs_ctx_synchronize(ctx);   // operands (caller memory) and results are settled on return
    if (rc == FHS_OK) rc = src;
    if (rc == FHS_OK && peer_window) rc = fhs_pool_peer_status(pool);
    if (stats) {
        float dev_ms = 0.f;
        if (!(timed && rc == FHS_OK && fhs_ctx_timer_ms(ctx, &dev_ms) == FHS_OK)) dev_ms = -1.f;
        stats[0] = (double)E->n_levels;
        stats[1] = (double)E->n_pbs;
        stats[2] = (double)(fhs_launch_count(ctx) - launches0);
        stats[3] = std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t_start).count();
        stats[4] = (double)dev_ms;
        stats[5] = plan_ms;
        stats[6] = (double)E->n_sharded;
        stats[7] = cached ? 1.0 : 0.0;
    }
    return rc;
}

// ------------------------------------------------------------------------------------------------
// Batched executor: M independent instances of one program (e.g. 1024 string comparisons, BASELINE
// config 5).  Instances are processed in chunks of Mc; within a chunk every level is ONE fhs_level_run
// over Mc * B_level evaluations (resident descriptors, generated once: slots are relative to the pool),
// and two pools are used in turn so that the H2D upload of chunk c+1 (side stream) overlaps with the
// levels of chunk c.  Pool layout: inputs [m][n_inputs]; group of node ids at base b: [m][ids.size()].
extern "C" int fhs_prog_run_many(const fhs_prog *p, fhs_ctx *ctx, size_t n_instances,
                                 const uint64_t *in_blocks, uint64_t *out_blocks, double stats[4]) {
    if (!p || !ctx || (!in_blocks && p->n_inputs) || !out_blocks) return FHS_ERR_INVALID;
    if (n_instances == 0) return FHS_OK;
    constexpr size_t kManyChunk = 24576;    // target LUT evaluations of the largest level of a chunk
    const size_t M = n_instances;
    const Graph &g = p->g;
    fhs_params prm;
    int rc = fhs_ctx_get_params(ctx, &prm);
    if (rc != FHS_OK) return rc;
    const size_t width = (size_t)prm.glwe_dim * prm.poly_size + 1;
    const uint64_t delta = (1ull << 63) / ((uint64_t)prm.message_modulus * prm.carry_modulus);
    const size_t N = g.nodes.size();
    const int D = g.depth();
    auto t_start = std::chrono::steady_clock::now();
    std::vector<std::vector<int32_t>> pbs_at(D + 1), lin_at(D + 1);
    size_t gmax = 1;
    for (size_t i = 0; i < N; i++) {
        if (g.nodes[i].kind == K_PBS) pbs_at[g.nodes[i].level].push_back((int32_t)i);
        else if (g.nodes[i].kind == K_LIN) lin_at[g.nodes[i].level].push_back((int32_t)i);
    }
    for (auto &v : pbs_at) gmax = std::max(gmax, v.size());
    // Chunk size: at most kManyChunk evaluations in the largest level, and among those sizes the one
    // that minimises (number of chunks) x (blind-rotation rounds per chunk).  A round is 4 ciphertexts
    // per SM (592 on a B200); a remainder of <= 74 goes to the pair kernel (0.28 of a round), of
    // <= 148 / <= 296 to the wide kernel (0.39 / 0.59), of <= 370 to wide + pair (0.89).  A short last chunk still runs whole levels, so equal chunks beat "full chunks + a rest".
    const size_t Mc_max = std::max<size_t>(1, std::min(M, kManyChunk / gmax));
    size_t Mc = Mc_max;
    {
        constexpr size_t kRound = 592;
        double best = -1.0;
        for (size_t cand = Mc_max; cand >= std::max<size_t>(1, Mc_max / 2); cand--) {
            double rounds = 0.0;
            for (auto &v : pbs_at) {
                if (v.empty()) continue;
                const size_t rows = v.size() * cand, rem = rows % kRound;
                rounds += (double)(rows / kRound) + (rem == 0 ? 0.0 : rem <= 74 ? 0.28 : rem <= 148 ? 0.39 : rem <= 296 ? 0.59 : rem <= 370 ? 0.89 : 1.0);
            }
            const double cost = rounds * (double)((M + cand - 1) / cand);
            if (best < 0.0 || cost < best) { best = cost; Mc = cand; }
            if (cand == 1) break;
        }
    }
    // ---- plan for one chunk of Mc instances: slot(node, m) = base[node] + m * stride[node] ----
    std::vector<size_t> base(N, 0), stride(N, 0);
    size_t n_in = 0;
    for (size_t i = 0; i < N; i++)
        if (g.nodes[i].kind == K_INPUT) base[i] = n_in++;
    if (n_in != p->n_inputs) return FHS_ERR_INVALID;
    for (size_t i = 0; i < N; i++)
        if (g.nodes[i].kind == K_INPUT) stride[i] = n_in;
    const size_t n_out = p->outputs.size();
    if ((N + n_out + 1) * Mc >= (1ull << 32)) return FHS_ERR_INVALID;
    struct Grp { size_t row0, rows, out_base; bool is_pbs; };
    std::vector<Grp> groups;
    std::vector<uint32_t> off{0}, idx, lut;
    std::vector<int32_t> coef;
    std::vector<uint64_t> cst;
    size_t next = n_in * Mc, max_rows = 1, pbs_per_instance = 0, n_levels = 0;
    auto plan_group = [&](const std::vector<int32_t> &ids, bool is_pbs) {
        if (ids.empty()) return;
        const size_t G = ids.size();
        groups.push_back(Grp{cst.size(), G * Mc, next, is_pbs});
        for (size_t m = 0; m < Mc; m++) {
            for (int32_t id : ids) {
                const Node &n = g.nodes[id];
                for (uint32_t t = n.t0; t < n.t0 + n.tn; t++) {
                    const int32_t src = g.terms[t].src;
                    idx.push_back((uint32_t)(base[src] + m * stride[src]));
                    coef.push_back(g.terms[t].coef);
                }
                off.push_back((uint32_t)idx.size());
                cst.push_back((uint64_t)n.c * delta);
                lut.push_back(is_pbs ? (uint32_t)n.lut : 0u);
            }
        }
        for (size_t q = 0; q < G; q++) { base[ids[q]] = next + q; stride[ids[q]] = G; }
        next += G * Mc;
        if (is_pbs) { max_rows = std::max(max_rows, G * Mc); pbs_per_instance += G; n_levels++; }
    };
    for (int l = 0; l <= D; l++) { plan_group(pbs_at[l], true); plan_group(lin_at[l], false); }
    // output gather: [m][n_out] consecutive slots -> one read-back per chunk
    Grp OG{cst.size(), n_out * Mc, next, false};
    for (size_t m = 0; m < Mc; m++) {
        for (size_t k = 0; k < n_out; k++) {
            if (p->outputs[k] >= 0) {
                const size_t node = (size_t)p->outputs[k];
                idx.push_back((uint32_t)(base[node] + m * stride[node]));
                coef.push_back(1);
                cst.push_back(0);
            } else {
                cst.push_back((uint64_t)(-1 - p->outputs[k]) * delta);
            }
            off.push_back((uint32_t)idx.size());
            lut.push_back(0u);
        }
    }
    next += n_out * Mc;
    std::vector<uint64_t> tables(g.luts.size() * 16);
    for (size_t i = 0; i < g.luts.size(); i++) memcpy(&tables[16 * i], g.luts[i].data(), 16 * sizeof(uint64_t));

    const size_t n_chunks = (M + Mc - 1) / Mc;
    fhs_pool *pool[2] = {nullptr, nullptr};
    const uint64_t launches0 = fhs_launch_count(ctx);
    do {
        for (int b = 0; b < (n_chunks > 1 ? 2 : 1) && rc == FHS_OK; b++) {
            if ((rc = fhs_pool_create(ctx, next ? next : 1, &pool[b])) != FHS_OK) break;
            rc = fhs_pool_load_program(pool[b], cst.size(), idx.size(), off.data(), idx.data(), coef.data(), cst.data(),
                                       lut.data(), tables.data(), g.luts.size(), max_rows);
        }
        if (rc != FHS_OK) break;
        auto upload = [&](size_t c) -> int {   // inputs of chunk c -> its pool, on the side stream
            const size_t m0 = c * Mc, mc = std::min(Mc, M - m0);
            return n_in ? fhs_pool_write_async(pool[c & 1], 0, n_in * mc, in_blocks + m0 * n_in * width) : FHS_OK;
        };
        if ((rc = upload(0)) != FHS_OK) break;
        for (size_t c = 0; c < n_chunks && rc == FHS_OK; c++) {
            fhs_pool *P = pool[c & 1];
            const size_t m0 = c * Mc, mc = std::min(Mc, M - m0);
            if ((rc = fhs_pool_wait_writes(P)) != FHS_OK) break;
            // a short last chunk runs the same rows; instances >= mc compute on stale inputs and are ignored
            for (const Grp &G : groups)
                if ((rc = fhs_level_run(P, G.row0, G.rows, G.is_pbs ? 1 : 0, G.out_base)) != FHS_OK) break;
            if (rc != FHS_OK) break;
            if (n_out && (rc = fhs_level_run(P, OG.row0, OG.rows, 0, OG.out_base)) != FHS_OK) break;
            if (c + 1 < n_chunks && (rc = upload(c + 1)) != FHS_OK) break;   // overlaps with the levels above
            if (n_out) rc = fhs_pool_read(P, OG.out_base, n_out * mc, out_blocks + m0 * n_out * width);
        }
    } while (0);
    fhs_ctx_synchronize(ctx);
    if (stats) {
        stats[0] = (double)n_levels;
        stats[1] = (double)(pbs_per_instance * M);
        stats[2] = (double)(fhs_launch_count(ctx) - launches0);
        stats[3] = std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t_start).count();
    }
    fhs_pool_destroy(pool[0]);
    fhs_pool_destroy(pool[1]);
    return rc;
}
