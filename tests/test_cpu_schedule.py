"""Level balancing (csrc/host/prog.cpp balance_levels): a level costs what its blind-rotation waves cost, so
evaluations with slack are moved to later levels when that saves a wave.  Checked on CPU for every op family:
the schedule the executor runs respects every data dependency, keeps depth and evaluation count, is never
modelled slower than the as-soon-as-possible one, is the same on every rank, and is a real gain on the ops
BASELINE names (split / replace)."""
import os
import subprocess
import sys

import numpy as np
import pytest

from fhe_string_b200.program import Program

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))

CASES = [
    ("eq", (65, 65), {}), ("lt", (65, 65), {}), ("contains", (65, 9), {}), ("find", (65, 9), {}),
    ("rfind", (65, 9), {}), ("trim", (65,), {}), ("split", (65, 9), {}), ("rsplit", (33, 3), {}),
    ("split", (257, 2), dict(num_blocks_usize=5)), ("to_uppercase", (257,), dict(num_blocks_usize=5)),
    ("replace", (65, 3, 2), dict(scalars=(64,))), ("strip_prefix", (33, 5), {}),
]


def _check(prog, world):
    d = prog.export()
    lvl, asap_ms, bal_ms = prog.schedule(world)
    kind, t0, tn, src = d["kind"], d["t0"], d["tn"], d["term_src"]
    assert bal_ms <= asap_ms + 1e-9
    assert lvl.max() <= prog.depth and np.all(lvl[kind == 0] == 0)
    for i in range(prog.n_nodes):
        s = src[t0[i]:t0[i] + tn[i]]
        if kind[i] == 1:
            assert np.all(lvl[s] < lvl[i]), f"LUT {i} at level {lvl[i]} reads level {lvl[s].max()}"
            assert lvl[i] >= d["level"][i]                    # never earlier than its inputs allow
        elif kind[i] == 2 and len(s):
            assert lvl[i] == lvl[s].max()
    assert int((kind == 1).sum()) == prog.pbs_count
    # the plan (what fhs_prog_run_ex executes and what the ranks split) is this schedule
    plan = prog.plan(0, world)
    sizes = np.bincount(lvl[kind == 1], minlength=prog.depth + 1)[1:]
    assert [int(x) for x in plan[:, 0]] == [int(x) for x in sizes if x]
    return asap_ms, bal_ms


@pytest.mark.parametrize("op,n_chars,kw", CASES)
def test_balanced_schedule_is_valid_and_not_slower(op, n_chars, kw):
    prog = Program(op, n_chars=n_chars, **kw)
    for world in (1, 2, 8):
        _check(prog, world)
    # every rank of a world derives the same levels (the pass does not look at the rank)
    a = prog.plan(0, 4)
    b = prog.plan(3, 4)
    assert np.array_equal(a[:, :2], b[:, :2])


def test_balancing_gains_on_the_baseline_ops():
    split64 = Program("split", n_chars=(65, 9))
    split256 = Program("split", n_chars=(257, 2), num_blocks_usize=5)
    for prog, world, want in ((split64, 1, 0.03), (split256, 1, 0.03), (split256, 2, 0.0), (split64, 8, 0.005)):
        asap_ms, bal_ms = _check(prog, world)
        assert bal_ms <= (1 - want) * asap_ms, (prog.op, world, asap_ms, bal_ms)


def test_balancing_can_be_switched_off():
    code = ("from fhe_string_b200.program import Program; p = Program('split', n_chars=(65, 9)); "
            "import numpy as np; l, a, b = p.schedule(1); print(int(np.array_equal(l, p.export()['level'])), a == b)")
    env = dict(os.environ, FHS_LEVEL_BALANCE="0", PYTHONPATH=ROOT)
    out = subprocess.run([sys.executable, "-c", code], capture_output=True, text=True, env=env, timeout=120)
    assert out.stdout.split() == ["1", "True"], (out.stdout, out.stderr[-500:])


@pytest.mark.parametrize("op,n_chars,kw", CASES)
def test_no_duplicate_and_no_dead_evaluations(op, n_chars, kw):
    """Graph::pbs merges equal evaluations (same table, same linear combination) at creation and fhs_prog_build
    prunes what no output depends on: the builders' helpers used to ask for the same zero masks / equality bits
    several times (35 % of split-256's evaluations, 10 % of split-64's)."""
    prog = Program(op, n_chars=n_chars, **kw)
    d = prog.export()
    kind, t0, tn, src, coef = d["kind"], d["t0"], d["tn"], d["term_src"], d["term_coef"]
    seen = set()
    for i in range(prog.n_nodes):
        if kind[i] == 0:
            continue
        key = (int(kind[i]), int(d["lut"][i]) if kind[i] == 1 else -1, int(d["cst"][i]), int(d["signed_in"][i]),
               tuple(sorted((int(src[t]), int(coef[t])) for t in range(t0[i], t0[i] + tn[i]))))
        assert key not in seen, f"node {i} repeats an earlier evaluation"
        seen.add(key)
    live = np.zeros(prog.n_nodes, dtype=bool)
    live[[int(o) for o in d["outputs"] if o >= 0]] = True
    for i in range(prog.n_nodes - 1, -1, -1):
        if live[i]:
            live[src[t0[i]:t0[i] + tn[i]]] = True
    assert np.all(live[kind != 0]), f"{int((~live[kind != 0]).sum())} evaluations reach no output"
    assert int((kind == 1).sum()) == prog.pbs_count


def test_cse_and_table_composition_shrink_split():
    """Repeated evaluations merged (5 997 -> 5 323, 10 260 -> 6 430), then a table on the output of a table folded
    into one table on the inner input (-> 5 204 / depth 31, 5 780 / depth 18)."""
    p64 = Program("split", n_chars=(65, 9))
    p256 = Program("split", n_chars=(257, 2), num_blocks_usize=5)
    assert p64.pbs_count <= 5250 and p64.depth <= 31
    assert p256.pbs_count <= 5800 and p256.depth <= 18


@pytest.mark.parametrize("op,n_chars,kw", CASES)
def test_no_table_directly_on_a_table(op, n_chars, kw):
    """After composition no unsigned evaluation reads exactly one other unsigned evaluation and nothing else."""
    prog = Program(op, n_chars=n_chars, **kw)
    d = prog.export()
    kind, t0, tn, src, sg = d["kind"], d["t0"], d["tn"], d["term_src"], d["signed_in"]
    for i in range(prog.n_nodes):
        if kind[i] == 1 and tn[i] == 1 and not sg[i]:
            s = src[t0[i]]
            assert not (kind[s] == 1 and not sg[s]), f"node {i} is a table on table {s}"


NOISE_CASES = CASES + [("len", (65,), {}), ("add", (33, 33), {}), ("insert", (33, 9), dict(n_usize_args=1, scalars=(64,))),
                       ("repeat", (9,), dict(n_usize_args=1, scalars=(4,))), ("replace", (257, 3, 2), dict(scalars=(256,), num_blocks_usize=5)),
                       ("splitn", (65, 9), dict(n_usize_args=1)), ("eq_ignore_ascii_case", (65, 65), {})]


@pytest.mark.parametrize("op,n_chars,kw", NOISE_CASES)
def test_chained_operations_stay_inside_the_noise_budget(op, n_chars, kw):
    """An operation treats its operands as fresh (variance multiplier 1), but the previous operation may hand it
    blocks that carry up to `output_noise` PBS-output variances.  Two bounds keep chains inside the budget of
    DESIGN.md section 4: outputs carry at most 12, and operand blocks enter a table with sum(coef^2) <= 34
    (the eq / bivariate packs), so a chained table input is at most 163 + 34 * 11 = 537 PBS-output variances =
    3.6e-7 on top of 4.74e-6 of keyswitch + modulus-switch noise (z 7.175 -> 6.92, 2^-37.7 per evaluation)."""
    prog = Program(op, n_chars=n_chars, **kw)
    assert prog.output_noise <= 12, prog.output_noise
    d = prog.export()
    kind, t0, tn, src, coef = d["kind"], d["t0"], d["tn"], d["term_src"], d["term_coef"].astype(np.int64)
    worst = 0
    for i in np.where(kind == 1)[0]:
        s, c = src[t0[i]:t0[i] + tn[i]], coef[t0[i]:t0[i] + tn[i]]
        worst = max(worst, int((c[kind[s] == 0] ** 2).sum()))
    assert worst <= 34, worst
    assert prog.max_noise <= 163
