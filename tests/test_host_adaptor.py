"""The C++ host adaptor (trajopt-1_b200/host) driven like the reference's constraints drive their evaluators."""
import os
import subprocess

import numpy as np
import pytest

from tests import helpers

capi = helpers.capi
HERE = os.path.join(helpers.ROOT, "tests", "host")


def _build(tmp_path):
    if not os.path.exists(helpers.pkg.LIB_PATH):
        subprocess.check_call(["make", "-C", os.path.join(helpers.ROOT, "trajopt-1_b200", "csrc"), "-s", "-j4"])
    subprocess.check_call(["make", "-C", HERE, "-s"])
    files = []
    for name in ("boxbot", "arm_around_table"):
        p = str(tmp_path / (name + ".scene"))
        helpers.save_scene_file(helpers.load_scene(name), p)
        files.append(p)
    return [os.path.join(HERE, "test_adaptor")] + files


def test_adaptor_builds_and_refuses_to_run_without_a_gpu(tmp_path):
    """plain g++ build against include/trajopt_cuda.h; on a box without a B200 the evaluators throw (no fallback)"""
    cmd = _build(tmp_path)
    try:
        import torch
        has_gpu = torch.cuda.is_available()
    except Exception:
        has_gpu = False
    if has_gpu:
        pytest.skip("GPU present: covered by the gpu test")
    r = subprocess.run(cmd, capture_output=True, text=True)
    assert r.returncode == 1 and "tc_init" in r.stderr, (r.returncode, r.stderr)


def _parse(line):
    tok = line.split()
    out = {"tag": tok[0], "term": int(tok[1]), "contacts": int(tok[3]), "sets": int(tok[5])}
    i = tok.index("values")
    j = tok.index("jac0")
    k = tok.index("jac1") if "jac1" in tok else len(tok)
    out["values"] = np.array([float(x) for x in tok[i + 1:j]])
    out["jac0"] = np.array([float(x) for x in tok[j + 1:k]])
    out["jac1"] = np.array([float(x) for x in tok[k + 1:]])
    return out


@pytest.mark.gpu
def test_adaptor_matches_known_answers_and_oracle(tmp_path, oracle):
    cmd = _build(tmp_path)
    r = subprocess.run(cmd, capture_output=True, text=True)
    assert r.returncode == 0, r.stderr + r.stdout
    lines = r.stdout.strip().splitlines()
    # reference KATs, trajopt_ifopt/test/collision_unit.cpp:103-158, EXPECT_NEAR 1e-6
    kat = [_parse(l) for l in lines if l.startswith("KAT")]
    want = [(-0.01, (0, 0)), (0.1, (1, 0)), (0.1, (0, -1)), (0.2, (1, 0))]
    assert len(kat) == 4
    for got, (val, jac) in zip(kat, want):
        assert abs(got["values"][0] - val) <= 1e-6, got
        assert np.abs(got["jac0"] - np.array(jac)).max() <= 1e-6, got
    assert "BADCONFIG 1" in lines
    assert any(l.startswith("NOSETS contacts 1 sets 0") for l in lines)
    assert "BATCH_PASSES 0 1" in lines and "BATCH_PASSES 1 1" in lines  # one tc_eval served every segment
    # fixed timesteps make tc_eval skip terms: slices must still be the right timestep / segment (one pass each)
    assert "BATCH_SKIPPED_TERMS 0 mismatches 0 skipped_refused 1 passes 1" in lines, [l for l in lines if "SKIPPED" in l]
    assert "BATCH_SKIPPED_TERMS 1 mismatches 0 skipped_refused 1 passes 1" in lines
    bd = [l for l in lines if l.startswith("BATCH_DISCRETE_FIXED")]
    assert len(bd) == 1 and bd[0].startswith("BATCH_DISCRETE_FIXED mismatches 0 contacts") and bd[0].endswith("passes 1"), bd
    assert int(bd[0].split()[4]) > 0
    # arm_around_table seed trajectory against the oracle, segment by segment
    s = helpers.load_scene("arm_around_table")
    so = oracle.scene(s, 0.025, 20.0, 0.01)
    traj = s["seed_traj"]
    T, n = traj.shape
    fixed = np.zeros(T, np.uint8)
    fixed[0] = fixed[-1] = 1
    for tag, ev in (("LVSC", capi.EVAL_LVS_CONTINUOUS), ("LVSD", capi.EVAL_LVS_DISCRETE)):
        got = [_parse(l) for l in lines if l.startswith(tag + " ")]
        assert len(got) == T - 1
        ro = oracle.eval(so, traj[None], ev, lvs=0.02, fixed=fixed)
        st = ro.sets(3)
        tp = ro.term_ptr()
        for t, g in enumerate(got):
            assert g["contacts"] == tp[t + 1] - tp[t], (tag, t)
            assert np.abs(g["values"] - st["values"][t]).max() <= 1e-6, (tag, t, g["values"], st["values"][t])
            for blk, key in ((0, "jac0"), (1, "jac1")):
                ref = st["jac"][t, :, blk].reshape(-1)
                err = np.abs(g[key] - ref).max() / max(np.abs(ref).max(), 1e-12)
                assert err <= 1e-5, (tag, t, key, err)
    # numerical-gradient constraints (A18) and trajectory check (N2) of the C++ adaptor against the same host logic
    # driven by the oracle (tests/test_constraints.py states the tolerances)
    cons = helpers.pkg.constraints
    num = [l.split() for l in lines if l.startswith("NUMKAT")]
    assert len(num) == 4 and "NUM_BADCONFIG 1" in lines
    for tok, (val, jac) in zip(num, want):
        assert abs(float(tok[3]) - val) <= 1e-6 and float(tok[5]) == 1.0
        assert np.abs(np.array([float(tok[7]), float(tok[8])]) - np.array(jac)).max() <= 2e-2, tok
    assert [int(tok[10]) for tok in num] == [1, 2, 3, 4]
    for tag, chk, ev in (("NUMC", "CHECKC", capi.EVAL_LVS_CONTINUOUS), ("NUMD", "CHECKD", capi.EVAL_LVS_DISCRETE)):
        tok = [l.split() for l in lines if l.startswith(tag + " ")][0]
        i, j, k = tok.index("values"), tok.index("jac"), tok.index("passes")
        vals = np.array([float(x) for x in tok[i + 1:j]])
        J = np.array([float(x) for x in tok[j + 1:k]]).reshape(3, 2, n)
        assert int(tok[k + 1]) == 1
        vo, co, Jo = cons.ContinuousCollisionNumericalConstraint(oracle, so, ev, 0.02, False, False, 3).evaluate(
            traj[1], traj[2])
        assert np.abs(vals - vo).max() <= 1e-6, (tag, vals, vo)
        assert np.abs(J - Jo).max() <= 2e-2, (tag, J, Jo)
        tok = [l.split() for l in lines if l.startswith(chk + " ")][0]
        o = cons.check_trajectories(oracle, so, traj[None], ev, lvs=0.02)
        assert int(tok[2]) == int(o["in_collision"][0]) == 1
        assert abs(float(tok[4]) - o["min_distance"][0]) <= 1e-6


def _parse_exprs(lines, tag):
    """{term: (n_exprs, n_contacts, n_vars, [constant, margin, coeff, value_at_x], dense coefficients)}"""
    out = {}
    for l in lines:
        tok = l.split()
        if tok[0] == tag:
            out[int(tok[1])] = dict(exprs=int(tok[3]), contacts=int(tok[5]), vars=int(tok[7]), head=[], dense=[])
        elif tok[0] == tag + "X":
            e = out[int(tok[1])]
            e["head"].append([float(v) for v in tok[3:7]])
            e["dense"].append([float(v) for v in tok[tok.index("coeffs") + 1:]])
    return out


def _check_exprs(got, ro, x, term, n_vars_total):
    """legacy seam: AffExpr i of a term == row i of the oracle's legacy rows for that term, after cleanupAff"""
    rows, tp = ro.rows(), ro.term_ptr()
    lo, hi = int(tp[term]), int(tp[term + 1])
    assert got["exprs"] == hi - lo and got["contacts"] == hi - lo, (term, got["exprs"], hi - lo)
    for i in range(lo, hi):
        a, b = int(rows["row_ptr"][i]), int(rows["row_ptr"][i + 1])
        dense = np.zeros(n_vars_total)
        np.add.at(dense, rows["col_idx"][a:b], rows["val"][a:b])
        dense[np.abs(dense) <= 1e-7] = 0.0
        const, margin, coeff, value = got["head"][i - lo]
        g = np.array(got["dense"][i - lo])
        # coefficients that sit within 1e-11 of the pruning threshold may fall either side of it
        edge = np.abs(np.abs(g + (g == 0) * dense) - 1e-7) < 1e-11
        err = np.abs(g - dense)[~edge].max() / max(np.abs(dense).max(), 1e-9)
        assert err <= 1e-5, (term, i, err)
        assert abs(margin - rows["margin"][i]) <= 1e-12 and abs(coeff - rows["coeff"][i]) <= 1e-12
        assert abs(const - rows["value"][i]) <= 1e-5 * max(1.0, abs(rows["value"][i])), (term, i, const, rows["value"][i])
        # the expression evaluated at the linearisation point is the signed distance
        assert abs(value - (rows["value"][i] + dense @ x)) <= 2e-6, (term, i)


@pytest.mark.gpu
def test_legacy_evaluators_match_oracle_rows(tmp_path, oracle):
    """seam B2: CalcDistExpressions / CalcCollisions of the legacy evaluator mirror against the oracle's legacy rows"""
    cmd = _build(tmp_path)
    r = subprocess.run(cmd, capture_output=True, text=True)
    assert r.returncode == 0, r.stderr + r.stdout
    lines = r.stdout.strip().splitlines()
    assert "LEGACY_BADCONFIG 1" in lines
    # boxbot, single timestep
    sb = helpers.load_scene("boxbot")
    ob = oracle.scene(sb, 0.1, 1.0, 0.01)
    legs = _parse_exprs(lines, "LEGS")
    states = [(-1.9, 0.01), (-1.0, 0.01), (0.01, 1.0), (-0.9, 0.01)]
    assert sorted(legs) == [0, 1, 2, 3]
    for k, st in enumerate(states):
        x = np.array(st)
        ro = oracle.eval(ob, x[None, None], capi.EVAL_DISCRETE, stack=capi.STACK_LEGACY, want=capi.WANT_ALL)
        assert legs[k]["vars"] == 2
        _check_exprs(legs[k], ro, x, 0, 2)
    assert legs[0]["exprs"] == 0 and legs[1]["exprs"] == 1
    # signed distance of the reference's known answers (error = margin - distance, collision_unit.cpp:103-158)
    for k, err in ((1, 0.1), (2, 0.1), (3, 0.2)):
        assert abs((0.1 - legs[k]["head"][0][3]) - err) <= 1e-6
    # arm_around_table seed trajectory, both two-state evaluators, segment by segment
    s = helpers.load_scene("arm_around_table")
    so = oracle.scene(s, 0.025, 20.0, 0.01)
    traj = s["seed_traj"]
    T, n = traj.shape
    fixed = np.zeros(T, np.uint8)
    fixed[0] = fixed[-1] = 1
    for tag, ev in (("LEGC", capi.EVAL_LVS_CONTINUOUS), ("LEGD", capi.EVAL_LVS_DISCRETE)):
        got = _parse_exprs(lines, tag)
        assert sorted(got) == list(range(T - 1))
        ro = oracle.eval(so, traj[None], ev, stack=capi.STACK_LEGACY, lvs=0.02, fixed=fixed, want=capi.WANT_ALL)
        assert sum(g["exprs"] for g in got.values()) > 0
        for t in range(T - 1):
            assert got[t]["vars"] == 2 * n
            _check_exprs(got[t], ro, traj.reshape(-1), t, T * n)
