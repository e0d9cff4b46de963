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
