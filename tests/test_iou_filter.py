"""The fp32 interval prefilter (yolo_v1_b200/csrc/iou_filter.cuh) may only answer where the reference's own
fp64 arithmetic gives the same answer.  The header is plain C++: tests/native/iou_filter_check.cpp compiles it
for the host and checks every verdict / interval against an fp64 restatement of metrics.py:317-341 and
inference.py:237-248 on random, degenerate, huge and threshold-steered box pairs.  CPU only."""
import json
import os
import subprocess

import pytest

HERE = os.path.dirname(os.path.abspath(__file__))
SRC = os.path.join(HERE, "native", "iou_filter_check.cpp")


@pytest.fixture(scope="module")
def checker(tmp_path_factory):
    exe = str(tmp_path_factory.mktemp("iou_filter") / "iou_filter_check")
    subprocess.check_call(["g++", "-O2", "-ffp-contract=off", "-std=c++17", "-o", exe, SRC])
    return exe


@pytest.mark.parametrize("seed", [1, 2, 3])
def test_prefilter_never_contradicts_the_reference(checker, seed):
    res = subprocess.run([checker, "4000000", str(seed)], capture_output=True, text=True)
    out = json.loads(res.stdout)
    assert res.returncode == 0 and out["wrong"] == 0 and out["interval_misses"] == 0, res.stderr
    assert out["decided"] > 0.5 * out["pairs"]            # it does answer most of the time
    assert out["steered"] > 0.15 * out["pairs"]           # and the near-threshold cases were exercised
