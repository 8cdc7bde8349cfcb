"""The CPU oracle against the golden fixtures the REFERENCE produced (oracle/make_golden.py).
Runs anywhere (no /root/reference needed): this is what keeps the oracle pinned on the GPU box."""
import glob
import os

import pytest

from oracle import synth
from oracle.driver import check_against_golden, golden_cfg
from dexrobot_isaac_b200.config import load_task_config

GOLDEN = sorted(glob.glob(os.path.join(os.path.dirname(__file__), "golden", "*.npz")))


def make_oracle(meta):
    from oracle.dexstep_oracle import OracleEnv

    cfg = golden_cfg(meta, load_task_config)
    layout = synth.BodyLayout(with_box=meta["task"] == "BlindGrasping")
    return OracleEnv(cfg, meta["n"], layout, synth.dof_props_tensor())


@pytest.mark.parametrize("path", GOLDEN, ids=[os.path.basename(p) for p in GOLDEN])
def test_oracle_matches_reference_golden(path):
    bad = check_against_golden(path, make_oracle, rtol=0.0, atol=0.0)
    assert not bad, "\n".join(bad)


def test_golden_fixtures_present():
    assert len(GOLDEN) >= 8
