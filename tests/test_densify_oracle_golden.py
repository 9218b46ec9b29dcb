"""CPU: pins oracle/densify_oracle.py (prune / opacity reset / add_new_gaussians incl. Adam-moment surgery) against
tests/golden/densify_*.npz -- outputs of the reference's own Python run unmodified on CPU by
tests/golden/make_densify_golden.py."""
import os

import numpy as np
import pytest
import torch

from util import GOLDEN_DIR
from oracle import densify_oracle as DO


def _same(a, b):      # surgery is pure data movement: bit-exact
    a, b = torch.as_tensor(a), torch.as_tensor(b)
    assert a.shape == b.shape, (a.shape, b.shape)
    assert torch.equal(a, b)


def _close(a, b, tol):
    a, b = torch.as_tensor(a).double(), torch.as_tensor(b).double()
    assert a.shape == b.shape, (a.shape, b.shape)
    scale = float(b.abs().max()) + 1e-30
    assert float((a - b).abs().max()) <= tol * scale


@pytest.mark.parametrize("case", DO.PRUNE_CASES, ids=lambda c: c["name"])
def test_prune_oracle_vs_reference_python(case):
    gold = np.load(os.path.join(GOLDEN_DIR, "densify_prune_%s.npz" % case["name"]))
    params, state, variables = DO.random_state(case["P"], case["iso"], case["seed"])
    params, state, variables = DO.prune_gaussians(params, state, variables, case["iter"], case["pd"])
    for k in DO.GAUSS_KEYS:
        _same(params[k], gold["p_" + k])
        _same(state[k]["exp_avg"], gold["m_" + k])
        _same(state[k]["exp_avg_sq"], gold["v_" + k])
    for k in DO.VAR_KEYS:
        _same(variables[k], gold["var_" + k])


@pytest.mark.parametrize("case", DO.ADD_CASES, ids=lambda c: c["name"])
def test_add_new_gaussians_oracle_vs_reference_python(case):
    gold = np.load(os.path.join(GOLDEN_DIR, "densify_add_%s.npz" % case["name"]))
    params, state, variables = DO.random_state(case["P"], case["iso"], case["seed"])
    im, gt_depth, depth_sil, intr = DO.frame(case["h"], case["w"], case["seed"])
    params, state, variables = DO.add_new_gaussians(params, state, variables, depth_sil, im, gt_depth, intr,
                                                    case["sil_thres"], case["time_idx"])
    for k in DO.GAUSS_KEYS:
        _close(params[k], gold["p_" + k], 1e-6)
        assert state[k]["exp_avg"].shape == params[k].shape
        assert float(state[k]["exp_avg"][case["P"]:].abs().max()) == 0.0
    for k in DO.VAR_KEYS:
        _same(variables[k], gold["var_" + k])
