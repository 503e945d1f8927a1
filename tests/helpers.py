"""Shared helpers for the parity tests: build a ModelParameter that IMPORTS the oracle's seeded weights
(the reference seeds its initialisers from Python's global `random`, so parity runs import weights on both sides,
SURVEY.md 8(a) quirk 2)."""
import copy

import numpy as np
import torch

from oracle import obst_oracle as orc


def make_params(config: dict, oracle_params: dict, **overrides):
    from obst.dataclass import ModelParameter, gpt_mode
    cfg = gpt_mode(copy.deepcopy(config))
    cfg.update(overrides)
    params = ModelParameter(cfg)
    sh = orc.Shapes(config)
    params.import_variables = [oracle_params[n] for n in orc.param_names(sh)]
    return params


def variables_by_oracle_name(params, sh):
    """creation order of the product's variables == orc.param_names(sh)."""
    names = orc.param_names(sh)
    variables = list(params.variable_cache.values())
    assert len(variables) == len(names), (len(variables), len(names))
    return dict(zip(names, variables))


def rel(a, b):
    a = np.asarray(a, dtype=np.float64)
    b = np.asarray(b, dtype=np.float64)
    return float(np.linalg.norm(a - b) / (np.linalg.norm(b) + 1e-30))


def to_np(t: torch.Tensor):
    return t.detach().float().cpu().numpy()
