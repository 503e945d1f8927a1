"""Work figures of a config: parameter count (the reference's analyze_model,
src/run/utils_run.py:65-113, counts the same variables) and the algorithmic FLOPs per token."""
from .dataclass import ModelParameter


def _sizes(params: ModelParameter):
    H, K = params.heads, params.features_per_head
    F = H * K
    I = params.intermediate[0].size
    K2 = K * params.group_linear_factor
    I0 = int(I * params.vocab_weight_factorization)
    return H, K, F, I, K2, I0, params.vocab_size, params.depth, params.sequence_length


def count_params(params: ModelParameter) -> int:
    H, K, F, I, K2, I0, V, L, S = _sizes(params)
    per_depth = F * I + I * H * K2 + H * K2 * K + 2 * H * K2 + 6 * H * K
    return L * per_depth + 2 * H * S * S + V * I0 + I0 * F + F * V


def flops_per_token_fwd(params: ModelParameter) -> float:
    """Causal-aware forward FLOPs per token; a training step is credited 3x this (forward + dgrad + wgrad); the
    reversible recompute is NOT credited."""
    H, K, F, I, K2, I0, V, L, S = _sizes(params)
    return L * (2 * (F * I + I * H * K2 + H * K2 * K) + 2 * 2 * F * (S + 1) / 2) + 2 * (I0 * F + F * V)
