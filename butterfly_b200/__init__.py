"""butterfly_b200: B200-native (sm_100a) partitioned transformer block forward.

Public API:
    make_plan(cfg, tp) -> PartitionPlan
    shard_weights(source, plan, rank) -> RankWeights
    Model(plan, rank, weights | SyntheticSource, ...).prefill(tokens) / .decode(tokens) / .decode_greedy(n)

The compute path is libbutterfly_b200.so (hand-written CUDA behind a C ABI, include/butterfly_b200.h).
There is no CPU fallback: constructing a Model without the built library or without a GPU raises.
"""
from .config import (ModelConfig, GPT2_SMALL, LLAMA3_8B, LLAMA3_70B, MIXTRAL_8X7B, PRESETS, tiny)  # noqa: F401
from .plan import PartitionPlan, RankPlan, PlanError, make_plan  # noqa: F401
from .weights import (SyntheticSource, DictSource, RankWeights, shard_weights, shard_spec,  # noqa: F401
                      synthetic_tokens)


from .checkpoint import load_checkpoint, from_hf_state_dict  # noqa: F401,E402


def __getattr__(name):
    if name == "Model":
        from .model import Model
        return Model
    raise AttributeError(name)
