"""pokeai_b200 -- B200-native decision hot path of select766/pokeai.

Host classes keep the reference's names and signatures (pokeai/ai/generic_move_model/*):
FeatureExtractor, MLPModel, Agent / AgentVal / AgentTrain, ReplayBuffer, Trainer.  All tensor
work is done by hand-written sm_100a kernels in pokeai_b200/_lib/libpokeai_b200.so behind the
C ABI of include/pokeai_b200.h; there is no CPU or eager-PyTorch path.
"""
__version__ = "0.1.0"

_LAZY = {
    "FeatureExtractor": ".feature_extractor", "MLPModel": ".mlp_model", "Agent": ".agent", "AgentVal": ".agent",
    "AgentTrain": ".agent", "ReplayBuffer": ".replay_buffer", "ReplayBufferItem": ".replay_buffer",
    "Trainer": ".trainer", "DQN_DEFAULT_PARAMS": ".trainer", "DecisionSession": ".session", "DecisionPipeline": ".session", "pinned_records": ".session",
    "PartyEvaluator": ".party_evaluator", "PartyEvaluatorQuick": ".party_evaluator",
    "LockStepDriver": ".lockstep", "SyntheticBattle": ".lockstep",
    "PackedRLPolicy": ".protocol", "pack_request": ".protocol", "get_possible_actions": ".protocol", "pack_requests": ".protocol",
}


def __getattr__(name):
    if name in _LAZY:
        import importlib
        return getattr(importlib.import_module(_LAZY[name], __name__), name)
    raise AttributeError(name)
