"""plaincv_b200 — the training-step hot path of GeorgTirp/plainCV's models/LM causal Transformer on B200 (sm_100a).

Python here is host plumbing over libplaincv_b200.so (hand-written CUDA, C ABI in include/plaincv_b200.h).
There is no CPU or PyTorch fallback; importing the compute modules without the built library raises.
"""
from .batch import doc_lengths_to_doc_start, doc_start_from_mask, prepare_batch, shard_for_rank, trim_last_token  # noqa: F401

__all__ = ["construct_model", "Transformer", "ModelConfig", "Engine", "prepare_batch"]


def __getattr__(name):  # lazy: keep `import plaincv_b200` cheap and CPU-importable for the host-logic tests
    if name in ("construct_model", "Transformer", "ModelConfig", "model_config_from_cfg", "ParamLayout"):
        from . import model
        return getattr(model, name)
    if name in ("Engine", "GradReducer"):
        from . import engine
        return getattr(engine, name)
    raise AttributeError(name)
