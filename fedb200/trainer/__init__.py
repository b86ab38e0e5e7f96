"""Local trainer / evaluator (C5, C6) and the HF-semantics AdamW (K13 host reference)."""
from .adamw import FlatAdamW, hf_adamw_reference_step  # noqa: F401
from .local import TorchRunner, make_runner, accuracy_from_logits  # noqa: F401
