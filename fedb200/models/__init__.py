"""Model families of the reference: BERT-base/BioBERT/BERT-large and ALBERT-base-v2 (C4)."""
from .bert import BertForSequenceClassification  # noqa: F401
from .albert import AlbertForSequenceClassification  # noqa: F401
from .arena import FlatArena, param_layout  # noqa: F401
from .factory import build_model, SequenceClassifierOutput  # noqa: F401
from .hf_io import save_pretrained, load_pretrained, dir_size_gib  # noqa: F401
