"""Mirror of the reference's `owl_wms/nn` package surface for the DiT-block hot path."""
from . import attn, embeddings, kv_cache, mlp, modulation, normalization, rope  # noqa: F401
from .attn import DiT, DiTBlock, Attn, FinalLayer, SkipConnection, UViT, get_block_mask, checkpoint  # noqa: F401
from .embeddings import (LearnedPosEnc, SinCosEmbed, TimestepEmbedding, StepEmbedding, ConditionEmbedding,  # noqa: F401
                         MouseEmbedding, ButtonEmbeddding, ControlEmbedding)
from .kv_cache import KVCache, SingleKVCache  # noqa: F401
from .mlp import MLP, MLPCustom  # noqa: F401
from .modulation import AdaLN, Gate, cond_adaln, cond_gate  # noqa: F401
from .normalization import layer_norm, rms_norm, RMSNorm, L2Norm, QKNorm  # noqa: F401
from .rope import RoPE, OrthoRoPE, MotionRoPE, Audio1DRoPE, get_rope_cls, cast_rope_buffers_to_fp32  # noqa: F401
