"""Drop-in for owl_wms/nn/attn.py: get_block_mask, Attn, DiTBlock, DiT, SkipConnection, UViT, FinalLayer.

Module names, constructor signatures, attributes and state-dict keys are the reference's
(`blocks.{i}.attn.qkv.*`, `.attn.out.*`, `.mlp.fc1/fc2.*`, `.adaln1/2.fc.*`, `.gate1/2.fc_c.*`; RoPE
cos/sin are non-persistent buffers), so checkpoints, Muon's key matching and the samplers work
unchanged.  What differs is underneath: `DiTBlock.forward` is ONE autograd node
(functional.DiTBlockFn) that sequences hand-written sm_100a kernels, and the BlockMask of
attn.py:24-62 is a closed-form descriptor (functional.FrameMask) evaluated inside the attention
kernel with tile-granular skipping, so nothing O(L^2/128^2) is built per forward (the reference
calls create_block_mask twice per DiT.forward, attn.py:179-182).
"""
import torch
from torch import nn
from torch.utils.checkpoint import checkpoint as torch_checkpoint

from .. import functional as OF
from ..functional import FrameMask
from .mlp import MLP
from .modulation import AdaLN, Gate
from .normalization import rms_norm  # noqa: F401  (re-exported like the reference module)
from .rope import get_rope_cls


def checkpoint(function, *args, **kwargs):
    kwargs.setdefault("use_reentrant", False)
    return torch_checkpoint(function, *args, **kwargs)


def get_block_mask(n_tokens: int, tokens_per_frame: int, window_len=None, doc_id=None, q_offset: int = 0,
                   is_causal: bool = True, device="cpu"):
    """attn.py:24-62.  Returns a FrameMask descriptor instead of a flex BlockMask; `.mask_mod(b,h,q,kv)`
    evaluates the same predicate for dense rendering."""
    assert 0 <= q_offset < n_tokens, "kv cache cannot exceed total tokens"
    if not is_causal:
        assert q_offset == 0, "kv caching not supported with bidirectional"
    window = 0 if window_len is None else int(window_len)  # None -> n_frames in the reference = never binding
    doc_ext = None
    if doc_id is not None:
        doc_id = doc_id.to(device=device, dtype=torch.int32).contiguous()
        if doc_id.is_cuda:
            # document extents once per mask (tiny kernel): tiles that hold only other documents are then skipped
            doc_ext = _doc_extents(doc_id)
    return FrameMask(tokens_per_frame=int(tokens_per_frame), window_frames=window, q_offset=int(q_offset),
                     causal=bool(is_causal), doc_id=doc_id, doc_ext=doc_ext)


_DOC_EXT_CACHE = {}


def _doc_extents(doc_id):
    """DiT.forward builds the local and the global mask from the same doc_id (attn.py:179-182): one launch serves both."""
    from .. import ops
    key = (doc_id.data_ptr(), doc_id._version, tuple(doc_id.shape))
    hit = _DOC_EXT_CACHE.get("last")
    if hit is not None and hit[0] == key and hit[1] is doc_id:
        return hit[2]
    ext = ops.attn_doc_extents(doc_id)
    _DOC_EXT_CACHE["last"] = (key, doc_id, ext)
    return ext


def _check_mask(block_mask):
    if block_mask is not None and not isinstance(block_mask, FrameMask):
        raise TypeError("block_mask must come from this package's get_block_mask / DiT.get_block_mask "
                        "(a FrameMask descriptor); torch flex BlockMask objects are not consumed by the sm_100a kernels")
    return block_mask


class Attn(nn.Module):
    def __init__(self, config, layer_idx, local=False):
        super().__init__()
        self.config = config
        self.layer_idx = layer_idx
        self.n_heads = config.n_heads

        self.qkv = nn.Linear(config.d_model, 3 * config.d_model)
        self.out = nn.Linear(config.d_model, config.d_model)
        self.rope = get_rope_cls(getattr(config, "rope_impl", "ortho"))(config)

        self.local = local
        self.local_offset = config.local_window * config.tokens_per_frame
        self._stage = OF.WeightStage()

    def plan(self, block_mask, kv_cache):
        return OF.AttnPlan(n_heads=self.n_heads, cos=self.rope.cos, sin=self.rope.sin, mask=_check_mask(block_mask),
                           local=self.local, local_offset=self.local_offset, kv_cache=kv_cache,
                           layer_idx=self.layer_idx)

    def forward(self, x, block_mask, kv_cache=None):
        """attn.py:79-113 as three autograd nodes: qkv GEMM, attention core, out GEMM."""
        qkv = OF.linear(x, self.qkv.weight, self.qkv.bias, self._stage, "qkv")
        o = OF.attention(qkv, self.plan(block_mask, kv_cache))
        return OF.linear(o, self.out.weight, self.out.bias, self._stage, "out")


class DiTBlock(nn.Module):
    def __init__(self, config, layer_idx, local=False):
        super().__init__()
        dim = config.d_model
        self.attn = Attn(config, layer_idx, local)
        self.mlp = MLP(config)
        self.adaln1 = AdaLN(dim)
        self.gate1 = Gate(dim)
        self.adaln2 = AdaLN(dim)
        self.gate2 = Gate(dim)
        self._stage = OF.WeightStage()

    def block_params(self):
        return (self.attn.qkv.weight, self.attn.qkv.bias, self.attn.out.weight, self.attn.out.bias,
                self.mlp.fc1.weight, self.mlp.fc1.bias, self.mlp.fc2.weight, self.mlp.fc2.bias,
                self.adaln1.fc.weight, self.adaln1.fc.bias, self.gate1.fc_c.weight, self.gate1.fc_c.bias,
                self.adaln2.fc.weight, self.adaln2.fc.bias, self.gate2.fc_c.weight, self.gate2.fc_c.bias)

    def cond_params(self):
        return self.block_params()[8:]

    def forward(self, x, cond, block_mask, kv_cache=None, scond=None, bank=None, token=None, light=False):
        """attn.py:130-143.  Extensions used by DiT.forward: `scond` = silu(cond) precomputed once per
        stack forward (the reference recomputes it in each of the block's four modulation layers,
        modulation.py:17,38); `bank`/`token` = the stack's batched conditioning projections
        (functional.CondBank) -- the block then reads its modulation vectors instead of projecting them."""
        return OF.DiTBlockFn.apply(x, cond, scond, token, self.attn.plan(block_mask, kv_cache), self._stage, bank, light,
                                   *self.block_params())


class DiT(nn.Module):
    def __init__(self, config):
        super().__init__()
        self.config = config
        if not hasattr(config, "local_idx"):
            config.local_idx = 4
        self.local_layers = [(layer_idx % config.local_idx != 0) for layer_idx in range(config.n_layers)]
        self.blocks = nn.ModuleList([DiTBlock(config, idx, local) for idx, local in enumerate(self.local_layers)])
        self.decoding = False
        self.cond_group = 4          # layers per conditioning-bank weight-gradient GEMM (functional.CondBank)
        self._stage = OF.WeightStage()

    def enable_decoding(self):
        self.decoding = True

    def disable_decoding(self):
        self.decoding = False

    def get_block_mask(self, seq_len, doc_id, window_len, q_offset, device):
        n_tokens = seq_len + q_offset
        return get_block_mask(n_tokens=n_tokens, tokens_per_frame=self.config.tokens_per_frame,
                              window_len=window_len, doc_id=doc_id, q_offset=q_offset,
                              is_causal=self.config.causal, device=device)

    def forward(self, x, cond, doc_id=None, kv_cache=None, local_block_mask=None, global_block_mask=None):
        """attn.py:175-191."""
        from .decode import RingKVCache, dit_decode_forward
        if isinstance(kv_cache, RingKVCache):
            # device-state ring cache: the gradient-free decode schedule (nn/decode.py); decode mode only (no mask)
            if not self.decoding:
                raise RuntimeError("RingKVCache serves decode mode only: call enable_decoding() first (attn.py:102-109)")
            return dit_decode_forward(self, x, cond, kv_cache)
        seq_len, device = x.size(1), x.device
        q_offset = kv_cache.length_at(0) if kv_cache is not None else 0

        if local_block_mask is None and not self.decoding:
            local_block_mask = self.get_block_mask(seq_len, doc_id, self.config.local_window, q_offset, device)
        if global_block_mask is None and not self.decoding:
            global_block_mask = self.get_block_mask(seq_len, doc_id, self.config.global_window, q_offset, device)

        scond = OF.silu(cond)  # layer-invariant: once per forward instead of 4x per block
        ckpt = self.training and getattr(self.config, "gradient_checkpointing", False)
        # gradient checkpointing (attn.py:187-188): `light` keeps block input + attention output + lse and re-runs the rest
        # of the block in backward WITHOUT the attention kernel (functional.DiTBlockFn); `torch` = torch.utils.checkpoint per
        # block as the reference does (recomputes the attention too)
        light = ckpt and getattr(self, "checkpoint_impl", "light") == "light" and kv_cache is None
        ckpt = ckpt and not light
        bank = token = None
        if not ckpt and self.cond_group > 0:
            # all 4 x n_layers conditioning projections as one GEMM (cond is the same for every block)
            bank, token = OF.cond_bank(scond, [b.cond_params() for b in self.blocks], self.config.d_model,
                                       self._stage, self.cond_group)
            # the blocks below form a linear chain on a bf16 residual stream: the passes either side of every layer
            # boundary fuse (functional.CondBank.pre / tail / pre_bwd)
            bank.chain = x.dtype == torch.bfloat16 and getattr(self, "chain_fusion", True) and not light
        for layer_idx, block in enumerate(self.blocks):
            block_mask = local_block_mask if self.local_layers[layer_idx] else global_block_mask
            if ckpt:
                x = checkpoint(block, x, cond, block_mask, kv_cache, scond)
            else:
                x = block(x, cond, block_mask, kv_cache, scond, bank, token, light)
        return x


class SkipConnection(nn.Module):
    def __init__(self, config):
        super().__init__()
        self.norm = AdaLN(config.d_model)
        self.proj = nn.Linear(config.d_model, config.d_model)
        self._stage = OF.WeightStage()

    def forward(self, x, prev, cond):
        x = x + prev
        x = self.norm(x, cond)
        return OF.linear(x, self.proj.weight, self.proj.bias, self._stage, "proj")


class UViT(nn.Module):
    """attn.py:206-260.  The reference class cannot be constructed at this commit (it builds
    `DiTBlock(config)` without layer_idx and calls get_block_mask with the wrong arity, SURVEY.md
    Appendix B); this version keeps the layout (blocks + skip_projs) with the obvious repairs."""

    def __init__(self, config):
        super().__init__()
        self.config = config
        self.tokens_per_frame = config.tokens_per_frame
        self.causal = config.causal
        self.blocks = nn.ModuleList([DiTBlock(config, i, False) for i in range(config.n_layers)])
        self.skip_projs = nn.ModuleList([SkipConnection(config) for _ in range(config.n_layers // 2)])

    def forward(self, x, cond, kv_cache=None):
        q_offset = kv_cache.length_at(0) if kv_cache is not None else 0
        block_mask = get_block_mask(x.size(1) + q_offset, self.tokens_per_frame, None, None, q_offset, self.causal,
                                    x.device)
        early = []
        n_blocks = len(self.blocks)
        mid = n_blocks // 2
        for i in range(mid):
            x = self.blocks[i](x, cond, block_mask, kv_cache)
            early.append(x)
        x = self.blocks[mid](x, cond, block_mask, kv_cache)
        for i in range(mid + 1, n_blocks):
            x = self.skip_projs[i - (mid + 1)](x, early[n_blocks - 1 - i], cond)
            x = self.blocks[i](x, cond, block_mask, kv_cache)
        return x


class FinalLayer(nn.Module):
    """attn.py:264-277: AdaLN -> SiLU -> Linear(d, channels)."""

    def __init__(self, sample_size, d_model, channels=3, patch_size=1):
        super().__init__()
        self.norm = AdaLN(d_model)
        self.act = nn.SiLU()
        self.proj = nn.Linear(d_model, channels * patch_size * patch_size)
        self._stage = OF.WeightStage()

    def forward(self, x, cond):
        x = self.norm(x, cond)
        return OF.linear(x, self.proj.weight, self.proj.bias, self._stage, "proj", silu_in=True)
