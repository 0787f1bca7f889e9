"""Drop-in for owl_wms/models/audiorft.py: proj_in -> DiT -> FinalLayer around the block stack, conditioned on the
per-latent timestep embedding.  Same module / parameter names (core.transformer, core.t_embed, core.proj_in,
core.proj_out) so reference checkpoints load with load_state_dict.  Every matmul is the tcgen05 GEMM, the noise
lerp and the MSE are plain elementwise torch (section 8f rank 4 is their fusion)."""
import torch
import torch.nn.functional as F
from torch import nn

from .. import functional as OF
from ..nn.attn import DiT, FinalLayer
from ..nn.embeddings import TimestepEmbedding


class AudioRFTCore(nn.Module):
    """audiorft.py:13-57."""

    def __init__(self, config):
        super().__init__()
        self.config = config
        assert config.backbone == "dit"
        self.transformer = DiT(config)
        self.t_embed = TimestepEmbedding(config.d_model)                        # unconditional: timestep only
        self.proj_in = nn.Linear(config.channels, config.d_model, bias=False)
        self.proj_out = FinalLayer(1, config.d_model, config.channels)         # one token per latent
        assert config.tokens_per_frame == 1
        self._stage = OF.WeightStage()

    def forward(self, x, t, doc_id=None, kv_cache=None, local_block_mask=None, global_block_mask=None):
        """x: [b, n_latents, channels] latents, t: [b, n_latents] timesteps -> [b, n_latents, channels]."""
        t_cond = self.t_embed(t)
        x = OF.linear(x, self.proj_in.weight, None, self._stage, "proj_in")
        x = self.transformer(x, t_cond, doc_id, kv_cache, local_block_mask, global_block_mask)
        return self.proj_out(x, t_cond)


class AudioRFT(nn.Module):
    """audiorft.py:60-97: rectified-flow training wrapper (per-latent sigmoid-normal timesteps, velocity target)."""

    def __init__(self, config):
        super().__init__()
        self.config = config
        self.core = AudioRFTCore(config)

    def noise(self, tensor, ts):
        z = torch.randn_like(tensor)
        lerpd, target = OF.rf_noise(tensor, z, ts.reshape(ts.shape[:2]))     # one pass for both (bf16 latents)
        return lerpd, target, z

    def forward(self, x, doc_id=None, return_dict=False):
        B, n_latents, _ = x.shape
        with torch.no_grad():
            ts = torch.randn(B, n_latents, device=x.device, dtype=x.dtype).sigmoid()
            lerpd_audio, target_audio, z_audio = self.noise(x, ts[:, :, None])
        pred_audio = self.core(lerpd_audio, ts, doc_id)
        loss = OF.mse_loss(pred_audio, target_audio)
        if not return_dict:
            return loss
        return {"diffusion_loss": loss, "audio_loss": loss, "lerpd_audio": lerpd_audio, "pred_audio": pred_audio,
                "ts": ts, "z_audio": z_audio}
