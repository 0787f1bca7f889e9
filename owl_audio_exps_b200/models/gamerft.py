"""Drop-in for owl_wms/models/gamerft.py: patchify -> proj_in -> DiT -> FinalLayer -> unpatchify, conditioned on
timestep (+ mouse/button controls unless `uncond`).  Module / parameter names are the reference's (core.transformer,
core.control_embed, core.t_embed, core.proj_in, core.proj_out)."""
import torch
import torch.nn.functional as F
from torch import nn

from .. import functional as OF
from ..nn.attn import DiT, FinalLayer
from ..nn.embeddings import ControlEmbedding, TimestepEmbedding


def cfg_dropout(has_controls, cfg_prob: float):
    """gamerft.py:69-91: drop controls from enough samples that at least `cfg_prob` of the batch is control-free.  Written
    without the reference's data-dependent Python branch (a device sync that cannot sit inside a CUDA-graph capture): when
    the batch already has enough control-free samples `need` is 0 and nothing is dropped."""
    have = has_controls.float().mean()
    need = (cfg_prob - (1.0 - have)).clamp_min(0.0)
    frac = need / have.clamp_min(1e-12)
    drop = (torch.rand(has_controls.shape[0], device=has_controls.device) <= frac) & has_controls & (need > 0)
    return has_controls & (~drop)


class GameRFTCore(nn.Module):
    """gamerft.py:13-60."""

    def __init__(self, config):
        super().__init__()
        self.config = config
        assert config.backbone == "dit"
        self.transformer = DiT(config)
        if not config.uncond:
            self.control_embed = ControlEmbedding(config.n_buttons, config.d_model)
        self.t_embed = TimestepEmbedding(config.d_model)
        self.proj_in = nn.Linear(config.channels, config.d_model, bias=False)
        self.proj_out = FinalLayer(config.sample_size, config.d_model, config.channels)
        assert config.tokens_per_frame == config.sample_size ** 2
        self.uncond = config.uncond
        self._stage = OF.WeightStage()

    def forward(self, x, t, mouse, btn, doc_id=None, has_controls=None, kv_cache=None, local_block_mask=None,
                global_block_mask=None):
        """x [b,n,c,h,w], t [b,n], mouse [b,n,2], btn [b,n,n_buttons] -> [b,n,c,h,w]."""
        b, n, c, h, w = x.shape
        cond = self.t_embed(t)
        if not self.uncond:
            ctrl = self.control_embed(mouse, btn)
            if has_controls is not None:
                ctrl = torch.where(has_controls[:, None, None], ctrl, torch.zeros_like(ctrl))
            cond = cond + ctrl
        x = x.permute(0, 1, 3, 4, 2).reshape(b, n * h * w, c)                  # 'b n c h w -> b (n h w) c'
        x = OF.linear(x, self.proj_in.weight, None, self._stage, "proj_in")
        x = self.transformer(x, cond, doc_id, kv_cache, local_block_mask, global_block_mask)
        x = self.proj_out(x, cond)
        return x.view(b, n, h, w, c).permute(0, 1, 4, 2, 3)                    # 'b (n h w) c -> b n c h w'


class GameRFT(nn.Module):
    """gamerft.py:63-125: rectified-flow training wrapper with classifier-free-guidance control dropout."""

    def __init__(self, config):
        super().__init__()
        self.config = config
        self.core = GameRFTCore(config)

    def handle_cfg(self, has_controls=None, cfg_prob=None):
        if cfg_prob is None:
            cfg_prob = self.config.cfg_prob
        if cfg_prob <= 0.0 or has_controls is None:
            return has_controls
        return cfg_dropout(has_controls, cfg_prob)

    def noise(self, tensor, ts):
        z = torch.randn_like(tensor)
        lerpd, target = OF.rf_noise(tensor, z, ts.reshape(ts.shape[:2]))     # one pass for both (bf16 latents)
        return lerpd, target, z

    def forward(self, x, mouse=None, btn=None, doc_id=None, return_dict=False, cfg_prob=None, has_controls=None):
        B, S = x.size(0), x.size(1)
        if has_controls is None:
            has_controls = torch.ones(B, device=x.device, dtype=torch.bool)
        if mouse is None or btn is None:
            has_controls = torch.zeros_like(has_controls)
        has_controls = self.handle_cfg(has_controls, cfg_prob)
        with torch.no_grad():
            ts = torch.randn(B, S, device=x.device, dtype=x.dtype).sigmoid()
            lerpd_video, target_video, z_video = self.noise(x, ts[:, :, None, None, None])
        pred_video = self.core(lerpd_video, ts, mouse, btn, doc_id, has_controls)
        loss = OF.mse_loss(pred_video, target_video)
        if not return_dict:
            return loss
        return {"diffusion_loss": loss, "video_loss": loss, "lerpd_video": lerpd_video, "pred_video": pred_video,
                "ts": ts, "z_video": z_video, "cfg_mask": has_controls}
