"""Drop-in for owl_wms/models/gamerft_audio.py: video patches + one audio token per frame through one DiT / UViT stack.

Per frame the `h*w` video tokens are followed by the audio token (tokens_per_frame = h*w + 1); after the stack the two
streams are split again, video goes through `FinalLayer(layer_norm(video), layer_norm(cond))`, audio through its own
`FinalLayer`.  Module / parameter names are the reference's.  The reference wraps the backbone in `torch.compile`, which
prefixes its state-dict keys with `transformer._orig_mod.`; load / save hooks keep that spelling so checkpoints move both
ways.  `backbone: mmdit` (owl_wms/nn/mmattn.py) is outside this path and raises."""
import torch
import torch.nn.functional as F
from torch import nn

from .. import functional as OF
from ..nn.attn import DiT, FinalLayer, UViT
from ..nn.embeddings import ControlEmbedding, TimestepEmbedding
from ..nn.normalization import layer_norm
from .gamerft import cfg_dropout

_COMPILED = "transformer._orig_mod."


class GameRFTAudioCore(nn.Module):
    """gamerft_audio.py:19-98."""

    def __init__(self, config):
        super().__init__()
        self.config = config
        if config.backbone == "dit":
            backbone_cls = DiT
        elif config.backbone == "uvit":
            backbone_cls = UViT
        elif config.backbone == "mmdit":
            raise NotImplementedError("backbone 'mmdit' (owl_wms/nn/mmattn.py) is outside the DiT-block path")
        else:
            raise ValueError(f"Invalid backbone: {config.backbone}")
        self.backbone = config.backbone
        self.transformer = backbone_cls(config)
        if not config.uncond:
            self.control_embed = ControlEmbedding(config.n_buttons, config.d_model)
        self.t_embed = TimestepEmbedding(config.d_model)
        self.proj_in = nn.Linear(config.channels, config.d_model, bias=False)
        self.proj_out = FinalLayer(config.sample_size, config.d_model, config.channels)
        self.audio_proj_in = nn.Linear(config.audio_channels, config.d_model, bias=False)
        self.audio_proj_out = FinalLayer(None, config.d_model, config.audio_channels)
        self.uncond = config.uncond
        self._stage = OF.WeightStage()
        self._register_load_state_dict_pre_hook(self._strip_compiled_prefix)
        self._register_state_dict_hook(self._add_compiled_prefix)

    @staticmethod
    def _strip_compiled_prefix(state_dict, prefix, *args):
        for k in [k for k in state_dict if k.startswith(prefix + _COMPILED)]:
            state_dict[prefix + "transformer." + k[len(prefix + _COMPILED):]] = state_dict.pop(k)

    @staticmethod
    def _add_compiled_prefix(module, state_dict, prefix, local_metadata):
        plain = prefix + "transformer."
        items = list(state_dict.items())
        state_dict.clear()                                   # rebuilt in place: key order is part of the contract
        for k, v in items:
            if k.startswith(plain) and not k.startswith(prefix + _COMPILED):
                k = prefix + _COMPILED + k[len(plain):]
            state_dict[k] = v
        return state_dict

    def forward(self, x, audio, t, mouse, btn, has_controls=None, kv_cache=None):
        """x [b,n,c,h,w], audio [b,n,ca], t [b,n], mouse [b,n,2], btn [b,n,n_buttons] -> (video [b,n,c,h,w], audio [b,n,ca])."""
        cond = self.t_embed(t)
        if not self.uncond:
            ctrl = self.control_embed(mouse, btn)
            if has_controls is not None:
                ctrl = torch.where(has_controls[:, None, None], ctrl, torch.zeros_like(ctrl))
            cond = cond + ctrl
        b, n, c, h, w = x.shape
        d = self.config.d_model
        vid = OF.linear(x.permute(0, 1, 3, 4, 2).reshape(b, n * h * w, c), self.proj_in.weight, None, self._stage, "proj_in")
        aud = OF.linear(audio, self.audio_proj_in.weight, None, self._stage, "audio_proj_in")
        tok = torch.cat([vid.view(b, n, h * w, d), aud.view(b, n, 1, d)], dim=2).reshape(b, n * (h * w + 1), d)
        if self.backbone == "dit":
            tok = self.transformer(tok, cond, None, kv_cache)     # the reference passes kv_cache in DiT's doc_id slot
        else:                                                     # (gamerft_audio.py:73); the cache is what is meant
            tok = self.transformer(tok, cond, kv_cache)
        tok = tok.view(b, n, h * w + 1, d)
        vid = tok[:, :, :-1].reshape(b, n * h * w, d)
        aud = tok[:, :, -1]
        vid = self.proj_out(layer_norm(vid), layer_norm(cond))
        vid = vid.reshape(b, n, h, w, c).permute(0, 1, 4, 2, 3)
        return vid, self.audio_proj_out(aud, cond)


class GameRFTAudio(nn.Module):
    """gamerft_audio.py:100-175: joint video + audio rectified-flow loss with CFG control dropout."""

    def __init__(self, config):
        super().__init__()
        self.core = GameRFTAudioCore(config)
        self.cfg_prob = config.cfg_prob

    def handle_cfg(self, has_controls=None, cfg_prob=None):
        if cfg_prob is None:
            cfg_prob = self.cfg_prob
        if cfg_prob <= 0.0 or has_controls is None:
            return has_controls
        return cfg_dropout(has_controls, cfg_prob)

    def forward(self, x, audio, mouse, btn, return_dict=False, cfg_prob=None, has_controls=None):
        b, n = x.shape[:2]
        if has_controls is None:
            has_controls = torch.ones(b, device=x.device, dtype=torch.bool)
        has_controls = self.handle_cfg(has_controls, cfg_prob)
        with torch.no_grad():
            ts = torch.randn(b, n, device=x.device, dtype=x.dtype).sigmoid()
            tv, ta = ts[:, :, None, None, None], ts.unsqueeze(-1)
            z_video, z_audio = torch.randn_like(x), torch.randn_like(audio)
            lerpd_video, target_video = OF.rf_noise(x, z_video, ts)
            lerpd_audio, target_audio = OF.rf_noise(audio, z_audio, ts)
        pred_video, pred_audio = self.core(lerpd_video, lerpd_audio, ts, mouse, btn, has_controls)
        video_loss = OF.mse_loss(pred_video, target_video)
        audio_loss = OF.mse_loss(pred_audio, target_audio)
        diff_loss = video_loss + audio_loss
        if not return_dict:
            return diff_loss, video_loss, audio_loss
        return {"diffusion_loss": diff_loss, "video_loss": video_loss, "audio_loss": audio_loss,
                "lerpd_video": lerpd_video, "lerpd_audio": lerpd_audio, "pred_video": pred_video,
                "pred_audio": pred_audio, "ts": ts, "z_video": z_video, "z_audio": z_audio, "cfg_mask": has_controls}
