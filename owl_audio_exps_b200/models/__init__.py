"""The callers either side of the DiT stack, mirroring owl_wms/models."""
from .audiorft import AudioRFT, AudioRFTCore  # noqa: F401
from .gamerft import GameRFT, GameRFTCore  # noqa: F401
from .gamerft_audio import GameRFTAudio, GameRFTAudioCore  # noqa: F401


def get_model_cls(model_id):
    """owl_wms/models/__init__.py:1-25 for the model families built on this path."""
    table = {"game_rft": GameRFT, "game_rft_audio": GameRFTAudio, "audio_rft": AudioRFT}
    if model_id not in table:
        raise NotImplementedError(f"model '{model_id}' is not built on the DiT-block path (game_mft_audio: mean-flow JVP)")
    return table[model_id]
