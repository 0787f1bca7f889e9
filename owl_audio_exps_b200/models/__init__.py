"""The callers either side of the DiT stack, mirroring owl_wms/models."""
from .audiorft import AudioRFT, AudioRFTCore  # noqa: F401
from .gamerft import GameRFT, GameRFTCore  # noqa: F401
