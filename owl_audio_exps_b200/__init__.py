"""owl_audio_exps_b200: B200-native (sm_100a) implementation of the owl_wms DiT-block hot path.

Layout: csrc/ (CUDA kernels + C ABI, built into libowl_dit.so), _lib.py (ctypes binding), ops.py
(tensor-level wrappers), nn/ (drop-in mirror of owl_wms/nn), engine.py (data-parallel step).
"""
__version__ = "0.1.0"
