"""llie-b200: B200-native MSRCR Retinex and Zero-DCE behind one operator, `enhance()`.

Python is the host language BASELINE.json#north_star names; all compute happens in the
hand-written sm_100a kernels of libllie.so, reached through the C ABI of include/llie.h.
PyTorch is used only for device memory, pinned staging and streams.
"""
from .api import (enhance, enhance_sharded, shard_last_ms, enhance_stream, Enhancer, default_msrcr_params, default_dce_params, init_dce_weights,
                  shard_slice, plan_query, msrcr_tiled, tile_rows, load_dce_state_dict)
from ._lib import LlieError, LlieCudaError

__all__ = ["enhance", "enhance_sharded", "shard_last_ms", "enhance_stream", "Enhancer", "default_msrcr_params", "default_dce_params", "init_dce_weights",
           "shard_slice", "plan_query", "msrcr_tiled", "tile_rows", "load_dce_state_dict", "LlieError",
           "LlieCudaError"]
__version__ = "0.1.0"
