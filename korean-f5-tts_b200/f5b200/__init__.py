"""f5b200 -- host-side mirror of the reference's Python interface for the CFM.sample + Vocos hot path, executing on
libf5b200.so (hand-written sm_100a CUDA).  Same names as `f5_tts.model` / `f5_tts.infer.utils_infer`:
CFM, DiT, UNetT, MelSpec, load_vocoder, lens_to_mask, list_str_to_idx, get_epss_timesteps."""
from ._lib import F5BError, LIB_PATH  # noqa: F401
from .backbones import DiT, UNetT  # noqa: F401
from .cfm import CFM  # noqa: F401
from .modules import MelSpec  # noqa: F401
from .utils import (get_epss_timesteps, lens_to_mask, list_str_to_idx, list_str_to_tensor,  # noqa: F401
                    seed_everything, sway_time_grid)
from .vocoder import Vocos, load_vocoder  # noqa: F401
from . import tokenizers  # noqa: F401
from .infer import infer_batch_process  # noqa: F401
from .checkpoint import get_tokenizer, load_checkpoint, load_model  # noqa: F401
from .api import F5TTS, chunk_text, infer_process  # noqa: F401
from . import serving  # noqa: F401
from .train import TrainStep  # noqa: F401

__all__ = ["CFM", "DiT", "UNetT", "MelSpec", "Vocos", "load_vocoder", "lens_to_mask", "list_str_to_idx",
           "list_str_to_tensor", "get_epss_timesteps", "sway_time_grid", "seed_everything", "F5BError", "infer_batch_process",
           "load_model", "load_checkpoint", "get_tokenizer", "F5TTS", "chunk_text", "infer_process"]
