"""B200-native synthesis hot path of sarulab-speech/visual-onoma-to-wave.

Drop-in mirrors of the reference's Python surface for this path only:

    hifigan.Generator / hifigan.AttrDict     (scripts/hifigan/models.py, scripts/hifigan/__init__.py)
    LengthRegulator, get_mask_from_lengths   (scripts/model/modules.py:126-159, scripts/utils/tools.py:164-171)
    get_vocoder, vocoder_infer               (scripts/utils/model.py:41-98)
    PostNet, ConvNorm, mel_linear            (scripts/transformer/Layers.py:32-137, scripts/model/vtts.py:106-107)
    VarianceAdaptor, VariancePredictor, Conv   (scripts/model/modules.py:15-124, 161-262)
    Decoder (+ FFTBlock, MultiHeadAttention, PositionwiseFeedForward holders)   (scripts/transformer/Models.py:134-197)

All compute runs in hand-written sm_100a CUDA kernels behind the C ABI of include/vo2w.h
(libvo2w.so, loaded with ctypes).  There is no CPU or PyTorch fallback.
"""
from . import hifigan  # noqa: F401
from .hifigan import AttrDict, Generator  # noqa: F401
from .length_regulator import LengthRegulator, get_mask_from_lengths, round_durations  # noqa: F401
from .vocoder import get_vocoder, vocoder_infer  # noqa: F401
from .postnet import PostNet, ConvNorm, mel_linear  # noqa: F401
from .variance_adaptor import VarianceAdaptor, VariancePredictor, Conv  # noqa: F401
from .acoustic_tail import AcousticTail  # noqa: F401
from .decoder import Decoder, FFTBlock, MultiHeadAttention, PositionwiseFeedForward  # noqa: F401
from .sharding import shard_bounds  # noqa: F401

__all__ = ["hifigan", "AttrDict", "Generator", "LengthRegulator", "get_mask_from_lengths", "round_durations",
           "get_vocoder", "vocoder_infer", "shard_bounds", "PostNet", "ConvNorm", "mel_linear",
           "AcousticTail", "VarianceAdaptor", "VariancePredictor", "Conv", "Decoder", "FFTBlock", "MultiHeadAttention", "PositionwiseFeedForward"]
