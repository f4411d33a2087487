"""Everything of ``vTTS.forward`` behind the encoder (scripts/model/vtts.py:84-107) on the B200 path:

    variance_adaptor -> decoder -> mel_linear -> postnet(+ residual)          [-> vocoder_infer]

``AcousticTail`` holds the four sub-modules under the reference's own attribute names, so the matching entries of a
vTTS checkpoint (``variance_adaptor.*``, ``decoder.*``, ``mel_linear.*``, ``postnet.*``) load with
``load_state_dict(..., strict=False)`` on the remaining ``encoder.*`` keys.  The text/visual encoder is out of scope
(SURVEY.md section 8): its output is this module's input.
"""
from __future__ import annotations

import torch
import torch.nn as nn

from .decoder import Decoder
from .postnet import PostNet, mel_linear
from .variance_adaptor import VarianceAdaptor


class AcousticTail(nn.Module):
    def __init__(self, preprocess_config, model_config):
        super().__init__()
        self.model_config = model_config
        self.variance_adaptor = VarianceAdaptor(preprocess_config, model_config)
        self.decoder = Decoder(model_config)
        self.mel_linear = nn.Linear(model_config["transformer"]["decoder_hidden"],
                                    preprocess_config["preprocessing"]["mel"]["n_mel_channels"])
        self.postnet = PostNet()

    def forward(self, encoder_output, src_masks, mel_masks=None, max_mel_len=None, e_targets=None, k_targets=None,
                d_targets=None, e_control=1.0, d_control=1.0):
        """encoder_output [B,T_src,H] fp32 CUDA (after the audiotype embedding add), src_masks [B,T_src] bool.
        Returns vTTS.forward's tuple from ``output`` to ``mel_lens`` without the two src entries the caller already
        has: (output, postnet_output, e_predictions, k_predictions, log_d_predictions, d_rounded, mel_masks, mel_lens)."""
        if self.training:
            raise RuntimeError("AcousticTail: the B200 path is inference only (call .eval())")
        nvtx = torch.cuda.nvtx.range             # a no-op unless a profiler is attached
        with nvtx("vo2w.variance_adaptor"):
            (output, e_predictions, k_predictions, log_d_predictions, d_rounded, mel_lens, mel_masks) = self.variance_adaptor(
                encoder_output, src_masks, mel_masks, max_mel_len, e_targets, k_targets, d_targets, e_control, d_control)
        with nvtx("vo2w.decoder"):
            output, mel_masks = self.decoder(output, mel_masks)
        with nvtx("vo2w.mel_linear_postnet"):
            output = mel_linear(output, self.mel_linear.weight, self.mel_linear.bias)
            postnet_output = self.postnet(output, add_input=True)        # postnet(output) + output, one pass
        return (output, postnet_output, e_predictions, k_predictions, log_d_predictions, d_rounded, mel_masks, mel_lens)
