"""Drop-in for the vocoder glue of the reference (scripts/utils/model.py:41-98)."""
from __future__ import annotations

import json
import os

import numpy as np
import torch

from . import hifigan


def get_vocoder(config, device, hifigan_config_path: str = "scripts/hifigan/config.json", ckpt_path: str | None = None,
                state_dict=None, precision: str | None = None):
    """scripts/utils/model.py:41-70 for ``config["vocoder"]["model"] == "HiFi-GAN"``.

    Same sequence: json -> AttrDict -> Generator -> load_state_dict(ckpt["generator"]) -> eval ->
    remove_weight_norm -> to(device).  ``state_dict`` (234-key g/v or 156-key folded) may be passed
    directly because the pretrained blob is not in the tree; MelGAN is out of scope.
    """
    name = config["vocoder"]["model"]
    speaker = config["vocoder"]["speaker"]
    if name != "HiFi-GAN":
        raise NotImplementedError(f"vocoder {name!r}: only HiFi-GAN is on the B200 path")
    with open(hifigan_config_path, "r") as f:
        h = hifigan.AttrDict(json.load(f))
    vocoder = hifigan.Generator(h, precision=precision)
    if state_dict is None:
        if ckpt_path is None:
            ckpt_path = {"LJSpeech": "scripts/hifigan/generator_LJSpeech.pth.tar",
                         "universal": "scripts/hifigan/generator_universal.pth.tar"}[speaker]
        state_dict = torch.load(ckpt_path, map_location="cpu")["generator"]
    folded = any(k.endswith(".weight") for k in state_dict)
    if folded:
        vocoder.eval()
        vocoder.remove_weight_norm()
        vocoder.load_state_dict(state_dict)
    else:
        vocoder.load_state_dict(state_dict)
        vocoder.eval()
        vocoder.remove_weight_norm()
    vocoder.to(device)
    return vocoder


def vocoder_infer(mels, vocoder, model_config, preprocess_config, lengths=None, Normalize=True, out=None):
    """scripts/utils/model.py:73-98.  ``mels`` [B,80,T] on the vocoder's device, or on the host.

    Host mels go through the C ABI's host-buffer entry (H2D, kernels, D2H in one call, pinned
    staging); device mels through ``vocoder(mels)`` then ``.cpu().numpy()`` as the reference does.
    ``out`` (beyond the reference; host mels only): a caller-owned, ideally pinned host tensor [B, T*hop] the
    waveforms are copied into directly; the returned arrays are then views of it.
    """
    name = model_config["vocoder"]["model"]
    if name != "HiFi-GAN":
        raise NotImplementedError(f"vocoder {name!r}: only HiFi-GAN is on the B200 path")
    scale = None
    if Normalize:   # same key path as the reference (KeyError with the ICASSP preprocess.yaml)
        scale = preprocess_config["preprocessing"]["audio"]["max_wav_value"]
    with torch.no_grad():
        if not mels.is_cuda:
            res = vocoder.infer_host(mels.float(), int16_scale=scale, out=out)
            wavs = res.numpy() if vocoder.host_out_owned else res.numpy().copy()
        else:
            wavs = vocoder(mels).squeeze(1)
            if Normalize:
                wavs = (wavs.cpu().numpy() * scale).astype("int16")
            else:
                wavs = wavs.cpu().numpy()
    wavs = [wav for wav in wavs]
    for i in range(len(mels)):
        if lengths is not None:
            wavs[i] = wavs[i][: lengths[i]]
    return wavs
