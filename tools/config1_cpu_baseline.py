"""configs[0] of BASELINE.json timed through the REFERENCE itself (build container only: /root/reference cannot travel):
single-utterance end-to-end synthesis on CPU, sample/ipaexg_24pt visual-onomatopoeia image + source id -> vTTS -> mel ->
HiFi-GAN V1 waveform, batch 1, random-init weights.  Follows prediction.ipynb cell 5 (on_click_synthesize, l.78-117) and
the recipe of SURVEY.md section 8(c): glyph widths [(106 + i) // 4 for i in range(4)], each padded to the dataset's
width with 255 (dataset.py:71-92), class whistle3, text "ピィピィ", d_control = 30 (with random-init weights the
duration predictor would otherwise give a 1-frame mel).  The number goes into BASELINE.md section 4.

    python tools/config1_cpu_baseline.py [--threads N] [--reps 5]
"""
import argparse
import json
import os
import sys
import time
from unittest.mock import MagicMock

import numpy as np
import torch

REF = "/root/reference"


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--threads", type=int, default=os.cpu_count())
    ap.add_argument("--reps", type=int, default=5)
    ap.add_argument("--d-control", type=float, default=30.0)
    args = ap.parse_args()
    torch.set_num_threads(args.threads)
    for m in ("matplotlib", "matplotlib.pyplot", "librosa", "librosa.util", "librosa.filters"):
        sys.modules.setdefault(m, MagicMock())
    os.chdir(REF)                                    # the reference uses cwd-relative paths
    sys.path.insert(0, os.path.join(REF, "scripts"))
    import yaml
    from PIL import Image
    import hifigan
    from utils.model import get_model, vocoder_infer
    from utils.tools import to_device
    from dataset import Dataset

    cp = yaml.load(open("config/ICASSP/preprocess.yaml"), Loader=yaml.FullLoader)
    cm = yaml.load(open("config/ICASSP/model.yaml"), Loader=yaml.FullLoader)
    ct = yaml.load(open("config/ICASSP/train.yaml"), Loader=yaml.FullLoader)
    dev = torch.device("cpu")
    torch.manual_seed(1234)
    model = get_model(0, (cp, cm, ct), dev)          # restore_step 0: random init (no checkpoint ships)
    h = hifigan.AttrDict(json.load(open("scripts/hifigan/config.json")))
    vocoder = hifigan.Generator(h)                   # get_vocoder's sequence (utils/model.py:56-68) minus the missing ckpt
    vocoder.eval()
    vocoder.remove_weight_norm()
    vocoder.to(dev)
    ds = Dataset("test.txt", cp, ct, cm)

    img = np.array(Image.open("sample/ipaexg_24pt_c3-whistle3-000-0271-517-repeat2.png").convert("L"))
    widths = [(img.shape[1] + i) // 4 for i in range(4)]
    padded = ds.character_padding_forinput(img, widths)
    pil = Image.fromarray(padded)
    text = "ピィピィ"

    def synth():
        batch = ([text], np.array([ds.audiotype_map["whistle3"]]), np.array([[ds.symbol_to_id[t] for t in text]]),
                 np.array([len(text)]), len(text), None, None, None, None, None, None, [pil], [None])
        b = to_device(batch, dev)
        out = model(*(b[1:]), ct["use_image"], d_control=args.d_control)
        wav = vocoder_infer(out[1].detach().transpose(1, 2), vocoder, cm, cp, Normalize=False)
        return out[1].shape[1], wav[0].shape[0]

    frames, samples = synth()                        # warm-up
    times = []
    for _ in range(args.reps):
        t0 = time.perf_counter()
        synth()
        times.append(time.perf_counter() - t0)
    audio_s = samples / 22050.0
    print(json.dumps({"workload": "configs[0]: sample PNG + class id -> vTTS -> HiFi-GAN V1, batch 1, random init, CPU", "threads": args.threads,
                      "cores": os.cpu_count(), "mel_frames": int(frames), "samples": int(samples), "audio_s": audio_s,
                      "wall_s_best": min(times), "wall_s_median": sorted(times)[len(times) // 2], "audio_s_per_s": audio_s / min(times),
                      "rtf": min(times) / audio_s, "torch": torch.__version__}))


if __name__ == "__main__":
    main()
