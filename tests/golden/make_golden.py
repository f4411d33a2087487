"""Generate tests/golden/*.npz by running the REFERENCE's own modules (build container only).

    python tests/golden/make_golden.py          # needs /root/reference (read-only)

The reference has no tests or golden vectors for this path, so these
fixtures are what pins the oracle: inputs are seeded here, outputs come from the imported
reference classes (hifigan.Generator, model.modules.LengthRegulator, utils.tools.*,
utils.model.vocoder_infer).  Weights are NOT stored (55 MB): they are regenerated from a
seed by visual_onoma_to_wave_b200.synthetic.seeded_state_dict and checked by sha256.
"""
import hashlib
import json
import os
import sys
from unittest.mock import MagicMock

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
REF = "/root/reference"
sys.path.insert(0, ROOT)

from visual_onoma_to_wave_b200.synthetic import seeded_state_dict, synthetic_mel  # noqa: E402


def import_reference():
    """cwd-relative paths + stub the missing plotting/audio deps."""
    for m in ("matplotlib", "matplotlib.pyplot", "librosa", "librosa.util", "librosa.filters"):
        sys.modules.setdefault(m, MagicMock())
    os.chdir(REF)
    sys.path.insert(0, os.path.join(REF, "scripts"))
    import hifigan                                   # noqa
    from model.modules import LengthRegulator        # noqa
    from utils.tools import get_mask_from_lengths, pad  # noqa
    from utils.model import vocoder_infer            # noqa
    return hifigan, LengthRegulator, get_mask_from_lengths, vocoder_infer


def sd_sha256(sd):
    h = hashlib.sha256()
    for k in sorted(sd):
        h.update(k.encode())
        h.update(sd[k].contiguous().numpy().tobytes())
    return h.hexdigest()


def main():
    torch.set_num_threads(8)
    hifigan, LengthRegulator, get_mask_from_lengths, vocoder_infer = import_reference()
    h = hifigan.AttrDict(json.load(open(os.path.join(REF, "scripts/hifigan/config.json"))))

    # ---- Generator, folded 156-key checkpoint -------------------------------------------
    seed = 1234
    sd = seeded_state_dict(seed)
    gen = hifigan.Generator(h)
    gen.eval()
    gen.remove_weight_norm()
    gen.load_state_dict(sd)
    for name, (B, T, kind) in {"a": (2, 12, "logmel"), "b": (1, 33, "randn"),
                               "c": (3, 1, "logmel")}.items():
        mel = synthetic_mel(B, T, seed=7 + B, kind=kind)
        with torch.no_grad():
            wav = gen(mel)
        np.savez_compressed(os.path.join(HERE, f"generator_{name}.npz"), seed=seed,
                            weights_sha256=sd_sha256(sd), mel=mel.numpy(), wav=wav.numpy())
        print("generator", name, tuple(wav.shape), float(wav.mean()), float(wav.std()))

    # ---- Generator, 234-key weight_norm checkpoint through get_vocoder's sequence ---------
    sd_wn = seeded_state_dict(4321, weight_norm=True)
    # make g differ from ||v|| so the fold is actually exercised
    for k in sd_wn:
        if k.endswith("weight_g"):
            sd_wn[k] = sd_wn[k] * (1.0 + 0.25 * torch.sin(torch.arange(sd_wn[k].numel()).float()
                                                            ).reshape(sd_wn[k].shape))
    gen2 = hifigan.Generator(h)
    gen2.load_state_dict(sd_wn)      # utils/model.py:65
    gen2.eval()
    gen2.remove_weight_norm()        # utils/model.py:67
    mel = synthetic_mel(2, 9, seed=11)
    lengths = [9 * 256, 5 * 256 + 17]
    wavs = vocoder_infer(mel, gen2, {"vocoder": {"model": "HiFi-GAN"}}, {}, lengths=lengths,
                         Normalize=False)      # utils/model.py:73-98
    folded = {k: v.detach().clone() for k, v in gen2.state_dict().items()}
    np.savez_compressed(os.path.join(HERE, "vocoder_infer_wn.npz"), seed=4321,
                        weights_sha256=sd_sha256(sd_wn), folded_sha256=sd_sha256(folded),
                        conv_pre_w=folded["conv_pre.weight"].numpy()[:4],
                        ups0_w=folded["ups.0.weight"].numpy()[:2],
                        mel=mel.numpy(), lengths=np.array(lengths),
                        wav0=wavs[0], wav1=wavs[1])
    print("vocoder_infer", [w.shape for w in wavs])

    # ---- LengthRegulator edge cases (modules.py:132-159) -------------------------------
    lr = LengthRegulator()
    g = torch.Generator().manual_seed(0)
    cases = {}
    x = torch.randn(4, 6, 8, generator=g)
    cases["int_basic"] = (x, torch.randint(0, 5, (4, 6), generator=g), None)
    cases["int_maxlen_pad"] = (x, torch.randint(0, 5, (4, 6), generator=g), 40)
    cases["int_maxlen_crop"] = (x, torch.randint(1, 6, (4, 6), generator=g), 7)
    cases["maxlen_zero_falsy"] = (x, torch.randint(0, 5, (4, 6), generator=g), 0)
    cases["float_trunc"] = (x, torch.tensor([[0.0, 0.9, 1.0, 1.99, 2.5, 3.0001],
                                             [-0.0, -0.5, -1.5, -3.0, 4.7, 1e-3],
                                             [2.0, 2.0, 2.0, 2.0, 2.0, 2.0],
                                             [0.5, 0.5, 0.5, 0.5, 0.5, 7.999]]), None)
    cases["all_zero"] = (x, torch.zeros(4, 6), None)
    cases["one_row_empty"] = (x, torch.tensor([[1, 2, 0, 0, 3, 1], [0, 0, 0, 0, 0, 0],
                                               [5, 0, 0, 0, 0, 0], [0, 0, 0, 0, 0, 9]]), None)
    cases["negative_int"] = (x, torch.tensor([[-1, 2, -3, 0, 3, 1]] * 4), None)
    xb = torch.randn(64, 32, 256, generator=torch.Generator().manual_seed(3))   # regenerated in tests
    cases["cfg3_int"] = (xb, torch.randint(0, 25, (64, 32), generator=g), None)
    logd = torch.randn(64, 32, generator=g)
    dur_f = torch.clamp(torch.round(torch.exp(logd) - 1) * 1.5, min=0)   # modules.py:110-113
    cases["cfg3_float_maxlen1000"] = (xb, dur_f, 1000)
    cases["bf16_passthrough"] = (x.to(torch.bfloat16), torch.randint(0, 5, (4, 6), generator=g), None)
    out = {}
    for name, (xx, dd, ml) in cases.items():
        y, mel_len = lr(xx, dd, ml)
        assert mel_len.dtype == torch.int64
        big = xx.numel() > 100000      # config-3 sized cases: store digests, not 60 MB of frames
        if not big:
            out[f"{name}.x"] = xx.float().numpy()
        out[f"{name}.x_is_bf16"] = np.array(xx.dtype == torch.bfloat16)
        out[f"{name}.dur"] = dd.numpy()
        out[f"{name}.max_len"] = np.array(-1 if ml is None else ml)
        if big:
            out[f"{name}.out_sha256"] = np.array(hashlib.sha256(y.contiguous().numpy().tobytes()).hexdigest())
            out[f"{name}.out_shape"] = np.array(y.shape)
        else:
            out[f"{name}.out"] = y.float().numpy()
        out[f"{name}.out_is_bf16"] = np.array(y.dtype == torch.bfloat16)
        out[f"{name}.mel_len"] = mel_len.numpy()
        if int(mel_len.max()) > 0:
            out[f"{name}.mask"] = get_mask_from_lengths(mel_len).numpy()   # tools.py:164-171
        print("LR", name, tuple(y.shape), mel_len.tolist()[:6])
    out["round.logd"] = logd.numpy()
    out["round.dur"] = dur_f.numpy()
    np.savez_compressed(os.path.join(HERE, "length_regulator.npz"), **out)


def postnet_golden():
    """PostNet + mel_linear tail, outputs of the reference's own classes."""
    from visual_onoma_to_wave_b200.synthetic import seeded_postnet_state_dict
    from transformer.Layers import PostNet   # reference
    sd = seeded_postnet_state_dict(4321)
    pn = PostNet()
    pn.load_state_dict(sd)                    # strict: the key set is the reference's
    pn.eval()
    g = torch.Generator().manual_seed(7)
    out = {"sd_sha256": np.frombuffer(sd_sha256(sd).encode(), dtype=np.uint8)}
    with torch.no_grad():
        for name, (B, L) in {"a": (2, 37), "b": (1, 172), "c": (3, 5)}.items():
            x = torch.randn(B, L, 80, generator=g) * 2.0 - 3.0          # log-mel-like range
            out[f"x_{name}"] = x.numpy()
            out[f"y_{name}"] = pn(x).numpy()
        lin = torch.nn.Linear(256, 80)
        lin.weight.data = torch.randn(80, 256, generator=g) / 16.0
        lin.bias.data = torch.randn(80, generator=g) * 0.1
        dec = torch.randn(2, 23, 256, generator=g)
        o = lin(dec)
        out["lin_w"], out["lin_b"], out["dec"] = lin.weight.data.numpy(), lin.bias.data.numpy(), dec.numpy()
        out["mel"] = o.numpy()
        out["postnet_mel"] = (pn(o) + o).numpy()                       # vtts.py:106-107
    np.savez_compressed(os.path.join(HERE, "postnet.npz"), **out)
    print("postnet.npz written")


def decoder_golden():
    """mel Decoder FFT-block stack, outputs of the reference's own class."""
    from visual_onoma_to_wave_b200.synthetic import seeded_decoder_state_dict, decoder_config
    for m in ("utils.symbols", "model.visual_feature_extractor"):    # Encoder-only imports of Models.py (not on this path)
        sys.modules.setdefault(m, MagicMock())
    from transformer.Models import Decoder    # reference
    out = {}
    g = torch.Generator().manual_seed(11)
    cases = {  # name: (config kwargs, seed, B, L, lengths)
        "full": (dict(), 2468, 2, 61, [61, 40]),                                   # the shipped model.yaml shape
        "small": (dict(n_layers=2, d_inner=256, kernels=(3, 1)), 97, 3, 70, [70, 1, 33]),
        "long": (dict(n_layers=1, d_inner=128, kernels=(9, 1), max_seq_len=32), 98, 2, 50, [50, 45]),   # L > max_seq_len
    }
    with torch.no_grad():
        for name, (kw, seed, B, L, lens) in cases.items():
            sd = seeded_decoder_state_dict(seed, **kw)
            dec = Decoder(decoder_config(**kw))
            dec.load_state_dict(sd)             # strict: the key set is the reference's
            dec.eval()
            x = torch.randn(B, L, 256, generator=g)
            mask = torch.arange(L)[None, :] >= torch.tensor(lens)[:, None]
            y, m2 = dec(x, mask)
            assert torch.equal(m2, mask)
            out[f"sha_{name}"] = np.frombuffer(sd_sha256(sd).encode(), dtype=np.uint8)
            out[f"x_{name}"], out[f"mask_{name}"], out[f"y_{name}"] = x.numpy(), mask.numpy(), y.numpy()
    np.savez_compressed(os.path.join(HERE, "decoder.npz"), **out)
    print("decoder.npz written")


def variance_golden():
    """VarianceAdaptor in front of LengthRegulator, outputs of the reference's own classes.
    Seeds are walked until every predicted value sits clear of a rounding / bin edge, so that a 1e-6 difference in a
    predictor output cannot flip an integer on the other side."""
    import tempfile
    from visual_onoma_to_wave_b200.synthetic import (seeded_variance_adaptor_state_dict, variance_model_config,
                                                     VARIANCE_STATS)
    import model.modules as mm               # reference
    mm.device = torch.device("cpu")          # module-global device (modules.py:13)
    import utils.tools as ut
    ut.device = torch.device("cpu")
    tmp = tempfile.mkdtemp()
    with open(os.path.join(tmp, "stats.json"), "w") as f:
        json.dump(VARIANCE_STATS, f)
    pre_cfg = {"path": {"preprocessed": tmp}}
    out = {}
    cases = {"energy": dict(is_energy=True, is_kurtosis=False), "both": dict(is_energy=True, is_kurtosis=True)}
    for name, kw in cases.items():
        lens = [29, 17, 5]
        B, T = len(lens), max(lens)
        src_mask = torch.arange(T)[None, :] >= torch.tensor(lens)[:, None]
        for seed in range(1357, 1457):
            sd = seeded_variance_adaptor_state_dict(seed, **kw)
            va = mm.VarianceAdaptor(pre_cfg, variance_model_config(**kw))
            va.load_state_dict(sd)           # strict: the key set is the reference's
            va.eval()
            x = torch.randn(B, T, 256, generator=torch.Generator().manual_seed(seed + 1))
            with torch.no_grad():
                r = va(x, src_mask, e_control=1.2, d_control=1.5)
                frac = torch.exp(r[3]) - 1
                ok = float(((frac - torch.floor(frac)) - 0.5).abs().min()) > 2e-3
                ok = ok and float((r[1][..., None] - sd["energy_bins"]).abs().min()) > 2e-4
                if kw["is_kurtosis"]:
                    ok = ok and float((r[2][..., None] - sd["kurt_bins"]).abs().min()) > 2e-4
            if ok:
                break
        assert ok, "no seed with safe margins"
        with torch.no_grad():
            et = torch.rand(B, T, generator=torch.Generator().manual_seed(3)) * 8 - 1
            dt = torch.randint(0, 7, (B, T), generator=torch.Generator().manual_seed(4))
            rt = va(x, src_mask, mel_mask=None, max_len=100, energy_target=et, duration_target=dt)
        out[f"seed_{name}"] = np.array(seed)
        out[f"sha_{name}"] = np.frombuffer(sd_sha256(sd).encode(), dtype=np.uint8)
        out[f"x_{name}"], out[f"mask_{name}"] = x.numpy(), src_mask.numpy()
        keys = ["out", "e_pred", "k_pred", "log_d", "d_round", "mel_len", "mel_mask"]
        for k, val in zip(keys, r):
            if val is not None:
                out[f"{k}_{name}"] = val.numpy()
        out[f"et_{name}"], out[f"dt_{name}"] = et.numpy(), dt.numpy()
        for k, val in zip(keys, rt):
            if val is not None:
                out[f"t_{k}_{name}"] = val.numpy()
    np.savez_compressed(os.path.join(HERE, "variance.npz"), **out)
    print("variance.npz written", {k: int(out[k]) for k in out if k.startswith("seed_")})


if __name__ == "__main__":
    main()
    postnet_golden()
    decoder_golden()
    variance_golden()
