"""The oracle against the fixtures produced by the reference itself (CPU, no GPU)."""
import os

import numpy as np
import pytest
import torch

from oracle import hifigan_oracle as ho
from oracle import length_regulator_oracle as lro
from oracle.emulate import generator_forward_emulated, snr_db
from visual_onoma_to_wave_b200.synthetic import seeded_state_dict, FLOP_PER_FRAME

from helpers import sd_sha256, lr_cases, check_lr_output


@pytest.fixture(scope="module")
def sd1234():
    return seeded_state_dict(1234)


@pytest.mark.parametrize("case", ["a", "b", "c"])
def test_generator_oracle_matches_reference_golden(golden_dir, sd1234, case):
    g = np.load(os.path.join(golden_dir, f"generator_{case}.npz"))
    assert sd_sha256(sd1234) == str(g["weights_sha256"])
    wav = ho.generator_forward(sd1234, torch.from_numpy(g["mel"]))
    assert wav.shape == g["wav"].shape
    # same library calls in the same order; allow for thread-count dependent summation order
    np.testing.assert_allclose(wav.numpy(), g["wav"], rtol=0, atol=2e-6)


def test_generator_oracle_fp64_arbiter(golden_dir, sd1234):
    g = np.load(os.path.join(golden_dir, "generator_a.npz"))
    wav64 = ho.generator_forward(sd1234, torch.from_numpy(g["mel"]).double())
    assert np.abs(wav64.numpy() - g["wav"]).max() < 1e-6


def test_weight_norm_fold_and_vocoder_infer_golden(golden_dir):
    g = np.load(os.path.join(golden_dir, "vocoder_infer_wn.npz"))
    sd = seeded_state_dict(4321, weight_norm=True)
    for k in sd:
        if k.endswith("weight_g"):
            sd[k] = sd[k] * (1.0 + 0.25 * torch.sin(torch.arange(sd[k].numel()).float()).reshape(sd[k].shape))
    assert sd_sha256(sd) == str(g["weights_sha256"])
    folded = ho.fold_weight_norm(sd)
    assert len(sd) == 234 and len(folded) == 156
    assert sd_sha256(folded) == str(g["folded_sha256"])          # bit-equal fold
    np.testing.assert_array_equal(folded["conv_pre.weight"].numpy()[:4], g["conv_pre_w"])
    np.testing.assert_array_equal(folded["ups.0.weight"].numpy()[:2], g["ups0_w"])
    wav = ho.generator_forward(sd, torch.from_numpy(g["mel"])).squeeze(1).numpy()
    lengths = g["lengths"]
    np.testing.assert_allclose(wav[0][: lengths[0]], g["wav0"], atol=2e-6, rtol=0)
    np.testing.assert_allclose(wav[1][: lengths[1]], g["wav1"], atol=2e-6, rtol=0)


def test_flop_count():
    assert ho.count_flops(1) == FLOP_PER_FRAME == 614_105_088
    assert ho.count_flops(172) == 105_626_075_136


def test_numpy_conv_restatement_matches_torch():
    """Pins what F.conv1d / F.conv_transpose1d mean, independently of torch (small case)."""
    from oracle.np_conv import conv1d_np, conv_transpose1d_np, conv_transpose1d_polyphase_np
    rng = np.random.default_rng(0)
    x = rng.standard_normal((2, 6, 37))
    w = rng.standard_normal((5, 6, 7))
    b = rng.standard_normal(5)
    for d in (1, 3, 5):
        ref = torch.nn.functional.conv1d(torch.from_numpy(x), torch.from_numpy(w), torch.from_numpy(b),
                                         padding=ho.get_padding(7, d), dilation=d).numpy()
        np.testing.assert_allclose(conv1d_np(x, w, b, ho.get_padding(7, d), d), ref, atol=1e-12)
    for (k, s) in ((16, 8), (4, 2)):
        wt = rng.standard_normal((6, 5, k))
        ref = torch.nn.functional.conv_transpose1d(torch.from_numpy(x), torch.from_numpy(wt), torch.from_numpy(b),
                                                   stride=s, padding=(k - s) // 2).numpy()
        np.testing.assert_allclose(conv_transpose1d_np(x, wt, b, s, (k - s) // 2), ref, atol=1e-12)
        np.testing.assert_allclose(conv_transpose1d_polyphase_np(x, wt, b, s), ref, atol=1e-12)


def test_length_regulator_oracle_matches_reference_golden(golden_dir):
    g = np.load(os.path.join(golden_dir, "length_regulator.npz"))
    n = 0
    for name, x, dur, ml in lr_cases(g):
        if x.dtype == torch.bfloat16:
            xn = x.view(torch.int16).numpy()          # bitwise copy semantics
            out, mel_len = lro.length_regulate(xn, dur, ml)
            out_t = torch.from_numpy(out).view(torch.bfloat16)
        else:
            out, mel_len = lro.length_regulate(x.numpy(), dur, ml)
            out_t = torch.from_numpy(out)
        check_lr_output(g, name, out_t, mel_len)
        if f"{name}.mask" in g.files:
            np.testing.assert_array_equal(lro.get_mask_from_lengths(mel_len), g[f"{name}.mask"])
        n += 1
    assert n == 11
    np.testing.assert_array_equal(lro.round_durations(g["round.logd"], 1.5), g["round.dur"])


def test_emulated_bf16_pipeline_is_inside_tolerance(golden_dir, sd1234):
    """The numerics plan (bf16 operands, fp32 accumulate + fp32 residual) clears 40 dB."""
    g = np.load(os.path.join(golden_dir, "generator_a.npz"))
    mel = torch.from_numpy(g["mel"])
    emu = generator_forward_emulated(sd1234, mel, mode="bf16", residual="fp32")
    assert snr_db(torch.from_numpy(g["wav"]), emu) > 40.0


# ------------------------------------------------------------------------------------------
# PostNet + mel_linear: the oracle against outputs of the reference classes
# ------------------------------------------------------------------------------------------
def test_postnet_oracle_matches_reference_golden(golden_dir):
    import hashlib
    from oracle import postnet_oracle as po
    from visual_onoma_to_wave_b200.synthetic import seeded_postnet_state_dict
    g = np.load(os.path.join(golden_dir, "postnet.npz"))
    sd = seeded_postnet_state_dict(4321)
    h = hashlib.sha256()
    for k in sorted(sd):
        h.update(k.encode())
        h.update(sd[k].contiguous().numpy().tobytes())
    assert h.hexdigest().encode() == bytes(g["sd_sha256"]), "seeded PostNet weights drifted from the fixture"
    for name in "abc":
        y = po.postnet_forward(sd, torch.from_numpy(g[f"x_{name}"]))
        assert float((y - torch.from_numpy(g[f"y_{name}"])).abs().max()) <= 1e-5
    out, pm = po.vtts_tail(sd, torch.from_numpy(g["lin_w"]), torch.from_numpy(g["lin_b"]), torch.from_numpy(g["dec"]))
    assert float((out - torch.from_numpy(g["mel"])).abs().max()) <= 1e-5
    assert float((pm - torch.from_numpy(g["postnet_mel"])).abs().max()) <= 1e-5


def test_postnet_dropin_has_the_reference_state_dict_keys():
    import visual_onoma_to_wave_b200 as v
    from visual_onoma_to_wave_b200.synthetic import seeded_postnet_state_dict
    pn = v.PostNet()
    sd = seeded_postnet_state_dict(4321)
    assert set(pn.state_dict().keys()) == set(sd.keys())
    pn.load_state_dict(sd)                      # strict
    pn.eval()
    with pytest.raises(RuntimeError):           # no CPU fallback
        pn(torch.zeros(1, 4, 80))
    pn.train()
    with pytest.raises(RuntimeError):           # inference only
        pn(torch.zeros(1, 4, 80))
    blob = pn.folded_blob()
    assert blob.numel() == 5 * (80 * 512 + 3 * 512 * 512 + 512 * 80) + 4 * 512 + 80


# ------------------------------------------------------------------------------------------
# mel Decoder FFT-block stack: the oracle against outputs of the reference class
# ------------------------------------------------------------------------------------------
DECODER_CASES = {"full": (dict(), 2468), "small": (dict(n_layers=2, d_inner=256, kernels=(3, 1)), 97),
                 "long": (dict(n_layers=1, d_inner=128, kernels=(9, 1), max_seq_len=32), 98)}


def _sha(sd):
    import hashlib
    h = hashlib.sha256()
    for k in sorted(sd):
        h.update(k.encode())
        h.update(sd[k].contiguous().numpy().tobytes())
    return h.hexdigest().encode()


@pytest.mark.parametrize("name", list(DECODER_CASES))
def test_decoder_oracle_matches_reference_golden(golden_dir, name):
    from oracle import decoder_oracle as do
    from visual_onoma_to_wave_b200.synthetic import seeded_decoder_state_dict
    g = np.load(os.path.join(golden_dir, "decoder.npz"))
    kw, seed = DECODER_CASES[name]
    sd = seeded_decoder_state_dict(seed, **kw)
    assert _sha(sd) == bytes(g[f"sha_{name}"]), "seeded Decoder weights drifted from the fixture"
    mask = torch.from_numpy(g[f"mask_{name}"])
    y, m2 = do.decoder_forward(sd, torch.from_numpy(g[f"x_{name}"]), mask)
    assert torch.equal(m2, mask)
    ref = torch.from_numpy(g[f"y_{name}"])
    assert float((y - ref).abs().max()) <= 2e-5
    assert float(y[mask].abs().max()) == 0.0          # padding frames are zeroed after every sub-layer


def test_sinusoid_table_matches_the_reference_formula():
    from oracle.decoder_oracle import sinusoid_table as slow
    from visual_onoma_to_wave_b200.synthetic import sinusoid_table as fast
    assert torch.equal(slow(67, 256), fast(67, 256))


def test_decoder_dropin_has_the_reference_state_dict_keys():
    import visual_onoma_to_wave_b200 as v
    from visual_onoma_to_wave_b200.synthetic import seeded_decoder_state_dict, decoder_config
    kw = dict(n_layers=2, d_inner=256, kernels=(3, 1))
    dec = v.Decoder(decoder_config(**kw))
    sd = seeded_decoder_state_dict(97, **kw)
    assert set(dec.state_dict().keys()) == set(sd.keys())
    dec.load_state_dict(sd)                     # strict
    dec.eval()
    mask = torch.zeros(1, 4, dtype=torch.bool)
    with pytest.raises(RuntimeError):           # no CPU fallback
        dec(torch.zeros(1, 4, 256), mask)
    dec.train()
    with pytest.raises(RuntimeError):           # inference only
        dec(torch.zeros(1, 4, 256), mask)
    D, H = 256, 256
    assert dec.weight_blob().numel() == 2 * (4 * D * D + 4 * D + 4 * D + H * D * 3 + H + D * H + D)


# ------------------------------------------------------------------------------------------
# VarianceAdaptor: the oracle against outputs of the reference classes
# ------------------------------------------------------------------------------------------
VARIANCE_CASES = {"energy": dict(is_energy=True, is_kurtosis=False), "both": dict(is_energy=True, is_kurtosis=True)}
VARIANCE_KEYS = ["out", "e_pred", "k_pred", "log_d", "d_round", "mel_len", "mel_mask"]


def _check_variance_tuple(got, g, name, prefix, tol):
    for k, val in zip(VARIANCE_KEYS, got):
        key = f"{prefix}{k}_{name}"
        if key not in g.files:
            assert val is None or k == "mel_mask", key      # mel_mask=None passes through on the target path
            continue
        ref = torch.from_numpy(g[key])
        val = val.cpu()
        assert val.shape == ref.shape, key
        if ref.dtype in (torch.int64, torch.bool) or k == "d_round":
            assert torch.equal(val.to(ref.dtype), ref), key
        else:
            assert float((val - ref).abs().max()) <= tol, key


@pytest.mark.parametrize("name", list(VARIANCE_CASES))
def test_variance_oracle_matches_reference_golden(golden_dir, name):
    from oracle import variance_oracle as vo
    from visual_onoma_to_wave_b200.synthetic import seeded_variance_adaptor_state_dict, VARIANCE_STATS
    g = np.load(os.path.join(golden_dir, "variance.npz"))
    kw = VARIANCE_CASES[name]
    sd = seeded_variance_adaptor_state_dict(int(g[f"seed_{name}"]), **kw)
    assert _sha(sd) == bytes(g[f"sha_{name}"]), "seeded VarianceAdaptor weights drifted from the fixture"
    x, mask = torch.from_numpy(g[f"x_{name}"]), torch.from_numpy(g[f"mask_{name}"])
    got = vo.variance_adaptor_forward(sd, VARIANCE_STATS, x, mask, e_control=1.2, d_control=1.5, **kw)
    _check_variance_tuple(got, g, name, "", 2e-5)
    got = vo.variance_adaptor_forward(sd, VARIANCE_STATS, x, mask, max_len=100, energy_target=torch.from_numpy(g[f"et_{name}"]),
                                      duration_target=torch.from_numpy(g[f"dt_{name}"]), **kw)
    _check_variance_tuple(got, g, name, "t_", 2e-5)


def test_variance_adaptor_dropin_has_the_reference_state_dict_keys(tmp_path):
    import json
    import visual_onoma_to_wave_b200 as v
    from visual_onoma_to_wave_b200.synthetic import (seeded_variance_adaptor_state_dict, variance_model_config,
                                                     VARIANCE_STATS)
    (tmp_path / "stats.json").write_text(json.dumps(VARIANCE_STATS))
    kw = dict(is_energy=True, is_kurtosis=True)
    va = v.VarianceAdaptor({"path": {"preprocessed": str(tmp_path)}}, variance_model_config(**kw))
    sd = seeded_variance_adaptor_state_dict(1, **kw)
    assert set(va.state_dict().keys()) == set(sd.keys())
    va.load_state_dict(sd)                      # strict
    assert torch.equal(va.energy_bins, sd["energy_bins"])      # the constructor's own edges are the reference's
    va.eval()
    with pytest.raises(RuntimeError):           # no CPU fallback
        va(torch.zeros(1, 4, 256), torch.zeros(1, 4, dtype=torch.bool))


def test_acoustic_tail_dropin_keys_and_errors(tmp_path):
    """vtts.py:36-52: the four sub-modules behind the encoder under the reference's attribute names."""
    import json
    import visual_onoma_to_wave_b200 as v
    from visual_onoma_to_wave_b200.synthetic import (seeded_variance_adaptor_state_dict, variance_model_config,
                                                     seeded_decoder_state_dict, decoder_config, seeded_postnet_state_dict,
                                                     VARIANCE_STATS)
    (tmp_path / "stats.json").write_text(json.dumps(VARIANCE_STATS))
    dkw = dict(n_layers=1, d_inner=64, kernels=(3, 1))
    model_config = {**variance_model_config(), **decoder_config(**dkw)}
    model_config["transformer"]["encoder_hidden"] = 256
    pre_cfg = {"path": {"preprocessed": str(tmp_path)}, "preprocessing": {"mel": {"n_mel_channels": 80}}}
    tail = v.AcousticTail(pre_cfg, model_config)
    want = {"variance_adaptor." + k for k in seeded_variance_adaptor_state_dict(1)} | \
           {"decoder." + k for k in seeded_decoder_state_dict(1, **dkw)} | \
           {"postnet." + k for k in seeded_postnet_state_dict(1)} | {"mel_linear.weight", "mel_linear.bias"}
    assert set(tail.state_dict().keys()) == want
    tail.eval()
    with pytest.raises(RuntimeError):           # no CPU fallback anywhere in the chain
        tail(torch.zeros(1, 4, 256), torch.zeros(1, 4, dtype=torch.bool))
    tail.train()
    with pytest.raises(RuntimeError):           # inference only
        tail(torch.zeros(1, 4, 256), torch.zeros(1, 4, dtype=torch.bool))
