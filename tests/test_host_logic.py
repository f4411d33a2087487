"""Host-side mirrors of the reference surface (CPU): state-dict layout, folding, sharding."""
import json
import os

import numpy as np
import pytest
import torch

import visual_onoma_to_wave_b200 as v
from visual_onoma_to_wave_b200.sharding import shard_bounds, shard_batch
from visual_onoma_to_wave_b200.synthetic import V1_CONFIG, seeded_state_dict, layer_shapes
from oracle import hifigan_oracle as ho

from helpers import sd_sha256


def make_gen():
    return v.Generator(v.AttrDict(V1_CONFIG))


def test_state_dict_layout_matches_reference_names():
    g = make_gen()
    sd = g.state_dict()
    assert len(sd) == 234                                   # {bias, weight_g, weight_v} x 78
    assert set(sd) == set(seeded_state_dict(0, weight_norm=True))
    assert sd["ups.0.weight_g"].shape == (512, 1, 1)        # norm over Cin for ConvTranspose1d
    assert sd["ups.0.weight_v"].shape == (512, 256, 16)
    g.eval()
    g.remove_weight_norm()
    sd = g.state_dict()
    assert len(sd) == 156 and set(sd) == set(seeded_state_dict(0))
    assert sum(p.numel() for p in g.parameters()) == 13_926_017
    with pytest.raises(ValueError):
        g.remove_weight_norm()                              # torch raises the second time too


def test_fold_is_bit_identical_to_reference_fold(golden_dir):
    gold = np.load(os.path.join(golden_dir, "vocoder_infer_wn.npz"))
    sd = seeded_state_dict(4321, weight_norm=True)
    for k in sd:
        if k.endswith("weight_g"):
            sd[k] = sd[k] * (1.0 + 0.25 * torch.sin(torch.arange(sd[k].numel()).float()).reshape(sd[k].shape))
    g = make_gen()
    g.load_state_dict(sd)          # get_vocoder's order: load, eval, remove_weight_norm (utils/model.py:65-67)
    g.eval()
    g.remove_weight_norm()
    folded = {k: t.detach() for k, t in g.state_dict().items()}
    assert sd_sha256(folded) == str(gold["folded_sha256"])
    # the blob handed to the C ABI: weight then bias for every layer in registration order
    blob = g.folded_blob()
    assert blob.numel() == 13_926_017
    off = 0
    ref = ho.fold_weight_norm(sd)
    for name, (shape, _) in layer_shapes().items():
        n = int(np.prod(shape))
        assert torch.equal(blob[off:off + n].reshape(shape), ref[name + ".weight"]), name
        off += n
        nb = ref[name + ".bias"].numel()
        assert torch.equal(blob[off:off + nb], ref[name + ".bias"]), name
        off += nb
    assert off == blob.numel()


def test_get_vocoder_sequence_accepts_both_checkpoint_layouts(tmp_path):
    cfg = tmp_path / "config.json"
    cfg.write_text(json.dumps(V1_CONFIG))
    for wn in (True, False):
        sd = seeded_state_dict(5, weight_norm=wn)
        ck = tmp_path / f"g{int(wn)}.pth.tar"
        torch.save({"generator": sd}, ck)
        voc = v.get_vocoder({"vocoder": {"model": "HiFi-GAN", "speaker": "universal"}}, "cpu",
                            hifigan_config_path=str(cfg), ckpt_path=str(ck))
        assert len(voc.state_dict()) == 156 and not voc.training
    with pytest.raises(NotImplementedError):
        v.get_vocoder({"vocoder": {"model": "MelGAN", "speaker": "universal"}}, "cpu")


def test_vocoder_infer_normalize_keeps_reference_keyerror():
    # scripts/utils/model.py:86 reads preprocess_config["preprocessing"]["audio"]; the ICASSP yaml has top-level "audio"
    with pytest.raises(KeyError):
        v.vocoder_infer(torch.zeros(1, 80, 2), make_gen(), {"vocoder": {"model": "HiFi-GAN"}},
                        {"audio": {"max_wav_value": 32768.0}}, Normalize=True)


def test_shard_bounds_partition():
    for n in (0, 1, 7, 16, 1024, 1025):
        for ws in (1, 2, 3, 4, 8):
            spans = [shard_bounds(n, ws, r) for r in range(ws)]
            assert spans[0][0] == 0 and spans[-1][1] == n
            assert all(a[1] == b[0] for a, b in zip(spans, spans[1:]))
            sizes = [hi - lo for lo, hi in spans]
            assert max(sizes) - min(sizes) <= 1
    x = torch.arange(10)
    assert torch.equal(torch.cat([shard_batch(x, 4, r) for r in range(4)]), x)
    with pytest.raises(ValueError):
        shard_bounds(4, 2, 2)


def test_round_durations_matches_golden(golden_dir):
    g = np.load(os.path.join(golden_dir, "length_regulator.npz"))
    got = v.round_durations(torch.from_numpy(g["round.logd"]), 1.5)
    np.testing.assert_array_equal(got.numpy(), g["round.dur"])


def test_pipeline_chunks_cover_the_batch_in_order():
    """hifigan.pipeline_chunks: the sub-batches of one host call tile [0, B) in order, none larger than the sub-batch, and
    for a long job the first and the last ones are the small ones (their copies are what the pipeline cannot hide)."""
    from visual_onoma_to_wave_b200.hifigan import pipeline_chunks
    for B in list(range(1, 70)) + [127, 128, 129, 255, 256, 300, 1000, 1024, 4096]:
        for sub in (1, 2, 3, 16, 63, 64, 128):
            ch = pipeline_chunks(B, sub)
            assert ch[0][0] == 0 and ch[-1][1] == B
            assert all(a[1] == b[0] for a, b in zip(ch, ch[1:]))
            assert all(0 < e - s <= sub for s, e in ch)
    sizes = [e - s for s, e in pipeline_chunks(1024, 128)]
    assert sizes[0] == 32 and sizes[-1] <= 32 and sizes.count(128) == 7 and sum(sizes) == 1024
    assert [e - s for s, e in pipeline_chunks(7, 3)] == [3, 3, 1]          # small sub-batches: plain walk
