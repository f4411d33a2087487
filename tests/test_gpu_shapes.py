"""Parity at the shapes BASELINE.json names and bench.py times ("Next round" item 1).

* configs[1] (B=16 x T=172) and the bench sub-batch (B=128 x T=172): the bf16 Generator is compared PER SAMPLE with the
  CPU restatement of its own dataflow (oracle/emulate.py::generator_forward_emulated_v3) and with the fp32 oracle, and per
  stage through the debug taps, so a wrong tile cannot hide under a total-power gate;
* configs[4] (B=8 x T=2584, 30 s): whole waveforms of two utterances against the fp32 oracle, both precision modes;
* the int16 host path on a re-used workspace (generator.cu:592), and two GPUs driven from one process.
"""
import numpy as np
import pytest
import torch

import visual_onoma_to_wave_b200 as v
from visual_onoma_to_wave_b200 import _lib
from visual_onoma_to_wave_b200.synthetic import V1_CONFIG, seeded_state_dict, synthetic_mel
from oracle import hifigan_oracle as ho
from oracle.emulate import generator_forward_emulated_v3, snr_db, ac_snr_db

pytestmark = pytest.mark.gpu

MAX_ABS_ERR_FP32 = 1e-4     # north_star: fp32/TF32 mode
MIN_SNR_DB_BF16 = 40.0      # north_star: bf16 mode, total-power SNR
# The synthetic log-mel inputs give a waveform with mean 0.0105 and std 0.0101: half of the total power is DC, so the
# same error shows 3 dB lower against the AC power alone (the CPU emulation of the bf16 dataflow measures 43.1 dB total,
# 40.0 dB AC).  The AC floor makes sure the DC term cannot carry the total-power gate.
MIN_AC_SNR_DB_BF16 = 37.0
SR = 22050


@pytest.fixture(scope="module")
def dev():
    assert torch.cuda.is_available(), "gpu tests need a CUDA device"
    assert _lib.lib().vo2w_device_check() == 0, "libvo2w needs an sm_100 device"
    return torch.device("cuda:0")


def make_gen(sd, precision, dev):
    g = v.Generator(v.AttrDict(V1_CONFIG), precision=precision)
    g.eval()
    g.remove_weight_norm()
    g.load_state_dict(sd)
    return g.to(dev)


@pytest.fixture(scope="module")
def gens(dev):
    sd = seeded_state_dict(1234)
    return {p: make_gen(sd, p, dev) for p in ("fp32", "bf16")}, sd


def _check_bf16_against_emulation(wav, emu, ref, what):
    """wav (GPU), emu (CPU restatement of the same dataflow), ref (fp32 oracle): [n,1,S].
    sigma = rms(emu - ref) is the rounding noise the bf16 dataflow itself has (1.0e-4 on these inputs).  The kernels differ
    from their restatement only by rounding decisions (packed-bf16 leaky-relu with slope bf16(0.1), bias as bf16 hi + lo,
    accumulation order), i.e. by noise of at most the same size; a wrong tile is an error of the signal's size, 100 sigma."""
    for b in range(wav.shape[0]):
        w, e, r = wav[b, 0].double(), emu[b, 0].double(), ref[b, 0].double()
        sigma = float((e - r).pow(2).mean().sqrt())
        d = w - e
        assert torch.isfinite(w).all(), what
        # measured on B200 (tools/gpu_stats.py, 16 utterances): rms 0.90 - 0.92 sigma, max 3.7 - 5.0 sigma, worst 128-sample
        # window 1.06 - 1.19 sigma
        rms, mx = float(d.pow(2).mean().sqrt()), float(d.abs().max())
        assert rms <= 1.5 * sigma, (what, b, rms, sigma)
        assert mx <= 10.0 * sigma, (what, b, mx, sigma)
        n = d.numel() // 128 * 128
        win = d[:n].reshape(-1, 128).pow(2).mean(1).sqrt()          # every 128-sample window
        assert float(win.max()) <= 3.0 * sigma, (what, b, float(win.max()), sigma)
        assert snr_db(r, w) >= MIN_SNR_DB_BF16, (what, b, snr_db(r, w))
        assert ac_snr_db(r, w) >= MIN_AC_SNR_DB_BF16, (what, b, ac_snr_db(r, w))


def _check_taps(gen, taps, n_utt, names):
    """Per-stage streams of the last forward against the emulation's: every (utterance, 128-sample window) of every tap
    within 3 % of the tap's rms (rounding flips measure 0.06 - 0.5 % on B200; a wrong tile is ~100 %), which localises a bad tile."""
    for name in names:
        got = gen.debug_tap(name)
        exp = taps[name][:n_utt]
        got = got.reshape(-1, *exp.shape[1:])[:n_utt]
        assert got.shape == exp.shape, (name, got.shape, exp.shape)
        rms = float(exp.double().pow(2).mean().sqrt())
        d = (got.double() - exp.double())
        L = d.shape[-1] // 128 * 128
        win = d[..., :L].reshape(d.shape[0], d.shape[1], -1, 128).pow(2).mean(dim=(1, 3)).sqrt()   # [utt, window]
        assert float(win.max()) <= 0.03 * rms, (name, float(win.max()), rms)
        if L < d.shape[-1]:
            tail = d[..., L:].pow(2).mean(dim=(1, 2)).sqrt()
            assert float(tail.max()) <= 0.03 * rms, (name, "tail", float(tail.max()), rms)


def test_bf16_generator_config2_per_sample_and_per_stage(dev, gens):
    """configs[1]: B=16 x T=172, bf16 mode -- every sample of every utterance against the restated dataflow, every stage
    through the taps (scripts/hifigan/models.py:149-165)."""
    gen, sd = gens
    mel = synthetic_mel(16, 172, seed=0)
    taps = {}
    emu = generator_forward_emulated_v3(sd, mel, taps=taps)
    ref = ho.generator_forward(sd, mel)
    wav = gen["bf16"](mel.to(dev)).cpu()
    assert wav.shape == ref.shape == (16, 1, 172 * 256)
    names = [f"ups.{i}" for i in range(4)] + [n for n in (f"mrf_act.{i}" for i in range(4)) if gen["bf16"].has_tap(n)]
    assert len(names) >= 7
    _check_taps(gen["bf16"], taps, 16, names)
    _check_bf16_against_emulation(wav, emu, ref, "B=16")


def test_bf16_generator_bench_sub_batch_per_sample(dev, gens):
    """The shape bench.py times: B=128 x T=172 per forward (about 11 000 tiles at C=32, 76 per persistent CTA).  16
    utterances spread over the batch -- first, last, and both sides of every 32-boundary -- are compared per sample."""
    gen, sd = gens
    mel = synthetic_mel(128, 172, seed=1)
    idx = [0, 1, 31, 32, 33, 63, 64, 65, 77, 95, 96, 97, 110, 125, 126, 127]
    sub = mel[idx].contiguous()
    emu = generator_forward_emulated_v3(sd, sub)
    ref = ho.generator_forward(sd, sub)
    wav = gen["bf16"](mel.to(dev))
    assert wav.shape == (128, 1, 172 * 256) and bool(torch.isfinite(wav).all())
    _check_bf16_against_emulation(wav[idx].cpu(), emu, ref, "B=128")
    # and the whole batch through a cheap invariant: the same utterances alone give the same bits (no cross-talk between
    # the tiles of different utterances that share a persistent CTA)
    alone = gen["bf16"](sub.to(dev))
    assert torch.equal(alone, wav[idx])


def test_fp32_generator_config2_and_sub_batch(dev, gens):
    """configs[1] in the exact mode: B=16 x T=172 against the fp32 oracle on every sample, and 8 utterances of a B=64
    forward (the sub-batch the fp32 bench figure runs)."""
    gen, sd = gens
    mel = synthetic_mel(16, 172, seed=0)
    ref = ho.generator_forward(sd, mel)
    wav = gen["fp32"](mel.to(dev)).cpu()
    assert float((wav - ref).abs().max()) <= MAX_ABS_ERR_FP32
    mel64 = synthetic_mel(64, 172, seed=2)
    idx = [0, 1, 31, 32, 33, 50, 62, 63]
    ref64 = ho.generator_forward(sd, mel64[idx].contiguous())
    wav64 = gen["fp32"](mel64.to(dev))[idx].cpu()
    assert float((wav64 - ref64).abs().max()) <= MAX_ABS_ERR_FP32


@pytest.mark.parametrize("precision", ["bf16", "fp32"])
def test_generator_config5_long_form_whole_waveforms(dev, gens, precision):
    """configs[4]: B=8 x T=2584 (30 s, 661 504 samples per utterance).  The WHOLE waveforms of the first and the last
    utterance against the fp32 oracle: max-abs-err in the exact mode; in bf16 mode the SNR per utterance and in every
    1-second window (total power and AC), so a bad stretch cannot be averaged away."""
    gen, sd = gens
    mel = synthetic_mel(8, 2584, seed=5)
    wav = gen[precision](mel.to(dev))
    assert wav.shape == (8, 1, 2584 * 256) and bool(torch.isfinite(wav).all())
    for b in (0, 7):
        ref = ho.generator_forward(sd, mel[b:b + 1])[0, 0]
        got = wav[b, 0].cpu()
        if precision == "fp32":
            assert float((got - ref).abs().max()) <= MAX_ABS_ERR_FP32, b
            continue
        assert snr_db(ref, got) >= MIN_SNR_DB_BF16, (b, snr_db(ref, got))
        assert ac_snr_db(ref, got) >= MIN_AC_SNR_DB_BF16, (b, ac_snr_db(ref, got))
        for s in range(0, ref.numel() - SR + 1, SR):
            r, g = ref[s:s + SR], got[s:s + SR]
            assert snr_db(r, g) >= MIN_SNR_DB_BF16, (b, s, snr_db(r, g))
            assert ac_snr_db(r, g) >= MIN_AC_SNR_DB_BF16, (b, s, ac_snr_db(r, g))
    del wav
    torch.cuda.empty_cache()


@pytest.mark.parametrize("precision", ["bf16", "fp32"])
def test_int16_host_path_leaves_the_workspace_intact(dev, gens, precision):
    """ADVICE r01 (generator.cu:592): the int16 PCM of the host path used to be staged over the head of the workspace,
    i.e. over zero pad rows the bf16 kernels rely on, so the SECOND forward on a cached (B,T) workspace read garbage as
    padding.  Repeated int16 calls, slot re-use in the pipeline (B >= 3 x sub-batch) and a device forward afterwards must
    all be bit-equal to a device forward on a fresh generator."""
    gen = gens[0][precision]
    mel = synthetic_mel(7, 48, seed=11)
    ref = gen(mel.to(dev)).squeeze(1).cpu()
    exp16 = torch.from_numpy((ref.numpy() * np.float32(32768.0)).astype("int16"))
    for _ in range(3):                                   # same (B,T): second and third call run on a used workspace
        got = gen.infer_host(mel, int16_scale=32768.0).clone()
        assert got.dtype == torch.int16 and torch.equal(got, exp16)
    assert torch.equal(gen(mel.to(dev)).squeeze(1).cpu(), ref)      # device forward on the workspace the int16 call used
    for sub in (2, 3):                                   # 7 utterances, sub-batches of 2 / 3: every slot is re-used
        for _ in range(2):
            piped = gen.infer_host(mel, int16_scale=32768.0, sub_batch=sub)
            assert torch.equal(piped, exp16)
        assert torch.equal(gen.infer_host(mel, sub_batch=sub), ref)  # float through the same slots afterwards


def test_two_gpus_in_one_process(gens):
    """ADVICE r01 (pair_umma.cu:849 and friends): kernel attributes (> 48 KB of dynamic shared memory) and the SM count
    are per device.  One process, generator moved to cuda:1 after cuda:0 was used: same bits."""
    if torch.cuda.device_count() < 2:
        pytest.skip("needs two GPUs in one box")
    gen, sd = gens
    mel = synthetic_mel(3, 40, seed=2)
    for precision in ("bf16", "fp32"):
        a = gen[precision](mel.to("cuda:0")).cpu()
        g1 = make_gen(sd, precision, torch.device("cuda:1"))
        b = g1(mel.to("cuda:1")).cpu()
        assert torch.equal(a, b), precision
    x = torch.randn(4, 6, 256)
    d = torch.randint(0, 5, (4, 6))
    o0, _ = v.LengthRegulator()(x.to("cuda:0"), d.to("cuda:0"), None)
    o1, _ = v.LengthRegulator()(x.to("cuda:1"), d.to("cuda:1"), None)
    assert torch.equal(o0.cpu(), o1.cpu())


def test_host_path_into_a_caller_owned_pinned_buffer(dev, gens):
    """vocoder_infer(..., out=pinned): every sub-batch's device->host copy lands in the caller's buffer (no staging copy, no
    allocation); the arrays returned are views of it and equal the device path, float and int16, one call and pipelined."""
    gen = gens[0]["bf16"]
    mel = synthetic_mel(7, 40, seed=4)
    ref = gen(mel.to(dev)).squeeze(1).cpu()
    cfg = {"vocoder": {"model": "HiFi-GAN"}}
    for sub in (8, 3):
        gen.host_sub_batch = sub
        out = torch.full((7, 40 * 256), float("nan")).pin_memory()
        wavs = v.vocoder_infer(mel, gen, cfg, {}, lengths=[100] * 7, Normalize=False, out=out)
        assert torch.equal(out, ref) and len(wavs) == 7 and wavs[3].shape == (100,)
        assert wavs[3].base is not None and np.shares_memory(wavs[3], out.numpy())
        out16 = torch.zeros((7, 40 * 256), dtype=torch.int16).pin_memory()
        pc = {"preprocessing": {"audio": {"max_wav_value": 32768.0}}}
        v.vocoder_infer(mel, gen, cfg, pc, Normalize=True, out=out16)
        assert torch.equal(out16, torch.from_numpy((ref.numpy() * np.float32(32768.0)).astype("int16")))
    gen.host_sub_batch = 64
    with pytest.raises(RuntimeError):
        gen.infer_host(mel, out=torch.empty(7, 3))


def test_back_to_back_forwards_are_bit_identical(dev, gens):
    """The 33 tensor-pipe launches of a forward are chained by programmatic dependent launch: launch n + 1 sets up (and
    its producer threads wait in griddepcontrol.wait) while launch n drains.  A launch that touched its input early, or a
    buffer the previous forward still reads, shows up as run-to-run differences: 24 forwards queued back to back without a
    host synchronisation, two shapes alternating on the same generator (its two cached workspaces), both precision modes,
    must each reproduce the first forward of their shape bit for bit; so must forwards issued from a second stream."""
    for precision in ("bf16", "fp32"):
        gen = gens[0][precision]
        mels = [synthetic_mel(16, 172, seed=21).to(dev), synthetic_mel(5, 61, seed=22).to(dev)]
        refs = [gen(m).clone() for m in mels]
        torch.cuda.synchronize()
        outs = []
        for i in range(24):
            outs.append(gen(mels[i & 1]).clone())        # clone: the next forward of this shape reuses the output buffer
        torch.cuda.synchronize()
        for i, o in enumerate(outs):
            assert torch.equal(o, refs[i & 1]), (precision, i)
        side = torch.cuda.Stream(dev)
        side.wait_stream(torch.cuda.current_stream(dev))
        with torch.cuda.stream(side):
            o2 = [gen(mels[0]).clone() for _ in range(4)]
        side.synchronize()
        for o in o2:
            assert torch.equal(o, refs[0]), precision


def test_fresh_workspace_on_dirty_memory(dev, gens):
    """A workspace that has not run a (B, T) before only gets the pad / tail rows of its operand buffers cleared (one launch;
    csrc/generator.cu zero_pad_rows), not the whole allocation.  Here the allocator is made to hand out memory full of NaN
    bit patterns (0xFF bytes) for a shape the generator has not seen: the result must equal the one computed on a
    zero-filled allocation of the same shape, bit for bit -- anything that read a row nobody wrote would be NaN."""
    gen, sd = gens
    for B, T in ((3, 37), (2, 130)):
        mel = synthetic_mel(B, T, seed=31).to(dev)
        g0 = make_gen(sd, "bf16", dev)
        torch.cuda.empty_cache()
        need = _lib.lib().vo2w_generator_workspace_bytes(g0._get_handle(dev), B, T)
        clean = torch.zeros(need + (64 << 20), dtype=torch.uint8, device=dev)
        del clean                                           # cached block, all zero: the workspace is carved from it
        ref = g0(mel).clone()
        assert bool(torch.isfinite(ref).all())
        del g0
        torch.cuda.empty_cache()
        g1 = make_gen(sd, "bf16", dev)
        dirty = torch.full((need + (64 << 20),), 0xFF, dtype=torch.uint8, device=dev)
        del dirty                                           # cached block of NaN patterns
        got = g1(mel)
        assert torch.equal(got, ref), (B, T)
        got16 = g1.infer_host(mel.cpu(), int16_scale=32768.0)   # the host int16 path on the same (now used) workspace
        assert torch.equal(got16, torch.from_numpy((ref.squeeze(1).cpu().numpy() * np.float32(32768.0)).astype("int16")))
        del g1
