"""Device timing of the acoustic tail in front of the vocoder at configs[2]'s shape:
B=64 utterances, T_src=32 tokens, durations ~U{0..24} => ~480 frames.  One JSON line.  "tflops" are fp32-equivalent
(the reference's FLOPs): the Linear / Conv1d layers run on the tensor pipe as 3 TF32 products per fp32 product
(conv_tf32.cu), attention / LayerNorm / the predictors on fp32 CUDA cores.  bench.py reports the same numbers under
"extra".acoustic_tail.

    python tools/bench_tail.py [--steps 5] [--warmup 3] [--batch 64]
"""
import argparse
import json
import os
import sys
import tempfile

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402
import visual_onoma_to_wave_b200 as v  # noqa: E402
from visual_onoma_to_wave_b200.synthetic import (VARIANCE_STATS, V1_CONFIG, decoder_config, seeded_decoder_state_dict,  # noqa: E402
                                                 seeded_postnet_state_dict, seeded_state_dict,
                                                 seeded_variance_adaptor_state_dict, variance_model_config)


def timed(fn, steps, warmup):
    for _ in range(warmup):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(steps):
        r = fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / steps, r


def measure(steps=5, warmup=3, batch=64, dev=None):
    class A:
        pass
    a = A()
    a.steps, a.warmup, a.batch = steps, warmup, batch
    dev = dev or torch.device("cuda:0")
    tmp = tempfile.mkdtemp()
    with open(os.path.join(tmp, "stats.json"), "w") as f:
        json.dump(VARIANCE_STATS, f)
    model_config = {**variance_model_config(), **decoder_config()}
    model_config["transformer"]["encoder_hidden"] = 256
    pre_cfg = {"path": {"preprocessed": tmp}, "preprocessing": {"mel": {"n_mel_channels": 80}}}
    tail = v.AcousticTail(pre_cfg, model_config)
    g = torch.Generator().manual_seed(1)
    sd = {**{"variance_adaptor." + k: t for k, t in seeded_variance_adaptor_state_dict(1).items()},
          **{"decoder." + k: t for k, t in seeded_decoder_state_dict(1).items()},
          **{"postnet." + k: t for k, t in seeded_postnet_state_dict(1).items()},
          "mel_linear.weight": torch.randn(80, 256, generator=g) / 16, "mel_linear.bias": torch.randn(80, generator=g) * 0.1}
    tail.load_state_dict(sd)
    tail.eval().to(dev)
    gen = v.Generator(v.AttrDict(V1_CONFIG), precision="bf16")
    gen.eval()
    gen.remove_weight_norm()
    gen.load_state_dict(seeded_state_dict(1234))
    gen.to(dev)

    B, T = a.batch, 32
    x = torch.randn(B, T, 256, generator=g).to(dev)
    src_mask = torch.zeros(B, T, dtype=torch.bool, device=dev)
    dur = torch.randint(0, 25, (B, T), generator=g).to(dev)
    va, dec, pn = tail.variance_adaptor, tail.decoder, tail.postnet
    stages = {}
    ms, r = timed(lambda: va(x, src_mask, duration_target=dur), a.steps, a.warmup)
    h, mel_len = r[0], r[5]
    L = h.shape[1]
    mel_mask = v.get_mask_from_lengths(mel_len, L)
    frames = B * L
    D, H = 256, 1024
    stages["variance_adaptor"] = {"ms": ms, "flop": 2 * B * T * 2 * (2 * 256 * 256 * 3 + 256),
                                  "what": "2 predictors + bucketize/embedding + LengthRegulator (duration targets)"}
    ms, r = timed(lambda: dec(h, mel_mask), a.steps, a.warmup)
    stages["decoder"] = {"ms": ms, "flop": 2 * frames * 6 * (4 * D * D + 2 * L * D + D * H * 9 + H * D)}
    d_out = r[0]
    ms, r = timed(lambda: pn(v.mel_linear(d_out, tail.mel_linear.weight, tail.mel_linear.bias), add_input=True), a.steps, a.warmup)
    stages["mel_linear_postnet"] = {"ms": ms, "flop": 2 * frames * (256 * 80 + 5 * (80 * 512 + 3 * 512 * 512 + 512 * 80))}
    mel = r.transpose(1, 2).contiguous()
    ms, _ = timed(lambda: gen(mel), a.steps, a.warmup)
    stages["generator_bf16"] = {"ms": ms, "flop": 614_105_088 * frames}
    ms_all, _ = timed(lambda: gen(tail(x, src_mask, mel_masks=mel_mask, max_mel_len=L, d_targets=dur)[1].transpose(1, 2).contiguous()), a.steps, a.warmup)
    for s in stages.values():
        s["tflops_fp32_equivalent"] = s["flop"] / s["ms"] / 1e9
    return {"workload": f"configs[2] shape: B={B}, T_src={T}, d~U{{0..24}} -> Lmax={L} ({frames} frames incl. padding)",
            "stages": stages, "chain_ms": ms_all, "audio_seconds": frames * 256 / 22050,
            "chain_audio_s_per_s": frames * 256 / 22050 / ms_all * 1e3,
            "chain": "VarianceAdaptor -> Decoder -> mel_linear -> PostNet -> Generator(bf16), device-resident input"}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--batch", type=int, default=64)
    a = ap.parse_args()
    print(json.dumps(measure(a.steps, a.warmup, a.batch)))


if __name__ == "__main__":
    main()
