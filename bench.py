#!/usr/bin/env python
"""Benchmark of the synthesis hot path (HiFi-GAN V1 vocoder) on B200 -- driver contract.

    python bench.py --gpus N --steps K --warmup W          # our arm (torchrun for N > 1)
    python bench.py --impl reference --gpus N --steps K --warmup W   # CPU arm (reference's algorithm)

Workload = BASELINE.json configs[3] ("batch-sharded vocoder inference, 1024 utterances x 2 s across
1/2/4/8 B200"), the configuration the metric "vocoder audio-sec/sec at 1/2/4/8 B200" is quoted on:
a step is one pass of Generator.forward over all 1024 synthetic [80,172] mels, the utterances split
contiguously over the ranks (no data-path collective), each rank walking its shard in
sub-batches.  One JSON line on rank 0.
"""
from __future__ import annotations

import argparse
import contextlib
import json
import os
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

import torch

from visual_onoma_to_wave_b200.synthetic import (V1_CONFIG, FLOP_PER_FRAME, PAIR_FLOP_PER_FRAME, seeded_state_dict,
                                                 synthetic_mel)
from visual_onoma_to_wave_b200.sharding import shard_bounds

N_UTT, T_FRAMES, HOP, SR = 1024, 172, 256, 22050
AUDIO_S_PER_UTT = T_FRAMES * HOP / SR
METRIC, UNIT = "vocoder_audio_seconds_per_second", "audio-s/s"


def load_peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        d = json.load(open(p))
        return {"hbm_gbs": d["hbm_gbs"], "bf16_tflops": d["bf16_tflops"],
                "bf16_tflops_sustained": d.get("bf16_tflops_sustained", d["bf16_tflops"]), "source": "measured"}
    return {"hbm_gbs": 6650.0, "bf16_tflops": 1590.0, "bf16_tflops_sustained": 1400.0, "source": "fallback"}


class ClockSampler(threading.Thread):
    """Samples SM clocks and throttle reasons of one GPU while the timed region runs (pynvml)."""

    def __init__(self, index: int, period: float = 0.1):
        super().__init__(daemon=True)
        self.index, self.period = index, period
        self.samples, self.reasons, self.max_mhz = [], set(), None
        self._halt = threading.Event()
        self.ok = False
        try:
            import pynvml
            pynvml.nvmlInit()
            self.nv = pynvml
            self.h = pynvml.nvmlDeviceGetHandleByIndex(index)
            self.max_mhz = pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM)
            self.ok = True
        except Exception as e:  # pragma: no cover
            self.err = str(e)

    def run(self):
        if not self.ok:
            return
        nv = self.nv
        names = {
            getattr(nv, "nvmlClocksEventReasonHwSlowdown", 0x8): "hw_slowdown",
            getattr(nv, "nvmlClocksEventReasonHwThermalSlowdown", 0x40): "hw_thermal_slowdown",
            getattr(nv, "nvmlClocksEventReasonSwThermalSlowdown", 0x20): "sw_thermal_slowdown",
            getattr(nv, "nvmlClocksEventReasonSwPowerCap", 0x4): "sw_power_cap",
        }
        while not self._halt.is_set():
            try:
                self.samples.append(nv.nvmlDeviceGetClockInfo(self.h, nv.NVML_CLOCK_SM))
                try:
                    mask = nv.nvmlDeviceGetCurrentClocksEventReasons(self.h)
                except Exception:
                    mask = nv.nvmlDeviceGetCurrentClocksThrottleReasons(self.h)
                for bit, name in names.items():
                    if mask & bit:
                        self.reasons.add(name)
            except Exception:
                pass
            time.sleep(self.period)

    def stop(self):
        self._halt.set()
        self.join(timeout=2)
        s = sorted(self.samples)
        return {"sm_mhz": s[len(s) // 2] if s else None, "sm_max_mhz": self.max_mhz, "reasons": sorted(self.reasons),
                "samples": len(s)}


# ------------------------------------------------------------------------------------------
# CPU arm: the reference's algorithm (oracle port: the same ATen calls hifigan.Generator makes)
# ------------------------------------------------------------------------------------------
def cpu_generator_rate(n_utt: int, reps: int, warm: int = 1):
    from oracle import hifigan_oracle as ho
    torch.set_num_threads(os.cpu_count() or 1)
    sd = ho.fold_weight_norm(seeded_state_dict(1234))
    mel = synthetic_mel(n_utt, T_FRAMES, seed=0)
    times = []
    for i in range(warm + reps):
        t0 = time.perf_counter()
        ho.generator_forward(sd, mel)
        if i >= warm:
            times.append(time.perf_counter() - t0)
    return n_utt * AUDIO_S_PER_UTT, times


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    n_utt = args.ref_utts
    audio_s, times = cpu_generator_rate(n_utt, args.steps, args.warmup)
    total = sum(times)
    value = audio_s * len(times) / total
    sample = f"{n_utt} of the {N_UTT} utterances x {T_FRAMES} frames per step (bounded sample of the same workload)"
    line = {
        "impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": 1e3 * total / len(times), "higher_is_better": True, "scaling": args.scaling,
        "vs_baseline": None, "dtype": "fp32", "data": "synthetic",
        "config": {"workload": "configs[3]: batch-sharded vocoder inference, 1024 utterances x 2 s (T=172), HiFi-GAN V1, random-init",
                   "sample": sample, "host_threads": torch.get_num_threads()},
        "cpu_baseline": {"value": value, "unit": UNIT, "cores": torch.get_num_threads(), "kind": "port", "sample": sample,
                         "note": "oracle port = the ATen calls (F.conv1d/conv_transpose1d/leaky_relu/tanh) the reference's "
                                 "hifigan.Generator dispatches on CPU; the Python reference itself cannot travel to the GPU box"},
        "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line))


# ------------------------------------------------------------------------------------------
# our arm
# ------------------------------------------------------------------------------------------
def lr_bandwidth(dev, peaks):
    """LengthRegulator achieved HBM GB/s on a working set >> L2."""
    import visual_onoma_to_wave_b200 as v
    B, T, H = 4096, 64, 256
    g = torch.Generator(device=dev).manual_seed(0)
    x = torch.randn(B, T, H, device=dev, generator=g)
    d = torch.randint(0, 25, (B, T), device=dev, generator=g)
    lr = v.LengthRegulator()
    Lmax = int(d.sum(1).max())
    for _ in range(3):
        out, ml = lr(x, d, Lmax)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    n = 10
    e0.record()
    for _ in range(n):
        out, ml = lr(x, d, Lmax)
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / n
    bytes_alg = B * T * H * 4 + B * T * 8 + B * Lmax * H * 4 + 8 * B
    gbs = bytes_alg / (ms * 1e-3) / 1e9
    return {"workload": f"B={B} T_src={T} H={H} fp32, d~U{{0..24}}, Lmax={Lmax} (output {B * Lmax * H * 4 / 1e9:.2f} GB)",
            "ms": ms, "achieved_gbs": gbs, "peak_gbs": peaks["hbm_gbs"], "frac": gbs / peaks["hbm_gbs"],
            "algorithmic_bytes": bytes_alg}


def run_ours(args):
    import torch.distributed as dist
    import visual_onoma_to_wave_b200 as v

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device (this path has no CPU fallback; use --impl reference for the CPU arm)")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    peaks = load_peaks()

    # weak scaling (default): every rank holds configs[3]'s 1024 utterances, the job is the 1024 * N utterance batch
    # split contiguously; --scaling strong keeps the job at 1024 utterances in total
    n_job = N_UTT * world if args.scaling == "weak" else N_UTT
    lo, hi = shard_bounds(n_job, world, rank)
    n_local = hi - lo
    sub = min(args.sub_batch, n_local)
    gen = v.Generator(v.AttrDict(V1_CONFIG), precision=args.precision)
    gen.eval()
    with contextlib.redirect_stdout(sys.stderr):     # the reference prints here; stdout carries the JSON line only
        gen.remove_weight_norm()
    gen.load_state_dict(seeded_state_dict(1234))
    gen.to(dev)
    gen.host_sub_batch = sub
    mel_host = synthetic_mel(N_UTT, T_FRAMES, seed=0)[[i % N_UTT for i in range(lo, hi)]].contiguous().pin_memory()
    mel_dev = mel_host.to(dev)
    chunks = [(a, min(a + sub, n_local)) for a in range(0, n_local, sub)]
    cfg_m = {"vocoder": {"model": "HiFi-GAN"}}

    def step_device():
        out = None
        for a, b in chunks:
            out = gen(mel_dev[a:b])
        return out

    def step_e2e():
        # public API, host buffers (utils/model.py:73-98): ONE vocoder_infer call over this rank's shard; the library
        # walks it in sub-batches on two streams (pinned H2D, kernels, D2H overlapped) and returns numpy waveforms
        wavs = v.vocoder_infer(mel_host, gen, cfg_m, {}, Normalize=False)
        return len(wavs)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def timed(fn, steps, warmup, sampler=None):
        for _ in range(warmup):
            fn()
        barrier()
        if sampler:
            sampler.start()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        t0 = time.perf_counter()
        e0.record()
        for _ in range(steps):
            fn()
        e1.record()
        barrier()
        wall = time.perf_counter() - t0
        clocks = sampler.stop() if sampler else None
        ms = torch.tensor([e0.elapsed_time(e1), wall * 1e3], device=dev, dtype=torch.float64)
        if world > 1:
            dist.all_reduce(ms, op=dist.ReduceOp.MAX)
        return float(ms[0]), float(ms[1]), clocks

    warm = max(args.warmup, 3)
    dev_ms, _, clocks = timed(step_device, args.steps, warm, ClockSampler(local))
    launches_per_step = gen.last_launches * len(chunks)
    _, e2e_wall_ms, _ = timed(step_e2e, args.steps, 1)

    total_audio_s = n_job * AUDIO_S_PER_UTT
    ms_per_step = dev_ms / args.steps
    value = total_audio_s / (ms_per_step * 1e-3)
    e2e_value = total_audio_s / (e2e_wall_ms / args.steps * 1e-3)

    # roofline of the dominant kernel, pair_umma_kernel (28 of the 35 launches of a forward): its launches are timed
    # live with CUDA events on the launching stream (vo2w_generator_set_timing), one sub-batch per pass, right after
    # the timed steps (GPU hot).  achieved = algorithmic FLOPs of those launches / their summed duration.
    a, b = chunks[0]
    pair_ms, all_ms, n_pair = 0.0, 0.0, 0
    passes = 3
    if args.precision == "bf16":
        for _ in range(passes):
            ms_l, kinds = gen.launch_times(mel_dev[a:b])
            pair_ms += sum(m for m, k in zip(ms_l, kinds) if k == 2)
            all_ms += sum(ms_l)
            n_pair += sum(1 for k in kinds if k == 2)
    if n_pair:
        flops_pair = PAIR_FLOP_PER_FRAME * (b - a) * T_FRAMES * passes
        achieved = flops_pair / (pair_ms * 1e-3) / 1e12
        per_launch = {"flops_per_launch": flops_pair / n_pair, "avg_launch_us": pair_ms * 1e3 / n_pair,
                      "share_of_forward_time": pair_ms / all_ms, "launches_timed": n_pair}
    else:   # fp32 mode: whole-step figure
        flops_local = FLOP_PER_FRAME * n_local * T_FRAMES
        achieved = flops_local / (ms_per_step * 1e-3) / 1e12
        per_launch = {"flops_per_launch": flops_local / launches_per_step, "avg_launch_us": ms_per_step * 1e3 / launches_per_step}
    peak = peaks["bf16_tflops_sustained"]
    roofline = {"bound": "tensor", "kernel": "pair_umma_kernel (28 of the 35 launches of a forward, 96.8 % of its FLOPs)" if args.precision == "bf16" else "conv1d_f32_kernel",
                "achieved": achieved, "peak": peak, "unit": "TFLOP/s", "frac": achieved / peak,
                "peak_source": f"bf16_tflops_sustained of {peaks['source']} MEASURED_PEAKS.json (kernel timed inside a long step)",
                "traffic": None, **per_launch}
    prof = os.path.join(ROOT, "profiles", "traffic.json")
    if os.path.exists(prof) and args.precision == "bf16":
        try:   # ncu --set full capture at batch tj["batch"]; DRAM bytes scale linearly with the utterances of a launch
            tj = json.load(open(prof))
            roofline["traffic"] = tj["pair_umma_kernel_dram_bytes_per_launch"] * (b - a) / tj["batch"]
            roofline["traffic_source"] = tj["source"]
        except Exception:
            pass

    extra = {}
    cpu_baseline = None
    if rank == 0:
        try:
            extra["length_regulator"] = lr_bandwidth(dev, peaks)
        except Exception as e:  # reported, not fatal for the vocoder metric
            extra["length_regulator"] = {"error": str(e)}
        if world == 1 and args.precision == "bf16":
            try:   # the exact mode (max-abs-err <= 1e-4 gate): same Generator on the tf32 hi/lo-split tensor-pipe kernel
                g32 = v.Generator(v.AttrDict(V1_CONFIG), precision="fp32")
                g32.eval()
                with contextlib.redirect_stdout(sys.stderr):
                    g32.remove_weight_norm()
                g32.load_state_dict(seeded_state_dict(1234))
                g32.to(dev)
                m32 = mel_dev[:64]
                for _ in range(2):
                    g32(m32)
                torch.cuda.synchronize(dev)
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                e0.record()
                for _ in range(3):
                    g32(m32)
                e1.record()
                torch.cuda.synchronize(dev)
                ms32 = e0.elapsed_time(e1) / 3
                nb = m32.shape[0]
                extra["fp32_mode"] = {"utterances": nb, "ms": ms32, "audio_s_per_s": nb * T_FRAMES * 256 / 22050 / ms32 * 1e3,
                                      "tflops_fp32_equivalent": FLOP_PER_FRAME * nb * T_FRAMES / ms32 / 1e9,
                                      "kernel": "tf32_conv_kernel: tcgen05 kind::tf32, 3 products per fp32 product"}
                del g32
            except Exception as e:
                extra["fp32_mode"] = {"error": str(e)}
        if world == 1:
            try:   # the acoustic tail in front of the vocoder, device timing per stage
                sys.path.insert(0, os.path.join(ROOT, "tools"))
                import bench_tail
                with torch.cuda.device(dev), contextlib.redirect_stdout(sys.stderr):
                    extra["acoustic_tail"] = bench_tail.measure(steps=3, warmup=2, batch=64, dev=dev)
            except Exception as e:
                extra["acoustic_tail"] = {"error": str(e)}
        if world == 1 and not args.no_cpu_baseline:
            n_cpu = 64
            audio_s, times = cpu_generator_rate(n_cpu, reps=2, warm=1)
            cpu_baseline = {"value": audio_s / min(times), "unit": UNIT, "cores": torch.get_num_threads(), "kind": "port",
                            "sample": f"{n_cpu} of the {N_UTT} utterances x {T_FRAMES} frames, best of 2 after 1 warm-up "
                                      f"({sum(times):.1f} s of CPU work)"}
    if rank == 0:
        line = {
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": warm,
            "ms_per_step": ms_per_step, "higher_is_better": True, "scaling": args.scaling, "vs_baseline": None,
            "dtype": args.precision, "data": "synthetic",
            "config": {"workload": "configs[3]: batch-sharded vocoder inference, 1024 utterances x 2 s (T=172 frames -> 44032 samples), "
                                   "HiFi-GAN V1 Generator.forward, random-init weights (seed 1234); " + ("weak scaling: 1024 utterances per GPU, the 1024 x N job split contiguously over ranks" if args.scaling == "weak" else "strong scaling: 1024 utterances in total split contiguously over ranks"),
                       "utterances": n_job, "utterances_per_gpu": n_local, "frames": T_FRAMES, "sub_batch": sub, "parallelism": f"utterance-shard x{world}",
                       "precision_mode": args.precision,
                       "l2": "no flush: each step streams >100 GB of activations per GPU through a 126 MB L2; inputs (56 MB) do not survive"},
            "e2e": {"value": e2e_value, "unit": UNIT, "h2d_bytes_per_step": n_job * 80 * T_FRAMES * 4,
                    "d2h_bytes_per_step": n_job * T_FRAMES * HOP * 4,
                    "api": "vocoder_infer(host mels of the whole shard) -> vo2w_generator_forward_host_async per sub-batch on two alternating streams (pinned H2D, kernels, D2H), numpy out, wall clock"},
            "gpu_launches": launches_per_step * args.steps,
            "clocks": clocks, "roofline": roofline, "cpu_baseline": cpu_baseline,
            "tflops": FLOP_PER_FRAME * n_job * T_FRAMES / (ms_per_step * 1e-3) / 1e12, "extra": extra,
        }
        print(json.dumps(line))
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--precision", default="bf16", choices=["bf16", "fp32"])
    ap.add_argument("--sub-batch", type=int, default=128)
    ap.add_argument("--ref-utts", type=int, default=16, help="utterances per step of the CPU arm (bounded sample)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--scaling", default="weak", choices=["weak", "strong"],
                    help="weak: 1024 utterances per GPU (job = 1024 x N); strong: 1024 utterances in total")
    args = ap.parse_args()
    if args.impl == "reference":
        run_reference(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()
