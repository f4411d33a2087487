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
from visual_onoma_to_wave_b200.sharding import shard_bounds, max_over_ranks

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


def make_generator(v, precision, dev, sub):
    gen = v.Generator(v.AttrDict(V1_CONFIG), precision=precision)
    gen.eval()
    with contextlib.redirect_stdout(sys.stderr):     # the reference prints here; stdout carries the JSON line only
        gen.remove_weight_norm()
    gen.load_state_dict(seeded_state_dict(1234))
    gen.to(dev)
    gen.host_sub_batch = sub
    return gen


def device_ms(fn, steps, warmup):
    """CUDA-event time per call on torch's current stream (the stream the library launches on)."""
    for _ in range(warmup):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(steps):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / steps


def tf32_peak(dev):
    """TF32 tensor peak measured the way MEASURED_PEAKS.json measures bf16: torch.matmul 8192^3 with allow_tf32, best of
    10 (burst) and back to back for 2 s (sustained).  BASELINE.md section 2 left this number to the builder."""
    old = torch.backends.cuda.matmul.allow_tf32
    torch.backends.cuda.matmul.allow_tf32 = True
    try:
        n = 8192
        a = torch.randn(n, n, device=dev)
        b = torch.randn(n, n, device=dev)
        flops = 2.0 * n ** 3
        for _ in range(3):
            a @ b
        torch.cuda.synchronize()
        best = 1e9
        for _ in range(10):
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            a @ b
            e1.record()
            torch.cuda.synchronize()
            best = min(best, e0.elapsed_time(e1))
        reps = max(10, int(2000.0 / best))
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(reps):
            a @ b
        e1.record()
        torch.cuda.synchronize()
        sus = e0.elapsed_time(e1) / reps
        return {"tf32_tflops": flops / best / 1e9, "tf32_tflops_sustained": flops / sus / 1e9,
                "how": f"torch.matmul fp32 {n}^3 with allow_tf32=True: best of 10 (burst), {reps} back to back (sustained)"}
    finally:
        torch.backends.cuda.matmul.allow_tf32 = old


def graph_ms(gen, mel, steps=10):
    """One Generator.forward captured in a CUDA graph (the library's launches are plain stream work): ms per replay."""
    side = torch.cuda.Stream(mel.device)
    side.wait_stream(torch.cuda.current_stream(mel.device))
    with torch.cuda.stream(side):
        for _ in range(2):
            gen(mel)                      # warm-up on the capture stream: its workspace exists and is initialised
    side.synchronize()
    graph = torch.cuda.CUDAGraph()
    with torch.cuda.graph(graph, stream=side):
        wav = gen(mel)
    for _ in range(2):
        graph.replay()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(steps):
        graph.replay()
    e1.record()
    torch.cuda.synchronize()
    del wav
    return e0.elapsed_time(e1) / steps


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
        # NCCL as the contract says; VO2W_BENCH_PG=gloo (experiments: the path has no data-path collective, the group only
        # carries the timing barrier and the max over ranks) keeps NCCL out of the process
        if os.environ.get("VO2W_BENCH_PG", "nccl") == "gloo":
            dist.init_process_group("gloo")
        else:
            dist.init_process_group("nccl", device_id=dev)
    peaks = load_peaks()
    cfg_m = {"vocoder": {"model": "HiFi-GAN"}}
    mel_all = synthetic_mel(N_UTT, T_FRAMES, seed=0)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def timed(fn, steps, warmup, sampler=None):
        # A barrier BEFORE the warm-up as well: the first collective on an NCCL communicator sets its connections up
        # (hundreds of ms with the GPU idle).  Left between warm-up and timing it hands the timed steps a cold, power-
        # credited GPU: measured at 2 ranks, 3 steps: 102.6 ms per step instead of the steady 107.8.
        barrier()
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
        ms = max_over_ranks([e0.elapsed_time(e1), wall * 1e3], None if os.environ.get("VO2W_BENCH_PG") == "gloo" else dev)      # the job's time is its slowest rank's
        return ms[0], ms[1], clocks

    def run_job(gen, n_job, sub_batch, steps, warm, sampler=None):
        """One job of n_job utterances split contiguously over the ranks: (device ms, e2e wall ms, clocks, shard info).
        device: inputs resident in HBM, CUDA events; e2e: ONE vocoder_infer call over the rank's host mels (pinned), the
        library walks them in sub-batches on two streams and copies every result straight into the caller's pinned
        result buffer (utils/model.py:73-98: host mel in, numpy out)."""
        lo, hi = shard_bounds(n_job, world, rank)
        n_local = hi - lo
        sub = max(1, min(sub_batch, n_local))
        if n_local <= sub_batch and n_local >= 64:
            sub = (n_local + 1) // 2     # a shard of one sub-batch is still walked as two, so that the copies of one half hide
                                         # under the kernels of the other (a forward of 64 utterances is as efficient as one of 128)
        gen.host_sub_batch = sub
        mel_host = mel_all[[i % N_UTT for i in range(lo, hi)]].contiguous().pin_memory()
        mel_dev = mel_host.to(dev)
        out_host = torch.empty(n_local, T_FRAMES * HOP).pin_memory()
        chunks = [(a, min(a + sub, n_local)) for a in range(0, n_local, sub)]

        def step_device():
            out = None
            for a, b in chunks:
                out = gen(mel_dev[a:b])
            return out

        def step_e2e():
            wavs = v.vocoder_infer(mel_host, gen, cfg_m, {}, Normalize=False, out=out_host)
            return len(wavs)

        dev_ms, _, clocks = timed(step_device, steps, warm, sampler)
        launches = gen.last_launches * len(chunks)
        _, e2e_ms, _ = timed(step_e2e, steps, 1)
        return {"dev_ms": dev_ms / steps, "e2e_ms": e2e_ms / steps, "clocks": clocks, "n_local": n_local, "sub": sub,
                "chunks": chunks, "launches": launches, "mel_dev": mel_dev}

    # weak scaling (default): every rank holds configs[3]'s 1024 utterances, the job is the 1024 * N utterance batch
    # split contiguously; --scaling strong keeps the job at 1024 utterances in total
    n_job = N_UTT * world if args.scaling == "weak" else N_UTT
    gen = make_generator(v, args.precision, dev, args.sub_batch)
    warm = max(args.warmup, 3)
    job = run_job(gen, n_job, args.sub_batch, args.steps, warm, ClockSampler(local))
    n_local, sub, chunks, mel_dev = job["n_local"], job["sub"], job["chunks"], job["mel_dev"]
    launches_per_step = job["launches"]
    total_audio_s = n_job * AUDIO_S_PER_UTT
    ms_per_step = job["dev_ms"]
    value = total_audio_s / (ms_per_step * 1e-3)
    e2e_value = total_audio_s / (job["e2e_ms"] * 1e-3)
    clocks = job["clocks"]

    extra = {}
    # configs[3] as BASELINE.json words it -- ONE 1024-utterance job split over the N GPUs (strong scaling) -- timed after
    # the weak region at N > 1 (at N = 1 it is the headline itself)
    if world > 1 and args.scaling == "weak":
        sj = run_job(gen, N_UTT, args.sub_batch, args.steps, 2)
        a_s = N_UTT * AUDIO_S_PER_UTT
        extra["strong_scaling"] = {"utterances": N_UTT, "utterances_per_gpu": sj["n_local"], "sub_batch": sj["sub"],
                                   "ms_per_step": sj["dev_ms"], "value": a_s / (sj["dev_ms"] * 1e-3), "unit": UNIT,
                                   "e2e": {"value": a_s / (sj["e2e_ms"] * 1e-3), "unit": UNIT,
                                           "h2d_bytes_per_step": N_UTT * 80 * T_FRAMES * 4, "d2h_bytes_per_step": N_UTT * T_FRAMES * HOP * 4},
                                   "note": "the fixed 1024-utterance job of configs[3] split contiguously over the ranks; device time and "
                                           "wall time are the max over ranks"}

    # roofline of the dominant kernel, pair_umma_kernel (28 of the 34 launches of a forward): its launches are timed
    # live with CUDA events on the launching stream (vo2w_generator_set_timing), one sub-batch per pass, right after
    # the timed steps (GPU hot).  achieved = algorithmic FLOPs of the kernel's launches in a step / (the step's device
    # time x the kernel's share of a forward); the event pass's own, serialised durations are reported next to it.
    a, b = chunks[0]
    pair_ms, all_ms, n_pair = 0.0, 0.0, 0
    passes = 3
    kind_ms = {}
    if args.precision == "bf16":
        for _ in range(passes):
            ms_l, kinds = gen.launch_times(mel_dev[a:b])
            pair_ms += sum(m for m, k in zip(ms_l, kinds) if k == 2)
            all_ms += sum(ms_l)
            n_pair += sum(1 for k in kinds if k == 2)
            for m, k in zip(ms_l, kinds):
                kind_ms[k] = kind_ms.get(k, 0.0) + m
    if n_pair:
        # One event per launch serialises the launches (an event between two kernels is a full dependency: no programmatic
        # dependent launch, every kernel's tail and set-up exposed), so the per-launch durations of this pass are ~5 - 9 %
        # longer than the same launches inside the timed steps.  What the pass gives reliably is the kernel's SHARE of a
        # forward (it agrees with the ncu launch list, profiles/r02_bench_launches.csv); the kernel's time inside the timed
        # region is that share of the region's device time.  Both figures are reported.
        flops_pair = PAIR_FLOP_PER_FRAME * (b - a) * T_FRAMES * passes
        share = pair_ms / all_ms
        serialised = flops_pair / (pair_ms * 1e-3) / 1e12
        launches_in_step = (n_pair // passes) * len(chunks)
        pair_ms_in_step = ms_per_step * share
        achieved = PAIR_FLOP_PER_FRAME * n_local * T_FRAMES / (pair_ms_in_step * 1e-3) / 1e12
        per_launch = {"flops_per_launch": flops_pair / n_pair, "avg_launch_us": pair_ms_in_step * 1e3 / launches_in_step,
                      "how": "CUDA events around the timed steps (ms_per_step) x the kernel's share of a forward, measured with one "
                             "CUDA event per launch on the launching stream right after the timed steps",
                      "serialised": {"avg_launch_us": pair_ms * 1e3 / n_pair, "achieved": serialised,
                                     "frac": serialised / peaks["bf16_tflops_sustained"],
                                     "what": "the per-launch event pass itself: launches serialised by the events between them"},
                      "share_of_forward_time": share, "launches_timed": n_pair,
                      "other_launches": {"single_conv_share_of_forward_time": kind_ms.get(1, 0.0) / all_ms,
                                         "single_conv_us_per_forward": kind_ms.get(1, 0.0) * 1e3 / passes,
                                         "what": "conv_pre + 4 upsamplers on stream_umma_kernel (conv_post rides the last fused launch)"}}
    else:   # fp32 mode: whole-step figure
        flops_local = FLOP_PER_FRAME * n_local * T_FRAMES
        achieved = flops_local / (ms_per_step * 1e-3) / 1e12
        per_launch = {"flops_per_launch": flops_local / launches_per_step, "avg_launch_us": ms_per_step * 1e3 / launches_per_step}
    peak = peaks["bf16_tflops_sustained"]
    roofline = {"bound": "tensor", "kernel": "pair_umma_kernel (28 of the 34 launches of a forward, 96.8 % of its FLOPs)" if args.precision == "bf16" else "tf32_conv_kernel",
                "achieved": achieved, "peak": peak, "unit": "TFLOP/s", "frac": achieved / peak,
                "peak_source": f"bf16_tflops_sustained of {peaks['source']} MEASURED_PEAKS.json (kernel timed inside a long step)",
                "frac_of_burst_peak": achieved / peaks["bf16_tflops"],
                "traffic": None, **per_launch}
    # DRAM bytes per launch come from an `ncu --set full` capture and are only reported while that capture is of the
    # kernel source being timed (sha256 of csrc/pair_umma.cu recorded next to it); otherwise null, never a stale constant
    prof = os.path.join(ROOT, "profiles", "traffic.json")
    if os.path.exists(prof) and args.precision == "bf16":
        try:
            import hashlib
            tj = json.load(open(prof))
            src = os.path.join(ROOT, "visual_onoma_to_wave_b200", "csrc", "pair_umma.cu")
            if tj.get("pair_umma_cu_sha256") == hashlib.sha256(open(src, "rb").read()).hexdigest():
                roofline["traffic"] = tj["pair_umma_kernel_dram_bytes_per_launch"] * (b - a) / tj["batch"]
                roofline["traffic_source"] = tj["source"]
            else:
                roofline["traffic_note"] = "null: the committed ncu capture (profiles/traffic.json) is of an older pair_umma.cu"
        except Exception:
            pass

    cpu_baseline = None
    if rank == 0:
        try:
            extra["length_regulator"] = lr_bandwidth(dev, peaks)
        except Exception as e:  # reported, not fatal for the vocoder metric
            extra["length_regulator"] = {"error": str(e)}
        if world == 1 and args.precision == "bf16" and not args.no_extras:
            try:   # the exact mode (max-abs-err <= 1e-4 gate) on the FULL configs[3]: tcgen05 kind::tf32, hi/lo split
                tp = tf32_peak(dev)
                g32 = make_generator(v, "fp32", dev, 64)
                j32 = run_job(g32, N_UTT, 64, 2, 1)
                a_s = N_UTT * AUDIO_S_PER_UTT
                eq = FLOP_PER_FRAME * N_UTT * T_FRAMES / (j32["dev_ms"] * 1e-3) / 1e12
                extra["fp32_mode"] = {"utterances": N_UTT, "sub_batch": 64, "ms_per_step": j32["dev_ms"], "value": a_s / (j32["dev_ms"] * 1e-3),
                                      "unit": UNIT, "e2e": {"value": a_s / (j32["e2e_ms"] * 1e-3), "unit": UNIT},
                                      "tflops_fp32_equivalent": eq, "tflops_tf32_issued": 3 * eq,
                                      "roofline": {"bound": "tensor", "achieved": 3 * eq, "peak": tp["tf32_tflops_sustained"], "unit": "TFLOP/s",
                                                   "frac": 3 * eq / tp["tf32_tflops_sustained"], "frac_of_burst_peak": 3 * eq / tp["tf32_tflops"],
                                                   "note": "whole forward; three TF32 products per fp32 product (hi.hi + lo.hi + hi.lo)"},
                                      "tf32_peak": tp, "kernel": "tf32_conv_kernel: tcgen05 kind::tf32, 3 products per fp32 product"}
                # configs[1]: B=16 x T=172, both modes, eager launches and one CUDA graph per forward
                m16 = mel_dev[:16].contiguous()
                c2 = {}
                for name, gg in (("bf16", gen), ("fp32", g32)):
                    ms_e = device_ms(lambda: gg(m16), 10, 3)
                    try:
                        ms_g = graph_ms(gg, m16)
                    except Exception as e:
                        ms_g = None
                        c2[name + "_graph_error"] = str(e)
                    a16 = 16 * AUDIO_S_PER_UTT
                    c2[name] = {"ms": ms_e, "audio_s_per_s": a16 / (ms_e * 1e-3), "ms_cuda_graph": ms_g,
                                "audio_s_per_s_cuda_graph": a16 / (ms_g * 1e-3) if ms_g else None, "launches": gg.last_launches}
                extra["config2"] = {"workload": "configs[1]: HiFi-GAN V1 Generator forward, batch 16 x 2 s (T=172), device-resident", **c2}
                del g32
                torch.cuda.empty_cache()
                # configs[4]: long-form, B=8 x T=2584 (30 s)
                m5 = synthetic_mel(8, 2584, seed=5).to(dev)
                ms5 = device_ms(lambda: gen(m5), 3, 2)
                ws5 = int(_ws_bytes(gen, 8, 2584))
                extra["config5"] = {"workload": "configs[4]: long-form synthesis, 30 s mels (T=2584) batch 8, bf16 mode, device-resident",
                                    "ms": ms5, "audio_s_per_s": 8 * 2584 * HOP / SR / (ms5 * 1e-3), "workspace_bytes": ws5,
                                    "tflops": FLOP_PER_FRAME * 8 * 2584 / (ms5 * 1e-3) / 1e12}
                del m5
                gen._ws.clear()
                torch.cuda.empty_cache()
            except Exception as e:
                extra["extras_error"] = repr(e)
        if world == 1 and not args.no_extras:
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
                    "api": "vocoder_infer(host mels of the whole shard, out=pinned result buffer) -> vo2w_generator_forward_host_async per sub-batch on two alternating streams (pinned H2D, kernels, D2H straight into the caller's buffer), numpy views out, wall clock"},
            "gpu_launches": launches_per_step * args.steps,
            "clocks": clocks, "roofline": roofline, "cpu_baseline": cpu_baseline,
            "tflops": FLOP_PER_FRAME * n_job * T_FRAMES / (ms_per_step * 1e-3) / 1e12, "extra": extra,
        }
        print(json.dumps(line))
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


def _ws_bytes(gen, B, T):
    from visual_onoma_to_wave_b200 import _lib
    return _lib.lib().vo2w_generator_workspace_bytes(gen._handle, B, T)


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
    ap.add_argument("--no-extras", action="store_true", help="skip extra.fp32_mode / config2 / config5 / acoustic_tail")
    ap.add_argument("--scaling", default="weak", choices=["weak", "strong"],
                    help="weak: 1024 utterances per GPU (job = 1024 x N); strong: 1024 utterances in total")
    args = ap.parse_args()
    if args.impl == "reference":
        run_reference(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()
