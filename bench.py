#!/usr/bin/env python
"""Benchmark of the Fbank + ConvAutoencoder (or FullyConnected) anonymization hot path.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--workload convae|fcae] [--impl ours|reference]

One "step" = one batch of `--batch` synthetic 10 s / 16 kHz utterances per GPU taken from a
resident pool of `--pool` utterances per GPU (default 360 = 1 h of audio, BASELINE.json configs[4]),
through Fbank -> clamp -> InputNormalization -> pad(x36) -> anonymizer (random-init weights, seed 8886).
Prints ONE JSON line (see the contract in the task statement / DESIGN.md section "Measurement").
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np
import torch

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests", "golden"))

SAMPLE_RATE, N_SAMPLES, UTT_SECONDS = 16000, 160000, 10.0
# algorithmic work per 10 s utterance
FLOP_CONVAE_UTT = 22.28e9          # 15.02 GFLOP autoencoder + 7.26 GFLOP TDNN classifier
BYTES_FBANK_UTT = 960320           # 640000 B PCM in + 320320 B features out
FLOP_FCAE_UTT = 36.8e6


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=12)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default="convae", choices=["convae", "fcae"])
    ap.add_argument("--batch", type=int, default=60)
    ap.add_argument("--pool", type=int, default=360)
    ap.add_argument("--precision", default="bf16", choices=["bf16", "fp32"])
    ap.add_argument("--cpu-sample", type=int, default=4, help="utterances in the CPU baseline sample")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    return ap.parse_args()


def peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        d = json.load(open(p))
        return {"hbm": d["hbm_gbs"], "tf_burst": d["bf16_tflops"], "tf_sust": d.get("bf16_tflops_sustained", d["bf16_tflops"]),
                "src": "measured"}
    return {"hbm": 6650.0, "tf_burst": 1590.0, "tf_sust": 1400.0, "src": "fallback"}


def real_norm_stats():
    g = np.load(os.path.join(ROOT, "tests", "golden", "golden_v1.npz"))
    return g["real_glob_mean"].astype(np.float32), g["real_glob_std"].astype(np.float32), int(g["real_count"])


def make_wavs(n, seed):
    g = torch.Generator().manual_seed(seed)
    return 0.1 * torch.randn(n, N_SAMPLES, generator=g)


# --------------------------------------------------------------------------- CPU arm (oracle port)
def cpu_anonymize(wav_np, weights, mean, std, workload, noise):
    """The reference's CPU path restated by oracle/ (the reference itself does not travel to the GPU
    box): fp64 numpy Fbank -> normalise -> pad -> torch.nn.functional anonymizer on all host threads."""
    from oracle import anonymizer_ref, fbank_np, normalize_np
    f = fbank_np.fbank(wav_np, n_mels=80)
    f = (f - mean.astype(np.float64)) / std.astype(np.float64)
    with torch.no_grad():
        if workload == "convae":
            f = normalize_np.pad_time(f)
            return anonymizer_ref.convae_forward(f.astype(np.float32), weights, noise)
        return anonymizer_ref.fcae_forward(f.astype(np.float32), weights, noise)


def cpu_weights(workload):
    import recipe
    shapes = recipe.CONVAE_SHAPES if workload == "convae" else recipe.FCAE_SHAPES
    return recipe.make_weights(shapes, 8886), shapes


def time_cpu(workload, n_utts, steps, warmup):
    import recipe
    torch.set_num_threads(os.cpu_count() or 1)
    w, _ = cpu_weights(workload)
    mean, std, _ = real_norm_stats()
    wav = make_wavs(n_utts, 99).numpy()
    noise = recipe.make_noise(1, n_utts, 128 if workload == "convae" else 40)
    for _ in range(warmup):
        cpu_anonymize(wav, w, mean, std, workload, noise)
    t0 = time.perf_counter()
    for _ in range(steps):
        cpu_anonymize(wav, w, mean, std, workload, noise)
    dt = (time.perf_counter() - t0) / steps
    return n_utts * UTT_SECONDS / dt, dt, torch.get_num_threads()


# --------------------------------------------------------------------------- clocks sampler
class ClockSampler(threading.Thread):
    Q = "clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown," \
        "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap"

    def __init__(self, index):
        super().__init__(daemon=True)
        self.index, self.rows, self.stop_flag = index, [], False

    def run(self):
        while not self.stop_flag:
            try:
                out = subprocess.run(["nvidia-smi", "-i", str(self.index), "--query-gpu=" + self.Q,
                                      "--format=csv,noheader,nounits"], capture_output=True, text=True, timeout=5).stdout
                parts = [p.strip() for p in out.strip().split(",")]
                if len(parts) >= 6:
                    self.rows.append(parts)
            except Exception:
                pass
            time.sleep(0.1)

    def summary(self):
        if not self.rows:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        sm = sorted(float(r[0]) for r in self.rows)
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = [n for i, n in enumerate(names) if any(r[2 + i].lower().startswith("active") for r in self.rows)]
        return {"sm_mhz": sm[len(sm) // 2], "sm_max_mhz": float(self.rows[0][1]), "reasons": reasons,
                "samples": len(self.rows)}
