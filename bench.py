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
HOST_PLACEMENT = None
# algorithmic work per 10 s utterance
FLOP_CONVAE_UTT = 22.28e9          # 15.02 GFLOP autoencoder + 7.26 GFLOP TDNN classifier
BYTES_FBANK_UTT = 960320           # 640000 B PCM in + 320320 B features out
FLOP_FCAE_UTT = 36.8e6


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=40)
    ap.add_argument("--ramp-seconds", type=float, default=0.5, help="untimed pre-warm before the warm-up steps")
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--sustained-steps", type=int, default=300, help="extra replays after the timed region, reported as `sustained`")
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default="convae", choices=["convae", "fcae", "train"])
    ap.add_argument("--train-batch", type=int, default=8, help="utterances per GPU and training step (BASELINE configs[3])")
    ap.add_argument("--batch", type=int, default=60)
    ap.add_argument("--pool", type=int, default=360)
    ap.add_argument("--precision", default="fp16", choices=["fp16", "fp32"])
    ap.add_argument("--cpu-sample", type=int, default=16, help="utterances in the CPU baseline sample")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-graphs", action="store_true", help="launch every kernel of a step from the host instead of replaying a CUDA graph")
    ap.add_argument("--no-extras", action="store_true", help="skip the timed scope-row-8(f) extras (ConvReconstruction, x-vector)")
    ap.add_argument("--chunk", type=int, default=0, help="utterances per L2-resident chunk of the conv stack (0 = library default)")
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
        # NVML is initialised HERE, on the caller's thread and before any timed region: nvmlInit takes driver locks that
        # kernel launches also need (an init inside the timed loop showed up as a one-off 15-75 ms stall of the launch loop)
        self.nv = self._nvml_handle()

    def _nvml_handle(self):
        """NVML in-process (no fork: spawning nvidia-smi from a process with a CUDA context can stall every thread for
        milliseconds, which starves the launch loop being timed).  Returns None if pynvml is unusable."""
        try:
            import pynvml
            pynvml.nvmlInit()
            try:
                uuid = str(torch.cuda.get_device_properties(self.index).uuid)
                h = pynvml.nvmlDeviceGetHandleByUUID(("GPU-" + uuid).encode())
            except Exception:
                h = pynvml.nvmlDeviceGetHandleByIndex(self.index)
            return pynvml, h
        except Exception:
            return None

    def run(self):
        nv = self.nv
        while nv is not None and not self.stop_flag:
            try:
                pynvml, h = nv
                sm = pynvml.nvmlDeviceGetClockInfo(h, pynvml.NVML_CLOCK_SM)
                mx = pynvml.nvmlDeviceGetMaxClockInfo(h, pynvml.NVML_CLOCK_SM)
                r = pynvml.nvmlDeviceGetCurrentClocksThrottleReasons(h)
                act = lambda bit: "Active" if r & bit else "Not Active"   # noqa: E731
                self.rows.append([str(sm), str(mx), act(0x8), act(0x40), act(0x20), act(0x4)])
            except Exception:
                nv = None
                break
            time.sleep(0.025)
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
                "samples": len(self.rows), "sm_mhz_min": sm[0]}


# --------------------------------------------------------------------------- helpers
# per-utterance algorithmic flops of each conv launch (SURVEY.md 8a / DESIGN.md "Kernels"), L = 80640
def convae_layer_flops(L=80640):
    mac = {"first_layer_kernel": L * 15 * 32,
           "tapgemm:enc2": (L // 2) * 32 * 5 * 64,
           # encoder.0 generated inside encoder.2's launch: counted at the flops of the two reference layers
           "tapgemm:enc02": L * 15 * 32 + (L // 2) * 32 * 5 * 64, "tapgemm:enc5": (L // 2) * 64 * 5 * 64,
           "tapgemm:enc8": (L // 4) * 64 * 5 * 128, "tapgemm:enc11": (L // 4) * 128 * 5 * 128,
           "tapgemm:tdnn0": (L // 4 - 4) * 128 * 5 * 128, "tapgemm:tdnn3": (L // 4 - 8) * 128 * 3 * 128,
           "tapgemm:tdnn6": (L // 4 - 14) * 128 * 3 * 128,
           # the three TDNN convolutions as one dataflow launch
           "tapgemm:tdnn036": (L // 4 - 4) * 128 * 5 * 128 + (L // 4 - 8) * 128 * 3 * 128 + (L // 4 - 14) * 128 * 3 * 128,
           "tapgemm:dec0": (L // 4) * 128 * 5 * 128, "tapgemm:dec1": (L // 4) * 128 * 5 * 64,
           "tapgemm:dec4": (L // 2) * 64 * 5 * 64, "tapgemm:dec5": (L // 2) * 64 * 5 * 32,
           # decoder.4 o decoder.5 run as one composed layer: counted at the ALGORITHMIC flops of the two reference layers
           "tapgemm:dec45": (L // 2) * 64 * 5 * 64 + (L // 2) * 64 * 5 * 32,
           "tapgemm:dec8": L * 32 * 15}
    return {k: 2.0 * v for k, v in mac.items()}


def bind_rank_to_cores(local, n_local):
    """Host-side placement for the end-to-end arm: every rank streams tens of MB per step through pinned host buffers, so
    the rank's thread (and, by first touch, its pinned pages) go onto a core set of its own inside the GPU's NUMA-local
    CPU set instead of all ranks sharing (and migrating over) one affinity mask.  Returns a description or None."""
    try:
        allowed = sorted(os.sched_getaffinity(0))
        near = allowed
        try:
            import pynvml
            pynvml.nvmlInit()
            h = pynvml.nvmlDeviceGetHandleByIndex(local)
            words = pynvml.nvmlDeviceGetCpuAffinity(h, (max(allowed) // 64) + 1)
            mask = [c for c in allowed if (words[c // 64] >> (c % 64)) & 1]
            if len(mask) >= n_local:
                near = mask
        except Exception:
            pass
        per = max(1, len(near) // max(n_local, 1))
        mine = near[local * per:(local + 1) * per] or near
        os.sched_setaffinity(0, set(mine))
        return {"cores": [mine[0], mine[-1]], "numa_local_set": len(near), "allowed": len(allowed)}
    except Exception as e:      # (not fatal: placement is an optimisation)
        return {"error": str(e)[:80]}


def dist_setup(n_gpus):
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    global HOST_PLACEMENT
    HOST_PLACEMENT = bind_rank_to_cores(local, int(os.environ.get("LOCAL_WORLD_SIZE", str(world)))) if world > 1 else None
    if world > 1:
        import torch.distributed as dist
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        os.environ.setdefault("MASTER_PORT", "29533")
        torch.cuda.set_device(local)
        # NCCL prints its version banner on stdout when the first communicator comes up; the contract is ONE JSON line
        # on stdout, so fd 1 points at stderr until the communicator exists
        sys.stdout.flush()
        saved = os.dup(1)
        os.dup2(2, 1)
        try:
            dist.init_process_group("nccl", device_id=torch.device("cuda", local))
            dist.barrier()
            torch.cuda.synchronize()
        finally:
            sys.stdout.flush()
            os.dup2(saved, 1)
            os.close(saved)
    elif n_gpus > 1:
        raise SystemExit("--gpus %d needs torchrun (one process per GPU)" % n_gpus)
    return world, rank, local


def barrier(world):
    if world > 1:
        import torch.distributed as dist
        dist.barrier()


def max_over_ranks(v, world, device):
    if world == 1:
        return v
    import torch.distributed as dist
    t = torch.tensor([v], dtype=torch.float64, device=device)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item())


def build_pipeline(workload, precision, device):
    import recipe
    import speech_anonymization_b200 as sab
    mean, std, count = real_norm_stats()
    fb = sab.Fbank(n_mels=80)
    norm = sab.InputNormalization(norm_type="global", update_until_epoch=4)
    norm._load_statistics_dict({"count": count, "glob_mean": torch.from_numpy(mean), "glob_std": torch.from_numpy(std),
                                "spk_dict_mean": {}, "spk_dict_std": {}, "spk_dict_count": {}})
    norm = norm.to(device).eval()
    if workload == "convae":
        model, shapes = sab.ConvAutoencoder(), recipe.CONVAE_SHAPES
        model.precision = precision
    else:
        model, shapes = sab.FullyConnectedAutoencoder(80, 64), recipe.FCAE_SHAPES
    w = recipe.make_weights(shapes, 8886)   # random-init weights, seed 8886 (convae.yaml:11)
    model.load_state_dict({k: torch.from_numpy(np.asarray(v)) for k, v in w.items()})
    model = model.to(device).eval()
    return sab.AnonymizerPipeline(fb, norm, model)


# --------------------------------------------------------------------------- our arm
def run_ours(a):
    import recipe
    import speech_anonymization_b200 as sab
    from speech_anonymization_b200 import _lib
    world, rank, local = dist_setup(a.gpus)
    if not torch.cuda.is_available():
        raise SystemExit("bench.py (impl=ours) needs a CUDA device: there is no CPU fallback")
    dev = torch.device("cuda", local)
    torch.cuda.set_device(dev)
    pk = peaks()
    pipe = build_pipeline(a.workload, a.precision, dev)
    if a.workload == "convae":
        pipe.model.chunk_utts = a.chunk
    B, pool_n = a.batch, max(a.pool, a.batch)
    n_batches = pool_n // B
    host_pool = make_wavs(n_batches * B, 1234 + rank).pin_memory()       # SURVEY 8(d) synthetic input
    dev_pool = host_pool.to(dev)
    nz = 128 if a.workload == "convae" else 40
    noise = torch.from_numpy(recipe.make_noise(3, B, nz)).to(dev)
    T = 1 + N_SAMPLES // 160
    Tp = pipe.padded_frames(T)

    # One CUDA graph per pool slice (AnonymizerPipeline.capture): a step is then ONE host call replaying the library's
    # whole launch sequence, so host-side hiccups (a descheduled launch thread showed up as single 20-90 ms steps on a
    # fresh box) cannot starve the GPU between kernels.  --no-graphs launches the ~29 kernels per step directly.
    graphs = None
    if not a.no_graphs:
        graphs = [pipe.capture(dev_pool[k * B:(k + 1) * B], pool_noise=noise) for k in range(n_batches)]

    def step_device(i):
        k = i % n_batches
        if graphs is not None:
            return graphs[k].replay()
        return pipe(dev_pool[k * B:(k + 1) * B], pool_noise=noise)

    # clock ramp: a fresh box idles at low clocks and the first tens of milliseconds of work run slow; run untimed
    # steps for a fixed wall-clock budget before the W warm-up steps so that neither arm is timed during the ramp
    sampler = ClockSampler(local)        # (initialises NVML now, outside every timed region)
    t_ramp = time.perf_counter()
    i = 0
    while time.perf_counter() - t_ramp < a.ramp_seconds:
        for _ in range(8):
            step_device(i)
            i += 1
        torch.cuda.synchronize()
    sampler.start()                      # (thread start-up and its first NVML calls happen during the warm-up)
    launches0 = _lib.lib().spanon_launch_count()
    for i in range(a.warmup):
        step_device(i)
    torch.cuda.synchronize()
    launches_w = _lib.lib().spanon_launch_count()
    barrier(world)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    import gc
    gc.collect()
    gc.disable()    # a generation-2 collection inside the launch loop is a 20 ms host stall = one starved step
    step_ev = [torch.cuda.Event(enable_timing=True) for _ in range(a.steps)]   # diagnostics only: per-step spread
    e0.record()
    t_host = time.perf_counter()
    for i in range(a.steps):
        out = step_device(a.warmup + i)
        step_ev[i].record()
    host_enqueue_ms = (time.perf_counter() - t_host) * 1e3 / a.steps     # host time to enqueue one step (no sync)
    e1.record()
    torch.cuda.synchronize()
    ms_total = max_over_ranks(e0.elapsed_time(e1), world, dev)
    per_step = sorted(a_.elapsed_time(b_) for a_, b_ in zip([e0] + step_ev[:-1], step_ev))
    step_spread = {"min": per_step[0], "median": per_step[len(per_step) // 2], "max": per_step[-1]}
    barrier(world)
    gpu_launches = int(_lib.lib().spanon_launch_count() - launches_w)
    if graphs is not None:     # replays do not pass through the library's launch counter: kernels per captured step x K
        gpu_launches = sum(graphs[(a.warmup + i) % n_batches].launches for i in range(a.steps))
    ms_step = ms_total / a.steps
    value = world * B * UTT_SECONDS / (ms_step * 1e-3)
    # ---- e2e: host PCM in (pinned), host features + log-probs out, through the public host-buffer API.
    # Streaming service form (spanon_pipe_*): every step copies its own 38.4 MB of PCM H2D and its own features D2H;
    # two batches are in flight so those copies overlap the kernels of the neighbouring step. The one-shot
    # synchronous call (spanon_anonymize_host) is timed next to it.
    depth = 2
    logp_h = [torch.empty((B, 2), dtype=torch.float32).pin_memory() for _ in range(depth + 1)]

    def time_pipe(pool_h, out_dtype):
        """Streaming host-buffer arm: `pool_h` pinned PCM (fp32 or int16), reconstruction returned as `out_dtype`."""
        rec_h = [torch.empty((B, Tp, 80), dtype=out_dtype).pin_memory() for _ in range(depth + 1)]
        if a.workload == "convae":
            pipe.model.recon_fp16 = out_dtype == torch.float16
        sp = pipe.stream_pipe(B, N_SAMPLES, dev, depth=depth)
        for i in range(a.warmup):
            sp.submit(pool_h[(i % n_batches) * B:(i % n_batches + 1) * B], rec_h[i % (depth + 1)], logp_h[i % (depth + 1)],
                      pool_noise=noise)
        sp.drain()
        barrier(world)
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        for i in range(a.steps):
            k = (a.warmup + i) % n_batches
            sp.submit(pool_h[k * B:(k + 1) * B], rec_h[i % (depth + 1)], logp_h[i % (depth + 1)], pool_noise=noise)
        sp.drain()
        ms = max_over_ranks((time.perf_counter() - t0) * 1e3, world, dev) / a.steps
        sp.close()
        if a.workload == "convae":
            pipe.model.recon_fp16 = False
        return ms, rec_h

    # headline e2e: the audio file's own sample format in (int16 PCM, converted by the Fbank kernel exactly like the
    # reference's read_audio does on the host) and, for the 16-bit ConvAutoencoder mode, its fp16 reconstruction out;
    # the fp32-in / fp32-out wire of the reference's tensors is timed next to it
    host_pool16 = torch.clamp(torch.round(host_pool * 32768.0), -32768, 32767).to(torch.int16).pin_memory()
    out16 = a.workload == "convae" and a.precision == "fp16"
    ms_e2e, rec_h = time_pipe(host_pool16, torch.float16 if out16 else torch.float32)
    ms_e2e_f32, rec_h = time_pipe(host_pool, torch.float32)
    ws = pipe.host_workspace(B, N_SAMPLES, dev)
    pipe.anonymize_host(host_pool[:B], rec_h[0], logp_h[0], ws, pool_noise=noise)
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    n_sync = max(2, a.steps // 2)
    for i in range(n_sync):
        k = (a.warmup + i) % n_batches
        pipe.anonymize_host(host_pool[k * B:(k + 1) * B], rec_h[0], logp_h[0], ws, pool_noise=noise)
    ms_sync = max_over_ranks((time.perf_counter() - t0) * 1e3, world, dev) / n_sync
    sampler.stop_flag = True
    sampler.join(timeout=2)
    sustained = None
    gc.enable()
    e2e = {"value": world * B * UTT_SECONDS / (ms_e2e * 1e-3), "unit": "audio-s/s", "ms_per_step": ms_e2e,
           "h2d_bytes_per_step": B * N_SAMPLES * 2, "d2h_bytes_per_step": B * Tp * 80 * (2 if out16 else 4) + B * 8,
           "wire": "int16 PCM in, %s features out" % ("fp16" if out16 else "fp32"),
           "api": "spanon_pipe_submit_i16/wait (pinned host buffers in/out, %d batches in flight over H2D|kernel|D2H streams)" % depth,
           "timer": "host wall clock between synchronisation points, max over ranks",
           "host_placement": HOST_PLACEMENT,
           "fp32_wire": {"api": "spanon_pipe_submit (fp32 PCM in, fp32 features out)", "ms_per_step": ms_e2e_f32,
                         "value": world * B * UTT_SECONDS / (ms_e2e_f32 * 1e-3), "h2d_bytes_per_step": B * N_SAMPLES * 4,
                         "d2h_bytes_per_step": B * Tp * 80 * 4 + B * 8},
           "one_shot_sync_call": {"api": "spanon_anonymize_host", "ms_per_step": ms_sync,
                                  "value": world * B * UTT_SECONDS / (ms_sync * 1e-3)}}

    # ---- per-kernel event timing (consecutive CUDA events on the launch stream, 2 extra steps)
    for i in range(8):          # put the host ahead of the GPU first, so that no interval contains host latency
        k = (a.warmup + a.steps + i) % n_batches
        pipe(dev_pool[k * B:(k + 1) * B], pool_noise=noise)
    _lib.profile_begin(dev)
    n_prof = 2
    for i in range(n_prof):     # direct launches: the library records its profile events at launch time
        k = (a.warmup + a.steps + i) % n_batches
        pipe(dev_pool[k * B:(k + 1) * B], pool_noise=noise)
    prof = _lib.profile_end()
    # ---- sustained: the device-resident step for another --sustained-steps replays, reported beside `value`, never instead
    # of it -- and run AFTER the per-kernel profile, so that the profile sees the clocks of the timed region.  Under the
    # board's power cap the SM clock sags from ~1.9 GHz in the first tens of milliseconds of load to ~1.55 GHz; a K-step
    # timed region of 40 - 80 ms sees mostly the former, a long-running service the latter.
    if a.sustained_steps > 0:
        sampler2 = ClockSampler(local)
        sampler2.start()
        s0, s1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        s0.record()
        for i in range(a.sustained_steps):
            step_device(a.warmup + a.steps + i)
        s1.record()
        torch.cuda.synchronize()
        sampler2.stop_flag = True
        sampler2.join(timeout=2)
        ms_s = max_over_ranks(s0.elapsed_time(s1), world, dev) / a.sustained_steps
        sustained = {"steps": a.sustained_steps, "ms_per_step": ms_s, "value": world * B * UTT_SECONDS / (ms_s * 1e-3),
                     "clocks": sampler2.summary()}
        barrier(world)
    kern = {k: {"launches_per_step": n / n_prof, "ms_per_step": ms / n_prof} for k, (n, ms) in prof.items()}
    prof_ms = sum(v["ms_per_step"] for v in kern.values())
    if a.workload == "convae":
        fl = convae_layer_flops(Tp * 80)
        conv_ms = sum(v["ms_per_step"] for k, v in kern.items() if k in fl)
        conv_launches = sum(v["launches_per_step"] for k, v in kern.items() if k in fl)
        # algorithmic flops of the launches that RAN (a fused launch is listed at the sum of the reference layers it
        # replaces, so the layers it replaced must not be counted again)
        flops_step = sum(fl[k] for k in kern if k in fl) * B
        if Tp * 80 == 80640:     # 22.28 GFLOP per 10 s utterance
            assert abs(flops_step / B - FLOP_CONVAE_UTT) < 2e7, flops_step / B
        achieved = flops_step / (conv_ms * 1e-3) / 1e12
        peak = pk["tf_sust"]
        traffic, tsrc = None, None       # DRAM bytes per launch from the committed ncu --set full capture (batch 60)
        tp = next((q for q in (os.path.join(ROOT, "profiles", n) for n in ("r02_traffic.json", "r01_traffic.json"))
                   if os.path.exists(q)), None)
        if tp:
            tj = json.load(open(tp))
            traffic = tj["conv_stack"]["traffic_mb_per_launch"] * 1e6 * B / 60.0
            tsrc = "profiles/" + os.path.basename(tp) + " (ncu dram__bytes_read.sum + dram__bytes_write.sum, avg over the %d conv launches of a step)" % tj["conv_stack"].get("launches_captured", 0)
        ms_launch = conv_ms / max(conv_launches, 1)
        # whole driver-timed step against the same peak (Fbank, pooling, classifier head and launch gaps included)
        step_tf = flops_step / (ms_step * 1e-3) / 1e12
        # algorithmic minimum bytes of the fused path = PCM in + padded features out (+ log-probs)
        alg_bytes_step = B * (N_SAMPLES * 4 + Tp * 80 * 4 + 8)
        roof = {"bound": "tensor", "achieved": achieved, "peak": peak, "unit": "TFLOP/s", "frac": achieved / peak,
                "step_achieved": step_tf, "step_frac": step_tf / peak, "step_frac_of_burst": step_tf / pk["tf_burst"],
                "algorithmic_flops_per_step": flops_step, "algorithmic_bytes_per_step": alg_bytes_step,
                "traffic": traffic, "traffic_source": tsrc, "kernel": "conv stack (first_layer + tap-GEMM launches)",
                "traffic_per_step": None if traffic is None else traffic * conv_launches,
                "traffic_over_algorithmic_bytes": None if traffic is None else traffic * conv_launches / alg_bytes_step,
                "flops_per_launch": flops_step / max(conv_launches, 1), "ms_per_launch": ms_launch,
                "share_of_step": conv_ms / prof_ms, "peak_source": pk["src"] + " bf16 sustained",
                # the fp16 planes of a 60-utterance chunk do not fit L2, so every launch also streams them through HBM
                "hbm_cross_check": None if traffic is None else {
                    "GB/s": traffic / (ms_launch * 1e-3) / 1e9, "peak": pk["hbm"],
                    "frac": traffic / (ms_launch * 1e-3) / 1e9 / pk["hbm"]}}
    else:
        km = kern.get("fbank400_kernel", {"ms_per_step": float("nan"), "launches_per_step": 1})
        achieved = BYTES_FBANK_UTT * B / (km["ms_per_step"] / km["launches_per_step"] * 1e-3) / 1e9
        roof = {"bound": "hbm", "achieved": achieved, "peak": pk["hbm"], "unit": "GB/s", "frac": achieved / pk["hbm"],
                "traffic": None, "kernel": "fbank400_kernel", "share_of_step": km["ms_per_step"] / prof_ms,
                "peak_source": pk["src"]}
    fb = kern.get("fbank400_kernel")
    stages = {}
    if fb:
        gbs = BYTES_FBANK_UTT * B / (fb["ms_per_step"] * 1e-3) / 1e9
        stages["fbank"] = {"ms_per_step": fb["ms_per_step"], "GB/s": gbs, "frac_hbm": gbs / pk["hbm"]}

    # ---- scope rows 8(f) built so far, timed next to the headline path (not part of `value`): the "endtoend"
    #      anonymizer and the external x-vector gender classifier that evaluates anonymized features (config 3)
    extras = None
    if rank == 0 and world == 1 and a.workload == "convae" and not a.no_extras:
        extras = time_extras(sab, _lib, recipe, dev, B, T)
        extras["library_gpu_convae_fwd"] = time_library_gpu(recipe, dev, B, T)
        extras.update(time_config_lines(sab, _lib, recipe, dev, dev_pool, pk))
        extras["config3_train_step_b8"] = time_train_steps(sab, recipe, dev, 1, 8, 5, 3, host_pool)

    cpu = None
    if rank == 0 and world == 1 and not a.no_cpu_baseline:
        v, dt, cores = time_cpu(a.workload, a.cpu_sample, 8, 1)
        cpu = {"value": v, "unit": "audio-s/s", "cores": cores, "kind": "port",
               "sample": "%d x 10 s utterances per pass, 1 warm-up + 8 timed passes (%.2f s/pass, ~10 s of CPU work), oracle/ port of the "
                         "reference's PyTorch-CPU path" % (a.cpu_sample, dt)}
    if rank == 0:
        line = {"metric": "audio-sec/sec anonymized (Fbank+%s fwd)" % ("ConvAE" if a.workload == "convae" else "FC-AE"),
                "value": value, "unit": "audio-s/s", "n_gpus": world, "steps": a.steps, "warmup": a.warmup,
                "ms_per_step": ms_step, "host_enqueue_ms_per_step": host_enqueue_ms, "step_ms_spread": step_spread,
                "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
                "dtype": a.precision if a.workload == "convae" else "fp32", "data": "synthetic",
                "config": workload_config(a, world),
                "clocks": sampler.summary(), "sustained": sustained, "e2e": e2e, "gpu_launches": gpu_launches, "roofline": roof,
                "cpu_baseline": cpu, "stages": stages, "kernels": kern, "extras": extras}
        print(json.dumps(line))
    if world > 1:
        import torch.distributed as dist
        dist.destroy_process_group()


def time_extras(sab, _lib, recipe, dev, B, T, iters=6):
    """ConvReconstruction forward and x-vector classification of B x T x 80 feature batches (device-resident)."""
    def load(m, shapes, seed):
        w = recipe.make_weights(shapes, seed)
        m.load_state_dict({k: torch.from_numpy(np.asarray(v)) for k, v in w.items()})
        return m.to(dev).eval()

    feats = [torch.from_numpy(recipe.make_feats(50 + i, B, T)).to(dev) for i in range(3)]   # 3 x 19 MB inputs
    cr = load(sab.ConvReconstruction(), recipe.CONVRECON_SHAPES, 8886)
    xv = sab.XvectorClassifier()
    load(xv.embedding_model, recipe.XVECTOR_EMB_SHAPES, 1230)
    load(xv.classifier, recipe.XVECTOR_CLF_SHAPES, 1230)
    xv = xv.to(dev).eval()
    noise = torch.from_numpy(recipe.make_noise(9, B, 1500)).to(dev)
    out = {}
    for name, fn in (("convrecon_fwd", lambda x: cr(x)), ("xvector_classify", lambda x: xv.classify_batch_feats(x, pool_noise=noise))):
        with torch.no_grad():
            for i in range(3):
                fn(feats[i % 3])
            torch.cuda.synchronize()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            for i in range(iters):
                fn(feats[i % 3])
            e1.record()
            torch.cuda.synchronize()
            ms = e0.elapsed_time(e1) / iters
            _lib.profile_begin(dev)
            fn(feats[0])
            prof = _lib.profile_end()
        out[name] = {"ms_per_batch": ms, "batch": B, "frames": T, "audio_s_per_s": B * UTT_SECONDS / (ms * 1e-3),
                     "kernels_ms": {k: v[1] for k, v in prof.items()}}
    # x-vector TDNN: 2 * T * sum(k * cin * cout) flops per utterance (Xvector.py:60-79 with evaluator_inference.yaml)
    macs = T * (5 * 80 * 512 + 3 * 512 * 512 * 2 + 512 * 512 + 512 * 1500)
    g = out["xvector_classify"]
    gemm_ms = sum(v for k, v in g["kernels_ms"].items() if k.startswith("xv_gemm"))
    g["tdnn_gflop_per_utt"] = 2 * macs / 1e9
    g["tdnn_tflops"] = 2 * macs * B / (gemm_ms * 1e-3) / 1e12 if gemm_ms > 0 else None
    return out


def _time_fn(fn, iters, warm=2):
    with torch.no_grad():
        for i in range(warm):
            fn(i)
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for i in range(iters):
            fn(warm + i)
        e1.record()
        torch.cuda.synchronize()
    return e0.elapsed_time(e1) / iters


def time_config_lines(sab, _lib, recipe, dev, dev_pool, pk):
    """The other lines SURVEY 8(d) / BASELINE.json name, device-resident, under `extras` (not part of `value`):
    the fp32 parity mode of the headline path, configs[1] (Fbank + FullyConnected, batch 64) and configs[2]
    (Fbank + ConvAutoencoder on 256 x 10 s, anonymized features classified by the external x-vector model)."""
    out = {}
    n_pool = dev_pool.shape[0]
    # -- fp32 parity mode (<= 1e-4): CUDA-core FMA convolutions, reported separately, never against the bf16 peak
    B = 60
    p32 = build_pipeline("convae", "fp32", dev)
    ms = _time_fn(lambda i: p32(dev_pool[(i * B) % (n_pool - B + 1):(i * B) % (n_pool - B + 1) + B]), 3, warm=1)
    out["convae_fp32_parity_mode"] = {"ms_per_step": ms, "batch": B, "audio_s_per_s": B * UTT_SECONDS / (ms * 1e-3),
                                      "dtype": "fp32", "tflops_fp32": FLOP_CONVAE_UTT * B / (ms * 1e-3) / 1e12,
                                      "what": "Fbank + ConvAutoencoder forward, fp32 activations and CUDA-core FMA "
                                              "(the <= 1e-4 parity mode); device-resident PCM pool"}
    del p32
    # -- configs[1]: FullyConnected anonymizer over Fbank features, batch 64 x 10 s (HBM-bound: 960 320 B per utterance)
    B = 64
    pfc = build_pipeline("fcae", "fp32", dev)
    nb = n_pool // B
    _lib.lib()
    ms = _time_fn(lambda i: pfc(dev_pool[(i % nb) * B:(i % nb + 1) * B]), 20, warm=3)
    _lib.profile_begin(dev)
    with torch.no_grad():
        pfc(dev_pool[:B])
    prof = _lib.profile_end()
    gbs = BYTES_FBANK_UTT * B / (ms * 1e-3) / 1e9
    out["config1_fbank_fcae_b64"] = {"ms_per_step": ms, "batch": B, "audio_s_per_s": B * UTT_SECONDS / (ms * 1e-3),
                                     "dtype": "fp32", "algorithmic_GB/s": gbs, "frac_hbm": gbs / pk["hbm"],
                                     "kernels_ms": {k: v[1] for k, v in prof.items()}}
    del pfc
    # -- configs[2]: ConvAutoencoder + external gender-classifier evaluation on 256 x 10 s utterances
    B = 256
    if n_pool >= B:
        pca = build_pipeline("convae", "fp16", dev)
        xv = sab.XvectorClassifier()
        for m, shapes in ((xv.embedding_model, recipe.XVECTOR_EMB_SHAPES), (xv.classifier, recipe.XVECTOR_CLF_SHAPES)):
            m.load_state_dict({k: torch.from_numpy(np.asarray(v)) for k, v in recipe.make_weights(shapes, 1230).items()})
        xv = xv.to(dev).eval()
        T = 1 + N_SAMPLES // 160
        noise = torch.from_numpy(recipe.make_noise(5, B, 128)).to(dev)
        xnoise = torch.from_numpy(recipe.make_noise(6, B, 1500)).to(dev)

        def cfg2(i):
            o = (i * 52) % (n_pool - B + 1)
            rec, logp = pca(dev_pool[o:o + B], pool_noise=noise)
            return xv.classify_batch_feats(rec[:, :T], pool_noise=xnoise)
        ms = _time_fn(cfg2, 4, warm=2)
        out["config2_convae_xvector_b256"] = {"ms_per_step": ms, "batch": B, "audio_s_per_s": B * UTT_SECONDS / (ms * 1e-3),
                                              "dtype": "fp16",
                                              "what": "Fbank -> ConvAutoencoder (recon + sex log-probs) -> x-vector "
                                                      "classify_batch_feats on the anonymized features, 256 x 10 s"}
        del pca, xv
    torch.cuda.empty_cache()
    return out


def time_library_gpu(recipe, dev, B, T, iters=4):
    """the library-kernel baseline -- the same ConvAutoencoder + classifier forward written with plain
    torch ops (cuDNN / ATen kernels, fp16 weights and activations) on the same device-resident feature batch.  None of
    this repo's kernels run here; it is reported under `extras`, it is not a bench value."""
    import torch.nn.functional as F
    w = {k: torch.from_numpy(np.asarray(v)).to(dev).half() for k, v in recipe.make_weights(recipe.CONVAE_SHAPES, 8886).items()}
    feats = torch.from_numpy(recipe.make_feats(50, B, T)).to(dev).half()
    noise = torch.from_numpy(recipe.make_noise(9, B, 128)).to(dev).half()
    act = lambda x: x * torch.sigmoid(x)
    inorm = lambda x, p: F.instance_norm(x, weight=w[p + ".weight"], bias=w[p + ".bias"], eps=1e-5)
    bnorm = lambda x, p: F.batch_norm(x, w[p + ".running_mean"], w[p + ".running_var"], w[p + ".weight"], w[p + ".bias"], False, 0.0, 1e-5)
    conv = lambda x, p, **kw: F.conv1d(x, w[p + ".weight"], w[p + ".bias"], **kw)
    convt = lambda x, p: F.conv_transpose1d(x, w[p + ".weight"], w[p + ".bias"], stride=2, padding=2, output_padding=1)

    def fwd(f):
        x = f.reshape(B, 1, -1)
        x = act(conv(x, "encoder.0", padding=7))
        x = act(inorm(conv(x, "encoder.2", stride=2, padding=2), "encoder.3"))
        x = act(inorm(conv(x, "encoder.5", padding=2), "encoder.6"))
        x = act(inorm(conv(x, "encoder.8", stride=2, padding=2), "encoder.9"))
        enc = act(inorm(conv(x, "encoder.11", padding=2), "encoder.12"))
        c = bnorm(enc, "sex_classifier.norm")
        for i, dil in ((0, 1), (3, 2), (6, 3)):
            c = bnorm(F.relu(conv(c, "sex_classifier.tdnn.%d" % i, dilation=dil)), "sex_classifier.tdnn.%d" % (i + 2))
        c = c.reshape(B, c.shape[2], c.shape[1]).float()
        st = torch.cat((c.mean(dim=1) + noise.float(), c.std(dim=1) + 1e-5), dim=1).half()
        st = bnorm(F.relu(F.linear(st, w["sex_classifier.classify.0.weight"], w["sex_classifier.classify.0.bias"])), "sex_classifier.classify.2")
        st = bnorm(F.relu(F.linear(st, w["sex_classifier.classify.3.weight"], w["sex_classifier.classify.3.bias"])), "sex_classifier.classify.5")
        logp = F.log_softmax(F.linear(st, w["sex_classifier.classify.6.weight"], w["sex_classifier.classify.6.bias"]).float(), 1)
        x = convt(conv(enc, "decoder.0", padding=2), "decoder.1")
        x = act(inorm(x, "decoder.2"))
        x = convt(conv(x, "decoder.4", padding=2), "decoder.5")
        x = act(inorm(x, "decoder.6"))
        return conv(x, "decoder.8", padding=7).reshape(B, T, -1), logp

    try:
        with torch.no_grad():
            for _ in range(2):
                fwd(feats)
            torch.cuda.synchronize()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            for _ in range(iters):
                fwd(feats)
            e1.record()
            torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / iters
        out = {"ms_per_batch": ms, "batch": B, "frames": T, "audio_s_per_s": B * UTT_SECONDS / (ms * 1e-3),
               "what": "ConvAutoencoder + classifier forward on device-resident features with torch ops "
                       "(cuDNN/ATen, fp16), no Fbank/normalisation; this repo's kernels are not involved"}
    except Exception as e:   # (cuDNN may refuse the 80 640-sample 1-D problem or run out of workspace)
        out = {"unavailable": "%s: %s" % (type(e).__name__, str(e)[:200])}
    del w, feats
    torch.cuda.empty_cache()
    return out


def workload_config(a, world):
    """The workload both arms are quoted on (BASELINE.json configs[4], the throughput sweep)."""
    B = a.batch
    n_batches = max(1, a.pool // B)
    return {"workload": "fbank+%s, %d x 10 s @16 kHz utterances per GPU per step from a resident pool of "
                        "%d utterances (1 h) per GPU" % (a.workload, B, n_batches * B),
            "batch_per_gpu": B, "pool_per_gpu": n_batches * B, "parallelism": "utterance-sharded x%d" % world,
            "l2": "inputs cycle through a %.0f MB pool (> 126 MB L2)" % (n_batches * B * N_SAMPLES * 4 / 1e6),
            "weights": "random-init seed 8886", "norm": "global stats from the reference's normalizer.ckpt",
            "launch": "direct kernel launches" if getattr(a, "no_graphs", False) else "one CUDA graph replay per step (device arm)"}


# --------------------------------------------------------------------------- training step (BASELINE configs[3])
def time_train_steps(sab, recipe, dev, world, B, steps, warmup, host_pool):
    """One step = speechbrain_convae_train.py:195-259 on B x 10 s utterances per GPU: Fbank -> normalise -> pad ->
    ConvAutoencoder in TRAIN mode -> recon (MSE) + sex (NLL) + cosine-similarity losses, backward on the training kernels,
    flat gradient all-reduce (+ SyncBatchNorm statistics) over NCCL, Adam step; the kNN-MI statistic of the batch's
    (gathered) sex log-probabilities is computed alongside (the reference's MI term is not differentiable)."""
    pipe = build_pipeline("convae", "fp32", dev)
    model = pipe.model.train()
    dp = sab.DataParallelTrainer(model)
    dp.prepare()
    opt = torch.optim.Adam(model.parameters(), lr=1e-4)
    cos = sab.CosineSimilarityLoss()
    n_b = host_pool.shape[0] // B
    pool = host_pool[:n_b * B].to(dev)
    T = 1 + N_SAMPLES // 160
    lens = torch.ones(B, device=dev)
    noise = torch.from_numpy(recipe.make_noise(11, B, 128)).to(dev)
    y = (torch.arange(B, device=dev) % 2).long()
    ev = []

    def step(i):
        wav = pool[(i % n_b) * B:(i % n_b + 1) * B]
        with torch.no_grad():                                     # compute_forward :58-63 (features carry no graph)
            feats = sab.pad_time(pipe.normalize(pipe.compute_features(wav), lens))
        rec, logp = model(feats, pool_noise=noise)
        loss = (torch.nn.functional.mse_loss(rec.view(B, -1), feats.view(B, -1)) + 0.8 * torch.nn.functional.nll_loss(logp, y)
                + 0.2 * cos(rec, feats) / B)
        e0, e1, e2 = (torch.cuda.Event(enable_timing=True) for _ in range(3))
        loss.backward()
        e0.record()
        dp.finish_backward()                                      # waits for the early bucket, all-reduces the encoder's
        e1.record()
        torch.nn.utils.clip_grad_norm_(model.parameters(), 5.0)   # core.py check_gradients
        opt.step()
        dp.zero_grad()
        Xg, yg = sab.gather_for_mi(logp.detach(), y)              # MI statistic of the global batch
        mi = sab.knn_mi(Xg, yg, k=3)[0] if Xg.shape[0] >= 8 else None
        e2.record()
        ev.append((e0, e1))
        return loss, mi

    for i in range(warmup):
        step(i)
    torch.cuda.synchronize()
    barrier(world)
    torch.cuda.synchronize()
    ev.clear()
    t0, t1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    t0.record()
    for i in range(steps):
        loss, mi = step(warmup + i)
    t1.record()
    torch.cuda.synchronize()
    ms = max_over_ranks(t0.elapsed_time(t1), world, dev) / steps
    comm = sum(a_.elapsed_time(b_) for a_, b_ in ev) / steps
    return {"ms_per_step": ms, "batch_per_gpu": B, "audio_s_per_s": world * B * UTT_SECONDS / (ms * 1e-3),
            "exposed_allreduce_ms_per_step": comm, "allreduce_share": comm / ms, "gradient_bytes": dp.grads.n_total * 4,
            "early_bucket_bytes": dp.grads.n_early * 4, "final_loss": float(loss),
            "dtype": "fp32", "what": "Fbank + normalise + pad + ConvAutoencoder TRAIN forward/backward (csrc/train.cu), recon + sex + "
                                     "cosine losses, flat-gradient all-reduce + SyncBatchNorm over NCCL, clip, Adam"}


def run_train(a):
    import recipe
    import speech_anonymization_b200 as sab
    world, rank, local = dist_setup(a.gpus)
    if not torch.cuda.is_available():
        raise SystemExit("bench.py (impl=ours) needs a CUDA device: there is no CPU fallback")
    dev = torch.device("cuda", local)
    torch.cuda.set_device(dev)
    B = a.train_batch
    host_pool = make_wavs(max(B, 4 * B), 1234 + rank)
    sampler = ClockSampler(local)
    sampler.start()
    r = time_train_steps(sab, recipe, dev, world, B, a.steps, max(a.warmup, 3), host_pool)
    sampler.stop_flag = True
    sampler.join(timeout=2)
    if rank == 0:
        line = {"metric": "audio-sec/sec trained (Fbank+ConvAE training step)", "value": r["audio_s_per_s"], "unit": "audio-s/s",
                "n_gpus": world, "steps": a.steps, "warmup": max(a.warmup, 3), "ms_per_step": r["ms_per_step"],
                "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "fp32", "data": "synthetic",
                "config": {"workload": "training step, %d x 10 s @16 kHz utterances per GPU (BASELINE configs[3]), data-parallel "
                                       "x%d with NCCL gradient all-reduce" % (B, world), "batch_per_gpu": B,
                           "parallelism": "data-parallel x%d" % world, "weights": "random-init seed 8886",
                           "l2": "per-step working set (saved activations) %.1f GB > L2" % (B * 80640 * 32 * 4 * 20 / 1e9)},
                "clocks": sampler.summary(), "train": r, "gpu_launches": None, "e2e": None, "roofline": None, "cpu_baseline": None}
        print(json.dumps(line))
    if world > 1:
        import torch.distributed as dist
        dist.destroy_process_group()


# --------------------------------------------------------------------------- reference arm
def run_reference(a):
    if a.workload == "train":
        a.workload = "convae"     # the CPU arm times the forward path
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    if rank != 0:
        return
    n = a.cpu_sample * 2
    # bound the whole run to a few minutes: probe one pass, shrink the sample if needed
    v, dt, cores = time_cpu(a.workload, n, 1, 1)
    budget = 150.0
    while n > 1 and dt * (a.steps + a.warmup) > budget:
        n = max(1, n // 2)
        v, dt, cores = time_cpu(a.workload, n, 1, 0)
    v, dt, cores = time_cpu(a.workload, n, a.steps, a.warmup)
    T = 1 + N_SAMPLES // 160
    line = {"impl": "reference",
            "metric": "audio-sec/sec anonymized (Fbank+%s fwd)" % ("ConvAE" if a.workload == "convae" else "FC-AE"),
            "value": v, "unit": "audio-s/s", "n_gpus": world, "steps": a.steps, "warmup": a.warmup,
            "ms_per_step": dt * 1e3, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "fp32",
            "data": "synthetic",
            "config": dict(workload_config(a, world), cpu_sample="each step = %d x 10 s utterances of that workload on the "
                           "host CPU (throughput is per audio-second, so the sample size does not bias it)" % n),
            "cpu_baseline": {"value": v, "unit": "audio-s/s", "cores": cores, "kind": "port",
                             "sample": "%d x 10 s utterances per step (oracle/ port of the reference's PyTorch-CPU "
                                       "path: the reference itself is not on the GPU box)" % n},
            "e2e": {"value": v, "unit": "audio-s/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    print(json.dumps(line))


if __name__ == "__main__":
    args = parse()
    if args.impl == "reference":
        run_reference(args)
    elif args.workload == "train":
        run_train(args)
    else:
        run_ours(args)
