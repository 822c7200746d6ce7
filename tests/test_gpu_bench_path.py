"""GPU parity tests of the EXACT path bench.py times: the 16-bit (fp16 tensor-core) mode through the host-buffer
entry points (spanon_anonymize_host, spanon_pipe_submit / _submit_i16 incl. the per-slot CUDA-graph replay from a slot's
second use on) and the device arm's shape (60 x 10 s, default chunk, one CUDA graph per step), against the CPU oracle.

Tolerance: BASELINE.json's 16-bit bar, anonymizer output <= 1e-2 rel. (rel = max|a-ref| / max|ref|); the sex
log-probabilities are held to 2e-3 abs. (measured ~1e-4: the pooled statistics average 2.5 M values)."""
import numpy as np
import pytest
import torch

import recipe
from conftest import rel_max
from oracle import anonymizer_ref, fbank_np, normalize_np

import speech_anonymization_b200 as sab

pytestmark = pytest.mark.gpu
DEV = "cuda"
TOL16, LOGP_TOL16 = 1e-2, 2e-3


def cu(x):
    return torch.from_numpy(np.ascontiguousarray(x)).to(DEV)


def _model(precision="fp16"):
    m = sab.ConvAutoencoder()
    w = recipe.make_weights(recipe.CONVAE_SHAPES, 8886)
    m.load_state_dict({k: torch.from_numpy(np.asarray(v)) for k, v in w.items()})
    m = m.to(DEV).eval()
    m.precision = precision
    return m, w


def _pipeline(model, golden):
    fb = sab.Fbank(n_mels=80)
    norm = sab.InputNormalization(norm_type="global", update_until_epoch=4)
    norm._load_statistics_dict({"count": int(golden["real_count"]),
                                "glob_mean": torch.from_numpy(golden["real_glob_mean"]),
                                "glob_std": torch.from_numpy(golden["real_glob_std"]),
                                "spk_dict_mean": {}, "spk_dict_std": {}, "spk_dict_count": {}})
    return sab.AnonymizerPipeline(fb, norm.to(DEV).eval(), model)


def _oracle(wav, w, noise, golden):
    f = fbank_np.fbank(wav, n_mels=80)
    f = (f - golden["real_glob_mean"].astype(np.float64)) / golden["real_glob_std"].astype(np.float64)
    f = normalize_np.pad_time(f).astype(np.float32)
    torch.set_num_threads(8)
    with torch.no_grad():
        r, l = anonymizer_ref.convae_forward(f, w, noise)
    return r.numpy(), l.numpy()


def test_fp16_host_entry_points_vs_oracle(golden):
    """anonymize_host, StreamPipe (first use = direct launches, later uses = graph replay), int16 wire: fp16 mode."""
    m, w = _model("fp16")
    pipe = _pipeline(m, golden)
    B, n = 4, 16000
    wav, _ = recipe.make_wav(71, B, n)
    noise = recipe.make_noise(71, B, 128)
    r_ref, l_ref = _oracle(wav, w, noise, golden)
    dev = torch.device(DEV)
    pcm = torch.from_numpy(wav).pin_memory()
    nz = cu(noise)
    # one-shot synchronous host call
    rec_h = torch.empty(r_ref.shape, dtype=torch.float32).pin_memory()
    logp_h = torch.empty((B, 2), dtype=torch.float32).pin_memory()
    ws = pipe.host_workspace(B, n, dev)
    pipe.anonymize_host(pcm, rec_h, logp_h, ws, pool_noise=nz)
    assert rel_max(rec_h.numpy(), r_ref) < TOL16
    assert np.max(np.abs(logp_h.numpy() - l_ref)) < LOGP_TOL16
    # streaming pipe, depth 2: submissions 0,1 launch directly, 2.. replay the slot's captured graph
    sp = pipe.stream_pipe(B, n, dev, depth=2)
    outs = []
    for i in range(5):
        r = torch.full(r_ref.shape, float("nan"), dtype=torch.float32).pin_memory()
        l = torch.full((B, 2), float("nan"), dtype=torch.float32).pin_memory()
        sp.submit(pcm, r, l, pool_noise=nz)
        outs.append((r, l))
    sp.drain()
    for r, l in outs:
        assert rel_max(r.numpy(), r_ref) < TOL16
        assert np.max(np.abs(l.numpy() - l_ref)) < LOGP_TOL16
        assert rel_max(r.numpy(), rec_h.numpy()) < 1e-5          # same kernels, same order: replay == direct
    # alternating logp NULL / non-NULL on the same slot (the packed classifier constants must exist either way)
    sp2 = pipe.stream_pipe(B, n, dev, depth=1)
    r = torch.empty(r_ref.shape, dtype=torch.float32).pin_memory()
    l = torch.full((B, 2), float("nan"), dtype=torch.float32).pin_memory()
    sp2.submit(pcm, r, None, pool_noise=nz)
    sp2.submit(pcm, r, l, pool_noise=nz)
    sp2.submit(pcm, r, None, pool_noise=nz)
    sp2.submit(pcm, r, l, pool_noise=nz)
    sp2.drain()
    assert np.max(np.abs(l.numpy() - l_ref)) < LOGP_TOL16 and rel_max(r.numpy(), r_ref) < TOL16
    sp2.close()
    # int16 wire format: oracle on the quantised samples x / 32768
    pcm16_np = np.clip(np.round(wav * 32768.0), -32768, 32767).astype(np.int16)
    r_ref16, l_ref16 = _oracle(pcm16_np.astype(np.float32) / 32768.0, w, noise, golden)
    pcm16 = torch.from_numpy(pcm16_np).pin_memory()
    for i in range(3):       # third submit reuses slot 0 -> graph replay with the int16 format
        r = torch.full(r_ref.shape, float("nan"), dtype=torch.float32).pin_memory()
        l = torch.full((B, 2), float("nan"), dtype=torch.float32).pin_memory()
        sp.submit(pcm16, r, l, pool_noise=nz)
        sp.drain()
        assert rel_max(r.numpy(), r_ref16) < TOL16
        assert np.max(np.abs(l.numpy() - l_ref16)) < LOGP_TOL16
    # weights_changed(): the pipe re-packs; doubling decoder.8's weight doubles (recon - bias)
    with torch.no_grad():
        m.decoder[8].weight.mul_(2.0)
    sp.weights_changed(m)
    r2 = torch.empty(r_ref.shape, dtype=torch.float32).pin_memory()
    sp.submit(pcm, r2, None, pool_noise=nz)
    sp.drain()
    bias = float(m.decoder[8].bias[0])
    assert rel_max(r2.numpy() - bias, 2.0 * (rec_h.numpy() - bias)) < 2e-3
    sp.close()


def test_fp16_bench_shape_vs_oracle(golden):
    """The device arm of bench.py: 60 x 10 s utterances, default chunk, fp16, one CUDA-graph replay per step;
    three sampled utterances against the oracle (utterances are independent on the inference path)."""
    m, w = _model("fp16")
    pipe = _pipeline(m, golden)
    B = 60
    g = torch.Generator().manual_seed(4321)
    wavs = (0.1 * torch.randn(B, 160000, generator=g))
    noise = recipe.make_noise(72, B, 128)
    dwav, dnoise = wavs.to(DEV), cu(noise)
    rec_d, logp_d = pipe(dwav, pool_noise=dnoise)
    step = pipe.capture(dwav, pool_noise=dnoise)
    rec_g, logp_g = step.replay()
    rec_g, logp_g = step.replay()
    torch.cuda.synchronize()
    assert rec_g.shape == (B, 1008, 80)
    assert rel_max(rec_g.cpu().numpy(), rec_d.cpu().numpy()) < 1e-5
    sel = [0, 31, 59]
    r_ref, l_ref = _oracle(wavs[sel].numpy(), w, noise[sel], golden)
    assert rel_max(rec_g[sel].cpu().numpy(), r_ref) < TOL16
    assert np.max(np.abs(logp_g[sel].cpu().numpy() - l_ref)) < LOGP_TOL16
    # a captured step notices a parameter update (re-captures with the new weights)
    with torch.no_grad():
        m.decoder[8].bias.add_(1.0)
    rec_g2, _ = step.replay()
    torch.cuda.synchronize()
    assert abs(float((rec_g2[sel] - 1.0).cpu().sub(torch.from_numpy(r_ref)).abs().max()) / np.max(np.abs(r_ref))) < TOL16


def test_fp16_reconstruction_output(golden):
    """recon_fp16: the 16-bit mode returns its reconstruction as fp16 (half the device-to-host bytes of bench.py's e2e arm):
    device call, one-shot host call and streaming pipe == the fp32-output result rounded to half, and within the mode's bar."""
    m, w = _model("fp16")
    pipe = _pipeline(m, golden)
    B, n = 3, 16000
    wav, _ = recipe.make_wav(81, B, n)
    noise = recipe.make_noise(81, B, 128)
    r_ref, l_ref = _oracle(wav, w, noise, golden)
    nz = cu(noise)
    rec32, logp32 = pipe(cu(wav), pool_noise=nz)
    m.recon_fp16 = True
    rec16, logp16 = pipe(cu(wav), pool_noise=nz)
    assert rec16.dtype == torch.float16 and torch.equal(rec16, rec32.half()) and torch.equal(logp16, logp32)
    assert rel_max(rec16.float().cpu().numpy(), r_ref) < TOL16
    dev = torch.device(DEV)
    pcm16 = torch.from_numpy(np.clip(np.round(wav * 32768.0), -32768, 32767).astype(np.int16)).pin_memory()
    r_ref16, _ = _oracle(pcm16.numpy().astype(np.float32) / 32768.0, w, noise, golden)
    rec_h = torch.empty(r_ref.shape, dtype=torch.float16).pin_memory()
    logp_h = torch.empty((B, 2), dtype=torch.float32).pin_memory()
    pipe.anonymize_host(pcm16, rec_h, logp_h, pipe.host_workspace(B, n, dev), pool_noise=nz)
    assert rel_max(rec_h.float().numpy(), r_ref16) < TOL16
    sp = pipe.stream_pipe(B, n, dev, depth=2)
    with pytest.raises(TypeError):
        sp.submit(pcm16, torch.empty(r_ref.shape, dtype=torch.float32).pin_memory(), logp_h, pool_noise=nz)
    outs = []
    for _ in range(4):
        r = torch.full(r_ref.shape, float("nan"), dtype=torch.float16).pin_memory()
        sp.submit(pcm16, r, logp_h, pool_noise=nz)
        outs.append(r)
    sp.drain()
    sp.close()
    for r in outs:
        assert torch.equal(r, rec_h)
    m.precision = "fp32"        # the fp32 parity mode keeps its fp32 output
    assert pipe(cu(wav), pool_noise=nz)[0].dtype == torch.float32
