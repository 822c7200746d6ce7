"""ORACLE (test infrastructure, never on the product path).

CPU restatement, in numpy float64, of the reference's SpeechBrain Fbank pipeline:

  STFT               speechbrain/speechbrain/processing/features.py:101-188
  spectral_magnitude speechbrain/speechbrain/processing/features.py:327-356
  Filterbank         speechbrain/speechbrain/processing/features.py:415-488 (ctor),
                     :490-560 (forward), :586-610 (triangular), :691-712 (dB)
  Fbank              speechbrain/speechbrain/lobes/features.py:78-143

Only ``tests/``, ``__graft_entry__.smoke()`` and ``bench.py``'s cpu_baseline /
``--impl reference`` legs may import this file.  It is pinned against the real
reference by ``tests/test_oracle_vs_reference.py`` (runs whenever /root/reference
is mounted) and by the committed golden vectors in ``tests/golden/`` that were
generated from the reference's own code (``tests/golden/make_golden.py``).
The arithmetic that lives outside /root/reference (torch.stft, torch.matmul,
torch.log10: PyTorch, reference used 1.10, container has 2.11) is restated from
its published definition: centred, zero padded, periodic Hamming window, one
sided DFT.
"""
import math
import numpy as np


def stft_params(sample_rate=16000, win_length_ms=25, hop_length_ms=10):
    """processing/features.py:124-129 -- ms -> samples with Python round()."""
    win = int(round((sample_rate / 1000.0) * win_length_ms))
    hop = int(round((sample_rate / 1000.0) * hop_length_ms))
    return win, hop


def num_frames(n_samples, hop=160):
    """torch.stft(center=True): T = 1 + floor(n_samples / hop) (features.py:161-172)."""
    return 1 + n_samples // hop


def hamming_window(n):
    """torch.hamming_window(n) (features.py:107,131): periodic, alpha=.54, beta=.46."""
    k = np.arange(n, dtype=np.float64)
    return 0.54 - 0.46 * np.cos(2.0 * math.pi * k / n)


def to_mel(hz):
    """features.py:563-572"""
    return 2595.0 * math.log10(1.0 + hz / 700.0)


def mel_filter_params(n_mels=80, f_min=0.0, f_max=8000.0):
    """features.py:465-473: centre frequencies and (left) band widths in Hz."""
    mel = np.linspace(to_mel(f_min), to_mel(f_max), n_mels + 2)
    hz = 700.0 * (10.0 ** (mel / 2595.0) - 1.0)          # features.py:575-584
    band = (hz[1:] - hz[:-1])[:-1]
    f_central = hz[1:-1]
    return f_central, band


def mel_matrix(n_mels=80, f_min=0.0, f_max=8000.0, n_fft=400, sample_rate=16000):
    """features.py:485-488 + :586-610 -> [n_fft//2+1, n_mels].

    NB the reference's triangle is *symmetric* and uses only the LEFT band width
    (slope = (f - f_c) / band; min(slope+1, -slope+1) clipped at 0).
    """
    f_central, band = mel_filter_params(n_mels, f_min, f_max)
    n_stft = n_fft // 2 + 1
    all_freqs = np.linspace(0.0, sample_rate // 2, n_stft)
    slope = (all_freqs[None, :] - f_central[:, None]) / band[:, None]
    tri = np.maximum(0.0, np.minimum(slope + 1.0, -slope + 1.0))
    return tri.T.copy()                                   # [n_stft, n_mels]


def mel_matrix_f32(n_mels=80, f_min=0.0, f_max=8000.0, n_fft=400, sample_rate=16000):
    """The SAME matrix as mel_matrix() but evaluated the way the reference evaluates it:
    in float32 with torch's elementwise kernels (features.py:465-488, :586-610).  The
    fp32 rounding of torch.linspace / torch.pow moves the filter edges by ~1e-3 Hz, i.e.
    the weights by ~1e-5 -- ten times the Fbank parity budget on a bin that dominates
    its filter -- so parity against the reference needs this fp32 evaluation order.
    torch.linspace / torch.pow are PyTorch arithmetic (not vendored in /root/reference;
    pinned here by execution, torch 2.11)."""
    import torch
    mel = torch.linspace(to_mel(f_min), to_mel(f_max), n_mels + 2)
    hz = 700 * (10 ** (mel / 2595) - 1)
    band = (hz[1:] - hz[:-1])[:-1]
    f_central = hz[1:-1]
    all_freqs = torch.linspace(0, sample_rate // 2, n_fft // 2 + 1)
    slope = (all_freqs[None, :] - f_central[:, None]) / band[:, None]
    tri = torch.max(torch.zeros(1), torch.min(slope + 1.0, -slope + 1.0))
    return tri.transpose(0, 1).contiguous().numpy()        # [n_stft, n_mels] float32


def stft_power(wav, n_fft=400, win=400, hop=160):
    """|STFT|^2, [B, T, n_fft//2+1] (features.py:161-186 then :347, power=1)."""
    wav = np.asarray(wav, dtype=np.float64)
    if wav.ndim == 1:
        wav = wav[None]
    assert win == n_fft, "reference configs use win_length == n_fft"
    B, n = wav.shape
    T = num_frames(n, hop)
    pad = n_fft // 2
    x = np.zeros((B, n + 2 * pad), dtype=np.float64)      # pad_mode="constant"
    x[:, pad:pad + n] = wav
    idx = (np.arange(T) * hop)[:, None] + np.arange(n_fft)[None, :]
    frames = x[:, idx] * hamming_window(win)[None, None, :]
    spec = np.fft.rfft(frames, n=n_fft, axis=-1)
    return spec.real ** 2 + spec.imag ** 2


def amplitude_to_db(x, amin=1e-10, ref_value=1.0, top_db=80.0, multiplier=10.0):
    """features.py:691-712.  Per-utterance max over (T, F) INCLUDING padded tails."""
    x_db = multiplier * np.log10(np.maximum(x, amin))
    x_db = x_db - multiplier * math.log10(max(amin, ref_value))
    floor = x_db.max(axis=(-2, -1), keepdims=True) - top_db
    return np.maximum(x_db, floor)


def fbank(wav, n_mels=80, sample_rate=16000, n_fft=400, f_min=0.0, f_max=None,
          win_length_ms=25, hop_length_ms=10, exact_mel=False):
    """lobes/features.py:126-143 with deltas=False, context=False -> [B, T, n_mels] f64.
    exact_mel=False (default) uses the reference's fp32-evaluated filter matrix."""
    if f_max is None:
        f_max = sample_rate / 2
    win, hop = stft_params(sample_rate, win_length_ms, hop_length_ms)
    power = stft_power(wav, n_fft, win, hop)
    mm = mel_matrix if exact_mel else mel_matrix_f32
    mel = power @ mm(n_mels, f_min, f_max, n_fft, sample_rate).astype(np.float64)
    return amplitude_to_db(mel)


def fbank_raw_db(wav, **kw):
    """Same as fbank() but WITHOUT the top_db clamp; returns (raw_db, utt_max)."""
    n_mels = kw.pop("n_mels", 80)
    sample_rate = kw.pop("sample_rate", 16000)
    n_fft = kw.pop("n_fft", 400)
    win, hop = stft_params(sample_rate, 25, 10)
    power = stft_power(wav, n_fft, win, hop)
    mel = power @ mel_matrix_f32(n_mels, 0.0, sample_rate / 2, n_fft, sample_rate).astype(np.float64)
    x_db = 10.0 * np.log10(np.maximum(mel, 1e-10))
    return x_db, x_db.max(axis=(-2, -1))
