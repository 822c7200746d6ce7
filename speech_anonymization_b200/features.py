"""Host-side mirror of the reference's feature modules, backed by libspanon's CUDA kernels.

  Fbank               speechbrain/speechbrain/lobes/features.py:18-143
  (STFT, spectral_magnitude, Filterbank fused into one kernel)
                      speechbrain/speechbrain/processing/features.py:50-188, 327-356, 359-712
  InputNormalization  speechbrain/speechbrain/processing/features.py:938-1234
                      (fork-modified: `epoch` / `update_until_epoch`, zero-stat fallback)

Constructor and forward signatures, defaults, state handling and error behaviour follow the
reference so the YAML configs (`compute_features: !new:...Fbank`, `normalize:
!new:...InputNormalization`) only need their class path swapped.  All arithmetic on the
feature tensors runs in CUDA kernels; PyTorch only owns memory, streams and the O(n_mels)
bookkeeping of the running statistics.
"""
import ctypes as C
import logging
import math

import torch

from . import _lib

logger = logging.getLogger(__name__)


def _to_mel(hz):
    """processing/features.py:563-572"""
    return 2595 * math.log10(1 + hz / 700)


def _to_hz(mel):
    """processing/features.py:575-584"""
    return 700 * (10 ** (mel / 2595) - 1)


class Fbank(torch.nn.Module):
    """Drop-in for speechbrain.lobes.features.Fbank (lobes/features.py:78-143).

    wav [B, n_samples] (or [B, n_samples, channels]) fp32 -- or int16 PCM, converted as x / 32768 on load like the
    reference's read_audio does on the host -- on a CUDA device -> [B, T, n_mels] log-mel
    features in dB with the per-utterance top_db (80 dB) clamp; T = 1 + n_samples // hop.
    """

    def __init__(self, deltas=False, context=False, requires_grad=False, sample_rate=16000, f_min=0,
                 f_max=None, n_fft=400, n_mels=40, filter_shape="triangular", param_change_factor=1.0,
                 param_rand_factor=0.0, left_frames=5, right_frames=5, win_length=25, hop_length=10):
        super().__init__()
        self.deltas = deltas
        self.context = context
        self.requires_grad = requires_grad
        if deltas or context:
            raise NotImplementedError(
                "Fbank(deltas/context=True) is outside the hot path (off in every reference config, "
                "SURVEY.md section 2.1 #1)")
        if requires_grad:
            raise NotImplementedError("learnable filterbank parameters (requires_grad=True) need a backward pass; "
                                      "every reference config freezes them")
        if f_max is None:
            f_max = sample_rate / 2
        self.sample_rate = sample_rate
        self.f_min, self.f_max = f_min, f_max
        self.n_fft, self.n_mels = n_fft, n_mels
        self.filter_shape = filter_shape
        self.param_change_factor = param_change_factor
        self.param_rand_factor = param_rand_factor
        self.left_frames, self.right_frames = left_frames, right_frames
        # STFT.__init__ (processing/features.py:124-131): ms -> samples, periodic Hamming window
        self.win_length = int(round((sample_rate / 1000.0) * win_length))
        self.hop_length = int(round((sample_rate / 1000.0) * hop_length))
        if self.win_length > n_fft:
            raise RuntimeError("win_length (%d samples) must be <= n_fft (%d), as in torch.stft"
                               % (self.win_length, n_fft))
        self.window = torch.hamming_window(self.win_length)
        # Filterbank.__init__ (processing/features.py:415-488)
        self.amin, self.ref_value, self.top_db = 1e-10, 1.0, 80.0
        self.multiplier = 10  # power_spectrogram == 2
        self.db_multiplier = math.log10(max(self.amin, self.ref_value))
        self.n_stft = n_fft // 2 + 1
        if f_min >= f_max:  # the reference only logs (features.py:457-462)
            logger.error("Require f_min: %f < f_max: %f" % (f_min, f_max), exc_info=True)
        mel = torch.linspace(_to_mel(f_min), _to_mel(f_max), n_mels + 2)
        hz = _to_hz(mel)
        band = hz[1:] - hz[:-1]
        self.band = band[:-1]
        self.f_central = hz[1:-1]
        all_freqs = torch.linspace(0, sample_rate // 2, self.n_stft)
        self.all_freqs_mat = all_freqs.repeat(self.f_central.shape[0], 1)
        self._dev_cache = {}
        # Opt-in (not a constructor argument: the signature is the reference's): keep the features of the last waveform
        # tensor.  compute_objectives recomputes Fbank on the very tensor compute_forward just used
        # (speechbrain_convae_train.py:58 and :82); with the cache the second call is a 19 MB copy instead of a pass.
        self.cache_features = False
        self._feat_cache = None

    # -- filter matrix, evaluated with the same fp32 torch ops as the reference (:586-689) so the
    #    weights are bit-identical, then compressed to the runs of non-zero taps the kernel uses
    def fbank_matrix(self, rand_change=None):
        f_central_mat = self.f_central.repeat(self.all_freqs_mat.shape[1], 1).transpose(0, 1)
        band_mat = self.band.repeat(self.all_freqs_mat.shape[1], 1).transpose(0, 1)
        if rand_change is not None:
            f_central_mat = f_central_mat * rand_change[0]
            band_mat = band_mat * rand_change[1]
        if self.filter_shape == "triangular":
            slope = (self.all_freqs_mat - f_central_mat) / band_mat
            left_side = slope + 1.0
            right_side = -slope + 1.0
            m = torch.max(torch.zeros(1), torch.min(left_side, right_side)).transpose(0, 1)
        elif self.filter_shape == "rectangular":
            low_hz = f_central_mat - band_mat
            high_hz = f_central_mat + band_mat
            m = (self.all_freqs_mat.ge(low_hz) * self.all_freqs_mat.le(high_hz)).float().transpose(0, 1)
        else:
            smooth = torch.tensor(2)
            m = torch.exp(-0.5 * ((self.all_freqs_mat - f_central_mat) / (band_mat / smooth)) ** 2).transpose(0, 1)
        return m.contiguous()  # [n_stft, n_mels]

    @staticmethod
    def _taps(matrix):
        """[n_stft, n_mels] -> (start, length, offset, packed weights) of each filter's support."""
        n_stft, n_mels = matrix.shape
        nz = matrix != 0
        start = torch.zeros(n_mels, dtype=torch.int32)
        length = torch.zeros(n_mels, dtype=torch.int32)
        offs = torch.zeros(n_mels, dtype=torch.int32)
        chunks, off = [], 0
        for m in range(n_mels):
            idx = torch.nonzero(nz[:, m]).flatten()
            if idx.numel() > 0:
                s, e = int(idx[0]), int(idx[-1]) + 1
                start[m], length[m], offs[m] = s, e - s, off
                chunks.append(matrix[s:e, m])
                off += e - s
        w = torch.cat(chunks) if chunks else torch.zeros(1)
        return start, length, offs, w.contiguous().float()

    def _device_state(self, device, rand_change=None):
        key = (device.type, device.index)
        if rand_change is None and key in self._dev_cache:
            return self._dev_cache[key]
        start, length, offs, w = self._taps(self.fbank_matrix(rand_change))
        st = {"window": self.window.to(device), "start": start.to(device), "len": length.to(device),
              "off": offs.to(device), "w": w.to(device)}
        cfg = _lib.FbankCfg(self.n_fft, self.win_length, self.hop_length, self.n_mels, self.amin,
                            float(self.multiplier), float(self.multiplier * self.db_multiplier), self.top_db,
                            st["window"].data_ptr(), st["start"].data_ptr(), st["len"].data_ptr(),
                            st["off"].data_ptr(), st["w"].data_ptr())
        st["cfg"] = cfg
        if rand_change is None:
            self._dev_cache[key] = st
        return st

    def __getstate__(self):  # no device handles in pickles / deepcopies
        d = self.__dict__.copy()
        d["_dev_cache"] = {}
        d["_feat_cache"] = None
        return d

    def num_frames(self, n_samples):
        return 1 + n_samples // self.hop_length

    def forward_raw(self, wav):
        """Unclamped dB features plus the per-utterance max keys (input of the fused consumers)."""
        i16 = isinstance(wav, torch.Tensor) and wav.dtype == torch.int16   # raw PCM: sample = x / 32768 (read_audio)
        _lib.require_cuda(wav, "wav", torch.int16 if i16 else torch.float32)
        if wav.dim() != 2:
            raise ValueError("forward_raw expects [batch, samples]")
        wav = wav.contiguous()
        rand_change = None
        if self.param_rand_factor != 0 and self.training:  # features.py:525-533
            rand_change = 1.0 + torch.rand(2) * 2 * self.param_rand_factor - self.param_rand_factor
        st = self._device_state(wav.device, rand_change)
        B, n = wav.shape
        T = self.num_frames(n)
        out = torch.empty((B, T, self.n_mels), dtype=torch.float32, device=wav.device)
        utt_max = torch.empty((max(B, 1),), dtype=torch.int32, device=wav.device)
        with torch.cuda.device(wav.device):
            fn = _lib.lib().spanon_fbank_i16 if i16 else _lib.lib().spanon_fbank_f32
            _lib.check(fn(_lib.ptr(wav), B, n, wav.stride(0) if B > 1 else n, _lib.ptr(out), _lib.ptr(utt_max),
                          C.byref(st["cfg"]), _lib.stream_ptr(wav.device)), "fbank")
        return out, utt_max

    def forward(self, wav):
        """Returns a set of features generated from the input waveforms (lobes/features.py:126-143)."""
        key = None
        if self.cache_features and not (self.training and self.param_rand_factor != 0):
            # same storage, same in-place version, same view: the tensor has not changed since the cached call
            key = (wav.data_ptr(), wav._version, tuple(wav.shape), tuple(wav.stride()), wav.dtype, str(wav.device))
            if self._feat_cache is not None and self._feat_cache[0] == key:
                return self._feat_cache[1].clone()
        multi = wav.dim() == 3
        if multi:  # STFT.forward :143-146 / Filterbank.forward :538-559: channels are folded into the batch
            B0, n0, ch = wav.shape
            wav = wav.transpose(1, 2).reshape(B0 * ch, n0)
        out, utt_max = self.forward_raw(wav)
        B, T, F = out.shape
        if B > 0:
            with torch.cuda.device(out.device):
                _lib.check(_lib.lib().spanon_fbank_clamp(_lib.ptr(out), _lib.ptr(utt_max), B, T, F, self.top_db,
                                                         _lib.stream_ptr(out.device)), "fbank_clamp")
        if multi:
            out = out.reshape(B0, ch, T, F).permute(0, 2, 3, 1)
        if key is not None:
            self._feat_cache = (key, out.clone())
        return out


class InputNormalization(torch.nn.Module):
    """Drop-in for speechbrain.processing.features.InputNormalization (features.py:938-1234).

    Differences from the reference are implementation only: the per-utterance statistics of the
    whole batch come from one kernel launch and no host synchronisation happens (the reference
    loops over the batch in Python and syncs on `torch.round(len*T).int()` per utterance).
    """

    def __init__(self, mean_norm=True, std_norm=True, norm_type="global", avg_factor=None,
                 requires_grad=False, update_until_epoch=3):
        super().__init__()
        self.mean_norm = mean_norm
        self.std_norm = std_norm
        self.norm_type = norm_type
        self.avg_factor = avg_factor
        self.requires_grad = requires_grad
        self.glob_mean = torch.tensor([0])
        self.glob_std = torch.tensor([0])
        self.spk_dict_mean = {}
        self.spk_dict_std = {}
        self.spk_dict_count = {}
        self.weight = 1.0
        self.count = 0
        self.eps = 1e-10
        self.update_until_epoch = update_until_epoch
        self.device_inp = torch.device("cpu")

    # -- kernels
    def batch_stats(self, x, lengths, utt_max=None, top_db=80.0):
        """Per-utterance (mean [B,F], std [B,F], valid frames [B]) -- features.py:1018-1029,1123-1148."""
        _lib.require_cuda(x, "x")
        if x.dim() != 3:
            raise ValueError("InputNormalization expects [batch, time, features]")
        x = x.contiguous()
        B, T, F = x.shape
        lengths = lengths.to(device=x.device, dtype=torch.float32).contiguous()
        if lengths.numel() != B:
            raise ValueError("lengths must have one entry per utterance")
        mean = torch.empty((B, F), dtype=torch.float32, device=x.device)
        std = torch.empty((B, F), dtype=torch.float32, device=x.device)
        n_valid = torch.empty((max(B, 1),), dtype=torch.int32, device=x.device)
        with torch.cuda.device(x.device):
            _lib.check(_lib.lib().spanon_norm_stats(_lib.ptr(x), _lib.ptr(lengths), B, T, F, _lib.ptr(utt_max),
                                                    top_db, int(bool(self.mean_norm)), int(bool(self.std_norm)),
                                                    self.eps, _lib.ptr(mean), _lib.ptr(std), _lib.ptr(n_valid),
                                                    _lib.stream_ptr(x.device)), "norm_stats")
        return mean, std, n_valid[:B]

    @staticmethod
    def apply_stats(x, mean, std, out=None, pad_to=None, utt_max=None, top_db=80.0):
        """(x - mean) / std with mean/std [F] or [B,F]; optional zero padding of T to `pad_to`."""
        B, T, F = x.shape
        Tp = T if pad_to is None else pad_to
        if out is None:
            out = torch.empty((B, Tp, F), dtype=torch.float32, device=x.device)
        mean = mean.to(device=x.device, dtype=torch.float32).contiguous()
        std = std.to(device=x.device, dtype=torch.float32).contiguous()
        if mean.numel() == 1:
            mean = mean.reshape(1).expand(F).contiguous()
        if std.numel() == 1:
            std = std.reshape(1).expand(F).contiguous()
        bstride = F if mean.dim() == 2 else 0
        if (std.dim() == 2) != (mean.dim() == 2):
            raise ValueError("mean and std must both be [F] or both [B,F]")
        with torch.cuda.device(x.device):
            _lib.check(_lib.lib().spanon_norm_apply(_lib.ptr(x), _lib.ptr(out), _lib.ptr(mean), _lib.ptr(std), bstride,
                                                    _lib.ptr(utt_max), top_db, B, T, Tp, F,
                                                    _lib.stream_ptr(x.device)), "norm_apply")
        return out

    def _glob_uninitialised(self, t):
        # features.py:1114-1117 (`shape[0] == 1 and not is_nonzero`): true only for the constructor's
        # sentinel torch.tensor([0]).  The sentinel is the only INTEGER tensor these attributes ever
        # hold (statistics are fp32), so the dtype identifies it without a device round trip.
        return isinstance(t, int) or (t.dim() == 1 and t.shape[0] == 1 and not t.is_floating_point())

    def forward(self, x, lengths, spk_ids=torch.tensor([]), epoch=0):
        """Returns the normalised tensor (features.py:998-1121)."""
        _lib.require_cuda(x, "x")
        x = x.contiguous()
        N_batches = x.shape[0]
        means, stds, _ = self.batch_stats(x, lengths)

        if self.norm_type == "sentence":
            return self.apply_stats(x, means, stds, out=x)  # the reference writes x[snt_id] in place

        if self.norm_type == "speaker":
            sel_mean, sel_std = [], []
            for snt_id in range(N_batches):
                spk_id = int(spk_ids[snt_id][0])
                current_mean, current_std = means[snt_id], stds[snt_id]
                if self.training:
                    if spk_id not in self.spk_dict_mean:
                        self.spk_dict_mean[spk_id] = current_mean
                        self.spk_dict_std[spk_id] = current_std
                        self.spk_dict_count[spk_id] = 1
                    else:
                        self.spk_dict_count[spk_id] = self.spk_dict_count[spk_id] + 1
                        if self.avg_factor is None:
                            self.weight = 1 / self.spk_dict_count[spk_id]
                        else:
                            self.weight = self.avg_factor
                        self.spk_dict_mean[spk_id] = ((1 - self.weight) * self.spk_dict_mean[spk_id]
                                                      + self.weight * current_mean)
                        self.spk_dict_std[spk_id] = ((1 - self.weight) * self.spk_dict_std[spk_id]
                                                     + self.weight * current_std)
                    speaker_mean = self.spk_dict_mean[spk_id].data
                    speaker_std = self.spk_dict_std[spk_id].data
                else:
                    if spk_id in self.spk_dict_mean:
                        speaker_mean = self.spk_dict_mean[spk_id].data
                        speaker_std = self.spk_dict_std[spk_id].data
                    else:
                        speaker_mean, speaker_std = current_mean.data, current_std.data
                sel_mean.append(speaker_mean.to(x.device))
                sel_std.append(speaker_std.to(x.device))
            return self.apply_stats(x, torch.stack(sel_mean), torch.stack(sel_std), out=x)

        if self.norm_type == "batch" or self.norm_type == "global":
            current_mean = torch.mean(means, dim=0)
            current_std = torch.mean(stds, dim=0)
            if self.norm_type == "batch":
                return self.apply_stats(x, current_mean, current_std)
            if self.training:
                if self.count == 0:
                    self.glob_mean = current_mean
                    self.glob_std = current_std
                elif epoch < self.update_until_epoch:
                    if self.avg_factor is None:
                        self.weight = 1 / (self.count + 1)
                    else:
                        self.weight = self.avg_factor
                    self.glob_mean = (1 - self.weight) * self.glob_mean.to(x.device) + self.weight * current_mean
                    self.glob_std = (1 - self.weight) * self.glob_std.to(x.device) + self.weight * current_std
                self.count = self.count + 1
            if self._glob_uninitialised(self.glob_mean):
                self.glob_mean = current_mean
            if self._glob_uninitialised(self.glob_std):
                self.glob_std = current_std
            return self.apply_stats(x, self.glob_mean.data, self.glob_std.data)

        return x  # unknown norm_type: the reference falls through and returns x unchanged

    # -- state handling (features.py:1150-1234)
    def _statistics_dict(self):
        return {"count": self.count, "glob_mean": self.glob_mean, "glob_std": self.glob_std,
                "spk_dict_mean": self.spk_dict_mean, "spk_dict_std": self.spk_dict_std,
                "spk_dict_count": self.spk_dict_count}

    def _load_statistics_dict(self, state):
        self.count = state["count"]
        self.glob_mean = state["glob_mean"]
        self.glob_std = state["glob_std"]
        self.spk_dict_mean = {spk: v.to(self.device_inp) for spk, v in state["spk_dict_mean"].items()}
        self.spk_dict_std = {spk: v.to(self.device_inp) for spk, v in state["spk_dict_std"].items()}
        self.spk_dict_count = state["spk_dict_count"]
        return state

    def to(self, device):
        self = super().to(device)
        self.glob_mean = self.glob_mean.to(device)
        self.glob_std = self.glob_std.to(device)
        for spk in self.spk_dict_mean:
            self.spk_dict_mean[spk] = self.spk_dict_mean[spk].to(device)
            self.spk_dict_std[spk] = self.spk_dict_std[spk].to(device)
        self.device_inp = torch.device(device) if not isinstance(device, torch.device) else device
        return self

    def _save(self, path):
        torch.save(self._statistics_dict(), path)

    def _load(self, path, end_of_epoch=False, device=None):
        del end_of_epoch
        stats = torch.load(path, map_location=device, weights_only=False)
        self._load_statistics_dict(stats)


try:  # register SpeechBrain's checkpoint hooks when the host framework is installed (features.py:938,1208-1219)
    from speechbrain.utils.checkpoints import mark_as_loader, mark_as_saver, mark_as_transfer, register_checkpoint_hooks

    InputNormalization._save = mark_as_saver(InputNormalization._save)
    InputNormalization._load = mark_as_transfer(mark_as_loader(InputNormalization._load))
    InputNormalization = register_checkpoint_hooks(InputNormalization)
except Exception:  # SpeechBrain absent: plain methods, same names
    pass


def pad_time(feats, multiple=36):
    """speechbrain_convae_train.py:62-63: zero-pad T up to the next multiple of 36."""
    T = feats.shape[1]
    if T % multiple == 0:
        return feats
    return torch.nn.functional.pad(feats, (0, 0, 0, multiple - T % multiple, 0, 0), "constant", 0)
