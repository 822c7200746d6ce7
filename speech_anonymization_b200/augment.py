"""Drop-in for speechbrain.lobes.augment.SpecAugment (speechbrain/speechbrain/lobes/augment.py:32-200), the feature-domain
augmentation of the anonymizer's TRAIN stage (speechbrain_convae_train.py:64-67, convae.yaml:273-283).

Same constructor, same `forward(x [B, T, F]) -> [B, T, F]`, and the SAME random draws: the module makes the reference's
torch.randint calls (same bounds, shapes, devices and order), so a seeded run picks the same warp and masks; the
resampling and masking run in libspanon's kernels (csrc/augment.cu).  Unlike the reference, which edits `x` in place,
the input is left untouched when a time warp is applied (the output is a new tensor either way)."""
import ctypes as C

import torch

from . import _lib


class SpecAugment(torch.nn.Module):
    def __init__(self, time_warp=True, time_warp_window=5, time_warp_mode="bicubic", freq_mask=True,
                 freq_mask_width=(0, 20), n_freq_mask=2, time_mask=True, time_mask_width=(0, 100), n_time_mask=2,
                 replace_with_zero=True):
        super().__init__()
        assert time_warp or freq_mask or time_mask, "at least one of time_warp, time_mask, or freq_mask should be applied"
        if time_warp and time_warp_mode != "bicubic":
            raise NotImplementedError("the CUDA time warp implements the bicubic mode (the default and convae.yaml's)")
        self.apply_time_warp = time_warp
        self.time_warp_window = time_warp_window
        self.time_warp_mode = time_warp_mode
        self.freq_mask = freq_mask
        if isinstance(freq_mask_width, int):
            freq_mask_width = (0, freq_mask_width)
        self.freq_mask_width = freq_mask_width
        self.n_freq_mask = n_freq_mask
        self.time_mask = time_mask
        if isinstance(time_mask_width, int):
            time_mask_width = (0, time_mask_width)
        self.time_mask_width = time_mask_width
        self.n_time_mask = n_time_mask
        self.replace_with_zero = replace_with_zero

    # -- the reference's draws (augment.py:138-140, 181-187)
    def draw_time_warp(self, time):
        window = self.time_warp_window
        if time - window <= window:
            return None
        c = torch.randint(window, time - window, (1,))[0]
        w = torch.randint(c - window, c + window, (1,))[0] + 1
        return int(c), int(w)

    def draw_masks(self, batch, D, width_range, n_mask, device):
        mask_len = torch.randint(width_range[0], width_range[1], (batch, n_mask), device=device).unsqueeze(2)
        mask_pos = torch.randint(0, max(1, D - mask_len.max()), (batch, n_mask), device=device).unsqueeze(2)
        return mask_len.squeeze(2), mask_pos.squeeze(2)

    # -- kernels
    @staticmethod
    def time_warp_with(x, c, w):
        B, T, F = x.shape
        out = torch.empty_like(x)
        with torch.cuda.device(x.device):
            _lib.check(_lib.lib().spanon_specaug_warp(_lib.ptr(x), _lib.ptr(out), B, T, F, int(c), int(w),
                                                      _lib.stream_ptr(x.device)), "specaug_warp")
        return out

    def mask_with(self, x, axis, mask_len, mask_pos):
        """In place on `x`; axis 1 = time, 2 = feature."""
        B, T, F = x.shape
        ml = mask_len.to(device=x.device, dtype=torch.int32).contiguous()
        mp = mask_pos.to(device=x.device, dtype=torch.int32).contiguous()
        scratch = torch.empty((1,), dtype=torch.float64, device=x.device)
        with torch.cuda.device(x.device):
            _lib.check(_lib.lib().spanon_specaug_mask(_lib.ptr(x), B, T, F, axis, ml.shape[1], _lib.ptr(ml), _lib.ptr(mp),
                                                      int(bool(self.replace_with_zero)), _lib.ptr(scratch),
                                                      _lib.stream_ptr(x.device)), "specaug_mask")
        return x

    def forward(self, x):
        _lib.require_cuda(x, "x")
        if x.dim() != 3:
            raise ValueError("SpecAugment expects [batch, time, features]")
        x = x.contiguous()
        B, T, F = x.shape
        fresh = False
        if self.apply_time_warp:
            cw = self.draw_time_warp(T)
            if cw is not None:
                x = self.time_warp_with(x, *cw)
                fresh = True
        if not fresh:
            x = x.clone()
        if self.freq_mask:
            ml, mp = self.draw_masks(B, F, self.freq_mask_width, self.n_freq_mask, x.device)
            self.mask_with(x, 2, ml, mp)
        if self.time_mask:
            ml, mp = self.draw_masks(B, T, self.time_mask_width, self.n_time_mask, x.device)
            self.mask_with(x, 1, ml, mp)
        return x
