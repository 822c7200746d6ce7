"""The composed hot path: waveform -> Fbank -> InputNormalization -> (pad T to x36) -> anonymizer.

Mirrors SexAnonymizationTraining.compute_forward (speechbrain_convae_train.py:42-69) and
GenderBrain.prepare_features (gender_classifier_train_recon.py:84-90).  With frozen (eval-mode,
initialised) global normalisation statistics the clamp, the normalisation and the padding are
folded into the anonymizer's first load, so the features cross HBM once.
"""
import ctypes as C

import torch

from . import _lib
from .features import Fbank, InputNormalization, pad_time
from .models import ConvAutoencoder, FullyConnectedAutoencoder


def shard_range(n_items, rank, world_size):
    """Contiguous slice [lo, hi) of a batch of independent utterances owned by `rank` (SURVEY 8e:
    inference shards by utterance with no collective)."""
    base, rem = divmod(n_items, world_size)
    lo = rank * base + min(rank, rem)
    return lo, lo + base + (1 if rank < rem else 0)


class AnonymizerPipeline(torch.nn.Module):
    def __init__(self, compute_features: Fbank, normalize: InputNormalization, model):
        super().__init__()
        self.compute_features = compute_features
        self.normalize = normalize
        self.model = model
        self.is_conv = isinstance(model, ConvAutoencoder)
        if not self.is_conv and not isinstance(model, FullyConnectedAutoencoder):
            raise TypeError("model must be a ConvAutoencoder or a FullyConnectedAutoencoder")

    def padded_frames(self, T):
        return (T + 35) // 36 * 36 if self.is_conv else T  # train.py:62 (skipped for fcae)

    def _frozen_stats(self):
        n = self.normalize
        return (n.norm_type == "global" and not n.training and n.mean_norm and n.std_norm
                and n.glob_mean.dim() == 1 and n.glob_mean.shape[0] == self.compute_features.n_mels)

    @torch.no_grad()
    def forward(self, wavs, wav_lens=None, epoch=0, pool_noise=None):
        """wavs [B, n_samples] on the GPU -> (anonymized feats [B, T', F], sex log-probs [B, 2])."""
        fb = self.compute_features
        if self._frozen_stats():
            raw, utt_max = fb.forward_raw(wavs)
            T = raw.shape[1]
            n = self.normalize
            if self.is_conv:
                fused = (n.glob_mean, n.glob_std, utt_max, fb.top_db, self.padded_frames(T))
            else:
                fused = (n.glob_mean, n.glob_std, utt_max, fb.top_db)
            return self.model(raw, pool_noise=pool_noise, fused_norm=fused)
        feats = fb(wavs)
        if wav_lens is None:
            wav_lens = torch.ones(wavs.shape[0], device=wavs.device)
        feats = self.normalize(feats, wav_lens, epoch=epoch)
        if self.is_conv:
            feats = pad_time(feats)
        return self.model(feats, pool_noise=pool_noise)

    def capture(self, wavs, wav_lens=None, pool_noise=None):
        """Captures one forward pass on `wavs` (a device tensor that stays alive and may be overwritten in place) into a
        CUDA graph: GraphedStep.replay() re-runs the whole launch sequence (Fbank, the anonymizer's ~28 kernels and its
        memsets) with ONE host call, so a busy or descheduled host thread cannot starve the GPU between kernels."""
        return GraphedStep(self, wavs, wav_lens, pool_noise)

    # -- host-buffer path through the single C entry point (bench.py's e2e leg)
    def host_workspace(self, B, n_samples, device):
        opts = self.model.opts() if self.is_conv else _lib.ConvAEOpts(0, 0, 0, 0, 0)
        nbytes = _lib.lib().spanon_anonymize_workspace_bytes(0 if self.is_conv else 1, B, n_samples,
                                                             self.compute_features.hop_length,
                                                             self.compute_features.n_mels, C.byref(opts))
        return torch.empty((nbytes,), dtype=torch.uint8, device=device)

    @torch.no_grad()
    def anonymize_host(self, pcm_host, recon_host, logp_host, workspace, pool_noise=None):
        """pcm_host [B, n] (pinned) host fp32 or int16 PCM -> recon_host [B, T', F], logp_host [B, 2] host tensors.
        H2D copy, all kernels, D2H copy and the stream sync happen inside spanon_anonymize_host."""
        if not self._frozen_stats():
            raise _lib.SpanonError("anonymize_host needs eval-mode InputNormalization with loaded global statistics")
        dev = workspace.device
        fb, n = self.compute_features, self.normalize
        st = fb._device_state(dev)
        mean = n.glob_mean.to(device=dev, dtype=torch.float32).contiguous()
        std = n.glob_std.to(device=dev, dtype=torch.float32).contiguous()
        w = self.model.weight_struct()
        opts = self.model.opts() if self.is_conv else _lib.ConvAEOpts(0, 0, 0, 0, 0)
        if recon_host.dtype != (torch.float16 if opts.recon_fp16 else torch.float32):
            raise TypeError("recon_host dtype does not match model.recon_fp16")
        B, ns = pcm_host.shape
        fn = _lib.lib().spanon_anonymize_host_i16 if pcm_host.dtype == torch.int16 else _lib.lib().spanon_anonymize_host
        with torch.cuda.device(dev):
            _lib.check(fn(
                0 if self.is_conv else 1, _lib.ptr(pcm_host), B, ns, C.byref(st["cfg"]), _lib.ptr(mean),
                _lib.ptr(std), C.cast(C.byref(w), C.c_void_p), _lib.ptr(pool_noise), C.byref(opts),
                _lib.ptr(recon_host), _lib.ptr(logp_host), _lib.ptr(workspace), workspace.numel(),
                _lib.stream_ptr(dev)), "anonymize_host")
        return recon_host, logp_host

    def stream_pipe(self, max_batch, n_samples, device, depth=2):
        """A StreamPipe (spanon_pipe_*) bound to this pipeline's frozen features, statistics and weights."""
        return StreamPipe(self, max_batch, n_samples, device, depth)


class GraphedStep:
    """One captured forward pass of an AnonymizerPipeline (torch.cuda.CUDAGraph around the C-ABI launches; the library
    only enqueues on the current stream and never synchronises, so its launch sequence is capturable as is).
    `recon`, `logp` are the static output tensors every replay() overwrites."""

    def __init__(self, pipeline, wavs, wav_lens=None, pool_noise=None, warmup=2):
        self.pipeline, self.wavs, self.wav_lens, self.pool_noise, self.warmup = pipeline, wavs, wav_lens, pool_noise, warmup
        self._capture()

    def _weights_key(self):
        m = self.pipeline.model
        return tuple((t.data_ptr(), t._version) for t in list(m.parameters()) + list(m.buffers()))

    def _capture(self):
        pipeline, wavs, wav_lens, pool_noise = self.pipeline, self.wavs, self.wav_lens, self.pool_noise
        # The graph bakes in raw pointers into the model's workspace (packed weights included): give the captured step a
        # PRIVATE workspace that this object keeps alive, so a later eager call with another shape cannot free or
        # repurpose the memory the graph replays into.
        model = pipeline.model
        saved = model.__dict__.pop("_ws_cache", None)
        side = torch.cuda.Stream(device=wavs.device)
        side.wait_stream(torch.cuda.current_stream(wavs.device))
        with torch.cuda.stream(side):
            for _ in range(self.warmup):   # first calls pack weights / set kernel attributes: keep them out of the graph
                pipeline(wavs, wav_lens, pool_noise=pool_noise)
        torch.cuda.current_stream(wavs.device).wait_stream(side)
        torch.cuda.synchronize(wavs.device)
        self.launches_before = _lib.lib().spanon_launch_count()
        self.graph = torch.cuda.CUDAGraph()
        with torch.cuda.graph(self.graph):
            self.recon, self.logp = pipeline(wavs, wav_lens, pool_noise=pool_noise)
        self.launches = int(_lib.lib().spanon_launch_count() - self.launches_before)   # kernels per replay
        self._ws = model.__dict__.pop("_ws_cache", None)     # owned by this graph from here on
        if saved is not None:
            model.__dict__["_ws_cache"] = saved
        self._key = self._weights_key()

    def replay(self, check_weights=True):
        """Replays the captured step.  The graph holds the weights as packed at capture time; with `check_weights` a
        changed parameter (load_state_dict, optimizer step, any versioned in-place write) triggers a re-capture.
        Writes through `.data` do not bump the version counter -- call recapture() after those."""
        if check_weights and self._key != self._weights_key():
            self._capture()
        self.graph.replay()
        return self.recon, self.logp

    def recapture(self):
        self._capture()


class StreamPipe:
    """`depth` batches in flight over H2D | kernel | D2H streams (include/spanon.h, spanon_pipe_*).

    submit() enqueues a batch of pinned host PCM and returns; wait() retires the oldest batch, after which its
    recon / logp host tensors are complete.  The device workspaces belong to the pipe."""

    def __init__(self, pipeline, max_batch, n_samples, device, depth=2):
        if not pipeline._frozen_stats():
            raise _lib.SpanonError("StreamPipe needs eval-mode InputNormalization with loaded global statistics")
        fb, n = pipeline.compute_features, pipeline.normalize
        self.device = torch.device(device)
        self._keep = []          # host / device tensors referenced by batches in flight or by the pipe itself
        st = fb._device_state(self.device)
        self._mean = n.glob_mean.to(device=self.device, dtype=torch.float32).contiguous()
        self._std = n.glob_std.to(device=self.device, dtype=torch.float32).contiguous()
        self._w = pipeline.model.weight_struct()
        opts = pipeline.model.opts() if pipeline.is_conv else _lib.ConvAEOpts(0, 0, 0, 0, 0)
        self.recon_dtype = torch.float16 if opts.recon_fp16 else torch.float32     # what submit()'s recon_host must be
        self._h = C.c_void_p()
        with torch.cuda.device(self.device):
            _lib.check(_lib.lib().spanon_pipe_create(
                C.byref(self._h), 0 if pipeline.is_conv else 1, max_batch, n_samples, C.byref(st["cfg"]),
                _lib.ptr(self._mean), _lib.ptr(self._std), C.cast(C.byref(self._w), C.c_void_p), C.byref(opts), depth),
                "pipe_create")
        self.depth = depth

    def submit(self, pcm_host, recon_host, logp_host, pool_noise=None):
        # int16 PCM (the audio file's own sample format) halves the host-to-device bytes; the Fbank kernel converts
        if recon_host.dtype != self.recon_dtype:
            raise TypeError("recon_host must be %s for this pipe (model.recon_fp16 at creation time), got %s"
                            % (self.recon_dtype, recon_host.dtype))
        fn = _lib.lib().spanon_pipe_submit_i16 if pcm_host.dtype == torch.int16 else _lib.lib().spanon_pipe_submit
        with torch.cuda.device(self.device):
            _lib.check(fn(self._h, _lib.ptr(pcm_host), pcm_host.shape[0], _lib.ptr(pool_noise), _lib.ptr(recon_host),
                          _lib.ptr(logp_host)), "pipe_submit")
        self._keep.append((pcm_host, recon_host, logp_host, pool_noise))
        del self._keep[:-self.depth]

    def wait(self):
        _lib.check(_lib.lib().spanon_pipe_wait(self._h), "pipe_wait")

    def weights_changed(self, model=None):
        """Call after the model's parameters changed (or pass a new model): drains the pipe and makes every slot re-pack
        its weights (and re-capture its graph) on next use.  The pipe otherwise keeps using the packed copies."""
        w = None
        if model is not None:
            self._w = model.weight_struct()
            w = C.cast(C.byref(self._w), C.c_void_p)
        _lib.check(_lib.lib().spanon_pipe_weights_changed(self._h, w), "pipe_weights_changed")

    def drain(self):
        _lib.check(_lib.lib().spanon_pipe_drain(self._h), "pipe_drain")

    def close(self):
        if self._h:
            _lib.lib().spanon_pipe_destroy(self._h)
            self._h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass
