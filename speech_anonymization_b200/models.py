"""Host-side mirror of the reference's anonymizer modules, backed by libspanon's CUDA kernels.

  ConvAutoencoder / TDNNSexClassifier / GLU / GradReverse   models/ConvAutoEncoder.py:12-69,113-200
  ConvReconstruction ("endtoend")                           models/EndToEnd.py:36-87
  Xvector / Classifier / EncoderClassifier (feature path)   speechbrain/speechbrain/lobes/models/Xvector.py:19-172,
                                                            speechbrain/speechbrain/pretrained/interfaces.py:762-829
  FullyConnectedAutoencoder / FullyConnSexClassifier        models/FullyConnected.py:65-159
  StatisticsPooling                                         speechbrain/speechbrain/nnet/pooling.py:229-326

The modules are built from the same torch.nn layer objects in the same order as the reference,
so `state_dict()` keys, default initialisation under a given seed and checkpoint loading are
identical; the layer objects are parameter containers only -- `forward` hands the parameter
storages to one C call (spanon_convae_fwd / spanon_fcae_fwd).  Forward only (inference /
evaluation, eval-mode BatchNorm): the backward pass is a later row of the scope table.
"""
import ctypes as C
import os

import torch
import torch.nn as nn

from . import _lib


class GradReverse(torch.autograd.Function):
    """models/ConvAutoEncoder.py:12-28 (identity forward, negated gradient)."""

    @staticmethod
    def forward(ctx, input):
        return input

    @staticmethod
    def backward(ctx, grad_output):
        return -grad_output.clone()

    def grad_reverse(input):
        return GradReverse.apply(input)


class GLU(nn.Module):
    """models/ConvAutoEncoder.py:113-120: x * sigmoid(x), no channel halving (marker module:
    the kernels fuse it into the consumer's operand load)."""

    def forward(self, input):
        return input * torch.sigmoid(input)


class StatisticsPooling(nn.Module):
    """Noise source of speechbrain.nnet.pooling.StatisticsPooling (pooling.py:313-326); the pooling
    itself runs inside the classifier kernels."""

    def __init__(self, return_mean=True, return_std=True):
        super().__init__()
        self.eps = 1e-5
        self.return_mean = return_mean
        self.return_std = return_std
        if not (self.return_mean or self.return_std):
            raise ValueError("both of statistics are equal to False \n"
                             "consider enabling mean and/or std statistic pooling")

    def _get_gauss_noise(self, shape_of_tensor, device="cpu"):
        gnoise = torch.randn(shape_of_tensor, device=device)
        gnoise -= torch.min(gnoise)
        gnoise /= torch.max(gnoise)
        gnoise = self.eps * ((1 - 9) * gnoise + 9)
        return gnoise


def _p(t):
    if t.dtype != torch.float32 or not t.is_cuda:
        raise _lib.SpanonError("module parameters must be fp32 CUDA tensors (got %s on %s); call .to('cuda')"
                               % (t.dtype, t.device))
    if not t.is_contiguous():
        raise _lib.SpanonError("module parameters must be contiguous")
    return t.data_ptr()


def _bn(m):
    return _lib.BatchNorm(_p(m.weight), _p(m.bias), _p(m.running_mean), _p(m.running_var))


def _forward_only_guard(module, inp):
    if torch.is_grad_enabled() and (inp.requires_grad or any(p.requires_grad for p in module.parameters())):
        raise NotImplementedError(
            "%s: only the forward pass runs on the CUDA kernels; call under torch.no_grad() "
            "(the backward pass is scope row 8(f)-1)" % type(module).__name__)
    if any(isinstance(m, nn.modules.batchnorm._BatchNorm) and m.training for m in module.modules()):
        raise NotImplementedError(
            "%s: train-mode BatchNorm (batch statistics) is not implemented; call .eval()"
            % type(module).__name__)


def _precision_from_env():
    v = os.environ.get("SPANON_PRECISION", "fp16").lower()
    if v not in ("fp16", "fp32"):
        raise ValueError("SPANON_PRECISION must be fp16 or fp32")
    return v


class TDNNSexClassifier(nn.Module):
    """models/ConvAutoEncoder.py:30-69.  Inside ConvAutoencoder.forward the classifier runs fused in spanon_convae_fwd;
    called on its own (forward(input [B, 128, L']) -> log-probs [B, 2], :57-69) it runs on the fp32 training kernels
    (train.tdnn_classifier_forward): batch statistics + autograd in train mode, running statistics under no_grad in eval."""

    def __init__(self, num_classes):
        super().__init__()
        self.tdnn = nn.Sequential(
            nn.Conv1d(in_channels=128, out_channels=128, kernel_size=5, dilation=1), nn.ReLU(), nn.BatchNorm1d(128),
            nn.Conv1d(in_channels=128, out_channels=128, kernel_size=3, dilation=2), nn.ReLU(), nn.BatchNorm1d(128),
            nn.Conv1d(in_channels=128, out_channels=128, kernel_size=3, dilation=3), nn.ReLU(), nn.BatchNorm1d(128),
        )
        self.norm = nn.BatchNorm1d(128)
        self.stats_pooling = StatisticsPooling()
        self.classify = nn.Sequential(
            nn.Linear(256, 128), nn.ReLU(), nn.BatchNorm1d(128),
            nn.Linear(128, 64), nn.ReLU(), nn.BatchNorm1d(64),
            nn.Linear(64, num_classes),
        )

    def forward(self, input, pool_noise=None):
        _lib.require_cuda(input, "input")
        if input.dim() != 3 or input.shape[1] != 128:
            raise RuntimeError("TDNNSexClassifier expects [batch, 128, time] (channels first), got %s" % (tuple(input.shape),))
        if input.shape[2] < 15:
            raise RuntimeError("TDNNSexClassifier: the three TDNN layers need at least 15 time steps")
        from . import train
        return train.tdnn_classifier_forward(self, input.float().transpose(1, 2).contiguous(), pool_noise,
                                             getattr(self, "sync_bn_group", None))


class ConvAutoencoder(nn.Module):
    """Drop-in for models.ConvAutoEncoder.ConvAutoencoder (models/ConvAutoEncoder.py:136-200).

    forward(input [B,T,F]) -> (reconstruction [B,T,F], sex log-probs [B,2]).
    `precision`: "fp16" (fp16 operands on the tcgen05 tensor cores, fp32 accumulation; the 16-bit mode
    with the <= 1e-2 parity bar) or "fp32" (<= 1e-4 rel.); default from $SPANON_PRECISION, else fp16.
    """

    def __init__(self):
        super().__init__()
        self.encoder = nn.Sequential(
            nn.Conv1d(in_channels=1, out_channels=32, kernel_size=15, stride=1, padding=7), GLU(),
            nn.Conv1d(in_channels=32, out_channels=64, kernel_size=5, stride=2, padding=2),
            nn.InstanceNorm1d(num_features=64, affine=True), GLU(),
            nn.Conv1d(in_channels=64, out_channels=64, kernel_size=5, stride=1, padding=2),
            nn.InstanceNorm1d(num_features=64, affine=True), GLU(),
            nn.Conv1d(in_channels=64, out_channels=128, kernel_size=5, stride=2, padding=2),
            nn.InstanceNorm1d(num_features=128, affine=True), GLU(),
            nn.Conv1d(in_channels=128, out_channels=128, kernel_size=5, stride=1, padding=2),
            nn.InstanceNorm1d(num_features=128, affine=True), GLU(),
        )
        self.decoder = nn.Sequential(
            nn.Conv1d(in_channels=128, out_channels=128, kernel_size=5, stride=1, padding=2),
            nn.ConvTranspose1d(in_channels=128, out_channels=64, kernel_size=5, stride=2, padding=2,
                               output_padding=1),
            nn.InstanceNorm1d(num_features=64, affine=True), GLU(),
            nn.Conv1d(in_channels=64, out_channels=64, kernel_size=5, stride=1, padding=2),
            nn.ConvTranspose1d(in_channels=64, out_channels=32, kernel_size=5, stride=2, padding=2,
                               output_padding=1),
            nn.InstanceNorm1d(num_features=32, affine=True), GLU(),
            nn.Conv1d(in_channels=32, out_channels=1, kernel_size=15, stride=1, padding=7),
        )
        self.sex_classifier = TDNNSexClassifier(2)
        self.precision = _precision_from_env()
        self.chunk_utts = 0  # 0 = library default (L2-resident chunks)
        self.recon_fp16 = False  # fp16 mode only: return the reconstruction as a half tensor (the mode's own precision)

    def weight_struct(self):
        e, d, s = self.encoder, self.decoder, self.sex_classifier
        w = _lib.ConvAEWeights()
        for i, k in enumerate((0, 2, 5, 8, 11)):
            w.enc_w[i], w.enc_b[i] = _p(e[k].weight), _p(e[k].bias)
        for i, k in enumerate((3, 6, 9, 12)):
            w.enc_in_w[i], w.enc_in_b[i] = _p(e[k].weight), _p(e[k].bias)
        for i, k in enumerate((0, 1, 4, 5, 8)):
            w.dec_w[i], w.dec_b[i] = _p(d[k].weight), _p(d[k].bias)
        for i, k in enumerate((2, 6)):
            w.dec_in_w[i], w.dec_in_b[i] = _p(d[k].weight), _p(d[k].bias)
        w.clf_norm = _bn(s.norm)
        for i, k in enumerate((0, 3, 6)):
            w.tdnn_w[i], w.tdnn_b[i] = _p(s.tdnn[k].weight), _p(s.tdnn[k].bias)
            w.tdnn_bn[i] = _bn(s.tdnn[k + 2])
        for i, k in enumerate((0, 3, 6)):
            w.cls_w[i], w.cls_b[i] = _p(s.classify[k].weight), _p(s.classify[k].bias)
        w.cls_bn[0], w.cls_bn[1] = _bn(s.classify[2]), _bn(s.classify[5])
        return w

    def __getstate__(self):   # the cached device workspace is not part of the module's state
        d = dict(self.__dict__)
        d.pop("_ws_cache", None)
        return d

    def opts(self, skip_classifier=False):
        return _lib.ConvAEOpts(_lib.PREC_FP32 if self.precision == "fp32" else _lib.PREC_FP16,
                               int(self.chunk_utts), int(skip_classifier), 0,
                               int(bool(self.__dict__.get("recon_fp16", False)) and self.precision != "fp32"))

    def forward(self, input, pool_noise=None, fused_norm=None):
        """`pool_noise` [B,128]: the tensor StatisticsPooling adds to the pooled mean (default: drawn
        like the reference, pooling.py:313-326).  `fused_norm` = (mean [F], std [F], utt_max [B],
        top_db, T_pad): `input` is then RAW Fbank dB [B,T,F] and clamp + normalisation + zero padding to
        T_pad run inside the first layer's load (used by AnonymizerPipeline)."""
        _lib.require_cuda(input, "input")
        if input.dim() != 3:
            raise ValueError("ConvAutoencoder expects [batch, time, features]")
        # TRAIN mode (batch-statistics BatchNorm in the sex classifier) or a graph is being recorded: the training kernels
        # (fp32, with a backward pass; train.py).  Otherwise the inference kernels below.
        bn_training = any(isinstance(m, nn.modules.batchnorm._BatchNorm) and m.training for m in self.modules())
        wants_grad = torch.is_grad_enabled() and (input.requires_grad or any(p.requires_grad for p in self.parameters()))
        if bn_training or wants_grad:
            if not bn_training:
                raise NotImplementedError("ConvAutoencoder: a backward pass through eval-mode BatchNorm is not built; call "
                                          ".train() (the reference trains in TRAIN mode) or use torch.no_grad()")
            if fused_norm is not None:
                raise NotImplementedError("the fused normalisation load is an inference path; pass normalised features")
            if (input.shape[1] * input.shape[2]) % 4 != 0:
                raise RuntimeError("shape '[%d, %d, %d]' is invalid: time*features must be a multiple of 4" % tuple(input.shape))
            from .train import convae_train_forward
            return convae_train_forward(self, input.to(torch.float32).contiguous(), pool_noise,
                                        on_encoder_grad=self.__dict__.get("_on_encoder_grad"))
        x = input.contiguous()
        B, T_in, F = x.shape
        T = T_in if fused_norm is None else int(fused_norm[4])
        if (T * F) % 4 != 0:  # the reference's final reshape fails for these shapes (:197-198)
            raise RuntimeError("shape '[%d, %d, %d]' is invalid for input of size %d: time*features must be a "
                               "multiple of 4" % (B, T, F, B * 4 * (-(-T * F // 4))))
        dev = x.device
        half_out = bool(self.__dict__.get("recon_fp16", False)) and self.precision != "fp32"
        recon = torch.empty((B, T, F), dtype=torch.float16 if half_out else torch.float32, device=dev)
        logp = torch.empty((B, 2), dtype=torch.float32, device=dev)
        if B == 0:
            return recon, logp
        if pool_noise is None:
            pool_noise = self.sex_classifier.stats_pooling._get_gauss_noise((B, 128), device=dev)
        pool_noise = pool_noise.to(device=dev, dtype=torch.float32).contiguous()
        w = self.weight_struct()
        opts = self.opts()
        L = _lib.lib()
        ws_bytes = L.spanon_convae_workspace_bytes(B, T, F, C.byref(opts))
        # The workspace is kept between calls so that the packed (UMMA-layout) weights in it can be reused while no
        # parameter changed: key = every parameter/buffer's (storage pointer, in-place version counter) + the shapes.
        key = (B, T, F, opts.precision, opts.chunk_utts, str(dev),
               tuple((t.data_ptr(), t._version) for t in list(self.parameters()) + list(self.buffers())))
        cache = self.__dict__.get("_ws_cache")
        if cache is None or cache[0].numel() != ws_bytes or cache[0].device != dev:
            cache = [torch.empty((ws_bytes,), dtype=torch.uint8, device=dev), None]
            self.__dict__["_ws_cache"] = cache
        ws = cache[0]
        opts.weights_prepared = 1 if cache[1] == key else 0
        cache[1] = key
        if fused_norm is None:
            mean = std = umax = None
            top_db, t_valid = 0.0, T
        else:
            mean, std, umax, top_db = fused_norm[0], fused_norm[1], fused_norm[2], float(fused_norm[3])
            mean = mean.to(device=dev, dtype=torch.float32).contiguous()
            std = std.to(device=dev, dtype=torch.float32).contiguous()
            t_valid = T_in
        with torch.cuda.device(dev):
            _lib.check(L.spanon_convae_fwd(_lib.ptr(x), C.byref(w), _lib.ptr(pool_noise), _lib.ptr(mean),
                                           _lib.ptr(std), _lib.ptr(umax), top_db, t_valid, _lib.ptr(recon),
                                           _lib.ptr(logp), _lib.ptr(ws), ws_bytes, B, T, F, C.byref(opts),
                                           _lib.stream_ptr(dev)), "convae_fwd")
        return recon, logp


class ConvReconstruction(nn.Module):
    """Drop-in for models.EndToEnd.ConvReconstruction (models/EndToEnd.py:36-87), the "endtoend" anonymizer.

    `encoder` has the reference's layer objects at the reference's indices ({0,3,6,9,12} convolutions, {1,4,7,10}
    InstanceNorms), so checkpoints load unchanged.  The reference builds its `sex_classifier` from a pretrained
    x-vector EncoderClassifier on disk (`EncoderClassifier.from_hparams`, :57-61); here it is any callable
    `feats [B,T,F] -> (log-probs, score, index)` handed to the constructor (None: forward returns None for the logits).

    forward(input [B,T,F]) -> (reconstruction [B,T,F], sex_classifier_logits).
    """

    def __init__(self, sex_classifier=None):
        super().__init__()
        self.encoder = nn.Sequential(
            nn.Conv1d(in_channels=1, out_channels=32, kernel_size=15, stride=1, padding=7),
            nn.InstanceNorm1d(num_features=32, affine=True), GLU(),
            nn.Conv1d(in_channels=32, out_channels=64, kernel_size=5, stride=2, padding=2),
            nn.InstanceNorm1d(num_features=64, affine=True), GLU(),
            nn.Conv1d(in_channels=64, out_channels=64, kernel_size=5, stride=1, padding=2),
            nn.InstanceNorm1d(num_features=64, affine=True), GLU(),
            nn.ConvTranspose1d(in_channels=64, out_channels=32, kernel_size=5, stride=2, padding=2, output_padding=1),
            nn.InstanceNorm1d(num_features=32, affine=True), GLU(),
            nn.Conv1d(in_channels=32, out_channels=1, kernel_size=15, stride=1, padding=7),
        )
        self.sex_classifier = sex_classifier
        self.precision = _precision_from_env()
        self.chunk_utts = 0

    def weight_struct(self):
        e = self.encoder
        w = _lib.ConvReconWeights()
        for i, k in enumerate((0, 3, 6, 9, 12)):
            w.conv_w[i], w.conv_b[i] = _p(e[k].weight), _p(e[k].bias)
        for i, k in enumerate((1, 4, 7, 10)):
            w.in_w[i], w.in_b[i] = _p(e[k].weight), _p(e[k].bias)
        return w

    def __getstate__(self):
        d = dict(self.__dict__)
        d.pop("_ws_cache", None)
        return d

    def opts(self):
        return _lib.ConvAEOpts(_lib.PREC_FP32 if self.precision == "fp32" else _lib.PREC_FP16, int(self.chunk_utts), 1, 0, 0)

    def forward(self, input, fused_norm=None):
        """`fused_norm` = (mean [F], std [F], utt_max [B], top_db, T_pad) as in ConvAutoencoder.forward."""
        _lib.require_cuda(input, "input")
        if torch.is_grad_enabled() and (input.requires_grad or any(p.requires_grad for p in self.encoder.parameters())):
            raise NotImplementedError("ConvReconstruction: only the forward pass runs on the CUDA kernels; call under "
                                      "torch.no_grad() (the backward pass is scope row 8(f)-1)")
        if input.dim() != 3:
            raise ValueError("ConvReconstruction expects [batch, time, features]")
        x = input.contiguous()
        B, T_in, F = x.shape
        T = T_in if fused_norm is None else int(fused_norm[4])
        if (T * F) % 2 != 0:   # the reference's reshape at EndToEnd.py:78 fails: the stride-2 pair returns T*F + 1 values
            raise RuntimeError("shape '[%d, %d, %d]' is invalid for input of size %d: time*features must be even"
                               % (B, T, F, B * (T * F + 1)))
        dev = x.device
        recon = torch.empty((B, T, F), dtype=torch.float32, device=dev)
        if B > 0:
            w = self.weight_struct()
            opts = self.opts()
            L = _lib.lib()
            ws_bytes = L.spanon_convrecon_workspace_bytes(B, T, F, C.byref(opts))
            # workspace (and the packed weights in it) kept between calls, as in ConvAutoencoder.forward
            key = (B, T, F, opts.precision, opts.chunk_utts, str(dev),
                   tuple((t.data_ptr(), t._version) for t in self.encoder.parameters()))
            cache = self.__dict__.get("_ws_cache")
            if cache is None or cache[0].numel() != ws_bytes or cache[0].device != dev:
                cache = [torch.empty((ws_bytes,), dtype=torch.uint8, device=dev), None]
                self.__dict__["_ws_cache"] = cache
            ws = cache[0]
            opts.weights_prepared = 1 if cache[1] == key else 0
            cache[1] = key
            if fused_norm is None:
                mean = std = umax = None
                top_db, t_valid = 0.0, T
            else:
                mean, std, umax, top_db = fused_norm[0], fused_norm[1], fused_norm[2], float(fused_norm[3])
                mean = mean.to(device=dev, dtype=torch.float32).contiguous()
                std = std.to(device=dev, dtype=torch.float32).contiguous()
                t_valid = T_in
            with torch.cuda.device(dev):
                _lib.check(L.spanon_convrecon_fwd(_lib.ptr(x), C.byref(w), _lib.ptr(mean), _lib.ptr(std), _lib.ptr(umax),
                                                  top_db, t_valid, _lib.ptr(recon), _lib.ptr(ws), ws_bytes, B, T, F,
                                                  C.byref(opts), _lib.stream_ptr(dev)), "convrecon_fwd")
        logits = None
        if self.sex_classifier is not None:   # EndToEnd.py:81
            logits, _score, _index = self.sex_classifier(recon)
        return recon, logits


class _SbConv1d(nn.Module):
    """Parameter container with the key layout of speechbrain.nnet.CNN.Conv1d (`conv.weight`, `conv.bias`)."""

    def __init__(self, cin, cout, k, dilation):
        super().__init__()
        self.conv = nn.Conv1d(cin, cout, k, dilation=dilation, padding=0)


class _SbBatchNorm1d(nn.Module):
    """speechbrain.nnet.normalization.BatchNorm1d (`norm.*`)."""

    def __init__(self, n):
        super().__init__()
        self.norm = nn.BatchNorm1d(n)


class _SbLinear(nn.Module):
    """speechbrain.nnet.linear.Linear (`w.*`)."""

    def __init__(self, n_in, n_out):
        super().__init__()
        self.w = nn.Linear(n_in, n_out)


class Xvector(nn.Module):
    """Parameter container for speechbrain.lobes.models.Xvector.Xvector (Xvector.py:45-93) in the configuration of
    evaluator_inference.yaml:34-41; `state_dict()` keys = the reference's embedding_model.ckpt (`blocks.{0,3,6,9,12}.conv.*`,
    `blocks.{2,5,8,11,14}.norm.*`, `blocks.16.w.*`).  forward() = the embedding part of spanon_xvector_classify."""

    def __init__(self, device="cpu", activation=nn.LeakyReLU, tdnn_blocks=5, tdnn_channels=[512, 512, 512, 512, 1500],
                 tdnn_kernel_sizes=[5, 3, 3, 1, 1], tdnn_dilations=[1, 2, 3, 1, 1], lin_neurons=512, in_channels=40):
        # (defaults = the reference's, Xvector.py:45-56, so a default-constructed module holds a default-constructed
        # reference checkpoint; evaluator_inference.yaml:34-41 passes in_channels=80, lin_neurons=128, which is what the
        # CUDA path computes -- XvectorClassifier checks that when it runs)
        super().__init__()
        if (activation is not nn.LeakyReLU or tdnn_blocks != 5 or list(tdnn_channels) != [512, 512, 512, 512, 1500]
                or list(tdnn_kernel_sizes) != [5, 3, 3, 1, 1] or list(tdnn_dilations) != [1, 2, 3, 1, 1]):
            raise NotImplementedError("the CUDA x-vector path is built for evaluator_inference.yaml's TDNN "
                                      "(512,512,512,512,1500; k 5,3,3,1,1; dilation 1,2,3,1,1; LeakyReLU)")
        self.lin_neurons, self.in_channels = lin_neurons, in_channels
        self.blocks = nn.ModuleList()
        cin = in_channels
        for c, k, d in zip(tdnn_channels, tdnn_kernel_sizes, tdnn_dilations):
            self.blocks.extend([_SbConv1d(cin, c, k, d), activation(), _SbBatchNorm1d(c)])
            cin = c
        self.blocks.append(StatisticsPooling())
        self.blocks.append(_SbLinear(2 * cin, lin_neurons))

    def __getstate__(self):
        d = dict(self.__dict__)
        d.pop("_ws_cache", None)
        return d

    @torch.no_grad()
    def forward(self, x, lens=None, pool_noise=None):
        """Xvector.py:95-108: features [B, T, 80] (+ relative lengths) -> x-vectors [B, 1, lin_neurons]."""
        return _xvector_run(self, self, None, x, lens, pool_noise)[0]


class _SbDNNBlock(nn.Module):
    def __init__(self, n_in, n_out, activation):
        super().__init__()
        self.linear = _SbLinear(n_in, n_out)
        self.act = activation()
        self.norm = _SbBatchNorm1d(n_out)


class _SbDNN(nn.Module):
    def __init__(self, n_in, n_out, activation):
        super().__init__()
        self.block_0 = _SbDNNBlock(n_in, n_out, activation)


class XvectorHead(nn.Module):
    """Parameter container for speechbrain.lobes.models.Xvector.Classifier (Xvector.py:111-172) as configured by
    evaluator_inference.yaml:43-48; keys = classifier.ckpt (`norm.norm.*`, `DNN.block_0.linear.w.*`,
    `DNN.block_0.norm.norm.*`, `out.w.*`)."""

    def __init__(self, input_shape=(None, None, 128), activation=nn.LeakyReLU, lin_blocks=1, lin_neurons=128,
                 out_neurons=2):
        super().__init__()
        if activation is not nn.LeakyReLU or lin_blocks != 1 or lin_neurons != 128 or out_neurons != 2 or input_shape[-1] != 128:
            raise NotImplementedError("the CUDA classifier head is built for evaluator_inference.yaml (128 -> 128 -> 2)")
        self.act = activation()
        self.norm = _SbBatchNorm1d(128)
        self.DNN = _SbDNN(128, 128, activation)
        self.out = _SbLinear(128, out_neurons)


def _xvector_weight_struct(e, c):
    """SpanonXvectorWeights of an Xvector `e` and (optionally) its Classifier head `c`."""
    w = _lib.XvectorWeights()
    for i in range(5):
        w.tdnn_w[i], w.tdnn_b[i] = _p(e.blocks[3 * i].conv.weight), _p(e.blocks[3 * i].conv.bias)
        w.tdnn_bn[i] = _bn(e.blocks[3 * i + 2].norm)
    w.emb_w, w.emb_b = _p(e.blocks[16].w.weight), _p(e.blocks[16].w.bias)
    if c is not None:
        w.clf_norm = _bn(c.norm.norm)
        w.dnn_w, w.dnn_b = _p(c.DNN.block_0.linear.w.weight), _p(c.DNN.block_0.linear.w.bias)
        w.dnn_bn = _bn(c.DNN.block_0.norm.norm)
        w.out_w, w.out_b = _p(c.out.w.weight), _p(c.out.w.bias)
    return w


def _xvector_run(owner, emb_model, classifier, feats, wav_lens, pool_noise):
    """spanon_xvector_classify for `owner` (the module that keeps the workspace / packed-weight cache): returns
    (embeddings [B,1,128], log-probs [B,2] or None when `classifier` is None)."""
    _lib.require_cuda(feats, "feats")
    _forward_only_guard(owner, feats)
    if feats.dim() != 3 or feats.shape[2] != 80:
        raise ValueError("%s expects [batch, time, 80] features" % type(owner).__name__)
    if emb_model.lin_neurons != 128 or emb_model.in_channels != 80:
        raise NotImplementedError("the CUDA x-vector path computes evaluator_inference.yaml's model "
                                  "(in_channels=80, lin_neurons=128)")
    x = feats.contiguous()
    B, T, F = x.shape
    dev = x.device
    emb = torch.empty((B, 1, 128), dtype=torch.float32, device=dev)
    logp = torch.empty((B, 2), dtype=torch.float32, device=dev) if classifier is not None else None
    if B == 0:
        return emb, logp
    if wav_lens is not None:
        wav_lens = wav_lens.to(device=dev, dtype=torch.float32).contiguous()
    if pool_noise is None:
        pool_noise = emb_model.blocks[15]._get_gauss_noise((B, 1500), device=dev)
    pool_noise = pool_noise.to(device=dev, dtype=torch.float32).contiguous()
    w = _xvector_weight_struct(emb_model, classifier)
    L = _lib.lib()
    ws_bytes = L.spanon_xvector_workspace_bytes(B, T)
    key = (B, T, str(dev), tuple((t.data_ptr(), t._version) for t in list(owner.parameters()) + list(owner.buffers())))
    cache = owner.__dict__.get("_ws_cache")
    if cache is None or cache[0].numel() != ws_bytes or cache[0].device != dev:
        cache = [torch.empty((ws_bytes,), dtype=torch.uint8, device=dev), None]
        owner.__dict__["_ws_cache"] = cache
    prepared = 1 if cache[1] == key else 0
    cache[1] = key
    with torch.cuda.device(dev):
        _lib.check(L.spanon_xvector_classify(_lib.ptr(x), _lib.ptr(wav_lens), C.byref(w), _lib.ptr(pool_noise),
                                             _lib.ptr(emb), _lib.ptr(logp), _lib.ptr(cache[0]), ws_bytes, B, T, F,
                                             prepared, _lib.stream_ptr(dev)), "xvector_classify")
    return emb, logp


class XvectorClassifier(nn.Module):
    """The external gender classifier of the evaluation configs: drop-in for the parts of
    speechbrain.pretrained.EncoderClassifier the reference uses on FEATURES (interfaces.py:762-776, 825-829) with
    `mods.embedding_model` = Xvector and `mods.classifier` = Classifier of evaluator_inference.yaml.

    classify_batch_feats(feats [B,T,80], wav_lens) -> (log-probs [B,2], score [B], index [B]);
    encode_batch_feat(feats, wav_lens) -> embeddings [B,1,128].  `pool_noise` [B,1500] fixes StatisticsPooling's noise.
    Usable as ConvReconstruction(sex_classifier=XvectorClassifier(...).classify_batch_feats)."""

    def __init__(self, embedding_model=None, classifier=None):
        super().__init__()
        self.embedding_model = (embedding_model if embedding_model is not None
                                else Xvector(in_channels=80, lin_neurons=128))   # evaluator_inference.yaml:34-41
        self.classifier = classifier if classifier is not None else XvectorHead()

    def weight_struct(self):
        return _xvector_weight_struct(self.embedding_model, self.classifier)

    def __getstate__(self):
        d = dict(self.__dict__)
        d.pop("_ws_cache", None)
        return d

    def _run(self, feats, wav_lens, pool_noise):
        return _xvector_run(self, self.embedding_model, self.classifier, feats, wav_lens, pool_noise)

    @torch.no_grad()
    def encode_batch_feat(self, feats, wav_lens=None, normalize=False, pool_noise=None):
        """interfaces.py:762-776 (normalize=True needs hparams.mean_var_norm_emb, which evaluator_inference.yaml lacks)."""
        if normalize:
            raise NotImplementedError("mean_var_norm_emb is not part of evaluator_inference.yaml")
        return self._run(feats, wav_lens, pool_noise)[0]

    @torch.no_grad()
    def classify_batch_feats(self, feat, wav_lens=None, pool_noise=None):
        """interfaces.py:825-829."""
        _, out_prob = self._run(feat, wav_lens, pool_noise)
        score, index = torch.max(out_prob, dim=-1)
        return out_prob, score, index

    def forward(self, feats, wav_lens=None):
        return self.classify_batch_feats(feats, wav_lens)


class FullyConnSexClassifier(nn.Module):
    """models/FullyConnected.py:65-104.  Inside FullyConnectedAutoencoder.forward the classifier runs fused in spanon_fcae_fwd;
    called on its own, forward(z [B, T, 20]) -> log-probs [B, 2] runs spanon_fcae_classify (eval-mode BatchNorm, forward only)."""

    def __init__(self, num_classes):
        super().__init__()
        self.initial = nn.Sequential(nn.Linear(20, 40), nn.ReLU(), nn.Linear(40, 40), nn.ReLU())
        self.norm = nn.BatchNorm1d(20)
        self.classify = nn.Sequential(
            nn.Linear(80, 40), nn.BatchNorm1d(40), nn.ReLU(),
            nn.Linear(40, 40), nn.ReLU(),
            nn.Linear(40, 20), nn.BatchNorm1d(20),
            nn.Linear(20, num_classes),
        )
        self.stats_pooling = StatisticsPooling()

    def fill_weight_struct(self, w):
        s = self
        w.clf_norm = _bn(s.norm)
        for i, k in enumerate((0, 2)):
            w.init_w[i], w.init_b[i] = _p(s.initial[k].weight), _p(s.initial[k].bias)
        for i, k in enumerate((0, 3, 5, 7)):
            w.cls_w[i], w.cls_b[i] = _p(s.classify[k].weight), _p(s.classify[k].bias)
        w.cls_bn1, w.cls_bn6 = _bn(s.classify[1]), _bn(s.classify[6])
        return w

    def forward(self, input, pool_noise=None):
        _lib.require_cuda(input, "input")
        _forward_only_guard(self, input)
        if input.dim() != 3 or input.shape[2] != 20:
            raise RuntimeError("FullyConnSexClassifier expects [batch, time, 20], got %s" % (tuple(input.shape),))
        if self.training:
            raise NotImplementedError("FullyConnSexClassifier: the CUDA path is the eval-mode forward (running statistics)")
        z = input.float().contiguous()
        B, T, _ = z.shape
        dev = z.device
        logp = torch.empty((B, 2), dtype=torch.float32, device=dev)
        if B == 0:
            return logp
        if T == 0:
            raise RuntimeError("FullyConnSexClassifier: empty time axis")
        if pool_noise is None:
            pool_noise = self.stats_pooling._get_gauss_noise((B, 40), device=dev)
        pool_noise = pool_noise.to(device=dev, dtype=torch.float32).contiguous()
        w = _lib.FcaeWeights()
        w.in_dim = 80
        self.fill_weight_struct(w)
        L = _lib.lib()
        ws_bytes = L.spanon_fcae_workspace_bytes(B, T)
        ws = torch.empty((ws_bytes,), dtype=torch.uint8, device=dev)
        with torch.cuda.device(dev):
            _lib.check(L.spanon_fcae_classify(_lib.ptr(z), C.byref(w), _lib.ptr(pool_noise), _lib.ptr(logp), _lib.ptr(ws),
                                              ws_bytes, B, T, _lib.stream_ptr(dev)), "fcae_classify")
        return logp


class FullyConnectedAutoencoder(nn.Module):
    """Drop-in for models.FullyConnected.FullyConnectedAutoencoder (models/FullyConnected.py:118-159).

    forward(input [B,T,mfcc_feature_dim]) -> (reconstruction [B,T,80], sex log-probs [B,2]); fp32.
    """

    def __init__(self, mfcc_feature_dim, batch_size):
        super().__init__()
        self.mfcc_feature_dim = mfcc_feature_dim
        self.batch_size = batch_size
        self.encoder = nn.Sequential(
            nn.Linear(in_features=self.mfcc_feature_dim, out_features=60), nn.ReLU(),
            nn.Linear(in_features=60, out_features=40), nn.ReLU(),
            nn.Linear(in_features=40, out_features=20),
        )
        self.decoder = nn.Sequential(
            nn.Linear(in_features=20, out_features=40), nn.ReLU(),
            nn.Linear(in_features=40, out_features=60), nn.ReLU(),
            nn.Linear(in_features=60, out_features=80),
        )
        self.sex_classifier = FullyConnSexClassifier(2)

    def weight_struct(self):
        e, d, s = self.encoder, self.decoder, self.sex_classifier
        w = _lib.FcaeWeights()
        w.in_dim = self.mfcc_feature_dim
        for i, k in enumerate((0, 2, 4)):
            w.enc_w[i], w.enc_b[i] = _p(e[k].weight), _p(e[k].bias)
            w.dec_w[i], w.dec_b[i] = _p(d[k].weight), _p(d[k].bias)
        return s.fill_weight_struct(w)

    def forward(self, input, pool_noise=None, fused_norm=None):
        """`fused_norm` = (mean [F], std [F], utt_max [B], top_db): input is RAW Fbank dB."""
        _lib.require_cuda(input, "input")
        _forward_only_guard(self, input)
        if input.dim() != 3 or input.shape[2] != self.mfcc_feature_dim:
            raise RuntimeError("mat1 and mat2 shapes cannot be multiplied (expected last dim %d, got %s)"
                               % (self.mfcc_feature_dim, tuple(input.shape)))
        if self.mfcc_feature_dim != 80:
            raise NotImplementedError("the fused FC-AE kernel covers mfcc_feature_dim == 80 (all reference configs)")
        x = input.contiguous()
        B, T, F = x.shape
        dev = x.device
        recon = torch.empty((B, T, 80), dtype=torch.float32, device=dev)
        logp = torch.empty((B, 2), dtype=torch.float32, device=dev)
        if B == 0 or T == 0:
            return recon, logp
        if pool_noise is None:
            pool_noise = self.sex_classifier.stats_pooling._get_gauss_noise((B, 40), device=dev)
        pool_noise = pool_noise.to(device=dev, dtype=torch.float32).contiguous()
        w = self.weight_struct()
        L = _lib.lib()
        ws_bytes = L.spanon_fcae_workspace_bytes(B, T)
        ws = torch.empty((ws_bytes,), dtype=torch.uint8, device=dev)
        if fused_norm is None:
            mean = std = umax = None
            top_db = 0.0
        else:
            mean = fused_norm[0].to(device=dev, dtype=torch.float32).contiguous()
            std = fused_norm[1].to(device=dev, dtype=torch.float32).contiguous()
            umax, top_db = fused_norm[2], float(fused_norm[3])
        with torch.cuda.device(dev):
            _lib.check(L.spanon_fcae_fwd(_lib.ptr(x), C.byref(w), _lib.ptr(pool_noise), _lib.ptr(mean), _lib.ptr(std),
                                         _lib.ptr(umax), top_db, _lib.ptr(recon), _lib.ptr(logp), _lib.ptr(ws),
                                         ws_bytes, B, T, _lib.stream_ptr(dev)), "fcae_fwd")
        return recon, logp
