"""Training-mode forward + backward of the ConvAutoencoder on libspanon's training kernels (csrc/train.cu).

The reference trains through PyTorch autograd over library kernels (`(loss / gradient_accumulation).backward()`,
speechbrain_convae_train.py:238-241, on the graph of models/ConvAutoEncoder.py:178-200).  Here every layer of that
graph is a torch.autograd.Function whose forward AND backward are C-ABI calls (include/spanon.h, "Training step"):
autograd only does the bookkeeping -- which is exactly what makes the reference's per-epoch `requires_grad` toggling
(speechbrain_convae_train.py:212-235) work unchanged: a frozen parameter's weight-gradient kernel is simply not launched
(`ctx.needs_input_grad`), and a branch nothing upstream of needs a gradient is never entered.

Tensors between the nodes are fp32, channels-last [utterance, row, channel] (the kernels' layout), internal to this file.
The only torch arithmetic on the path is the reference's own GradReverse node and the classifier's three-layer head on
the pooled [B, 256] statistics (models/ConvAutoEncoder.py:47-55: 41 k of the 533 k parameters, B x 256 inputs).
"""
import ctypes as C

import torch
import torch.nn.functional as F

from . import _lib


def _call(fn, *args, what=""):
    _lib.check(fn(*args), what)


def _ws(nbytes, dev):
    return torch.empty((max(int(nbytes), 256),), dtype=torch.uint8, device=dev)


class _TapConv(torch.autograd.Function):
    """Conv1d / stride-2 Conv1d / stride-2 ConvTranspose1d as a tap GEMM (spanon_tapconv_fwd / _bwd).
    Returns (out, stats): `stats` [B, cout, 2] fp64 sums of the stored values for the normalisation that follows
    (non-differentiable), or an empty tensor.  relu = 1 stores ReLU(conv + bias); its consumer's backward masks the
    gradient (spanon_norm_act_bwd relu_mask), so the gradient arriving here is already the one in front of the ReLU."""

    @staticmethod
    def forward(ctx, x, w, b, cv, relu, want_stats):
        L = _lib.lib()
        B, rows_in, cin = x.shape
        spec = _lib.TapConv(*cv)
        kind, _, cout, k, dil, pad = cv
        rows_out = rows_in + 2 * pad - dil * (k - 1) if kind == 0 else rows_in // 2 if kind == 1 else rows_in * 2
        dev = x.device
        out = torch.empty((B, rows_out) if cout == 1 else (B, rows_out, cout), dtype=torch.float32, device=dev)
        stats = torch.zeros((B, cout, 2) if want_stats else (0,), dtype=torch.float64, device=dev)
        ws = _ws(L.spanon_tapconv_workspace_bytes(C.byref(spec)), dev)
        with torch.cuda.device(dev):
            _call(L.spanon_tapconv_fwd, _lib.ptr(x), _lib.ptr(w), _lib.ptr(b), _lib.ptr(out),
                  _lib.ptr(stats) if want_stats else None, int(relu), _lib.ptr(ws), ws.numel(), B, rows_in, C.byref(spec),
                  _lib.stream_ptr(dev), what="tapconv_fwd")
        ctx.save_for_backward(x, w)
        ctx.cv = cv
        ctx.has_bias = b is not None
        ctx.mark_non_differentiable(stats)
        return out, stats

    @staticmethod
    def backward(ctx, gout, _gstats):
        L = _lib.lib()
        x, w = ctx.saved_tensors
        B, rows_in, _ = x.shape
        spec = _lib.TapConv(*ctx.cv)
        dev = x.device
        gout = gout.contiguous()
        gx = torch.empty_like(x) if ctx.needs_input_grad[0] else None
        gw = torch.empty_like(w) if ctx.needs_input_grad[1] else None
        gb = torch.empty((ctx.cv[2],), dtype=torch.float32, device=dev) if (ctx.has_bias and ctx.needs_input_grad[2]) else None
        if gx is not None or gw is not None or gb is not None:
            ws = _ws(L.spanon_tapconv_workspace_bytes(C.byref(spec)), dev)
            with torch.cuda.device(dev):
                _call(L.spanon_tapconv_bwd, _lib.ptr(x), _lib.ptr(w), _lib.ptr(gout), _lib.ptr(gx), _lib.ptr(gw), _lib.ptr(gb),
                      _lib.ptr(ws), ws.numel(), B, rows_in, C.byref(spec), _lib.stream_ptr(dev), what="tapconv_bwd")
        return gx, gw, gb, None, None, None


class _Conv1(torch.autograd.Function):
    """encoder.0: Conv1d(1 -> 32, k15, p7) on the flattened signal (spanon_conv1_fwd / _bwd); raw output."""

    @staticmethod
    def forward(ctx, x, w, b):
        L = _lib.lib()
        B, n = x.shape
        out = torch.empty((B, n, 32), dtype=torch.float32, device=x.device)
        with torch.cuda.device(x.device):
            _call(L.spanon_conv1_fwd, _lib.ptr(x), _lib.ptr(w), _lib.ptr(b), _lib.ptr(out), B, n, _lib.stream_ptr(x.device),
                  what="conv1_fwd")
        ctx.save_for_backward(x)
        return out

    @staticmethod
    def backward(ctx, gout):
        if ctx.needs_input_grad[0]:
            raise NotImplementedError("the feature signal needs no gradient on this path (the reference computes Fbank "
                                      "without a graph); encoder.0's input gradient is not built")
        L = _lib.lib()
        (x,) = ctx.saved_tensors
        B, n = x.shape
        gout = gout.contiguous()
        gw = torch.empty((32, 1, 15), dtype=torch.float32, device=x.device) if ctx.needs_input_grad[1] else None
        gb = torch.empty((32,), dtype=torch.float32, device=x.device) if ctx.needs_input_grad[2] else None
        if gw is not None or gb is not None:
            scratch = torch.empty((32,), dtype=torch.float64, device=x.device)
            with torch.cuda.device(x.device):
                _call(L.spanon_conv1_bwd, _lib.ptr(x), _lib.ptr(gout), _lib.ptr(gw), _lib.ptr(gb), _lib.ptr(scratch), B, n,
                      _lib.stream_ptr(x.device), what="conv1_bwd")
        return None, gw, gb


class _NormAct(torch.autograd.Function):
    """InstanceNorm1d(affine) (per_batch = 0) or train-mode BatchNorm1d (per_batch = 1) from the producer's sums, then
    SiLU (act = 1) or nothing; gamma None: activation only (spanon_norm_act_fwd / _bwd).
    `group` (per_batch only): a torch.distributed process group -> SyncBatchNorm semantics (speechbrain/core.py:1156,
    `SyncBatchNorm.convert_sync_batchnorm` under DDP): the per-utterance sums are all-gathered in the forward, the two
    backward sums all-reduced, so the statistics are those of the GLOBAL batch.  Ranks must hold equal batch shapes."""

    @staticmethod
    def forward(ctx, z, stats, gamma, beta, per_batch, act, relu_mask, running_mean, running_var, momentum, eps, group):
        L = _lib.lib()
        B, rows, Cc = z.shape
        dev = z.device
        out = torch.empty_like(z)
        stats_B, count, world = 0, 0.0, 1
        if gamma is not None:
            n_stat = Cc if per_batch else B * Cc
            mean = torch.empty((n_stat,), dtype=torch.float32, device=dev)
            rstd = torch.empty((n_stat,), dtype=torch.float32, device=dev)
            if per_batch and group is not None:
                import torch.distributed as dist
                world = dist.get_world_size(group)
                if world > 1:
                    gathered = torch.empty((world * B, Cc, 2), dtype=torch.float64, device=dev)
                    dist.all_gather_into_tensor(gathered, stats.contiguous(), group=group)
                    stats, stats_B, count = gathered, world * B, float(world * B * rows)
        else:
            mean = rstd = None
        with torch.cuda.device(dev):
            _call(L.spanon_norm_act_fwd, _lib.ptr(z), _lib.ptr(stats) if gamma is not None else None, _lib.ptr(gamma),
                  _lib.ptr(beta), int(per_batch), int(act), float(eps), _lib.ptr(running_mean), _lib.ptr(running_var),
                  float(momentum), _lib.ptr(out), _lib.ptr(mean), _lib.ptr(rstd), B, rows, Cc, stats_B, count,
                  _lib.stream_ptr(dev), what="norm_act_fwd")
        ctx.save_for_backward(z, gamma, beta, mean, rstd)
        ctx.cfg = (int(per_batch), int(act), int(relu_mask))
        ctx.group, ctx.world = (group if world > 1 else None), world
        return out

    @staticmethod
    def backward(ctx, gout):
        L = _lib.lib()
        z, gamma, beta, mean, rstd = ctx.saved_tensors
        per_batch, act, relu_mask = ctx.cfg
        B, rows, Cc = z.shape
        dev = z.device
        gout = gout.contiguous()
        gz = torch.empty_like(z)
        has = gamma is not None
        gg = torch.empty((Cc,), dtype=torch.float32, device=dev) if has and ctx.needs_input_grad[2] else None
        gbeta = torch.empty((Cc,), dtype=torch.float32, device=dev) if has and ctx.needs_input_grad[3] else None
        scratch = torch.empty((B, Cc, 2), dtype=torch.float64, device=dev) if has else None

        def run(phase, count):
            with torch.cuda.device(dev):
                _call(L.spanon_norm_act_bwd, _lib.ptr(z), _lib.ptr(gout), _lib.ptr(gamma), _lib.ptr(beta), _lib.ptr(mean),
                      _lib.ptr(rstd), per_batch, act, relu_mask, _lib.ptr(gz), _lib.ptr(gg), _lib.ptr(gbeta),
                      _lib.ptr(scratch), B, rows, Cc, phase, count, _lib.stream_ptr(dev), what="norm_act_bwd")
        if ctx.group is None:
            run(0, 0.0)
        else:       # SyncBatchNorm: this rank's two sums -> all-reduce -> element-wise pass with the global count
            import torch.distributed as dist
            run(1, 0.0)
            dist.all_reduce(scratch[0], group=ctx.group)
            run(2, float(ctx.world * B * rows))
        return gz, None, gg, gbeta, None, None, None, None, None, None, None, None


class _PoolReinterp(torch.autograd.Function):
    """reshape(B, L, C) reinterpretation + StatisticsPooling (ConvAutoEncoder.py:61-65; spanon_pool_reinterp_fwd / _bwd)."""

    @staticmethod
    def forward(ctx, x, noise):
        L = _lib.lib()
        B, rows, Cc = x.shape
        dev = x.device
        out = torch.empty((B, 2 * Cc), dtype=torch.float32, device=dev)
        save = torch.empty((B, 2 * Cc), dtype=torch.float32, device=dev)
        scratch = torch.empty((B, Cc, 2), dtype=torch.float64, device=dev)
        with torch.cuda.device(dev):
            _call(L.spanon_pool_reinterp_fwd, _lib.ptr(x), _lib.ptr(noise), _lib.ptr(out), _lib.ptr(save), _lib.ptr(scratch),
                  B, rows, Cc, _lib.stream_ptr(dev), what="pool_reinterp_fwd")
        ctx.save_for_backward(x, save)
        return out

    @staticmethod
    def backward(ctx, gout):
        L = _lib.lib()
        x, save = ctx.saved_tensors
        B, rows, Cc = x.shape
        gx = torch.empty_like(x)
        gout = gout.contiguous()
        with torch.cuda.device(x.device):
            _call(L.spanon_pool_reinterp_bwd, _lib.ptr(x), _lib.ptr(save), _lib.ptr(gout), _lib.ptr(gx), B, rows, Cc,
                  _lib.stream_ptr(x.device), what="pool_reinterp_bwd")
        return gx, None


def colstats(x):
    B, rows, Cc = x.shape
    stats = torch.empty((B, Cc, 2), dtype=torch.float64, device=x.device)
    with torch.cuda.device(x.device):
        _call(_lib.lib().spanon_colstats, _lib.ptr(x), _lib.ptr(stats), B, rows, Cc, _lib.stream_ptr(x.device), what="colstats")
    return stats


def _conv(x, m, kind, relu=0, want_stats=False):
    """m: nn.Conv1d / nn.ConvTranspose1d parameter container."""
    if kind == 2:
        cin, cout = m.in_channels, m.out_channels
    else:
        cin, cout = m.in_channels, m.out_channels
    cv = (kind, cin, cout, m.kernel_size[0], m.dilation[0], m.padding[0])
    return _TapConv.apply(x, m.weight, m.bias, cv, relu, want_stats)


def _inorm_silu(z, stats, m):
    return _NormAct.apply(z, stats, m.weight, m.bias, 0, 1, 0, None, None, 0.0, m.eps, None)


def _bnorm(z, stats, m, relu_mask, group=None):
    """train-mode BatchNorm1d on a channels-last tensor (running statistics and num_batches_tracked updated like torch)."""
    if m.momentum is None:
        raise NotImplementedError("BatchNorm1d(momentum=None) (cumulative average) is not used by the reference models")
    out = _NormAct.apply(z, stats, m.weight, m.bias, 1, 0, relu_mask, m.running_mean if m.track_running_stats else None,
                         m.running_var if m.track_running_stats else None, m.momentum, m.eps, group)
    if m.track_running_stats and m.num_batches_tracked is not None:
        m.num_batches_tracked.add_(1)
    return out


def _classify_head(classify, pooled, group):
    """The three-layer head on the pooled [B, 256] statistics (ConvAutoEncoder.py:47-55).  With a process group its two
    BatchNorm1d layers use global-batch statistics like the SyncBatchNorm the reference's DDP wrap converts them to."""
    if group is None:
        return classify(pooled)
    import torch.distributed as dist
    if dist.get_world_size(group) == 1:
        return classify(pooled)
    x = pooled
    for m in classify:
        if isinstance(m, torch.nn.BatchNorm1d):
            x = _sync_bn_rows(x, m, group)
        else:
            x = m(x)
    return x


class _AllReduceSum(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, group):
        import torch.distributed as dist
        ctx.group = group
        y = x.clone()
        dist.all_reduce(y, group=group)
        return y

    @staticmethod
    def backward(ctx, g):
        import torch.distributed as dist
        g = g.clone()
        dist.all_reduce(g, group=ctx.group)
        return g, None


def _sync_bn_rows(x, m, group):
    """train-mode BatchNorm1d on [B, C] rows with the statistics of all ranks' rows (tiny: B x 128 values)."""
    import torch.distributed as dist
    world = dist.get_world_size(group)
    n = x.shape[0] * world
    s = _AllReduceSum.apply(torch.stack((x.sum(0), (x * x).sum(0))), group)
    mean = s[0] / n
    var = (s[1] / n - mean * mean).clamp_min(0.0)
    if m.track_running_stats:
        with torch.no_grad():
            m.running_mean.mul_(1 - m.momentum).add_(m.momentum * mean)
            m.running_var.mul_(1 - m.momentum).add_(m.momentum * var * n / max(n - 1, 1))
            m.num_batches_tracked.add_(1)
    return (x - mean) * torch.rsqrt(var + m.eps) * m.weight + m.bias


def _bnorm_frozen(z, m):
    """eval-mode BatchNorm1d on a channels-last tensor: the running statistics enter spanon_norm_act_fwd as the sums of ONE
    pseudo-sample per channel (sum = mean, sum of squares = var + mean^2, count 1), so the same kernels normalise with them."""
    rm, rv = m.running_mean.double(), m.running_var.double()
    stats = torch.stack((rm, rv + rm * rm), dim=1).reshape(1, -1, 2).contiguous()
    B, rows, Cc = z.shape
    out = torch.empty_like(z)
    mean = torch.empty((Cc,), dtype=torch.float32, device=z.device)
    rstd = torch.empty((Cc,), dtype=torch.float32, device=z.device)
    with torch.cuda.device(z.device):
        _call(_lib.lib().spanon_norm_act_fwd, _lib.ptr(z), _lib.ptr(stats), _lib.ptr(m.weight), _lib.ptr(m.bias), 1, 0, float(m.eps),
              None, None, 0.0, _lib.ptr(out), _lib.ptr(mean), _lib.ptr(rstd), B, rows, Cc, 1, 1.0, _lib.stream_ptr(z.device),
              what="norm_act_fwd")
    return out


def tdnn_classifier_forward(s, enc, pool_noise=None, group=None, _taps=None):
    """TDNNSexClassifier.forward (models/ConvAutoEncoder.py:57-69) on a CHANNELS-LAST bottleneck `enc` [B, L', 128]:
    GradReverse -> BatchNorm -> 3 x (conv, ReLU, BatchNorm) -> reshape + StatisticsPooling -> MLP -> log_softmax.
    Train mode: batch statistics, attached to the autograd graph.  Eval mode: running statistics, forward only."""
    from .models import GradReverse
    B = enc.shape[0]
    if s.training:
        c = GradReverse.apply(enc)
        c = _bnorm(c, colstats(c.detach()), s.norm, 0, group)
    else:
        if torch.is_grad_enabled() and (enc.requires_grad or any(p.requires_grad for p in s.parameters())):
            raise NotImplementedError("TDNNSexClassifier: a backward pass through eval-mode BatchNorm is not built; call "
                                      ".train() for a training step or run under torch.no_grad()")
        c = _bnorm_frozen(enc.contiguous(), s.norm)
    if _taps is not None:      # (debugging aid: intermediate tensors, channels-last)
        _taps["enc"], _taps["norm"] = enc, c
    for i in (0, 3, 6):
        r, st = _conv(c, s.tdnn[i], 0, relu=1, want_stats=s.training)
        c = _bnorm(r, st, s.tdnn[i + 2], 1, group) if s.training else _bnorm_frozen(r, s.tdnn[i + 2])
        if _taps is not None:
            _taps["relu%d" % i], _taps["bn%d" % (i + 2)] = r, c
    if pool_noise is None:
        pool_noise = s.stats_pooling._get_gauss_noise((B, 128), device=enc.device)
    pooled = _PoolReinterp.apply(c, pool_noise.to(device=enc.device, dtype=torch.float32).contiguous())
    return F.log_softmax(_classify_head(s.classify, pooled, group if s.training else None), 1)


def convae_train_forward(model, feats, pool_noise=None, _taps=None, on_encoder_grad=None):
    """ConvAutoencoder.forward in TRAIN mode (batch-statistics BatchNorm in the sex classifier) with a backward pass.
    feats [B, T, F] fp32 CUDA -> (reconstruction [B, T, F], sex log-probs [B, 2]), both attached to the autograd graph."""
    B, T, Fd = feats.shape
    Lf = T * Fd
    e, d, s = model.encoder, model.decoder, model.sex_classifier
    x = feats.reshape(B, Lf).contiguous()
    # ---- encoder (models/ConvAutoEncoder.py:141-157)
    z = _Conv1.apply(x, e[0].weight, e[0].bias)
    a = _NormAct.apply(z, None, None, None, 0, 1, 0, None, None, 0.0, 0.0, None)    # GLU = x * sigmoid(x)
    z, st = _conv(a, e[2], 1, want_stats=True)
    a = _inorm_silu(z, st, e[3])
    z, st = _conv(a, e[5], 0, want_stats=True)
    a = _inorm_silu(z, st, e[6])
    z, st = _conv(a, e[8], 1, want_stats=True)
    a = _inorm_silu(z, st, e[9])
    z, st = _conv(a, e[11], 0, want_stats=True)
    enc = _inorm_silu(z, st, e[12])                                                    # [B, L/4, 128]
    # ---- adversarial sex classifier (:57-69): GradReverse -> BatchNorm -> 3 x (conv, ReLU, BatchNorm) -> pooling -> MLP
    group = getattr(model, "sync_bn_group", None)      # set by distributed.DataParallelTrainer (SyncBatchNorm under DDP)
    if on_encoder_grad is not None and enc.requires_grad:
        # fires when both branches' gradients have reached the bottleneck: every decoder / classifier gradient is final
        enc.register_hook(lambda g: (on_encoder_grad(), g)[1])
    logp = tdnn_classifier_forward(s, enc, pool_noise, group, _taps)
    # ---- decoder (:160-172)
    z, _ = _conv(enc, d[0], 0)
    z, st = _conv(z, d[1], 2, want_stats=True)
    a = _inorm_silu(z, st, d[2])
    z, _ = _conv(a, d[4], 0)
    z, st = _conv(z, d[5], 2, want_stats=True)
    a = _inorm_silu(z, st, d[6])
    out, _ = _conv(a, d[8], 0)                                                         # [B, L]
    return out.reshape(B, T, Fd), logp
