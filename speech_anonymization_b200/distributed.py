"""Data-parallel training plumbing of the ConvAutoencoder path: one process per GPU, torch.distributed (NCCL over
NVLink / NVSwitch on the GPU box, gloo in the CPU tests).

What the reference does (speechbrain/core.py:1149-1162, utils/distributed.py:107-184): every trainable module is wrapped
in SyncBatchNorm.convert_sync_batchnorm + DistributedDataParallel, i.e. (C1) a bucketed all-reduce of the 533 123 fp32
gradients (2.13 MB) inside loss.backward(), (C2) BatchNorm batch statistics of the sex classifier exchanged between the
ranks in train mode.  Here:

  * FlatGradients    all gradients live in ONE flat fp32 buffer (`p.grad` are views into it), ordered so that the part
                     that is final first during backward -- decoder + classifier, complete when the gradient reaches the
                     bottleneck -- is a contiguous prefix.
  * DataParallelTrainer  arms the model's bottleneck hook: the prefix is all-reduced asynchronously (NCCL's own stream)
                     while the encoder's backward kernels still run, the encoder part right after backward; AVG over
                     ranks like DDP.  With gradient accumulation (speechbrain_convae_train.py:241-247) only the last
                     micro-batch of an optimizer step communicates (`no_sync`).  Sets `model.sync_bn_group`, which makes
                     the training kernels' BatchNorm use global-batch statistics (train.py).
  * gather_for_mi    all-gather of the (features, labels) the MI estimators consume, for a global-batch estimate
                     (the reference computes MI on the local batch, SURVEY.md 8(e) iv).

Messages are tiny (2.1 MB, ~1 KB of BatchNorm sums), i.e. latency-bound: no custom collective kernel is warranted;
torch.distributed's NCCL all-reduce over NVLink is the transport.
"""
import contextlib

import torch
import torch.distributed as dist


def _early_final(name):
    """Parameters whose gradient is final once backward has passed the bottleneck (ConvAutoEncoder.py:178-200)."""
    return not name.startswith("encoder.")


class FlatGradients:
    """One flat fp32 gradient buffer for `model`; `p.grad` become views into it (autograd accumulates in place)."""

    def __init__(self, model):
        named = [(n, p) for n, p in model.named_parameters()]
        early = [(n, p) for n, p in named if _early_final(n)]
        late = [(n, p) for n, p in named if not _early_final(n)]
        self.order = early + late
        self.n_early = sum(p.numel() for _, p in early)
        self.n_total = sum(p.numel() for _, p in self.order)
        dev = named[0][1].device
        self.flat = torch.zeros((self.n_total,), dtype=torch.float32, device=dev)
        off = 0
        self.views = {}
        for n, p in self.order:
            v = self.flat[off:off + p.numel()].view_as(p)
            self.views[n] = v
            p.grad = v
            off += p.numel()

    def early(self):
        return self.flat[:self.n_early]

    def late(self):
        return self.flat[self.n_early:]

    def reattach(self):
        """After optimizer.zero_grad(set_to_none=True) (torch's default): make the views the gradients again."""
        for n, p in self.order:
            if p.grad is None or p.grad.data_ptr() != self.views[n].data_ptr():
                if p.grad is not None:
                    self.views[n].copy_(p.grad)
                p.grad = self.views[n]

    def prepare(self):
        """Call after changing `requires_grad` flags (the reference toggles them per epoch parity,
        speechbrain_convae_train.py:212-235) and before the forward pass: trainable parameters accumulate into their view
        of the flat buffer, frozen ones have grad None so that the optimizer skips them exactly like the reference's."""
        for n, p in self.order:
            if p.requires_grad:
                if p.grad is None or p.grad.data_ptr() != self.views[n].data_ptr():
                    if p.grad is not None:
                        self.views[n].copy_(p.grad)
                    p.grad = self.views[n]
            else:
                p.grad = None

    def zero(self):
        self.flat.zero_()
        self.prepare()


class DataParallelTrainer:
    """Gradient synchronisation of one ConvAutoencoder replica per rank.

        dp = DataParallelTrainer(model)            # after dist.init_process_group
        (toggle requires_grad per epoch parity); dp.prepare()
        for micro-batch i of an optimizer step:
            with dp.no_sync() if i < accumulation - 1 else contextlib.nullcontext():
                loss = ...; (loss / accumulation).backward()
                dp.finish_backward()               # (no-op under no_sync)
        optimizer.step(); dp.zero_grad()
    """

    def __init__(self, model, group=None, sync_bn=True, overlap=True):
        self.model = model
        self.group = group if group is not None else (dist.group.WORLD if dist.is_initialized() else None)
        self.world = dist.get_world_size(self.group) if self.group is not None else 1
        self.grads = FlatGradients(model)
        self.overlap = overlap
        self._sync = True
        self._early_work = None
        self.comm_events = []          # (start, end) CUDA events of every all-reduce, for bench.py's all-reduce share
        if sync_bn and self.world > 1:
            model.sync_bn_group = self.group
        model.__dict__["_on_encoder_grad"] = self._on_encoder_grad

    # ---- hooks
    def _all_reduce(self, t, async_op):
        if self.world == 1:
            return None
        return dist.all_reduce(t, op=dist.ReduceOp.AVG if t.is_cuda else dist.ReduceOp.SUM, group=self.group, async_op=async_op)

    def _on_encoder_grad(self):
        # decoder + classifier gradients are final: start their all-reduce now, it overlaps the encoder's backward
        if self._sync and self.overlap and self.world > 1 and self._early_work is None:
            self._early_work = self._all_reduce(self.grads.early(), async_op=True)

    @contextlib.contextmanager
    def no_sync(self):
        old, self._sync = self._sync, False
        try:
            yield
        finally:
            self._sync = old

    def finish_backward(self):
        """Call after loss.backward(): all-reduces what has not been sent yet and waits for the transfers."""
        if not self._sync or self.world == 1:
            return
        g = self.grads
        g.reattach()
        if self._early_work is None:          # the hook did not fire (frozen encoder: nothing upstream needs a gradient)
            w1 = self._all_reduce(g.early(), async_op=True)
        else:
            w1 = self._early_work
        w2 = self._all_reduce(g.late(), async_op=True)
        w1.wait()
        w2.wait()
        if not g.flat.is_cuda:                # gloo has no AVG
            g.flat.div_(self.world)
        self._early_work = None

    def prepare(self):
        self.grads.prepare()

    def zero_grad(self):
        self.grads.zero()


def gather_for_mi(X, y, group=None):
    """All-gather of the rows the MI estimators consume (utils/mi_loss.py:10-16: X [B, D], labels y [B]) so that
    ClusterMI / GroupSamplingMI can run on the global batch.  Equal B on every rank."""
    if not dist.is_initialized() or dist.get_world_size(group) == 1:
        return X, y
    world = dist.get_world_size(group)
    Xg = torch.empty((world * X.shape[0],) + tuple(X.shape[1:]), dtype=X.dtype, device=X.device)
    yg = torch.empty((world * y.shape[0],) + tuple(y.shape[1:]), dtype=y.dtype, device=y.device)
    if X.is_cuda:
        dist.all_gather_into_tensor(Xg, X.contiguous(), group=group)
        dist.all_gather_into_tensor(yg, y.contiguous(), group=group)
    else:
        dist.all_gather(list(Xg.chunk(world)), X.contiguous(), group=group)
        dist.all_gather(list(yg.chunk(world)), y.contiguous(), group=group)
    return Xg, yg
