"""Each training primitive of csrc/train.cu through the C ABI (the autograd nodes of speech_anonymization_b200/train.py)
against the same operation written with torch ops in DOUBLE precision on the GPU, at the sizes of a 10 s utterance
(rows up to 80 640), where every reduction spans several CTAs.  Tolerance 2e-4 on outputs and gradients
(max|a - ref| / max|ref|): fp32 kernels against an fp64 reference."""
import numpy as np
import pytest
import torch
import torch.nn.functional as F

from speech_anonymization_b200 import train as T

pytestmark = pytest.mark.gpu
DEV = "cuda"
TOL = 2e-4


def rel(a, ref):
    return float((a.double() - ref.double()).abs().max() / ref.double().abs().max().clamp_min(1e-30))


def rnd(*shape, seed=0, scale=1.0):
    g = torch.Generator().manual_seed(seed)
    return (scale * torch.randn(*shape, generator=g)).to(DEV)


@pytest.mark.parametrize("kind,cin,cout,k,dil,pad,rows", [
    (0, 64, 64, 5, 1, 2, 40320), (0, 128, 128, 5, 1, 2, 20160), (0, 128, 128, 3, 2, 0, 20156), (0, 128, 128, 3, 3, 0, 20152),
    (1, 32, 64, 5, 1, 2, 80640), (1, 64, 128, 5, 1, 2, 40320), (2, 128, 64, 5, 1, 2, 20160), (2, 64, 32, 5, 1, 2, 40320),
    (0, 32, 1, 15, 1, 7, 80640), (0, 64, 64, 5, 1, 2, 37)])
def test_tapconv_fwd_bwd_vs_torch(kind, cin, cout, k, dil, pad, rows):
    B = 3
    x = rnd(B, rows, cin, seed=1).requires_grad_(True)
    wshape = (cin, cout, k) if kind == 2 else (cout, cin, k)
    w = rnd(*wshape, seed=2, scale=(cin * k) ** -0.5).requires_grad_(True)
    b = rnd(cout, seed=3, scale=0.1).requires_grad_(True)
    out, stats = T._TapConv.apply(x, w, b, (kind, cin, cout, k, dil, pad), 0, cout > 1)
    xd, wd, bd = (t.detach().double().requires_grad_(True) for t in (x, w, b))
    xc = xd.transpose(1, 2)
    if kind == 0:
        ref = F.conv1d(xc, wd, bd, padding=pad, dilation=dil)
    elif kind == 1:
        ref = F.conv1d(xc, wd, bd, stride=2, padding=2)
    else:
        ref = F.conv_transpose1d(xc, wd, bd, stride=2, padding=2, output_padding=1)
    ref_cl = ref.transpose(1, 2)
    got = out if cout > 1 else out.unsqueeze(2)
    assert got.shape == ref_cl.shape
    assert rel(got, ref_cl) < TOL
    if cout > 1:
        assert rel(stats[..., 0], ref_cl.sum(1)) < TOL and rel(stats[..., 1], (ref_cl ** 2).sum(1)) < TOL
    g = rnd(*got.shape, seed=4)
    (got * g).sum().backward()
    (ref_cl * g.double()).sum().backward()
    assert rel(x.grad, xd.grad) < TOL
    assert rel(w.grad, wd.grad) < TOL
    assert rel(b.grad, bd.grad) < TOL


def test_conv1_fwd_bwd_vs_torch():
    B, L = 3, 80640
    x = rnd(B, L, seed=1)
    w = rnd(32, 1, 15, seed=2, scale=0.25).requires_grad_(True)
    b = rnd(32, seed=3, scale=0.1).requires_grad_(True)
    out = T._Conv1.apply(x, w, b)
    wd, bd = w.detach().double().requires_grad_(True), b.detach().double().requires_grad_(True)
    ref = F.conv1d(x.double().unsqueeze(1), wd, bd, padding=7).transpose(1, 2)
    assert rel(out, ref) < TOL
    g = rnd(B, L, 32, seed=4)
    (out * g).sum().backward()
    (ref * g.double()).sum().backward()
    assert rel(w.grad, wd.grad) < TOL and rel(b.grad, bd.grad) < TOL


@pytest.mark.parametrize("per_batch,act,relu_mask,C,rows", [(0, 1, 0, 64, 40320), (0, 1, 0, 32, 80640), (0, 1, 0, 128, 20160),
                                                            (1, 0, 0, 128, 20160), (1, 0, 1, 128, 20156), (0, 1, 0, 64, 50)])
def test_norm_act_fwd_bwd_vs_torch(per_batch, act, relu_mask, C, rows):
    B = 4
    z0 = rnd(B, rows, C, seed=5) * 1.3 + 0.2
    gamma = (1.0 + 0.1 * rnd(C, seed=6)).requires_grad_(True)
    beta = (0.1 * rnd(C, seed=7)).requires_grad_(True)
    z = (F.relu(z0) if relu_mask else z0).detach().requires_grad_(True)     # relu_mask: z is a stored ReLU output
    stats = T.colstats(z.detach())
    rm, rv = torch.zeros(C, device=DEV), torch.ones(C, device=DEV)
    out = T._NormAct.apply(z, stats, gamma, beta, per_batch, act, relu_mask, rm if per_batch else None, rv if per_batch else None, 0.1, 1e-5, None)
    pre = z0.detach().double().requires_grad_(True)
    zd = F.relu(pre) if relu_mask else pre
    gd, bd = gamma.detach().double().requires_grad_(True), beta.detach().double().requires_grad_(True)
    zc = zd.transpose(1, 2)
    if per_batch:
        rmd, rvd = torch.zeros(C, device=DEV, dtype=torch.float64), torch.ones(C, device=DEV, dtype=torch.float64)
        y = F.batch_norm(zc, rmd, rvd, gd, bd, True, 0.1, 1e-5)
        assert rel(rm, rmd) < TOL and rel(rv, rvd) < TOL
    else:
        y = F.instance_norm(zc, weight=gd, bias=bd, eps=1e-5)
    if act:
        y = y * torch.sigmoid(y)
    ref = y.transpose(1, 2)
    assert rel(out, ref) < TOL
    g = rnd(B, rows, C, seed=8)
    (out * g).sum().backward()
    (ref * g.double()).sum().backward()
    assert rel(z.grad, pre.grad) < TOL          # relu_mask: the returned gradient is the one in FRONT of the ReLU
    assert rel(gamma.grad, gd.grad) < TOL and rel(beta.grad, bd.grad) < TOL
    # sums over rows of the input gradient (what a bias gradient in front of this layer is made of)
    assert float((z.grad.double().sum((0, 1)) - pre.grad.sum((0, 1))).abs().max()) < TOL * float(pre.grad.abs().sum((0, 1)).max())


def test_silu_only_fwd_bwd_vs_torch():
    z = rnd(2, 5000, 32, seed=9).requires_grad_(True)
    out = T._NormAct.apply(z, None, None, None, 0, 1, 0, None, None, 0.0, 0.0, None)
    zd = z.detach().double().requires_grad_(True)
    ref = zd * torch.sigmoid(zd)
    g = rnd(2, 5000, 32, seed=10)
    (out * g).sum().backward()
    (ref * g.double()).sum().backward()
    assert rel(out, ref) < TOL and rel(z.grad, zd.grad) < TOL


@pytest.mark.parametrize("rows", [20146, 706, 130])
def test_pool_reinterp_fwd_bwd_vs_torch(rows):
    B, C = 3, 128
    x = (rnd(B, rows, C, seed=11) + 0.3).requires_grad_(True)
    noise = 1e-5 * rnd(B, C, seed=12).abs()
    out = T._PoolReinterp.apply(x, noise)
    xd = x.detach().double().requires_grad_(True)
    r = xd.transpose(1, 2).contiguous().reshape(B, rows, C)        # ConvAutoEncoder.py:61 on the channels-first tensor
    ref = torch.cat((r.mean(dim=1) + noise.double(), r.std(dim=1) + 1e-5), dim=1)
    assert rel(out, ref) < TOL
    g = rnd(B, 2 * C, seed=13)
    (out * g).sum().backward()
    (ref * g.double()).sum().backward()
    assert rel(x.grad, xd.grad) < TOL
