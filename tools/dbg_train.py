import os, sys
import numpy as np, torch
import torch.nn.functional as F
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests", "golden")); sys.path.insert(0, os.path.join(ROOT, "tests"))
import recipe
import speech_anonymization_b200 as sab
from speech_anonymization_b200.models import GradReverse
from speech_anonymization_b200.train import convae_train_forward
torch.backends.cudnn.allow_tf32 = False; torch.backends.cuda.matmul.allow_tf32 = False

def torch_forward(m, x, noise, taps):
    B, T, Fd = x.shape
    h = m.encoder(x.reshape(B, 1, T * Fd))
    s = m.sex_classifier
    taps["enc"] = h
    c = s.norm(GradReverse.apply(h)); taps["norm"] = c
    for i in (0, 3, 6):
        r = s.tdnn[i + 1](s.tdnn[i](c)); c = s.tdnn[i + 2](r)
        taps["relu%d" % i], taps["bn%d" % (i + 2)] = r, c
    c = c.reshape(c.shape[0], c.shape[2], c.shape[1])
    st = torch.cat((c.mean(dim=1) + noise, c.std(dim=1) + 1e-5), dim=1)
    logp = F.log_softmax(s.classify(st), 1)
    return m.decoder(h).squeeze(1).reshape(B, T, Fd), logp

seed, B, T, wr, ws = 23, 4, 1008, 1.0, 0.8
w = recipe.make_weights(recipe.CONVAE_SHAPES, 8886)
def fresh(dtype):
    m = sab.ConvAutoencoder(); m.load_state_dict({k: torch.from_numpy(np.asarray(v)) for k, v in w.items()}); return m.cuda().to(dtype).train()
x = torch.from_numpy(recipe.make_feats(seed, B, T)).cuda()
noise = torch.from_numpy(recipe.make_noise(seed, B, 128)).cuda()
y = torch.from_numpy((np.arange(B) + seed) % 2).long().cuda()
T_ = {}
for tag in ("ours", "t32", "t64"):
    dt = torch.float64 if tag == "t64" else torch.float32
    m = fresh(dt); taps = {}
    if tag == "ours": rec, logp = convae_train_forward(m, x, noise, _taps=taps)
    else: rec, logp = torch_forward(m, x.to(dt), noise.to(dt), taps)
    for v in taps.values(): v.retain_grad()
    loss = ws * torch.nn.NLLLoss()(logp, y)          # classifier branch only
    loss.backward()
    cl = (lambda v: v) if tag == "ours" else (lambda v: v.transpose(1, 2))
    T_[tag] = {k: (cl(v.detach()).double(), cl(v.grad).double()) for k, v in taps.items()}
for k in T_["ours"]:
    fo, go = T_["ours"][k]; f3, g3 = T_["t32"][k]; f6, g6 = T_["t64"][k]
    rel = lambda a, b: float((a - b).abs().max() / b.abs().max())
    # per-channel sum of the gradient (what a bias gradient is made of)
    so, s3, s6 = go.sum((0, 1)), g3.sum((0, 1)), g6.sum((0, 1))
    print("%-6s fwd ours %.1e t32 %.1e | grad ours %.1e t32 %.1e | grad colsum ours %.1e t32 %.1e" % (
        k, rel(fo, f6), rel(f3, f6), rel(go, g6), rel(g3, g6), float((so - s6).abs().max() / s6.abs().max()), float((s3 - s6).abs().max() / s6.abs().max())))
print("---- gradient in FRONT of the ReLU (ours: stored; torch: relu.grad * (relu > 0))")
for k in ("relu0", "relu3", "relu6"):
    fo, go = T_["ours"][k]; f3, g3 = T_["t32"][k]; f6, g6 = T_["t64"][k]
    m3, m6 = g3 * (f3 > 0), g6 * (f6 > 0)
    so, s3, s6 = go.sum((0, 1)), m3.sum((0, 1)), m6.sum((0, 1))
    sc = s6.abs().max()
    eo, e3 = (so - s6).abs() / sc, (s3 - s6).abs() / sc
    flips_o = int(((fo > 0) != (f6 > 0)).sum()); flips_3 = int(((f3 > 0) != (f6 > 0)).sum())
    print(k, "colsum err ours max %.1e (ch %d) t32 max %.1e | elem err ours %.1e t32 %.1e | mask flips vs t64: ours %d t32 %d" % (
        float(eo.max()), int(eo.argmax()), float(e3.max()), float((go - m6).abs().max() / m6.abs().max()), float((m3 - m6).abs().max() / m6.abs().max()), flips_o, flips_3))
    c = int(eo.argmax())
    print("   ch %d: active frac %.4f, colsum t64 %.4e ours %.4e t32 %.4e ; sum|g| %.3e" % (c, float((f6[..., c] > 0).double().mean()), float(s6[c]), float(so[c]), float(s3[c]), float(m6[..., c].abs().sum())))
    d = (go - m6)[..., c]
    print("   err elems: n>1e-9: %d  sum err %.3e  max %.3e ; err where masks agree: sum %.3e" % (int((d.abs() > 1e-9).sum()), float(d.sum()), float(d.abs().max()), float(d[(fo[..., c] > 0) == (f6[..., c] > 0)].sum())))
