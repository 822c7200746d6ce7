"""Design study (CPU, torch): where does 16-bit error in the ConvAutoencoder come from?
Emulates the kernel's rounding points on top of the oracle's functional model:
  W   : conv weights rounded to `wt`
  A   : the MMA A operand (post InstanceNorm + SiLU activations) rounded to `at`
  P   : the stored raw conv outputs (pre-norm planes) rounded to `pt`
and reports max|err|/max|ref| of the reconstruction (BASELINE.json's parity metric)."""
import sys, os
import numpy as np, torch, torch.nn.functional as F
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests", "golden"))
import recipe
from oracle import anonymizer_ref as R

def rnd(x, t):
    return x if t is None else x.to(t).to(torch.float32)

def fwd(x, w, wt, at, pt):
    W = lambda k: rnd(w[k], wt) if k.endswith("weight") and w[k].dim() == 3 else w[k]
    inorm = lambda x, p: R.instance_norm(x, w, p)
    a = rnd(R.silu_glu(F.conv1d(x, w["encoder.0.weight"], w["encoder.0.bias"], padding=7)), at)
    p = rnd(F.conv1d(a, W("encoder.2.weight"), w["encoder.2.bias"], stride=2, padding=2), pt)
    a = rnd(R.silu_glu(inorm(p, "encoder.3")), at)
    p = rnd(F.conv1d(a, W("encoder.5.weight"), w["encoder.5.bias"], padding=2), pt)
    a = rnd(R.silu_glu(inorm(p, "encoder.6")), at)
    p = rnd(F.conv1d(a, W("encoder.8.weight"), w["encoder.8.bias"], stride=2, padding=2), pt)
    a = rnd(R.silu_glu(inorm(p, "encoder.9")), at)
    p = rnd(F.conv1d(a, W("encoder.11.weight"), w["encoder.11.bias"], padding=2), pt)
    a = rnd(R.silu_glu(inorm(p, "encoder.12")), at)
    p = rnd(F.conv1d(a, W("decoder.0.weight"), w["decoder.0.bias"], padding=2), pt)
    p = rnd(F.conv_transpose1d(rnd(p, at), W("decoder.1.weight"), w["decoder.1.bias"], stride=2, padding=2, output_padding=1), pt)
    a = rnd(R.silu_glu(inorm(p, "decoder.2")), at)
    p = rnd(F.conv1d(a, W("decoder.4.weight"), w["decoder.4.bias"], padding=2), pt)
    p = rnd(F.conv_transpose1d(rnd(p, at), W("decoder.5.weight"), w["decoder.5.bias"], stride=2, padding=2, output_padding=1), pt)
    a = rnd(R.silu_glu(inorm(p, "decoder.6")), at)
    return F.conv1d(a, W("decoder.8.weight"), w["decoder.8.bias"], padding=7)

if __name__ == "__main__":
    torch.set_num_threads(8)
    bf, hf = torch.bfloat16, torch.float16
    for wseed, T in ((8886, 108), (8886, 1008), (1, 108)):
        w = {k: torch.as_tensor(v) for k, v in recipe.make_weights(recipe.CONVAE_SHAPES, wseed).items()}
        x = torch.from_numpy(recipe.make_feats(77, 2, T)).reshape(2, 1, -1)
        with torch.no_grad():
            ref = fwd(x, w, None, None, None)
            for name, cfg in (("all bf16", (bf, bf, bf)), ("W,A bf16 / planes fp16", (bf, bf, hf)), ("all fp16", (hf, hf, hf)),
                              ("only W bf16", (bf, None, None)), ("only A bf16", (None, bf, None)), ("only P bf16", (None, None, bf)),
                              ("only P fp16", (None, None, hf))):
                out = fwd(x, w, *cfg)
                print("seed %d T %4d  %-24s rel max err %.3e" % (wseed, T, name, float((out - ref).abs().max() / ref.abs().max())))
