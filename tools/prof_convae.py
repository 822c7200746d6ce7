"""Small driver for ncu: runs the ConvAutoencoder forward on B utterances of 10 s (fp16 mode)."""
import os, sys
import numpy as np, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests", "golden"))
import recipe
import speech_anonymization_b200 as sab
B = int(sys.argv[1]) if len(sys.argv) > 1 else 6
iters = int(sys.argv[2]) if len(sys.argv) > 2 else 2
m = sab.ConvAutoencoder()
w = recipe.make_weights(recipe.CONVAE_SHAPES, 8886)
m.load_state_dict({k: torch.from_numpy(np.asarray(v)) for k, v in w.items()})
m = m.cuda().eval(); m.precision = "fp16"; m.chunk_utts = B
x = torch.from_numpy(recipe.make_feats(1, B, 1008)).cuda()
noise = torch.from_numpy(recipe.make_noise(1, B, 128)).cuda()
with torch.no_grad():
    for _ in range(iters):
        rec, logp = m(x, pool_noise=noise)
torch.cuda.synchronize()
print("ok", float(rec.abs().mean()))
