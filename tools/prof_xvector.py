"""Small driver for ncu: x-vector classification of B x 1001 x 80 features (scope row 8(f)-2)."""
import os, sys
import numpy as np, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests", "golden"))
import recipe
import speech_anonymization_b200 as sab
B = int(sys.argv[1]) if len(sys.argv) > 1 else 60
iters = int(sys.argv[2]) if len(sys.argv) > 2 else 3
m = sab.XvectorClassifier()
for mod, shp in ((m.embedding_model, recipe.XVECTOR_EMB_SHAPES), (m.classifier, recipe.XVECTOR_CLF_SHAPES)):
    mod.load_state_dict({k: torch.from_numpy(np.asarray(v)) for k, v in recipe.make_weights(shp, 1230).items()})
m = m.cuda().eval()
x = torch.from_numpy(recipe.make_feats(1, B, 1001)).cuda()
noise = torch.from_numpy(recipe.make_noise(1, B, 1500)).cuda()
for _ in range(iters):
    out, score, index = m.classify_batch_feats(x, pool_noise=noise)
torch.cuda.synchronize()
print("ok", out[0].tolist())
