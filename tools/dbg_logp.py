"""logp of the golden ConvAE cases vs the stored reference under the A/B switches (debugging aid)."""
import sys
import numpy as np, torch
sys.path.insert(0, "tests/golden")
import recipe
import speech_anonymization_b200 as sab
g = np.load("tests/golden/golden_v1.npz")
m = sab.ConvAutoencoder()
w = recipe.make_weights(recipe.CONVAE_SHAPES, 8886)
m.load_state_dict({k: torch.from_numpy(np.asarray(v)) for k, v in w.items()})
m = m.cuda().eval(); m.precision = "fp16"
for name, (seed, B, T) in {"convae_a": (5, 2, 36), "convae_b": (6, 1, 72)}.items():
    x = torch.from_numpy(recipe.make_feats(seed, B, T)).cuda()
    nz = torch.from_numpy(recipe.make_noise(seed, B, 128)).cuda()
    with torch.no_grad():
        for rep in range(3):
            rec, logp = m(x, pool_noise=nz)
            print(name, rep, float(np.abs(logp.cpu().numpy() - g[name + "_logp"]).max()), logp.cpu().numpy().ravel())
