"""SPANON_TC_TRACE=1 timeline of the tcgen05 launches on 60 x 10 s (prints to stderr; grep the layer tag)."""
import sys
import numpy as np, torch
sys.path.insert(0, "tests/golden")
import recipe
import speech_anonymization_b200 as sab
m = sab.ConvAutoencoder()
w = recipe.make_weights(recipe.CONVAE_SHAPES, 8886)
m.load_state_dict({k: torch.from_numpy(np.asarray(v)) for k, v in w.items()})
m = m.cuda().eval(); m.precision = "fp16"
x = torch.randn(60, 1008, 80, device="cuda")
with torch.no_grad():
    for _ in range(2):
        r, l = m(x)
torch.cuda.synchronize()
