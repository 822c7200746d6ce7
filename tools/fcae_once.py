"""Three FullyConnectedAutoencoder forwards on 64 x 1001 x 80 features (BASELINE configs[1] shape) -- workload for ncu."""
import sys
import numpy as np, torch
sys.path.insert(0, "tests/golden")
import recipe
import speech_anonymization_b200 as sab
m = sab.FullyConnectedAutoencoder(80, 64)
w = recipe.make_weights(recipe.FCAE_SHAPES, 8886)
m.load_state_dict({k: torch.from_numpy(np.asarray(v)) for k, v in w.items()})
m = m.cuda().eval()
x = torch.randn(64, 1001, 80, device="cuda")
with torch.no_grad():
    for _ in range(3):
        r, l = m(x)
torch.cuda.synchronize()
print(r.shape, l.shape)
