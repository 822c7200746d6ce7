"""Per-kernel CUDA-event times of one ConvAutoencoder training step (8 x 10 s, fp32 training kernels)."""
import sys
import numpy as np, torch
sys.path.insert(0, "tests/golden")
import recipe
import speech_anonymization_b200 as sab
from speech_anonymization_b200 import _lib
B, T = 8, 1008
m = sab.ConvAutoencoder()
w = recipe.make_weights(recipe.CONVAE_SHAPES, 8886)
m.load_state_dict({k: torch.from_numpy(np.asarray(v)) for k, v in w.items()})
m = m.cuda().train()
x = torch.randn(B, T, 80, device="cuda")
y = (torch.arange(B) % 2).cuda()
opt = torch.optim.Adam(m.parameters(), lr=1e-4)
def step():
    rec, logp = m(x)
    loss = torch.nn.functional.mse_loss(rec, x) + 0.8 * torch.nn.functional.nll_loss(logp, y)
    opt.zero_grad(set_to_none=True)
    loss.backward()
    opt.step()
for _ in range(3):
    step()
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(5):
    step()
e1.record(); torch.cuda.synchronize()
print("ms/step", e0.elapsed_time(e1) / 5)
_lib.profile_begin(torch.device("cuda"))
step()
prof = _lib.profile_end(256)
tot = sum(v[1] for v in prof.values())
print("profiled kernel ms", round(tot, 2))
for k, v in sorted(prof.items(), key=lambda kv: -kv[1][1])[:25]:
    print("%-34s n=%3d  %7.3f ms  %4.1f%%" % (k, v[0], v[1], 100 * v[1] / tot))
