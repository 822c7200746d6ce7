"""Three Fbank calls on the bench shape (60 x 10 s) -- the workload tools/ncu_fbank.sh profiles."""
import torch
import speech_anonymization_b200 as sab
fb = sab.Fbank(n_mels=80)
x = (0.1 * torch.randn(60, 160000)).cuda()
for _ in range(3):
    y = fb(x)
torch.cuda.synchronize()
print(y.shape)
