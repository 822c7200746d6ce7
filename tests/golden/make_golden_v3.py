"""Generates tests/golden/golden_v3.npz (training step, scope rows f-1 / a14 / a5-speaker / f-4) by EXECUTING THE
REFERENCE'S OWN CODE.

    PYTHONDONTWRITEBYTECODE=1 python tests/golden/make_golden_v3.py

  * ConvAutoencoder in TRAIN mode (models/ConvAutoEncoder.py:136-200, unmodified class): outputs, the loss
    `recon_w * MSE(recon, feats) + sex_w * NLL(sex_logits, labels)` (speechbrain_convae_train.py:105-128 with
    convae.yaml:78-84's losses) and its gradients w.r.t. all 533 123 parameters by torch autograd on CPU, plus the
    BatchNorm running statistics after the step.  Full gradients for the smallest case; for the other two every
    parameter's gradient norm and its first 64 values.
  * the fit_batch schedule (speechbrain_convae_train.py:195-259): epoch-parity freezing, gradient accumulation 3,
    clipping at 5.0 (core.py check_gradients) -- six steps, losses and a few parameters afterwards.  The optimizer is
    plain SGD instead of convae.yaml's Adam: Adam's first update is lr * sign(g), which turns the rounding noise of
    every near-zero gradient element into a +-lr move and makes element-wise comparison of two implementations meaningless.  The ASR utility
    term (weight 0.2) needs the absent pretrained ASR; the reconstruction term takes its place in the even epochs.
  * InputNormalization(norm_type="speaker") in train / eval mode (processing/features.py:1031-1080).
  * SimilarityMetricsStats append / peek / summarize (utils/utility_similarity_aggregator.py:4-52).
  * SpecAugment on features (speechbrain/lobes/augment.py:32-200, convae.yaml:273-283) with the random draws recorded.
Inputs / weights: tests/golden/recipe.py (numpy PCG64), rebuilt bit-identically by the tests.
"""
import importlib
import os
import sys
import types
import warnings

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
sys.path.insert(0, HERE)
warnings.filterwarnings("ignore")

from oracle import ref_shim  # noqa: E402
import recipe  # noqa: E402


def t(x):
    return torch.from_numpy(np.ascontiguousarray(x))


TRAIN_CASES = {"train_a": (21, 8, 36, 1.0, 0.8), "train_b": (22, 6, 72, 0.5, 0.5), "train_c": (23, 4, 1008, 1.0, 0.8)}


def labels_for(seed, B):
    return (np.arange(B) + seed) % 2


def build_model(R, fixed):
    R["StatisticsPooling"]._get_gauss_noise = lambda self, shape, device="cpu": fixed["noise"]
    w = recipe.make_weights(recipe.CONVAE_SHAPES, 8886)
    m = R["ConvAutoencoder"]()
    m.load_state_dict({k: t(v) for k, v in w.items()})
    return m.train()


def main():
    R = ref_shim.load()
    torch.set_num_threads(8)
    out = {}
    fixed = {}
    # ---------------- gradients of one training step ----------------
    # The reference's modules are run in DOUBLE precision here (same code, `.double()`): the stored gradients are then
    # free of the fp32 summation noise of any particular backend, and the test measures the CUDA path's own error.
    for name, (seed, B, T, w_rec, w_sex) in TRAIN_CASES.items():
        m = build_model(R, fixed).double()
        x = t(recipe.make_feats(seed, B, T)).double()
        fixed["noise"] = t(recipe.make_noise(seed, B, 128)).double()
        y = torch.from_numpy(labels_for(seed, B)).long()
        rec, logp = m(x)
        recon_loss = torch.nn.MSELoss()(rec.view(B, -1), x.view(B, -1))
        sex_loss = torch.nn.NLLLoss()(logp, y)
        loss = w_rec * recon_loss + w_sex * sex_loss
        loss.backward()
        out[name + "_recon"] = (rec.detach().numpy() if T <= 72 else rec.detach().numpy()[:, :8]).astype(np.float32)
        out[name + "_logp"] = logp.detach().numpy().astype(np.float32)
        out[name + "_loss"] = np.array([float(loss), float(recon_loss), float(sex_loss)])
        for k, p in m.named_parameters():
            g = p.grad.detach().numpy()
            if name == "train_a":
                out["%s_grad:%s" % (name, k)] = g.astype(np.float32)
            else:
                out["%s_gnorm:%s" % (name, k)] = np.array(np.linalg.norm(g.astype(np.float64)))
                out["%s_ghead:%s" % (name, k)] = g.reshape(-1)[:64].astype(np.float32)
        for k, b in m.named_buffers():
            if "running" in k and T <= 72:
                out["%s_buf:%s" % (name, k)] = b.detach().numpy().astype(np.float32)
        # yardstick: how far the reference's OWN fp32 run (torch CPU kernels) is from its double-precision run
        m32 = build_model(R, fixed)
        fixed["noise"] = t(recipe.make_noise(seed, B, 128))
        x32 = t(recipe.make_feats(seed, B, T))
        rec32, logp32 = m32(x32)
        (w_rec * torch.nn.MSELoss()(rec32.view(B, -1), x32.view(B, -1)) + w_sex * torch.nn.NLLLoss()(logp32, y)).backward()
        dev = {}
        for (k, p), (_, p32) in zip(m.named_parameters(), m32.named_parameters()):
            g, g32 = p.grad.numpy(), p32.grad.numpy().astype(np.float64)
            dev[k] = float(np.max(np.abs(g32 - g)) / max(float(np.max(np.abs(g))), 1e-5))
        out[name + "_fp32_ref_dev"] = np.array([dev[k] for k, _ in m.named_parameters()])
        print(name, "fp32 reference vs fp64 reference: worst", max(dev.items(), key=lambda kv: kv[1] if "bias" not in kv[0] else 0))
    # ---------------- fit_batch schedule ----------------
    m = build_model(R, fixed)
    opt = torch.optim.SGD(m.parameters(), lr=1e-2)
    losses = []
    accum = 3
    for step in range(1, 7):
        epoch = 0 if step <= 3 else 1
        if epoch % 2 == 0:
            w_rec, w_sex = 1.0, 0.8
            for n, p in m.named_parameters():
                p.requires_grad = "sex_classifier" not in n
        else:
            w_rec, w_sex = 0.0, 0.5
            for n, p in m.named_parameters():
                p.requires_grad = "sex_classifier" in n
        B, T = 8, 36
        x = t(recipe.make_feats(100 + step, B, T))
        fixed["noise"] = t(recipe.make_noise(100 + step, B, 128))
        y = torch.from_numpy(labels_for(step, B)).long()
        rec, logp = m(x)
        loss = w_rec * torch.nn.MSELoss()(rec.view(B, -1), x.view(B, -1)) + w_sex * torch.nn.NLLLoss()(logp, y)
        (loss / accum).backward()
        if step % accum == 0:
            torch.nn.utils.clip_grad_norm_((p for p in m.parameters()), 5.0)
            opt.step()
            opt.zero_grad()
        losses.append(float(loss))
    out["fit_losses"] = np.array(losses)
    sd = m.state_dict()
    for k in ("encoder.0.weight", "encoder.11.weight", "encoder.12.bias", "decoder.1.weight", "decoder.8.weight",
              "sex_classifier.tdnn.3.weight", "sex_classifier.classify.6.bias", "sex_classifier.tdnn.8.running_var",
              "sex_classifier.norm.running_mean"):
        out["fit_param:" + k] = sd[k].detach().numpy().copy()
    # ---------------- InputNormalization(norm_type="speaker") ----------------
    feats = t(recipe.make_feats(31, 4, 51))
    lens = t(np.array([1.0, 0.8, 0.6, 0.9], np.float32))
    spk = torch.tensor([[3], [5], [3], [7]])
    n = R["InputNormalization"](norm_type="speaker")
    n.train()
    out["spk_train_call1"] = n(feats.clone(), lens, spk_ids=spk).numpy()
    feats2 = t(recipe.make_feats(32, 4, 51))
    out["spk_train_call2"] = n(feats2.clone(), lens, spk_ids=torch.tensor([[5], [5], [9], [3]])).numpy()
    out["spk_mean_3"] = n.spk_dict_mean[3].numpy().copy()
    out["spk_std_5"] = n.spk_dict_std[5].numpy().copy()
    out["spk_count"] = np.array([n.spk_dict_count[k] for k in (3, 5, 7, 9)])
    n.eval()
    out["spk_eval_call3"] = n(feats.clone(), lens, spk_ids=torch.tensor([[3], [11], [9], [5]])).numpy()   # 11 is unseen
    # ---------------- SimilarityMetricsStats ----------------
    metric_stats = types.ModuleType("speechbrain.utils.metric_stats")

    class MetricStats:      # base class stub: the reference class overrides every method it uses
        pass
    metric_stats.MetricStats = MetricStats
    sys.modules["speechbrain.utils.metric_stats"] = metric_stats
    agg = importlib.import_module("spanon_ref_utils.utility_similarity_aggregator")
    st = agg.SimilarityMetricsStats()
    rng = np.random.Generator(np.random.PCG64(77))
    s1, s2 = rng.random(5).astype(np.float32), rng.random(3).astype(np.float32)
    st.append(t(s1))
    out["sim_peek1"] = np.array(float(st.peek()))
    st.append(t(s2))
    out["sim_peek2"] = np.array(float(st.peek()))
    out["sim_summary"] = np.array(float(st.summarize()))
    out["sim_scores"] = np.concatenate([s1, s2])
    # ---------------- SpecAugment on features (lobes/augment.py:32-200; convae.yaml:273-283) ----------------
    sb = sys.modules["speechbrain"]
    for name in ("speechbrain.utils.data_utils", "speechbrain.processing.speech_augmentation", "speechbrain.utils.torch_audio_backend"):
        stub = types.ModuleType(name)      # augment.py imports these for its OTHER classes; SpecAugment uses only torch
        for attr in ("download_file", "SpeedPerturb", "DropFreq", "DropChunk", "AddBabble", "AddNoise", "AddReverb"):
            setattr(stub, attr, object)
        stub.check_torchaudio_backend = lambda: None
        sys.modules[name] = stub
    if "torchaudio" not in sys.modules:
        try:
            import torchaudio  # noqa: F401
        except Exception:
            sys.modules["torchaudio"] = types.ModuleType("torchaudio")
    aug_mod = importlib.import_module("speechbrain.lobes.augment")
    real_randint = torch.randint
    for name, (seed, B, T, kw) in {
            "specaug_yaml": (7, 3, 108, dict(time_warp=True, time_warp_window=5, time_warp_mode="bicubic", freq_mask=True, n_freq_mask=2,
                                             time_mask=True, n_time_mask=2, replace_with_zero=False, freq_mask_width=30, time_mask_width=40)),
            "specaug_default": (8, 2, 1008, dict())}.items():
        draws = []

        def logged(*a, **k):
            r = real_randint(*a, **k)
            draws.append(r.clone().reshape(-1).to(torch.int64))
            return r
        torch.randint = logged
        try:
            torch.manual_seed(seed)
            aug = aug_mod.SpecAugment(**kw)
            x = t(recipe.make_feats(seed, B, T))
            y = aug(x.clone())
        finally:
            torch.randint = real_randint
        out[name + "_out"] = y.numpy()
        out[name + "_draws"] = torch.cat(draws).numpy()        # c, w, freq len [B*n], freq pos, time len, time pos
    np.savez_compressed(os.path.join(HERE, "golden_v3.npz"), **out)
    print("wrote golden_v3.npz with %d arrays, %.1f MB" % (len(out), os.path.getsize(os.path.join(HERE, "golden_v3.npz")) / 1e6))


if __name__ == "__main__":
    main()
