"""CPU-only checks of the drop-in boundary: the C-ABI library loads and exports every symbol
include/spanon.h declares, argument validation works without a device, and the host-side
mirror of the reference interface (state_dict keys, filter matrix, samplers, sharding) is right."""
import ctypes as C
import os
import re

import numpy as np
import pytest
import torch

import recipe
import speech_anonymization_b200 as sab
from speech_anonymization_b200 import _lib

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared():
    text = open(os.path.join(ROOT, "include", "spanon.h")).read()
    return sorted(set(re.findall(r"SPANON_API[^;(]*?\b(spanon_\w+)\s*\(", text)))


def test_library_exports_every_declared_symbol():
    names = _declared()
    assert len(names) >= 18
    handle = C.CDLL(_lib.LIB_PATH)
    for n in names:
        assert hasattr(handle, n), n
    assert sorted(_lib.SIGNATURES) == names          # the ctypes binding covers the whole header
    assert _lib.lib().spanon_abi_version() == 3
    assert _lib.lib().spanon_strerror(-4).decode().startswith("workspace")


def test_c_abi_argument_validation_without_device():
    L = _lib.lib()
    assert L.spanon_num_frames(160000, 160) == 1001
    assert L.spanon_num_frames(0, 160) == 1
    assert L.spanon_num_frames(5, 0) == -1
    assert L.spanon_fbank_f32(None, 1, 10, 10, None, None, None, None) == -2       # NULL cfg
    cfg = _lib.FbankCfg(400, 400, 160, 80, 1e-10, 10.0, 0.0, 80.0, 0, 0, 0, 0, 0)
    assert L.spanon_fbank_f32(None, 1, 10, 5, None, None, C.byref(cfg), None) == -1  # stride < n
    assert L.spanon_fbank_f32(None, 0, 10, 10, None, None, C.byref(cfg), None) == 0  # empty batch
    assert L.spanon_fbank_f32(None, 1, 10, 10, None, None, C.byref(cfg), None) == -2
    assert L.spanon_fbank_i16(None, 1, 10, 5, None, None, C.byref(cfg), None) == -1  # int16 wire format: same checks
    assert L.spanon_fbank_i16(None, 0, 10, 10, None, None, C.byref(cfg), None) == 0
    assert L.spanon_fbank_i16(None, 1, 10, 10, None, None, C.byref(cfg), None) == -2
    assert L.spanon_pipe_submit_i16(None, None, 1, None, None, None) == -2
    assert L.spanon_norm_apply(None, None, None, None, 3, None, 80.0, 1, 4, 4, 80, None) == -1
    opts = _lib.ConvAEOpts(_lib.PREC_FP16, 0, 0, 0, 0)
    assert L.spanon_convae_workspace_bytes(8, 1008, 80, C.byref(opts)) > 8 * 80640 * 32 * 2
    assert L.spanon_convae_workspace_bytes(0, 1008, 80, C.byref(opts)) == 0
    assert L.spanon_convae_fwd(None, None, None, None, None, None, 0.0, 0, None, None, None, 0, 1, 36, 80,
                               C.byref(opts), None) == -2
    w = _lib.ConvAEWeights()
    assert L.spanon_convae_fwd(None, C.byref(w), None, None, None, None, 0.0, 0, None, None, None, 0, 1, 37, 81,
                               C.byref(opts), None) == -1                            # L % 4 != 0
    rw = _lib.ConvReconWeights()
    assert L.spanon_convrecon_fwd(None, C.byref(rw), None, None, None, 0.0, 0, None, None, 0, 1, 37, 81,
                                  C.byref(opts), None) == -1                         # L odd (EndToEnd.py:78 reshape)
    assert L.spanon_convrecon_fwd(None, C.byref(rw), None, None, None, 0.0, 0, None, None, 0, 0, 36, 80,
                                  C.byref(opts), None) == 0                          # empty batch
    assert L.spanon_convrecon_fwd(None, C.byref(rw), None, None, None, 0.0, 0, None, None, 0, 1, 36, 80,
                                  C.byref(opts), None) == -2
    assert 0 < L.spanon_convrecon_workspace_bytes(8, 1008, 80, C.byref(opts)) < L.spanon_convae_workspace_bytes(8, 1008, 80, C.byref(opts))
    xw = _lib.XvectorWeights()
    assert L.spanon_xvector_classify(None, None, C.byref(xw), None, None, None, None, 0, 1, 101, 40, 0, None) == -1  # F != 80
    assert L.spanon_xvector_classify(None, None, C.byref(xw), None, None, None, None, 0, 1, 3, 80, 0, None) == -1    # T <= reflect pad
    assert L.spanon_xvector_classify(None, None, C.byref(xw), None, None, None, None, 0, 0, 101, 80, 0, None) == 0
    assert L.spanon_xvector_classify(None, None, C.byref(xw), None, None, None, None, 0, 1, 101, 80, 0, None) == -2
    assert L.spanon_xvector_workspace_bytes(64, 1001) > 64 * 1001 * 1536 * 2
    assert L.spanon_knn_mi_stats(None, None, None, 1, 2, 4, 2, 3, None, None, None, None) == -5   # k+1 > N
    assert L.spanon_anonymize_workspace_bytes(0, 64, 160000, 160, 80, C.byref(opts)) > 64 * 160000 * 4


def test_state_dict_keys_match_reference():
    torch.manual_seed(8886)
    m = sab.ConvAutoencoder()
    sd = m.state_dict()
    assert set(sd) == set(recipe.CONVAE_SHAPES)
    for k, shp in recipe.CONVAE_SHAPES.items():
        assert tuple(sd[k].shape) == tuple(shp), k
    fc = sab.FullyConnectedAutoencoder(80, 8)
    sd = fc.state_dict()
    assert set(sd) == set(recipe.FCAE_SHAPES)
    for k, shp in recipe.FCAE_SHAPES.items():
        assert tuple(sd[k].shape) == tuple(shp), k
    cr = sab.ConvReconstruction()
    sd = cr.state_dict()
    assert set(sd) == set(recipe.CONVRECON_SHAPES)                   # models/EndToEnd.py:40-54
    for k, shp in recipe.CONVRECON_SHAPES.items():
        assert tuple(sd[k].shape) == tuple(shp), k
    xv = sab.XvectorClassifier()                                     # evaluator_inference.yaml checkpoints
    assert set(xv.embedding_model.state_dict()) == set(recipe.XVECTOR_EMB_SHAPES)
    assert set(xv.classifier.state_dict()) == set(recipe.XVECTOR_CLF_SHAPES)
    for k, shp in recipe.XVECTOR_EMB_SHAPES.items():
        assert tuple(xv.embedding_model.state_dict()[k].shape) == tuple(shp), k
    assert sum(p.numel() for p in m.parameters()) == 533123
    m.load_state_dict({k: torch.from_numpy(np.asarray(v)) for k, v in recipe.make_weights(recipe.CONVAE_SHAPES, 1).items()})


def test_fbank_host_logic(golden):
    fb = sab.Fbank(n_mels=80)
    assert (fb.win_length, fb.hop_length, fb.n_stft) == (400, 160, 201)
    mm = fb.fbank_matrix()
    assert np.array_equal(mm.numpy(), golden["mel_matrix80"])          # bit-exact filter weights
    start, length, off, w = fb._taps(mm)
    assert w.numel() == 387 and int(length.max()) == 13
    rebuilt = np.zeros((201, 80), np.float32)
    for m in range(80):
        s, n, o = int(start[m]), int(length[m]), int(off[m])
        rebuilt[s:s + n, m] = w[o:o + n].numpy()
    assert np.array_equal(rebuilt, golden["mel_matrix80"])
    assert np.array_equal(sab.Fbank().fbank_matrix().numpy(), golden["mel_matrix40"])
    assert np.array_equal(fb.window.numpy(), torch.hamming_window(400).numpy())
    for n in (0, 1, 159, 160, 161, 400, 16000, 160000):
        assert fb.num_frames(n) == 1 + n // 160
    with pytest.raises(NotImplementedError):
        sab.Fbank(deltas=True)


def test_no_cpu_fallback():
    fb = sab.Fbank(n_mels=80)
    with pytest.raises(sab.SpanonError):
        fb(torch.zeros(1, 800))
    with pytest.raises(sab.SpanonError):
        sab.ConvAutoencoder().eval()(torch.zeros(1, 36, 80))
    with pytest.raises(sab.SpanonError):
        sab.InputNormalization()(torch.zeros(1, 6, 1), torch.ones(1))
    with pytest.raises(sab.SpanonError):
        sab.CosineSimilarityLoss()(torch.zeros(1, 2, 3), torch.zeros(1, 2, 3))


def test_group_sampler_matches_pandas_semantics():
    import pandas as pd
    groups = ["b", "a", "c", "a", "b", "c", "a", "b"]
    gs = sab.GroupSampler(groups, 2)
    ref = {k: np.array(list(v)) for k, v in pd.DataFrame(groups, columns=["groups"]).groupby("groups").groups.items()}
    assert list(gs.group_samples) == list(ref)
    for k in ref:
        assert np.array_equal(gs.group_samples[k], ref[k])
    np.random.seed(3)
    a = gs.sample_groups()
    np.random.seed(3)
    b = []
    for g in ref:
        b.extend(ref[g][np.random.choice(len(ref[g]), 2, replace=False)])
    assert a == b


def test_shard_range_partitions_batch():
    for n in (0, 1, 7, 8, 64, 360):
        for ws in (1, 2, 3, 8):
            spans = [sab.shard_range(n, r, ws) for r in range(ws)]
            assert spans[0][0] == 0 and spans[-1][1] == n
            assert all(spans[i][1] == spans[i + 1][0] for i in range(ws - 1))
            sizes = [hi - lo for lo, hi in spans]
            assert max(sizes) - min(sizes) <= 1


def _gloo_worker(rank, world, port, out):
    import torch.distributed as dist
    dist.init_process_group("gloo", init_method="tcp://127.0.0.1:%d" % port, rank=rank, world_size=world)
    lo, hi = sab.shard_range(13, rank, world)
    owned = torch.zeros(13)
    owned[lo:hi] = 1
    dist.all_reduce(owned)                      # test-only collective: every utterance owned exactly once
    t = torch.tensor([float(hi - lo) * (rank + 1)])
    dist.all_reduce(t, op=dist.ReduceOp.MAX)    # bench.py's max-over-ranks timing reduction
    if rank == 0:
        out.put((owned.tolist(), float(t)))
    dist.destroy_process_group()


def test_two_rank_gloo_sharding():
    import torch.multiprocessing as mp
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = 29500 + os.getpid() % 2000
    procs = [ctx.Process(target=_gloo_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    owned, tmax = q.get(timeout=120)
    for p in procs:
        p.join(timeout=120)
        assert p.exitcode == 0
    assert owned == [1.0] * 13 and tmax == 12.0


def test_similarity_metrics_stats_vs_reference():
    """utils/utility_similarity_aggregator.py:4-52 (golden_v3.npz: the reference class run by make_golden_v3.py)."""
    g = dict(np.load(os.path.join(ROOT, "tests", "golden", "golden_v3.npz")))
    scores = torch.from_numpy(g["sim_scores"])
    st = sab.SimilarityMetricsStats()
    st.append(scores[:5])
    assert abs(float(st.peek()) - float(g["sim_peek1"])) < 1e-7
    st.append(scores[5:])
    assert abs(float(st.peek()) - float(g["sim_peek2"])) < 1e-7
    assert abs(float(st.summarize()) - float(g["sim_summary"])) < 1e-7
    assert st.summary["average"] == st.summarize()
    st.clear()
    assert st.scores == [] and st.denom == 0 and st.value == 0


def _dp_worker(rank, world, port, q):
    import torch.distributed as dist
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    torch.manual_seed(0)

    class Toy(torch.nn.Module):     # parameter names like the ConvAutoencoder's: "encoder.*" is final last in backward
        def __init__(self):
            super().__init__()
            self.encoder = torch.nn.Linear(4, 4)
            self.decoder = torch.nn.Linear(4, 2)

        def forward(self, x, hook=None):
            h = torch.tanh(self.encoder(x))
            if hook is not None:
                h.register_hook(lambda g: (hook(), g)[1])
            return self.decoder(h)
    m = Toy()
    from speech_anonymization_b200.distributed import DataParallelTrainer, gather_for_mi
    dp = DataParallelTrainer(m, sync_bn=False)
    assert dp.grads.n_early == 4 * 2 + 2 and dp.grads.n_total == 30      # decoder first, then the encoder
    g = torch.Generator().manual_seed(5)
    xs = torch.randn(3, 2 * world, 4, generator=g)
    dp.prepare()
    fired = []
    for i in range(3):                              # two accumulation micro-batches without communication, then one with
        x = xs[i, rank * 2:(rank + 1) * 2]
        with (dp.no_sync() if i < 2 else __import__("contextlib").nullcontext()):
            (m(x, hook=lambda: (fired.append(i), dp._on_encoder_grad())).pow(2).mean() / 3).backward()
            dp.finish_backward()
    got = dp.grads.flat.clone()
    ref = Toy()
    ref.load_state_dict(m.state_dict())
    for i in range(3):
        (ref(xs[i]).pow(2).mean() / 3).backward()    # mean over the global batch == average of the ranks' means
    want = torch.cat([ref.decoder.weight.grad.reshape(-1), ref.decoder.bias.grad.reshape(-1),
                      ref.encoder.weight.grad.reshape(-1), ref.encoder.bias.grad.reshape(-1)])
    m.encoder.weight.requires_grad = False           # the reference's freeze schedule: a frozen parameter has grad None
    dp.zero_grad()
    X, y = gather_for_mi(torch.full((2, 3), float(rank)), torch.full((2,), rank))
    q.put((rank, float((got - want).abs().max()), fired, m.encoder.weight.grad is None, m.decoder.weight.grad is not None,
           float(dp.grads.flat.abs().max()), X[:, 0].tolist(), y.tolist()))
    dist.destroy_process_group()


def test_data_parallel_trainer_gloo_world2():
    """Host logic of the training collectives on CPU (gloo, world size 2): flat gradient buffer, accumulation without
    communication, early bucket on the bottleneck hook, averaged gradients == global-batch gradients, freeze handling,
    MI gather."""
    import torch.multiprocessing as mp
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = 29400 + os.getpid() % 500
    procs = [ctx.Process(target=_dp_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    res = sorted(q.get(timeout=120) for _ in procs)
    for p in procs:
        p.join(timeout=60)
    for rank, err, fired, frozen_none, dec_has, flat_max, Xcol, y in res:
        assert err < 1e-6, err
        assert fired == [0, 1, 2] and frozen_none and dec_has and flat_max == 0.0
        assert Xcol == [0.0, 0.0, 1.0, 1.0] and y == [0, 0, 1, 1]


def test_pdl_kernels_keep_invariant_loads_behind_the_wait():
    """Programmatic dependent launch (csrc/common.cuh): a kernel launched that way may start before its predecessor has
    finished and blocks in griddepcontrol.wait (ACQBULK in the SASS).  nvcc hoists `const __restrict__` / __ldg loads
    (LDG...CONSTANT, "invariant" to it) above that wait, memory clobber or not -- which once made clf_head_kernel read the
    pooled sums before tdnn.6 had written them.  Guard: in every kernel of the library that waits, no invariant load sits
    ahead of the wait (constants fetched early on purpose use plain LDG)."""
    import shutil
    import subprocess
    cuobjdump = shutil.which("cuobjdump") or "/usr/local/cuda/bin/cuobjdump"
    if not os.path.exists(cuobjdump):
        pytest.skip("cuobjdump not available")
    so = os.path.join(os.path.dirname(sab.__file__), "libspanon.so")
    sass = subprocess.run([cuobjdump, "-sass", so], capture_output=True, text=True, check=True).stdout
    name, early, seen, waits, bad = None, [], False, 0, []
    def close():
        nonlocal waits
        if name and seen:
            waits += 1
            if early:
                bad.append((name, early[:4]))
    for line in sass.splitlines():
        m = re.search(r"Function : (\S+)", line)
        if m:
            close()
            name, early, seen = m.group(1), [], False
            continue
        if "ACQBULK" in line:
            seen = True
        elif not seen and re.search(r"\bLDG\S*\.CONSTANT\b", line):
            early.append(line.strip()[:80])
    close()
    assert waits >= 10, "expected the tcgen05 tap-GEMM kernels, the finalize and the classifier head to wait (found %d)" % waits
    assert not bad, "invariant loads ahead of griddepcontrol.wait: %r" % bad
