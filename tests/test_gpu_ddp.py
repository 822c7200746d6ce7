"""2-rank NCCL test of the training collectives (scope row e'): SyncBatchNorm statistics + the flat gradient all-reduce
of DataParallelTrainer reproduce the single-GPU gradients of the concatenated batch (plan item iv;
reference: DDP + SyncBatchNorm wrap, speechbrain/core.py:1149-1162).  Needs two GPUs; skipped on a one-GPU box."""
import json
import os
import subprocess
import sys

import pytest
import torch

pytestmark = pytest.mark.gpu
HERE = os.path.dirname(os.path.abspath(__file__))


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs 2 GPUs (run with gpurun --gpus 2)")
def test_two_rank_gradients_equal_single_gpu(tmp_path):
    out = str(tmp_path / "ddp.json")
    port = 29650 + os.getpid() % 200
    procs = [subprocess.Popen([sys.executable, os.path.join(HERE, "ddp_worker.py"), str(r), "2", str(port), out]) for r in range(2)]
    rcs = [p.wait(timeout=300) for p in procs]
    assert rcs == [0, 0], rcs
    res = json.load(open(out))
    print(res)
    assert res["ranks_agree"] == 0.0                     # the all-reduce leaves identical buffers on every rank
    assert res["gather_ok"]
    # fp32 kernels on 4 + 4 utterances vs 8 utterances at once: summation order differs, nothing else
    assert res["max_abs_diff_over_max"] < 1e-4 and res["worst_param_rel"] < 2e-3
    assert res["running_stats_rel"] < 1e-4               # SyncBatchNorm's running statistics == the global batch's
