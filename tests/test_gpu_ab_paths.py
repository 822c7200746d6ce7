"""The A/B switches keep the older kernel forms alive (SPANON_FBANK, SPANON_FCAE, SPANON_TC2, SPANON_POOL_FUSE, SPANON_PDL,
SPANON_FUSE_ENC0, SPANON_COMPOSE_DEC, SPANON_ENC5_PAIR, SPANON_DEC45_PAIR, ring depths): every form must still meet the parity bars, or the measurements that
compare them mean nothing.  The library reads the switches once per process, so each setting runs in its own process."""
import json
import os
import subprocess
import sys

import pytest

pytestmark = pytest.mark.gpu
HERE = os.path.dirname(os.path.abspath(__file__))

SETTINGS = [
    {},                                                     # the defaults
    {"SPANON_FBANK": "stockham"}, {"SPANON_FBANK": "warp"},
    {"SPANON_FCAE": "1"},
    {"SPANON_POOL_FUSE": "0"}, {"SPANON_PDL": "0"}, {"SPANON_TC2": "0"},
    {"SPANON_FUSE_ENC0": "0"}, {"SPANON_COMPOSE_DEC": "0"}, {"SPANON_FUSED_STATS": "0"},
    {"SPANON_T2_STAGES": "16", "SPANON_TC_STAGES": "32"}, {"SPANON_TC2_STATS3": "1"},
    {"SPANON_TOEP_EPI8": "0"}, {"SPANON_ENC5_PAIR": "0"}, {"SPANON_DEC45_PAIR": "1"}, {"SPANON_DEC45_PAIR": "1", "SPANON_TC2_UNFOLD4": "0"},
]


@pytest.mark.parametrize("env", SETTINGS, ids=lambda e: ",".join("%s=%s" % kv for kv in e.items()) or "defaults")
def test_every_switchable_kernel_form_meets_the_parity_bars(env):
    e = dict(os.environ)
    e.update(env)
    e["PYTHONPATH"] = os.path.dirname(HERE) + os.pathsep + e.get("PYTHONPATH", "")
    r = subprocess.run([sys.executable, os.path.join(HERE, "ab_paths_worker.py")], env=e, capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, r.stderr[-2000:]
    d = json.loads(r.stdout.strip().splitlines()[-1])
    for k, v in d.items():
        if k.startswith("fbank"):
            assert v < 1e-4, (k, v)
        elif k.startswith("fcae"):
            assert v < 1e-4, (k, v)
        elif "_rec_" in k:
            assert v < 1e-2, (k, v)          # fp16 tensor-core mode: BASELINE's 16-bit bar
        elif "_logp_" in k:
            assert v < 2e-3, (k, v)
    assert len(d) >= 13
