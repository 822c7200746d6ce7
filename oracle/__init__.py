"""ORACLE -- TEST INFRASTRUCTURE ONLY, never on the product path.

CPU restatements of the reference's hot path, each function citing the reference
file:line it follows:

  fbank_np.py        STFT / power / mel / dB           (numpy float64)
  normalize_np.py    InputNormalization, x36 padding   (numpy float64)
  anonymizer_ref.py  ConvAutoencoder, FullyConnectedAutoencoder, sex classifiers, StatisticsPooling
                     (torch.nn.functional on explicit weight dicts -- the library the reference itself calls)
  losses_np.py       CosineSimilarityLoss, kNN mutual information
  ref_shim.py        imports the UNMODIFIED reference from /root/reference (this container only)

Parity is PINNED: tests/test_oracle_golden.py checks these files against golden vectors produced by executing
the reference's own code (tests/golden/make_golden.py -> tests/golden/golden_v1.npz) and against the reference's
known-answer tests (speechbrain/tests/unittests/test_features.py:60-88, :100-119);
tests/test_oracle_vs_reference.py re-checks them live against the reference wherever /root/reference is mounted.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may import this package.
speech_anonymization_b200/ never does (tests/test_host_cpu.py::test_no_cpu_fallback enforces it).
"""
