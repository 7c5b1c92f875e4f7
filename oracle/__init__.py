"""TEST INFRASTRUCTURE ONLY.

CPU restatement (plain torch fp32 / numpy fp64) of the reference's client-local
ASR update hot path and FedAvg.  Only tests/, __graft_entry__.smoke() and
bench.py's cpu_baseline / --impl reference legs may import this package.  The
product path (privacy-preserve-federated-asr_b200/b200asr) never does.

Parity pin: the reference has no tests or golden vectors for this path
 so the oracle is pinned against the reference's own code
executed in the build container: tests/golden/make_golden.py imports
/root/reference/federated/src/models.py (and HF transformers, the un-vendored
dependency that holds the encoder arithmetic) and writes tests/golden/*.npz;
tests/test_oracle_golden.py checks this restatement against those files.
"""
