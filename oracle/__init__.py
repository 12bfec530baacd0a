"""TEST INFRASTRUCTURE ONLY — CPU oracle for the ECHR caption-generation hot path.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may import
this package; the product (echr_b200/) never does.  Parity status: PINNED for greedy decode, the
relation encoder and teacher-forced log-probabilities (checked against the reference's own modules
run in the authoring container, tests/test_oracle_vs_reference.py, and frozen into tests/golden/);
UNPINNED for beam search, which the reference does not implement (models/OldModel_NEW.py:144-146 is
a vacuous assert that falls through to greedy) — the beam spec is defined in echr_oracle.beam_search
on top of the pinned step function.
"""
