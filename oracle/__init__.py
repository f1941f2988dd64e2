"""CPU oracle for the L-MSFC rate-distortion training step.

TEST INFRASTRUCTURE ONLY.  Nothing under ``oracle/`` is product code: it may be
imported by ``tests/``, by ``__graft_entry__.smoke()`` and by ``bench.py``'s
``cpu_baseline`` / ``--impl reference`` legs, and nowhere else.  The product
path (``l-msfc-v2-training_b200/``) never imports it and fails loudly when the
CUDA extension is missing.

What it is: a plain PyTorch-fp32 / C restatement of the reference's algorithm
for every function SURVEY.md §8(a) lists (CompressAI 1.2.4 vendored in
herok97/L-MSFC-V2-Training), each function citing the reference file:line it
follows.  Parity status:

* per-module maths (LowerBound, NonNegativeParametrizer, GDN/IGDN,
  EntropyBottleneck, GaussianConditional, RateDistortionLoss, net/aux optimiser
  split, pmf->quantised CDF): PINNED.  ``tests/golden/make_golden.py`` imports
  the real reference (built out-of-tree from /root/reference) and writes the
  fixtures in ``tests/golden/*.npz`` that ``tests/test_oracle_golden.py`` checks
  this restatement against, together with the reference's own known-answer
  tests (tests/test_ops.py:104-118, tests/test_layers.py:135-172).
* the L-MSFC *topology* (how P2-P5 are fused): PARITY UNPINNED at topology
  level.  The reference checkout contains no L-MSFC model; the
  topology is frozen by ``spec.py`` of the product package and the oracle is
  that same graph assembled from the pinned restated modules.
"""
