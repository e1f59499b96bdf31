"""CPU oracle for the cor-asv-ann neural hot path.  TEST INFRASTRUCTURE ONLY.

Only ``tests/``, ``__graft_entry__.smoke()`` and ``bench.py``'s ``cpu_baseline`` /
``--impl reference`` leg may import this package; the product
(``cor_asv_ann_b200``) never does and fails loudly without its CUDA extension.

PARITY UNPINNED: the reference's arithmetic lives in Keras 2.3.* / TensorFlow 1.15.*
(``/root/reference/requirements.txt:3,5``), which are neither vendored nor
installable here, the trained ``.h5`` models are absent, and the reference's own
tests hold no numeric fixtures for this path (``tests/test_all.py:21-104`` only
greps for "Aufklärung").  This oracle is therefore a restatement of
``ocrd_cor_asv_ann/lib/seq2seq.py`` and ``lib/attention.py`` plus the published
Keras semantics; it is cross-checked by an independent
torch float64 restatement in ``tests/test_oracle_independent.py``.
"""
