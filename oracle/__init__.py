"""CPU oracle for the cor-asv-ann neural hot path.  TEST INFRASTRUCTURE ONLY.

Only ``tests/``, ``__graft_entry__.smoke()`` and ``bench.py``'s ``cpu_baseline`` /
``--impl reference`` leg may import this package; the product
(``cor_asv_ann_b200``) never does and fails loudly without its CUDA extension.

PARITY: the decoders, the ``Node`` ordering, vectorisation, the attention formula and the re-alignment are PINNED
bit for bit against the reference's own Python, executed from ``/root/reference`` by ``tests/ref_harness.py``
(``tests/test_reference_pin.py``, frozen in ``tests/golden/ref_*.npz``).  The Keras layer arithmetic inside
``predict_on_batch`` (``LSTMCell``, ``Bidirectional``, softmax; Keras 2.3.* / TensorFlow 1.15.*,
``/root/reference/requirements.txt:3,5``) is UNPINNED: neither package is vendored or installable here, the
trained ``.h5`` models are absent, and the reference's own tests hold no numeric fixtures for this path
(``tests/test_all.py:21-104`` only greps for "Aufklärung").  That part is a restatement of
``ocrd_cor_asv_ann/lib/seq2seq.py`` / ``lib/attention.py`` plus the published Keras semantics (SURVEY.md
Appendix A), cross-checked by an independently written ``torch.nn.LSTM`` float64 restatement
(``tests/restatement_torch.py``, used by ``tests/test_oracle.py``).  ``oracle/train.py`` (the training graph, torch
float64 autograd) is unpinned too and tied to ``oracle/model.py`` instead (``tests/test_train_oracle.py``).
"""
