"""CPU oracle for the Rafiki serving hot path.  TEST INFRASTRUCTURE ONLY.

This package restates, on the CPU, the arithmetic of the reference path
``InferenceWorker -> TfFeedForward.predict -> ensemble_predictions`` so the CUDA
path can be checked against it.  Nothing under ``rafiki_b200/`` may import it;
the only allowed users are ``tests/``, ``__graft_entry__.smoke()`` and the
``cpu_baseline`` / ``--impl reference`` legs of ``bench.py``.

Parity status (see DESIGN.md "Oracle"):

* ``ensemble_predictions`` / ``resize_as_images`` -- PINNED.  The reference's own
  functions (``rafiki/predictor/ensemble.py:6-34``, ``rafiki/model/dataset.py:68-78``)
  were imported from ``/root/reference`` in the build container by
  ``oracle/make_golden.py`` and their outputs committed under ``tests/golden/``;
  ``tests/test_oracle.py`` holds the restatement to those vectors bit for bit.
* Keras forward pass (``TfFeedForward.py:141-157`` run by ``:95``) -- PARITY
  UNPINNED.  The arithmetic lives in tensorflow==1.12.0 (``tf.keras``), a
  third-party dependency that is neither under ``/root/reference`` nor installable
  here, and the reference has no test or golden vector for it (SURVEY.md F2, F3,
  F12).  It is restated from the published Keras layer definitions and
  cross-checked by a second, independent torch-CPU implementation
  (``oracle/torch_ref.py``) plus the known-answer tests of SURVEY.md Appendix A.4.
"""
