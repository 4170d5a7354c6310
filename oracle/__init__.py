"""CPU oracle for the MOPO model-rollout hot path.  TEST INFRASTRUCTURE ONLY.

This package is a NumPy restatement of the reference algorithm
(xionghuichen/mopo: mopo/models/bnn.py, fc.py, utils.py, fake_env.py,
mopo/static/*.py, softlearning/replay_pools/flexible_replay_pool.py,
mopo/algorithms/mopo.py:723-765).  Only ``tests/``, ``__graft_entry__.smoke()``
and ``bench.py``'s ``cpu_baseline`` / ``--impl reference`` legs may import it,
and there only as the checker or the timed CPU baseline.  The product package
``mopo_b200`` never imports it and has no CPU fallback.

Parity pinning status
---------------------
* ``oracle.fake_env`` / ``oracle.static_fns`` / ``oracle.pool`` are pinned
  against the reference's OWN ``FakeEnv.step``, ``mopo.static`` and
  ``FlexibleReplayPool`` run in the build container (they are pure NumPy and
  import with a stub ``tensorflow`` module): ``tests/golden/make_golden.py``
  generated ``tests/golden/*.npz`` from the real reference code, and
  ``tests/test_oracle_golden.py`` replays them against this restatement.
* ``oracle.ensemble.forward`` restates ``BNN._compile_outputs`` whose arithmetic
  lives in TensorFlow 1.14 (``tensorflow_gpu==1.14.0``,
  environment/requirements.txt:90), which is absent from the image and the
  reference ships no tests or golden vectors for it: for the forward pass
  **parity is unpinned** beyond closed-form known-answer tests (SURVEY.md §4
  items 5-6) and float64-vs-float32 self-consistency.
"""
