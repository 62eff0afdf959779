"""TEST INFRASTRUCTURE ONLY -- CPU oracle for the (candidate x CV-fold) sweep.

Nothing under ``oracle/`` is product code.  Only ``tests/``,
``__graft_entry__.smoke()`` and ``bench.py``'s ``cpu_baseline`` / ``--impl
reference`` legs may import it, and only as the checker or the timed CPU
baseline.  The product path (``cs230_distributed_machine_learning_b200``) never
imports it and fails loudly when its CUDA library is missing.

Parity status: the reference (sanjita2911/CS230-distributed-machine-learning)
has no tests and pins no numbers for this path ("parity unpinned" by the
reference itself).  The arithmetic lives in scikit-learn 1.9.0 /
scipy 1.18.1 (third-party, not vendored in the reference; the reference leaves
scikit-learn unpinned in aws-prod/docker/requirements_worker.txt:1).  The
oracle is therefore pinned two ways:
  * ``reference_path.py`` runs the reference's own call sequence
    (aws-prod/worker/worker.py:289-363) on the installed scikit-learn, and was
    checked against the *imported* reference ``worker.train_model`` and
    ``task_handler.create_subtasks`` in this container -- the fixtures under
    ``tests/golden/`` were produced by ``tests/golden/make_golden.py`` from those
    imports.
  * full-size pins: ``tests/golden/cfg2_full.npz`` (scikit-learn on float64, all 320 fits of BASELINE.json config 2,
    made by ``tests/golden/make_cfg_full.py``) and ``tests/golden/cfg3_full.npz`` (``ridge.py``'s solver from float64
    moments of all 4M rows, anchored by one scikit-learn Ridge fit, ``tests/golden/make_cfg3_full.py``).
  * ``lbfgs.py`` / ``ridge.py`` / ``folds.py`` restate the published algorithms
    (L-BFGS-B + MINPACK-2 dcsrch, Cholesky ridge, (Stratified)KFold) in numpy
    fp64 and are checked against scikit-learn / scipy in ``tests/test_oracle_*``.
"""
