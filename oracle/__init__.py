"""CPU oracle for the kube_gpu_stats hot path.  TEST INFRASTRUCTURE ONLY.

Nothing under ``oracle/`` is part of the product.  Only ``tests/``,
``__graft_entry__.smoke()`` and the ``cpu_baseline`` / ``--impl reference`` legs of
``bench.py`` may import it, and there only as the checker / the timed CPU arm.
The product (``kube_gpu_stats_b200``) never imports this package and fails
loudly when its CUDA library is missing.

Parity status
-------------
* ``oracle.ref_gpu_util_stats`` / ``oracle.ref_who_use_gpu`` restate the two
  reference scripts (Python 2, cannot run here: SURVEY.md section 8c).  The
  reference ships **no tests, fixtures or golden vectors**, and Python 2 is not
  installed, so the restatement is pinned only by known-answer vectors written
  from a reading of the source (``tests/golden/ref_rules_*.json``, every rule of
  SURVEY.md appendix A).  **Parity unpinned** against an executing reference.
* ``oracle.window_stats`` holds the *declared* extensions (max, percentiles,
  tumbling/sliding windows, per-GPU / per-node keys, the (node,pid)->pod join).
  They have no reference counterpart at all; the rules are declared in that
  module's docstring and frozen by ``tests/golden/window_kat_*.json``.
* ``oracle.prom_answer`` states, one byte at a time, which parts of a Prometheus
  answer's text are series, value arrays and ``[ts, "v"]`` pairs (what
  ``kgs_matrix_parse`` reports as offsets and byte spans).  The numbers of an
  answer are checked against the reference's own calls, ``json.loads`` and
  ``float()`` (gpu_util_stats.py:25,:48,:74) - for this row the checker is the
  reference's code itself.
"""
