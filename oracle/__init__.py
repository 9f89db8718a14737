"""TEST INFRASTRUCTURE ONLY — CPU restatement of the reference's batched NLLS path.

Nothing under ``oracle/`` is product code.  Only ``tests/``, ``__graft_entry__.smoke()``
and the ``cpu_baseline`` / ``--impl reference`` legs of ``bench.py`` may import it, and
only as the checker / timed CPU baseline.  ``torchimize_b200`` never imports it.
"""
