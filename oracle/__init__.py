"""CPU oracle for the PacedNegatives LCE hot path — TEST INFRASTRUCTURE ONLY.

Nothing under ``oracle/`` is product code.  Only ``tests/``, ``__graft_entry__.smoke()`` and the
``cpu_baseline`` / ``--impl reference`` legs of ``bench.py`` may import it, and there only as the
checker or the timed CPU baseline.  ``pacednegatives_b200`` never imports it.

Each function restates one reference function (file:line under the reference tree, cited in its
docstring) in numpy / torch-CPU.  Parity status: the reference ships no tests, golden vectors or
fixtures, so the restatements are pinned against the reference itself, run in the build
container: ``tests/golden/make_golden.py`` imports the reference modules from /root/reference, runs
them on seeded inputs and commits inputs+outputs as ``tests/golden/*.npz``; ``tests/test_oracle_golden.py``
replays those through this package.
"""
