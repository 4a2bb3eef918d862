"""CPU oracle (test infrastructure).  See ``oracle/ba_oracle.py`` for scope and parity status.

Only ``tests/``, ``__graft_entry__.smoke()`` and the ``cpu_baseline`` / ``--impl reference`` legs of
``bench.py`` may import this package; the product (``bundle_adjustment_b200``) never does.
"""
