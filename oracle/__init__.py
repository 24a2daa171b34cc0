"""ORACLE package — test infrastructure only (CPU restatement of the reference hot path).

Importable only from ``tests/``, ``__graft_entry__.smoke()`` and ``bench.py``'s cpu_baseline / ``--impl reference``
legs.  The product (``superresolution_b200``) never imports it.
"""
