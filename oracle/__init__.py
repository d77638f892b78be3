"""TEST INFRASTRUCTURE ONLY: CPU oracle of the decode hot path (see oracle/ref_ops.py)."""
