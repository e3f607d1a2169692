"""CPU oracle for the flat-kNN hot path. TEST INFRASTRUCTURE ONLY.

Nothing under the product package may import from here; only tests/,
__graft_entry__.smoke() and bench.py's CPU-baseline legs do.
"""
