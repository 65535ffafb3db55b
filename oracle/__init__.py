"""Test-only oracle package (CPU restatement + compiled reference).

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline /
--impl reference legs may import this.  The product path never does.
"""
