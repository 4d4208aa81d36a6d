"""TEST INFRASTRUCTURE ONLY.

CPU restatement ("oracle") of the ggpm hierarchical message-passing encoder path.
Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference
legs may import this package.  The product (ggpm_b200/) never imports it and has
no CPU fallback.
"""
