"""CPU oracle for the low-precision Linear path. TEST INFRASTRUCTURE ONLY.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / reference arm
may import this package; the product path (transformerengine_b200) never does.
"""
