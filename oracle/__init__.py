"""CPU oracle for the signals HSCN/ASCN HMM hot path -- TEST INFRASTRUCTURE ONLY.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference
legs may import this package.  The product (signals_b200/) must never do so.
Parity status: UNPINNED (the reference has no golden vectors and cannot run here);
see oracle/oracle.c header and DESIGN.md.
"""
