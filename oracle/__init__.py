"""CPU oracle for the signals HSCN/ASCN HMM hot path -- TEST INFRASTRUCTURE ONLY.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference
legs may import this package.  The product (signals_b200/) must never do so.
Parity status: decode side (viterbi, logspace_add) pinned against oracle/_ref -- the reference's own
src/code.cpp compiled verbatim against a mock Rcpp header; emission side UNPINNED (R code, R absent, no
golden vectors in the reference); see oracle/oracle.c header and DESIGN.md.
"""
