"""Test oracle for the flownet2-tf hot-path ops.  TEST INFRASTRUCTURE ONLY -- see oracle/fn2_oracle.c.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs import this package.
"""
