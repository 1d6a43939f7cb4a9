"""Parity oracle — TEST INFRASTRUCTURE.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may import
this package.  The product (stereo_matching_b200) never does: it has no CPU path at all.
"""
