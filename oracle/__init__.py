"""TEST INFRASTRUCTURE.  CPU checkers for the batched-FK hot path.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference leg may
import this package.  The product (tinyfk_b200) never does.
"""
