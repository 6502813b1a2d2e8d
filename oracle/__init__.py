"""CPU ORACLE -- test infrastructure only.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may
import this package.  The product (alphago_b200/) never does.
"""
