"""TEST INFRASTRUCTURE — CPU restatement of the reference's algorithms for the hot path.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may
import this package, and only as the checker (or the timed CPU baseline) — never on the
product path.  Parity pinning: see oracle/README.md and tests/golden/.
"""
