"""Prefixes every line read from stdin with the seconds since start (for `margin polish -a info` timelines)."""
import sys
import time

t0 = time.perf_counter()
for line in sys.stdin:
    sys.stdout.write("%8.3f %s" % (time.perf_counter() - t0, line))
    sys.stdout.flush()
