"""CPU oracle for the b200 JPEG decode path -- TEST INFRASTRUCTURE ONLY.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference
legs may import this package.  See oracle/jpeg_oracle.c for what it restates.
"""
from .oracle import *  # noqa: F401,F403
