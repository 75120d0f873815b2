"""CPU oracle for the ORB extract + Hamming match path — TEST INFRASTRUCTURE ONLY.

Only tests/, __graft_entry__.smoke() and bench.py's CPU-baseline leg may import this package,
and only as the checker or the timed CPU baseline; the product (visionx_slam_b200) never does.
See oracle/orb_oracle.c for what is restated and how it is pinned.
"""
from .oracle import *  # noqa: F401,F403
