"""CPU oracle for the Rosy Packager texture stage - TEST INFRASTRUCTURE ONLY.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may
import this package.  See oracle/rosy_oracle.h for what it restates and what is (un)pinned.
"""
from .binding import *  # noqa: F401,F403
