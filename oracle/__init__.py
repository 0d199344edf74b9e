"""oracle/ -- CPU checkers for the B200 k-mer hash-table engine.  TEST INFRASTRUCTURE ONLY.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs
may import this package.  The product package (kmerhash_b200/) never does.

  oracle.port  -> libkmerhash_oracle.so  (plain-C restatement, kmerhash_oracle.c)
  oracle.ref   -> _ref/libkmerhash_ref.so (the reference's own headers compiled from
                  /root/reference by oracle/Makefile; None when it has not been built)
"""
from .bindings import port, ref, build, HASH_KINDS, MAP_KINDS  # noqa: F401
