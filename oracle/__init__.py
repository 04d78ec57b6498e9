"""CPU oracle for the B200 multigrid engine -- TEST INFRASTRUCTURE ONLY.

Nothing under `evostencils_b200/` may import this package.  See mg_oracle.py."""
