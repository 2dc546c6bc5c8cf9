"""TEST INFRASTRUCTURE ONLY - CPU oracle for the --make-grm path (see oracle/grm_oracle.c)."""
