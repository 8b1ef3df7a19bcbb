"""CPU oracle for the magicsoup hot path.  TEST INFRASTRUCTURE ONLY (see ms_oracle.c)."""
