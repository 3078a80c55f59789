"""TEST INFRASTRUCTURE ONLY: CPU oracle of the reference's stepping path (see casim_oracle.c)."""
