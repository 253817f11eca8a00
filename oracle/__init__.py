"""CPU oracle of the rfarl hot path — test infrastructure only (see asv_oracle.h)."""
