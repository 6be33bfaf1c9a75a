"""TEST INFRASTRUCTURE ONLY: CPU restatement of the reference (see oracle/lara_oracle.h)."""
