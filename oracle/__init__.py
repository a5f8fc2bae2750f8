"""CPU oracle (test infrastructure only; see oracle/vit_oracle.py)."""
