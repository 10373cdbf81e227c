"""CPU oracle (test infrastructure). See ppls_ref.py and ppls_oracle.c."""
