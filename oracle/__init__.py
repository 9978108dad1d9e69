"""CPU oracle (test infrastructure).  See oracle/bmhrl_oracle.py."""
