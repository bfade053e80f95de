"""ORACLE package -- test infrastructure only (see oracle/acct_ref.c and
oracle/quantity.py headers).  Never imported by the product (`arena_b200/`)."""
