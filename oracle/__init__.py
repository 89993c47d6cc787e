"""ORACLE -- test infrastructure only (see oracle/oracle.py)."""
