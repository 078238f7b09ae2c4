"""Test infrastructure: CPU restatement of the reference path (see oracle/oracle.py)."""
