"""Benchmark harness of bench.py (repo root; NOT part of the product package: the CPU legs and the parity checks here use oracle/)."""
