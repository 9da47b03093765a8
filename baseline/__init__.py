"""Baseline harness: runs the UNMODIFIED reference (baseline/_ref) for `bench.py --impl reference`."""
