"""Synthetic workloads of the BASELINE shapes: test and bench infrastructure (not part of the product package)."""
