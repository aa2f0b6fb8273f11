"""Test infrastructure only (oracles); see rp_oracle.h / ref_harness.cpp."""
