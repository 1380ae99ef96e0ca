"""Test infrastructure only (see binding.py, Makefile): nothing under oracle/ is on the product path."""
