"""Stub of `isaacgym` (closed binary, not installable); test infrastructure only."""
