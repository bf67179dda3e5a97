"""Reference-parity utility layer (dist / hub / weights / layers)."""
