"""monitor subsystem."""
