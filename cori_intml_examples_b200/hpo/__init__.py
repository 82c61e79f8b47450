"""hpo subsystem."""
