"""Benchmark drivers shared by the entry-point scripts under collectives/."""
