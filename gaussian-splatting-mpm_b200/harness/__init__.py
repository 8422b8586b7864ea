"""Drivers around the hot path: synthetic particle clouds, config loading and a mirror of the
reference's main.py `simulate()` loop.  Not part of the solver."""
