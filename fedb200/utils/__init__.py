"""Utilities: checkpoints, progress/metrics, timers, synthetic data."""
