"""Evaluation metrics (reference package `metrics/`), CUDA-only."""
