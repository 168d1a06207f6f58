"""Training-time operator and losses (reference package `NewBP_model/`), CUDA-only."""
