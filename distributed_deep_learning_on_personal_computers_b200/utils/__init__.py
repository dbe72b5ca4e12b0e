"""Data, metrics/logging, checkpointing and timing helpers."""
