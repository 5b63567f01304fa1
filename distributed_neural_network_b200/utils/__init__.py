"""Data, logging, checkpointing, CLI helpers."""
