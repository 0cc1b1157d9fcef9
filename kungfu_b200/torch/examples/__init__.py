"""Self-checking mini examples: ``python -m kungfu_b200.torch.examples`` (also under kungfu-run)."""
