"""Command-line helpers (``python -m psb200.tools.<name>``)."""
