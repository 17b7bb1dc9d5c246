"""Small utilities: checkpoints, device timing, clock sampling."""
