"""GPU-backed single-game facades with the reference `Environment` interface."""
