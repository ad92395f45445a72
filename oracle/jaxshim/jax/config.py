def update(*_a, **_k):
    """jax.config.update(...) -- nothing to configure in the NumPy shim (x64 is always disabled)."""
