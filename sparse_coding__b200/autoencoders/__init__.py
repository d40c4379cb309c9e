"""Host-side mirror of the reference's `autoencoders` package (only the hot-path modules)."""
