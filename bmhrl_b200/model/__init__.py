"""Drop-in counterparts of the reference's `model/` package (same module, class and parameter names)."""
