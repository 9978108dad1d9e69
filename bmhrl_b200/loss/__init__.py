"""Drop-in counterparts of the reference's `loss/` package."""
