"""tadnn bootstrap (temporary minimal init; replaced below)."""
