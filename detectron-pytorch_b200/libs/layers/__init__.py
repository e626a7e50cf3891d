"""Mirror of the reference package libs/layers (hot-path modules only)."""
