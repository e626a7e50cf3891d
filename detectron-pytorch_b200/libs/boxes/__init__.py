"""Mirror of the reference package libs/boxes (hot-path modules only)."""
