"""Mirror of the reference package libs/nets (hot-path modules only)."""
