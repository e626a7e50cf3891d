"""Mirror of the reference package libs/configs (hot-path modules only)."""
