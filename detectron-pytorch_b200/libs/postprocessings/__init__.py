"""Mirror of the reference package libs/postprocessings (hot-path modules only)."""
