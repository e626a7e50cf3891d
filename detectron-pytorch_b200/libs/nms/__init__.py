"""Mirror of the reference package libs/nms (hot-path modules only)."""
