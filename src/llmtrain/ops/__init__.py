"""Python faces of the hand-written sm_100a kernels (thin ctypes wrappers, no fallbacks)."""
