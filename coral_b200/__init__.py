"""coral_b200 — the CoRAL reconstruct hot path on B200 (drop-in modules + the C ABI binding)."""
import os as _os

# A ctx runs its scan, its two pass lanes and its side copies on separate streams, and the multi-GPU exchanges wait inside
# kernels for peers: streams that share a hardware queue would serialise behind such a wait.  Read by CUDA when its
# context is created, so it only takes effect when this package is imported before CUDA starts.
_os.environ.setdefault("CUDA_DEVICE_MAX_CONNECTIONS", "32")
