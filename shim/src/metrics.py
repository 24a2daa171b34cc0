"""reference src/metrics.py -> libsrb drop-in (Metrics, ssim_torch)."""
from superresolution_b200.metrics import Metrics, ssim_torch  # noqa: F401
