"""reference src/losses.py -> libsrb drop-in (CombinedLoss: weighted L1, value + analytic gradient in one kernel)."""
from superresolution_b200.losses import CombinedLoss  # noqa: F401
