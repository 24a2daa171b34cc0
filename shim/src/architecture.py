"""reference src/architecture.py -> libsrb drop-in (same class names, constructor, forward(x), state-dict keys)."""
from superresolution_b200.architecture import AntiCheckerboardLayer, HybridSuperResolutionModel  # noqa: F401
