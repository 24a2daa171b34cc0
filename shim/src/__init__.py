"""Drop-in replacements for the reference's ``src`` package modules that sit on the hot path.

Copy ``architecture.py``, ``losses.py`` and ``metrics.py`` of this directory over the files of the same name in the
reference checkout's ``src/`` (``src/dataset.py`` and ``src/env_setup.py`` stay as they are), or put this directory's
parent first on ``sys.path``: ``from src.architecture import HybridSuperResolutionModel`` (reference
scripts/Modello_5.py:23, scripts/Modello_supporto.py:21) then resolves to the libsrb-backed module with no edit in the
calling scripts, and importing it no longer needs models/BasicSR or models/HAT (reference src/env_setup.py:29-57).
"""
