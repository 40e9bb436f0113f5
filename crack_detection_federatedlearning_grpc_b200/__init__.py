"""Alias of the ``crackfl`` package under the long project name (the hyphenated directory name in the task
statement is not an importable identifier).  ``models/ ops/ parallel/ utils/`` resolve to crackfl's."""
import importlib
import sys

import crackfl

for _sub in ("models", "ops", "parallel", "utils", "protocol", "io", "data", "postproc", "cli", "config", "trainer"):
    try:
        sys.modules[f"{__name__}.{_sub}"] = importlib.import_module(f"crackfl.{_sub}")
    except Exception:  # optional sub-packages
        pass
__version__ = crackfl.__version__
