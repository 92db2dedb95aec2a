"""Import alias used by tests/test_reference_python_suite.py: programs written against the reference's Python package
(`import pccl`) run against this repository's implementation when this directory is first on PYTHONPATH."""
import sys
from pathlib import Path

sys.path.insert(0, str(Path(__file__).resolve().parent.parent.parent.parent))
from pccl_b200 import *  # noqa: E402,F401,F403
from pccl_b200 import __all__ as _names  # noqa: E402
import pccl_b200 as _impl  # noqa: E402

__all__ = list(_names)
__version__ = getattr(_impl, "__version__", "0")
