"""TEST INFRASTRUCTURE ONLY -- imports the real reference (icyveins7/pydsproutines) from
/root/reference so that (a) the numpy restatement in oracle/xcorr_oracle.py can be validated
against it and (b) golden vectors can be generated (oracle/make_golden.py).

/root/reference exists only in the build container, never on the GPU box: nothing under
tests/ (-m gpu), bench.py or __graft_entry__.smoke() may import this module at run time.

Recipe: the reference is a flat bag of modules; `import xcorrRoutines`
fails on `musicRoutines.py:11` (matplotlib) and the GroupXcorrFFT class is defined inside a
`try: import cupy` block (`xcorrRoutines.py:1042-1047`), so both are stubbed with MagicMock.
"""
import os
import sys
from unittest import mock

REFERENCE_ROOT = "/root/reference"


def available() -> bool:
    return os.path.isfile(os.path.join(REFERENCE_ROOT, "xcorrRoutines.py"))


def load():
    """Returns the reference's xcorrRoutines module (CPU functions only)."""
    if not available():
        raise RuntimeError("reference tree not present at %s" % REFERENCE_ROOT)
    sys.dont_write_bytecode = True  # the reference tree is read-only
    for name in ("matplotlib", "matplotlib.pyplot", "matplotlib.colors", "matplotlib.cm",
                 "cupy", "cupyx", "cupyx.scipy", "cupyx.scipy.signal",
                 "pyqtgraph", "pyqtgraph.opengl", "PySide6", "PySide6.QtWidgets",
                 "PySide6.QtCore", "PySide6.QtGui"):
        if name not in sys.modules:
            sys.modules[name] = mock.MagicMock()
    if REFERENCE_ROOT not in sys.path:
        sys.path.insert(0, REFERENCE_ROOT)
    import xcorrRoutines  # noqa: E402  (the reference's own module)
    return xcorrRoutines
