"""Import the unmodified reference (``rtldavis.dsp`` / ``crc`` / ``protocol`` / ``worker``).

TEST INFRASTRUCTURE: used by tests/, by ``bench.py``'s CPU legs and by ``oracle/gen_golden.py``
only.  The modules come from ``oracle/_ref`` (byte-compiled from /root/reference by
``oracle/stage_ref.py``; present on the GPU box) or, failing that, from /root/reference itself
(build container).  ``available()`` is False when neither exists.
"""
from __future__ import annotations

import importlib
import importlib.machinery
import importlib.util
import os
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
_REF_DIR = os.path.join(HERE, "_ref")
_SRC_DIR = "/root/reference/src"


def _ref_hook(path):
    """Path hook for oracle/_ref only: byte-code modules with the suffix .rpyc."""
    if not os.path.abspath(path).startswith(_REF_DIR) or not os.path.isdir(path):
        raise ImportError
    return importlib.machinery.FileFinder(path, (importlib.machinery.SourcelessFileLoader, [".rpyc"]))


def _staged_ok() -> bool:
    """oracle/_ref holds byte-code THIS interpreter can run.  Byte-code is tied to the Python
    minor version that wrote it (the build container's); on a box with another Python the
    import would die with "bad magic number", so such files count as absent and the callers
    skip or fall back to the port instead of crashing."""
    try:
        with open(os.path.join(_REF_DIR, "rtldavis", "dsp.rpyc"), "rb") as fh:
            return fh.read(4) == importlib.util.MAGIC_NUMBER
    except OSError:
        return False


def _root():
    if _staged_ok():
        return _REF_DIR
    if os.path.exists(os.path.join(_SRC_DIR, "rtldavis", "dsp.py")):
        return _SRC_DIR
    return None


def available() -> bool:
    return _root() is not None


def kind() -> str:
    """'ref' when the unmodified reference can be imported, else 'port'."""
    return "ref" if available() else "port"


def load():
    """(dsp, crc, protocol) modules of the unmodified reference."""
    root = _root()
    if root is None:
        raise ImportError("the reference is neither staged in oracle/_ref nor at /root/reference")
    if root == _REF_DIR and _ref_hook not in sys.path_hooks:
        sys.path_hooks.insert(0, _ref_hook)
        sys.path_importer_cache.pop(_REF_DIR, None)
    if root not in sys.path:
        sys.path.insert(0, root)
    dsp = importlib.import_module("rtldavis.dsp")
    crc = importlib.import_module("rtldavis.crc")
    protocol = importlib.import_module("rtldavis.protocol")
    return dsp, crc, protocol
