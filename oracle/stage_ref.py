"""Build oracle/_ref: the UNMODIFIED reference, byte-compiled, for the GPU box.

TEST INFRASTRUCTURE.  The reference is pure Python, so "compiling it from the sources
where they lie" (/root/reference/src/rtldavis) means byte-compiling: every module the
demodulation path and its production caller need is compiled with ``py_compile`` straight
from the read-only tree into sourceless byte-code files under ``oracle/_ref/rtldavis/``
(suffix ``.rpyc``: the gpurun snapshot leaves ``*.pyc`` behind; ``oracle/ref_loader.py``
registers an importer for the suffix).
No reference source is copied into the repository; ``oracle/_ref/`` is git-ignored (not
gpurun-ignored), so the build output travels to the GPU box like the built ``.so`` files.

Modules: ``dsp`` (src/rtldavis/dsp.py), ``crc`` (crc.py), ``protocol`` (protocol.py: the
``Parser`` that owns the ``Demodulator`` and calls it, protocol.py:115-116,282-337),
``worker`` (worker.py:49-50, the production loop), and what ``protocol`` imports
(``sensor_classes``, ``decoders/*``, ``version``).

Run:  python oracle/stage_ref.py   (or ``make -C oracle ref``); a no-op when
/root/reference is absent (the GPU box uses the prebuilt files).
"""
from __future__ import annotations

import importlib.util
import os
import py_compile
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
REF_SRC = os.environ.get("RTLDAVIS_REFERENCE", "/root/reference/src/rtldavis")
OUT = os.path.join(HERE, "_ref", "rtldavis")

TOP = ["__init__", "version", "crc", "dsp", "protocol", "sensor_classes", "worker"]


def stage(verbose: bool = False) -> bool:
    """True when oracle/_ref holds the byte-compiled reference afterwards."""
    if not os.path.isdir(REF_SRC):
        return os.path.exists(os.path.join(OUT, "dsp.rpyc"))
    jobs = [(os.path.join(REF_SRC, m + ".py"), os.path.join(OUT, m + ".rpyc")) for m in TOP]
    dec = os.path.join(REF_SRC, "decoders")
    for name in sorted(os.listdir(dec)):
        if name.endswith(".py") and not name.endswith("_test.py"):
            jobs.append((os.path.join(dec, name), os.path.join(OUT, "decoders", name[:-3] + ".rpyc")))
    for src, dst in jobs:
        os.makedirs(os.path.dirname(dst), exist_ok=True)
        # dfile: what tracebacks name; unchecked-hash pycs are valid without their source
        py_compile.compile(src, cfile=dst, dfile="rtldavis/" + os.path.relpath(src, REF_SRC),
                           doraise=True,
                           invalidation_mode=py_compile.PycInvalidationMode.UNCHECKED_HASH)
        if verbose:
            print(dst)
    with open(os.path.join(HERE, "_ref", "PROVENANCE.txt"), "w") as fh:
        fh.write(f"byte-compiled by oracle/stage_ref.py from {REF_SRC} with Python "
                 f"{sys.version.split()[0]} (byte-code magic {importlib.util.MAGIC_NUMBER.hex()}, which "
                 f"oracle/ref_loader.py checks before importing); unmodified 2bitoperations/rtldavis modules\n")
    return True


if __name__ == "__main__":
    ok = stage(verbose="-v" in sys.argv)
    print("oracle/_ref ready" if ok else "no reference tree and no prebuilt oracle/_ref")
