"""Build ``oracle/_ref/``: the UNMODIFIED reference, byte-compiled from where it lies.

Run in the build container only (needs /root/reference):

    python oracle/make_ref.py

The reference is six Python files without package metadata, so there is nothing to ``pip
install``.  This recipe byte-compiles the three modules of the learner hot path
(``replay_memory.py``, ``model.py``, ``learner.py``) straight from ``/root/reference`` into
``oracle/_ref/*.ref`` (CPython bytecode; not named ``.pyc`` because snapshot tools skip those) --
build outputs only, no source is copied; ``oracle/_ref/`` is git-ignored and travels to the GPU
box with the snapshot like the built ``.so``.  ``oracle/ref_harness.py`` loads these modules
(``SourcelessFileLoader``) so that ``bench.py --impl reference`` times the literal
reference (``cpu_baseline.kind = "reference"``); when ``oracle/_ref`` is absent the arm falls
back to the oracle port.  ``actor.py`` / ``env.py`` are left out: they import gym + ALE, which
the image does not have.

TEST / BASELINE INFRASTRUCTURE ONLY -- see ``oracle/__init__.py``.
"""
from __future__ import annotations

import os
import py_compile
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
OUT = os.path.join(HERE, "_ref")
MODULES = ("replay_memory", "model", "learner")


def build(src="/root/reference", quiet=False) -> bool:
    """Returns True when oracle/_ref was (re)built, False when the reference is not present."""
    if not os.path.isdir(src):
        return False
    os.makedirs(OUT, exist_ok=True)
    for m in MODULES:
        py_compile.compile(os.path.join(src, m + ".py"), cfile=os.path.join(OUT, m + ".ref"),
                           dfile=f"<reference>/{m}.py", doraise=True)
    with open(os.path.join(OUT, "BUILT_FROM"), "w") as f:
        f.write(f"{src} python {sys.version.split()[0]}\n")
    if not quiet:
        print("oracle/_ref:", ", ".join(m + ".ref" for m in MODULES))
    return True


if __name__ == "__main__":
    if not build():
        sys.exit("no /root/reference here: oracle/_ref can only be built in the build container")
