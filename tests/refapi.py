"""Test-side access to (a) the compiled reference (oracle/_ref/libyref.so) and (b) the host emulator build of
the kernels (tests/emu/libyb200_emu.so), both through the same Python interface class as the product.
TEST INFRASTRUCTURE: never imported by core_b200."""
import os
import subprocess

from core_b200 import capi
from core_b200.interface import YafrayInterface

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REF_SO = os.path.join(ROOT, "oracle", "_ref", "libyref.so")
EMU_SO = os.path.join(ROOT, "tests", "emu", "libyb200_emu.so")
REF_STRICT_SO = os.path.join(ROOT, "oracle", "_ref", "strict", "libyref.so")    # same sources without -ffast-math (oracle/Makefile)
_ref = None
_ref_strict = None
_emu = None


def have_ref():
    return os.path.exists(REF_SO)


def ref_lib():
    global _ref
    if _ref is None:
        _ref = capi.CLib(REF_SO, prefix="yref_", product=False)
    return _ref


def RefInterface():
    return YafrayInterface(clib=ref_lib())


def have_ref_strict():
    return os.path.exists(REF_STRICT_SO)


def StrictRefInterface():
    """The reference compiled without -ffast-math: separates the release build's own float noise from real differences."""
    global _ref_strict
    if _ref_strict is None:
        _ref_strict = capi.CLib(REF_STRICT_SO, prefix="yref_", product=False)
    return YafrayInterface(clib=_ref_strict)


def render_in_subprocess(which, builder, **kw):
    """Frame of core_b200.scenes.<builder>(**kw) rendered by the `release` or the `strict` build of the reference in its own
    process (tests/ref_subprocess.py)."""
    import json
    import sys
    import tempfile
    import numpy as np
    with tempfile.TemporaryDirectory() as d:
        out = os.path.join(d, "frame.npy")
        subprocess.run([sys.executable, os.path.join(ROOT, "tests", "ref_subprocess.py"), which, builder, json.dumps(kw), out],
                       check=True, stdout=subprocess.DEVNULL)
        return np.load(out)


def have_emu():
    if not os.path.exists(EMU_SO):
        subprocess.run(["make", "-C", os.path.join(ROOT, "core_b200", "csrc"), "emu"], check=False,
                       stdout=subprocess.DEVNULL, stderr=subprocess.DEVNULL)
    return os.path.exists(EMU_SO)


def emu_lib():
    global _emu
    if _emu is None:
        _emu = capi.CLib(EMU_SO, prefix="yb200_", product=True)
    return _emu


def EmuInterface():
    return YafrayInterface(clib=emu_lib())
