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
_ref = None
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
