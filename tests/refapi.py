"""Test-side access to the compiled reference (oracle/_ref/libyref.so) through the same Python
interface class as the product.  TEST INFRASTRUCTURE: never imported by core_b200."""
import os

from core_b200 import capi
from core_b200.interface import YafrayInterface

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REF_SO = os.path.join(ROOT, "oracle", "_ref", "libyref.so")
_ref = None


def have_ref():
    return os.path.exists(REF_SO)


def ref_lib():
    global _ref
    if _ref is None:
        _ref = capi.CLib(REF_SO, prefix="yref_", product=False)
    return _ref


def RefInterface():
    return YafrayInterface(clib=ref_lib())
