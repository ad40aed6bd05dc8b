"""Render one scene with a build of the reference in a process of its own:
    python tests/ref_subprocess.py <release|strict> <scene builder in core_b200.scenes> '<json kwargs>' <out.npy>
Two builds of the reference cannot share a process (same SONAMEs: the plugins of the second would bind to the core library
of the first), so the strict-IEEE build (oracle/_ref/strict, see oracle/Makefile) is always driven through this script.
TEST INFRASTRUCTURE."""
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))


def main():
    which, builder, kwj, out = sys.argv[1:5]
    from core_b200 import capi, scenes
    from core_b200.interface import YafrayInterface
    import refapi
    so = refapi.REF_STRICT_SO if which == "strict" else refapi.REF_SO
    yi = YafrayInterface(clib=capi.CLib(so, prefix="yref_", product=False))
    yi.setVerbosityMute()
    kw = json.loads(kwj)
    kw = {k: (tuple(v) if isinstance(v, list) else v) for k, v in kw.items()}
    getattr(scenes, builder)(yi, **kw)
    np.save(out, yi.render(kw["width"], kw["height"]))


if __name__ == "__main__":
    main()
