"""TEST INFRASTRUCTURE: ctypes loader of the CPU oracle libraries (oracle/_ref when the reference's C core was compiled
in place, else the plain-C port under oracle/_build). The oracle exports the product's C ABI under the prefix ``fbo_``, so
the product's own ctypes binding class is reused to call it."""
from __future__ import annotations

import os

from binless_b200._abi import FastBinlessLib

_HERE = os.path.dirname(os.path.abspath(__file__))


def oracle_lib(kind: str = "auto") -> FastBinlessLib:
    ref = os.path.join(_HERE, "_ref", "libfbo_ref.so")
    port = os.path.join(_HERE, "_build", "libfbo_port.so")
    if kind == "auto":
        kind = "reference" if os.path.exists(ref) else "port"
    path = ref if kind == "reference" else port
    lib = FastBinlessLib(path, "fbo_")
    lib.kind = kind
    return lib
