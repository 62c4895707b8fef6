"""Start the CUDA driver before anything else is imported.

Creating the primary context takes ~0.7 s on a B200 box -- more than parsing an 830 MB profile with
librthc_hostio.so -- so the CLI calls start() as its first statement: a daemon thread loads
libreportree_hc.so with ctypes (no numpy needed) and calls rthc_prewarm(device), which releases the
GIL, while the main thread imports numpy, parses argv and reads the profile.  Failures are ignored
here; hc_backend.HCContext raises the real error later.
"""
import ctypes
import os
import threading

_LIB = os.path.join(os.path.dirname(os.path.abspath(__file__)), "libreportree_hc.so")


def start(device=0):
    def run():
        try:
            lib = ctypes.CDLL(_LIB)
            lib.rthc_prewarm.argtypes = [ctypes.c_int]
            lib.rthc_prewarm(int(device))
        except Exception:                                   # noqa: BLE001 -- reported by HCContext later
            pass
    t = threading.Thread(target=run, daemon=True, name="rthc-prewarm")
    t.start()
    return t
