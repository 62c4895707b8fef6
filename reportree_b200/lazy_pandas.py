"""pandas costs ~1 s to import; the CLI needs it late (loci report, metadata filter) or never on
the C++ text path.  `pandas` here is a proxy that imports the real module on first attribute
access; `preload()` starts that import on a helper thread so that it overlaps the C++ profile
parse and the CUDA context set-up (both release the GIL)."""
import importlib
import threading


class _LazyModule:
    def __init__(self, name):
        self.__dict__["_name"] = name
        self.__dict__["_mod"] = None

    def _load(self):
        if self.__dict__["_mod"] is None:
            self.__dict__["_mod"] = importlib.import_module(self.__dict__["_name"])
        return self.__dict__["_mod"]

    def __getattr__(self, attr):
        return getattr(self._load(), attr)


pandas = _LazyModule("pandas")


def preload():
    t = threading.Thread(target=pandas._load, daemon=True)
    t.start()
    return t
