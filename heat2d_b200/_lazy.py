"""Import-on-first-use module proxy.

The CPU backend (oracle, dump writers, CLI with ``--backend cpu``) needs neither PyTorch
nor a GPU; importing torch costs seconds, which is most of the run time of a reference-size
job (the reference binaries start in milliseconds).  Modules that only *sometimes* need
torch bind the name to a ``LazyModule`` instead of importing it at the top."""
from __future__ import annotations

import importlib


class LazyModule:
    def __init__(self, name: str):
        self.__dict__["_name"] = name
        self.__dict__["_mod"] = None

    def _load(self):
        if self.__dict__["_mod"] is None:
            self.__dict__["_mod"] = importlib.import_module(self.__dict__["_name"])
        return self.__dict__["_mod"]

    def __getattr__(self, attr):
        return getattr(self._load(), attr)

    def __repr__(self):
        state = "loaded" if self.__dict__["_mod"] is not None else "not loaded"
        return f"<LazyModule {self.__dict__['_name']} ({state})>"
