"""ctypes binding of libnpfk.so, generated from include/npfk.h so the two cannot drift apart.

There is no CPU fallback: if the library is missing or a call fails, a RuntimeError is raised.
"""
from __future__ import annotations

import ctypes
import os
import re
from typing import Dict, List, Tuple

HERE = os.path.dirname(os.path.abspath(__file__))
HEADER = os.path.join(os.path.dirname(HERE), "include", "npfk.h")
LIB_PATH = os.path.join(HERE, "libnpfk.so")

_CTYPES = {
    "int": ctypes.c_int,
    "int64_t": ctypes.c_int64,
    "float": ctypes.c_float,
    "double": ctypes.c_double,
}


def parse_header(path: str = HEADER) -> Dict[str, Tuple[str, List[Tuple[str, str]]]]:
    """Return {symbol: (return_type, [(arg_type, arg_name), ...])} for every prototype in the header."""
    with open(path) as f:
        text = f.read()
    text = re.sub(r"/\*.*?\*/", " ", text, flags=re.S)
    text = re.sub(r"//[^\n]*", " ", text)
    protos = {}
    for m in re.finditer(r"\b(int64_t|int|const char\s*\*)\s+(npfk_\w+)\s*\(([^;{]*?)\)\s*;", text, flags=re.S):
        ret, name, args = m.group(1).strip(), m.group(2), m.group(3).strip()
        parsed: List[Tuple[str, str]] = []
        if args and args != "void":
            for a in args.split(","):
                a = " ".join(a.split())
                mm = re.match(r"^(.*?)(\w+)$", a)
                parsed.append((mm.group(1).strip(), mm.group(2)))
        protos[name] = (ret, parsed)
    return protos


def _ctype_of(t: str):
    if "*" in t:
        return ctypes.c_void_p
    t = t.replace("const", "").strip()
    return _CTYPES[t]


class NpfkError(RuntimeError):
    pass


class _Lib:
    def __init__(self):
        self._dll = None
        self.protos = parse_header()

    def load(self):
        if self._dll is not None:
            return self._dll
        if not os.path.exists(LIB_PATH):
            raise NpfkError(
                f"{LIB_PATH} is missing: build it with `python -m npf_gwwaveform_b200.build` "
                "(there is no CPU or PyTorch fallback for the npf hot path)."
            )
        dll = ctypes.CDLL(LIB_PATH)
        for name, (ret, args) in self.protos.items():
            fn = getattr(dll, name)   # AttributeError here = header/library mismatch
            fn.argtypes = [_ctype_of(t) for t, _ in args]
            fn.restype = ctypes.c_char_p if "char" in ret else (ctypes.c_int64 if ret == "int64_t" else ctypes.c_int)
        self._dll = dll
        return dll

    def call(self, name: str, *args):
        dll = self.load()
        rc = getattr(dll, name)(*args)
        if rc != 0:
            msg = dll.npfk_last_error().decode(errors="replace")
            raise NpfkError(f"{name} failed with code {rc}: {msg}")

    def launch_count(self) -> int:
        return int(self.load().npfk_launch_count())

    def version(self) -> int:
        return int(self.load().npfk_version())


LIB = _Lib()
