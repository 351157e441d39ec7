"""ctypes mirrors of the C structs in include/mjs_model.h.

The header is the single source of truth; this module parses its simple
"one declaration per line" grammar so the Python model flattener, the CUDA
library and the oracle can never disagree about the layout.
"""
from __future__ import annotations

import ctypes
import re
from pathlib import Path

import numpy as np

HEADER = Path(__file__).resolve().parents[2] / "include" / "mjs_model.h"


def _parse(text: str):
    macros: dict[str, int] = {}
    for m in re.finditer(r"^#define\s+(MJS_\w+)\s+(-?\d+)\s*(?:/\*.*)?$", text, re.M):
        macros[m.group(1)] = int(m.group(2))
    structs: dict[str, list] = {}
    for sm in re.finditer(r"typedef struct (\w+) \{(.*?)\} \1;", text, re.S):
        fields = []
        body = re.sub(r"/\*.*?\*/", "", sm.group(2), flags=re.S)
        for line in body.splitlines():
            line = line.strip()
            if not line:
                continue
            fm = re.match(r"^(int|double)\s+(\w+)(?:\[(.+)\])?;$", line)
            if not fm:
                raise ValueError(f"mjs_model.h: cannot parse line {line!r}")
            ctype = ctypes.c_int if fm.group(1) == "int" else ctypes.c_double
            n = None
            if fm.group(3):
                n = 1
                for tok in fm.group(3).split("*"):
                    tok = tok.strip()
                    n *= macros[tok] if tok in macros else int(tok)
            fields.append((fm.group(2), ctype, n))
        structs[sm.group(1)] = fields
    return macros, structs


MACROS, _STRUCTS = _parse(HEADER.read_text())


def _make(name: str):
    fields = [(f, (t * n) if n else t) for f, t, n in _STRUCTS[name]]
    info = {f: (t, n) for f, t, n in _STRUCTS[name]}

    class S(ctypes.Structure):
        _fields_ = fields
        _info = info

        def set(self, key, value):
            t, n = self._info[key]
            if n is None:
                setattr(self, key, t(value).value)
                return
            arr = np.asarray(value, dtype=np.int32 if t is ctypes.c_int else np.float64).ravel()
            if arr.size > n:
                raise ValueError(f"{name}.{key}: {arr.size} values exceed capacity {n}")
            dst = getattr(self, key)
            for i, v in enumerate(arr.tolist()):
                dst[i] = v

        def get(self, key):
            t, n = self._info[key]
            v = getattr(self, key)
            if n is None:
                return v
            return np.ctypeslib.as_array(v).copy()

        def as_dict(self):
            return {k: self.get(k) for k in self._info}

    S.__name__ = name
    return S


MjsModel = _make("MjsModel")
MjsTask = _make("MjsTask")
MjsEnvState = _make("MjsEnvState")
MjsStepOut = _make("MjsStepOut")
MjsVisual = _make("MjsVisual")
MjsCamera = _make("MjsCamera")
