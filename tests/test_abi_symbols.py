"""The C-ABI library loads without a GPU and exports every symbol include/*.h declares."""
import ctypes
import glob
import os
import re

from spatial_queue_b200 import _lib

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
DECL = re.compile(r"^\s*(?:const\s+)?[A-Za-z_][A-Za-z0-9_\s\*]*?[\s\*]([A-Za-z_][A-Za-z0-9_]*)\s*\(", re.M)


def declared_symbols():
    names = set()
    for h in glob.glob(os.path.join(ROOT, "include", "*.h")):
        text = open(h).read()
        text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
        body = text[text.index('extern "C"'):]
        for stmt in body.split(";"):
            stmt = stmt.strip()
            if "(" not in stmt or stmt.startswith("typedef") or stmt.startswith("#"):
                continue
            m = DECL.search(stmt)
            if m:
                names.add(m.group(1))
    return names


def test_library_exports_every_declared_symbol():
    L = ctypes.CDLL(_lib.LIB_PATH)
    syms = declared_symbols()
    assert {"dijkstra", "creategraph", "update_edge", "distance", "parent", "clear", "origin", "simplegraph",
            "readgraph", "writegraph", "sp_sssp_batch", "sp_sssp_batch_device"} <= syms
    missing = [s for s in sorted(syms) if not hasattr(L, s)]
    assert not missing, missing


def test_compute_fails_loudly_without_gpu():
    import torch
    if torch.cuda.is_available():
        return
    L = _lib.lib()
    g = L.simplegraph(True)
    assert L.sp_sync(g) != 0
    assert b"no CPU fallback" in L.sp_last_error()
