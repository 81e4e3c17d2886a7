"""The codec's dictionary as DATA (it must be imported from the build whose decoder will read
the streams, never regenerated).

A ``.mndict`` file holds what ``compressed::CompressionContextFast`` holds
(CompressionLib/inc/CompressedImage.h:46-55): BaseDict (n_base x 64) and, per channel, the DetailBasis matrices
concatenated in atom order. It is written once per reference build by ``oracle/export_dictionary.py``
(``__graft_entry__.build()`` runs it) and read here.

Layout (little endian): b"MNDICT1\\0" | i32 n_base | i32 cols | i32 detail_rows[n_base] | f32 base[n_base*cols]
| f32 detailY | f32 detailU | f32 detailV.
"""
from __future__ import annotations

import hashlib
import os

import numpy as np

MAGIC = b"MNDICT1\0"
DEFAULT_PATH = os.path.join(os.path.dirname(os.path.abspath(__file__)), "data", "dict_bs8.mndict")


class Dictionary:
    def __init__(self, base: np.ndarray, detail_rows: np.ndarray, detail):
        self.base = np.ascontiguousarray(base, np.float32)
        self.detail_rows = np.ascontiguousarray(detail_rows, np.int32)
        self.detail = [np.ascontiguousarray(d, np.float32) for d in detail]
        n_base, cols = self.base.shape
        if self.detail_rows.shape != (n_base,) or len(self.detail) != 3:
            raise ValueError("inconsistent dictionary")
        for d in self.detail:
            if d.shape != (int(self.detail_rows.sum()), cols):
                raise ValueError("inconsistent detail table")

    @property
    def n_base(self) -> int:
        return self.base.shape[0]

    def sha256(self) -> str:
        h = hashlib.sha256()
        for a in (self.detail_rows, self.base, *self.detail):
            h.update(a.tobytes())
        return h.hexdigest()

    def save(self, path: str) -> None:
        os.makedirs(os.path.dirname(os.path.abspath(path)), exist_ok=True)
        with open(path, "wb") as f:
            f.write(MAGIC)
            f.write(np.array([self.base.shape[0], self.base.shape[1]], np.int32).tobytes())
            f.write(self.detail_rows.tobytes())
            f.write(self.base.tobytes())
            for d in self.detail:
                f.write(d.tobytes())

    @staticmethod
    def build(blockSize: int = 8, threads: int = 0) -> "Dictionary":
        """The dictionary computed here instead of imported (mnenc_build_dictionary, csrc/dictionary_host.cpp):
        shapes and segment atoms are those of the reference bit for bit, the KLT refinement rows come from a pinned eigen
        solver (double Jacobi) and therefore define their OWN codec revision - streams written with this dictionary decode with
        this dictionary, not with a reference build's."""
        import ctypes as C

        from . import _lib
        L = _lib.load()
        cols = blockSize * blockSize
        base = np.zeros((512, cols), np.float32)
        rows = np.zeros(512, np.int32)
        det = [np.zeros((512 * cols, cols), np.float32) for _ in range(3)]
        n = C.c_int(0)
        p = lambda a: a.ctypes.data_as(C.c_void_p)
        _lib.check(L.mnenc_build_dictionary(blockSize, threads, p(base), C.byref(n), p(rows), p(det[0]), p(det[1]), p(det[2])))
        nb = n.value
        total = int(rows[:nb].sum())
        return Dictionary(base[:nb].copy(), rows[:nb].copy(), [d[:total].copy() for d in det])

    @staticmethod
    def load(path: str = DEFAULT_PATH) -> "Dictionary":
        if not os.path.exists(path):
            raise FileNotFoundError(f"{path} is missing: it is exported from the reference build by __graft_entry__.build() "
                                    "(oracle/export_dictionary.py)")
        raw = np.fromfile(path, np.uint8)
        if raw[:8].tobytes() != MAGIC:
            raise ValueError("not a .mndict file")
        n_base, cols = (int(x) for x in raw[8:16].view(np.int32))
        o = 16
        rows = raw[o:o + 4 * n_base].view(np.int32).copy()
        o += 4 * n_base
        base = raw[o:o + 4 * n_base * cols].view(np.float32).reshape(n_base, cols).copy()
        o += 4 * n_base * cols
        total = int(rows.sum())
        det = []
        for _ in range(3):
            det.append(raw[o:o + 4 * total * cols].view(np.float32).reshape(total, cols).copy())
            o += 4 * total * cols
        if o != raw.size:
            raise ValueError("trailing bytes in .mndict file")
        return Dictionary(base, rows, det)
