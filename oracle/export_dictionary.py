"""Exports the dictionary of the reference build in oracle/_ref (createCompressionContextFast,
CompressionLib/src/CompressedImage.cpp:317-341) to imageexperiments_b200/data/dict_bs8.mndict, the data file the
encoder imports. Run by __graft_entry__.build(); needs oracle/_ref/libmnref.so."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

from imageexperiments_b200.dictionary import DEFAULT_PATH, Dictionary  # noqa: E402
from oracle.bindings import RefLib  # noqa: E402


def main(path: str = DEFAULT_PATH) -> str:
    R = RefLib()
    ctx = R.create(32, 8, 3.5)  # the dictionary does not depend on K or the quality
    d = R.dictionary(ctx)
    R.destroy(ctx)
    out = Dictionary(d.base, d.detail_rows, d.detail)
    out.save(path)
    return out.sha256()


if __name__ == "__main__":
    print(main())
