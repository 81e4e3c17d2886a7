"""Generates tests/golden/golden_decode.npz from the UNMODIFIED reference build (oracle/_ref/libmnref.so): `.mn` streams
written by the reference's encodeImageFast and what the reference's own readCompressedFast / decodeImageFast make of them.
Run in the build container:  python tests/golden/make_golden_decode.py

Contents (case names in `names`)
  mn_<n>            the stream (reference encoder; K, quality and image listed in CASES)
  rgb_<n>           decodeImageFast(mn)                                   CompressedImage.cpp:837-889
  len_<n>, sym_<n>  readCompressedFast(mn): lengths, concatenated streams CompressedImage.cpp:709-782
  sizes_<n>         the 6K stream sizes
  fixture_sha256    sha256 of decodeImageFast(Data/r0c1de5e1t_3_5.mn) pixels (the file itself stays in the reference tree)
  fixture_dims      its (height, width)
"""
import hashlib
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from imageexperiments_b200.synth import synthetic_rgb  # noqa: E402
from oracle.bindings import RefLib  # noqa: E402


def cases():
    rng = np.random.default_rng(77)
    half = np.zeros((40, 96, 3), np.uint8)
    half[:, 56:] = synthetic_rgb(40, 40, 9)
    return [  # name, K, quality (None = all-ones tables, Compression.cpp:104-114), image
        ("synth", 32, 3.5, synthetic_rgb(72, 56, 21)),
        ("ragged", 32, 3.5, synthetic_rgb(61, 45, 22)),        # partial blocks on both edges
        ("one_block", 32, 3.5, synthetic_rgb(8, 8, 23)),
        ("thin", 32, 5.0, synthetic_rgb(3, 50, 24)),
        ("black", 32, 3.5, np.zeros((24, 40, 3), np.uint8)),   # every stream empty
        ("gray", 32, 3.5, np.full((320, 328, 3), 77, np.uint8)),  # run-length coded streams
        ("half", 32, 3.5, half),
        ("noise", 32, 8.0, rng.integers(0, 256, (40, 48, 3)).astype(np.uint8)),
        ("coarse", 32, 1.0, synthetic_rgb(64, 64, 25)),
        ("k8", 8, 3.5, synthetic_rgb(80, 48, 26)),
        ("max", 32, None, synthetic_rgb(24, 16, 27)),
    ]


def main():
    R = RefLib()
    out, names = {}, []
    for name, K, q, img in cases():
        ctx = R.create(K, 8, 3.5 if q is None else q)
        if q is None:
            R.set_quant(ctx, np.ones(3 * K, np.float32))
        data = R.encode(ctx, img)
        R.destroy(ctx)
        p = R.parse(data)
        names.append(name)
        out[f"mn_{name}"] = np.frombuffer(data, np.uint8)
        out[f"rgb_{name}"] = R.decode(data)
        out[f"len_{name}"] = p["lengths"]
        out[f"sizes_{name}"] = np.array([len(s) for s in p["streams"]], np.int64)
        out[f"sym_{name}"] = np.concatenate(p["streams"]) if p["streams"] else np.zeros(0, np.uint16)
    out["names"] = np.array(names)
    with open("/root/reference/Data/r0c1de5e1t_3_5.mn", "rb") as f:
        dec = R.decode(f.read())
    out["fixture_sha256"] = np.frombuffer(hashlib.sha256(dec.tobytes()).hexdigest().encode(), np.uint8)
    out["fixture_dims"] = np.array(dec.shape[:2])
    np.savez_compressed(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden_decode.npz"), **out)
    print("wrote golden_decode.npz:", names, dec.shape)


if __name__ == "__main__":
    main()
