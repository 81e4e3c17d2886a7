"""Generates tests/golden/golden.npz from the UNMODIFIED reference build (oracle/_ref/libmnref.so, which
oracle/Makefile compiles from /root/reference). Run in the build container:  python tests/golden/make_golden.py

Contents
  quant_q / quant_tables      createQuantizationTables (float) for a list of qualities   CompressedImage.cpp:168-210
  fixture_header              first 206 bytes of the reference's own Data/r0c1de5e1t_3_5.mn (magic, dims, K, bs, 96 quantisers)
  ent_*                       entropy cases: input u16 arrays, writeHuffmanOrGolomb bits, runLengthEncode output
  img_* / sym_* / file_*      small images: reference symbols (CalcMPDynamicFast per block) and encodeImageFast bytes
  dict_sha256                 checksum of the dictionary the image cases were produced with
"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from imageexperiments_b200.dictionary import Dictionary  # noqa: E402
from imageexperiments_b200.synth import synthetic_rgb  # noqa: E402
from oracle.bindings import RefLib  # noqa: E402


def entropy_cases():
    rng = np.random.default_rng(2024)
    cases = [np.zeros(0, np.uint16), np.array([0], np.uint16), np.array([7] * 9, np.uint16), np.array([65535, 0, 65535, 1], np.uint16),
             np.array([1, 1, 1, 1, 1, 1, 1, 1, 1, 1], np.uint16),  # HuffmanTest.cpp:51-78: all-ones symbol vs pseudo-EOF
             np.arange(10, dtype=np.uint16), np.array([5, 5, 5, 5, 7], np.uint16), np.full(32770, 9, np.uint16),
             np.full(50000, 3, np.uint16), np.concatenate([np.full(40000, 0, np.uint16), np.arange(100, dtype=np.uint16)])]
    for part in ([10], [9, 1], [5, 5], [4, 3, 2, 1], [2, 2, 2, 2, 2], [1] * 10, [3, 3, 3, 1], [6, 2, 1, 1]):  # HuffmanTest.cpp:16-49
        cases.append(np.concatenate([np.full(c, i * 3, np.uint16) for i, c in enumerate(part)]))
    cases.append(rng.integers(0, 4, 500).astype(np.uint16))
    cases.append((rng.geometric(0.05, 3000) - 1).astype(np.uint16))
    cases.append(rng.integers(0, 2000, 4000).astype(np.uint16))
    cases.append(np.minimum(rng.pareto(1.0, 2000) * 50, 60000).astype(np.uint16))
    cases.append(rng.integers(0, 65536, 20000).astype(np.uint16))  # HuffmanTest.cpp:93-111 (scaled down)
    return cases


def main():
    R = RefLib()
    out = {}
    qs = np.array([0.5, 1.0, 2.0, 3.0, 3.5, 4.0, 5.0, 6.0, 7.0, 8.0], np.float64)
    out["quant_q"] = qs
    out["quant_tables"] = np.stack([R.quant_tables(32, 8, q) for q in qs])
    fixture = "/root/reference/Data/r0c1de5e1t_3_5.mn"
    with open(fixture, "rb") as f:
        out["fixture_header"] = np.frombuffer(f.read(206), np.uint8)
    cases = entropy_cases()
    out["ent_count"] = np.array(len(cases))
    for i, c in enumerate(cases):
        blob, nbits = R.huffman_or_golomb(c)
        out[f"ent_{i}_in"] = c
        out[f"ent_{i}_bits"] = np.frombuffer(blob, np.uint8)
        out[f"ent_{i}_nbits"] = np.array(nbits)
        out[f"ent_{i}_rle"] = R.rle(c)
    ctx = R.create(32, 8, 3.5)
    d = R.dictionary(ctx)
    out["dict_sha256"] = np.frombuffer(Dictionary(d.base, d.detail_rows, d.detail).sha256().encode(), np.uint8)
    imgs = [("a", 40, 24, 3.5, 1), ("b", 37, 29, 3.5, 2), ("c", 64, 48, 8.0, 3), ("d", 48, 32, 1.0, 4)]
    out["img_names"] = np.array([n for n, *_ in imgs])
    for name, w, h, q, seed in imgs:
        R.set_quant(ctx, R.quant_tables(32, 8, q))
        img = synthetic_rgb(w, h, seed)
        counts, pairs = R.symbols(ctx, img)
        out[f"img_{name}"] = img
        out[f"q_{name}"] = np.array(q)
        out[f"sym_{name}_counts"] = counts
        out[f"sym_{name}_pairs"] = pairs
        out[f"file_{name}"] = np.frombuffer(R.encode(ctx, img), np.uint8)
    np.savez_compressed(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden.npz"), **out)
    print("wrote golden.npz with", len(out), "arrays")


if __name__ == "__main__":
    main()
