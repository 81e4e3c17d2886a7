"""The Linux `Compression` front end (cli/compression.cpp; reference: Compression/Compression.cpp:96-182, modes -c / -d / -n)
on the C ABI. CPU: argument handling, PPM reading, loud failure without a GPU. GPU: files and printed figures equal what the
reference's own encoder / decoder / calculatePSNR give."""
import os
import re
import subprocess

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
EXE = os.path.join(ROOT, "imageexperiments_b200", "Compression")


def run(*args):
    return subprocess.run([EXE, *args], capture_output=True, text=True, timeout=300)


def write_ppm(path, img, comment=False):
    h, w, _ = img.shape
    with open(path, "wb") as f:
        f.write(b"P6\n" + (b"# made by a test\n" if comment else b"") + f"{w} {h}\n255\n".encode() + img.tobytes())


def read_ppm(path):
    raw = open(path, "rb").read()
    m = re.match(rb"P6\s+(\d+)\s+(\d+)\s+255\s", raw)
    w, h = int(m.group(1)), int(m.group(2))
    return np.frombuffer(raw[m.end():], np.uint8).reshape(h, w, 3)


@pytest.fixture(scope="module")
def exe():
    if not os.path.exists(EXE):
        pytest.skip("Compression not built (run __graft_entry__.build())")
    return EXE


def test_usage_and_argument_errors(exe, tmp_path):
    for args in ([], ["-x"], ["-c", "3.5"], ["-d", "a"], ["-n", "3.5"], ["-g", "a"], ["-g", "a", "b", "c"]):
        r = run(*args)
        assert r.returncode == 255 and "usage: Compression -c" in r.stdout  # return -1, Compression.cpp:92-95
    r = run("-c", "3.5", str(tmp_path / "missing.ppm"), str(tmp_path / "o.mn"))
    assert r.returncode == 1 and "cannot open" in r.stderr
    r = run("-d", str(tmp_path), str(tmp_path / "o.ppm"))  # a folder is not a stream
    assert r.returncode == 1 and "cannot open" in r.stderr
    (tmp_path / "bad.ppm").write_bytes(b"P5\n4 4\n255\n" + bytes(16))
    r = run("-c", "3.5", str(tmp_path / "bad.ppm"), str(tmp_path / "o.mn"))
    assert r.returncode == 1 and "not a binary PPM" in r.stderr
    (tmp_path / "short.ppm").write_bytes(b"P6\n4 4\n255\n" + bytes(10))
    r = run("-c", "3.5", str(tmp_path / "short.ppm"), str(tmp_path / "o.mn"))
    assert r.returncode == 1 and "truncated" in r.stderr


def test_fails_loudly_without_a_gpu(exe, tmp_path):
    import torch
    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    from imageexperiments_b200.synth import synthetic_rgb
    write_ppm(tmp_path / "i.ppm", synthetic_rgb(16, 16, 1), comment=True)
    r = run("-c", "3.5", str(tmp_path / "i.ppm"), str(tmp_path / "o.mn"))
    assert r.returncode == 1 and "mnenc_create" in r.stderr and not (tmp_path / "o.mn").exists()


def test_sweep_without_images_needs_no_gpu(exe, tmp_path):
    """-g (Compression.cpp:303-357): recursive walk, only image files are processed, the CSV header is always written."""
    (tmp_path / "in" / "sub").mkdir(parents=True)
    (tmp_path / "in" / "notes.txt").write_text("not an image")
    (tmp_path / "in" / "sub" / "x.mn").write_bytes(b"MN24")
    r = run("-g", str(tmp_path / "in"), str(tmp_path / "stats.csv"))
    assert r.returncode == 0, r.stderr
    assert (tmp_path / "stats.csv").read_text() == "File, Mode, Quality, Size, BPP, PSNR\n"
    r = run("-g", str(tmp_path / "missing"), str(tmp_path / "stats.csv"))
    assert r.returncode == 1 and "cannot read folder" in r.stderr


def test_sweep_fails_loudly_without_a_gpu(exe, tmp_path):
    import torch
    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    from imageexperiments_b200.synth import synthetic_rgb
    (tmp_path / "in").mkdir()
    write_ppm(tmp_path / "in" / "I.PPM", synthetic_rgb(16, 16, 1))  # extension compared in lower case (:313-315)
    r = run("-g", str(tmp_path / "in"), str(tmp_path / "stats.csv"))
    assert r.returncode == 1 and "mnenc_create" in r.stderr


@pytest.mark.gpu
def test_modes_equal_the_reference(exe, tmp_path, reflib):
    from imageexperiments_b200.synth import synthetic_rgb
    img = synthetic_rgb(200, 136, 7)
    src, mnf, dec = str(tmp_path / "i.ppm"), str(tmp_path / "o.mn"), str(tmp_path / "d.ppm")
    write_ppm(src, img, comment=True)
    ctx = reflib.create(32, 8, 3.5)
    want = reflib.encode(ctx, img)
    reflib.destroy(ctx)
    want_px = reflib.decode(want)
    r = run("-c", "3.5", src, mnf)
    assert r.returncode == 0, r.stderr
    assert open(mnf, "rb").read() == want
    lines = r.stdout.strip().splitlines()
    assert lines[0] == "create basis function" and lines[1] == f"start processing image {src} using largest 32 coefficients."
    assert lines[2] == f"bpp {8 * len(want) / (200 * 136)!r} total KB {len(want) // 1024}"
    r = run("-d", mnf, dec)
    assert r.returncode == 0, r.stderr
    assert (read_ppm(dec) == want_px).all()
    r = run("-n", "3.5", src, dec)
    assert r.returncode == 0, r.stderr
    m = re.search(r"PSNR (\S+) bpp (\S+) total KB (\d+)", r.stdout)
    assert m and abs(float(m.group(1)) - reflib.psnr(img, want_px)) < 1e-9 and int(m.group(3)) == len(want) // 1024
    assert (read_ppm(dec) == want_px).all()
    ctx = reflib.create(32, 8, 3.5)
    reflib.set_quant(ctx, np.ones(96, np.float32))
    want_max = reflib.encode(ctx, img[:40, :48].copy())
    reflib.destroy(ctx)
    write_ppm(src, img[:40, :48].copy())
    assert run("-c", "max", src, mnf).returncode == 0 and open(mnf, "rb").read() == want_max
    r = run("-d", src, dec)  # not a .mn stream
    assert r.returncode == 1 and "mnenc_decode_rgb8" in r.stderr


@pytest.mark.gpu
def test_builtin_dictionary_round_trip(exe, tmp_path):
    """MNENC_DICT=builtin: the front end without any file from a reference build (mnenc_build_dictionary). The stream is this
    library's own codec revision: it must decode with the same setting to the quality the reference dictionary gives."""
    from imageexperiments_b200.synth import synthetic_rgb
    img = synthetic_rgb(200, 136, 7)
    src, mnf, dec = str(tmp_path / "i.ppm"), str(tmp_path / "o.mn"), str(tmp_path / "d.ppm")
    write_ppm(src, img)
    env = dict(os.environ, MNENC_DICT="builtin")
    r = subprocess.run([EXE, "-n", "3.5", src, dec], capture_output=True, text=True, timeout=300, env=env)
    assert r.returncode == 0, r.stderr
    m = re.search(r"PSNR (\S+) bpp (\S+)", r.stdout)
    r2 = run("-n", "3.5", src, dec)
    m2 = re.search(r"PSNR (\S+) bpp (\S+)", r2.stdout)
    assert m and m2 and abs(float(m.group(1)) - float(m2.group(1))) < 0.25 and abs(float(m.group(2)) - float(m2.group(2))) < 0.05 * float(m2.group(2))
    assert subprocess.run([EXE, "-c", "3.5", src, mnf], capture_output=True, text=True, timeout=300, env=env).returncode == 0
    assert subprocess.run([EXE, "-d", mnf, dec], capture_output=True, text=True, timeout=300, env=env).returncode == 0
    err = read_ppm(dec).astype(np.float64) - img
    assert 10 * np.log10(255.0 ** 2 / np.mean(err ** 2)) > 35.0
