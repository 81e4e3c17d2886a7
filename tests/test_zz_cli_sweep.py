"""`Compression -g` (cli/compression.cpp; reference: Compression/Compression.cpp:303-357, the "MP" rows): every row of the
report equals what the reference's own createCompressionContextFast / encodeImageFast / decodeImageFast / calculatePSNR give
for that image and quality. (Kept in its own file, after the others: the sweep is the newest mode of the front end.)"""
import os

import pytest

from test_cli import EXE, run, write_ppm


def fmt(x: float) -> str:
    """std::format("{}", double): shortest round-trip representation, integral values without ".0"."""
    return repr(int(x)) if x == int(x) and abs(x) < 1e15 else repr(x)


@pytest.mark.gpu
def test_sweep_rows_equal_the_reference(tmp_path, reflib):
    if not os.path.exists(EXE):
        pytest.skip("Compression not built (run __graft_entry__.build())")
    from imageexperiments_b200.synth import synthetic_rgb
    folder = tmp_path / "in"
    (folder / "sub").mkdir(parents=True)
    images = {str(folder / "a.ppm"): synthetic_rgb(72, 40, 3), str(folder / "sub" / "b.PPM"): synthetic_rgb(33, 50, 4)}
    for path, img in images.items():
        write_ppm(path, img)
    (folder / "readme.txt").write_text("skipped")
    r = run("-g", str(folder), str(tmp_path / "stats.csv"))
    assert r.returncode == 0, r.stderr
    rows = (tmp_path / "stats.csv").read_text().splitlines()
    assert rows[0] == "File, Mode, Quality, Size, BPP, PSNR" and len(rows) == 1 + 8 * len(images)
    k = 1
    for path in sorted(images):
        img = images[path]
        for q in (8.0, 7.0, 6.0, 5.0, 4.0, 3.0, 2.0, 1.0):
            ctx = reflib.create(32, 8, q)
            want = reflib.encode(ctx, img)
            reflib.destroy(ctx)
            psnr = reflib.psnr(img, reflib.decode(want))
            f, mode, quality, size, bpp, p = rows[k].split(", ")
            assert (f, mode, quality, int(size)) == (path, "MP", fmt(q), len(want)), rows[k]
            assert float(bpp) == 8 * len(want) / (img.shape[0] * img.shape[1])
            assert abs(float(p) - psnr) < 1e-9, (rows[k], psnr)
            assert f"MP Quality {fmt(q)} PSNR {p} bpp {bpp} total KB {len(want)}" in r.stdout
            k += 1
