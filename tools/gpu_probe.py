"""Developer probe for a GPU box: parity vs oracle on small images, stage timings on the RAISE-sized image."""
import os, sys, time
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import imageexperiments_b200 as mn
from imageexperiments_b200.synth import synthetic_rgb
from oracle.bindings import Dictionary as ODict, OracleLib, RefLib, ref_available

def main():
    sizes = [(64, 64), (200, 136), (512, 512)]
    q = float(os.environ.get("Q", "3.5"))
    ctx = mn.createCompressionContextFast(32, 8, q, device=0)
    d = ctx.dictionary
    oc = OracleLib().create(ODict(32, 8, d.base, d.detail_rows, d.detail, ctx.Quant))
    for (w, h) in sizes:
        img = synthetic_rgb(w, h, seed=w + h)
        t = time.time(); data = mn.encodeImageFast(img, ctx); dt = time.time() - t
        counts, pairs = mn.lastSymbols(ctx, w, h)
        oc_counts, oc_pairs = oc.symbols(img)
        same_c = (counts == oc_counts); same_p = (pairs == oc_pairs).all(axis=(2, 3))
        bc_ok = (same_c & same_p)
        print(f"{w}x{h}: gpu {len(data)} B in {dt*1e3:.1f} ms; block-channel parity {bc_ok.mean()*100:.3f}% counts mean {counts.mean(0)}")
        if not bc_ok.all():
            bad = np.argwhere(~bc_ok)[:5]
            for b, c in bad:
                print("  mismatch blk", b, "ch", c, "gpu", counts[b, c], pairs[b, c, :4].tolist(), "oracle", oc_counts[b, c], oc_pairs[b, c, :4].tolist())
        ref = oc.lib.write_file(32, 8, w, h, ctx.Quant, counts, pairs)
        print("   back end bit-exact given GPU symbols:", ref == data, len(ref), len(data))
        if ref != data:
            n = min(len(ref), len(data)); a = np.frombuffer(ref[:n], np.uint8); b = np.frombuffer(data[:n], np.uint8)
            diff = np.nonzero(a != b)[0]
            print("   first differing byte", diff[:5], "of", n)
        ms, launches = ctx.last_timing()
        print("   stage ms", np.round(ms, 3).tolist(), "launches", launches)
    if os.environ.get("BIG", "1") == "1":
        w, h = 4928, 3264
        img = synthetic_rgb(w, h, seed=12345)
        for it in range(4):
            t = time.time(); data = mn.encodeImageFast(img, ctx); dt = time.time() - t
            ms, launches = ctx.last_timing()
            fl, ns, nb = ctx.last_work()
            print(f"C2 iter {it}: {dt*1e3:.1f} ms -> {w*h/1e6/dt:.1f} MP/s; {len(data)} B; stage ms {np.round(ms,3).tolist()} launches {launches} flops {fl/1e9:.1f} G symbols {ns}")
        counts, pairs = mn.lastSymbols(ctx, w, h)
        ref = oc.lib.write_file(32, 8, w, h, ctx.Quant, counts, pairs)
        print("C2 back end bit-exact given GPU symbols:", ref == data)
        # oracle symbols on a sample of block columns
        nbx = w // 8
        cols = [0, 100, 333, nbx - 1]
        nby = h // 8
        ok = tot = 0
        for bx in cols:
            occ, ocp = oc.symbols(img, bx, bx + 1)
            g_c = counts[bx * nby:(bx + 1) * nby]; g_p = pairs[bx * nby:(bx + 1) * nby]
            good = (g_c == occ) & (g_p == ocp).all(axis=(2, 3))
            ok += good.sum(); tot += good.size
        print(f"C2 sampled block-channel parity {ok}/{tot}")
        if ref_available():
            R = RefLib(); t = time.time(); dec = R.decode(data); print("ref decode s", time.time() - t, "PSNR", R.psnr(img, dec))

if __name__ == "__main__":
    main()
