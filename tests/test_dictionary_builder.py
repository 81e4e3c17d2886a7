"""Dictionary builder (csrc/dictionary_host.cpp) against the dictionary exported from the reference build
(oracle/export_dictionary.py -> imageexperiments_b200/data/dict_bs8.mndict): CPU only.

Shapes, group heights and the 510 segment atoms must equal the reference's bit for bit. The KLT refinement rows come from a
pinned solver (double Jacobi) instead of Eigen's float SelfAdjointEigenSolver, so they are checked through what both must
satisfy: per group, orthonormal rows that vanish outside their segment, sum to zero inside it and span the same space as the
reference's rows; the well-conditioned leading rows of the luma model also agree individually."""
import numpy as np
import pytest

from imageexperiments_b200.dictionary import Dictionary


@pytest.fixture(scope="module")
def built():
    return Dictionary.build()


@pytest.fixture(scope="module")
def ref():
    return Dictionary.load()


def test_shapes_heights_and_segment_atoms_are_bit_exact(built, ref):
    assert built.n_base == ref.n_base == 510
    assert np.array_equal(built.detail_rows, ref.detail_rows)
    assert sorted(set(built.detail_rows.tolist())) == [62, 63]
    assert np.array_equal(built.base.view(np.uint32), ref.base.view(np.uint32))
    # atoms 0 and 509 are the two constant shapes (+1/8 and -1/8), every other atom has zero mean and unit norm
    assert np.all(built.base[0] == 0.125) and np.all(built.base[-1] == -0.125)
    assert np.abs(built.base[1:-1].sum(1)).max() < 1e-5 and np.abs(np.linalg.norm(built.base, axis=1) - 1).max() < 1e-5


def test_build_is_deterministic_and_independent_of_the_thread_count(built):
    again = Dictionary.build(threads=3)
    assert again.sha256() == built.sha256()


def test_built_dictionary_is_the_pinned_codec_revision(built):
    """The computed dictionary defines its own codec revision: a stream written with it only decodes with the very same floats.
    Its hash is pinned here, so that drift in the host's libm (the covariance model calls exp) or in the builder shows up as a
    test failure instead of as silently wrong pictures. After a DELIBERATE change of the builder: regenerate
    tests/golden/built_dictionary.sha256 with Dictionary.build().sha256() and say in DESIGN.md 6 that old streams are orphaned."""
    import os
    path = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "built_dictionary.sha256")
    assert built.sha256() == open(path).read().strip()


def test_refinement_rows_have_the_structure_of_the_reference(built, ref):
    off = np.concatenate([[0], np.cumsum(built.detail_rows)])
    rng = np.random.default_rng(0)
    atoms = sorted(set([0, 1, 254, 255, 508, 509] + rng.choice(510, 40, replace=False).tolist()))
    for ch in range(3):
        for a in atoms:
            g = built.detail[ch][off[a]:off[a + 1]].astype(np.float64)
            r = ref.detail[ch][off[a]:off[a + 1]].astype(np.float64)
            # orthonormal rows
            assert np.abs(g @ g.T - np.eye(len(g))).max() < 5e-5, (ch, a)
            # every row lives on ONE of the two segments of the shape (two disjoint supports) and sums to zero there
            sup = np.abs(g) > 1e-7
            first = (sup & sup[0]).any(axis=1)
            union_a, union_b = sup[first].any(axis=0), sup[~first].any(axis=0)
            assert not (union_a & union_b).any(), (ch, a)
            assert np.abs(g.sum(1)).max() < 1e-5, (ch, a)
            # same span as the reference's rows (projectors agree)
            assert np.abs(g.T @ g - r.T @ r).max() < 2e-4, (ch, a)


def test_leading_luma_rows_agree_with_the_reference(built, ref):
    """The luma model's leading eigenvalues are well separated: there the two solvers must find the same vectors."""
    off = np.concatenate([[0], np.cumsum(built.detail_rows)])
    dots = []
    for a in range(1, 509):
        g, r = built.detail[0][off[a]:off[a] + 4], ref.detail[0][off[a]:off[a] + 4]
        dots.append(np.abs((g * r).sum(1)))
    dots = np.asarray(dots)
    assert np.median(dots) > 0.99999 and dots.min() > 0.999
