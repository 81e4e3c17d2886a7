import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a B200 (run with -m gpu on the GPU box)")


@pytest.fixture(scope="session")
def golden():
    return np.load(os.path.join(ROOT, "tests", "golden", "golden.npz"))


@pytest.fixture(scope="session")
def reflib():
    from oracle.bindings import RefLib, ref_available
    if not ref_available():
        pytest.skip("oracle/_ref/libmnref.so not built (needs /root/reference; run __graft_entry__.build())")
    return RefLib()


@pytest.fixture(scope="session")
def oraclelib():
    from oracle.bindings import OracleLib, oracle_available
    if not oracle_available():
        pytest.skip("oracle/libmnoracle.so not built (run __graft_entry__.build())")
    return OracleLib()


@pytest.fixture(scope="session")
def dictionary(golden):
    from imageexperiments_b200.dictionary import DEFAULT_PATH, Dictionary
    if not os.path.exists(DEFAULT_PATH):
        pytest.skip("dictionary data file missing (exported from the reference build by __graft_entry__.build())")
    d = Dictionary.load()
    # refuse to compare against golden vectors produced with another dictionary build
    assert d.sha256() == golden["dict_sha256"].tobytes().decode(), "dictionary differs from the one the golden vectors were made with"
    return d


def make_oracle_ctx(oraclelib, dictionary, quant, K=32):
    from oracle.bindings import Dictionary as ODict
    return oraclelib.create(ODict(K, 8, dictionary.base, dictionary.detail_rows, dictionary.detail, quant))


@pytest.fixture(scope="session")
def oracle_ctx_factory(oraclelib, dictionary):
    return lambda quant, K=32: make_oracle_ctx(oraclelib, dictionary, quant, K)


def bits_of(blob: bytes, nbits: int) -> np.ndarray:
    return np.unpackbits(np.frombuffer(blob, np.uint8))[:nbits]
