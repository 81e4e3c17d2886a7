"""imageexperiments_b200 — B200-native encoder hot path (and its decoder) of mnesbit/ImageExperiments' CompressionLib (`.mn` format).

Only the path BASELINE.json names lives here: csrc/ (CUDA kernels + C ABI, built into libmnenc.so) and the
host-side mirror of the reference's encoder interface (compressed.py; stripes.py shards one image over the GPUs of a
process group).
"""
from .compressed import (calculatePSNR, CompressionContextFast, MnBuffer, createCompressionContextFast, createQuantizationTables, encodeImageBatch, encodeImageFast,
                         decodeImageBatch, decodeImageFast, parseStream, streamInfo, lastSymbols, packSymbols, planSequenceHost)
from .dictionary import Dictionary
from .stripes import STRIPE_PHASES, TorchComm, ThreadComm, encodeImageStripes, encodeImageStriped, initStripeComm, stripeColumns
from ._lib import MnencError

__all__ = ["calculatePSNR", "CompressionContextFast", "MnBuffer", "createCompressionContextFast", "createQuantizationTables", "encodeImageFast", "encodeImageBatch", "decodeImageBatch", "decodeImageFast", "parseStream", "streamInfo", "lastSymbols",
           "packSymbols", "planSequenceHost", "Dictionary", "MnencError", "TorchComm", "ThreadComm", "encodeImageStripes", "encodeImageStriped", "initStripeComm", "STRIPE_PHASES",
           "stripeColumns"]
