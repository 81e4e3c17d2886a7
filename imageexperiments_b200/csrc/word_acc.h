// Per-thread accumulator for MSB-first bit packing (BitBuffer::WriteBits, BitBuffer.cpp:78-110): codes that fall into the
// same 32-bit word of the output are OR-ed together in a register and leave the thread as ONE update of that word. A thread
// that packs consecutive symbols (mean code ≈ 4.6 bits) touches a new word every ≈7 codes, so the sink — an atomicOr on
// the device — runs ≈4x less often than with one update per code. Positions need not be contiguous (a thread's items may
// straddle two sequences): bits of different codes are disjoint, so OR-ing in any grouping gives the same stream.
//
// Plain C++ so that the host build of the tests (tests/native/word_acc_test.cpp) runs the very code the kernel runs.
#pragma once
#include <cstdint>

#if defined(__CUDACC__)
#define MNENC_HD __host__ __device__ __forceinline__
#else
#define MNENC_HD inline
#endif

namespace mnenc {

// Logical big-endian word -> the little-endian value whose bytes are the stream's bytes 4w..4w+3.
MNENC_HD uint32_t streamWord(uint32_t bigEndianBits) {
#if defined(__CUDA_ARCH__)
    return __byte_perm(bigEndianBits, 0, 0x0123);
#else
    return (bigEndianBits >> 24) | ((bigEndianBits >> 8) & 0xFF00u) | ((bigEndianBits << 8) & 0xFF0000u) | (bigEndianBits << 24);
#endif
}

template <class Sink>  // sink(wordIndex, valueToOr): valueToOr is already in stream byte order
struct WordAcc {
    Sink sink;
    unsigned long long word = ~0ull;  // index of the pending word
    uint32_t bits = 0;                // its pending bits, logical big-endian

    MNENC_HD explicit WordAcc(Sink s) : sink(s) {}
    MNENC_HD void flush() {
        if (bits) sink(word, streamWord(bits));
        bits = 0;
    }
    // the low `width` (<= 32) bits of `value`, MSB first, at absolute bit position `pos`; higher bits of `value` must be 0
    MNENC_HD void put(unsigned long long pos, uint32_t value, int width) {
        if (width == 0) return;
        const unsigned long long w = pos >> 5;
        const int off = static_cast<int>(pos & 31);
        const unsigned long long v = static_cast<unsigned long long>(value) << (64 - off - width);
        if (w != word) {
            flush();
            word = w;
        }
        bits |= static_cast<uint32_t>(v >> 32);
        if (off + width > 32) {
            flush();
            word = w + 1;
            bits = static_cast<uint32_t>(v);
        }
    }
    // `quo` ones, a zero, then the truncated-binary remainder of writeGolombCode (BitBuffer.cpp:228-241); returns the end position
    MNENC_HD unsigned long long putGolomb(unsigned long long pos, uint32_t sym, uint32_t M, int bitWidthM) {
        uint32_t quo = sym / M;
        const uint32_t rem = sym - quo * M;
        while (quo >= 32) {
            put(pos, 0xFFFFFFFFu, 32);
            pos += 32;
            quo -= 32;
        }
        put(pos, ((1u << quo) - 1u) << 1, static_cast<int>(quo) + 1);
        pos += quo + 1;
        const uint32_t limit = (1u << (bitWidthM + 1)) - M;
        if (rem < limit) {
            put(pos, rem, bitWidthM);
            return pos + bitWidthM;
        }
        put(pos, rem + limit, bitWidthM + 1);
        return pos + bitWidthM + 1;
    }
};

// Which sequence of the concatenation an element belongs to: start[0 .. nSeq] ascending, empty sequences repeat their
// successor's start, the owner of element g is the LARGEST q < nSeq with start[q] <= g. A thread that walks consecutive
// elements searches once and then only steps forward.
struct SeqWalk {
    const uint32_t* start;
    int nSeq;
    int q;
    uint32_t seqStart, seqEnd;  // elements [seqStart, seqEnd) belong to q

    MNENC_HD SeqWalk(const uint32_t* s, int n, uint32_t g) : start(s), nSeq(n), q(0), seqStart(0), seqEnd(0) {
        int lo = 0, hi = n;
        while (hi - lo > 1) {
            const int mid = (lo + hi) >> 1;
            if (start[mid] <= g) lo = mid;
            else hi = mid;
        }
        q = lo;
        load();
    }
    MNENC_HD void load() {
        seqStart = start[q];
        seqEnd = q + 1 < nSeq ? start[q + 1] : 0xFFFFFFFFu;
    }
    // move to the owner of g (g not below the previous element); true when the sequence changed
    MNENC_HD bool advance(uint32_t g) {
        if (g < seqEnd) return false;
        while (q + 1 < nSeq && start[q + 1] <= g) ++q;
        load();
        return true;
    }
};

}  // namespace mnenc
