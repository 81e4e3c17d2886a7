// Host side of the decoder: the `.mn` bit stream back into `lengths` and the 6K symbol streams.
// Replaces, on the decode path,
//   compressed::readCompressedFast      CompressionLib/src/CompressedImage.cpp:709-782
//   compressed::readHuffmanOrGolomb     CompressionLib/src/CompressedImage.cpp:392-401
//   huffman::huffmanDecode              CompressionLib/src/Huffman.cpp:173-244
//   huffman::runLengthDecode            CompressionLib/src/Huffman.cpp:281-307
//   bitbuffer::readGolombCode           CompressionLib/src/BitBuffer.cpp:243-258
//   bitbuffer::readEliasFanoSequenceCode CompressionLib/src/BitBuffer.cpp:319-343
//
// Why this stage stays on the host: the format has no stream offsets — a Huffman payload ends at its pseudo-EOF code, so
// sequence i + 1 starts wherever decoding sequence i stopped. The bit stream of one image is one serial chain; a GPU
// thread would walk it an order of magnitude slower than a CPU core. The reference walks it bit by bit through an
// unordered_map (Huffman.cpp:228-241); here canonical codes are resolved by a 12-bit table plus first-code compares,
// symbols go straight into their final place in pinned memory, and the pixels are rebuilt on the GPU (decode_kernels.cu).
//
// Every length and count that the reference trusts is checked: a damaged stream gives an error, never an out-of-range access.
#include <sched.h>

#include <algorithm>
#include <atomic>
#include <bit>
#include <cmath>
#include <cstring>
#include <exception>
#include <functional>
#include <mutex>
#include <string>
#include <vector>

#include "host_pool.h"
#include "spin_team.h"
#include "mnenc_internal.h"

namespace mnenc {
namespace {

struct Bad {
    const char* why;
};

// MSB-first reader over the caller's bytes. `buf` holds `cnt` valid bits, left aligned; reading past the end yields zeros
// (BitBuffer::PeekBits clamps the same way) and is reported by overrun().
class BitReader {
public:
    BitReader(const uint8_t* p, size_t n) : base_(p), ptr_(p), end_(p + n) {}
    uint64_t peek(int n) {  // n in 1..56
        if (cnt_ < n) refill();
        return buf_ >> (64 - n);
    }
    void skip(int n) {
        buf_ <<= n;
        cnt_ -= n;
    }
    uint64_t read(int n) {  // n in 0..56
        if (n == 0) return 0;
        const uint64_t v = peek(n);
        skip(n);
        return v;
    }
    uint32_t read32() {
        const uint32_t hi = static_cast<uint32_t>(read(16));
        return hi << 16 | static_cast<uint32_t>(read(16));
    }
    bool overrun() const { return cnt_ < 0; }
    void seek(uint64_t bitPos) {  // absolute position; must lie inside the data
        ptr_ = base_ + (bitPos >> 3);
        buf_ = 0, cnt_ = 0;
        if (bitPos & 7) {
            refill();
            skip(static_cast<int>(bitPos & 7));
        }
    }
    uint64_t position() const { return static_cast<uint64_t>(ptr_ - base_) * 8 - static_cast<uint64_t>(cnt_); }
    // number of leading one bits (unary quotient of a Golomb code), consumed together with the terminating zero
    uint32_t readUnaryOnes() {
        uint32_t q = 0;
        for (;;) {
            if (cnt_ < 32) refill();
            if (cnt_ <= 0) throw Bad{"unary code runs past the end of the stream"};
            const int valid = cnt_ < 56 ? cnt_ : 56;
            const uint64_t inv = ~buf_ | ((1ull << (64 - valid)) - 1ull);  // a zero bit, or the end of the valid window
            const int ones = std::countl_zero(inv);
            if (ones < valid) {
                skip(ones + 1);
                return q + static_cast<uint32_t>(ones);
            }
            skip(valid);
            q += static_cast<uint32_t>(valid);
            if (q > (1u << 20)) throw Bad{"unary code too long"};
        }
    }
    // number of leading zero bits before the next one bit (Elias-Fano bucket advance), consumed together with the one
    uint32_t readUnaryZeros() {
        uint32_t q = 0;
        for (;;) {
            if (cnt_ < 32) refill();
            if (cnt_ <= 0) throw Bad{"bucket code runs past the end of the stream"};
            const int valid = cnt_ < 56 ? cnt_ : 56;
            const uint64_t w = buf_ & ~((1ull << (64 - valid)) - 1ull);
            const int zeros = w ? std::countl_zero(w) : 64;
            if (zeros < valid) {
                skip(zeros + 1);
                return q + static_cast<uint32_t>(zeros);
            }
            skip(valid);
            q += static_cast<uint32_t>(valid);
            if (q > 65536u) throw Bad{"bucket code too long"};
        }
    }

    void refill() {
        if (cnt_ < 0) return;  // already past the end: keep delivering zeros
        if (end_ - ptr_ >= 8) {
            uint64_t w;
            std::memcpy(&w, ptr_, 8);
            w = __builtin_bswap64(w);
            buf_ |= cnt_ ? w >> cnt_ : w;
            const int take = (63 - cnt_) >> 3;  // whole bytes that fit
            ptr_ += take;
            cnt_ += take * 8;
            if (cnt_ < 64) buf_ &= ~((1ull << (64 - cnt_)) - 1ull);  // drop the partial byte; it is read again next time
        } else {
            while (cnt_ <= 56 && ptr_ < end_) {
                buf_ |= static_cast<uint64_t>(*ptr_++) << (56 - cnt_);
                cnt_ += 8;
            }
        }
    }
    // state (public: HuffTable::decode keeps it in registers across its fast loop)
    const uint8_t* base_;
    const uint8_t* ptr_;
    const uint8_t* end_;
    uint64_t buf_ = 0;
    int cnt_ = 0;
};

// BitBuffer.cpp:345-354 (decision bound, as in the encoder)
inline uint32_t eliasFanoBound(uint32_t n, uint32_t maxSymbol) {
    if (n == 0) return 0;
    const uint32_t m = std::bit_width(maxSymbol), nb = std::bit_width(n), lb = m >= nb ? m - nb : 0;
    return n * (1 + lb) + (1u << (m - lb)) - 1;
}

inline int32_t unzigzag(uint32_t x) { return static_cast<int32_t>((x >> 1) ^ -(static_cast<int64_t>(x) & 1)); }  // BitBuffer.h:112-114

constexpr int kLutBits = 12;
constexpr int kMaxCodeLen = 48;

// Canonical Huffman table of one sequence (Huffman.cpp:176-226): counts per length, symbols in (length, symbol) order, the
// last one being the pseudo-EOF written as the all-ones `mask`.
struct HuffTable {
    int maxLen = 0;
    uint32_t total = 0;
    std::vector<uint16_t> symbols;           // [total]; index total-1 = EOF
    uint64_t firstCode[kMaxCodeLen + 2] = {};  // first code of each length
    uint32_t firstIndex[kMaxCodeLen + 2] = {};
    uint32_t count[kMaxCodeLen + 2] = {};
    std::vector<uint32_t> lut;               // [1 << kLutBits]: index << 8 | length, 0 = not resolved by the table
    std::vector<uint32_t> fast;              // same slots: symbol << 8 | length, or kSpecial (pseudo-EOF, long code, no code)
    std::vector<uint64_t> pair;              // same slots, up to THREE codes: bits 0-7 total length, 8-15 symbols (0 = special),
                                             // 16-31 first symbol, 32-47 second, 48-63 third

    void read(BitReader& in) {
        maxLen = static_cast<int>(in.read(8));
        if (maxLen > kMaxCodeLen) throw Bad{"Huffman code longer than 48 bits"};
        total = 0;
        for (int len = 1; len <= maxLen; ++len) total += (count[len] = static_cast<uint32_t>(in.read(16)));
        if (total == 0 || total > 65536u) throw Bad{"Huffman table without pseudo-EOF or with too many symbols"};
        const int width = static_cast<int>(in.read(8));
        if (width > 16) throw Bad{"symbol width above 16 bits"};
        const uint32_t mask = (1u << width) - 1u;
        symbols.resize(total);
        uint32_t at = 0;
        for (int len = 1; len <= maxLen; ++len) {
            const uint32_t n = count[len];
            if (eliasFanoBound(n, mask) < n * static_cast<uint32_t>(width)) {  // Huffman.cpp:187-190
                const uint32_t m = std::bit_width(mask), nb = std::bit_width(n), lb = m >= nb ? m - nb : 0;
                uint32_t bucket = 0;
                for (uint32_t i = 0; i < n; ++i) {
                    bucket += in.readUnaryZeros();
                    if (bucket > 65535u) throw Bad{"Elias-Fano bucket out of range"};
                    symbols[at + i] = static_cast<uint16_t>(bucket << lb);
                }
                for (uint32_t i = 0; i < n; ++i) symbols[at + i] |= static_cast<uint16_t>(in.read(static_cast<int>(lb)));
            } else {
                for (uint32_t i = 0; i < n; ++i) symbols[at + i] = static_cast<uint16_t>(in.read(width));
            }
            at += n;
        }
        if (symbols[total - 1] != mask) throw Bad{"last table symbol is not the pseudo-EOF"};  // Huffman.cpp:212-215
        if (in.overrun()) throw Bad{"Huffman table runs past the end of the stream"};
        // canonical codes: code <<= (length - prevLength) before each entry, ++code after (Huffman.cpp:200-226)
        uint64_t code = 0;
        int prevLen = 0;
        uint32_t index = 0;
        lut.assign(size_t{1} << kLutBits, 0u);
        fast.assign(size_t{1} << kLutBits, kSpecial);
        for (int len = 1; len <= maxLen; ++len) {
            if (count[len] == 0) continue;
            code <<= (len - prevLen);
            prevLen = len;
            firstCode[len] = code;
            firstIndex[len] = index;
            if (code + count[len] > (1ull << len)) throw Bad{"Huffman code lengths oversubscribed"};
            if (len <= kLutBits) {
                for (uint32_t i = 0; i < count[len]; ++i) {
                    const uint32_t entry = (index + i) << 8 | static_cast<uint32_t>(len);
                    const size_t lo = static_cast<size_t>(code + i) << (kLutBits - len);
                    std::fill(lut.begin() + lo, lut.begin() + lo + (size_t{1} << (kLutBits - len)), entry);
                    if (index + i != total - 1)
                        std::fill(fast.begin() + lo, fast.begin() + lo + (size_t{1} << (kLutBits - len)),
                                  static_cast<uint32_t>(symbols[index + i]) << 8 | static_cast<uint32_t>(len));
                }
            }
            code += count[len];
            index += count[len];
        }
        // second code of a slot: whatever single code starts in the bits the first one leaves over
        pair.assign(size_t{1} << kLutBits, 0ull);
        for (uint32_t slot = 0; slot < (1u << kLutBits); ++slot) {
            const uint32_t e0 = fast[slot];
            if (e0 >= kSpecial) continue;
            const uint32_t l0 = e0 & 0xFF;
            uint64_t v = static_cast<uint64_t>(e0 >> 8) << 16 | 1u << 8 | l0;
            const uint32_t e1 = fast[(slot << l0) & ((1u << kLutBits) - 1u)];
            if (e1 < kSpecial && l0 + (e1 & 0xFF) <= static_cast<uint32_t>(kLutBits)) {
                const uint32_t l01 = l0 + (e1 & 0xFF);
                v = static_cast<uint64_t>(e1 >> 8) << 32 | static_cast<uint64_t>(e0 >> 8) << 16 | 2u << 8 | l01;
                const uint32_t e2 = fast[(slot << l01) & ((1u << kLutBits) - 1u)];
                if (e2 < kSpecial && l01 + (e2 & 0xFF) <= static_cast<uint32_t>(kLutBits))
                    v = static_cast<uint64_t>(e2 >> 8) << 48 | static_cast<uint64_t>(e1 >> 8) << 32 | static_cast<uint64_t>(e0 >> 8) << 16 | 3u << 8 |
                        (l01 + (e2 & 0xFF));
            }
            pair[slot] = v;
        }
    }

    // Payload codes up to the pseudo-EOF into out[n0, cap); returns the number of symbols (Huffman.cpp:227-243).
    // fastRun: one unconditional 8-byte refill, then up to four table hits (4 x 12 <= 56 valid bits), each giving one to three
    // symbols, without any end-of-stream or capacity test; everything else (long codes, the pseudo-EOF, the last bytes, the
    // last symbols) goes through step().
    size_t decode(BitReader& in, uint16_t* out, size_t cap, size_t n0 = 0) const {
        size_t n = n0;
        for (;;) {
            fastRun(in, out, n, cap, ~0ull);
            if (step(in, out, n, cap)) return n;
        }
    }

    // One sequence split over the threads of `pool` (see decodeParallel below).
    struct Chunk {
        uint64_t from = 0, until = 0;   // bit range the chunk's worker is responsible for
        std::vector<uint16_t> syms;     // its symbols (chunk 0 writes straight into the output instead)
        std::vector<uint64_t> bnd;      // bit position at which each of its first kBoundaries codes starts
        size_t count = 0;               // symbols decoded
        uint64_t endPos = 0;            // where it stopped: a code boundary of ITS walk at or after `until`
        bool eof = false;               // it stopped at a pseudo-EOF code ...
        uint64_t eofEnd = 0;            // ... which ends here
        bool failed = false;            // chunk 0 only: invalid code, capacity exceeded
        size_t lastBad = 0;             // blind workers: symbols decoded when they last met bits that are no code (0 = never)
    };
    static constexpr size_t kBoundaries = 2048;
    static constexpr uint64_t kMinChunkBits = 1u << 16;

    // Huffman payloads self-synchronise: a decoder started at an arbitrary bit usually falls into step with the true code
    // boundaries after a few dozen codes. The payload's bit length is estimated from the table (sum over the codes of
    // length x 2^-length, i.e. the symbol probabilities the lengths imply) and the expected symbol count; it is cut into
    // chunks; chunk 0 is the true walk and writes into `out`, every other worker starts blind, keeps its symbols aside and
    // records where its first codes begin. The chunks are then stitched in order: the true walk continues serially from the
    // end of the previous chunk until it stands on one of the next worker's recorded boundaries — from there on that worker's
    // symbols ARE the true walk's. A chunk that never falls into step (fixed-length-like codes) is walked serially; a wrong
    // estimate costs a serial tail or some discarded work. The result is the serial decoder's, symbol for symbol; on
    // anything unexpected the sequence is simply decoded again serially from its start (which also raises the proper error).
    size_t decodeParallel(BitReader& in, uint16_t* out, size_t cap, size_t expect, SpinTeam& pool, std::vector<Chunk>& chunks) const {
        const uint64_t start = in.position(), totalBits = static_cast<uint64_t>(in.end_ - in.base_) * 8;
        double num = 0.0, den = 0.0;
        int minLen = 0;
        for (int l = 1; l <= maxLen; ++l) {
            if (!count[l]) continue;
            if (!minLen) minLen = l;
            const double p = std::ldexp(1.0, -l);
            num += count[l] * p * l, den += count[l] * p;
        }
        if (den <= 0.0 || in.cnt_ < 0) return decode(in, out, cap);
        const uint64_t est = std::min<uint64_t>(static_cast<uint64_t>(static_cast<double>(expect) * (num / den) * 1.02) + 64, totalBits - std::min(totalBits, start));
        const size_t nChunks = static_cast<size_t>(std::min<uint64_t>(2 * static_cast<uint64_t>(pool.size()), est / kMinChunkBits));
        if (nChunks < 2) return decode(in, out, cap);
        const uint64_t chunkBits = (est + nChunks - 1) / nChunks;
        const size_t symCap = static_cast<size_t>(chunkBits / static_cast<uint64_t>(minLen)) + 64;
        if (symCap * nChunks > (size_t{64} << 20)) return decode(in, out, cap);
        if (chunks.size() < nChunks) chunks.resize(nChunks);
        for (size_t k = 0; k < nChunks; ++k) {
            chunks[k].from = start + k * chunkBits;
            chunks[k].until = std::min(start + (k + 1) * chunkBits, totalBits);
            if (k && chunks[k].syms.size() < symCap) chunks[k].syms.resize(symCap);
        }
        const BitReader proto = in;
        pool.forEach(static_cast<int>(nChunks), [&](int k) {
            Chunk& c = chunks[static_cast<size_t>(k)];
            if (k == 0)
                runChunk(proto, c, out, cap, 0);
            else
                runChunk(proto, c, c.syms.data(), symCap, kBoundaries);
        });
        // stitch
        const Chunk& c0 = chunks[0];
        if (c0.failed) {
            in.seek(start);
            return decode(in, out, cap);
        }
        size_t n = c0.count;
        uint64_t pos = c0.endPos;
        bool done = c0.eof, positioned = false;
        uint64_t endBit = c0.eofEnd;
        for (size_t k = 1; k < nChunks && !done; ++k) {
            const Chunk& c = chunks[k];
            size_t bi = 0;
            bool synced = false;
            while (pos < c.until) {
                while (bi < c.bnd.size() && c.bnd[bi] < pos) ++bi;
                if (bi < c.bnd.size() && c.bnd[bi] == pos && bi <= c.count) {
                    synced = true;
                    break;
                }
                if (!positioned) in.seek(pos), positioned = true;  // the true walk, one code
                if (step(in, out, n, cap)) {
                    done = true;
                    endBit = in.position();
                    break;
                }
                pos = in.position();
            }
            if (synced) {
                if (c.lastBad > bi + 1) {  // bits that are no code AFTER the walks fell into step: the stream is damaged;
                    in.seek(start);        // the serial decoder reports it
                    return decode(in, out, cap);
                }
                const size_t add = c.count - bi;
                if (n + add > cap) throw Bad{"sequence longer than the block counts allow"};
                std::memcpy(out + n, c.syms.data() + bi, add * sizeof(uint16_t));
                n += add;
                pos = c.endPos;
                positioned = false;
                if (c.eof) done = true, endBit = c.eofEnd;
            }
        }
        if (done) {
            in.seek(endBit);
            return n;
        }
        in.seek(pos);  // estimate too short: the rest serially
        return decode(in, out, cap, n);
    }

    static constexpr uint32_t kSpecial = 1u << 24;  // `fast` entry: pseudo-EOF or a code longer than the table

private:
    // Table hits only, while at least 8 input bytes and 8 output slots remain and the position is below `untilBit`.
    void fastRun(BitReader& in, uint16_t* out, size_t& n, size_t cap, uint64_t untilBit) const {
        while (in.end_ - in.ptr_ >= 8 && n + 16 <= cap && in.cnt_ >= 0 &&
               static_cast<uint64_t>(in.ptr_ - in.base_) * 8 - static_cast<uint64_t>(in.cnt_) < untilBit) {
            uint64_t w;
            std::memcpy(&w, in.ptr_, 8);
            uint64_t buf = in.buf_ | (__builtin_bswap64(w) >> in.cnt_);
            in.ptr_ += (63 - in.cnt_) >> 3;
            int cnt = in.cnt_ | 56;
            bool special = false;
            for (int k = 0; k < 4; ++k) {
                const uint64_t e = pair[buf >> (64 - kLutBits)];
                const uint32_t got = static_cast<uint32_t>(e >> 8) & 0xFF;
                if (got == 0) {
                    special = true;
                    break;
                }
                const uint64_t syms = e >> 16;  // one 8-byte store: the symbols beyond `got` are junk, overwritten by the next hit
                std::memcpy(out + n, &syms, 8);
                n += got;
                buf <<= (e & 0xFF);
                cnt -= static_cast<int>(e & 0xFF);
            }
            in.buf_ = cnt < 64 ? buf & ~((1ull << (64 - cnt)) - 1ull) : buf;  // drop the partial byte (BitReader's invariant)
            in.cnt_ = cnt;
            if (special) break;
        }
    }

    // Code at the head of `window` (kMaxCodeLen bits) -> symbol index and length; false when the bits are no code.
    bool resolve(uint64_t window, uint32_t& index, uint32_t& len) const {
        const uint32_t entry = lut[window >> (kMaxCodeLen - kLutBits)];
        if (entry) {
            index = entry >> 8, len = entry & 0xFF;
            return true;
        }
        for (int l = kLutBits + 1; l <= maxLen; ++l) {
            const uint64_t v = window >> (kMaxCodeLen - l);
            if (count[l] && v >= firstCode[l] && v - firstCode[l] < count[l]) {
                index = firstIndex[l] + static_cast<uint32_t>(v - firstCode[l]), len = static_cast<uint32_t>(l);
                return true;
            }
        }
        return false;
    }

    // One symbol, all checks. Returns true at the pseudo-EOF.
    bool step(BitReader& in, uint16_t* out, size_t& n, size_t cap) const {
        uint32_t index = 0, len = 0;
        if (!resolve(in.peek(kMaxCodeLen), index, len)) throw Bad{"bit pattern is not a Huffman code"};
        in.skip(static_cast<int>(len));
        if (in.overrun()) throw Bad{"Huffman payload without pseudo-EOF"};
        if (index == total - 1) return true;
        if (n >= cap) throw Bad{"sequence longer than the block counts allow"};
        out[n++] = symbols[index];
        return false;
    }

    // Worker of decodeParallel: walks [c.from, c.until) and stops at the first code boundary at or after c.until, at a
    // pseudo-EOF, or at the end of the data. `record` > 0: a blind start — the first `record` code starts are noted, and bits
    // that are no code are skipped one at a time (the walk is not in step yet). record == 0: the true walk (chunk 0).
    void runChunk(const BitReader& proto, Chunk& c, uint16_t* dst, size_t dstCap, size_t record) const {
        BitReader r = proto;
        r.seek(c.from);
        const uint64_t totalBits = static_cast<uint64_t>(r.end_ - r.base_) * 8;
        size_t n = 0;
        c.bnd.clear();
        c.eof = c.failed = false;
        c.lastBad = 0;
        for (;;) {
            if (n >= record) fastRun(r, dst, n, dstCap, c.until);
            const uint64_t pos = r.position();
            if (pos >= c.until || pos >= totalBits) break;
            uint32_t index = 0, len = 0;
            if (!resolve(r.peek(kMaxCodeLen), index, len)) {
                if (!record) {
                    c.failed = true;
                    break;
                }
                r.skip(1);
                c.lastBad = n + 1;  // the bad bits lie before the code that yields symbol n
                continue;
            }
            r.skip(static_cast<int>(len));
            if (r.overrun()) {
                c.failed = !record;
                break;
            }
            if (index == total - 1) {
                c.eof = true;
                c.eofEnd = r.position();
                if (n < record) c.bnd.push_back(pos);  // the pseudo-EOF's own start: the true walk may arrive exactly here
                break;
            }
            if (n >= dstCap) {
                c.failed = !record;
                break;
            }
            if (n < record) c.bnd.push_back(pos);
            dst[n++] = symbols[index];
        }
        c.count = n;
        c.endPos = r.position();
    }
};

constexpr size_t kParallelMinSymbols = 32768;  // below this a sequence is decoded serially

// readHuffmanOrGolomb (CompressedImage.cpp:392-401): `expect` symbols when Golomb coded; a Huffman payload carries its own end.
size_t readSequence(BitReader& in, HuffTable& table, size_t expect, uint16_t* out, size_t cap, SpinTeam* pool,
                    std::vector<HuffTable::Chunk>& chunks) {
    if (in.read(1) == 0) {
        table.read(in);
        if (pool && pool->size() > 1 && expect >= kParallelMinSymbols) return table.decodeParallel(in, out, cap, expect, *pool, chunks);
        return table.decode(in, out, cap);
    }
    const uint32_t M = static_cast<uint32_t>(in.read(16));
    if (M == 0) throw Bad{"Golomb parameter 0"};
    if (expect > cap) throw Bad{"sequence longer than the block counts allow"};
    const uint32_t b = std::bit_width(M), limit = (1u << (b + 1)) - M;  // BitBuffer.cpp:248-256
    for (size_t i = 0; i < expect; ++i) {
        const uint32_t q = in.readUnaryOnes();
        const uint32_t rem1 = static_cast<uint32_t>(in.read(static_cast<int>(b)));
        const uint32_t rem = rem1 < limit ? rem1 : (rem1 << 1) + static_cast<uint32_t>(in.read(1)) - limit;
        out[i] = static_cast<uint16_t>(q * M + rem);
        if (in.overrun()) throw Bad{"Golomb payload runs past the end of the stream"};
    }
    return expect;
}

// runLengthDecode (Huffman.cpp:281-307) from `in` into out[0, cap).
size_t expandRuns(const uint16_t* in, size_t n, uint16_t* out, size_t cap) {
    size_t o = 0;
    uint16_t prev = 0;
    bool pending = false, newRun = true;
    for (size_t i = 0; i < n; ++i) {
        const uint16_t v = in[i];
        if (pending) {
            if (o + v > cap) throw Bad{"run-length coded stream longer than the block counts allow"};
            std::fill(out + o, out + o + v, prev);
            o += v;
            pending = false;
            newRun = true;
        } else {
            if (o >= cap) throw Bad{"run-length coded stream longer than the block counts allow"};
            out[o++] = v;
            if (v == prev && !newRun) pending = true;
            newRun = false;
            prev = v;
        }
    }
    return o;
}

}  // namespace

struct MnParser::Impl {
    BitReader in;
    HuffTable table;
    std::vector<uint16_t> runs;
    HostPool* pool;
    SpinTeam* team = nullptr;  // set while readAll runs inside the pool
    std::vector<HuffTable::Chunk> chunks;
    Impl(const uint8_t* p, size_t n, HostPool* hp) : in(p, n), pool(hp) {}
};

MnParser::MnParser(const uint8_t* bytes, size_t n, HostPool* pool) : impl_(new Impl(bytes, n, pool)), nbytes_(n) {}
MnParser::~MnParser() { delete impl_; }

int MnParser::readHeader(std::string& err, bool headerOnly) {
    BitReader& in = impl_->in;
    if (nbytes_ < 14) return err = "stream shorter than its header", -1;
    if (in.read32() != 0x4D4E3234u) return err = "not a .mn stream (magic)", -1;  // CompressedImage.cpp:721-723
    width = in.read32();
    height = in.read32();
    K = static_cast<int>(in.read(8));
    blockSize = static_cast<int>(in.read(8));
    if (K < 1 || K > kMaxK) return err = "K outside 1..32", -1;
    if (blockSize != 8) return err = "block size is not 8", -1;
    if (width == 0 || height == 0 || width > (1u << 20) || height > (1u << 20)) return err = "image dimensions out of range", -1;
    if (nbytes_ < 14 + 6 * static_cast<size_t>(K)) return err = "stream shorter than its header", -1;
    for (int i = 0; i < 3 * K; ++i) quant[i] = static_cast<float>(in.read(16));  // :727-736
    nblocks = static_cast<size_t>((width + 7) / 8) * ((height + 7) / 8);
    // every `lengths` symbol costs at least one bit: bounds what a caller allocates for a stream that claims absurd dimensions
    if (!headerOnly && 3 * nblocks > 8 * nbytes_) return err = "fewer bits than blocks", -1;
    return 0;
}

int MnParser::readLengths(uint16_t* counts, std::string& err) {
    try {
        const size_t n = readSequence(impl_->in, impl_->table, 3 * nblocks, counts, 3 * nblocks, impl_->team, impl_->chunks);
        if (n != 3 * nblocks) throw Bad{"`lengths` does not hold one count per block and channel"};
    } catch (const Bad& b) {
        return err = b.why, -1;
    }
    // stream (layer, depth) holds one entry per block with count > depth (CompressedImage.cpp:751-758)
    size_t hist[3][kMaxK + 1] = {};
    for (size_t b = 0; b < nblocks; ++b)
        for (int ch = 0; ch < 3; ++ch) {
            const uint16_t c = counts[3 * b + ch];
            if (c > K) return err = "block count above K", -1;
            ++hist[ch][c];
        }
    size_t at = 0;
    for (int ch = 0; ch < 3; ++ch) {
        size_t deeper[kMaxK + 1] = {};
        for (int d = K - 1; d >= 0; --d) deeper[d] = (d + 1 < K ? deeper[d + 1] : 0) + hist[ch][d + 1];
        for (int d = 0; d < K; ++d)
            for (int part = 0; part < 2; ++part) {
                streamOffset[2 * (ch * K + d) + part] = at;
                at += deeper[d];
            }
    }
    streamOffset[6 * K] = at;
    if (at > 0xFFFFFFF0ull) return err = "more than 2^32 symbols", -1;
    return 0;
}

int MnParser::withTeam(const std::function<int()>& body) {
    HostPool* pool = impl_->pool;
    if (!pool || pool->size() < 2) return body();
    SpinTeam team(pool->size());
    int rc = 0;
    pool->withHelpers([&] { team.helper(); }, [&] {
        struct Release {  // on every path out of body(), an exception included: the helpers spin until finish()
            Impl* impl;
            SpinTeam* team;
            ~Release() {
                impl->team = nullptr;
                team->finish();
            }
        } release{impl_, &team};
        impl_->team = &team;
        rc = body();
    });
    return rc;
}

int MnParser::readStreams(uint16_t* arena, std::string& err) {
    BitReader& in = impl_->in;
    try {
        for (int s = 0; s < 6 * K; ++s) {
            uint16_t* dst = arena + streamOffset[s];
            const size_t want = streamOffset[s + 1] - streamOffset[s];
            size_t got;
            if (in.read(1) == 1) {  // run-length coded, CompressedImage.cpp:744-750
                const size_t codeLength = in.read32();
                if (codeLength > want + 4) throw Bad{"run-length coded stream longer than the plain one"};
                impl_->runs.resize(codeLength + 1);
                const size_t n = readSequence(in, impl_->table, codeLength, impl_->runs.data(), codeLength, impl_->team, impl_->chunks);
                got = expandRuns(impl_->runs.data(), n, dst, want);
            } else {
                got = readSequence(in, impl_->table, want, dst, want, impl_->team, impl_->chunks);
            }
            if (got != want) throw Bad{"stream size differs from the block counts"};
        }
    } catch (const Bad& b) {
        return err = b.why, -1;
    }
    for (int ch = 0; ch < 3; ++ch) {  // undo the DC delta chains, :766-781
        uint16_t* dc = arena + streamOffset[2 * ch * K + 1];
        const size_t n = streamOffset[2 * ch * K + 2] - streamOffset[2 * ch * K + 1];
        int32_t acc = 0;
        for (size_t i = 0; i < n; ++i) {
            acc = unzigzag(dc[i]) + acc;
            dc[i] = static_cast<uint16_t>(acc);
        }
    }
    bitsRead = in.position();
    return 0;
}

}  // namespace mnenc
