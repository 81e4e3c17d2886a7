// TEST INFRASTRUCTURE — NOT PART OF THE PRODUCT PATH.
//
// CPU restatement ("port") of the reference's `.mn` encoder hot path and of its decoder, written from the
// reference's behaviour, scalar and single-threaded per call. Only tests/, __graft_entry__.smoke() and bench.py's
// cpu_baseline / --impl reference legs may load this library; the product (libmnenc.so) never does.
//
// Parity status: PINNED. tests/test_oracle.py (encoder) and tests/test_oracle_decode.py (decoder) check every
// function below against the unmodified reference compiled into oracle/_ref/libmnref.so (this container and the
// GPU box), and tests/golden/ holds vectors generated from that reference build (make_golden.py,
// make_golden_decode.py) plus the known answers of the reference's own fixture Data/r0c1de5e1t_3_5.mn: its header
// (quantiser tables) and the sha256 of its decoded pixels.
//
// All file:line citations are relative to the reference repository root.
//
// Floating point: this file must be compiled with -ffp-contract=off and without -march flags so that every
// f32/f64 operation is individually rounded, as in the canonical oracle build.
#include <algorithm>
#include <bit>
#include <cmath>
#include <cstdint>
#include <cstdlib>
#include <cstring>
#include <queue>
#include <unordered_map>
#include <vector>

namespace {

// ---------------------------------------------------------------------------------------------------
// Bit writer: MSB-first bit stream; equals BitBuffer::WriteBits + Save (BitBuffer.cpp:78-110,183-202).
// ---------------------------------------------------------------------------------------------------
struct Bits {
    std::vector<uint8_t> bytes;
    uint64_t nbits = 0;
    void put(uint64_t value, int width) {
        for (int i = width - 1; i >= 0; --i) {
            if ((nbits & 7) == 0) bytes.push_back(0);
            if ((value >> i) & 1ULL) bytes.back() |= static_cast<uint8_t>(0x80u >> (nbits & 7));
            ++nbits;
        }
    }
    void append(const Bits& o) {
        for (uint64_t i = 0; i < o.nbits; ++i) put((o.bytes[i >> 3] >> (7 - (i & 7))) & 1u, 1);
    }
};

// BitBuffer.h:112-118
inline uint32_t zz(int32_t x) { return (static_cast<uint32_t>(x) << 1) ^ static_cast<uint32_t>(x >> 31); }

// BitBuffer.cpp:260-269 — note b = bit_width(M), not ceil(log2 M); limit = 2^(b+1) - M.
inline uint32_t golombLen(uint32_t v, uint32_t M) {
    uint32_t q = v / M, rem = v % M;
    uint32_t b = std::bit_width(M);
    uint32_t limit = (1u << (b + 1)) - M;
    return rem < limit ? b + q + 1 : b + q + 2;
}
// BitBuffer.cpp:228-241
inline void golombPut(Bits& out, uint32_t v, uint32_t M) {
    uint32_t q = v / M, rem = v % M;
    for (uint32_t i = 0; i < q; ++i) out.put(1, 1);
    out.put(0, 1);
    uint32_t b = std::bit_width(M);
    uint32_t limit = (1u << (b + 1)) - M;
    if (rem < limit) out.put(rem, b);
    else out.put(static_cast<uint64_t>(rem) + limit, b + 1);
}

// BitBuffer.cpp:345-354 — the *upper bound* used for the raw-vs-Elias-Fano decision.
inline uint32_t eliasFanoBound(size_t n, uint16_t maxSymbol) {
    if (n == 0) return 0;
    uint32_t m = std::bit_width(maxSymbol);
    uint32_t nb = std::bit_width(static_cast<uint32_t>(n));
    uint32_t lb = (m >= nb) ? m - nb : 0;
    return static_cast<uint32_t>(n) * (1 + lb) + (1u << (m - lb)) - 1;
}
// BitBuffer.cpp:292-317
inline void eliasFanoPut(Bits& out, const std::vector<uint16_t>& seq, uint16_t maxSymbol) {
    if (seq.empty()) return;
    uint32_t m = std::bit_width(maxSymbol);
    uint32_t nb = std::bit_width(static_cast<uint32_t>(seq.size()));
    uint32_t lb = (m >= nb) ? m - nb : 0;
    uint32_t bucket = 0;
    for (uint16_t s : seq) {
        uint32_t want = static_cast<uint32_t>(s) >> lb;
        while (bucket != want) {
            out.put(0, 1);
            ++bucket;
        }
        out.put(1, 1);
    }
    for (uint16_t s : seq) out.put(s, static_cast<int>(lb));
}

// Huffman.cpp:246-279 — stated on maximal runs: run of v with length L is cut into chunks of 32769
// ([v,v,32767] each); the remainder r emits [v] (r==1) or [v,v,r-2] (r>=2).
std::vector<uint16_t> rle(const uint16_t* d, size_t n) {
    std::vector<uint16_t> out;
    size_t i = 0;
    while (i < n) {
        size_t j = i;
        while (j < n && d[j] == d[i]) ++j;
        size_t L = j - i;
        while (L >= 32769) {
            out.push_back(d[i]);
            out.push_back(d[i]);
            out.push_back(32767);
            L -= 32769;
        }
        if (L == 1) out.push_back(d[i]);
        else if (L >= 2) {
            out.push_back(d[i]);
            out.push_back(d[i]);
            out.push_back(static_cast<uint16_t>(L - 2));
        }
        i = j;
    }
    return out;
}

// ---------------------------------------------------------------------------------------------------
// Canonical Huffman with pseudo-EOF (Huffman.cpp:46-163). The code lengths depend on libstdc++'s
// unordered_map iteration order and priority_queue tie handling, so the same containers
// are driven through the same operation sequence: insert symbols in first-occurrence order, then EOF;
// push leaves in map iteration order; pop two / push merged until one node is left.
// ---------------------------------------------------------------------------------------------------
struct Leaf {
    uint32_t symbol;
    uint64_t freq;
    uint32_t len;
    uint32_t code;
    int parent;
};
struct Node {
    uint64_t freq;
    int id;
    uint32_t depth;
};
struct NodeGreater {
    bool operator()(const Node& a, const Node& b) const { return a.freq > b.freq; }  // Huffman.cpp:22-26
};
const uint32_t EOF_SYM = 0xFFFFFFFFu;  // Huffman.cpp:38

void huffman(Bits& out, const uint16_t* data, size_t n) {
    std::unordered_map<uint32_t, int> index;  // symbol -> slot in `nodes`
    std::vector<Leaf> leaves;
    uint32_t symWidth = 1;
    for (size_t i = 0; i < n; ++i) {
        uint32_t s = data[i];
        symWidth = std::max<uint32_t>(symWidth, std::bit_width(s));
        auto it = index.find(s);
        if (it == index.end()) {
            index[s] = static_cast<int>(leaves.size());
            leaves.push_back(Leaf{s, 0, 0, 0, -1});
            it = index.find(s);
        }
        ++leaves[it->second].freq;
    }
    index[EOF_SYM] = static_cast<int>(leaves.size());
    leaves.push_back(Leaf{EOF_SYM, 0, n == 0 ? 1u : 0u, 0, -1});

    // parents[] holds the merge tree; ids < leaves.size() are leaves.
    std::vector<int> parents(leaves.size(), -1);
    std::priority_queue<Node, std::vector<Node>, NodeGreater> heap;
    for (const auto& kv : index) heap.push(Node{leaves[kv.second].freq, kv.second, 0});
    while (heap.size() > 1) {
        Node a = heap.top();
        heap.pop();
        Node b = heap.top();
        heap.pop();
        int id = static_cast<int>(parents.size());
        parents.push_back(-1);
        parents[a.id] = id;
        parents[b.id] = id;
        heap.push(Node{a.freq + b.freq, id, std::max(a.depth, b.depth) + 1});
    }
    for (size_t i = 0; i < leaves.size(); ++i)
        for (int p = parents[i]; p != -1; p = parents[p]) ++leaves[i].len;  // one ++ per merge above the leaf (:94-96)
    uint32_t maxLen = std::max<uint32_t>(heap.top().depth, 1);  // :110

    std::vector<Leaf> sorted(leaves);
    std::sort(sorted.begin(), sorted.end(), [](const Leaf& a, const Leaf& b) {  // :28-36
        return a.len == b.len ? a.symbol < b.symbol : a.len < b.len;
    });
    out.put(maxLen, 8);
    std::vector<uint32_t> counts;  // per length 1..; :112-135
    uint32_t prevLen = 0, code = 0, count = 0;
    for (Leaf& e : sorted) {
        if (e.len != prevLen) {
            if (prevLen != 0) counts.push_back(count);
            for (uint32_t l = prevLen + 1; l < e.len; ++l) counts.push_back(0);
            count = 0;
            code <<= (e.len - prevLen);
        }
        e.code = code++;
        ++count;
        prevLen = e.len;
        leaves[index[e.symbol]].code = e.code;
    }
    counts.push_back(count);
    for (uint32_t c : counts) out.put(c & 0xFFFFu, 16);
    out.put(symWidth, 8);
    uint16_t mask = static_cast<uint16_t>((1u << symWidth) - 1u);
    size_t pos = 0;
    for (uint32_t c : counts) {  // :139-156
        uint16_t rangeLen = static_cast<uint16_t>(c);
        std::vector<uint16_t> range;
        for (size_t k = 0; k < rangeLen; ++k) range.push_back(static_cast<uint16_t>(sorted[pos + k].symbol) & mask);
        if (eliasFanoBound(rangeLen, mask) < static_cast<uint32_t>(rangeLen) * symWidth) eliasFanoPut(out, range, mask);
        else
            for (uint16_t s : range) out.put(s, static_cast<int>(symWidth));
        pos += rangeLen;
    }
    for (size_t i = 0; i < n; ++i) {  // :157-161
        const Leaf& e = leaves[index[data[i]]];
        out.put(e.code, static_cast<int>(e.len));
    }
    const Leaf& e = leaves[index[EOF_SYM]];
    out.put(e.code, static_cast<int>(e.len));  // :162
}

// CompressedImage.cpp:359-390
void huffmanOrGolomb(Bits& out, const uint16_t* data, size_t n) {
    Bits tmp;
    huffman(tmp, data, n);
    uint64_t best = tmp.nbits;
    int bestM = -1;
    for (int m = 1; m < 2048;) {
        uint64_t est = 16;
        for (size_t i = 0; i < n; ++i) est += golombLen(data[i], static_cast<uint32_t>(m));
        if (est < best) {
            best = est;
            bestM = m;
        }
        m = (m & 1) ? m + 1 : (m << 1) - 1;
    }
    if (bestM < 0) {
        out.put(0, 1);
        out.append(tmp);
    } else {
        out.put(1, 1);
        out.put(static_cast<uint32_t>(bestM), 16);
        for (size_t i = 0; i < n; ++i) golombPut(out, data[i], static_cast<uint32_t>(bestM));
    }
}

uint8_t* giveBits(const Bits& b, uint64_t* nbits, uint64_t* nbytes) {
    *nbits = b.nbits;
    *nbytes = b.bytes.size();
    uint8_t* p = static_cast<uint8_t*>(std::malloc(b.bytes.size() + 1));
    std::memcpy(p, b.bytes.data(), b.bytes.size());
    return p;
}

}  // namespace

// ---------------------------------------------------------------------------------------------------
// Matching pursuit (MatchingPursuit.cpp:27-37,76-107; CompressedImage.cpp:252-289)
// ---------------------------------------------------------------------------------------------------
namespace {

struct Dict {
    int K, bs, nBase, n;                 // n = bs*bs
    const float* base;                   // [nBase][n]
    std::vector<int> rows;               // detail rows per base atom
    std::vector<int64_t> rowOffset;      // first detail row of each group in the concatenated table
    const float* detail[3];              // per channel [sum(rows)][n]
    float quant[3][64];
};

// Eigen col-major GEMV for < 128 columns == per row: acc = fl(acc + fl(d*r)), j ascending
// (eigen/Eigen/src/Core/products/GeneralMatrixVector.h:108-215 without EIGEN_VECTORIZE_FMA).
inline float dotSeq(const float* d, const float* r, int n) {
    float acc = 0.0f;
    for (int j = 0; j < n; ++j) {
        float p = d[j] * r[j];
        acc = acc + p;
    }
    return acc;
}

// One block-channel. pairs[2*i] = deltaId, pairs[2*i+1] = intCoeff for i < K (entries >= count left untouched
// except entry `count`, which the reference also writes before returning: MatchingPursuit.cpp:91-101).
int mpBlock(const Dict& D, int ch, const float* in, uint16_t* pairs) {
    const int n = D.n;
    float r[256];
    std::memcpy(r, in, sizeof(float) * n);
    int groups[64];
    int ngroups = 0;
    int prev = 0;
    for (int reps = 0; reps < D.K; ++reps) {
        int best = 0;
        float bestP = 0.0f;
        const float* bestRow = nullptr;
        int idx = 0;
        for (int i = 0; i < D.nBase; ++i, ++idx) {
            float p = dotSeq(D.base + static_cast<size_t>(i) * n, r, n);
            if (idx == 0 || std::fabs(p) > std::fabs(bestP)) {  // strict > : first maximum wins (Visitor.h:227)
                bestP = p;
                best = idx;
                bestRow = D.base + static_cast<size_t>(i) * n;
            }
        }
        for (int g = 0; g < ngroups; ++g) {
            const float* grp = D.detail[ch] + D.rowOffset[groups[g]] * n;
            for (int i = 0; i < D.rows[groups[g]]; ++i, ++idx) {
                float p = dotSeq(grp + static_cast<size_t>(i) * n, r, n);
                if (std::fabs(p) > std::fabs(bestP)) {
                    bestP = p;
                    best = idx;
                    bestRow = grp + static_cast<size_t>(i) * n;
                }
            }
        }
        pairs[2 * reps] = reps ? static_cast<uint16_t>(zz(best - prev)) : static_cast<uint16_t>(best);
        prev = best;
        float ratio = bestP / D.quant[ch][reps];
        int q = static_cast<int>(std::round(ratio));  // half away from zero
        pairs[2 * reps + 1] = static_cast<uint16_t>(zz(q));
        if (q == 0) return reps;
        float c = D.quant[ch][reps] * static_cast<float>(q);
        for (int j = 0; j < n; ++j) {
            float t = c * bestRow[j];
            r[j] = r[j] - t;
        }
        if (best < D.nBase) groups[ngroups++] = best;  // CompressedImage.cpp:266-269,281-286
    }
    return D.K;
}

// CompressedImage.cpp:595-611 + ImageHelper/src/misc.cpp:12-21 (double math, then cast to float).
void gatherBlock(const uint8_t* rgb, size_t w, size_t h, size_t stride, size_t x0, size_t y0, int bs, float* Y, float* U, float* V) {
    const double uScale = 0.436 / (1.0 - 0.114);
    const double vScale = 0.615 / (1.0 - 0.299);
    for (int dy = 0; dy < bs; ++dy)
        for (int dx = 0; dx < bs; ++dx) {
            size_t u = x0 + dx, v = y0 + dy;
            float fy = 0.f, fu = 0.f, fv = 0.f;
            if (u < w && v < h) {
                const uint8_t* p = rgb + v * stride + 3 * u;
                double R = p[0], G = p[1], B = p[2];
                double yy = 0.299 * R + 0.587 * G + 0.114 * B;
                fy = static_cast<float>(yy);
                fu = static_cast<float>(uScale * (B - yy));
                fv = static_cast<float>(vScale * (R - yy));
            }
            Y[dx + bs * dy] = fy;
            U[dx + bs * dy] = fu;
            V[dx + bs * dy] = fv;
        }
}

}  // namespace

extern "C" {

struct mno_dict {
    Dict d;
};

void* mno_dict_create(int K, int bs, int nBase, const float* base, const int* detailRows, const float* detY, const float* detU,
                      const float* detV, const float* quant3K) {
    if (K > 64 || bs * bs > 256) return nullptr;
    mno_dict* h = new mno_dict;
    Dict& D = h->d;
    D.K = K;
    D.bs = bs;
    D.nBase = nBase;
    D.n = bs * bs;
    D.base = base;
    D.rows.assign(detailRows, detailRows + nBase);
    D.rowOffset.resize(nBase + 1);
    int64_t o = 0;
    for (int i = 0; i < nBase; ++i) {
        D.rowOffset[i] = o;
        o += detailRows[i];
    }
    D.rowOffset[nBase] = o;
    D.detail[0] = detY;
    D.detail[1] = detU;
    D.detail[2] = detV;
    for (int c = 0; c < 3; ++c)
        for (int i = 0; i < K; ++i) D.quant[c][i] = quant3K[c * K + i];
    return h;
}
void mno_dict_set_quant(void* h, const float* quant3K) {
    Dict& D = static_cast<mno_dict*>(h)->d;
    for (int c = 0; c < 3; ++c)
        for (int i = 0; i < D.K; ++i) D.quant[c][i] = quant3K[c * D.K + i];
}
void mno_dict_destroy(void* h) { delete static_cast<mno_dict*>(h); }
void mno_free(void* p) { std::free(p); }

void mno_gather_block(const uint8_t* rgb, size_t w, size_t h, size_t stride, size_t x0, size_t y0, int bs, float* yuv3n) {
    gatherBlock(rgb, w, h, stride, x0, y0, bs, yuv3n, yuv3n + bs * bs, yuv3n + 2 * bs * bs);
}
int mno_mp_block(void* h, int ch, const float* in, uint16_t* pairs) { return mpBlock(static_cast<mno_dict*>(h)->d, ch, in, pairs); }

// Block columns [bx0,bx1) of the encodeImageFast walk. Same output layout as mnref_symbols:
//   counts[3*blk+ch], pairs[((3*blk+ch)*K+i)*2+{0,1}], entries >= count zeroed.
void mno_symbols(void* h, const uint8_t* rgb, size_t w, size_t hgt, size_t stride, size_t bx0, size_t bx1, uint16_t* counts,
                 uint16_t* pairs) {
    const Dict& D = static_cast<mno_dict*>(h)->d;
    const int K = D.K, bs = D.bs;
    float blk[3][256];
    uint16_t tmp[128];
    size_t b = 0;
    for (size_t bx = bx0; bx < bx1; ++bx)
        for (size_t y = 0; y < hgt; y += bs, ++b) {
            gatherBlock(rgb, w, hgt, stride, bx * bs, y, bs, blk[0], blk[1], blk[2]);
            for (int ch = 0; ch < 3; ++ch) {
                int count = mpBlock(D, ch, blk[ch], tmp);
                counts[3 * b + ch] = static_cast<uint16_t>(count);
                uint16_t* p = pairs + (3 * b + ch) * K * 2;
                for (int i = 0; i < K; ++i) {
                    p[2 * i] = i < count ? tmp[2 * i] : 0;
                    p[2 * i + 1] = i < count ? tmp[2 * i + 1] : 0;
                }
            }
        }
}

// Entropy primitives.
uint16_t* mno_rle(const uint16_t* d, size_t n, size_t* nout) {
    std::vector<uint16_t> r = rle(d, n);
    *nout = r.size();
    uint16_t* p = static_cast<uint16_t*>(std::malloc(r.size() * 2 + 2));
    std::memcpy(p, r.data(), r.size() * 2);
    return p;
}
uint8_t* mno_huffman(const uint16_t* d, size_t n, uint64_t* nbits, uint64_t* nbytes) {
    Bits b;
    huffman(b, d, n);
    return giveBits(b, nbits, nbytes);
}
uint8_t* mno_huffman_or_golomb(const uint16_t* d, size_t n, uint64_t* nbits, uint64_t* nbytes) {
    Bits b;
    huffmanOrGolomb(b, d, n);
    return giveBits(b, nbits, nbytes);
}
uint32_t mno_golomb_length(uint32_t v, uint32_t m) { return golombLen(v, m); }
uint8_t* mno_golomb(const uint32_t* v, size_t n, uint32_t m, uint64_t* nbits, uint64_t* nbytes) {
    Bits b;
    for (size_t i = 0; i < n; ++i) golombPut(b, v[i], m);
    return giveBits(b, nbits, nbytes);
}
uint32_t mno_elias_fano_bound(size_t n, uint16_t maxSymbol) { return eliasFanoBound(n, maxSymbol); }
uint8_t* mno_elias_fano(const uint16_t* s, size_t n, uint16_t maxSymbol, uint64_t* nbits, uint64_t* nbytes) {
    Bits b;
    eliasFanoPut(b, std::vector<uint16_t>(s, s + n), maxSymbol);
    return giveBits(b, nbits, nbytes);
}
uint32_t mno_zigzag(int32_t x) { return zz(x); }

// writeCompressed (CompressedImage.cpp:462-519) from raw symbols in the mno_symbols layout for the whole image.
uint8_t* mno_write_file(int K, int bs, size_t w, size_t h, const float* quant3K, const uint16_t* counts, const uint16_t* pairs,
                        uint64_t* nbytes) {
    const size_t nblocks = ((w + bs - 1) / bs) * ((h + bs - 1) / bs);
    std::vector<std::vector<uint16_t>> codes(6 * K);
    for (size_t b = 0; b < nblocks; ++b)  // stream scatter, CompressedImage.cpp:612-629
        for (int ch = 0; ch < 3; ++ch) {
            const uint16_t* p = pairs + (3 * b + ch) * K * 2;
            for (int i = 0; i < counts[3 * b + ch]; ++i) {
                codes[2 * K * ch + 2 * i].push_back(p[2 * i]);
                codes[2 * K * ch + 2 * i + 1].push_back(p[2 * i + 1]);
            }
        }
    Bits out;
    out.put(0x4D4E3234u, 32);
    out.put(static_cast<uint32_t>(w), 32);
    out.put(static_cast<uint32_t>(h), 32);
    out.put(static_cast<uint8_t>(K), 8);
    out.put(static_cast<uint8_t>(bs), 8);
    for (int i = 0; i < 3 * K; ++i) out.put(static_cast<uint16_t>(quant3K[i]), 16);
    for (int ch = 0; ch < 3; ++ch) {  // :487-505
        int32_t prev = 0;
        for (uint16_t& c : codes[2 * K * ch + 1]) {
            int32_t diff = static_cast<int32_t>(c) - prev;
            prev = c;
            c = static_cast<uint16_t>(zz(diff));
        }
    }
    huffmanOrGolomb(out, counts, 3 * nblocks);  // :506
    for (int i = 0; i < 6 * K; ++i) {           // :507-517
        std::vector<uint16_t> r = rle(codes[i].data(), codes[i].size());
        if (r.size() + 4 < codes[i].size()) {
            out.put(1, 1);
            out.put(static_cast<uint32_t>(r.size()), 32);
            huffmanOrGolomb(out, r.data(), r.size());
        } else {
            out.put(0, 1);
            huffmanOrGolomb(out, codes[i].data(), codes[i].size());
        }
    }
    uint64_t nbits;
    return giveBits(out, &nbits, nbytes);
}

// encodeImageFast (CompressedImage.cpp:578-633), whole path.
uint8_t* mno_encode(void* h, const uint8_t* rgb, size_t w, size_t hgt, size_t stride, uint64_t* nbytes) {
    const Dict& D = static_cast<mno_dict*>(h)->d;
    const size_t nbx = (w + D.bs - 1) / D.bs, nby = (hgt + D.bs - 1) / D.bs;
    std::vector<uint16_t> counts(3 * nbx * nby), pairs(3 * nbx * nby * D.K * 2);
    mno_symbols(h, rgb, w, hgt, stride, 0, nbx, counts.data(), pairs.data());
    float q[192];
    for (int c = 0; c < 3; ++c)
        for (int i = 0; i < D.K; ++i) q[c * D.K + i] = D.quant[c][i];
    return mno_write_file(D.K, D.bs, w, hgt, q, counts.data(), pairs.data(), nbytes);
}

}  // extern "C"

// ---------------------------------------------------------------------------------------------------
// Decoder: restates readCompressedFast (CompressedImage.cpp:709-782), decodeImageFast (:837-889),
// FromCoeffsDynamicFast (MatchingPursuit.cpp:130-147), huffmanDecode (Huffman.cpp:173-244), runLengthDecode (:281-307),
// readGolombCode / readEliasFanoSequenceCode (BitBuffer.cpp:243-258, 319-343), RGBFromYUV (ImageHelper/src/misc.cpp:28-36).
// Pinned against the reference's own reader/decoder in tests/test_oracle.py (symbols and pixels, bit-exact).
// ---------------------------------------------------------------------------------------------------
namespace {

struct DecodeError {};

// BitBuffer.h:112-114
inline int32_t unzz(uint32_t x) { return static_cast<int32_t>((x >> 1) ^ -(static_cast<int64_t>(x) & 1)); }

// BitBuffer::Load + ReadBits (BitBuffer.cpp:112-156): MSB-first; a read past the end is clamped to the bits that remain.
struct BitsIn {
    const uint8_t* p;
    uint64_t n, pos = 0;
    uint64_t remaining() const { return n - pos; }
    uint64_t read(int width) {
        if (static_cast<uint64_t>(width) > remaining()) width = static_cast<int>(remaining());
        uint64_t v = 0;
        for (int i = 0; i < width; ++i, ++pos) v = (v << 1) | ((p[pos >> 3] >> (7 - (pos & 7))) & 1u);
        return v;
    }
};

// BitBuffer.cpp:243-258
uint32_t golombGet(BitsIn& in, uint32_t M) {
    uint32_t q = 0;
    while (in.read(1) != 0) {
        ++q;
        if (in.remaining() == 0) throw DecodeError{};  // (the reference would spin on a truncated stream)
    }
    const uint32_t b = std::bit_width(M), limit = (1u << (b + 1)) - M;
    const uint32_t rem1 = static_cast<uint32_t>(in.read(b));
    const uint32_t rem = rem1 < limit ? rem1 : (rem1 << 1) + static_cast<uint32_t>(in.read(1)) - limit;
    return q * M + rem;
}

// BitBuffer.cpp:319-343
std::vector<uint16_t> eliasFanoGet(BitsIn& in, size_t n, uint16_t maxSymbol) {
    std::vector<uint16_t> r(n);
    if (n == 0) return r;
    const uint32_t m = std::bit_width(maxSymbol), nb = std::bit_width(static_cast<uint32_t>(n)), lb = m >= nb ? m - nb : 0;
    uint16_t bucket = 0;
    for (uint16_t& s : r) {
        if (in.remaining() == 0) throw DecodeError{};
        while (in.read(1) == 0) {
            ++bucket;
            if (in.remaining() == 0) throw DecodeError{};
        }
        s = static_cast<uint16_t>(bucket << lb);
    }
    for (uint16_t& s : r) s |= static_cast<uint16_t>(in.read(lb));
    return r;
}

// Huffman.cpp:173-244
void huffmanGet(BitsIn& in, std::vector<uint16_t>& out) {
    const uint8_t maxLength = static_cast<uint8_t>(in.read(8));
    std::vector<uint16_t> lengthCounts;
    uint16_t total = 0;
    for (int len = 1; len <= maxLength; ++len) {
        const uint16_t c = static_cast<uint16_t>(in.read(16));
        total = static_cast<uint16_t>(total + c);
        lengthCounts.push_back(c);
    }
    const uint8_t width = static_cast<uint8_t>(in.read(8));
    const uint16_t mask = static_cast<uint16_t>((1u << width) - 1u);
    std::vector<uint16_t> symbols;
    for (uint16_t rangeLength : lengthCounts) {
        if (eliasFanoBound(rangeLength, mask) < static_cast<uint32_t>(rangeLength) * width) {
            std::vector<uint16_t> part = eliasFanoGet(in, rangeLength, mask);
            symbols.insert(symbols.end(), part.begin(), part.end());
        } else {
            for (int i = 0; i < rangeLength; ++i) symbols.push_back(static_cast<uint16_t>(in.read(width)));
        }
    }
    struct Entry {
        uint32_t symbol;
        uint8_t len;
    };
    std::unordered_map<uint32_t, Entry> codeMap;
    uint8_t length = 1, prevLength = 0;
    uint32_t code = 0;
    for (uint16_t k = 0; k < total; ++k) {
        while (lengthCounts.at(static_cast<size_t>(length) - 1) == 0) ++length;
        --lengthCounts[static_cast<size_t>(length) - 1];
        uint32_t symbol = symbols.at(k);
        if (length != prevLength) code <<= (length - prevLength);
        if (k == total - 1) {
            if (symbol != mask) throw DecodeError{};
            symbol = 0xFFFFFFFFu;
        }
        codeMap[code] = Entry{symbol, length};
        ++code;
        prevLength = length;
    }
    uint32_t decode = 0;
    uint8_t bitcount = 0;
    while (in.remaining() > 0) {
        decode = (decode << 1) | static_cast<uint32_t>(in.read(1));
        ++bitcount;
        auto it = codeMap.find(decode);
        if (it != codeMap.end() && it->second.len == bitcount) {
            if (it->second.symbol == 0xFFFFFFFFu) return;
            out.push_back(static_cast<uint16_t>(it->second.symbol));
            decode = 0;
            bitcount = 0;
        }
    }
    throw DecodeError{};
}

// CompressedImage.cpp:392-401
void huffmanOrGolombGet(BitsIn& in, size_t length, std::vector<uint16_t>& out) {
    if (in.read(1) == 0) {
        huffmanGet(in, out);
    } else {
        const uint32_t m = static_cast<uint32_t>(in.read(16));
        if (m == 0) throw DecodeError{};  // (the reference divides by zero in the encoder only; m = 0 is never written)
        for (size_t i = 0; i < length; ++i) out.push_back(static_cast<uint16_t>(golombGet(in, m)));
    }
}

// Huffman.cpp:281-307
std::vector<uint16_t> unrle(const std::vector<uint16_t>& d) {
    std::vector<uint16_t> r;
    uint16_t prev = 0;
    bool pending = false, newRun = true;
    for (uint16_t v : d) {
        if (pending) {
            r.insert(r.end(), v, prev);
            pending = false;
            newRun = true;
        } else {
            r.push_back(v);
            if (v == prev && !newRun) pending = true;
            newRun = false;
            prev = v;
        }
    }
    return r;
}

struct Parsed {
    size_t K = 0, bs = 0, w = 0, h = 0;
    float quant[3][64] = {};
    std::vector<uint16_t> lengths;
    std::vector<std::vector<uint16_t>> codes;
};

// CompressedImage.cpp:709-782
void parseFile(const uint8_t* bytes, size_t n, Parsed& P) {
    BitsIn in{bytes, 8ull * n};
    if (in.read(32) != 0x4D4E3234u) throw DecodeError{};
    P.w = in.read(32);
    P.h = in.read(32);
    P.K = in.read(8);
    P.bs = in.read(8);
    if (P.K == 0 || P.K > 64 || P.bs == 0) throw DecodeError{};
    for (int c = 0; c < 3; ++c)
        for (size_t i = 0; i < P.K; ++i) P.quant[c][i] = static_cast<float>(in.read(16));
    const size_t K = P.K, patch = 3 * ((P.w + P.bs - 1) / P.bs) * ((P.h + P.bs - 1) / P.bs);
    huffmanOrGolombGet(in, patch, P.lengths);
    P.codes.assign(6 * K, {});
    for (size_t i = 0; i < 6 * K; ++i) {
        if (in.read(1) == 1) {
            const size_t codeLength = in.read(32);
            std::vector<uint16_t> comp;
            huffmanOrGolombGet(in, codeLength, comp);
            P.codes[i] = unrle(comp);
        } else {
            const size_t layer = (i / 2) / K, depth = (i / 2) % K;
            size_t layerLength = 0;
            for (size_t o = 0; o < P.lengths.size() / 3; ++o) layerLength += P.lengths[3 * o + layer] > depth;
            huffmanOrGolombGet(in, layerLength, P.codes[i]);
        }
    }
    for (int ch = 0; ch < 3; ++ch) {  // undo the DC delta chains, :766-781
        int32_t acc = 0;
        for (uint16_t& c : P.codes[2 * K * ch + 1]) {
            acc = unzz(c) + acc;
            c = static_cast<uint16_t>(acc);
        }
    }
}

}  // namespace

extern "C" {

// Same output layout as mnref_parse (oracle/ref_driver.cpp). Returns 0, or -1 on an invalid stream.
int mno_parse(const uint8_t* bytes, size_t n, size_t* K, size_t* bs, size_t* w, size_t* h, uint16_t** lengths, size_t* nLengths,
              size_t* streamSizes, uint16_t** streams) {
    Parsed P;
    try {
        parseFile(bytes, n, P);
    } catch (...) {
        return -1;
    }
    *K = P.K, *bs = P.bs, *w = P.w, *h = P.h;
    *nLengths = P.lengths.size();
    *lengths = static_cast<uint16_t*>(std::malloc(P.lengths.size() * 2 + 2));
    std::memcpy(*lengths, P.lengths.data(), P.lengths.size() * 2);
    size_t total = 0;
    for (size_t i = 0; i < P.codes.size(); ++i) total += (streamSizes[i] = P.codes[i].size());
    *streams = static_cast<uint16_t*>(std::malloc(total * 2 + 2));
    size_t o = 0;
    for (const auto& s : P.codes) {
        std::memcpy(*streams + o, s.data(), s.size() * 2);
        o += s.size();
    }
    return 0;
}

// decodeImageFast (CompressedImage.cpp:837-889) with the dictionary of `h` (block size and atoms; the quantisation tables
// come from the file, as in readCompressedFast :727-736). rgbOut: malloc'ed packed RGB8, row-major.
int mno_decode(void* h, const uint8_t* bytes, size_t n, uint8_t** rgbOut, size_t* wOut, size_t* hOut) {
    const Dict& D = static_cast<mno_dict*>(h)->d;
    Parsed P;
    try {
        parseFile(bytes, n, P);
    } catch (...) {
        return -1;
    }
    if (static_cast<int>(P.bs) != D.bs) return -2;
    const size_t K = P.K, bs = P.bs, N = bs * bs, w = P.w, hgt = P.h;
    uint8_t* out = static_cast<uint8_t*>(std::calloc(w * hgt * 3 + 1, 1));
    std::vector<size_t> offsets(3 * K, 0);
    size_t lengthOffset = 0;
    std::vector<float> dec(3 * N);
    try {
        for (size_t x = 0; x < w; x += bs)
            for (size_t y = 0; y < hgt; y += bs) {
                for (int ch = 0; ch < 3; ++ch) {
                    const int count = P.lengths.at(lengthOffset++);
                    float* r = dec.data() + ch * N;
                    std::fill(r, r + N, 0.0f);
                    // FromCoeffsDynamicFast (MatchingPursuit.cpp:130-147) on the dictionary dynamicBasisFast builds from ALL `count`
                    // picks (CompressedImage.cpp:252-289): base atoms, then the detail group of every pick < nBase, in pick order.
                    int choice = 0;
                    std::vector<int> groups;
                    std::vector<int> abs(count);
                    for (int i = 0; i < count; ++i) {
                        const uint16_t deltaId = P.codes[2 * K * ch + 2 * i].at(offsets[K * ch + i]);
                        choice = i ? choice + unzz(deltaId) : static_cast<int>(deltaId);
                        abs[i] = choice;
                        if (choice < D.nBase) groups.push_back(choice);  // a negative index makes the reference read out of range
                    }
                    for (int i = 0; i < count; ++i) {
                        const uint16_t intCoeff = P.codes[2 * K * ch + 2 * i + 1].at(offsets[K * ch + i]++);
                        const float coeff = P.quant[ch][i] * static_cast<float>(unzz(intCoeff));
                        const float* row = nullptr;
                        int idx = abs[i];
                        if (idx < 0) throw DecodeError{};
                        if (idx < D.nBase) {
                            row = D.base + static_cast<size_t>(idx) * N;
                        } else {
                            idx -= D.nBase;
                            for (int g : groups) {
                                if (idx < D.rows[g]) {
                                    row = D.detail[ch] + (D.rowOffset[g] + idx) * N;
                                    break;
                                }
                                idx -= D.rows[g];
                            }
                            if (!row) throw DecodeError{};
                        }
                        for (size_t j = 0; j < N; ++j) r[j] = r[j] + coeff * row[j];  // results += coeff * dictionary.row(choice)
                    }
                }
                for (size_t dx = 0; dx < bs; ++dx)
                    for (size_t dy = 0; dy < bs; ++dy) {
                        const size_t u = x + dx, v = y + dy;
                        if (u >= w || v >= hgt) continue;
                        const double yy = dec[dx + bs * dy], uu = dec[N + dx + bs * dy], vv = dec[2 * N + dx + bs * dy];
                        uint8_t* px = out + (v * w + u) * 3;  // misc.cpp:28-36
                        px[0] = static_cast<uint8_t>(std::clamp(std::round(yy + 1.13983 * vv), 0.0, 255.0));
                        px[1] = static_cast<uint8_t>(std::clamp(std::round(yy - 0.39466 * uu - 0.58060 * vv), 0.0, 255.0));
                        px[2] = static_cast<uint8_t>(std::clamp(std::round(yy + 2.03211 * uu), 0.0, 255.0));
                    }
            }
    } catch (...) {
        std::free(out);
        return -1;
    }
    *rgbOut = out, *wOut = w, *hOut = hgt;
    return 0;
}

}  // extern "C"
