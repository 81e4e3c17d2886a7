// Host side of the entropy back end: everything that the reference derives per coded sequence and that depends
// only on the symbol histogram and the first-occurrence order:
//   huffman::huffmanEncode table construction + header   CompressionLib/src/Huffman.cpp:65-156
//   compressed::writeHuffmanOrGolomb choice              CompressionLib/src/CompressedImage.cpp:359-390
//   bitbuffer::golombCodeLength                          CompressionLib/src/BitBuffer.cpp:260-269
//   bitbuffer::eliasFanoSequenceCodeLength / write       CompressionLib/src/BitBuffer.cpp:292-317,345-354
//   compressed::createQuantizationTables (float)         CompressionLib/src/CompressedImage.cpp:168-210
// The per-symbol work (frequency counting, code emission) runs on the GPU; this file never sees the symbols.
//
// Huffman tie-breaking in the reference is a property of libstdc++'s std::unordered_map iteration order and
// std::priority_queue sift order. The same containers are therefore driven with the same
// operation sequence: keys inserted in first-occurrence order then the pseudo-EOF, leaves pushed in map
// iteration order, pop-two/push-merged. Only (frequency) takes part in heap comparisons, so light nodes with
// parent links replace the reference's child vectors (O(n log n) instead of O(n * depth) vector churn); the heap itself is
// libstdc++'s algorithm restated on 8-byte nodes (FreqHeap below).
#include <algorithm>
#include <bit>
#include <cmath>
#include <memory_resource>
#include <numeric>
#include <unordered_map>

#include "mnenc_internal.h"
#include "quant_data.h"

namespace mnenc {

void BitString::put(uint64_t value, int width) {
    // MSB-first. The open byte (if any) takes the leading bits, whole bytes follow, a trailing partial byte is opened.
    if (width <= 0) return;
    if (width < 64) value &= (1ULL << width) - 1ULL;
    const int used = static_cast<int>(nbits & 7);
    nbits += static_cast<uint64_t>(width);
    if (used) {
        const int room = 8 - used;
        if (width <= room) {
            bytes.back() |= static_cast<uint8_t>(value << (room - width));
            return;
        }
        bytes.back() |= static_cast<uint8_t>(value >> (width - room));
        width -= room;
    }
    while (width >= 8) {
        width -= 8;
        bytes.push_back(static_cast<uint8_t>(value >> width));
    }
    if (width) bytes.push_back(static_cast<uint8_t>(value << (8 - width)));
}

void BitString::append(const BitString& other) {
    if (other.nbits == 0) return;
    const int used = static_cast<int>(nbits & 7);
    const size_t nOther = static_cast<size_t>((other.nbits + 7) / 8);  // unused bits of its last byte are zero
    if (used == 0) {
        bytes.insert(bytes.end(), other.bytes.begin(), other.bytes.begin() + static_cast<std::ptrdiff_t>(nOther));
        nbits += other.nbits;
        return;
    }
    const uint64_t total = nbits + other.nbits;
    bytes.reserve(static_cast<size_t>((total + 7) / 8));
    for (size_t i = 0; i < nOther; ++i) {  // every source byte straddles two destination bytes
        const uint8_t v = other.bytes[i];
        bytes.back() |= static_cast<uint8_t>(v >> used);
        bytes.push_back(static_cast<uint8_t>(v << (8 - used)));
    }
    nbits = total;
    bytes.resize(static_cast<size_t>((total + 7) / 8));  // the last split may have opened a byte that holds no bit
}

namespace {

constexpr uint32_t kEof = 0xFFFFFFFFu;  // Huffman.cpp:38

struct Sym {
    uint32_t symbol;
    uint32_t len;
    uint32_t code;
    int origin;  // index in the caller's stats vector, -1 for EOF
};

// order[] = 0..n-1 sorted by key(i), stable: LSD radix passes of 11 bits (the sequences have up to 65 536 distinct symbols;
// comparison sorts were a fifth of the table build).
template <class Key>
void radixOrder(size_t n, Key key, std::vector<int>& order) {
    order.resize(n);
    std::iota(order.begin(), order.end(), 0);
    uint32_t maxKey = 0;
    bool sorted = true;
    for (size_t i = 0; i < n; ++i) {
        maxKey = std::max(maxKey, key(static_cast<int>(i)));
        sorted = sorted && (i == 0 || key(static_cast<int>(i - 1)) <= key(static_cast<int>(i)));
    }
    if (sorted) return;
    if (n < 48) {
        std::stable_sort(order.begin(), order.end(), [&](int a, int b) { return key(a) < key(b); });
        return;
    }
    std::vector<int> tmp(n);
    uint32_t bucket[2049];
    for (int shift = 0; shift < 32 && (maxKey >> shift) != 0; shift += 11) {
        std::fill(std::begin(bucket), std::end(bucket), 0u);
        for (size_t i = 0; i < n; ++i) ++bucket[((key(order[i]) >> shift) & 2047u) + 1];
        for (int b = 0; b < 2048; ++b) bucket[b + 1] += bucket[b];
        for (size_t i = 0; i < n; ++i) tmp[bucket[(key(order[i]) >> shift) & 2047u]++] = order[i];
        order.swap(tmp);
    }
}

// libstdc++'s binary heap (bits/stl_heap.h: __push_heap, __adjust_heap, as driven by std::priority_queue::push / pop) on
// packed nodes: frequency << 20 | node id, compared on the frequency alone — ties must fall exactly as they do in the
// reference's std::priority_queue<HuffmanNode, ..., NodeCompare>, so the sift loops are restated
// move for move; only the node shrinks from the reference's vectors to 8 bytes.
struct FreqHeap {
    std::vector<uint64_t> a;
    static bool greater(uint64_t x, uint64_t y) { return (x >> 20) > (y >> 20); }  // comp(x, y) of a min-heap on frequency
    void siftUp(size_t hole, size_t top, uint64_t value) {  // __push_heap
        while (hole > top) {
            const size_t parent = (hole - 1) / 2;
            if (!greater(a[parent], value)) break;
            a[hole] = a[parent];
            hole = parent;
        }
        a[hole] = value;
    }
    void push(uint64_t v) {
        a.push_back(v);
        siftUp(a.size() - 1, 0, v);
    }
    uint64_t top() const { return a.front(); }
    void pop() {  // pop_heap + pop_back
        const size_t len = a.size() - 1;
        if (len > 0) {
            const uint64_t value = a[len];
            size_t hole = 0, child = 0;
            while (child < (len - 1) / 2) {  // __adjust_heap(first, 0, len, value)
                child = 2 * (child + 1);
                if (greater(a[child], a[child - 1])) --child;
                a[hole] = a[child];
                hole = child;
            }
            if ((len & 1) == 0 && child == (len - 2) / 2) {
                child = 2 * (child + 1);
                a[hole] = a[child - 1];
                hole = child - 1;
            }
            siftUp(hole, 0, value);
        }
        a.pop_back();
    }
};

}  // namespace

void planSequence(const std::vector<SymbolStat>& stats, SequencePlan& plan) {
    const size_t n = stats.size();
    std::vector<int> order;
    radixOrder(n, [&](int i) { return stats[i].firstPos; }, order);

    // --- frequency map in the reference's insertion order (Huffman.cpp:47-71) ---
    // Same libstdc++ hashtable as the reference's std::unordered_map (only the node allocator differs: one arena per
    // call instead of one malloc per symbol; the allocator has no influence on bucket or iteration order).
    std::pmr::monotonic_buffer_resource arena((n + 2) * 48 + 1024);
    std::pmr::unordered_map<uint32_t, uint32_t> slot(&arena);  // symbol -> leaf id
    uint32_t symWidth = 1;
    uint64_t dataCount = 0;
    std::vector<uint64_t> leafFreq;
    leafFreq.reserve(n + 1);
    for (int i : order) {
        symWidth = std::max<uint32_t>(symWidth, std::bit_width(stats[i].symbol));
        slot[stats[i].symbol] = static_cast<uint32_t>(leafFreq.size());
        leafFreq.push_back(stats[i].freq);
        dataCount += stats[i].freq;
    }
    slot[kEof] = static_cast<uint32_t>(leafFreq.size());
    leafFreq.push_back(0);
    const size_t nLeaves = leafFreq.size();

    // --- code lengths (Huffman.cpp:72-103) ---
    std::vector<int32_t> parent(nLeaves, -1);
    parent.reserve(2 * nLeaves);
    std::vector<uint8_t> height(nLeaves, 0);  // HuffmanNode::depth: max(children) + 1 (Huffman.cpp:96)
    height.reserve(2 * nLeaves);
    FreqHeap heap;
    heap.a.reserve(nLeaves);
    for (const auto& kv : slot) heap.push(leafFreq[kv.second] << 20 | kv.second);
    while (heap.a.size() > 1) {
        const uint64_t a = heap.top();
        heap.pop();
        const uint64_t b = heap.top();
        heap.pop();
        const uint32_t ia = static_cast<uint32_t>(a & 0xFFFFFu), ib = static_cast<uint32_t>(b & 0xFFFFFu);
        const uint32_t id = static_cast<uint32_t>(parent.size());
        parent.push_back(-1);
        height.push_back(static_cast<uint8_t>(std::max(height[ia], height[ib]) + 1));
        parent[ia] = static_cast<int32_t>(id);
        parent[ib] = static_cast<int32_t>(id);
        heap.push(((a >> 20) + (b >> 20)) << 20 | id);
    }
    const uint32_t maxLen = std::max<uint32_t>(height[heap.top() & 0xFFFFFu], 1);  // :110
    // depth of every node from the root down (ids are created bottom-up, so parents have larger ids)
    std::vector<uint32_t> depth(parent.size(), 0);
    for (size_t i = parent.size(); i-- > 0;)
        if (parent[i] >= 0) depth[i] = depth[parent[i]] + 1;

    // sort by (length, symbol) (:28-36): by symbol first (radix, usually already in order), then a stable counting sort by length
    std::vector<int> bySymbol;
    radixOrder(n, [&](int k) { return stats[order[k]].symbol; }, bySymbol);  // k = leaf id
    const uint32_t eofLen = dataCount == 0 ? 1u : depth[n];  // empty input: EOF gets length 1 (:66-71)
    uint32_t longestLen = eofLen;
    for (size_t k = 0; k < n; ++k) longestLen = std::max(longestLen, depth[k]);
    std::vector<uint32_t> lenStart(longestLen + 2, 0);
    for (size_t k = 0; k < n; ++k) ++lenStart[depth[k] + 1];
    ++lenStart[eofLen + 1];
    for (uint32_t l = 0; l <= longestLen; ++l) lenStart[l + 1] += lenStart[l];
    std::vector<Sym> sorted(nLeaves);
    for (int k : bySymbol) sorted[lenStart[depth[k]]++] = Sym{stats[order[k]].symbol, depth[k], 0, order[k]};
    sorted[lenStart[eofLen]++] = Sym{kEof, eofLen, 0, -1};  // 0xFFFFFFFF: the largest symbol of its length

    // --- canonical codes + header (Huffman.cpp:107-156) ---
    BitString huff;
    huff.put(maxLen, 8);
    std::vector<uint32_t> counts;
    uint32_t prevLen = 0, code = 0, count = 0;
    for (Sym& e : sorted) {
        if (e.len != prevLen) {
            if (prevLen != 0) counts.push_back(count);
            for (uint32_t l = prevLen + 1; l < e.len; ++l) counts.push_back(0);
            count = 0;
            code = (e.len - prevLen) >= 32 ? 0u : code << (e.len - prevLen);
        }
        e.code = code++;
        ++count;
        prevLen = e.len;
    }
    counts.push_back(count);
    for (uint32_t c : counts) huff.put(c & 0xFFFFu, 16);
    huff.put(symWidth, 8);
    const uint32_t mask = (1u << symWidth) - 1u;
    size_t pos = 0;
    for (uint32_t c : counts) {
        const uint32_t rangeLen = c & 0xFFFFu;
        if (rangeLen == 0) continue;
        // Elias-Fano iff its *upper bound* beats raw symbols (Huffman.cpp:148; BitBuffer.cpp:345-354)
        const uint32_t m = std::bit_width(mask);
        const uint32_t nb = std::bit_width(rangeLen);
        const uint32_t lb = (m >= nb) ? m - nb : 0;
        const uint32_t bound = rangeLen * (1 + lb) + (1u << (m - lb)) - 1;
        if (bound < rangeLen * symWidth) {
            uint32_t bucket = 0;
            for (uint32_t k = 0; k < rangeLen; ++k) {  // BitBuffer.cpp:301-313
                const uint32_t want = (sorted[pos + k].symbol & mask & 0xFFFFu) >> lb;
                while (bucket != want) {
                    const uint32_t z = std::min<uint32_t>(want - bucket, 32);
                    huff.put(0, static_cast<int>(z));
                    bucket += z;
                }
                huff.put(1, 1);
            }
            for (uint32_t k = 0; k < rangeLen; ++k) huff.put(sorted[pos + k].symbol & mask & 0xFFFFu, static_cast<int>(lb));
        } else {
            for (uint32_t k = 0; k < rangeLen; ++k) huff.put(sorted[pos + k].symbol & mask & 0xFFFFu, static_cast<int>(symWidth));
        }
        pos += rangeLen;
    }

    plan.code.assign(n, 0);
    plan.len.assign(n, 0);
    uint64_t huffPayload = 0;
    uint32_t longest = 0;
    for (const Sym& e : sorted) {
        longest = std::max(longest, e.len);
        if (e.origin >= 0) {
            plan.code[e.origin] = e.code;
            plan.len[e.origin] = static_cast<uint8_t>(e.len);
            huffPayload += static_cast<uint64_t>(stats[e.origin].freq) * e.len;
        } else {
            plan.eofCode = e.code;
            plan.eofLen = static_cast<uint8_t>(e.len);
        }
    }
    plan.maxCodeLen = longest;
    const uint64_t huffBits = huff.nbits + huffPayload + plan.eofLen;

    // --- Huffman vs Golomb (CompressedImage.cpp:362-379): first strictly smaller estimate wins ---
    uint64_t best = huffBits;
    int bestM = -1;
    uint64_t bestPayload = 0;
    for (int m = 1; m < 2048;) {
        // golombCodeLength (BitBuffer.cpp:260-269) summed over the histogram. v / M by reciprocal multiplication:
        // exact for v < 2^16, M < 2^11 (27 significant bits needed, 32 available).
        const uint32_t M = static_cast<uint32_t>(m);
        const uint32_t b = std::bit_width(M);
        const uint32_t limit = (1u << (b + 1)) - M;
        const uint64_t inv = (1ull << 32) / M + 1;
        uint64_t est = 16;
        for (size_t i = 0; i < n && est < best;) {  // the estimate only grows: stop once it can no longer win
            const size_t stop = std::min(n, i + 64);
            for (; i < stop; ++i) {
                const SymbolStat& s = stats[i];
                const uint32_t q = s.symbol < 65536u ? static_cast<uint32_t>((s.symbol * inv) >> 32) : s.symbol / M;
                const uint32_t rem = s.symbol - q * M;
                est += static_cast<uint64_t>(s.freq) * (b + q + 1 + (rem >= limit ? 1u : 0u));
            }
        }
        if (est < best) {
            best = est;
            bestM = m;
            bestPayload = est - 16;
        }
        m = (m & 1) ? m + 1 : (m << 1) - 1;
    }
    plan.header = BitString();
    if (bestM < 0) {
        plan.golomb = false;
        plan.golombM = 0;
        plan.header.put(0, 1);
        plan.header.append(huff);
        plan.payloadBits = huffPayload;
    } else {
        plan.golomb = true;
        plan.golombM = static_cast<uint32_t>(bestM);
        plan.header.put(1, 1);
        plan.header.put(static_cast<uint32_t>(bestM), 16);
        plan.payloadBits = bestPayload;
    }
}

int quantTables(int K, int blockSize, double bpp, float* out) {
    if (K < 1 || K > 32 || blockSize < 1) return -1;
    float allocated = 0.0f;
    std::vector<float> alloc(3 * K), variances(3 * K);
    for (int i = 0; i < K; ++i) {
        variances[i] = static_cast<float>(kVarY[i]);
        variances[i + K] = static_cast<float>(kVarU[i]);
        variances[i + 2 * K] = static_cast<float>(kVarV[i]);
    }
    const float bs = static_cast<float>(blockSize);
    while ((allocated / static_cast<double>(blockSize * blockSize)) < bpp) {
        const ptrdiff_t index = std::max_element(variances.cbegin(), variances.cend()) - variances.cbegin();
        alloc[index] += 1.0f;
        variances[index] /= 2.0f;
        allocated += 1.0f;
        const float floorQ = (index == 0 || index == K || index == 2 * K) ? bs : 1.0f;
        if (255.0f * bs * std::pow(0.5f, alloc[index]) < floorQ) variances[index] = 0.0f;
    }
    const float decay[3] = {static_cast<float>(kDecayY), static_cast<float>(kDecayU), static_cast<float>(kDecayV)};
    for (int c = 0; c < 3; ++c)
        for (int i = 0; i < K; ++i) {
            const float floorQ = i == 0 ? bs : 1.0f;
            out[c * K + i] = std::max(std::ceil(255.0f * bs * std::pow(decay[c], static_cast<float>(i)) * std::pow(0.5f, alloc[i + c * K])), floorQ);
        }
    return 0;
}

}  // namespace mnenc
