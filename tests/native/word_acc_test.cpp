// Host build of csrc/word_acc.h (the accumulator packApplyKernelV2 runs per thread) against a bit-by-bit writer:
// random code widths 1..32, Golomb codes with long unary prefixes, items of one "thread" straddling a jump in position
// (sequence boundary), many threads OR-ing into one zeroed buffer in shuffled order. Exit code 0 = identical streams.
#include <algorithm>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <bit>
#include <random>
#include <vector>

#include "../../imageexperiments_b200/csrc/word_acc.h"

struct Item {
    unsigned long long pos;
    uint32_t value;  // code bits, or the symbol for a Golomb item
    int width;       // 0 = Golomb
    uint32_t M;
};

static void naiveBits(std::vector<uint8_t>& out, unsigned long long pos, unsigned long long value, int width) {
    for (int i = 0; i < width; ++i, ++pos)
        if ((value >> (width - 1 - i)) & 1) out[pos >> 3] |= static_cast<uint8_t>(0x80u >> (pos & 7));
}

int main() {
    std::mt19937 rng(20261104);
    for (int round = 0; round < 200; ++round) {
        const int nItems = 1 + static_cast<int>(rng() % 5000);
        const int style = round % 4;  // 0 short codes, 1 any width, 2 Golomb heavy, 3 mixed with position jumps
        std::vector<Item> items;
        unsigned long long pos = rng() % 97;
        for (int i = 0; i < nItems; ++i) {
            if (style == 3 && rng() % 64 == 0) pos += rng() % 300;  // header of the next sequence sits in between
            Item it{pos, 0, 0, 0};
            const bool golomb = style == 2 ? (rng() % 2 == 0) : (style == 3 && rng() % 8 == 0);
            if (golomb) {
                static const uint32_t ms[] = {1, 2, 3, 4, 7, 8, 15, 16, 31, 32, 63, 64, 127, 128, 255, 256, 511, 512, 1023, 1024, 2047};
                it.M = ms[rng() % 21];
                it.value = (rng() % 16 == 0) ? rng() % 65536 : rng() % (8 * it.M);
                const int b = std::bit_width(it.M);
                pos += it.value / it.M + 1 + b;
            } else {
                it.width = style == 0 ? 1 + static_cast<int>(rng() % 9) : 1 + static_cast<int>(rng() % 32);
                it.value = it.width == 32 ? rng() : rng() & ((1u << it.width) - 1u);
                pos += it.width;
            }
            items.push_back(it);
        }
        const size_t bytes = ((pos + 31) / 32 + 2) * 4;
        std::vector<uint8_t> want(bytes, 0);
        for (const Item& it : items) {
            if (it.width) {
                naiveBits(want, it.pos, it.value, it.width);
            } else {  // writeGolombCode, BitBuffer.cpp:228-241
                unsigned long long p = it.pos;
                const uint32_t q = it.value / it.M, rem = it.value % it.M;
                for (uint32_t k = 0; k < q; ++k) naiveBits(want, p++, 1, 1);
                naiveBits(want, p++, 0, 1);
                const int b = std::bit_width(it.M);
                const uint32_t limit = (1u << (b + 1)) - it.M;
                if (rem < limit) naiveBits(want, p, rem, b);
                else naiveBits(want, p, static_cast<unsigned long long>(rem) + limit, b + 1);
            }
        }
        std::vector<uint32_t> got(bytes / 4, 0);
        size_t sinkCalls = 0;
        auto sink = [&](unsigned long long w, uint32_t v) {
            if (w >= got.size()) {
                std::printf("round %d: word %llu out of range\n", round, w);
                std::exit(1);
            }
            got[w] |= v;
            ++sinkCalls;
        };
        // "threads" of 8 consecutive items, run in shuffled order
        std::vector<int> order((nItems + 7) / 8);
        for (size_t t = 0; t < order.size(); ++t) order[t] = static_cast<int>(t);
        std::shuffle(order.begin(), order.end(), rng);
        for (int t : order) {
            mnenc::WordAcc<decltype(sink)> acc(sink);
            for (int k = 8 * t; k < std::min(nItems, 8 * t + 8); ++k) {
                const Item& it = items[k];
                if (it.width) {
                    acc.put(it.pos, it.value, it.width);
                    continue;
                }
                const int b = std::bit_width(it.M);
                if (acc.putGolomb(it.pos, it.value, it.M, b) != it.pos + it.value / it.M + 1 + b) {
                    std::printf("round %d: Golomb end position wrong at item %d\n", round, k);
                    return 1;
                }
            }
            acc.flush();
        }
        if (std::memcmp(got.data(), want.data(), bytes) != 0) {
            std::printf("round %d (style %d, %d items): streams differ\n", round, style, nItems);
            return 1;
        }
        if (style == 0 && nItems > 1000 && sinkCalls * 2 > static_cast<size_t>(nItems)) {
            std::printf("round %d: %zu sink calls for %d short codes — the accumulator does not merge\n", round, sinkCalls, nItems);
            return 1;
        }
    }
    // SeqWalk: one search, then forward steps, against "largest q with start[q] <= g" computed the slow way
    for (int round = 0; round < 2000; ++round) {
        const int nSeq = 1 + static_cast<int>(rng() % 193);
        std::vector<uint32_t> start(nSeq + 1, 0);
        for (int q = 0; q < nSeq; ++q) {
            const uint32_t len = (rng() % 3 == 0) ? 0u : (rng() % 4 == 0 ? rng() % 40 : rng() % 4);  // many empty and tiny sequences
            start[q + 1] = start[q] + len;
        }
        const uint32_t total = start[nSeq];
        if (total == 0) continue;
        for (int t = 0; t < 64; ++t) {
            const uint32_t first = rng() % total;
            mnenc::SeqWalk walk(start.data(), nSeq, first);
            for (uint32_t g = first; g < std::min(total, first + 8); ++g) {
                walk.advance(g);
                int want = 0;
                for (int q = 0; q < nSeq; ++q)
                    if (start[q] <= g) want = q;
                if (walk.q != want || walk.seqStart != start[want] || g < walk.seqStart || g >= walk.seqEnd) {
                    std::printf("SeqWalk round %d: element %u -> sequence %d, want %d\n", round, g, walk.q, want);
                    return 1;
                }
            }
        }
    }
    std::puts("word_acc OK");
    return 0;
}
