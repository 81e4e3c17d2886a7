// TEST INFRASTRUCTURE — NOT PART OF THE PRODUCT PATH.
//
// C-ABI wrapper around the UNMODIFIED reference (mnesbit/ImageExperiments) so that Python
// tests / bench.py's cpu_baseline leg can call the reference's own CPU encoder, decoder and
// entropy primitives. The reference sources are compiled where they lie under $(REF)
// (default /root/reference) by oracle/Makefile; this file is the only source of ours that
// goes into oracle/_ref/libmnref.so. Nothing here is copied from the reference: it only
// *calls* the reference's public functions:
//   compressed::createCompressionContextFast   CompressionLib/src/CompressedImage.cpp:317
//   compressed::encodeImageFast                CompressionLib/src/CompressedImage.cpp:578
//   compressed::decodeImageFast                CompressionLib/src/CompressedImage.cpp:837
//   compressed::readCompressedFast             CompressionLib/src/CompressedImage.cpp:709
//   compressed::writeHuffmanOrGolomb           CompressionLib/src/CompressedImage.cpp:359
//   compressed::calculatePSNR                  CompressionLib/src/CompressedImage.cpp:343
//   matching::CalcMPDynamicFast                CompressionLib/src/MatchingPursuit.cpp:76
//   huffman::runLengthEncode / huffmanEncode   CompressionLib/src/Huffman.cpp:246 / 46
//   bitbuffer::golombCodeLength etc.           CompressionLib/src/BitBuffer.cpp:228-354
#include "CompressionLib/inc/CompressedImage.h"
#include "CompressionLib/inc/Huffman.h"
#include "CompressionLib/inc/BitBuffer.h"
#include "ImageHelper/inc/misc.h"

#include <cstdlib>
#include <cstring>
#include <iostream>
#include <mutex>
#include <sstream>
#include <stdexcept>
#include <string>

// Functions with external linkage that the reference headers do not declare.
namespace compressed {
void writeHuffmanOrGolomb(const std::vector<uint16_t>& data, bitbuffer::BitBuffer& buffer);
void createQuantizationTables(const size_t K, const size_t blockSize, const double bppAllocation,
                              Eigen::VectorXf& quantY, Eigen::VectorXf& quantU, Eigen::VectorXf& quantV);
std::unique_ptr<CompressionContextFast> readCompressedFast(const uint8_t bytes[], size_t byteSize, size_t& K,
                                                           size_t& blockSize, size_t& width, size_t& height,
                                                           std::vector<uint16_t>& lengths,
                                                           std::vector<std::vector<uint16_t>>& codes);
}  // namespace compressed

namespace {

thread_local std::string g_err;

// The reference prints a progress line per block column (CompressedImage.cpp:593,852). std::cout is switched once,
// process-wide, to a stateless discarding buffer (safe when several threads encode at the same time).
struct NullBuf : std::streambuf {
    int overflow(int c) override { return traits_type::not_eof(c); }
    std::streamsize xsputn(const char*, std::streamsize n) override { return n; }
};
struct CoutMute {
    CoutMute() {
        static NullBuf sink;
        static std::once_flag once;
        std::call_once(once, [] { std::cout.rdbuf(&sink); });
    }
};

template <class F>
int guarded(F&& f) {
    try {
        f();
        return 0;
    } catch (std::exception* e) {  // the reference throws pointers
        g_err = e->what();
        return -1;
    } catch (const std::exception& e) {
        g_err = e.what();
        return -2;
    } catch (...) {
        g_err = "unknown exception";
        return -3;
    }
}

std::unique_ptr<img::image<img::rgb>> wrapImage(const uint8_t* rgb, size_t w, size_t h) {
    auto im = std::make_unique<img::image<img::rgb>>(w, h, false);
    std::memcpy(im->data, rgb, w * h * 3);
    return im;
}

using Ctx = compressed::CompressionContextFast;

compressed::ChannelContextFast& channel(Ctx* c, int ch) { return ch == 0 ? c->Y : (ch == 1 ? c->U : c->V); }

uint8_t* saveBits(bitbuffer::BitBuffer& bits, size_t* nbits, size_t* nbytes) {
    *nbits = bits.Size();
    size_t n = 0;
    std::unique_ptr<uint8_t[]> data = bits.Save(n);
    *nbytes = n;
    uint8_t* out = static_cast<uint8_t*>(std::malloc(n ? n : 1));
    std::memcpy(out, data.get(), n);
    return out;
}

}  // namespace

extern "C" {

const char* mnref_last_error() { return g_err.c_str(); }
void mnref_free(void* p) { std::free(p); }

void* mnref_create(int K, int blockSize, double bpp) {
    Ctx* out = nullptr;
    guarded([&] { out = compressed::createCompressionContextFast(K, blockSize, bpp).release(); });
    return out;
}
void mnref_destroy(void* c) { delete static_cast<Ctx*>(c); }

// Quantisation tables only (cheap; no dictionary build).
int mnref_quant_tables(int K, int blockSize, double bpp, float* out3K) {
    return guarded([&] {
        Eigen::VectorXf y, u, v;
        compressed::createQuantizationTables(K, blockSize, bpp, y, u, v);
        for (int i = 0; i < K; ++i) {
            out3K[i] = y[i];
            out3K[K + i] = u[i];
            out3K[2 * K + i] = v[i];
        }
    });
}

void mnref_get_quant(void* c, float* out3K) {
    Ctx* ctx = static_cast<Ctx*>(c);
    const int K = static_cast<int>(ctx->K);
    for (int i = 0; i < K; ++i) {
        out3K[i] = ctx->Y.Quant[i];
        out3K[K + i] = ctx->U.Quant[i];
        out3K[2 * K + i] = ctx->V.Quant[i];
    }
}
void mnref_set_quant(void* c, const float* in3K) {
    Ctx* ctx = static_cast<Ctx*>(c);
    const int K = static_cast<int>(ctx->K);
    for (int i = 0; i < K; ++i) {
        ctx->Y.Quant[i] = in3K[i];
        ctx->U.Quant[i] = in3K[K + i];
        ctx->V.Quant[i] = in3K[2 * K + i];
    }
}

// Dictionary export. detailRows[n_base] receives the row count of each detail group.
int mnref_dict_shape(void* c, int* nBase, int* cols, int* detailRows) {
    Ctx* ctx = static_cast<Ctx*>(c);
    *nBase = static_cast<int>(ctx->BaseDict.rows());
    *cols = static_cast<int>(ctx->BaseDict.cols());
    if (detailRows) {
        for (size_t i = 0; i < ctx->Y.DetailBasis.size(); ++i) detailRows[i] = static_cast<int>(ctx->Y.DetailBasis[i].rows());
    }
    return 0;
}
void mnref_get_base(void* c, float* rowMajor) {
    Ctx* ctx = static_cast<Ctx*>(c);
    const auto& m = ctx->BaseDict;
    for (Eigen::Index r = 0; r < m.rows(); ++r)
        for (Eigen::Index j = 0; j < m.cols(); ++j) rowMajor[r * m.cols() + j] = m(r, j);
}
// All detail groups of one channel, concatenated in atom order, row-major.
void mnref_get_detail(void* c, int ch, float* rowMajor) {
    Ctx* ctx = static_cast<Ctx*>(c);
    size_t o = 0;
    for (const auto& m : channel(ctx, ch).DetailBasis)
        for (Eigen::Index r = 0; r < m.rows(); ++r)
            for (Eigen::Index j = 0; j < m.cols(); ++j) rowMajor[o++] = m(r, j);
}
// The same dictionary, but obtained the way a drop-in shim must: through the opaque callback
// (CompressedImage.cpp:331-339). prev = -1 -> dynamic(0,{}); else dynamic(1,{BasisChoice{prev,0}}).
int mnref_dynamic_probe(void* c, int ch, int prev, float* rowMajor, int* rows) {
    Ctx* ctx = static_cast<Ctx*>(c);
    return guarded([&] {
        std::vector<matching::BasisChoice> ch0(ctx->K);
        int n = 0;
        if (prev >= 0) {
            ch0[0].deltaId = static_cast<unsigned short>(prev);
            ch0[0].intCoeff = 0;
            n = 1;
        }
        Eigen::MatrixXf d = channel(ctx, ch).Dynamic(n, ch0);
        *rows = static_cast<int>(d.rows());
        if (rowMajor)
            for (Eigen::Index r = 0; r < d.rows(); ++r)
                for (Eigen::Index j = 0; j < d.cols(); ++j) rowMajor[r * d.cols() + j] = d(r, j);
    });
}

// encodeImageFast as-is.
int mnref_encode(void* c, const uint8_t* rgb, size_t w, size_t h, uint8_t** out, size_t* outBytes) {
    Ctx* ctx = static_cast<Ctx*>(c);
    return guarded([&] {
        CoutMute mute;
        auto im = wrapImage(rgb, w, h);
        size_t n = 0;
        std::unique_ptr<uint8_t[]> bytes = compressed::encodeImageFast(im.get(), ctx->K, ctx->BlockSize, ctx->Y.Quant, ctx->U.Quant,
                                                                       ctx->V.Quant, ctx->Y.Dynamic, ctx->U.Dynamic, ctx->V.Dynamic, n);
        *out = static_cast<uint8_t*>(std::malloc(n ? n : 1));
        std::memcpy(*out, bytes.get(), n);
        *outBytes = n;
    });
}

// Block walk of encodeImageFast (CompressedImage.cpp:592-631) restricted to block columns [bx0,bx1),
// calling the reference's own YUVFromRGB and CalcMPDynamicFast, returning the raw symbols:
//   counts[3*blk + ch], pairs[((3*blk + ch)*K + i)*2 + {0:deltaId,1:intCoeff}]   (blk in column-major order from bx0)
int mnref_symbols(void* c, const uint8_t* rgb, size_t w, size_t h, size_t bx0, size_t bx1, uint16_t* counts, uint16_t* pairs) {
    Ctx* ctx = static_cast<Ctx*>(c);
    return guarded([&] {
        const size_t K = ctx->K, bs = ctx->BlockSize;
        std::vector<matching::BasisChoice> choices(K);
        Eigen::VectorXf block[3] = {Eigen::VectorXf(bs * bs), Eigen::VectorXf(bs * bs), Eigen::VectorXf(bs * bs)};
        size_t blk = 0;
        for (size_t bx = bx0; bx < bx1; ++bx) {
            for (size_t y = 0; y < h; y += bs, ++blk) {
                for (size_t dx = 0; dx < bs; ++dx)
                    for (size_t dy = 0; dy < bs; ++dy) {
                        size_t u = bx * bs + dx, v = y + dy;
                        float fy = 0.f, fu = 0.f, fv = 0.f;
                        if (u < w && v < h) {
                            const uint8_t* p = rgb + 3 * (v * w + u);
                            img::rgb pt{p[0], p[1], p[2]};
                            img::yuv col = img::YUVFromRGB(pt);
                            fy = static_cast<float>(col.y);
                            fu = static_cast<float>(col.u);
                            fv = static_cast<float>(col.v);
                        }
                        block[0][dx + bs * dy] = fy;
                        block[1][dx + bs * dy] = fu;
                        block[2][dx + bs * dy] = fv;
                    }
                for (int ch = 0; ch < 3; ++ch) {
                    auto& cc = channel(ctx, ch);
                    int count = matching::CalcMPDynamicFast(static_cast<int>(K), cc.Quant, choices, block[ch], cc.Dynamic);
                    counts[3 * blk + ch] = static_cast<uint16_t>(count);
                    uint16_t* p = pairs + (3 * blk + ch) * K * 2;
                    for (size_t i = 0; i < K; ++i) {
                        p[2 * i] = i < static_cast<size_t>(count) ? choices[i].deltaId : 0;
                        p[2 * i + 1] = i < static_cast<size_t>(count) ? choices[i].intCoeff : 0;
                    }
                }
            }
        }
    });
}

// Parse a .mn stream back into the symbol streams with the reference's reader (DC chains undone).
// lengths: 3*nblocks u16 ; streamSizes: 6K ; streams: concatenation of the 6K streams (malloc'ed).
int mnref_parse(const uint8_t* bytes, size_t n, size_t* K, size_t* bs, size_t* w, size_t* h, uint16_t** lengths, size_t* nLengths,
                size_t* streamSizes, uint16_t** streams) {
    return guarded([&] {
        CoutMute mute;
        std::vector<uint16_t> len;
        std::vector<std::vector<uint16_t>> codes;
        auto ctx = compressed::readCompressedFast(bytes, n, *K, *bs, *w, *h, len, codes);
        *nLengths = len.size();
        *lengths = static_cast<uint16_t*>(std::malloc(len.size() * 2 + 2));
        std::memcpy(*lengths, len.data(), len.size() * 2);
        size_t total = 0;
        for (size_t i = 0; i < codes.size(); ++i) {
            streamSizes[i] = codes[i].size();
            total += codes[i].size();
        }
        *streams = static_cast<uint16_t*>(std::malloc(total * 2 + 2));
        size_t o = 0;
        for (const auto& s : codes) {
            std::memcpy(*streams + o, s.data(), s.size() * 2);
            o += s.size();
        }
    });
}

int mnref_decode(const uint8_t* bytes, size_t n, uint8_t** rgbOut, size_t* w, size_t* h) {
    return guarded([&] {
        CoutMute mute;
        auto im = compressed::decodeImageFast(bytes, n);
        *w = im->width();
        *h = im->height();
        *rgbOut = static_cast<uint8_t*>(std::malloc(*w * *h * 3));
        std::memcpy(*rgbOut, im->data, *w * *h * 3);
    });
}

double mnref_psnr(const uint8_t* a, const uint8_t* b, size_t w, size_t h) {
    auto ia = wrapImage(a, w, h);
    auto ib = wrapImage(b, w, h);
    return compressed::calculatePSNR(ia.get(), ib.get());
}

// Entropy primitives for per-stream parity tests.
int mnref_write_huffman_or_golomb(const uint16_t* data, size_t n, uint8_t** out, size_t* nbits, size_t* nbytes) {
    return guarded([&] {
        std::vector<uint16_t> v(data, data + n);
        bitbuffer::BitBuffer bits;
        compressed::writeHuffmanOrGolomb(v, bits);
        *out = saveBits(bits, nbits, nbytes);
    });
}
int mnref_huffman_encode(const uint16_t* data, size_t n, uint8_t** out, size_t* nbits, size_t* nbytes) {
    return guarded([&] {
        std::vector<uint16_t> v(data, data + n);
        bitbuffer::BitBuffer bits;
        huffman::huffmanEncode(v, bits);
        *out = saveBits(bits, nbits, nbytes);
    });
}
int mnref_huffman_decode(const uint8_t* bytes, size_t nbytes, uint16_t** out, size_t* n) {
    return guarded([&] {
        std::vector<uint16_t> v = huffman::huffmanDecode(bytes, nbytes);
        *n = v.size();
        *out = static_cast<uint16_t*>(std::malloc(v.size() * 2 + 2));
        std::memcpy(*out, v.data(), v.size() * 2);
    });
}
int mnref_rle_encode(const uint16_t* data, size_t n, uint16_t** out, size_t* nout) {
    return guarded([&] {
        std::vector<uint16_t> v(data, data + n);
        std::vector<uint16_t> r = huffman::runLengthEncode(v);
        *nout = r.size();
        *out = static_cast<uint16_t*>(std::malloc(r.size() * 2 + 2));
        std::memcpy(*out, r.data(), r.size() * 2);
    });
}
int mnref_rle_decode(const uint16_t* data, size_t n, uint16_t** out, size_t* nout) {
    return guarded([&] {
        std::vector<uint16_t> v(data, data + n);
        std::vector<uint16_t> r = huffman::runLengthDecode(v);
        *nout = r.size();
        *out = static_cast<uint16_t*>(std::malloc(r.size() * 2 + 2));
        std::memcpy(*out, r.data(), r.size() * 2);
    });
}
uint32_t mnref_golomb_length(uint32_t value, uint32_t m) { return bitbuffer::golombCodeLength(value, m); }
int mnref_golomb_write(const uint32_t* values, size_t n, uint32_t m, uint8_t** out, size_t* nbits, size_t* nbytes) {
    return guarded([&] {
        bitbuffer::BitBuffer bits;
        for (size_t i = 0; i < n; ++i) bitbuffer::writeGolombCode(values[i], m, bits);
        *out = saveBits(bits, nbits, nbytes);
    });
}
uint32_t mnref_elias_fano_length(size_t n, uint16_t maxSymbol) { return bitbuffer::eliasFanoSequenceCodeLength(n, maxSymbol); }
int mnref_elias_fano_write(const uint16_t* seq, size_t n, uint16_t maxSymbol, uint8_t** out, size_t* nbits, size_t* nbytes) {
    return guarded([&] {
        std::vector<uint16_t> v(seq, seq + n);
        bitbuffer::BitBuffer bits;
        bitbuffer::writeEliasFanoSequenceCode(v, maxSymbol, bits);
        *out = saveBits(bits, nbits, nbytes);
    });
}
uint32_t mnref_zigzag_encode(int32_t x) { return bitbuffer::zigzagEncode(x); }
int32_t mnref_zigzag_decode(uint32_t x) { return bitbuffer::zigzagDecode(x); }

}  // extern "C"
