/*
 * mnenc — C ABI of the B200-native `.mn` encoder (drop-in for the CompressionLib encoder hot path of
 * mnesbit/ImageExperiments). Plain pointers and sizes only; no C++/torch types cross this boundary.
 *
 * Every entry point names the reference interface it replaces (paths relative to the reference root).
 * Return value: 0 = OK, negative = error (text via mnenc_last_error()). No exception crosses the ABI.
 * A context is bound to one CUDA device and must be used by one host thread at a time.
 *
 * There is no CPU fallback: every call that needs the GPU fails with MNENC_ERR_CUDA when no sm_100 device
 * is usable.
 */
#ifndef MNENC_H_
#define MNENC_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define MNENC_OK 0
#define MNENC_ERR_ARG (-1)      /* bad argument (K, block size, null pointer, ...) */
#define MNENC_ERR_CUDA (-2)     /* CUDA runtime error or no usable device */
#define MNENC_ERR_STATE (-3)    /* call order (e.g. encode before set_quant) */
#define MNENC_ERR_INTERNAL (-4) /* consistency check failed inside the encoder */

typedef struct mnenc_ctx mnenc_t;

/* Thread-local text of the last error on this thread. */
const char* mnenc_last_error(void);

/*
 * Replaces the dictionary half of compressed::createCompressionContextFast
 * (CompressionLib/src/CompressedImage.cpp:317-341): the dictionary is IMPORTED, never regenerated
 * (it is build-dependent float data). A caller holding a CompressionContextFast passes
 *   base_rowmajor  = BaseDict              (n_base x block_size^2, row-major; CompressedImage.h:49)
 *   detail_rows[i] = DetailBasis[i].rows() (CompressedImage.h:41)
 *   detailY/U/V    = the channel's DetailBasis matrices concatenated in atom order, row-major
 * or the same floats recovered through the opaque Dynamic callbacks (CompressedImage.cpp:331-339), see
 * INTEGRATION.md. device = CUDA ordinal. Only block_size == 8, 1 <= K <= 32, n_base <= 512 are supported.
 */
int mnenc_create(mnenc_t** out, int device, int K, int block_size, const float* base_rowmajor, int n_base,
                 const int32_t* detail_rows, const float* detailY, const float* detailU, const float* detailV);
void mnenc_destroy(mnenc_t* ctx);

/*
 * Replaces compressed::createQuantizationTables, float overload (CompressedImage.cpp:168-210): the
 * `bppAllocation` quality parameter -> 3*K quantisers (Y then U then V). Host only, no GPU needed.
 */
int mnenc_quant_tables(int K, int block_size, double bpp_allocation, float* out_3K);

/* quantY/quantU/quantV arguments of encodeImageFast (CompressedImage.h:67-73), concatenated Y,U,V. */
int mnenc_set_quant(mnenc_t* ctx, const float* quant_3K);

/*
 * Replaces compressed::encodeImageFast (CompressedImage.cpp:578-633): RGB8 image -> complete `.mn` byte
 * stream that compressed::decodeImageFast (CompressedImage.cpp:837) round-trips.
 *   rgb: HOST pointer, packed R,G,B bytes, row y at rgb + y*row_stride_bytes (img::image<rgb>, image.h:33-104).
 *   *out: malloc'ed by the library, release with mnenc_free(); *out_bytes = outputByteSize.
 * The host->device copy of the pixels and the device->host copy of the stream are part of the call.
 */
int mnenc_encode_rgb8(mnenc_t* ctx, const uint8_t* rgb, size_t width, size_t height, size_t row_stride_bytes,
                      uint8_t** out, size_t* out_bytes);
/* Same, with the pixels already resident in device memory of the context's GPU. */
int mnenc_encode_rgb8_device(mnenc_t* ctx, const uint8_t* rgb_dev, size_t width, size_t height, size_t row_stride_bytes,
                             uint8_t** out, size_t* out_bytes);
void mnenc_free(uint8_t* p);

/*
 * Debug taps for parity tests (the reference has no equivalent call; they expose the intermediates of
 * encodeImageFast, CompressedImage.cpp:588-631):
 *   counts[3*blk + ch]                         = return value of CalcMPDynamicFast (MatchingPursuit.cpp:76)
 *   pairs[((3*blk + ch)*K + i)*2 + {0,1}]      = BasisChoice{deltaId,intCoeff} (MatchingPursuit.h:13-17), 0 for i >= count
 * blk runs in the reference's column-major block order. Valid after an encode call on this context.
 */
int mnenc_get_symbols(mnenc_t* ctx, uint16_t* counts, uint16_t* pairs);
/*
 * Back end only: replaces compressed::writeCompressed (CompressedImage.cpp:462-519) fed with the given
 * symbols (same layout as mnenc_get_symbols) instead of the matching-pursuit kernel's — "entropy coding is
 * bit-exact given identical symbols".
 */
int mnenc_pack_symbols(mnenc_t* ctx, size_t width, size_t height, const uint16_t* counts, const uint16_t* pairs,
                       uint8_t** out, size_t* out_bytes);

/*
 * Timing / accounting of the last encode on this context (CUDA events on the context's stream).
 * ms[0]=H2D pixels, ms[1]=matching pursuit kernel, ms[2]=stream scatter+DC+RLE, ms[3]=histogram,
 * ms[4]=host table build (wall), ms[5]=bit packing, ms[6]=D2H + header merge (wall), ms[7]=total wall.
 * Returns the number of kernels launched by that encode (>0), or a negative error.
 */
int mnenc_last_timing(mnenc_t* ctx, float* ms8);
/* Algorithmic work of the last encode: flops of the matching pursuit, symbols, payload bits. */
int mnenc_last_work(mnenc_t* ctx, double* mp_flops, uint64_t* n_symbols, uint64_t* n_bits);

/*
 * Host-logic test hook (used by the CPU-only tests; never called by the encode entry points): runs the host
 * table builder that replaces huffman::huffmanEncode's table construction (Huffman.cpp:65-156) and
 * compressed::writeHuffmanOrGolomb's choice (CompressedImage.cpp:359-390) on one sequence described by its
 * histogram and first-occurrence positions. Outputs: golomb_m (0 = Huffman), per-symbol codes/lens in input
 * order, EOF code, the header bits that precede the payload (mode flag + table, or flag + m), payload bits.
 */
int mnenc_debug_plan_sequence(const uint32_t* symbols, const uint32_t* freqs, const uint32_t* first_pos, size_t n,
                              uint32_t* golomb_m, uint32_t* codes, uint8_t* lens, uint32_t* eof_code, uint8_t* eof_len,
                              uint8_t* header_bytes, size_t header_cap, uint64_t* header_bits, uint64_t* payload_bits);

/*
 * Host-logic test hook of the multi-GPU stripe path (used by the CPU-only world_size-2 gloo tests): run stitching
 * of huffman::runLengthEncode (Huffman.cpp:246-279) across ranks. records = [nranks][n_streams][5] u32
 * {n, first value, first run length, last value, last run length} as all-gathered by the ranks; outputs for `rank`:
 * carry_in[s] / extend_out[s] = elements of its first / last run that live on earlier / later ranks.
 * mnenc_debug_rle_chunk_size = RLE symbols a rank emits for `len` elements of a run with `carry` elements before them
 * and `total` elements overall (chunks of 32 769 counted from the run's global start).
 */
int mnenc_debug_stitch_runs(const uint32_t* records, int nranks, int n_streams, int rank, uint32_t* carry_in, uint32_t* extend_out);
uint32_t mnenc_debug_rle_chunk_size(uint32_t carry, uint32_t len, uint64_t total);

#ifdef __cplusplus
}
#endif
#endif /* MNENC_H_ */
