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
#define MNENC_ERR_STREAM (-5)   /* decoder: not a `.mn` stream, or a damaged one (the reference throws range_error("Invalid ...")) */

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
 * Batch / throughput mode (BASELINE.json configs[4]): n_images independent images -> n_images independent `.mn` streams,
 * what a caller of the reference gets from encodeImageFast in a loop (Compression.cpp:117-123), each stream byte-identical
 * to the one-at-a-time result. Inside, the images are pipelined over lanes (own stream + buffers, one shared dictionary in
 * HBM; MNENC_LANES, default 9 with >= 12 host threads else 6, at most 16) so that host stages, H2D and D2H of one image overlap the matching pursuit of another.
 * rgb[i]: host pointers (on_device = 0) or pointers into the context's GPU (on_device = 1); out[i] from mnenc_free().
 * On error every out[i] is NULL. mnenc_last_timing/work afterwards describe the whole batch (ms[1] = summed kernel time,
 * ms[7] = wall time of the batch).
 */
int mnenc_encode_batch_rgb8(mnenc_t* ctx, size_t n_images, const uint8_t* const* rgb, int on_device, const size_t* width,
                            const size_t* height, const size_t* row_stride_bytes, uint8_t** out, size_t* out_bytes);

/*
 * Decoder. Replaces
 *   std::unique_ptr<image<rgb>> compressed::decodeImageFast(const uint8_t bytes[], size_t byteSize)
 *       CompressionLib/inc/CompressedImage.h:75, CompressionLib/src/CompressedImage.cpp:837-889
 * (readCompressedFast :709-782, huffmanDecode Huffman.cpp:173-244, runLengthDecode :281-307, readGolombCode
 * BitBuffer.cpp:243-258, FromCoeffsDynamicFast MatchingPursuit.cpp:130-147, RGBFromYUV misc.cpp:28-36).
 *   mn, mn_bytes: the stream, host memory. The quantisation tables are the stream's; the dictionary is the context's
 *                 (the reference rebuilds it in readCompressedFast :726 — same floats by construction);
 *                 K of the stream may differ from the context's.
 *   *rgb:         packed RGB8, row-major, 3 * width bytes per row (image<rgb>::data), from the library: mnenc_free().
 * The pixels are those of the reference's decoder, bit for bit. A stream the reference would reject (bad magic, no
 * pseudo-EOF) and any stream whose lengths are inconsistent — where the reference reads out of range — give
 * MNENC_ERR_STREAM; block sizes other than 8 and K > 32 give MNENC_ERR_STREAM as well (unsupported).
 * The entropy decode runs on the host (decode_host.cpp), the block loop on the GPU. By default the context's host threads
 * help the calling thread with large Huffman payloads (speculative chunk starts, same result); they busy-spin for the length
 * of the call, so on a host shared with other work set the environment variable MNENC_DECODE_THREADS=1, which keeps the
 * whole read on the calling thread (about 1.7x slower per image). mnenc_last_timing afterwards: ms[0] = header + `lengths`, ms[1] = symbol streams, ms[2] = uploads + kernels +
 * download (host clock), ms[3] = gather + reconstruct kernels (CUDA events), ms[7] = whole call.
 */
int mnenc_decode_rgb8(mnenc_t* ctx, const uint8_t* mn, size_t mn_bytes, uint8_t** rgb, size_t* width, size_t* height);
/* Same, writing into device memory of the context's GPU (capacity_bytes >= row_stride_bytes * (height - 1) + 3 * width). */
int mnenc_decode_rgb8_device(mnenc_t* ctx, const uint8_t* mn, size_t mn_bytes, uint8_t* rgb_dev, size_t capacity_bytes,
                             size_t row_stride_bytes, size_t* width, size_t* height);
/*
 * Batch / throughput mode of the decoder: n_images independent streams -> n_images images, what decodeImageFast gives in a
 * loop; pipelined over the same lanes as mnenc_encode_batch_rgb8 (one host thread per lane runs the entropy decode of its
 * stream while the other lanes' kernels and copies proceed). rgb[i] from mnenc_free(); on error every rgb[i] is NULL.
 */
int mnenc_decode_batch_rgb8(mnenc_t* ctx, size_t n_images, const uint8_t* const* mn, const size_t* mn_bytes, uint8_t** rgb,
                            size_t* width, size_t* height);
/* Header only (host, no GPU): CompressedImage.cpp:721-736. Any out pointer may be NULL; quant receives 3 * K floats (<= 96). */
int mnenc_stream_info(const uint8_t* mn, size_t mn_bytes, size_t* width, size_t* height, int* K, int* block_size, float* quant);
/*
 * Debug tap (host, no GPU): readCompressedFast's result. lengths[3 * nblocks]; stream_offsets[6K + 1] (stream s =
 * symbols[stream_offsets[s], stream_offsets[s + 1])); symbols may be NULL to obtain the sizes first. Run lengths are
 * expanded and the DC chains undone, as in the reference (:744-781). threads > 1: large Huffman payloads are decoded by that
 * many host threads (the decoder's parallel reader); the result does not depend on it.
 */
int mnenc_debug_parse(const uint8_t* mn, size_t mn_bytes, uint16_t* lengths, size_t lengths_capacity, uint64_t* stream_offsets,
                      uint16_t* symbols, size_t symbols_capacity, int threads);

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
/* Choices of compressed::writeHuffmanOrGolomb (CompressedImage.cpp:359-390) and of the run-length test (:508-516) in the last
 * single-image encode: coded sequences (`lengths` + 6K streams) written with a Huffman table / as Golomb codes, and symbol
 * streams that were run-length coded first. */
int mnenc_last_plan(mnenc_t* ctx, int* n_huffman, int* n_golomb, int* n_rle);

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
 * Stripe protocol: ONE image sharded by block columns over `nranks` contexts (one process per GPU; SURVEY.md §8e,
 * BASELINE.json configs[3]). Replaces compressed::encodeImageFast (CompressedImage.cpp:578-633) for images too large or
 * too urgent for one GPU. The reference walks blocks column-major (CompressedImage.cpp:592-594), so the rank holding
 * block columns [bx0, bx1) owns the matching contiguous slice of every symbol stream. Matching pursuit needs no
 * exchange; the entropy back end (compressed::writeCompressed, CompressedImage.cpp:462-519) needs the five exchanges
 * below, which the CALLER performs (torch.distributed / NCCL in imageexperiments_b200/stripes.py, or any transport):
 *
 *   mnenc_stripe_begin      -> rec1[3K+3]            all-gather  -> all_rec1[nranks][3K+3]
 *   mnenc_stripe_rle        -> rec2[6K][6]           all-gather  -> all_rec2[nranks][6K][6]
 *   mnenc_stripe_histogram  -> device buffers        all-reduce in place over the ranks (as int32): SUM on hist,
 *                                                    MIN on first_pos, MAX on max_sym
 *   mnenc_stripe_plan       -> seq_bits[6K+1]        all-gather  -> all_seq_bits[nranks][6K+1]   (offset scan)
 *   mnenc_stripe_pack       -> device words          reduce SUM to rank 0 in place (as int32; the ranks' bits are
 *                                                    disjoint, so the sum concatenates the per-GPU bitstream segments)
 *   mnenc_stripe_finish     -> rank 0: the complete `.mn` stream (mnenc_free it); other ranks: *out = NULL
 *
 * rgb_stripe points at pixel (x = 8*bx0, y = 0): host memory (on_device = 0, copied inside the call) or memory of the
 * context's GPU (on_device = 1). bx0 == bx1 (a rank without blocks) is allowed. Calls out of order return
 * MNENC_ERR_STATE. The returned device pointers stay valid until the next call on the context; the context's stream is
 * idle when each call returns, so the caller's collectives need no further ordering.
 */
int mnenc_stripe_begin(mnenc_t* ctx, const uint8_t* rgb_stripe, int on_device, size_t width, size_t height, size_t row_stride_bytes,
                       int rank, int nranks, uint32_t bx0, uint32_t bx1, uint32_t* rec1);
int mnenc_stripe_rle(mnenc_t* ctx, const uint32_t* all_rec1, uint32_t* rec2);
int mnenc_stripe_histogram(mnenc_t* ctx, const uint32_t* all_rec2, void** hist_dev, void** first_pos_dev, size_t* hist_elems,
                           void** max_sym_dev, size_t* max_sym_elems);
int mnenc_stripe_plan(mnenc_t* ctx, uint64_t* seq_bits);
int mnenc_stripe_pack(mnenc_t* ctx, const uint64_t* all_seq_bits, void** out_dev, size_t* out_words, size_t* out_bytes);
int mnenc_stripe_finish(mnenc_t* ctx, uint8_t** out, size_t* out_bytes);

/*
 * Stripes with the exchanges INSIDE the library (NCCL over NVLink / NVSwitch, loaded at run time from libnccl.so.2; SURVEY.md
 * §8b "multi-GPU handled inside the context"). Replaces compressed::encodeImageFast (CompressedImage.cpp:578-633) for the
 * reference-side callers (Compression.cpp:117-123) when one image is to be spread over the GPUs of a box.
 *
 * One process per GPU: one rank calls mnenc_comm_unique_id and hands the 128 bytes to the others (any transport), every rank
 * calls mnenc_comm_init_rank on its context, then all ranks call mnenc_encode_striped_rgb8 collectively with their block
 * columns [bx0, bx1) (arguments as mnenc_stripe_begin). Rank 0 receives the complete `.mn` stream (mnenc_free it), the others
 * *out = NULL. The result is byte-identical to mnenc_encode_rgb8 of the whole image on one GPU.
 * Errors are collective: each exchange also carries the result code of every rank's local work since the previous one, so when
 * one rank fails (bad argument, out of memory, ...) ALL ranks return an error from the same point - the failing rank its own
 * code, the others MNENC_ERR_STATE naming it - and the communicator stays usable. Not covered: a rank that dies or whose CUDA
 * context is broken (sticky error); its peers then wait in NCCL until the process group is torn down.
 * mnenc_last_stripe_timing: host milliseconds of the last striped encode on this rank — ms[0] begin (H2D, matching pursuit,
 * scatter), [1] gather of sizes, [2] DC + runs, [3] gather of run records, [4] stitched RLE + local histogram, [5] histogram
 * exchange (sparse all-gather + merge), [6] code tables (dealt over the ranks) + payload bits, [7] offset scan gather,
 * [8] pack, [9] reduce of the segments to rank 0, [10] D2H + headers on rank 0.
 */
int mnenc_comm_unique_id(uint8_t* id128);
int mnenc_comm_init_rank(mnenc_t* ctx, int rank, int nranks, const uint8_t* id128);
int mnenc_comm_destroy(mnenc_t* ctx);
int mnenc_encode_striped_rgb8(mnenc_t* ctx, const uint8_t* rgb_stripe, int on_device, size_t width, size_t height,
                              size_t row_stride_bytes, uint32_t bx0, uint32_t bx1, uint8_t** out, size_t* out_bytes);
int mnenc_last_stripe_timing(mnenc_t* ctx, float* ms16);
/*
 * One process, several GPUs: a context over n_devices GPUs (ncclCommInitAll), dictionary arguments as mnenc_create.
 * mnenc_set_quant reaches every GPU; mnenc_encode_rgb8 on such a context shards the host image by block columns over the GPUs
 * (one host thread per GPU) and returns the same bytes as a single-GPU context. mnenc_destroy releases all of it.
 */
int mnenc_create_multi(mnenc_t** out, const int* devices, int n_devices, int K, int block_size, const float* base_rowmajor, int n_base,
                       const int32_t* detail_rows, const float* detailY, const float* detailU, const float* detailV);
/*
 * The dictionary, computed instead of imported (host only, no GPU): what the dictionary half of
 * createCompressionContextFast builds at start-up (CompressedImage.cpp:317-341 -> BasisSet.cpp:204-297 distinctLineShapes,
 * :430-511 createSegmentDictionaryFast, :618-721 createIntraSegmentDictionaryFast). Outputs in the layout mnenc_create takes:
 * base_rowmajor [n_base][64] (capacity 512 rows), detail_rows [n_base] (capacity 512), detailY/U/V: the groups concatenated in
 * atom order (capacity 512 * 64 rows of 64 each). Shapes, heights and segment atoms equal the reference's bit for bit; the KLT
 * rows come from a pinned solver (cyclic Jacobi in double) instead of Eigen's float eigensolver, whose rounding is
 * build-dependent - so a stream written with this dictionary decodes with this dictionary, not with a reference build's.
 * threads <= 0: all host threads. Only block_size 8.
 */
int mnenc_build_dictionary(int block_size, int threads, float* base_rowmajor, int* n_base, int32_t* detail_rows, float* detailY,
                           float* detailU, float* detailV);
/*
 * Test hook: the next encode on this context fails with MNENC_ERR_INTERNAL right after its histogram kernel has run - an ordinary
 * error return with the histogram bins populated. The encode after that must be unaffected (the context re-clears its tables).
 * In a striped encode (mnenc_encode_striped_rgb8) it fails the histogram phase of THIS rank only: every rank of the group must
 * then return an error from the same exchange (the failing rank MNENC_ERR_INTERNAL, its peers MNENC_ERR_STATE) instead of
 * waiting in NCCL, and the group must encode correctly afterwards.
 */
int mnenc_debug_fail_next_backend(mnenc_t* ctx);
/*
 * Host-logic test hook of the in-library stripe exchange (CPU only): the 193 code-table builds dealt over `nranks` ranks,
 * serialised, and applied on rank 0 (whose peers hold the same entries in another order). Entries of sequence q are
 * [seq_start[q], seq_start[q + 1]). Must give exactly what mnenc_debug_plan_sequence gives per sequence.
 */
int mnenc_debug_plan_sharded(const uint32_t* symbols, const uint32_t* freqs, const uint32_t* first_pos, const uint32_t* seq_start, int n_seq,
                             int nranks, uint32_t* codes, uint8_t* lens, uint64_t* payload_bits, uint32_t* golomb_m);
/*
 * mnenc_create or mnenc_create_multi, chosen by the environment - what a host with no device argument of its own calls (the shim
 * behind encodeImageFast, CompressedImage.cpp:578; the Compression front end): MNENC_DEVICES="0,1,2" or "all" names the GPUs of a
 * multi-GPU context, otherwise MNENC_DEVICE (default 0) names the one GPU.
 */
int mnenc_create_auto(mnenc_t** out, int K, int block_size, const float* base_rowmajor, int n_base, const int32_t* detail_rows,
                      const float* detailY, const float* detailU, const float* detailV);

/*
 * Host-logic test hooks of the stripe path (used by the CPU-only world_size-2/3 gloo tests): run stitching of
 * huffman::runLengthEncode (Huffman.cpp:246-279) across ranks. records = [nranks][n_streams][6] u32
 * {n, first value, first run length, last value, last run length, RLE symbols of the interior runs} as all-gathered
 * by the ranks (rec2 above). mnenc_debug_stitch_runs, for `rank`: carry_in[s] / extend_out[s] = elements of its
 * first / last run that live on earlier / later ranks. mnenc_debug_rle_chunk_size = RLE symbols a rank emits for `len`
 * elements of a run with `carry` elements before them and `total` elements overall (chunks of 32 769 counted from the
 * run's global start). mnenc_debug_plan_stripes = what every rank derives from the records: emitted[nranks][n_streams]
 * RLE symbols per rank, and per stream the whole-stream sizes and the RLE-or-raw decision of
 * compressed::writeCompressed (CompressedImage.cpp:508-516).
 */
int mnenc_debug_stitch_runs(const uint32_t* records, int nranks, int n_streams, int rank, uint32_t* carry_in, uint32_t* extend_out);
uint32_t mnenc_debug_rle_chunk_size(uint32_t carry, uint32_t len, uint64_t total);
int mnenc_debug_plan_stripes(const uint32_t* records, int nranks, int n_streams, uint32_t* emitted, uint8_t* use_rle, uint64_t* raw_len,
                             uint64_t* rle_len);

/*
 * The symbol-stream scatter for callers that cannot allocate GPU memory. This entry point has no counterpart in the
 * reference; it exposes one step of the encoder (the stream scatter of encodeImageFast) in a form that never calls
 * cudaMalloc, cudaFree or any other allocating routine.
 *
 * Who needs it: a plugin inside a host program that owns all of HBM and hands out every byte itself (a foreign cudaMalloc
 * there either fails or counts as a leak), code on a latency-critical path where an allocator call would synchronise the
 * device, and processes that gave the whole GPU to one pool at start-up. All other GPU entry points of this header grow
 * per-context device buffers when they need them and are therefore unusable for such callers.
 *
 * Inputs and outputs are the caller's buffers, all in memory of the CURRENT device (cudaSetDevice) except stream_offsets:
 *   counts_dev[3 * n_blocks]              counts, layout of mnenc_get_symbols; every value must be <= K
 *   pairs_dev[3 * n_blocks * K * 2]       BasisChoice pairs, layout of mnenc_get_symbols
 *   streams_dev[streams_capacity]         out: the 6K streams back to back, u16 symbols (2 * sum(counts) <= 6 * K * n_blocks)
 *   stream_offsets[6K + 1]                out, HOST memory: stream s = streams_dev[stream_offsets[s], stream_offsets[s + 1]),
 *                                         s = 2 * (layer * K + depth) + {0: deltaId, 1: intCoeff}; same numbering and same
 *                                         contents as mnenc_debug_parse reports for the finished `.mn` file
 *   cuda_stream                           a cudaStream_t (NULL = default stream); the call returns after the work is complete
 * The per-tile offsets the kernels need in between live in a fixed 1.5 MB __device__ array inside the library (present on
 * every GPU that loads the library, until the process exits). Hence the limit: n_blocks <= 1048576 (8192 x 8192 pixels);
 * more gives MNENC_ERR_ARG, as does a streams_capacity below stream_offsets[6K] (stream_offsets is still filled in, so the
 * caller learns the size). Host or foreign-GPU pointers are refused with MNENC_ERR_ARG.
 * All calls on one GPU share that array, and it keeps the previous call's contents (each call overwrites what it reads, so
 * nothing needs clearing): calls that overlap on the same GPU - from two threads or two streams - race and must be
 * serialised by the caller. The first call on each device looks up the array's device address and makes the CUDA runtime
 * load the library's kernels there; make that first call during start-up, before any stream capture begins.
 */
int mnenc_scatter_streams_static(const uint16_t* counts_dev, const uint16_t* pairs_dev, int K, size_t n_blocks, uint16_t* streams_dev,
                                 size_t streams_capacity, uint64_t* stream_offsets, void* cuda_stream);

#ifdef __cplusplus
}
#endif
#endif /* MNENC_H_ */
