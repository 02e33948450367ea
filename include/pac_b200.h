/* pac_b200.h -- C ABI of libpac_b200.so, the B200 (sm_100a) implementation of the
 * Music-422 perceptual coder hot path (Abhipray/audio-codec, coder/*.py).
 *
 * The reference is pure Python with no FFI of its own: its boundary is the
 * Python function surface.  Every entry point below names the
 * reference function(s) it replaces; audio_codec_b200/*.py binds them with
 * ctypes under the reference's own module / function names.
 *
 * Conventions: plain pointers and sizes only; `stream` is a cudaStream_t passed
 * as void* (NULL = default stream); buffers named d_* are DEVICE pointers owned
 * by the caller, h_* are HOST pointers; every function returns 0 or a negative
 * pac_status; nothing throws.  One context per GPU; the codec configuration of a
 * context is immutable after pac_create.
 */
#ifndef PAC_B200_H
#define PAC_B200_H
#ifdef __cplusplus
extern "C" {
#endif

typedef struct pac_ctx pac_ctx;

enum pac_dtype { PAC_F32 = 0, PAC_F64 = 1 };
enum pac_status {
    PAC_OK = 0,
    PAC_ERR_ARG = -1,        /* bad argument */
    PAC_ERR_ALLOC = -2,      /* device/host allocation failed */
    PAC_ERR_UNSUPPORTED = -3,/* e.g. block floating point + SBR (crashes in the reference too) */
    PAC_ERR_SOFT = -4,       /* per-frame soft error flagged on device (see pac_encode flags) */
    PAC_ERR_TOO_LARGE = -5   /* a *_static (no-allocation) entry point was asked for more than its fixed scratch holds */
    /* <= -1000: -(cudaError_t) - 1000 */
};
/* window shapes: coder/window.py:14 SineWindow, :61 StartWindow, :73 StopWindow,
 * :82 StartStopWindow, :45 KBDWindow (a Kaiser window, alpha = 4), none */
enum pac_window { PAC_W_SINE = 0, PAC_W_START = 1, PAC_W_STOP = 2, PAC_W_STARTSTOP = 3, PAC_W_KAISER = 4, PAC_W_NONE = 5,
                  PAC_W_HANN = 6 /* coder/window.py:29 HanningWindow; pac_window only */ };

/* The fields of the reference's CodingParams bag the path reads
 * (coder/pacfile.py:157-174, :699-707). */
typedef struct pac_config {
    int sample_rate;            /* Hz */
    int n_channels;
    int n_mdct_lines;           /* 1024 (long); short blocks are always 128 */
    int n_scale_bits;           /* 4 */
    int n_mant_size_bits;       /* 12 */
    int use_sbr;                /* coder/pacfile.py:704 */
    int use_vq;                 /* coder/pacfile.py:705 */
    double target_bits_per_sample; /* rate_kbps / (fs/1000), per channel (pacfile.py:702) */
    int precision;              /* PAC_F64: the reference's fp64 arithmetic, bit-exact .pac / PCM (default);
                                 * PAC_F32: fast path - transforms and psychoacoustic model in fp32 (MDCT and
                                 * SMRs within 1e-5, decoded PCM within 1 LSB); quantisers stay fp64 + integer */
} pac_config;

/* pac_create returns PAC_ERR_UNSUPPORTED when 2^n_scale_bits - 1 + max mantissa bits exceeds 62 (the
 * device quantisers keep magnitudes in 64-bit integers; the reference uses Python bigints there):
 * n_scale_bits <= 5 with the default 16-bit mantissa cap. */
int pac_create(int device, const pac_config* cfg, pac_ctx** out);
void pac_destroy(pac_ctx* ctx);
/* Tuning knobs / test hooks; no result depends on them.  BATCH_CB: channel-blocks per launch of the
 * thread-per-band kernels (default 131072; small values exercise the batch loop).  SMEM_PAD: extra
 * dynamic shared memory of the frame kernels (occupancy experiments).  HOST_CHUNK_MB: smallest chunk
 * (MB of PCM) of the pac_encode_host / pac_decode_host copy pipeline (default 24). */
enum pac_option { PAC_OPT_BATCH_CB = 0, PAC_OPT_SMEM_PAD = 1, PAC_OPT_HOST_CHUNK_MB = 2,
                  PAC_OPT_VARIANT = 3 /* bit flags: run the previous form of a stage (same results; A/B timing, tests) */ };
int pac_set_option(pac_ctx* ctx, int option, double value);
const char* pac_strerror(int status);
/* band layout as ScaleFactorBands(AssignMDCTLinesFromFreqLimits(..)) builds it
 * (coder/psychoac.py:106-160); which = 0 long, 1 short.  Returns nBands. */
int pac_get_bands(const pac_ctx* ctx, int which, int* n_lines, int* omitted);

/* mdct.MDCT / window.* (coder/mdct.py:63-69, coder/codec.py:303-305), batched.
 * d_x: n_rows rows of (n_frames+1)*N/2 samples, row pitch row_stride elements;
 * frame f of a row is samples [f*N/2, f*N/2+N).  d_coefs: [n_rows][n_frames][N/2].
 * N is 2048 or 256. */
int pac_mdct_forward(pac_ctx* ctx, int dtype, const void* d_x, long long row_stride, int n_rows,
                     int n_frames, int N, int window, void* d_coefs, void* stream);
/* mdct.IMDCT + synthesis window + overlap-add (coder/mdct.py:56-62,
 * coder/codec.py:88-89, coder/pacfile.py:289-295).  d_y: n_rows rows of
 * (n_frames+1)*N/2 samples. */
int pac_imdct_ola(pac_ctx* ctx, int dtype, const void* d_coefs, int n_rows, int n_frames, int N,
                  int window, void* d_y, long long row_stride, void* stream);

/* mdct.MDCT / mdct.IMDCT / mdct.MDCTslow for ANY block split a, b with a + b even (coder/mdct.py:11-76;
 * the reference's own self-test uses a = b = 4 and a = b = 10, :86-107).  Direct O(N^2 / 2) evaluation, one
 * frame per row: d_in [n_rows][a+b] -> d_out [n_rows][(a+b)/2], or the reverse when inverse != 0.  The
 * codec's sizes (a == b, N = 2048 / 256) go through pac_mdct_forward / pac_imdct_ola above. */
int pac_mdct_direct(pac_ctx* ctx, int dtype, const void* d_in, long long n_rows, int a, int b, int inverse,
                    void* d_out, void* stream);

/* ---- whole-stream encode: replaces the "Encode" leg of pacfile.encode_decode_test
 * (coder/pacfile.py:674-756): PCMFile.ReadDataBlock's int16 -> float mapping
 * (coder/pcmfile.py:89-98), parTransientDetect (coder/detect_transients.py:5-23),
 * PACFile.WriteDataBlock (coder/pacfile.py:449-610) -> codec.Encode / Encode_SBR
 * (coder/codec.py:225-555: window, MDCT, ScaleFactor, CalcSMRs, BitAlloc[_SBR],
 * quantize_gain_shape | ScaleFactor+vMantissa), PackedBits.WriteBits, the length
 * prefixes, WriteFileHeader and Close().
 *
 * d_pcm: int16 planar, block-padded: stream s owns h_nblocks[s] blocks of 1024
 * samples per channel at element offset (sum_{t<s} nblocks[t] * C + c * nblocks[s]) * 1024.
 * h_hdr_samples[s]: sample count the header's numSamples is derived from.
 * d_out: caller buffer of out_capacity bytes (>= pac_encode_bound()); stream s'
 * .pac image is d_out[h_stream_off[s] .. h_stream_off[s+1]).  *h_err receives
 * the OR of per-frame soft-error flags (inputs whose behaviour is undefined in
 * the reference); the call then returns PAC_ERR_SOFT.  Synchronises `stream`. */
int pac_header_bytes(const pac_ctx* ctx);
long long pac_encode_bound(const pac_ctx* ctx, const int* h_nblocks, int n_streams);
int pac_encode_streams(pac_ctx* ctx, const short* d_pcm, const int* h_nblocks, const int* h_hdr_samples,
                       int n_streams, unsigned char* d_out, long long out_capacity,
                       long long* h_stream_off, int* h_err, void* stream);

/* ---- whole-stream decode: replaces the "Decode" leg of pacfile.encode_decode_test:
 * PACFile.ReadDataBlock / getDecodedBlock (coder/pacfile.py:177-298: length-prefix
 * walk, flag bits, scale / bit-allocation fields), codec.Decode / Decode_SBR
 * (coder/codec.py:47-222: dequantize_gain_shape | vDequantize, SBR envelope +
 * transposition, IMDCT, synthesis window), the short-block and block overlap-add,
 * and PCMFile.WriteDataBlock's 16-bit quantiser (coder/pcmfile.py:118-147).
 *
 * d_pac: the .pac images back to back, stream s at [h_stream_off[s], h_stream_off[s+1]);
 * every stream must have been coded with this context's configuration (header
 * fields are the caller's to check).  Output: int16 planar; stream s channel c is
 * d_pcm[h_pcm_off[s] + c*n_s .. + n_s), n_s = (h_pcm_off[s+1]-h_pcm_off[s])/C samples
 * (all written blocks plus the final overlap flush, i.e. what the reference's WAV
 * writer receives).  pac_decode_plan only walks the length prefixes and fills
 * h_pcm_off so the caller can size d_pcm.  Both synchronise `stream`. */
int pac_decode_plan(pac_ctx* ctx, const unsigned char* d_pac, const long long* h_stream_off, int n_streams,
                    long long* h_pcm_off, void* stream);
int pac_decode_streams(pac_ctx* ctx, const unsigned char* d_pac, const long long* h_stream_off, int n_streams,
                       short* d_pcm, long long pcm_capacity, long long* h_pcm_off, int* h_err, void* stream);

/* Host-buffer forms of the two calls above (ordinary or pinned host memory in
 * and out; H2D copy, device pipeline, D2H copy on `stream`).  This is what a host
 * driver standing in for pacfile.encode_decode_test calls.  pac_decode_host with
 * h_pcm == NULL only fills h_pcm_off. */
int pac_encode_host(pac_ctx* ctx, const short* h_pcm, const int* h_nblocks, const int* h_hdr_samples,
                    int n_streams, unsigned char* h_out, long long out_capacity, long long* h_stream_off,
                    int* h_err, void* stream);
int pac_decode_host(pac_ctx* ctx, const unsigned char* h_pac, const long long* h_stream_off, int n_streams,
                    short* h_pcm, long long pcm_capacity, long long* h_pcm_off, int* h_err, void* stream);

/* Measurement aid (no reference counterpart): device milliseconds of the dominant
 * stage (k_encode .. k_pack / k_hdr .. k_decode) of the last encode / decode call on this context,
 * from CUDA events recorded on the caller's stream; -1 if none yet. */
int pac_last_kernel_ms(pac_ctx* ctx, float* enc_ms, float* dec_ms);
/* The same per kernel: device milliseconds summed over the launches (one per batch of channel-blocks)
 * of kind 0 k_encode, 1 k_pvq, 2 k_dequant, 3 k_decode in the last completed encode / decode call,
 * and how many launches that was. */
int pac_kernel_stats(pac_ctx* ctx, int kind, float* ms, int* launches);
/* Kernels launched by the library on this context since pac_create. */
long long pac_kernel_launches(pac_ctx* ctx);


/* ======================================================================================
 * Per-function ops: the reference's Python function surface, one batched call each.
 * All pointers are DEVICE pointers unless named h_*; rows are contiguous.
 * ====================================================================================== */

/* window.SineWindow / HanningWindow / KBDWindow(alpha) / StartWindow / StopWindow /
 * StartStopWindow (coder/window.py:14-92): d_out[r][n] = w[n] * d_x[r][n], any N; n_short is
 * N_short of the Edler shapes (N_long = N); alpha only for PAC_W_KAISER. */
int pac_window(pac_ctx* ctx, int dtype, const void* d_x, long long n_rows, int N, int window, int n_short,
               double alpha, void* d_out, void* stream);

/* detect_transients.parTransientDetect(block, thresh, axis=1) (coder/detect_transients.py:5-23)
 * on n_blocks blocks of [n_channels][len] doubles; d_flags[b] = 1 if transient. */
int pac_transient_detect(pac_ctx* ctx, const double* d_blocks, long long n_blocks, int n_channels, int len,
                         double thresh, unsigned char* d_flags, void* stream);

/* psychoac.CalcSMRs / getMaskedThreshold (coder/psychoac.py:163-291) for n_frames frames of
 * N (2048 | 256) samples at the context's sample rate.  d_mdct: MDCT lines scaled by 2^d_scale[f].
 * Bands as ScaleFactorBands gives them (lowerLine, nLines).  d_smr [n][n_bands] and/or
 * d_mask [n][N/2] (masked threshold in dB SPL at the MDCT lines); either may be NULL.
 * dtype: arithmetic of the model (PAC_F64 = the reference's; PAC_F32 = fast path, SMRs within
 * 1e-5 of their range); the buffers are fp64 either way. */
int pac_calc_smr(pac_ctx* ctx, int dtype, const double* d_data, const double* d_mdct, const int* d_scale, long long n_frames,
                 int N, int n_bands, const int* h_lower, const int* h_nlines, double* d_smr, double* d_mask,
                 void* stream);

/* psychoac.SPL (array form) / Intensity / Bark / Thresh (coder/psychoac.py:10-48), elementwise;
 * op: 0 SPL, 1 Intensity, 2 Bark, 3 Thresh; 4 / 5: SPL / Bark in the straight-line form the stream encoder
 * runs (same bits as 0 / 2 for non-negative finite input; tests/test_ops_gpu.py). */
int pac_psy_map(pac_ctx* ctx, int op, const double* d_x, long long n, double* d_out, void* stream);

/* PCMFile.ReadDataBlock's sample mapping (coder/pcmfile.py:89-98): int16 code -> sign * 2|c| / 65535
 * (-32768 -> 0.0), elementwise, as the stream encoder applies it. */
int pac_pcm_to_float(pac_ctx* ctx, const short* d_pcm, long long n, double* d_out, void* stream);

/* Self-test (no reference counterpart): the quantiser kernels form x / d for many x and one d (a vector norm,
 * sqrt(2)) from one correctly rounded reciprocal and a residual correction (pac_div, csrc/pac_common.cuh), which
 * equals the IEEE quotient.  Runs n_pairs pseudo-random pairs on the device and returns how many quotients differ
 * from the division instruction (expected 0), or a negative pac_status. */
long long pac_selftest_division(pac_ctx* ctx, long long n_pairs, unsigned long long seed, void* stream);

/* bitalloc.BitAlloc(bitBudget, maxMantBits, nBands, nLines, SMR) (coder/bitalloc.py:62-121);
 * BitAlloc_SBR is the same call with nLines[omitted] = 1 (coder/bitalloc.py:123-145). */
int pac_bit_alloc(pac_ctx* ctx, const double* d_budget, int max_mant_bits, int n_bands, const int* h_nlines,
                  const double* d_smr, long long n_frames, int* d_bits, void* stream);

/* quantize.vQuantizeUniform / vDequantizeUniform / ScaleFactor / vMantissa / vDequantize
 * (coder/quantize.py:61-125, :229-274), elementwise over n values. */
int pac_quantize_uniform(pac_ctx* ctx, const double* d_x, long long n, int n_bits, long long* d_code, void* stream);
int pac_dequantize_uniform(pac_ctx* ctx, const long long* d_code, long long n, int n_bits, double* d_x, void* stream);
int pac_scale_factor(pac_ctx* ctx, const double* d_x, long long n, int n_scale_bits, int n_mant_bits,
                     long long* d_scale, void* stream);
int pac_mantissa(pac_ctx* ctx, const double* d_x, long long n, int scale, int n_scale_bits, int n_mant_bits,
                 long long* d_mant, void* stream);
int pac_dequantize(pac_ctx* ctx, int scale, const long long* d_mant, long long n, int n_scale_bits, int n_mant_bits,
                   double* d_x, void* stream);

/* gain_shape_quantize.quantize_gain_shape(x, bit_alloc, k_fine=0) (coder/gain_shape_quantize.py:476-512)
 * for n_vec vectors of length L with budgets d_bit_alloc[v] (<= max_bit_alloc): the (index, width)
 * fields in stream order go to d_idx/d_bits [v][max_fields], their number to d_nfields[v].
 * Synchronises; *h_err receives soft-error flags (inputs undefined in the reference). */
int pac_gain_shape_quantize(pac_ctx* ctx, const double* d_x, int L, long long n_vec, const int* d_bit_alloc,
                            int max_bit_alloc, unsigned long long* d_idx, int* d_bits, int* d_nfields,
                            int max_fields, int* h_err, void* stream);
/* gain_shape_quantize.dequantize_gain_shape(pb, bit_alloc, L, k_fine=0) (:515-538): vector v is
 * parsed from the shared byte buffer starting at bit d_bitpos[v]; the cursor after its fields is
 * returned in d_bitpos_out[v]. */
int pac_gain_shape_dequantize(pac_ctx* ctx, const unsigned char* d_bytes, long long n_bytes, const long long* d_bitpos,
                              const int* d_bit_alloc, int L, long long n_vec, double* d_out, long long* d_bitpos_out,
                              int* h_err, void* stream);
/* NO COUNTERPART IN THE REFERENCE.  pac_gain_shape_dequantize for
 * callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid
 * it): same arguments, same results, but it calls no cudaMalloc / cudaFree or any allocating routine.  Its
 * workspaces (3 x pow2ceil(L) doubles per vector) and its soft-error word live in one file-scope device array of
 * 4 MiB + 256 B that every GPU loading the library holds until exit; its device address is resolved at
 * pac_create, once per device.  Limit: n_vec <= pac_gain_shape_dequantize_static_max_vectors(L)
 * (= 4 MiB / (24 * pow2ceil(L)): 341 vectors at L = 363, 174 762 at L = 1); beyond it the call returns
 * PAC_ERR_TOO_LARGE and does nothing.  All calls on a device share the array: calls that overlap on the GPU
 * (other streams, threads or contexts) race.  Synchronises `stream`. */
long long pac_gain_shape_dequantize_static_max_vectors(int L);
int pac_gain_shape_dequantize_static(pac_ctx* ctx, const unsigned char* d_bytes, long long n_bytes,
                                     const long long* d_bitpos, const int* d_bit_alloc, int L, long long n_vec,
                                     double* d_out, long long* d_bitpos_out, int* h_err, void* stream);

/* bitpack.PackedBits.WriteBits / ReadBits (coder/bitpack.py:37-175): row r packs its
 * d_count[r] fields (value, width <= 64) MSB-first into row_bytes bytes (multiple of 4). */
int pac_pack_bits(pac_ctx* ctx, const unsigned long long* d_val, const int* d_width, const int* d_count,
                  int max_fields, long long n_rows, int row_bytes, unsigned char* d_out, int* h_err, void* stream);
int pac_unpack_bits(pac_ctx* ctx, const unsigned char* d_in, int row_bytes, const long long* d_start_bit,
                    const int* d_width, const int* d_count, int max_fields, long long n_rows,
                    unsigned long long* d_val, void* stream);

/* PACFile.WriteDataBlock for n channel-blocks (coder/pacfile.py:449-610): d_frames [n][2048] =
 * priorBlock || data as signed fractions, d_flags[n] = last | cur<<1 | next<<2.  d_rec
 * [n][pac_record_stride()] receives the packed payload (flag bits, frame(s), zero padding),
 * d_rec_bytes[n] the declared size (getNumBytesNeeded), d_zero_sub[n] = 1 if a short sub-frame
 * is all zero (the block is then dropped by the reference, :503-506). */
int pac_record_stride(void);
int pac_encode_blocks(pac_ctx* ctx, const double* d_frames, const unsigned char* d_flags, long long n,
                      unsigned char* d_rec, int* d_rec_bytes, unsigned char* d_zero_sub, int* h_err, void* stream);
/* codec.EncodeSingleChannel / EncodeSingleChannel_SBR (coder/codec.py:266-380, :426-555) for n
 * frames of N samples: overall scale, bit allocation [n][32], BFP scale factors [n][32], bits
 * emitted per band [n][32], the frame's packed fields (d_rec, d_rec_bits = bit count) and, for
 * VQ, the per-band field log d_log_val/width [n][32][log_cap], d_log_count [n][32]. */
int pac_encode_frames(pac_ctx* ctx, const double* d_data, int N, const unsigned char* d_flags, int sbr, long long n,
                      int* d_scale, int* d_bit_alloc, int* d_scale_factor, int* d_band_bits, unsigned char* d_rec,
                      int* d_rec_bits, unsigned long long* d_log_val, int* d_log_width, int* d_log_count, int log_cap,
                      int* h_err, void* stream);
/* PACFile.ReadDataBlock's per-channel work (coder/pacfile.py:256-287): payload records ->
 * 2048 windowed samples each, before the block overlap-add. */
int pac_decode_blocks(pac_ctx* ctx, const unsigned char* d_rec, int rec_stride, const int* d_rec_bytes, long long n,
                      double* d_out, int* h_err, void* stream);
/* codec.Decode / Decode_SBR (coder/codec.py:47-222): one frame per record, fields from bit 0
 * (overall scale, bit allocations, [scale factors,] band payloads) -> N windowed samples;
 * d_bits_consumed[n] = bit cursor after the frame. */
int pac_decode_frames(pac_ctx* ctx, const unsigned char* d_rec, int rec_stride, const int* d_rec_bytes, int N,
                      const unsigned char* d_flags, long long n, double* d_out, int* d_bits_consumed, int* h_err,
                      void* stream);

#ifdef __cplusplus
}
#endif
#endif
