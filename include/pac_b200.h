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
 * pac_status; nothing throws.  One context per GPU; a context is immutable after
 * pac_create and calls on distinct streams/buffers are re-entrant.
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
    PAC_ERR_SOFT = -4        /* per-frame soft error flagged on device (see pac_encode flags) */
    /* <= -1000: -(cudaError_t) - 1000 */
};
/* window shapes: coder/window.py:14 SineWindow, :61 StartWindow, :73 StopWindow,
 * :82 StartStopWindow, :45 KBDWindow (a Kaiser window, alpha = 4), none */
enum pac_window { PAC_W_SINE = 0, PAC_W_START = 1, PAC_W_STOP = 2, PAC_W_STARTSTOP = 3, PAC_W_KAISER = 4, PAC_W_NONE = 5 };

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
} pac_config;

int pac_create(int device, const pac_config* cfg, pac_ctx** out);
void pac_destroy(pac_ctx* ctx);
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
 * kernel (k_encode / k_decode) of the last encode / decode call on this context,
 * from CUDA events recorded on the caller's stream; -1 if none yet. */
int pac_last_kernel_ms(pac_ctx* ctx, float* enc_ms, float* dec_ms);

#ifdef __cplusplus
}
#endif
#endif
