// ggml-b200.h -- C-ABI of the B200 (sm_100a) hot path.
//
// Two surfaces, one shared library (libggml-b200.so):
//
//  (1) The ggml backend plugin entry points, exactly what the reference's registry dlsym()s
//      (reference: ggml/src/ggml-backend-reg.cpp:255-272, ggml/src/ggml-backend-impl.h:215-251):
//          ggml_backend_reg_t ggml_backend_init(void);
//          int                ggml_backend_score(void);
//      plus the in-tree registration symbol ggml_backend_b200_reg() (registry constructor,
//      ggml/src/ggml-backend-reg.cpp:172-212).  Through them llama.cpp, llama-bench and
//      test-backend-ops reach supports_op / graph_compute with no other glue.
//
//  (2) A flat operator API over device pointers (b200_*), one call per reference operator:
//          b200_quantize_q8_1   <- quantize_row_q8_1        ggml/src/ggml-cpu/arch/x86/quants.c:388-490
//          b200_mul_mat         <- ggml_compute_forward_mul_mat      ggml/src/ggml-cpu/ggml-cpu.c:1226-1418
//          b200_mul_mat_id      <- ggml_compute_forward_mul_mat_id   ggml/src/ggml-cpu/ggml-cpu.c:1500-1700
//          b200_flash_attn_ext  <- ggml_compute_forward_flash_attn_ext ggml/src/ggml-cpu/ops.cpp:8047-8374
//      The plugin's graph_compute is a thin translation ggml_tensor -> b200_tensor -> these
//      calls; parity tests and bench.py bind the same functions through ctypes.
//
// Tensors follow ggml conventions: ne[0] is the fastest dimension, nb[] are byte strides,
// quantized types are stored as GGUF blocks (block_q4_0 / block_q8_0 / block_q4_K / block_q5_K /
// block_q6_K, reference ggml/src/ggml-common.h:170-344).
// All functions return 0 on success and a negative b200_status on failure; b200_last_error()
// gives the message.  There is no CPU fallback: without a usable sm_100 device every call fails.
#pragma once
#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#if defined(_WIN32)
#  define B200_API __declspec(dllexport)
#else
#  define B200_API __attribute__((visibility("default")))
#endif

enum b200_status {
    B200_OK               =  0,
    B200_ERR_NO_DEVICE    = -1,   // no CUDA device / not sm_100
    B200_ERR_CUDA         = -2,   // a CUDA runtime call failed
    B200_ERR_UNSUPPORTED  = -3,   // type / shape combination outside the hot path
    B200_ERR_INVALID      = -4,   // malformed arguments
    B200_ERR_ALLOC        = -5,
};

// same numeric values as enum ggml_type (ggml/include/ggml.h:389-431)
enum b200_type {
    B200_TYPE_F32 = 0, B200_TYPE_F16 = 1, B200_TYPE_Q4_0 = 2, B200_TYPE_Q8_0 = 8, B200_TYPE_Q8_1 = 9,
    B200_TYPE_Q4_K = 12, B200_TYPE_Q5_K = 13, B200_TYPE_Q6_K = 14, B200_TYPE_I32 = 26, B200_TYPE_BF16 = 30,
};

// q8_1 rounding variants
enum b200_q8_1_variant {
    B200_Q8_1_X86  = 0,  // ggml-cpu x86 AVX2: id = 127/amax, round-half-even, s = fp16(d*sum q)   (default, what libggml-cpu runs)
    B200_Q8_1_REF  = 1,  // quantize_row_q8_1_ref: id = 1/(amax/127), roundf (half away), s = fp16(sum*d)
};

// device-resident strided view, ggml layout
typedef struct b200_tensor {
    void *  data;
    int32_t type;       // enum b200_type
    int32_t flags;      // bit 0: data is in the B200 repacked (SoA) weight layout
    int64_t ne[4];
    size_t  nb[4];
} b200_tensor;
#define B200_TENSOR_REPACKED 1

typedef struct b200_ctx b200_ctx;   // one device + one stream + scratch arena

// ---- device / context -------------------------------------------------------------------
B200_API int          b200_device_count(void);
B200_API const char * b200_last_error(void);
B200_API b200_ctx *   b200_ctx_create(int device);             // NULL on failure
B200_API void         b200_ctx_destroy(b200_ctx * ctx);
B200_API int          b200_ctx_device(const b200_ctx * ctx);
B200_API void *       b200_ctx_stream(const b200_ctx * ctx);   // cudaStream_t
B200_API int          b200_ctx_synchronize(b200_ctx * ctx);
B200_API uint64_t     b200_ctx_launch_count(const b200_ctx * ctx);   // kernels launched so far through this ctx
B200_API uint64_t     b200_ctx_scratch_generation(const b200_ctx * ctx);   // bumped when a scratch block moves: CUDA graphs recorded before are stale
// The ctx keeps the q8_1 image of the last quantized activation tensor so that matmuls sharing src1 (wq/wk/wv, gate/up)
// quantize it once.  Writes made through this API invalidate it; call this after writing activations by other means.
B200_API int          b200_ctx_invalidate(b200_ctx * ctx);
/* tell the context that [ptr, ptr + nbytes) was written by something other than this API (a glue kernel of the caller):
 * cached activation images / mask tile flags derived from that memory are dropped */
B200_API int          b200_ctx_note_write(b200_ctx * ctx, const void * ptr, size_t nbytes);

// ---- measurement helpers (bench.py): CUDA events on the ctx stream, per-kernel-kind event profile, CUDA graphs ----
// kernel kinds: 0 quantize_q8_1, 1 gemv, 2 gemv_id, 3 gemm, 4 fattn, 5 fattn_combine, 6 matmul_float, 7 glue, 8 mmid_sort, 9 convert
B200_API uint64_t b200_ctx_launch_count_kind(const b200_ctx * ctx, int kind);
B200_API int      b200_ctx_profile(b200_ctx * ctx, int enable);      // (re)start or stop bracketing every launch with events
B200_API int      b200_ctx_profile_get(b200_ctx * ctx, int kind, double * ms_total, uint64_t * launches);   // synchronizes
B200_API void *   b200_event_create(void);
B200_API int      b200_event_destroy(void * ev);
B200_API int      b200_event_record(b200_ctx * ctx, void * ev);
B200_API int      b200_event_elapsed_ms(void * ev0, void * ev1, float * ms);     // waits for ev1
// capture everything issued on the ctx stream between begin/end into a graph; launch replays it on the ctx stream
/* measurement only: every CTA of the decode GEMV kernels appends a 32 x u64 timeline record to `buf` (device memory,
 * (1 + cap) * 256 bytes, word 0 = record counter, zeroed by the caller); buf = NULL switches it off */
B200_API int      b200_debug_trace(b200_ctx * ctx, void * buf, int cap);
B200_API int      b200_graph_begin(b200_ctx * ctx);
B200_API void *   b200_graph_end(b200_ctx * ctx);                    // cudaGraphExec_t, NULL on failure
B200_API int      b200_graph_launch(b200_ctx * ctx, void * graph_exec);
B200_API int      b200_graph_destroy(void * graph_exec);

// ---- memory ------------------------------------------------------------------------------
B200_API void * b200_malloc(b200_ctx * ctx, size_t nbytes);    // cudaMalloc on ctx's device
B200_API int    b200_free(b200_ctx * ctx, void * dptr);
B200_API void * b200_host_alloc(size_t nbytes);                // pinned host memory
B200_API int    b200_host_free(void * hptr);
B200_API int    b200_memcpy_h2d(b200_ctx * ctx, void * dst, const void * src, size_t nbytes);   // async on ctx stream
B200_API int    b200_memcpy_d2h(b200_ctx * ctx, void * dst, const void * src, size_t nbytes);   // async on ctx stream
B200_API int    b200_memset(b200_ctx * ctx, void * dst, int value, size_t nbytes);

// ---- operators ----------------------------------------------------------------------------
// f32 rows [nrows][k] (row stride x_row_stride bytes) -> block_q8_1 [nrows][k/32], k % 32 == 0.
B200_API int b200_quantize_q8_1(b200_ctx * ctx, const float * x, size_t x_row_stride, void * y,
                                int64_t k, int64_t nrows, int variant);

// dst[ne01, ne11, ne12, ne13] (f32) = w[ne00, ne01, ne02, ne03]^T * x[ne00, ne11, ne12, ne13] (f32),
// w broadcast over dims 2,3 (ggml/src/ggml.c:3168-3203).  w: f32/f16/bf16/q4_0/q8_0/q4_K/q5_K/q6_K.
B200_API int b200_mul_mat(b200_ctx * ctx, const b200_tensor * w, const b200_tensor * x, b200_tensor * dst);

// n MUL_MATs sharing the activation tensor x (wq/wk/wv, ffn_gate/ffn_up in src/models/llama.cpp:40-75): same result as n
// b200_mul_mat calls; served by one launch of the streaming GEMV when x is a single column and the weights allow it.
B200_API int b200_mul_mat_multi(b200_ctx * ctx, const b200_tensor * const * w, int n, const b200_tensor * x, b200_tensor * const * dst);
/* gate = wg x, up = wu x, glu = swiglu(gate, up), one activation column, one launch: ffn_gate + ffn_up + GGML_OP_GLU (SWIGLU, split;
 * ggml-cpu/ops.cpp ggml_compute_forward_glu) of a llama layer (src/llama-graph.cpp build_ffn, LLM_FFN_SILU / LLM_FFN_PAR).
 * B200_ERR_UNSUPPORTED unless both weights are raw q4_K served by the tensor-core GEMV: use b200_mul_mat_multi + the GLU op then. */
B200_API int b200_mul_mat_gate_up_swiglu(b200_ctx * ctx, const b200_tensor * wg, const b200_tensor * wu, const b200_tensor * x,
                                         b200_tensor * gate, b200_tensor * up, b200_tensor * glu);
/* b200_mul_mat for callers that may not allocate (CUDA graph capture, tight serving loops, harnesses that forbid cudaMalloc).  NO
 * counterpart in the reference.  Same result as b200_mul_mat's general small-batch path
 * (q8_1 image + one-warp-per-row GEMV), but the q8_1 image of the activations lives in a static device array of
 * B200_NOALLOC_SCRATCH_BYTES that every GPU loading the library holds until exit; the call performs no cudaMalloc / cudaFree / event
 * creation.  Limits: quantized w (q4_0 / q8_0 / q4_K / q5_K / q6_K), x->ne[1] <= 8 columns, and K * ncol * 9 / 8 bytes
 * (ncol = ne[1] * ne[2] * ne[3] of x) must fit the array, else B200_ERR_INVALID with a message naming both sizes.  All calls on a device
 * share the array: calls that overlap on the GPU (two contexts on one device) race; calls on one context are stream-ordered. */
#define B200_NOALLOC_SCRATCH_BYTES (1u << 20)
B200_API int    b200_mul_mat_noalloc(b200_ctx * ctx, const b200_tensor * w, const b200_tensor * x, b200_tensor * dst);
B200_API size_t b200_noalloc_scratch_bytes(void);
/* The persistent decode chain: up to B200_CHAIN_MAX_PHASES dependent MUL_MAT groups of ONE token and the glue between them as ONE kernel
 * launch (b200_chain.cu).  Replaces the node sequence of a llama layer between two attention operators -- wo, ADD, RMS_NORM,
 * MUL, ffn_gate / ffn_up, GLU(SWIGLU), ffn_down, ADD, RMS_NORM, MUL, wq / wk / wv (src/models/llama.cpp:40-130) -- which the reference
 * computes node by node (ggml_compute_forward_mul_mat ggml/src/ggml-cpu/ggml-cpu.c:1226-1418, rms_norm / glu ggml/src/ggml-cpu/ops.cpp:3560-3640).
 * Phase p:   x = in0                                    (B200_CHAIN_X_PLAIN)
 *            x = rms_norm(in0 + in1, eps) * normw       (B200_CHAIN_X_ADD_NORM_MUL; the three nodes' values also go to out_add / out_norm / out_mul if not NULL)
 *            x = rms_norm(in0, eps) * normw             (B200_CHAIN_X_NORM_MUL)
 *            dst[i] = w[i] x, i < n_w <= 3              (B200_CHAIN_Y_PLAIN)
 *            + glu = swiglu(dst[0], dst[1]), n_w == 2   (B200_CHAIN_Y_SWIGLU)
 * A phase may read what earlier phases of the same call wrote.  All activations are contiguous 16-byte aligned f32 columns; weights are
 * raw q4_K or repacked q6_K with K % 2048 == 0 and M % 4 == 0.  B200_ERR_UNSUPPORTED otherwise: run the nodes one by one then. */
#define B200_CHAIN_MAX_PHASES 6
enum b200_chain_prologue { B200_CHAIN_X_PLAIN = 0, B200_CHAIN_X_ADD_NORM_MUL = 1, B200_CHAIN_X_NORM_MUL = 2 };
enum b200_chain_epilogue { B200_CHAIN_Y_PLAIN = 0, B200_CHAIN_Y_SWIGLU = 1 };
/* What follows result i of the chain's LAST phase in a llama layer's decode graph (one token): ROPE of the Q / K head rows
 * (ggml_compute_forward_rope_f32, ggml/src/ggml-cpu/ops.cpp:5524-5790; the adjacent-pair form, no YaRN mixing) and the KV-cache write of
 * the K / V rows (SET_ROWS f32 -> f16, src/llama-kv-cache.cpp:1008-1058), finished by the CTA that owns the rows.  kind: */
enum b200_chain_tail { B200_CHAIN_T_NONE = 0, B200_CHAIN_T_ROPE = 1, B200_CHAIN_T_ROPE_CACHE = 2, B200_CHAIN_T_CACHE = 3 };
typedef struct b200_chain_rope {
    int32_t         kind;
    int32_t         skip_dst;                               /* dst[i] itself is dead and its memory belongs to another result's ROPE by now: do not store it */
    int32_t         head_dim, n_dims;                       /* row length of the rope tensor; rotated dims (even, <= head_dim) */
    float           freq_scale, theta_scale, attn_factor;   /* angle of pair j: pos * theta_scale^j / freq_factors[j] * freq_scale; cos / sin scaled by attn_factor */
    const int32_t * pos;                                    /* device: the token's position */
    const float *   freq_factors;                           /* device, n_dims / 2 entries, or NULL */
    float *         out;                                    /* device: the ROPE node's result column (may be dst[i] itself) */
    const int64_t * cache_idx;                              /* device: the cache row SET_ROWS names (one entry) */
    void *          cache;                                  /* f16 cache tensor; its row r starts at cache + r * cache_nb1 */
    size_t          cache_nb1;
} b200_chain_rope;
typedef struct b200_chain_phase {
    int32_t prologue, epilogue, n_w;
    float   eps;
    const b200_tensor * w[3];
    b200_tensor *       dst[3];
    const b200_tensor * in0, * in1, * normw;
    b200_tensor *       out_add, * out_norm, * out_mul, * glu;
    const b200_chain_rope * tail[3];                        /* last phase only, B200_CHAIN_Y_PLAIN: NULL or what happens to dst[i] afterwards */
} b200_chain_phase;
B200_API int b200_decode_chain(b200_ctx * ctx, const b200_chain_phase * phases, int n_phases);
B200_API int b200_decode_chain_supported(b200_ctx * ctx, const b200_chain_phase * phases, int n_phases);   // 1 / 0
/* The chain ROW-SPLIT over the GPUs of one node, one process per GPU (the decode form of the reference's split-mode row,
 * ggml/src/ggml-cuda/ggml-cuda.cu:1451-1769, where every MUL_MAT is followed by a copy of the dst slices back to the main device): every
 * rank launches the same program; its weights w[i] hold rows [row_lo[p][i], row_lo[p][i] + w[i]->ne[1]) (multiples of 32) of result i of
 * phase p, dst[i] / glu / in0 are the FULL columns.  The all-gather is the hand-over between the phases itself: a result row is stored as
 * a (value, tag) pair into EVERY rank's hand-over area (NVLink peer stores), each rank's CTAs poll their own copy.  ll[r] = rank r's area
 * (b200_decode_chain_split_bytes(rows of the last phase's results) bytes, zeroed once, peer-mapped in this process); after the launch
 * every rank holds the full last-phase results, the full in0 columns, and its own rows of everything else.  No ROPE / KV-cache tail.
 * world == 1 is allowed (same arithmetic on one GPU: what the split results are compared with, bit for bit).  The tags that validate the
 * records count the launches through THESE areas (the counter is the last word block of the rank's own area): every rank must run the
 * same sequence of launches over a given set of areas, and nothing else it launches matters. */
#define B200_CHAIN_SPLIT_MAX_PEERS 8
typedef struct b200_chain_split {
    int32_t  world, rank;
    void *   ll[B200_CHAIN_SPLIT_MAX_PEERS];
    uint64_t ll_bytes;
    int64_t  row_lo[B200_CHAIN_MAX_PHASES][3];
} b200_chain_split;
B200_API int    b200_decode_chain_split(b200_ctx * ctx, const b200_chain_phase * phases, int n_phases, const b200_chain_split * split);
B200_API int    b200_decode_chain_split_supported(b200_ctx * ctx, const b200_chain_phase * phases, int n_phases, const b200_chain_split * split);
B200_API size_t b200_decode_chain_split_bytes(int64_t last_rows);
/* every wait inside the chain kernel is bounded (2 s): 0, or the code of the first wait that ran out since the last call (synchronizes) */
B200_API int b200_decode_chain_error(b200_ctx * ctx);
// as[K, M, n_expert], b[K, n_used|1, n_tokens] f32, ids[n_used, n_tokens] i32 -> dst[M, n_used, n_tokens]
// (ggml/src/ggml.c:3205-3241).
B200_API int b200_mul_mat_id(b200_ctx * ctx, const b200_tensor * as, const b200_tensor * b,
                             const b200_tensor * ids, b200_tensor * dst);

// q[DK, N, H, S] f32; k[DK, KV, Hk, Sk], v[DV, KV, Hk, Sk] f16 (strided views of the KV cache);
// mask[KV, >=N, 1|H, 1|S] f16 or NULL; sinks[H] f32 or NULL; dst[DV, H, N, S] f32
// (ggml/src/ggml.c:5254-5330, ggml/src/ggml-cpu/ops.cpp:8047-8272).
B200_API int b200_flash_attn_ext(b200_ctx * ctx, const b200_tensor * q, const b200_tensor * k,
                                 const b200_tensor * v, const b200_tensor * mask, const b200_tensor * sinks,
                                 b200_tensor * dst, float scale, float max_bias, float logit_softcap);

// The same operator over a PAGED KV cache (intended semantics: src/llama-kv-cache-paged.cpp:137-196 -- a layer's cache is a set of
// fixed-size pages of `page_size` consecutive positions, allocated on demand, not one contiguous buffer).  k / v describe the page POOL:
// data = first row of physical page 0, nb[1] = stride between consecutive token rows inside the pool, nb[2] = head stride, ne[1] = n_kv
// (the LOGICAL number of keys attended).  Logical key t lives in pool row table[t / page_tokens] * page_tokens + t % page_tokens.
// Query batches of at most 8 tokens (the decode kernels); larger batches: B200_ERR_UNSUPPORTED.  table == identity gives exactly
// b200_flash_attn_ext's result (same kernels, same order of operations).
typedef struct b200_kv_pages {
    const int32_t * table;        // device array [n_pages]: logical page -> physical page of the pool
    int32_t         n_pages;      // >= ceil(n_kv / page_tokens)
    int32_t         page_tokens;  // power of two, >= 16
    int64_t         pool_tokens;  // token rows the pool holds (bounds for the TMA descriptors)
} b200_kv_pages;
B200_API int b200_flash_attn_ext_paged(b200_ctx * ctx, const b200_tensor * q, const b200_tensor * k, const b200_tensor * v,
                                       const b200_kv_pages * pages, const b200_tensor * mask, const b200_tensor * sinks,
                                       b200_tensor * dst, float scale, float max_bias, float logit_softcap);

// ---- load-time weight repack (DESIGN.md "HBM layout"): in place, size- and stride-preserving, exactly invertible ------
// q6_K rows whose length is a multiple of 2048 are permuted per group of 8 blocks into a 16-byte aligned field-major
// layout; the caller then sets B200_TENSOR_REPACKED in b200_tensor.flags.  Other types are consumed in GGUF layout.
B200_API int b200_repack_supported(int type, int64_t k);
B200_API int b200_repack(b200_ctx * ctx, int type, void * data, int64_t k, int64_t nrows, size_t row_stride);
B200_API int b200_unrepack(b200_ctx * ctx, int type, void * data, int64_t k, int64_t nrows, size_t row_stride);

// ---- split-mode=row across processes: MUL_MAT fused with its all-gather over NVLink peer memory ------------------------------
// ---- weight load path: file -> pinned staging ring -> device, repack overlapped, per-GPU sharded reads --------------------
// Replaces the upload loop of llama_model_loader::load_all_data (src/llama-model-loader.cpp:1136-1186: one host thread alternating between
// read() into a 1 MiB pinned buffer and ggml_backend_tensor_set_async) and, for split-mode=row, ggml_backend_cuda_split_buffer_set_tensor
// (ggml/src/ggml-cuda/ggml-cuda.cu:900-937: the whole tensor arrives on the host, every device's row slice is copied out of it).
// An item is a byte range of the file and where it goes inside a device tensor; a reader thread fills a ring of n_buffers pinned buffers of
// chunk_bytes while the caller's thread issues the host-to-device copies on a copy stream of its own; an item with repack != 0 (a whole
// tensor of a type b200_repack_supported) is re-laid out on ctx's stream as soon as its last chunk has been issued, overlapping the next
// tensor's transfer, and gets B200_TENSOR_REPACKED set.  Returns after everything has landed.  chunk_bytes 0 = 4 MiB, n_buffers 0 = 8; reader threads: GGML_B200_LOAD_THREADS (default 4).
typedef struct b200_load_item {
    int64_t       file_offset;   // first byte of the range in the file
    size_t        nbytes;
    b200_tensor * dst;           // device tensor; bytes go to dst->data + dst_offset
    size_t        dst_offset;
    int32_t       repack;
} b200_load_item;
typedef struct b200_load_stats { double seconds, read_seconds; uint64_t bytes, chunks; } b200_load_stats;   // wall time, time inside pread, bytes uploaded
B200_API int b200_load_file(b200_ctx * ctx, const char * path, const b200_load_item * items, int n_items, size_t chunk_bytes, int n_buffers,
                            b200_load_stats * stats);
// the item of rows [row_lo, row_hi) of a [ne0, nrows] tensor of `type` stored at file_offset: what one device of a row split reads
// (row_lo / row_hi from b200_split_row_range); the caller fills dst (a tensor of row_hi - row_lo rows) and repack
B200_API int b200_load_row_slice(int type, int64_t ne0, int64_t file_offset, int64_t row_lo, int64_t row_hi, b200_load_item * item);

// One process per GPU.  Every rank allocates an identical "exchange arena" (b200_malloc), publishes it with b200_peer_handle
// (a cudaIpcMemHandle, 64 bytes, exchanged by the host program) and maps the others with b200_peer_open.  Arena layout:
//   [0, ctrl_off)                         tensors (dst of the row-split matmuls), same offsets on every rank
//   [ctrl_off, ctrl_off + 256)            control block (uint32 expected @32, cta counter @36, error @40), zeroed
//   [ctrl_off + 256, ctrl_off + 256 + 2*ctrl_off)   low-latency mirror: 8 bytes (f32 value, uint32 tag) per f32 of the tensor area
// A row-split MUL_MAT computes rows [row_lo, row_lo + w->ne[1]) of dst and its epilogue stores each of them as a (value, tag)
// pair straight into EVERY rank's mirror (peer stores over NVLink; an 8-byte store is one transaction, so the tag validates the
// value: no fence, no flag).  The SAME kernel then receives: its threads poll the rank's own mirror until every row of dst carries
// the tag and write the plain f32 column, so kernels after it on the stream see the complete dst (a rank that owns no rows of a
// tensor runs a small receive-only kernel instead).  One launch per exchange; no NCCL call and no host involvement on the path.
// Replaces the reference's copy-compute-copy loop over split buffers (ggml/src/ggml-cuda/ggml-cuda.cu:1451-1769).
#define B200_MAX_PEERS 8
typedef struct b200_exchange {
    int32_t world, rank;
    void *  base[B200_MAX_PEERS];   // arena base address of every rank as mapped in THIS process (base[rank] = own arena)
    size_t  ctrl_off;               // size of the tensor area = offset of the control block; arena size = 3 * ctrl_off + 256
} b200_exchange;
B200_API int b200_peer_handle(b200_ctx * ctx, void * dptr, void * handle64);
B200_API int b200_peer_open(b200_ctx * ctx, const void * handle64, void ** dptr);
B200_API int b200_peer_close(b200_ctx * ctx, void * dptr);
// w: this rank's rows (a [K, rows] tensor), x: the full activation column, dst: the FULL result column inside the arena
B200_API int b200_mul_mat_rowsplit(b200_ctx * ctx, const b200_tensor * w, const b200_tensor * x, b200_tensor * dst, int64_t row_lo,
                                   const b200_exchange * ex);
// n <= 3 row-split MUL_MATs of the same activation column (wq/wk/wv, gate/up): runs of equal weight type share one launch
B200_API int b200_mul_mat_rowsplit_multi(b200_ctx * ctx, const b200_tensor * const * w, int n, const b200_tensor * x, b200_tensor * const * dst,
                                         const int64_t * row_lo, const b200_exchange * ex);
B200_API int b200_exchange_error(b200_ctx * ctx, const b200_exchange * ex);     // nonzero after a receive timed out (a peer died)

// ---- row split (split-mode=row) helpers; reference ggml-cuda.cu:777-803 --------------------
// rows [*row_low, *row_high) of an nrows-row weight that rank `rank` of `world` owns under `tensor_split`
// (NULL = even), boundaries rounded down to `rounding` rows.
B200_API int b200_split_row_range(int64_t nrows, const float * tensor_split, int world, int rank,
                                  int64_t rounding, int64_t * row_low, int64_t * row_high);

// ---- layer distribution over GPUs (host only); reference src/llama-multi-gpu.{h,cpp} -------------------------------------------
// llama_multi_gpu_strategy (llama-multi-gpu.h:84-91); values other than the three below plan like MEMORY_BALANCED (llama-multi-gpu.cpp:170-185)
enum b200_layer_strategy { B200_LAYERS_ROUND_ROBIN = 0, B200_LAYERS_MEMORY_BALANCED = 1, B200_LAYERS_TENSOR_PARALLEL = 4 };
// llama_layer_assignment (llama-multi-gpu.h:75-81); split_gpu_ids[0 .. n_split) are the GPUs a split layer spans, the rest -1
/* sizeof of a public struct as this library was compiled: what a binding (ctypes, cgo, JNI) checks its own mirror against before the first call.
 * which: 0 b200_tensor, 1 b200_chain_rope, 2 b200_chain_phase, 3 b200_chain_split, 4 b200_kv_pages, 5 b200_load_item, 6 b200_load_stats,
 * 7 b200_exchange, 8 b200_layer_assignment; anything else: 0 */
B200_API size_t b200_abi_sizeof(int which);

typedef struct b200_layer_assignment {
    int32_t  layer_id, gpu_id;
    uint64_t memory_required;
    int32_t  is_split, n_split;
    int32_t  split_gpu_ids[B200_MAX_PEERS];
} b200_layer_assignment;
// llama_multi_gpu_manager::distribute_layers (llama-multi-gpu.cpp:156-332): plans n_layers layers of layer_sizes[i] bytes (0 beyond n_sizes)
// over n_gpus GPUs with gpu_free_memory[g] bytes free.  Writes out[0 .. n_layers) and, if not NULL, the bytes charged to each GPU into
// gpu_allocated[n_gpus].  Returns the number of assignments (0 without GPUs) or a negative b200_status.
B200_API int b200_distribute_layers(int strategy, int tensor_parallel_size, int n_layers, const uint64_t * layer_sizes, int n_sizes,
                                    int n_gpus, const uint64_t * gpu_free_memory, b200_layer_assignment * out, uint64_t * gpu_allocated);
// the plan in the form llama.cpp consumes: llama_model_params::tensor_split[n_gpus] (include/llama.h; `-ts`), the share of layers
// (-sm layer) or of every split matrix's rows (-sm row; what b200_split_row_range takes) per GPU
B200_API int b200_plan_tensor_split(const b200_layer_assignment * plan, int n, int n_gpus, float * tensor_split);

// ---- ggml plugin entry points ----------------------------------------------------------------
struct ggml_backend_reg;
B200_API struct ggml_backend_reg * ggml_backend_init(void);       // ggml-backend-reg.cpp:255-272
B200_API int                       ggml_backend_score(void);
B200_API struct ggml_backend_reg * ggml_backend_b200_reg(void);   // for a static in-tree registration

#ifdef __cplusplus
}
#endif
