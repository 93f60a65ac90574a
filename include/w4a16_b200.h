/*
 * w4a16_b200.h -- C ABI of libw4a16_b200.so: the B200-native (sm_100a) W4A16
 * quantized-linear forward that replaces vllm-rust's AWQ / AWQ-Marlin / GPTQ
 * CUDA path behind `trait QuantizedLinear`.
 *
 * Every entry point names the reference interface it replaces (file:line under
 * crates/core/src/quantization/ of vasia123/vllm-rust).  INTEGRATION.md shows the
 * Rust `extern "C"` block + `impl QuantizedLinear for B200W4A16Linear` that binds
 * these symbols from crates/core.
 *
 * Conventions
 *   - Plain pointers and sizes only.  No torch / candle types.
 *   - Every function returns 0 on success, non-zero on failure; the message of
 *     the last failure on the calling thread is w4a16_last_error()
 *     (mirrors candle_core::Error::Msg via bail!, marlin_cuda.rs:546-564).
 *   - All device pointers must live on the handle's device.  x / y / checkpoint
 *     tensors are caller-owned; the handle owns the repacked weights and a fixed
 *     workspace.  w4a16_forward_into never allocates and never synchronises, so
 *     it can be captured in a CUDA graph (replaces OutputPool-backed
 *     marlin_gemm_pooled, marlin_cuda.rs:1655-1753).
 *   - `stream` is a cudaStream_t passed as void* (0 = legacy default stream).
 *   - Re-entrancy: the weights are read-only after load, and every scratch a forward
 *     uses belongs either to the handle or to the (device, stream) pair, so different
 *     handles may run concurrently on different streams.  ONE handle runs one forward
 *     at a time (its calls must be stream-ordered), like the reference's MarlinLinear
 *     with its workspace tensor (marlin.rs:586-640).  Scratch is grown outside stream
 *     capture only and an outgrown buffer lives as long as its handle, so a captured
 *     graph never sees a freed address; inside a capture a too-small scratch is an
 *     error that names w4a16_reserve.  The split-K scratch of the (device, stream)
 *     pair (mid-M and prefill kernels, 11 <= M <= 256) is the exception: it is
 *     created by the first un-captured forward on that stream, and a captured forward
 *     that finds none runs unsplit / on the kernel family that needs none instead
 *     of failing -- run one forward on the stream before capturing.
 *   - There is no CPU fallback.  If no sm_100 device is present every call that
 *     needs one fails loudly.
 */
#ifndef W4A16_B200_H
#define W4A16_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define W4A16_ABI_VERSION 1

typedef struct w4a16_linear *w4a16_handle_t;
typedef void *w4a16_stream_t; /* cudaStream_t */
typedef void *w4a16_comm_t;   /* ncclComm_t   */

/* Checkpoint formats.  SURVEY.md 8a rows U1/D1 (AWQ), G1 (GPTQ, zeros as stored,
 * no "+1"), G3 (symmetric q-8, qzeros ignored: marlin_gemm.cu:43-48). */
enum w4a16_format {
    W4A16_FORMAT_AWQ = 0,      /* qweight [K,N/8] + qzeros [G,N/8], AWQ nibble interleave (awq.rs:20-49) */
    W4A16_FORMAT_GPTQ_ZP = 1,  /* qweight [K/8,N] + qzeros [G,N/8], sequential nibbles (gptq_dequant.cu:6-14) */
    W4A16_FORMAT_GPTQ_SYM = 2  /* qweight [K/8,N], W = (q-8)*s (MarlinScalarType::Uint4b8, marlin.rs:52-56) */
};

enum w4a16_dtype { W4A16_DTYPE_F16 = 0, W4A16_DTYPE_BF16 = 1, W4A16_DTYPE_F32 = 2 };

/* Tensor-parallel shard of one linear (distributed/parallel_layers.rs:36-324). */
enum w4a16_tp_mode {
    W4A16_TP_NONE = 0,
    W4A16_TP_COLUMN = 1, /* rank r keeps output columns [r*N/tp, (r+1)*N/tp); no collective          */
    W4A16_TP_ROW = 2     /* rank r keeps input rows    [r*K/tp, (r+1)*K/tp); forward ends in all-reduce */
};

/* -- errors ---------------------------------------------------------------- */
const char *w4a16_last_error(void);
int w4a16_abi_version(void);

/* -- shape gate -------------------------------------------------------------
 * check_marlin_supports_shape (marlin.rs:308-341): N % 64 == 0, K % 128 == 0,
 * group_size in {-1, 32, 64, 128} (marlin.rs:48) and K % group_size == 0.     */
int w4a16_check_shape(int64_t out_features, int64_t in_features, int group_size);

/* -- construction -----------------------------------------------------------
 * QuantizationConfig::create_linear(in, out, bias, device) (config.rs:125-131);
 * AwqMarlinLinear::new (awq_marlin.rs:431-446); GptqLinear::new (gptq.rs:254-).
 * in_features / out_features are the FULL (unsharded) sizes; w4a16_create is
 * w4a16_create_shard with W4A16_TP_NONE.                                       */
int w4a16_create(w4a16_handle_t *out, int64_t in_features, int64_t out_features, int group_size,
                 int format, int has_bias, int device);
int w4a16_create_shard(w4a16_handle_t *out, int64_t in_features, int64_t out_features,
                       int group_size, int format, int has_bias, int device, int tp_mode,
                       int tp_rank, int tp_size);
int w4a16_destroy(w4a16_handle_t h);

/* -- load-time repack -------------------------------------------------------
 * QuantizedLinear::load_weights(&HashMap) with keys "qweight","qzeros","scales",
 * "bias","g_idx" in CHECKPOINT layout; the repack happens inside
 * (AwqMarlinLinear::load_weights awq_marlin.rs:580-667 = R2+R3+R1+S1;
 *  GptqLinear::load_weights gptq.rs:435-454).
 * All pointers are DEVICE pointers to the full (unsharded) checkpoint tensors;
 * a sharded handle reads only its slice.  scales may be F16/BF16/F32 and are
 * stored as BF16 (round-to-nearest-even, awq_marlin.rs:649-656).  bias may be
 * NULL when has_bias == 0.  The call is asynchronous on `stream`; the
 * checkpoint tensors may be freed once the stream has passed it.              */
int w4a16_load_awq(w4a16_handle_t h, const uint32_t *qweight /*[K,N/8]*/,
                   const uint32_t *qzeros /*[G,N/8]*/, const void *scales /*[G,N]*/,
                   int scales_dtype, const void *bias /*[N] or NULL*/, int bias_dtype,
                   w4a16_stream_t stream);
/* g_idx: int32 [K] or NULL.  With g_idx the canonical act-order semantics
 * W[k,n] = (q[k,n] - z[g_idx[k],n]) * s[g_idx[k],n] are implemented by a K
 * permutation at load plus an activation gather at forward.  (The reference loads
 * g_idx and never uses it: gptq.rs:349-366, marlin_gemm.cu:79-83.)  g_idx must
 * assign exactly group_size rows to every group.  g_idx is read on the HOST side
 * of the call (it synchronises `stream` once, at load).
 * ROW shards: the rows are sorted by group over the FULL K and then
 * sharded, so rank r keeps whole groups; its forward expects x already in that
 * order (w4a16_row_perm gives it) -- there is no gather inside a ROW shard.   */
int w4a16_load_gptq(w4a16_handle_t h, const uint32_t *qweight /*[K/8,N]*/,
                    const uint32_t *qzeros /*[G,N/8] or NULL for GPTQ_SYM*/,
                    const void *scales /*[G,N]*/, int scales_dtype,
                    const int32_t *g_idx /*[K] device, or NULL*/, const void *bias, int bias_dtype,
                    w4a16_stream_t stream);

/* -- fused sibling projections and the SwiGLU epilogue -------
 * The reference runs the sibling projections of one input as separate quantized linears:
 * q_proj / k_proj / v_proj (qwen3_quantized.rs:84-96 create them one by one, :179-181 run
 * them), gate_proj / up_proj followed by fused_swiglu or silu * mul
 * (qwen3_quantized.rs:532-551, swiglu.cu:14-46).  One handle of width sum(n_part) loaded
 * from the siblings' separate checkpoints replaces the 2-3 launches by one: part p fills
 * output columns [n_begin, n_begin + n_part) from a checkpoint that is n_part columns wide
 * (the same tensors w4a16_load_awq / _gptq take for a linear of that width).  n_begin and
 * n_part are multiples of 16; the handle is unsharded (W4A16_TP_NONE; sharded siblings are
 * sharded by the caller: each rank loads its column slice of every sibling as a part of a
 * local handle).  The handle counts as loaded once the parts cover every column exactly
 * once; loading the same part again overwrites it in place.  GPTQ parts must bring the same
 * g_idx (they share x, so one gather serves all).  bias: the part's own [n_part] vector.
 * The result is bit-identical to one linear loaded from the column-wise concatenated
 * checkpoint (tests/test_fused_gpu.py).                                        */
int w4a16_load_awq_part(w4a16_handle_t h, int64_t n_begin, int64_t n_part, const uint32_t *qweight,
                        const uint32_t *qzeros, const void *scales, int scales_dtype, const void *bias,
                        int bias_dtype, w4a16_stream_t stream);
int w4a16_load_gptq_part(w4a16_handle_t h, int64_t n_begin, int64_t n_part, const uint32_t *qweight,
                         const uint32_t *qzeros, const void *scales, int scales_dtype,
                         const int32_t *g_idx, const void *bias, int bias_dtype, w4a16_stream_t stream);
/* W4A16_EPILOGUE_SWIGLU: the handle holds gate_proj (part at n_begin = 0) and up_proj (part at
 * n_begin = I) of one MLP, I = out_features / 2 at creation, I % 8 == 0, and its forward
 * writes y[M, I] = bf16( silu_f32(gate) * f32(up) ), gate and up being the two projections
 * ROUNDED TO BF16 first (with their bias, if any) -- the tensors the reference hands to
 * fused_swiglu_bf16 (swiglu.cu:40-45: silu(x) = x / (1 + expf(-x)) in fp32, one bf16
 * rounding of the product).  The resident layout interleaves the two projections (gate
 * column c and up column c share a 16-row block), so the pair meets in one thread's
 * registers and the [M, 2I] intermediate never exists.  Call before any load; loads go
 * through w4a16_load_*_part only.  w4a16_out_features then reports I, w4a16_weight_rows
 * 2I; the parity hooks answer in [gate | up] column order.  Forward: w4a16_forward_into
 * only (no ROW/fused all-reduce: the handle is unsharded).                    */
enum w4a16_epilogue { W4A16_EPILOGUE_NONE = 0, W4A16_EPILOGUE_SWIGLU = 1 };
int w4a16_set_epilogue(w4a16_handle_t h, int epilogue);

/* -- cache of the resident layout -----------------------------
 * The reference repacks at every start (AwqMarlinLinear::load_weights,
 * awq_marlin.rs:580-667).  save_tiled writes a loaded handle's resident bytes (header +
 * qw | meta | bias | perm, checksummed; written to path.tmp and renamed) and load_tiled
 * reads them into a handle created with the SAME geometry, format, bias flag, shard and
 * epilogue: every header field is compared and a mismatch is an error naming the field.
 * fingerprint: the caller's identity of the source checkpoint (e.g. a hash of the
 * safetensors header); 0 on load skips that one check.  After load_tiled the handle is
 * bit-identical to a freshly repacked one (tests/test_cache_gpu.py).           */
int w4a16_save_tiled(w4a16_handle_t h, const char *path, uint64_t fingerprint);
int w4a16_load_tiled(w4a16_handle_t h, const char *path, uint64_t fingerprint, w4a16_stream_t stream);

/* -- forward ----------------------------------------------------------------
 * QuantizedLinear::forward(&Tensor) (config.rs:149) as it is executed by
 * MarlinLinear::forward -> marlin_gemm_pooled / marlin_gemm (marlin.rs:741-754,
 * marlin_cuda.rs:843-953, 1655-1753) and GptqLinear::forward -> gptq_gemm
 * (gptq.rs:413-433, gptq_cuda.rs:322).
 * x: [M, K_local] bf16, contiguous, 16-byte aligned.  y: [M, N_local] bf16; every
 * element is written, no zero-init needed (marlin_cuda.rs:431-440).  Leading
 * dims are flattened by the caller (awq_marlin.rs:545-571).  M >= 1.
 * For a ROW shard y is the rank's partial sum WITHOUT bias.
 * Non-finite activations: an Inf or NaN anywhere in a token's row makes that token's
 * outputs non-finite, as the reference's fp32 FMA chain does (awq_gemv.cu:597-664).
 * The M <= 4 kernels, which round activations to per-group scaled integers, return NaN
 * for every output of such a token (the reference may return +-Inf for some of them);
 * other tokens of the batch are unaffected.  Finite inputs: |x| down to 2^-6 of its
 * group's largest magnitude is exact, smaller ones carry an absolute error of at most
 * 2^-14 of that magnitude (tests/test_parity_gpu.py::test_imma_*).            */
int w4a16_forward_into(w4a16_handle_t h, const void *x_bf16, void *y_bf16, int64_t M,
                       w4a16_stream_t stream);

/* Same call with HOST buffers: H2D copy of x, forward, D2H copy of y, stream
 * synchronise.  Staging buffers are owned by the handle and grown on demand
 * (so this entry point may allocate; it is the end-to-end path bench.py times). */
int w4a16_forward_host(w4a16_handle_t h, const void *x_bf16_host, void *y_bf16_host, int64_t M,
                       w4a16_stream_t stream);

/* No counterpart in the reference: added for callers that
 * cannot call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid it).
 * The decode forward for 1 <= M <= 32 tokens through the register-ring mma.sync kernel, with the
 * kernel's split-K scratch (fp32 partial fragments + arrival tickets) in a static device array
 * of the library instead of the handle's workspace: 16 KiB of tickets + 8 MiB of partials, held
 * by every GPU that loads the library until the process exits.  The call allocates nothing and
 * synchronises nothing (one cudaMemsetAsync of the tickets + one kernel on `stream`).
 * x, y, bias and row-shard conventions as w4a16_forward_into.  Fails with a message when M > 32,
 * when the shape's scratch need -- split x (N / 16) x ceil(M / 8 to 1, 2 or 4) x 512 bytes, the
 * split chosen as for w4a16_forward_into -- exceeds 8 MiB, when N > 65536, or for an act-order
 * handle that gathers x itself (that gather buffer is allocated on demand).
 * Every call on a device shares the one array and it keeps the last call's contents (the tickets
 * are cleared on entry, partials are rewritten before they are read): calls that may overlap on
 * the GPU -- different streams, or graph branches -- race; serialise them.                 */
int w4a16_forward_decode_static(w4a16_handle_t h, const void *x_bf16, void *y_bf16, int64_t M,
                                w4a16_stream_t stream);

/* RowParallelLinear::forward (parallel_layers.rs:275-313): local GEMM, then
 * all_reduce(Sum) (nccl.rs:1338-1364), then + bias.  `comm` is the caller's
 * ncclComm_t.  tp_size == 1 short-circuits the collective (nccl.rs:1339-1341). */
int w4a16_forward_row_allreduce(w4a16_handle_t h, const void *x_bf16, void *y_bf16, int64_t M,
                                w4a16_comm_t comm, w4a16_stream_t stream);

/* Fused row-parallel forward for decode sizes: replaces the ncclAllReduce of
 * RowParallelLinear::forward (parallel_layers.rs:295-304, nccl.rs:1338-1364) and the dormant
 * kernels/custom_allreduce.cu.  Every rank owns one symmetric buffer of fp32 words that hold a sentinel
 * (-0.0f) when empty; the GEMV's epilogue stores its fp32 partial sums into every peer's buffer over
 * NVLink (P2P stores), polls its own buffer until every source's word has arrived (the data words are
 * their own flags: no fence, no flag round trip), sums the world's partials in rank order (+ bias) and
 * puts the sentinel back.  One kernel, no NCCL launch; results are bit-identical on every rank.  One
 * peer group serves row-parallel linears of different N and M (every launch leaves its buffer clean).
 *   peer_create : allocate this rank's buffer on the current device (max_partial_floats >=
 *                 N_out * 4 of the largest row-parallel linear);
 *   peer_export : 64-byte IPC handle of the buffer (HOST memory), to be all-gathered out of band;
 *   peer_attach : all_handles = world_size x 64 bytes in rank order; opens every peer's buffer.
 * forward_row_fused falls back to w4a16_forward_row_allreduce (same result up to the reduction's
 * rounding) when M > 4, the shape is not eligible, or peer is NULL.  The launch counter the ranks
 * agree on lives in device memory, so the fused call can be captured into a CUDA graph.
 * All ranks must make the same sequence of calls (SPMD), like any collective.               */
typedef struct w4a16_peer *w4a16_peer_t;
#define W4A16_PEER_HANDLE_BYTES 64
int w4a16_peer_create(w4a16_peer_t *out, int rank, int world_size, int64_t max_partial_floats);
int w4a16_peer_export(w4a16_peer_t peer, void *handle_out);
int w4a16_peer_attach(w4a16_peer_t peer, const void *all_handles);
int w4a16_peer_destroy(w4a16_peer_t peer);
int w4a16_forward_row_fused(w4a16_handle_t h, const void *x_bf16, void *y_bf16, int64_t M,
                            w4a16_peer_t peer, w4a16_comm_t comm, w4a16_stream_t stream);

/* -- a chain of decode linears in ONE launch ------------------
 * The reference's decode step calls its quantized linears one after the other in a fixed
 * order -- q/k/v and o around attention (qwen3_quantized.rs:704-709), gate/up -> silu*mul
 * -> down (:532-551), layer after layer (:890-925) -- and replays that sequence as a CUDA
 * graph (engine/cuda_graph.rs).  A chain is that fixed sequence for M <= 4 tokens as one
 * object: linear i reads x[i] and writes y[i], in order, and linear i+1 may read what
 * linear i wrote (x[i+1] == y[i], or any overlap).  Where the shapes allow it (one group
 * size of 32 / 64 / 128 for all linears, M <= 4 and the staged activations leaving room
 * for the weight ring; act-order linears included: their gather happens inside the
 * launch) chain_run is ONE persistent launch: one CTA per SM streams every linear's weights
 * through one shared-memory ring that stays full across the boundaries between linears, and
 * consecutive linears are separated by a grid-wide barrier.  Otherwise chain_run makes the
 * n w4a16_forward_into calls itself (w4a16_chain_kernel tells which).  Results: those of
 * w4a16_forward_into up to the summation order of the fp32 partial sums (same integer
 * arithmetic, same bounds, same non-finite policy).
 * x / y pointers are fixed at creation, like the arguments of a captured graph; SwiGLU
 * handles and biases are honoured; a ROW shard writes its partial sum (no collective
 * inside a chain).  chain_run never allocates or synchronises and can be captured in a
 * CUDA graph.  The persistent launch needs every SM of the device: all its CTAs must be
 * resident together (it waits for earlier kernels, but must not share the device with a
 * kernel that waits for IT).  One chain must not run concurrently with itself.          */
typedef struct w4a16_chain *w4a16_chain_t;
int w4a16_chain_create(w4a16_chain_t *out, int n, const w4a16_handle_t *handles,
                       const void *const *x_bf16, void *const *y_bf16, int64_t M);
/* Prologue of a linear inside a chain: the memory-bound ops the reference runs between its
 * linears -- residual add (add_bf16, activations.cu:99-113; half_add_pooled, qwen3_quantized.rs:1062-1068) and
 * RMSNorm (rms_norm_bf16, layernorm.cu:59-126; input_layernorm / post_attention_layernorm,
 * qwen3_quantized.rs:1043-1075) -- done by the chain kernel while it stages the linear's activations:
 *     h      = bf16(f32(x) + f32(residual))     (h = x when there is no residual)
 *     normed = bf16(f32(h) * rsqrt(mean(f32(h)^2) + eps) * f32(norm_weight))     -> the linear's input
 * residual is residual_in (caller's memory) if given, else -- with add_running_residual -- the h of the previous
 * prologue of the same launch, which the kernel keeps in shared memory.  residual_out / normed_out, if given,
 * receive h / normed ([M, in_features] bf16).  The sum of squares is taken over the ROUNDED h in fp32 like the
 * reference's separate kernels do; results equal theirs up to the order of that sum (a bf16 ulp on isolated
 * elements, tests/test_chain_gpu.py).  norm_weight == NULL: the linear has no prologue.  Chains with prologues
 * need the persistent kernel (creation fails otherwise) and M <= 4. */
typedef struct w4a16_chain_prologue {
    const void *norm_weight;  /* bf16 [in_features], device */
    const void *residual_in;  /* bf16 [M, in_features] or NULL */
    void *residual_out;       /* bf16 [M, in_features] or NULL */
    void *normed_out;         /* bf16 [M, in_features] or NULL (parity hook) */
    float eps;
    int add_running_residual;
} w4a16_chain_prologue;
int w4a16_chain_create_ex(w4a16_chain_t *out, int n, const w4a16_handle_t *handles,
                          const void *const *x_bf16, void *const *y_bf16, int64_t M,
                          const w4a16_chain_prologue *prologues /* [n] or NULL */);
/* Tensor-parallel chains: with a peer group attached (w4a16_peer_*, see below), every ROW shard of the chain finishes
 * with the all-reduce of RowParallelLinear::forward (parallel_layers.rs:295-311) INSIDE the launch: its epilogue stores
 * the rank's fp32 partial sums into every peer's symmetric buffer over NVLink, polls its own buffer until every rank's
 * words have arrived (the data are their own flags), sums them in rank order (bit-identical on every rank), adds the
 * bias and hands the bf16 result to the next linear through the chain's mailbox -- no NCCL call, no kernel boundary.
 * Without a peer group a ROW shard writes its partial sum.  All ranks must build the same chain and run it the same
 * number of times (SPMD); call once after creation, before the first run.  Not combined with prologues yet. */
int w4a16_chain_set_peer(w4a16_chain_t chain, w4a16_peer_t peer);
int w4a16_chain_run(w4a16_chain_t chain, w4a16_stream_t stream);
/* chain_run with HOST buffers at the two ends: H2D copy of the first linear's x ([M, in_features] bf16), the chain,
 * D2H copy of the last linear's y, stream synchronise -- what one decode step moves between host and device (the
 * activations in between never leave the GPU, in the reference neither).  The end-to-end call bench.py times. */
int w4a16_chain_run_host(w4a16_chain_t chain, const void *x0_bf16_host, void *y_last_bf16_host,
                         w4a16_stream_t stream);
int w4a16_chain_destroy(w4a16_chain_t chain);
/* "chain_imma (<m> mailboxes, <b> barriers)" (persistent launch) or "per_linear (<reason>)".
 * Ordering inside the persistent launch: where x[i] lies inside the y of the latest earlier
 * linear that wrote it (the normal case: x[i] == y[j], or at M = 1 any 16-byte aligned
 * sub-range of it), that y is also stored into a chain-owned mailbox whose words hold a
 * sentinel until written and linear i polls the mailbox -- no grid barrier, the data are
 * their own flags.  Any other overlap between the buffers of two linears puts a grid-wide
 * barrier in front of the later one.  Linears that share nothing are not ordered at all. */
const char *w4a16_chain_kernel(w4a16_chain_t chain);
/* Profiling hook (W4A16_GV_DEBUG=1 at chain_create): %globaltimer stamps
 * [n][CTAs][8] = {linear started, activations staged, weights streamed, y written, arrival hint
 * seen, first activations valid, re-reads of those, -} of the last run; synchronises the device. */
int w4a16_chain_debug_stamps(w4a16_chain_t chain, uint64_t *out, int64_t words);

/* Communicator bootstrap for callers that do not already hold an ncclComm_t
 * (distributed/launcher.rs:132-262: ncclGetUniqueId on rank 0, out-of-band
 * broadcast of the 128-byte id, ncclCommInitRank).  NCCL is dlopen'ed as
 * libnccl.so.2, like nccl.rs:479-485.  unique_id is a 128-byte HOST buffer. */
#define W4A16_NCCL_UNIQUE_ID_BYTES 128
int w4a16_nccl_unique_id(void *unique_id_out);
int w4a16_nccl_comm_init(w4a16_comm_t *comm_out, int world_size, const void *unique_id, int rank);
int w4a16_nccl_comm_destroy(w4a16_comm_t comm);

/* Reserve per-handle scratch for up to max_tokens rows (act-order gather buffer,
 * host staging).  Call before CUDA-graph capture.                             */
int w4a16_reserve(w4a16_handle_t h, int64_t max_tokens);

/* -- parity hooks (bit-exact checks against the oracle) ----------------------
 * dequant: gptq_dequant_int4_bf16 / awq_marlin_dequant_int4_bf16 semantics,
 * out[k,n] = bf16_rn(f32(q - z) * f32(bf16 scale)), dense [K_local, N_local]
 * row-major, in ORIGINAL k order (act-order un-permuted).
 * unpack: q as one byte per nibble [K_local, N_local], z [G_local, N_local].
 * scales: the resident bf16 scales [G_local, N_local].                        */
int w4a16_dequant_into(w4a16_handle_t h, void *w_bf16, w4a16_stream_t stream);
/* The reference's other grid ("D1", awq_dequantize_cpu, awq.rs:56-166): f16_rn(f32(q - z) * f32(scale)) with the
 * checkpoint's ORIGINAL scales (the resident layout keeps only their bf16 rounding, awq_marlin.rs:649-656, so the
 * checkpoint's [G,N] scales tensor is passed again; device pointer, full/unsharded like at load).  out: dense
 * [K_local, N_local] f16.  A parity hook: no kernel computes on this grid.   */
int w4a16_dequant_d1_into(w4a16_handle_t h, const void *scales, int scales_dtype, void *w_f16,
                          w4a16_stream_t stream);
/* Checkpoint row held by every local (tiled) row: identity + k_off without act-order.  For an act-order ROW
 * shard this is the slice of the global sort-by-group permutation the shard was cut from, and the order in which
 * the caller must present the shard's x columns (x_local[:, i] = x_full[:, perm[i]]).  perm_out: HOST, K_local. */
int w4a16_row_perm(w4a16_handle_t h, int32_t *perm_out_host);
int w4a16_unpack_into(w4a16_handle_t h, uint8_t *q, uint8_t *z, w4a16_stream_t stream);
int w4a16_scales_into(w4a16_handle_t h, void *scales_bf16, w4a16_stream_t stream);

/* -- introspection ----------------------------------------------------------- */
int64_t w4a16_in_features(w4a16_handle_t h);   /* local K */
int64_t w4a16_out_features(w4a16_handle_t h);  /* local N: the width of y (I for a SwiGLU handle) */
int64_t w4a16_weight_rows(w4a16_handle_t h);   /* columns of W held (2I for a SwiGLU handle, else = out_features) */
int w4a16_has_bias(w4a16_handle_t h);
int64_t w4a16_weight_bytes(w4a16_handle_t h);  /* resident bytes of the repacked layout */
/* Kernels launched by this library in this process so far (bench.py's gpu_launches). */
uint64_t w4a16_kernel_launches(void);
/* Name of the kernel the dispatcher picks for M tokens ("gemv_stream_imma2", "decode_mma_nt1", "midm_mma", "prefill_umma", ...). */
const char *w4a16_kernel_for(w4a16_handle_t h, int64_t M);
/* Override dispatch for tests: "auto" | "decode" (M <= 32 kernels) | "prefill" (tcgen05 kernel).  Non-zero if unknown. */
int w4a16_set_dispatch(const char *mode);

/* Profiling hook (W4A16_GV_DEBUG=1 in the environment): the decode kernel stamps %globaltimer per CTA
 * {start, dependency wait passed, activations staged, weights streamed, CTA combined, end, -, -} into a
 * ring of 4 launches x 1024 CTAs x 8 words; this copies the ring out (synchronises the device). */
int w4a16_debug_stamps(uint64_t *out, int64_t words, uint64_t *launches);

#ifdef __cplusplus
}
#endif
#endif /* W4A16_B200_H */
