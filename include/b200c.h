/*
 * b200c.h -- C ABI of libb200c.so: the B200-native (sm_100a) tensor-parallel
 * communication hot path behind vLLM's device-communicator API.
 *
 * Plain C, plain pointers and sizes, no torch types.  Every entry point returns 0
 * on success or a negative B200C_E* code; b200c_last_error() gives the message of
 * the last failure on the calling thread.  No entry point synchronises the host
 * with the device unless its comment says so; kernels are enqueued on the given
 * stream.  Nothing touches CUDA at dlopen() time (the reference's local workers
 * are fork()ed, /root/reference/src/launch.py:153-190).
 *
 * Citations: REF/ = /root/reference/, VLLM/ = site-packages/vllm (0.22.0),
 * FI/ = site-packages/flashinfer (0.6.11).  The reference repo itself is a pure
 * Python launcher; the native interface this library replaces is the op set of
 * its un-vendored dependency, reached from REF/src/launch.py:290-292 (init_device)
 * and :322-349 (execute_model).
 */
#ifndef B200C_H_
#define B200C_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define B200C_MAX_RANKS 8
#define B200C_IPC_HANDLE_BYTES 64

/* dtype codes */
#define B200C_BF16 0
#define B200C_F16 1
#define B200C_F32 2

/* all-reduce variants */
#define B200C_AUTO 0
#define B200C_ONESHOT 1 /* every rank pushes its whole message to every peer */
#define B200C_TWOSHOT 2 /* push-scatter, reduce own chunk, push-gather */
#define B200C_NVLS 3    /* in-switch reduction (multimem.ld_reduce / multimem.st); needs b200c_attach_nvls; \
                           tolerance-parity: the switch's summation order is unspecified */

/* error codes */
#define B200C_OK 0
#define B200C_EINVAL (-1)   /* bad argument (shape, alignment, dtype, rank)   */
#define B200C_ECUDA (-2)    /* a CUDA runtime call failed                       */
#define B200C_ETOOBIG (-3)  /* message exceeds the arena the ctx was built for */
#define B200C_ETIMEOUT (-4) /* a device-side wait on a peer timed out (sticky)  */
#define B200C_ENOMEM (-5)

/* tuning keys for b200c_set_tuning / b200c_get_tuning */
#define B200C_TUNE_ONESHOT_MAX_BYTES 1 /* AUTO picks one-shot when bytes <= this   */
#define B200C_TUNE_ONESHOT_BLOCKS 2    /* max CTAs for the one-shot kernels         */
#define B200C_TUNE_TWOSHOT_BLOCKS 3    /* CTAs for the two-shot kernels             */
#define B200C_TUNE_SPIN_TIMEOUT_MS 4   /* device-side wait timeout, 0 = wait forever */
#define B200C_TUNE_COPY_BLOCKS 5       /* CTAs for all-gather / send / recv          */
#define B200C_TUNE_LINK_BLOCKS 6       /* two-shot CTAs doing link work; 0 = half the grid */
#define B200C_TUNE_PDL 7               /* programmatic dependent launch on/off (default on) */
#define B200C_TUNE_ONESHOT_SPLIT 8     /* CTAs (cluster size) per row, one-shot fused: 0 auto, 1 none */
#define B200C_TUNE_DEBUG_EPOCH 9       /* test hook, host-synchronous: read / jump the call counter     */
#define B200C_TUNE_TRAP_ON_TIMEOUT 10  /* 0 (default): a timed-out wait sets the sticky error word; 1: it also \
                                          executes `trap`, so the next CUDA call of the process fails    */

#define B200C_TUNE_NVLS_MIN_BYTES 11   /* AUTO picks NVLS at or above this size; < 0 = never (default: 48 MiB at 8 ranks, never below) */
#define B200C_TUNE_NVLS_LINK_BLOCKS 12 /* NVLS CTAs doing the in-switch reduce + broadcast; 0 = grid/6    */
#define B200C_TUNE_NVLS_GROUPS 13      /* NVLS pipeline groups (1..14); 0 = by size (one per 8 MiB, <= 8) */
#define B200C_TUNE_NVLS_COPY_BLOCKS 14 /* retired (every non-reduce CTA copies in, then consumes): accepted, ignored */
#define B200C_TUNE_NVLS_TEAMS 15       /* NVLS reduce CTAs work in this many teams (1..14), each on its own group; 0 = one per group */
#define B200C_TUNE_NVLS_DEBUG 16       /* measurement only, results are garbage: 1 no copy-in, 2 no ld_reduce, \
                                          4 no multimem.st, 8 no consumer */

typedef struct b200c_ctx b200c_ctx;

int b200c_version(void);
const char* b200c_last_error(void);

/* ---- arenas ------------------------------------------------------------------
 * One peer-visible arena per rank holds everything peers write into: control
 * flags, three rotating receive regions for the all-reduce, two staging regions
 * for the all-gather and the point-to-point mailboxes.  The caller owns it
 * ("this class does not own any device memory", as the reference's CustomAllreduce,
 * FI/data/include/flashinfer/comm/vllm_custom_all_reduce.cuh:333-335).
 */

/* Bytes one rank's arena needs.  max_bytes: largest all-reduce message;
 * ag_bytes: largest all-gather OUTPUT; p2p_bytes: largest single send.
 * Replaces meta_size() (VLLM/_custom_ops.py:2892-2893). */
int b200c_arena_bytes(int world, size_t max_bytes, size_t ag_bytes, size_t p2p_bytes,
                      size_t* out_bytes);

/* cudaMalloc + cudaIpcGetMemHandle on the current device.
 * Replaces allocate_shared_buffer_and_handle (VLLM/_custom_ops.py:2906-2907,
 * FI/data/csrc/vllm_custom_all_reduce.cu:120-140). */
int b200c_alloc_shared(size_t bytes, void** dptr, uint8_t handle[B200C_IPC_HANDLE_BYTES]);

/* cudaIpcOpenMemHandle.  Replaces open_mem_handle (VLLM/_custom_ops.py:2910-2911). */
int b200c_open_handle(const uint8_t handle[B200C_IPC_HANDLE_BYTES], void** dptr);
int b200c_close_handle(void* dptr);

/* cudaFree.  Replaces free_shared_buffer (VLLM/_custom_ops.py:2914-2915). */
int b200c_free_shared(void* dptr);

/* ---- context -------------------------------------------------------------------
 * arena_ptrs[r] is rank r's arena as mapped in THIS process (own arena at
 * [rank]).  Fills the own arena with the "not arrived" pattern and zeroes its
 * flags (host-synchronous), so the caller must run a host barrier across the
 * group between b200c_init and the first collective.  world == 1 is allowed
 * (arena_ptrs may be NULL): only the norm entry points do work.
 * Replaces init_custom_ar + register_buffer (VLLM/_custom_ops.py:2867-2876,
 * 2896-2897; VLLM/distributed/device_communicators/custom_all_reduce.py:175-196).
 */
int b200c_init(int rank, int world, void* const* arena_ptrs, size_t max_bytes, size_t ag_bytes,
               size_t p2p_bytes, b200c_ctx** ctx);

/* Replaces dispose (VLLM/_custom_ops.py:2888-2889). Host-synchronous. */
int b200c_destroy(b200c_ctx* ctx);

/* ---- NVLS (NVSwitch multicast) buffer ----------------------------------------------
 * The B200C_NVLS variant reduces inside the switch, which can only read and write a
 * buffer that is bound to a multicast object on every rank.  The caller provides it
 * (e.g. torch.distributed._symmetric_memory: empty + rendezvous gives both mappings):
 * uc_ptr = this rank's copy through its ordinary mapping, mc_ptr = the multicast mapping
 * of the same offset range, bytes >= b200c_nvls_bytes(max_bytes).  attach arms the
 * rank's own copy (host-synchronous); run a host barrier across the group before the
 * first NVLS call.  The ctx does not own the buffer.
 * Replaces the symmetric-memory set-up of SymmMemCommunicator.__init__
 * (VLLM/distributed/device_communicators/symm_mem.py:88-112) and the multicast
 * workspace of FlashInfer's MNNVL all-reduce (FI/comm/trtllm_mnnvl_ar.py). */
int b200c_nvls_bytes(size_t max_bytes, size_t* out_bytes);
int b200c_attach_nvls(b200c_ctx* ctx, void* uc_ptr, void* mc_ptr, size_t bytes);

int b200c_set_tuning(b200c_ctx* ctx, int key, long long value);
int b200c_get_tuning(b200c_ctx* ctx, int key, long long* value);

/* Host-synchronous read of the sticky device-side error word (0 = healthy,
 * B200C_ETIMEOUT if a wait on a peer timed out; results since then are garbage;
 * B200C_ECUDA if the one-shot cluster kernel found itself running with a cluster
 * shape other than the one it was compiled for and wrote nothing).
 * The reference's barrier spins forever on a dead peer
 * (vllm_custom_all_reduce.cuh:190-211); this library bounds the wait, and with
 * B200C_TUNE_TRAP_ON_TIMEOUT it turns a timeout into a dead CUDA context -- the
 * same "worker is gone" signal the reference launcher acts on
 * (REF/src/launch.py:130-144, 386-388) -- instead of wrong results. */
int b200c_query_error(b200c_ctx* ctx);

/* ---- the hot path -----------------------------------------------------------------
 * All pointers are device pointers on ctx's device; `in` is NOT required to be in
 * the arena (pushed design: ranks read their own input locally and write into the
 * peers' arenas, so no buffer registration is needed, eagerly or under CUDA-graph
 * capture -- contrast register_graph_buffers, VLLM/_custom_ops.py:2900-2903).
 * All calls of one ctx must be issued on one stream at a time, in the same order
 * with the same sizes on every rank.
 */

/* out = sum over ranks of in, fp32 accumulation in rank order 0..N-1, one rounding:
 * bit-identical on every rank.  numel*sizeof(T) must be a multiple of 16, in/out
 * 16-byte aligned, in != out.  Replaces all_reduce (VLLM/_custom_ops.py:2878-2885;
 * kernels cross_device_reduce_1stage/2stage, vllm_custom_all_reduce.cuh:223-288). */
int b200c_all_reduce(b200c_ctx* ctx, const void* in, void* out, size_t numel, int dtype,
                     int variant, void* stream);

/* The fused decoder-layer boundary:
 *   s            = T(sum_r in_r)                       (as b200c_all_reduce)
 *   x32          = f32(s) + f32(residual_in)
 *   residual_out = T(x32)
 *   norm_out     = T( T(x32 * rsqrt(mean(x32^2) + eps)) * gamma )
 * in, residual_in: [tokens, hidden]; gamma: [hidden]; residual_out may alias
 * residual_in, norm_out may alias in.  hidden*sizeof(T) % 16 == 0.
 * world == 1 degenerates to fused_add_rms_norm.
 * Replaces the pair tensor_model_parallel_all_reduce + fused_add_rms_norm
 * (VLLM/model_executor/layers/linear.py:1559 + VLLM/ir/ops/layernorm.py:39-58) and
 * the fused op flashinfer_trtllm_fused_allreduce_norm
 * (VLLM/compilation/passes/fusion/allreduce_rms_fusion.py:121-245;
 * FI/data/include/flashinfer/comm/trtllm_allreduce_fusion.cuh:1336-1467). */
int b200c_all_reduce_add_rmsnorm(b200c_ctx* ctx, const void* in, const void* residual_in,
                                 void* residual_out, void* norm_out, const void* gamma, float eps,
                                 int tokens, int hidden, int dtype, int variant, void* stream);

/* In-place pair used at TP=1 and by the eager (unfused) path:
 *   residual += x ; x = rmsnorm(residual) * gamma.      No ctx state is used.
 * Replaces _C.fused_add_rms_norm (VLLM/_custom_ops.py:322-327, VLLM/kernels/vllm_c.py:54-91). */
int b200c_fused_add_rms_norm(void* x_inout, void* residual_inout, const void* gamma, float eps,
                             int tokens, int hidden, int dtype, void* stream);

/* out = rmsnorm(x) * gamma.  Replaces _C.rms_norm (VLLM/_custom_ops.py:316-319). */
int b200c_rms_norm(void* out, const void* x, const void* gamma, float eps, int tokens, int hidden,
                   int dtype, void* stream);

/* lm_head logits gather: out[row, r*cols + c] = in_r[row, c]   (bit-exact byte move).
 * row_bytes = bytes of one rank's row (cols * sizeof(T)), a multiple of 4: 16-byte vectors when
 * it allows, 8- or 4-byte words otherwise (the [batch, 2] fp32 pairs of vLLM's vocab-parallel
 * argmax, VLLM/model_executor/layers/logits_processor.py:106-156, travel this way).
 * Replaces tensor_model_parallel_all_gather(dim=-1)
 * (VLLM/distributed/device_communicators/base_device_communicator.py:184-207,
 * called from VLLM/model_executor/layers/logits_processor.py:75-87). */
int b200c_all_gather_lastdim(b200c_ctx* ctx, const void* in, void* out, int rows, size_t row_bytes,
                             void* stream);

/* Pipeline-stage hop: stream-ordered send of `bytes` (multiple of 16, <= p2p_bytes)
 * to group rank dst / receive from group rank src.  Sends and receives between a
 * pair must match in order.  Replaces CudaCommunicator.send/recv (NCCL p2p,
 * VLLM/distributed/device_communicators/cuda_communicator.py send/recv) under
 * GroupCoordinator.isend/irecv_tensor_dict (VLLM/distributed/parallel_state.py:871-1050). */
int b200c_send(b200c_ctx* ctx, int dst, const void* buf, size_t bytes, void* stream);
int b200c_recv(b200c_ctx* ctx, int src, void* buf, size_t bytes, void* stream);

/* ---- single-GPU rank emulation (bring-up and the 1-GPU parity tests) ----------------
 * No counterpart in the reference.  All `world` ranks live on ONE device (ctxs[r]
 * built by b200c_init with rank r over the same arena_ptrs); ONE cooperative
 * launch runs every rank's CTAs of the SAME device code as the real entry points
 * (blockIdx.y selects the rank), so ranks that wait on each other are guaranteed
 * co-resident.  Arrays are indexed by rank.
 */
int b200c_emu_all_reduce(b200c_ctx* const* ctxs, int world, const void* const* ins,
                         void* const* outs, size_t numel, int dtype, int variant, void* stream);
int b200c_emu_all_reduce_add_rmsnorm(b200c_ctx* const* ctxs, int world, const void* const* ins,
                                     const void* const* residual_ins, void* const* residual_outs,
                                     void* const* norm_outs, const void* gamma, float eps,
                                     int tokens, int hidden, int dtype, int variant, void* stream);
int b200c_emu_all_gather_lastdim(b200c_ctx* const* ctxs, int world, const void* const* ins,
                                 void* const* outs, int rows, size_t row_bytes, void* stream);
/* one matched send (on ctxs[src]) + recv (on ctxs[dst]) */
int b200c_emu_sendrecv(b200c_ctx* const* ctxs, int world, int src, int dst, const void* sendbuf,
                       void* recvbuf, size_t bytes, void* stream);

/* ---- the one entry point that allocates nothing ---------------------------------------------------
 * No counterpart in the reference.  Same result as b200c_emu_all_reduce, for callers that
 * must not allocate device memory when they call: code running inside a serving process after
 * the engine has given all free HBM to its KV-cache pool (a cudaMalloc can fail, and it
 * escapes the engine's memory accounting and its sleep-mode allocator), or while another
 * thread is capturing a CUDA graph (cudaMalloc is illegal then).  It needs no ctx and no
 * arenas: all scratch (every emulated rank's receive regions and state) lives in a
 * file-scope device array of about 3.6 MiB that each GPU loading this library holds until
 * the process exits.
 *   - limit: world <= B200C_MAX_RANKS and numel*sizeof(T) <= B200C_NOALLOC_MAX_BYTES per
 *     rank (one-shot additionally needs world*bytes <= 2*B200C_NOALLOC_MAX_BYTES); anything
 *     larger returns B200C_ETOOBIG and launches nothing;
 *   - the first call on a device resolves the array's device address; make it before any
 *     stream capture begins;
 *   - every call re-arms the shared array on `stream` before it runs, because the array
 *     keeps the previous call's contents; calls whose GPU work may overlap (different
 *     streams) race on it and must be serialised by the caller;
 *   - it calls no cudaMalloc, cudaFree or other allocating routine, on any call. */
#define B200C_NOALLOC_MAX_BYTES (64 * 1024)
int b200c_emu_all_reduce_noalloc(int world, const void* const* ins, void* const* outs,
                                 size_t numel, int dtype, int variant, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* B200C_H_ */
