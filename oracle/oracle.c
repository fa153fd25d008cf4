/*
 * oracle.c -- CPU restatement of the reference's TP hot path.
 *
 * TEST INFRASTRUCTURE ONLY.  Nothing under oracle/ is on the product path:
 * only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / reference
 * legs may load this library, and only as the checker or the timed CPU
 * baseline.  The product (vllm_distributed_b200/) never imports it.
 *
 * Parity status: PINNED.  The functions below are checked against golden
 * vectors produced by running the reference's own dependency code (vLLM
 * 0.22.0, unmodified, gloo backend) in the build container -- see
 * oracle/gen_golden.py and tests/test_oracle_golden.py.  The reference repo
 * itself (koush/vllm-distributed) holds no arithmetic and no tests; its call
 * sites for this path are /root/reference/src/launch.py:290-292 (bring-up),
 * :322-349 (execute_model) and :524-538 (run_worker dispatch).
 *
 * VLLM/ = site-packages/vllm (0.22.0), FI/ = site-packages/flashinfer.
 *
 * All 16-bit floats travel as raw uint16_t bit patterns.
 */
#include <math.h>
#include <stddef.h>
#include <stdint.h>
#include <string.h>

#ifdef _OPENMP
#include <omp.h>
#endif

/* ---- scalar conversions -------------------------------------------------- */

static inline float bf16_to_f32(uint16_t h) {
  uint32_t u = (uint32_t)h << 16;
  float f;
  memcpy(&f, &u, 4);
  return f;
}

/* round-to-nearest-even, NaN kept quiet: same rule torch and __float2bfloat16_rn use */
static inline uint16_t f32_to_bf16(float f) {
  uint32_t u;
  memcpy(&u, &f, 4);
  if ((u & 0x7fffffffu) > 0x7f800000u) return (uint16_t)((u >> 16) | 0x0040u);
  u += 0x7fffu + ((u >> 16) & 1u);
  return (uint16_t)(u >> 16);
}

static inline float f16_to_f32(uint16_t h) {
  _Float16 v;
  memcpy(&v, &h, 2);
  return (float)v;
}

static inline uint16_t f32_to_f16(float f) {
  _Float16 v = (_Float16)f; /* RNE */
  uint16_t h;
  memcpy(&h, &v, 2);
  return h;
}

enum { DT_BF16 = 0, DT_F16 = 1, DT_F32 = 2 };

static inline float load_elt(const void* p, size_t i, int dt) {
  switch (dt) {
    case DT_BF16: return bf16_to_f32(((const uint16_t*)p)[i]);
    case DT_F16: return f16_to_f32(((const uint16_t*)p)[i]);
    default: return ((const float*)p)[i];
  }
}

static inline void store_elt(void* p, size_t i, int dt, float v) {
  switch (dt) {
    case DT_BF16: ((uint16_t*)p)[i] = f32_to_bf16(v); break;
    case DT_F16: ((uint16_t*)p)[i] = f32_to_f16(v); break;
    default: ((float*)p)[i] = v; break;
  }
}

/* ---- a5: all-reduce --------------------------------------------------------
 * Follows vLLM's custom all-reduce reduction rule, packed_reduce():
 *   FI/data/include/flashinfer/comm/vllm_custom_all_reduce.cuh:213-221
 *   (tmp = upcast(ptrs[0][i]); for r in 1..N-1: tmp += upcast(ptrs[r][i]);
 *    out = downcast(tmp)) -- fp32 accumulation in FIXED rank order 0..N-1, one
 *   rounding at the end, so every rank holds the bit-identical result
 *   (cross_device_reduce_1stage, :223-238).
 */
int oracle_all_reduce(const void* const* xs, int world, size_t numel, int dtype, void* out) {
  if (world < 1 || world > 64) return -1;
#pragma omp parallel for schedule(static)
  for (size_t i = 0; i < numel; ++i) {
    float acc = load_elt(xs[0], i, dtype);
    for (int r = 1; r < world; ++r) acc += load_elt(xs[r], i, dtype);
    store_elt(out, i, dtype, acc);
  }
  return 0;
}

/* ---- a6: fused residual add + RMSNorm --------------------------------------
 * Follows the semantic (native) definition VLLM/ir/ops/layernorm.py:39-58:
 *   x32 = f32(x) + f32(residual)
 *   residual_out = T(x32)
 *   var = mean(x32^2)                    (on the UNROUNDED fp32 sum)
 *   y = T( T(x32 * rsqrt(var + eps)) * gamma )
 * The row sum of squares is accumulated in double and rounded to fp32 once;
 * torch's CPU reduction order is an implementation detail, so norm_out is
 * compared under vLLM's own tolerance (VLLM/ir/tolerances.py:12-36), while
 * residual_out involves no row reduction and is bit-exact.
 * residual may be NULL (plain rms_norm, layernorm.py:10-21).  gamma may be NULL.
 */
int oracle_fused_add_rms_norm(const void* x, const void* residual, const void* gamma, float eps,
                              int tokens, int hidden, int dtype, void* norm_out,
                              void* residual_out) {
#pragma omp parallel for schedule(static)
  for (int t = 0; t < tokens; ++t) {
    size_t base = (size_t)t * (size_t)hidden;
    double ss = 0.0;
    for (int h = 0; h < hidden; ++h) {
      float v = load_elt(x, base + h, dtype);
      if (residual) v = v + load_elt(residual, base + h, dtype);
      ss += (double)v * (double)v;
    }
    float var = (float)(ss / (double)hidden);
    float rs = 1.0f / sqrtf(var + eps); /* torch CPU rsqrt = 1/sqrt, IEEE */
    for (int h = 0; h < hidden; ++h) {
      float v = load_elt(x, base + h, dtype);
      if (residual) {
        v = v + load_elt(residual, base + h, dtype);
        if (residual_out) store_elt(residual_out, base + h, dtype, v);
      }
      float n = v * rs;
      if (gamma) {
        /* x.to(weight.dtype) * weight : round, multiply (fp32 inside torch), round */
        if (dtype == DT_BF16) n = bf16_to_f32(f32_to_bf16(n));
        else if (dtype == DT_F16) n = f16_to_f32(f32_to_f16(n));
        n = n * load_elt(gamma, (size_t)h, dtype);
      }
      store_elt(norm_out, base + h, dtype, n);
    }
  }
  return 0;
}

/* ---- a1+a6 chained: what one decoder-layer boundary computes ---------------
 * RowParallelLinear.forward -> tensor_model_parallel_all_reduce
 *   (VLLM/model_executor/layers/linear.py:1540-1566, :1559), then
 * RMSNorm(h, residual) -> fused_add_rms_norm (VLLM/model_executor/models/llama.py:316-333).
 * The all-reduce result is rounded to T before the add, exactly as the unfused
 * reference does (the AR returns a T tensor).
 */
int oracle_all_reduce_add_rms_norm(const void* const* xs, int world, const void* residual,
                                   const void* gamma, float eps, int tokens, int hidden, int dtype,
                                   void* scratch_ar, void* norm_out, void* residual_out) {
  int rc = oracle_all_reduce(xs, world, (size_t)tokens * (size_t)hidden, dtype, scratch_ar);
  if (rc) return rc;
  return oracle_fused_add_rms_norm(scratch_ar, residual, gamma, eps, tokens, hidden, dtype,
                                   norm_out, residual_out);
}

/* ---- a10: all-gather along the last dim --------------------------------------
 * VLLM/distributed/device_communicators/base_device_communicator.py:184-207:
 * dim-0 gather, movedim(0, dim), reshape  ==  out[row, r*cols + c] = slab_r[row, c].
 * Pure byte movement: bit-exact for any payload.
 */
int oracle_all_gather_lastdim(const void* const* slabs, int world, int rows, int cols,
                              int elt_bytes, void* out) {
  size_t rb = (size_t)cols * (size_t)elt_bytes;
#pragma omp parallel for schedule(static)
  for (int row = 0; row < rows; ++row)
    for (int r = 0; r < world; ++r)
      memcpy((char*)out + ((size_t)row * world + r) * rb, (const char*)slabs[r] + (size_t)row * rb,
             rb);
  return 0;
}

/* ---- a12: PP hop with TP slicing ---------------------------------------------
 * VLLM/distributed/parallel_state.py:913-916 (send side) and :1014-1038 (recv
 * side): when numel % tp == 0 each TP rank t sends tensor.reshape(tp, -1)[t];
 * the receiver of the same TP rank gets that slice and the stage re-assembles
 * the tensor with a dim-0 all-gather.  Byte movement only.
 */
int oracle_pp_slice(const void* full, size_t bytes, int tp, int tp_rank, void* slice) {
  if (tp < 1 || bytes % (size_t)tp) return -1;
  size_t sb = bytes / (size_t)tp;
  memcpy(slice, (const char*)full + sb * (size_t)tp_rank, sb);
  return 0;
}

int oracle_pp_join(const void* const* slices, size_t bytes, int tp, void* full) {
  if (tp < 1 || bytes % (size_t)tp) return -1;
  size_t sb = bytes / (size_t)tp;
  for (int t = 0; t < tp; ++t) memcpy((char*)full + sb * (size_t)t, slices[t], sb);
  return 0;
}

int oracle_num_threads(void) {
#ifdef _OPENMP
  return omp_get_max_threads();
#else
  return 1;
#endif
}

void oracle_set_num_threads(int n) {
#ifdef _OPENMP
  omp_set_num_threads(n);
#else
  (void)n;
#endif
}
