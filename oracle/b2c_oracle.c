/* oracle/b2c_oracle.c — TEST INFRASTRUCTURE, NOT PRODUCT.
 *
 * CPU restatement of the arithmetic on ray_lightning's gradient-sync path.  The
 * reference (/root/reference/ray_lightning) owns no arithmetic; the path resolves to
 * third-party code that is not under /root/reference:
 *   - torch c10d Reducer + ProcessGroupGloo/NCCL (torch unpinned in the reference,
 *     setup.py:12; 2.11.0 installed here), entered from ray_ddp.py:218-219 and
 *     ray_ddp.py:232-237;
 *   - Horovod (git master, .github/workflows/test.yaml:39), entered from
 *     ray_horovod.py:149-159.  Horovod is absent from this image: its semantics are
 *     restated from its published behaviour (Average = sum / size).
 * Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline leg may use this file.
 *
 * PINNING: the reference has no golden vectors for this path.  This
 * oracle is pinned against outputs of the real upstream implementation run in this
 * container — torch.distributed gloo all_reduce / broadcast / DDP — stored under
 * tests/golden/ by tests/golden/make_golden.py (tests/test_oracle.py checks them).
 *
 * Build: make -C oracle   (gcc only; output oracle/_build/liboracle.so)
 */
#include <math.h>
#include <stddef.h>
#include <stdint.h>
#include <string.h>

/* ---- bf16 helpers (round-to-nearest-even, NaN kept quiet) ---------------- */
static inline float bf16_to_f32(uint16_t h) {
  uint32_t u = (uint32_t)h << 16;
  float f;
  memcpy(&f, &u, 4);
  return f;
}
static inline uint16_t f32_to_bf16_rne(float f) {
  uint32_t u;
  memcpy(&u, &f, 4);
  if ((u & 0x7fffffffu) > 0x7f800000u) return 0x7fff; /* NaN -> canonical (cvt.rn.bf16.f32) */
  uint32_t lsb = (u >> 16) & 1u;
  u += 0x7fffu + lsb;
  return (uint16_t)(u >> 16);
}

/* ---- C3, reference semantics -------------------------------------------------
 * DDP default comm hook: torch:distributed/algorithms/ddp_comm_hooks/default_hooks.py:18-32
 *   tensor.div_(world_size); all_reduce(tensor, SUM)
 * (without a hook the Reducer pre-multiplies by 1/W while copying into the bucket,
 *  torch:nn/parallel/distributed.py:2021-2023 — same arithmetic).
 * Summation order inside gloo/NCCL is not rank order in general; `order` lets a test
 * pick the order (NULL = 0..W-1).  For W = 2 every order gives the same bits. */
void oracle_ddp_allreduce_f32(const float* const* g, int W, size_t n, const int* order, float* out) {
  for (size_t i = 0; i < n; ++i) {
    int r0 = order ? order[0] : 0;
    float acc = g[r0][i] / (float)W;
    for (int k = 1; k < W; ++k) {
      int r = order ? order[k] : k;
      float t = g[r][i] / (float)W;
      acc = acc + t;
    }
    out[i] = acc;
  }
}

/* bf16 gradients through the same hook: div_ rounds to bf16, every pairwise add rounds to bf16 */
void oracle_ddp_allreduce_bf16(const uint16_t* const* g, int W, size_t n, const int* order, uint16_t* out) {
  for (size_t i = 0; i < n; ++i) {
    int r0 = order ? order[0] : 0;
    uint16_t acc = f32_to_bf16_rne(bf16_to_f32(g[r0][i]) / (float)W);
    for (int k = 1; k < W; ++k) {
      int r = order ? order[k] : k;
      uint16_t t = f32_to_bf16_rne(bf16_to_f32(g[r][i]) / (float)W);
      acc = f32_to_bf16_rne(bf16_to_f32(acc) + bf16_to_f32(t));
    }
    out[i] = acc;
  }
}

/* ---- C6, Horovod Average (ray_horovod.py:159 -> hvd.allreduce_async_(op=Average)):
 * sum over ranks then divide by size. */
void oracle_hvd_average_f32(const float* const* g, int W, size_t n, float* out) {
  for (size_t i = 0; i < n; ++i) {
    float acc = g[0][i];
    for (int r = 1; r < W; ++r) acc = acc + g[r][i];
    out[i] = acc / (float)W;
  }
}

/* ---- the contract the CUDA kernels implement (include/b2c.h b2c_allreduce):
 * every term is multiplied by `scale` first (its own rounding, exactly tensor.div_(W) when
 * W is a power of two — same overflow / underflow behaviour as the reference), then the
 * terms are accumulated in fp32 in rank order 0..W-1; one final rounding to the output type. */
void oracle_b2c_allreduce_f32(const float* const* g, int W, size_t n, float scale, float* out) {
  for (size_t i = 0; i < n; ++i) {
    float acc = g[0][i] * scale;
    for (int r = 1; r < W; ++r) {
      float t = g[r][i] * scale;
      acc = acc + t;
    }
    out[i] = acc;
  }
}
void oracle_b2c_allreduce_bf16(const uint16_t* const* g, int W, size_t n, float scale, uint16_t* out) {
  for (size_t i = 0; i < n; ++i) {
    float acc = bf16_to_f32(g[0][i]) * scale;
    for (int r = 1; r < W; ++r) {
      float t = bf16_to_f32(g[r][i]) * scale;
      acc = acc + t;
    }
    out[i] = f32_to_bf16_rne(acc);
  }
}
/* fp32 gradients on a bf16 wire (bf16_compress_hook analogue, default_hooks.py:57-93):
 * cast to bf16, scale each term, accumulate fp32, round to bf16, widen to fp32 */
void oracle_b2c_allreduce_f32_bf16wire(const float* const* g, int W, size_t n, float scale, float* out) {
  for (size_t i = 0; i < n; ++i) {
    float acc = bf16_to_f32(f32_to_bf16_rne(g[0][i])) * scale;
    for (int r = 1; r < W; ++r) {
      float t = bf16_to_f32(f32_to_bf16_rne(g[r][i])) * scale;
      acc = acc + t;
    }
    out[i] = bf16_to_f32(f32_to_bf16_rne(acc));
  }
}

/* exact answer in fp64: removes "whose rounding is right" from tolerance checks */
void oracle_exact_mean_f64(const float* const* g, int W, size_t n, double* out) {
  for (size_t i = 0; i < n; ++i) {
    double acc = 0.0;
    for (int r = 0; r < W; ++r) acc += (double)g[r][i];
    out[i] = acc / (double)W;
  }
}
void oracle_exact_mean_bf16_f64(const uint16_t* const* g, int W, size_t n, double* out) {
  for (size_t i = 0; i < n; ++i) {
    double acc = 0.0;
    for (int r = 0; r < W; ++r) acc += (double)bf16_to_f32(g[r][i]);
    out[i] = acc / (double)W;
  }
}

/* ---- C2 / C4 / C5: broadcast is a byte copy of root's buffer
 * (_sync_module_states -> _broadcast_coalesced, torch:distributed/utils.py:289-328) */
void oracle_broadcast(const unsigned char* root_buf, size_t bytes, unsigned char* const* outs, int W) {
  for (int r = 0; r < W; ++r)
    if (outs[r] != root_buf) memcpy(outs[r], root_buf, bytes);
}

void oracle_f32_to_bf16(const float* in, size_t n, uint16_t* out) {
  for (size_t i = 0; i < n; ++i) out[i] = f32_to_bf16_rne(in[i]);
}
void oracle_bf16_to_f32(const uint16_t* in, size_t n, float* out) {
  for (size_t i = 0; i < n; ++i) out[i] = bf16_to_f32(in[i]);
}
