// b2c device code: cross-GPU reduce / broadcast kernels for sm_100a.
//
// One symmetric region per rank (exported with CUDA IPC, mapped by every peer):
//
//   [0, 4096)        control words (local use only): epoch, call counter, exit ticket
//   [4096, 12288)    flag page: flags[block][src_rank] (u32, monotonic epochs,
//                    written remotely by `src_rank`, spun on locally)
//   [16384, +slot)   arena slot 0   } double buffered by call parity, so a rank that
//   [.., +slot)      arena slot 1   } runs one call ahead never clobbers data a peer still reads
//
// Every CTA is an independent channel: CTA b of rank r only ever exchanges data
// with CTA b of the peers (same chunk -> CTA mapping on every rank), so the only
// cross-GPU synchronisation is a per-CTA flag barrier; there is no grid-wide sync.
//
// Data that another GPU may have written is always read with ld.relaxed.sys
// (never the non-coherent path, never a possibly-stale L1 line); flags use
// st.release.sys / ld.acquire.sys.  Sums are accumulated in fp32 in FIXED rank
// order 0..W-1 on every rank, so replicas stay bitwise identical.
#pragma once
#include <cuda_runtime.h>
#include <cuda_bf16.h>
#include <stdint.h>

#include "../../include/b2c.h"

#define B2C_CTL_BYTES 4096
#define B2C_FLAG_BYTES 8192
#define B2C_HEADER_BYTES 16384
#define B2C_CTL_EPOCH 0
#define B2C_CTL_CALLS 1
#define B2C_CTL_TICKET 2
#define B2C_CTL_PHASE_SEQ 3   /* kernels that stamped the phase ring so far (channel 0 only) */
#define B2C_CTL_PHASE_U64 8   /* u64 index: start, after barrier A, after barrier B, end (channel 0) of the LAST kernel */
#define B2C_CTL_MC_EPOCH 64   /* u32 index: per-channel epoch of the multicast flag counters (256 words) */
#define B2C_CTL_RING_BYTE 2048 /* byte offset: ring of the last B2C_PHASE_RING kernels' phase stamps (4 x u64 each) */
#define B2C_PHASE_RING 32
/* head of the multicast arena: one 128-byte line per channel holding a monotonic arrival counter that
 * every rank bumps with ONE multimem.red (the switch adds it into all W copies) and spins on locally */
#define B2C_MC_FLAG_STRIDE 128
#define B2C_MC_FLAG_BYTES (B2C_MAX_BLOCKS * B2C_MC_FLAG_STRIDE)
#define B2C_MC_HEAD_BYTES 65536

static_assert(B2C_CTL_MC_EPOCH * 4 + B2C_MAX_BLOCKS * 4 <= B2C_CTL_RING_BYTE && B2C_CTL_RING_BYTE + B2C_PHASE_RING * 32 <= B2C_CTL_BYTES, "control page layout");
static_assert(B2C_MC_FLAG_BYTES <= B2C_MC_HEAD_BYTES, "multicast flag head too small");
static_assert(B2C_MAX_BLOCKS * B2C_MAX_WORLD * 4 <= B2C_FLAG_BYTES && B2C_CTL_BYTES + B2C_FLAG_BYTES <= B2C_HEADER_BYTES, "flag page too small");

enum { MODE_F32 = 0, MODE_BF16 = 1, MODE_F32_BF16 = 2, MODE_U8 = 3 };

struct B2cKArgs {
  void* buf[B2C_MAX_WORLD];        // caller buffer per rank (real mode: only [rank] is valid)
  char* region[B2C_MAX_WORLD];     // rank q's symmetric region as mapped in this process
  uint32_t* status[B2C_MAX_WORLD]; // sticky error words (host-mapped) per rank (real: only [rank])
  size_t count;                    // caller elements (bytes for MODE_U8)
  size_t slot_bytes;
  unsigned long long timeout_ns;
  float scale;
  int rank;                        // >= 0: this launch is one rank; -1: rank = blockIdx.y (local group)
  int world;
  int root;
  int variant;
  uint32_t chunk_vecs;             // ownership granularity (16-byte vectors), multiple of blockDim.x
  // NVLS path only
  char* nvls_uc;                   // local unicast view of the multicast arena
  char* nvls_mc;                   // multicast view (multimem.* only)
  size_t nvls_slot_bytes;          // staged mode: arena is two slots, picked by call parity
  size_t nvls_off;                 // in-place: offset of the caller's tensor inside the arena; staged: of slot 0
  int nvls_inplace;
};

// ---------------------------------------------------------------- PTX helpers
__device__ __forceinline__ void st_release_sys(uint32_t* p, uint32_t v) {
  asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ uint32_t ld_acquire_sys(const uint32_t* p) {
  uint32_t v;
  asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ uint32_t ld_volatile_u32(const uint32_t* p) {
  uint32_t v;
  asm volatile("ld.volatile.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
// 16-byte load that is coherent with writes from other GPUs (bypasses L1)
__device__ __forceinline__ uint4 ld_sys_v4(const void* p) {
  uint4 v;
  asm volatile("ld.relaxed.sys.global.v4.u32 {%0,%1,%2,%3}, [%4];"
               : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w)
               : "l"(p));
  return v;
}
// 16-byte store, no L1 allocation (streaming; may target a peer GPU)
__device__ __forceinline__ void st_v4(void* p, uint4 v) {
  asm volatile("st.global.L1::no_allocate.v4.u32 [%0], {%1,%2,%3,%4};" ::"l"(p), "r"(v.x),
               "r"(v.y), "r"(v.z), "r"(v.w)
               : "memory");
}
__device__ __forceinline__ unsigned long long globaltimer_ns() {
  unsigned long long t;
  asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
  return t;
}

// ---------------------------------------------------------------- per-mode traits
template <int MODE> struct Mode;
template <> struct Mode<MODE_F32> { static constexpr int VW = 4; static constexpr int UB = 4; };
template <> struct Mode<MODE_BF16> { static constexpr int VW = 8; static constexpr int UB = 2; };
template <> struct Mode<MODE_F32_BF16> { static constexpr int VW = 8; static constexpr int UB = 4; };
template <> struct Mode<MODE_U8> { static constexpr int VW = 16; static constexpr int UB = 1; };

__device__ __forceinline__ uint32_t pack_bf16x2(float lo, float hi) {
  __nv_bfloat162 h = __floats2bfloat162_rn(lo, hi);  // round-to-nearest-even, once
  return *reinterpret_cast<uint32_t*>(&h);
}
__device__ __forceinline__ float bf16_lo(uint32_t u) { return __uint_as_float(u << 16); }
__device__ __forceinline__ float bf16_hi(uint32_t u) { return __uint_as_float(u & 0xffff0000u); }

// caller buffer -> one 16-byte wire vector (vector g covers caller elements [g*VW, g*VW+VW))
template <int MODE>
__device__ __forceinline__ uint4 load_user(const void* buf, size_t g, size_t count, bool fast) {
  constexpr int VW = Mode<MODE>::VW;
  const size_t e0 = g * VW;
  if (fast && e0 + VW <= count) {
    if (MODE == MODE_F32_BF16) {
      const uint4* p = reinterpret_cast<const uint4*>(static_cast<const float*>(buf) + e0);
      uint4 a = p[0], b = p[1];
      uint4 r;
      r.x = pack_bf16x2(__uint_as_float(a.x), __uint_as_float(a.y));
      r.y = pack_bf16x2(__uint_as_float(a.z), __uint_as_float(a.w));
      r.z = pack_bf16x2(__uint_as_float(b.x), __uint_as_float(b.y));
      r.w = pack_bf16x2(__uint_as_float(b.z), __uint_as_float(b.w));
      return r;
    } else {
      return *reinterpret_cast<const uint4*>(static_cast<const char*>(buf) + e0 * Mode<MODE>::UB);
    }
  }
  // ragged tail or misaligned base: element by element, zero filled past `count`
  uint32_t w[4] = {0u, 0u, 0u, 0u};
  if (MODE == MODE_F32) {
    const float* p = static_cast<const float*>(buf);
#pragma unroll
    for (int k = 0; k < 4; ++k)
      if (e0 + k < count) w[k] = __float_as_uint(p[e0 + k]);
  } else if (MODE == MODE_BF16) {
    const unsigned short* p = static_cast<const unsigned short*>(buf);
#pragma unroll
    for (int k = 0; k < 8; ++k)
      if (e0 + k < count) w[k >> 1] |= (uint32_t)p[e0 + k] << ((k & 1) * 16);
  } else if (MODE == MODE_F32_BF16) {
    const float* p = static_cast<const float*>(buf);
    float f[8];
#pragma unroll
    for (int k = 0; k < 8; ++k) f[k] = (e0 + k < count) ? p[e0 + k] : 0.f;
#pragma unroll
    for (int k = 0; k < 4; ++k) w[k] = pack_bf16x2(f[2 * k], f[2 * k + 1]);
  } else {
    const unsigned char* p = static_cast<const unsigned char*>(buf);
#pragma unroll
    for (int k = 0; k < 16; ++k)
      if (e0 + k < count) w[k >> 2] |= (uint32_t)p[e0 + k] << ((k & 3) * 8);
  }
  return make_uint4(w[0], w[1], w[2], w[3]);
}

// one 16-byte wire vector -> caller buffer (never writes past `count`)
template <int MODE>
__device__ __forceinline__ void store_user(void* buf, size_t g, size_t count, bool fast, uint4 v) {
  constexpr int VW = Mode<MODE>::VW;
  const size_t e0 = g * VW;
  if (fast && e0 + VW <= count) {
    if (MODE == MODE_F32_BF16) {
      uint4* p = reinterpret_cast<uint4*>(static_cast<float*>(buf) + e0);
      p[0] = make_uint4(v.x << 16, v.x & 0xffff0000u, v.y << 16, v.y & 0xffff0000u);
      p[1] = make_uint4(v.z << 16, v.z & 0xffff0000u, v.w << 16, v.w & 0xffff0000u);
    } else {
      *reinterpret_cast<uint4*>(static_cast<char*>(buf) + e0 * Mode<MODE>::UB) = v;
    }
    return;
  }
  const uint32_t w[4] = {v.x, v.y, v.z, v.w};
  if (MODE == MODE_F32) {
    float* p = static_cast<float*>(buf);
#pragma unroll
    for (int k = 0; k < 4; ++k)
      if (e0 + k < count) p[e0 + k] = __uint_as_float(w[k]);
  } else if (MODE == MODE_BF16) {
    unsigned short* p = static_cast<unsigned short*>(buf);
#pragma unroll
    for (int k = 0; k < 8; ++k)
      if (e0 + k < count) p[e0 + k] = (unsigned short)(w[k >> 1] >> ((k & 1) * 16));
  } else if (MODE == MODE_F32_BF16) {
    float* p = static_cast<float*>(buf);
#pragma unroll
    for (int k = 0; k < 8; ++k)
      if (e0 + k < count)
        p[e0 + k] = (k & 1) ? bf16_hi(w[k >> 1]) : bf16_lo(w[k >> 1]);
  } else {
    unsigned char* p = static_cast<unsigned char*>(buf);
#pragma unroll
    for (int k = 0; k < 16; ++k)
      if (e0 + k < count) p[e0 + k] = (unsigned char)(w[k >> 2] >> ((k & 3) * 8));
  }
}

// fp32 accumulator over one wire vector.  Every term is scaled first (its own rounding:
// exactly tensor.div_(W) of the reference hook when W is a power of two, so overflow /
// underflow behave like the reference), terms are added in the order given (rank order),
// and the result is rounded once to the wire type.  No FMA contraction (__fmul_rn/__fadd_rn).
template <int MODE> struct Acc {
  static constexpr int VW = Mode<MODE>::VW;
  float a[VW];
  __device__ __forceinline__ void unpack(uint4 v, float* f) const {
    if (MODE == MODE_F32) {
      f[0] = __uint_as_float(v.x); f[1] = __uint_as_float(v.y);
      f[2] = __uint_as_float(v.z); f[3] = __uint_as_float(v.w);
    } else {
      const uint32_t w[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
      for (int k = 0; k < 4; ++k) { f[2 * k] = bf16_lo(w[k]); f[2 * k + 1] = bf16_hi(w[k]); }
    }
  }
  __device__ __forceinline__ void init(uint4 v, float scale) {  // first term: no add to +0 (keeps -0)
    float f[VW];
    unpack(v, f);
#pragma unroll
    for (int k = 0; k < VW; ++k) a[k] = __fmul_rn(f[k], scale);
  }
  __device__ __forceinline__ void add(uint4 v, float scale) {
    float f[VW];
    unpack(v, f);
#pragma unroll
    for (int k = 0; k < VW; ++k) a[k] = __fadd_rn(a[k], __fmul_rn(f[k], scale));
  }
  __device__ __forceinline__ uint4 finish() const {
    if (MODE == MODE_F32) {
      return make_uint4(__float_as_uint(a[0]), __float_as_uint(a[1]), __float_as_uint(a[2]), __float_as_uint(a[3]));
    } else {
      return make_uint4(pack_bf16x2(a[0], a[1]), pack_bf16x2(a[2], a[3]),
                        pack_bf16x2(a[4], a[5]), pack_bf16x2(a[6], a[7]));
    }
  }
};

// ---------------------------------------------------------------- per-CTA context
struct Ctx {
  int rank;
  int b;            // channel (blockIdx.x)
  int nb;           // channels (gridDim.x)
  uint32_t epoch;   // flags reach epoch+1 / epoch+2 in this call
  uint32_t calls;
  size_t slot_off;  // byte offset of the active arena slot in every region
  uint32_t* ctl;
  bool dead;        // a barrier timed out: skip the remaining phases
};

__device__ __forceinline__ int* dead_flag() {
  __shared__ int s_dead;  // sticky for the whole launch; zeroed once in prologue()
  return &s_dead;
}

__device__ __forceinline__ Ctx prologue(const B2cKArgs& a) {
  Ctx c;
  if (threadIdx.x == 0) *dead_flag() = 0;
  __syncthreads();
  c.rank = a.rank >= 0 ? a.rank : (int)blockIdx.y;
  c.b = blockIdx.x;
  c.nb = gridDim.x;
  c.ctl = reinterpret_cast<uint32_t*>(a.region[c.rank]);
  c.epoch = ld_volatile_u32(c.ctl + B2C_CTL_EPOCH);
  c.calls = ld_volatile_u32(c.ctl + B2C_CTL_CALLS);
  c.slot_off = (size_t)B2C_HEADER_BYTES + (size_t)(c.calls & 1u) * a.slot_bytes;
  c.dead = false;
  if (blockIdx.x == 0 && threadIdx.x == 0)  // phase clock of channel 0 (read back by b2c_get_phase_ns)
    reinterpret_cast<unsigned long long*>(c.ctl)[B2C_CTL_PHASE_U64] = globaltimer_ns();
  return c;
}

// channel 0, one thread, at the end of a kernel: keep the stamps of the last B2C_PHASE_RING kernels so
// that a training step can be attributed bucket by bucket afterwards without any host sync in between
__device__ __forceinline__ void phase_commit(uint32_t* ctl) {
  unsigned long long* last = reinterpret_cast<unsigned long long*>(ctl) + B2C_CTL_PHASE_U64;
  const uint32_t seq = ld_volatile_u32(ctl + B2C_CTL_PHASE_SEQ);
  unsigned long long* slot = reinterpret_cast<unsigned long long*>(reinterpret_cast<char*>(ctl) + B2C_CTL_RING_BYTE) +
                             (size_t)(seq % B2C_PHASE_RING) * 4;
  slot[0] = last[0]; slot[1] = last[1]; slot[2] = last[2]; slot[3] = last[3];
  ctl[B2C_CTL_PHASE_SEQ] = seq + 1u;
}

// last CTA out advances the epoch / call counters for the next launch on this stream
__device__ __forceinline__ void epilogue(const Ctx& c) {
  __syncthreads();
  if (blockIdx.x == 0 && threadIdx.x == 0) {
    reinterpret_cast<unsigned long long*>(c.ctl)[B2C_CTL_PHASE_U64 + 3] = globaltimer_ns();
    phase_commit(c.ctl);
  }
  if (threadIdx.x == 0) {
    __threadfence();
    unsigned t = atomicAdd(c.ctl + B2C_CTL_TICKET, 1u);
    if (t == (unsigned)c.nb - 1u) {
      c.ctl[B2C_CTL_TICKET] = 0u;
      c.ctl[B2C_CTL_EPOCH] = c.epoch + 2u;
      c.ctl[B2C_CTL_CALLS] = c.calls + 1u;
      __threadfence();
    }
  }
}

// Per-channel cross-GPU barrier.  Thread q (< world) publishes `target` into peer q's
// flag page (release: orders every write this CTA made before the __syncthreads) and
// then spins on the word peer q writes into OUR page (acquire).  Bounded spin: on
// timeout the sticky status word is set and the CTA skips the rest of the call.
__device__ __forceinline__ void channel_barrier(const B2cKArgs& a, Ctx& c, uint32_t target) {
  int* s_dead = dead_flag();
  __syncthreads();
  if ((int)threadIdx.x < a.world) {
    const int q = threadIdx.x;
    uint32_t* theirs = reinterpret_cast<uint32_t*>(a.region[q] + B2C_CTL_BYTES) + c.b * B2C_MAX_WORLD + c.rank;
    const uint32_t* mine = reinterpret_cast<const uint32_t*>(a.region[c.rank] + B2C_CTL_BYTES) + c.b * B2C_MAX_WORLD + q;
    st_release_sys(theirs, target);
    unsigned long long t0 = 0;
    unsigned spins = 0;
    while ((int)(ld_acquire_sys(mine) - target) < 0) {
      if ((++spins & 0xffu) == 0u) {
        unsigned long long now = globaltimer_ns();
        if (t0 == 0) t0 = now;
        else if (now - t0 > a.timeout_ns) {
          volatile uint32_t* st = a.status[c.rank];
          if (st[0] == 0u) {
            st[1] = target; st[2] = (uint32_t)c.b; st[3] = (uint32_t)q;
            st[0] = (uint32_t)(-B2C_ETIMEOUT);
            __threadfence_system();
          }
          *s_dead = 1;
          break;
        }
      }
    }
  }
  __syncthreads();
  if (*s_dead) c.dead = true;
  if (blockIdx.x == 0 && threadIdx.x == 0)
    reinterpret_cast<unsigned long long*>(c.ctl)[B2C_CTL_PHASE_U64 + ((target - c.epoch) & 3u)] = globaltimer_ns();
}

// iterate vectors i of [0, n) owned by channel c.b, U at a time per thread.
// BODY(idx[U], pred[U]) must be a callable; chunks of a.chunk_vecs are dealt round-robin.
#define B2C_FOR_OWNED(n_, U_, ...)                                                            \
  for (size_t c0_ = (size_t)c.b * a.chunk_vecs; c0_ < (n_); c0_ += (size_t)c.nb * a.chunk_vecs) { \
    const size_t c1_ = (c0_ + a.chunk_vecs < (n_)) ? c0_ + a.chunk_vecs : (n_);                \
    for (size_t i_ = c0_ + threadIdx.x; i_ < c1_; i_ += (size_t)blockDim.x * (U_)) {          \
      size_t idx[U_]; bool pred[U_];                                                         \
      _Pragma("unroll") for (int u = 0; u < (U_); ++u) {                                      \
        idx[u] = i_ + (size_t)u * blockDim.x; pred[u] = idx[u] < c1_;                         \
      }                                                                                      \
      __VA_ARGS__                                                                            \
    }                                                                                        \
  }

template <int W> struct Unroll { static constexpr int U = (W >= 8) ? 2 : (W >= 4) ? 4 : (W >= 2) ? 8 : 8; };
template <> struct Unroll<3> { static constexpr int U = 4; };
template <> struct Unroll<5> { static constexpr int U = 3; };
template <> struct Unroll<6> { static constexpr int U = 2; };
template <> struct Unroll<7> { static constexpr int U = 2; };

// ================================================================= one-shot allreduce
// stage own bucket -> arena; barrier; every rank pulls all W copies of the vectors its
// channel owns, adds them in rank order, scales, rounds once, writes its own buffer.
template <int MODE, int W>
__global__ void __launch_bounds__(512, 1) k_allreduce_oneshot(const __grid_constant__ B2cKArgs a) {
  constexpr int VW = Mode<MODE>::VW;
  constexpr int U = Unroll<W>::U;
  Ctx c = prologue(a);
  void* buf = a.buf[c.rank];
  const size_t count = a.count;
  const size_t nvec = (count + VW - 1) / VW;
  const bool fast = ((reinterpret_cast<uintptr_t>(buf) & 15u) == 0);

  if (W == 1) {
    // single rank: out = in * scale (rounded once); nothing crosses a link, no arena pass
    B2C_FOR_OWNED(nvec, 8, {
      uint4 v[8];
      _Pragma("unroll") for (int u = 0; u < 8; ++u) if (pred[u]) v[u] = load_user<MODE>(buf, idx[u], count, fast);
      _Pragma("unroll") for (int u = 0; u < 8; ++u) if (pred[u]) {
        Acc<MODE> acc; acc.init(v[u], a.scale);
        store_user<MODE>(buf, idx[u], count, fast, acc.finish());
      }
    })
    epilogue(c);
    return;
  }

  char* my_slot = a.region[c.rank] + c.slot_off;
  B2C_FOR_OWNED(nvec, 8, {
    uint4 v[8];
    _Pragma("unroll") for (int u = 0; u < 8; ++u) if (pred[u]) v[u] = load_user<MODE>(buf, idx[u], count, fast);
    _Pragma("unroll") for (int u = 0; u < 8; ++u) if (pred[u]) st_v4(my_slot + idx[u] * 16, v[u]);
  })
  channel_barrier(a, c, c.epoch + 1u);
  if (!c.dead) {
    const char* src[W];
#pragma unroll
    for (int q = 0; q < W; ++q) src[q] = a.region[q] + c.slot_off;
    B2C_FOR_OWNED(nvec, U, {
      uint4 v[U][W];
      _Pragma("unroll") for (int u = 0; u < U; ++u)
        _Pragma("unroll") for (int q = 0; q < W; ++q)
          if (pred[u]) v[u][q] = ld_sys_v4(src[q] + idx[u] * 16);
      _Pragma("unroll") for (int u = 0; u < U; ++u) if (pred[u]) {
        Acc<MODE> acc; acc.init(v[u][0], a.scale);
        _Pragma("unroll") for (int q = 1; q < W; ++q) acc.add(v[u][q], a.scale);
        store_user<MODE>(buf, idx[u], count, fast, acc.finish());
      }
    })
  }
  epilogue(c);
}

// ================================================================= two-shot allreduce
// stage -> barrier A -> rank r reduces shard r from all W arenas (reduce-scatter) and
//   variant 0: pushes the finished vectors into every peer's arena (in place; only rank r
//              ever reads shard r of any arena), writes its own buffer directly
//   variant 1: writes them into its own arena + buffer
// -> barrier B -> copy the W-1 foreign shards into the caller buffer from the local arena
//   (variant 0) or pulled from the owning peer (variant 1).
template <int MODE, int W>
__global__ void __launch_bounds__(512, 1) k_allreduce_twoshot(const __grid_constant__ B2cKArgs a) {
  constexpr int VW = Mode<MODE>::VW;
  constexpr int U = Unroll<W>::U;
  Ctx c = prologue(a);
  void* buf = a.buf[c.rank];
  const size_t count = a.count;
  const size_t nvec = (count + VW - 1) / VW;
  const size_t nvs = (nvec + W - 1) / W;  // vectors per shard (last shards may be short / empty)
  const bool fast = ((reinterpret_cast<uintptr_t>(buf) & 15u) == 0);
  char* my_slot = a.region[c.rank] + c.slot_off;

  // phase 0: stage. Channel b stages slice b of EVERY shard, because channel b of rank r'
  // will read exactly slice b of shard r' from us.
  for (int r = 0; r < W; ++r) {
    const size_t s0 = (size_t)r * nvs;
    if (s0 >= nvec) break;
    const size_t sl = (s0 + nvs <= nvec) ? nvs : nvec - s0;
    B2C_FOR_OWNED(sl, 8, {
      uint4 v[8];
      _Pragma("unroll") for (int u = 0; u < 8; ++u) if (pred[u]) v[u] = load_user<MODE>(buf, s0 + idx[u], count, fast);
      _Pragma("unroll") for (int u = 0; u < 8; ++u) if (pred[u]) st_v4(my_slot + (s0 + idx[u]) * 16, v[u]);
    })
  }
  channel_barrier(a, c, c.epoch + 1u);

  // phase 1: reduce my shard
  if (!c.dead) {
    const size_t s0 = (size_t)c.rank * nvs;
    const size_t sl = (s0 >= nvec) ? 0 : ((s0 + nvs <= nvec) ? nvs : nvec - s0);
    char* slot[W];
#pragma unroll
    for (int q = 0; q < W; ++q) slot[q] = a.region[q] + c.slot_off + s0 * 16;
    const int variant = a.variant;
    B2C_FOR_OWNED(sl, U, {
      uint4 v[U][W];
      _Pragma("unroll") for (int u = 0; u < U; ++u)
        _Pragma("unroll") for (int q = 0; q < W; ++q)
          if (pred[u]) v[u][q] = ld_sys_v4(slot[q] + idx[u] * 16);
      _Pragma("unroll") for (int u = 0; u < U; ++u) if (pred[u]) {
        Acc<MODE> acc; acc.init(v[u][0], a.scale);
        _Pragma("unroll") for (int q = 1; q < W; ++q) acc.add(v[u][q], a.scale);
        const uint4 res = acc.finish();
        store_user<MODE>(buf, s0 + idx[u], count, fast, res);
        if (variant == 0) {
          _Pragma("unroll") for (int k = 1; k < W; ++k) {  // rotate targets so peers are hit evenly
            int q = c.rank + k; if (q >= W) q -= W;
            st_v4(a.region[q] + c.slot_off + (s0 + idx[u]) * 16, res);
          }
        } else {
          st_v4(my_slot + (s0 + idx[u]) * 16, res);
        }
      }
    })
  }
  channel_barrier(a, c, c.epoch + 2u);

  // phase 2: gather the other shards into the caller buffer
  if (!c.dead) {
    for (int k = 1; k < W; ++k) {
      int r = c.rank + k; if (r >= W) r -= W;
      const size_t s0 = (size_t)r * nvs;
      if (s0 >= nvec) continue;
      const size_t sl = (s0 + nvs <= nvec) ? nvs : nvec - s0;
      const char* src = (a.variant == 0 ? my_slot : a.region[r] + c.slot_off) + s0 * 16;
      B2C_FOR_OWNED(sl, 8, {
        uint4 v[8];
        _Pragma("unroll") for (int u = 0; u < 8; ++u) if (pred[u]) v[u] = ld_sys_v4(src + idx[u] * 16);
        _Pragma("unroll") for (int u = 0; u < 8; ++u) if (pred[u]) store_user<MODE>(buf, s0 + idx[u], count, fast, v[u]);
      })
    }
  }
  epilogue(c);
}

// ================================================================= broadcast (bit exact)
// root stages its bytes into its arena; barrier; every other rank pulls them.
__global__ void __launch_bounds__(512, 1) k_broadcast(const __grid_constant__ B2cKArgs a) {
  Ctx c = prologue(a);
  void* buf = a.buf[c.rank];
  const size_t count = a.count;
  const size_t nvec = (count + 15) / 16;
  const bool fast = ((reinterpret_cast<uintptr_t>(buf) & 15u) == 0);
  char* root_slot = a.region[a.root] + c.slot_off;
  if (c.rank == a.root) {
    B2C_FOR_OWNED(nvec, 8, {
      uint4 v[8];
      _Pragma("unroll") for (int u = 0; u < 8; ++u) if (pred[u]) v[u] = load_user<MODE_U8>(buf, idx[u], count, fast);
      _Pragma("unroll") for (int u = 0; u < 8; ++u) if (pred[u]) st_v4(root_slot + idx[u] * 16, v[u]);
    })
  }
  channel_barrier(a, c, c.epoch + 1u);
  if (!c.dead && c.rank != a.root) {
    B2C_FOR_OWNED(nvec, 8, {
      uint4 v[8];
      _Pragma("unroll") for (int u = 0; u < 8; ++u) if (pred[u]) v[u] = ld_sys_v4(root_slot + idx[u] * 16);
      _Pragma("unroll") for (int u = 0; u < 8; ++u) if (pred[u]) store_user<MODE_U8>(buf, idx[u], count, fast, v[u]);
    })
  }
  epilogue(c);
}

// ================================================================= barrier
__global__ void __launch_bounds__(32, 1) k_barrier(const __grid_constant__ B2cKArgs a) {
  Ctx c = prologue(a);
  channel_barrier(a, c, c.epoch + 1u);
  epilogue(c);
}

// ================================================================= zero-copy two-shot allreduce
// The caller's buffers themselves are peer-mapped (registered, or a local group), so there is
// no staging pass and no copy-out:  barrier A (every rank's gradients are final: stream order
// before its launch) -> rank r pulls shard r from all W buffers, reduces, and writes the result
// back IN PLACE into all W buffers (only rank r ever touches shard r between the barriers)
// -> barrier B.  HBM traffic per rank: S/W * W read-serves + S writes; NVLink 2(W-1)/W * S.
// Needs 16-byte aligned buffers; the ragged tail (< one vector) is reduced element-wise by
// rank 0 / channel 0.
template <int MODE> struct Elem;
template <> struct Elem<MODE_F32> {
  static __device__ __forceinline__ float ld(const void* base, size_t e) {
    uint32_t v;
    asm volatile("ld.relaxed.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(static_cast<const float*>(base) + e));
    return __uint_as_float(v);
  }
  static __device__ __forceinline__ void st(void* base, size_t e, float f) { static_cast<float*>(base)[e] = f; }
};
template <> struct Elem<MODE_BF16> {
  static __device__ __forceinline__ float ld(const void* base, size_t e) {
    unsigned short v;
    asm volatile("ld.relaxed.sys.global.u16 %0, [%1];" : "=h"(v) : "l"(static_cast<const unsigned short*>(base) + e));
    return __uint_as_float((uint32_t)v << 16);
  }
  static __device__ __forceinline__ void st(void* base, size_t e, float f) {
    __nv_bfloat16 h = __float2bfloat16_rn(f);
    static_cast<unsigned short*>(base)[e] = *reinterpret_cast<unsigned short*>(&h);
  }
};

template <int MODE, int W, int UU = 0>
__global__ void __launch_bounds__(512, 1) k_allreduce_twoshot_zc(const __grid_constant__ B2cKArgs a) {
  constexpr int VW = Mode<MODE>::VW;
  constexpr int U = UU > 0 ? UU : Unroll<W>::U;
  Ctx c = prologue(a);
  const size_t count = a.count;
  const size_t nfull = count / VW;          // whole vectors
  const size_t nvs = (nfull + W - 1) / W;   // vectors per shard
  // in a local group every virtual rank shares one address space: buf[q] is rank q's buffer;
  // in IPC mode buf[q] is rank q's buffer as mapped in THIS process.
  channel_barrier(a, c, c.epoch + 1u);
  if (!c.dead) {
    const size_t s0 = (size_t)c.rank * nvs;
    const size_t sl = (s0 >= nfull) ? 0 : ((s0 + nvs <= nfull) ? nvs : nfull - s0);
    char* p[W];
#pragma unroll
    for (int q = 0; q < W; ++q) p[q] = static_cast<char*>(a.buf[q]) + s0 * 16;
    B2C_FOR_OWNED(sl, U, {
      uint4 v[U][W];
      _Pragma("unroll") for (int u = 0; u < U; ++u)
        _Pragma("unroll") for (int q = 0; q < W; ++q)
          if (pred[u]) v[u][q] = ld_sys_v4(p[q] + idx[u] * 16);
      _Pragma("unroll") for (int u = 0; u < U; ++u) if (pred[u]) {
        Acc<MODE> acc; acc.init(v[u][0], a.scale);
        _Pragma("unroll") for (int q = 1; q < W; ++q) acc.add(v[u][q], a.scale);
        const uint4 res = acc.finish();
        _Pragma("unroll") for (int k = 0; k < W; ++k) {  // own copy first, then peers rotated
          int q = c.rank + k; if (q >= W) q -= W;
          st_v4(static_cast<char*>(a.buf[q]) + (s0 + idx[u]) * 16, res);
        }
      }
    })
    if (c.rank == 0 && c.b == 0) {
      const size_t e = nfull * VW + threadIdx.x;
      if (e < count) {
        float acc = __fmul_rn(Elem<MODE>::ld(a.buf[0], e), a.scale);
#pragma unroll
        for (int q = 1; q < W; ++q) acc = __fadd_rn(acc, __fmul_rn(Elem<MODE>::ld(a.buf[q], e), a.scale));
#pragma unroll
        for (int q = 0; q < W; ++q) Elem<MODE>::st(a.buf[q], e, acc);
      }
    }
  }
  channel_barrier(a, c, c.epoch + 2u);
  epilogue(c);
}


// ================================================================= NVLS allreduce (in-switch reduction)
// multimem.ld_reduce on a multicast address returns the SUM over every GPU's copy (the NVSwitch
// adds, fp32 accumulate); multimem.st writes one value into every GPU's copy.  Each rank reduces
// its own shard and broadcasts it, so a GPU moves ~S bytes per direction instead of 2(W-1)/W*S.
//   staged  : stage bucket -> local arena slot; barrier A; ld_reduce / scale / multimem.st (+ own
//             bucket directly); barrier B; copy the other shards from the local arena into the bucket
//   in place: the tensor lives in the arena: barrier A; ld_reduce / scale / multimem.st; barrier B
// World size is a runtime value here (the switch does the W-way add).
__device__ __forceinline__ uint4 multimem_ld_reduce_f32(const void* mc) {
  uint4 v;
  asm volatile("multimem.ld_reduce.relaxed.sys.global.add.v4.f32 {%0,%1,%2,%3}, [%4];"
               : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(mc) : "memory");
  return v;
}
__device__ __forceinline__ uint4 multimem_ld_reduce_bf16(const void* mc) {
  uint4 v;
  asm volatile("multimem.ld_reduce.relaxed.sys.global.add.acc::f32.v4.bf16x2 {%0,%1,%2,%3}, [%4];"
               : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(mc) : "memory");
  return v;
}
__device__ __forceinline__ void multimem_st_v4(void* mc, uint4 v) {
  asm volatile("multimem.st.relaxed.sys.global.v4.f32 [%0], {%1,%2,%3,%4};" ::"l"(mc), "r"(v.x), "r"(v.y),
               "r"(v.z), "r"(v.w) : "memory");
}

template <int MODE>
__device__ __forceinline__ uint4 nvls_scale(uint4 v, float scale) {
  if (MODE == MODE_F32) {
    return make_uint4(__float_as_uint(__fmul_rn(__uint_as_float(v.x), scale)), __float_as_uint(__fmul_rn(__uint_as_float(v.y), scale)),
                      __float_as_uint(__fmul_rn(__uint_as_float(v.z), scale)), __float_as_uint(__fmul_rn(__uint_as_float(v.w), scale)));
  } else {
    const uint32_t w[4] = {v.x, v.y, v.z, v.w};
    uint32_t o[4];
#pragma unroll
    for (int k = 0; k < 4; ++k) o[k] = pack_bf16x2(__fmul_rn(bf16_lo(w[k]), scale), __fmul_rn(bf16_hi(w[k]), scale));
    return make_uint4(o[0], o[1], o[2], o[3]);
  }
}

template <int MODE, int NU = 8>
__global__ void __launch_bounds__(512, 1) k_allreduce_nvls(const __grid_constant__ B2cKArgs a) {
  constexpr int VW = Mode<MODE>::VW;
  Ctx c = prologue(a);
  const int W = a.world;
  void* buf = a.buf[c.rank];
  const size_t count = a.count;
  const size_t nvec = (count + VW - 1) / VW;
  const size_t nvs = (nvec + W - 1) / W;
  const bool inplace = a.nvls_inplace != 0;
  const bool fast = ((reinterpret_cast<uintptr_t>(buf) & 15u) == 0);
  const size_t base = inplace ? a.nvls_off : a.nvls_off + (size_t)(c.calls & 1u) * a.nvls_slot_bytes;
  char* uc = a.nvls_uc + base;
  char* mc = a.nvls_mc + base;

  if (!inplace) {
    for (int r = 0; r < W; ++r) {
      const size_t s0 = (size_t)r * nvs;
      if (s0 >= nvec) break;
      const size_t sl = (s0 + nvs <= nvec) ? nvs : nvec - s0;
      B2C_FOR_OWNED(sl, 8, {
        uint4 v[8];
        _Pragma("unroll") for (int u = 0; u < 8; ++u) if (pred[u]) v[u] = load_user<MODE>(buf, s0 + idx[u], count, fast);
        _Pragma("unroll") for (int u = 0; u < 8; ++u) if (pred[u]) st_v4(uc + (s0 + idx[u]) * 16, v[u]);
      })
    }
  }
  channel_barrier(a, c, c.epoch + 1u);
  if (!c.dead) {
    const size_t s0 = (size_t)c.rank * nvs;
    const size_t sl = (s0 >= nvec) ? 0 : ((s0 + nvs <= nvec) ? nvs : nvec - s0);
    B2C_FOR_OWNED(sl, NU, {
      uint4 v[NU];
      _Pragma("unroll") for (int u = 0; u < NU; ++u) if (pred[u])
        v[u] = (MODE == MODE_F32) ? multimem_ld_reduce_f32(mc + (s0 + idx[u]) * 16) : multimem_ld_reduce_bf16(mc + (s0 + idx[u]) * 16);
      _Pragma("unroll") for (int u = 0; u < NU; ++u) if (pred[u]) {
        const uint4 res = nvls_scale<MODE>(v[u], a.scale);
        multimem_st_v4(mc + (s0 + idx[u]) * 16, res);
        if (!inplace) store_user<MODE>(buf, s0 + idx[u], count, fast, res);
      }
    })
  }
  channel_barrier(a, c, c.epoch + 2u);
  if (!c.dead && !inplace) {
    for (int k = 1; k < W; ++k) {
      int r = c.rank + k; if (r >= W) r -= W;
      const size_t s0 = (size_t)r * nvs;
      if (s0 >= nvec) continue;
      const size_t sl = (s0 + nvs <= nvec) ? nvs : nvec - s0;
      B2C_FOR_OWNED(sl, 8, {
        uint4 v[8];
        _Pragma("unroll") for (int u = 0; u < 8; ++u) if (pred[u]) v[u] = ld_sys_v4(uc + (s0 + idx[u]) * 16);
        _Pragma("unroll") for (int u = 0; u < 8; ++u) if (pred[u]) store_user<MODE>(buf, s0 + idx[u], count, fast, v[u]);
      })
    }
  }
  epilogue(c);
}


// ================================================================= broadcast through the switch
// With the multicast arena up the root sends its bytes ONCE: multimem.st replicates every 16-byte vector
// into all W copies of an arena range (a staging slot picked by call parity, or the tensor's own place
// when it lives in the arena at the same offset on every rank), one flag barrier, and every other rank
// copies slot -> buffer locally at HBM rate.  Root egress S instead of (W-1)*S.
__global__ void __launch_bounds__(512, 1) k_broadcast_mc(const __grid_constant__ B2cKArgs a) {
  Ctx c = prologue(a);
  void* buf = a.buf[c.rank];
  const size_t count = a.count;
  const size_t nvec = (count + 15) / 16;
  const bool inplace = a.nvls_inplace != 0;
  const bool fast = ((reinterpret_cast<uintptr_t>(buf) & 15u) == 0);
  const size_t base = inplace ? a.nvls_off : a.nvls_off + (size_t)(c.calls & 1u) * a.nvls_slot_bytes;
  char* mc = a.nvls_mc + base;
  const char* uc = a.nvls_uc + base;
  // In place the root writes straight into the peers' tensors, so it has to know that every peer has reached
  // this call (whatever wrote that memory earlier on the peer's stream is done): entry barrier.  A staging slot
  // is only ever written by broadcast kernels and is double buffered, so the staged form needs none.
  if (inplace) channel_barrier(a, c, c.epoch + 1u);
  if (c.rank == a.root && !c.dead) {
    B2C_FOR_OWNED(nvec, 8, {
      uint4 v[8];
      _Pragma("unroll") for (int u = 0; u < 8; ++u) if (pred[u]) v[u] = load_user<MODE_U8>(buf, idx[u], count, fast);
      _Pragma("unroll") for (int u = 0; u < 8; ++u) if (pred[u]) multimem_st_v4(mc + idx[u] * 16, v[u]);
    })
  }
  channel_barrier(a, c, c.epoch + (inplace ? 2u : 1u));
  if (!c.dead && c.rank != a.root && !inplace) {
    B2C_FOR_OWNED(nvec, 8, {
      uint4 v[8];
      _Pragma("unroll") for (int u = 0; u < 8; ++u) if (pred[u]) v[u] = ld_sys_v4(uc + idx[u] * 16);
      _Pragma("unroll") for (int u = 0; u < 8; ++u) if (pred[u]) store_user<MODE_U8>(buf, idx[u], count, fast, v[u]);
    })
  }
  epilogue(c);
}


// ================================================================= NVLS in place, lean (the per-step hot kernel)
// The tensor lives in the multicast arena at the same offset on every rank (DDP / hvd gradient buckets do).
// Differences from k_allreduce_nvls, all about the FIXED cost of a call:
//   * the cross-GPU barrier is a counter in the multicast arena's head: one thread issues ONE
//     multimem.red.release (the switch adds 1 into all W copies) and spins on its local copy until it
//     reaches W * epoch — instead of W unicast release-stores and W spinning lanes;
//   * epochs are per channel, so a CTA never waits for the other CTAs of its own grid: no exit ticket,
//     no grid-wide counters (channel b only ever meets channel b of the peers);
//   * small argument block (launch latency grows with the parameter bytes);
//   * griddepcontrol: launched with programmatic stream serialization, the next kernel of a bucket
//     train becomes resident while this one drains (its prologue waits for our completion).
struct B2cNvlsArgs {
  char* uc;                 // local unicast view of the multicast arena
  char* mc;                 // multicast view
  uint32_t* ctl;            // this rank's control page
  uint32_t* status;         // sticky error words (host mapped)
  size_t off;               // byte offset of the tensor inside the arena
  size_t nvec;              // whole 16-byte vectors
  unsigned long long timeout_ns;
  float scale;
  int rank, world;
  uint32_t chunk_vecs;
  int acq_poll;             // flag spin: 1 = ld.acquire.sys polls, 0 = relaxed polls + one fence at the end
};

__device__ __forceinline__ void multimem_red_add_release_u32(void* mc, uint32_t v) {
  asm volatile("multimem.red.release.sys.global.add.u32 [%0], %1;" ::"l"(mc), "r"(v) : "memory");
}
__device__ __forceinline__ uint32_t ld_relaxed_sys_u32(const uint32_t* p) {
  uint32_t v;
  asm volatile("ld.relaxed.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ void fence_acq_rel_sys() { asm volatile("fence.acq_rel.sys;" ::: "memory"); }
__device__ __forceinline__ void pdl_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }
__device__ __forceinline__ void pdl_launch_dependents() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }

// one thread: arrive (+1 in every rank's copy) and wait until all W ranks have arrived `target / W` times.
// returns false on timeout (sticky status word set).
template <bool ACQ_POLL>
__device__ __forceinline__ bool mc_arrive_wait(char* mc_flag, const uint32_t* uc_flag, uint32_t target,
                                               unsigned long long timeout_ns, uint32_t* status, int channel) {
  multimem_red_add_release_u32(mc_flag, 1u);
  unsigned long long t0 = 0;
  unsigned spins = 0;
  while ((int)((ACQ_POLL ? ld_acquire_sys(uc_flag) : ld_relaxed_sys_u32(uc_flag)) - target) < 0) {
    if ((++spins & 0x3ffu) == 0u) {
      unsigned long long now = globaltimer_ns();
      if (t0 == 0) t0 = now;
      else if (now - t0 > timeout_ns) {
        volatile uint32_t* st = status;
        if (st[0] == 0u) {
          st[1] = target; st[2] = (uint32_t)channel; st[3] = 0xffffffffu;
          st[0] = (uint32_t)(-B2C_ETIMEOUT);
          __threadfence_system();
        }
        return false;
      }
    }
  }
  if (!ACQ_POLL) fence_acq_rel_sys();
  return true;
}

template <int MODE, int NU, int THREADS, int MINB>
__global__ void __launch_bounds__(THREADS, MINB) k_nvls_inplace(const __grid_constant__ B2cNvlsArgs a) {
  __shared__ uint32_t s_epoch;
  __shared__ int s_dead;
  const int b = blockIdx.x;
  char* mc_flag = a.mc + (size_t)b * B2C_MC_FLAG_STRIDE;
  const uint32_t* uc_flag = reinterpret_cast<const uint32_t*>(a.uc + (size_t)b * B2C_MC_FLAG_STRIDE);
  unsigned long long* ph = reinterpret_cast<unsigned long long*>(a.ctl) + B2C_CTL_PHASE_U64;
  const uint32_t W = (uint32_t)a.world;
  // everything before this line overlaps the tail of the previous kernel on the stream (PDL); the epoch
  // word below is written by that kernel's last instructions
  pdl_wait();
  if (threadIdx.x == 0) {
    if (b == 0) ph[0] = globaltimer_ns();
    const uint32_t e = ld_volatile_u32(a.ctl + B2C_CTL_MC_EPOCH + b);
    s_epoch = e;
    // barrier A: every rank's gradients are final (stream order before its launch)
    const bool okA = a.acq_poll ? mc_arrive_wait<true>(mc_flag, uc_flag, (e + 1u) * W, a.timeout_ns, a.status, b)
                                : mc_arrive_wait<false>(mc_flag, uc_flag, (e + 1u) * W, a.timeout_ns, a.status, b);
    s_dead = okA ? 0 : 1;
    if (b == 0) ph[1] = globaltimer_ns();
  }
  __syncthreads();
  pdl_launch_dependents();
  const bool dead = s_dead != 0;
  if (!dead) {
    const size_t nvs = (a.nvec + W - 1) / W;
    const size_t s0 = (size_t)a.rank * nvs;
    const size_t sl = (s0 >= a.nvec) ? 0 : ((s0 + nvs <= a.nvec) ? nvs : a.nvec - s0);
    char* mc = a.mc + a.off + s0 * 16;
    const size_t step = (size_t)gridDim.x * a.chunk_vecs;
    for (size_t c0 = (size_t)b * a.chunk_vecs; c0 < sl; c0 += step) {
      const size_t c1 = (c0 + a.chunk_vecs < sl) ? c0 + a.chunk_vecs : sl;
      for (size_t i = c0 + threadIdx.x; i < c1; i += (size_t)THREADS * NU) {
        uint4 v[NU];
#pragma unroll
        for (int u = 0; u < NU; ++u) {
          const size_t j = i + (size_t)u * THREADS;
          if (j < c1) v[u] = (MODE == MODE_F32) ? multimem_ld_reduce_f32(mc + j * 16) : multimem_ld_reduce_bf16(mc + j * 16);
        }
#pragma unroll
        for (int u = 0; u < NU; ++u) {
          const size_t j = i + (size_t)u * THREADS;
          if (j < c1) multimem_st_v4(mc + j * 16, nvls_scale<MODE>(v[u], a.scale));
        }
      }
    }
  }
  __syncthreads();  // every multimem.st of this CTA is issued; the release below orders them before the arrival
  if (threadIdx.x == 0) {
    const uint32_t e = s_epoch;
    // barrier B: every rank's results have landed in every copy
    if (!dead) {
      if (a.acq_poll) mc_arrive_wait<true>(mc_flag, uc_flag, (e + 2u) * W, a.timeout_ns, a.status, b);
      else mc_arrive_wait<false>(mc_flag, uc_flag, (e + 2u) * W, a.timeout_ns, a.status, b);
    }
    if (b == 0) {
      const unsigned long long t = globaltimer_ns();
      ph[2] = t; ph[3] = t;
      phase_commit(a.ctl);
    }
    a.ctl[B2C_CTL_MC_EPOCH + b] = e + 2u;
  }
}


// ================================================================= fused optimizer step
// The step on the far side of the path: instead of "allreduce the gradients, then let every rank run the same
// optimizer over all of them", the rank that reduces a shard also APPLIES the optimizer to that shard — fp32
// state held for the owned shard only (1/W of the momentum / Adam moments per rank) — and what travels back
// is the updated WEIGHTS, written into every rank's parameter bucket.  Same NVLink bytes as the allreduce,
// minus a full read/write pass over gradients, parameters and state on every rank and the optimizer's launches.
// Arithmetic follows torch.optim.SGD / Adam / AdamW (single-tensor form) operation by operation in fp32.
enum { B2C_OPT_NONE = 0, B2C_OPT_SGD = 1, B2C_OPT_ADAM = 2, B2C_OPT_ADAMW = 3 };

// Every derived scalar is computed by the host in double and rounded to float ONCE, the way torch's Python
// optimizers hand their scalars to the kernels (1 - beta2 formed in float would be off by 1e-5 relative).
struct B2cOptK {
  int kind, nesterov, first;     // first: momentum buffers are created from the first gradient (torch: buf = clone(g))
  float lr, momentum, one_m_dampening, weight_decay;
  float one_m_beta1, beta2, one_m_beta2, eps, step_size /* lr / bias_correction1 */, bc2_sqrt, adamw_keep /* 1 - lr * wd */;
};

template <int KIND>
__device__ __forceinline__ float opt_apply(float g, float w, float& s1, float& s2, const B2cOptK& o) {
  if (KIND == B2C_OPT_SGD) {
    if (o.weight_decay != 0.f) g = __fmaf_rn(o.weight_decay, w, g);
    if (o.momentum != 0.f) {
      s1 = o.first ? g : __fmaf_rn(o.one_m_dampening, g, __fmul_rn(o.momentum, s1));   // buf.mul_(mom).add_(g, alpha=1-damp)
      g = o.nesterov ? __fmaf_rn(o.momentum, s1, g) : s1;
    }
    return __fmaf_rn(-o.lr, g, w);
  } else {
    if (KIND == B2C_OPT_ADAMW) w = __fmul_rn(w, o.adamw_keep);                  // param.mul_(1 - lr * wd)
    else if (o.weight_decay != 0.f) g = __fmaf_rn(o.weight_decay, w, g);
    s1 = __fmaf_rn(o.one_m_beta1, __fsub_rn(g, s1), s1);                        // exp_avg.lerp_(g, 1 - beta1)
    s2 = __fmaf_rn(__fmul_rn(o.one_m_beta2, g), g, __fmul_rn(o.beta2, s2));     // exp_avg_sq.mul_(b2).addcmul_(g, g, 1 - b2)
    const float denom = __fadd_rn(__fdiv_rn(__fsqrt_rn(s2), o.bc2_sqrt), o.eps);
    return __fmaf_rn(-o.step_size, __fdiv_rn(s1, denom), w);                   // param.addcdiv_(exp_avg, denom, -step_size)
  }
}

// one 16-byte vector of reduced fp32 gradients -> updated weights (fp32 parameters and state)
template <int KIND>
__device__ __forceinline__ uint4 opt_vec(uint4 g, uint4 w, float4* s1p, float4* s2p, const B2cOptK& o) {
  float4 s1 = (KIND == B2C_OPT_SGD && (o.momentum == 0.f || o.first)) ? make_float4(0.f, 0.f, 0.f, 0.f) : *s1p;
  float4 s2 = (KIND == B2C_OPT_SGD) ? make_float4(0.f, 0.f, 0.f, 0.f) : *s2p;
  uint4 r;
  r.x = __float_as_uint(opt_apply<KIND>(__uint_as_float(g.x), __uint_as_float(w.x), s1.x, s2.x, o));
  r.y = __float_as_uint(opt_apply<KIND>(__uint_as_float(g.y), __uint_as_float(w.y), s1.y, s2.y, o));
  r.z = __float_as_uint(opt_apply<KIND>(__uint_as_float(g.z), __uint_as_float(w.z), s1.z, s2.z, o));
  r.w = __float_as_uint(opt_apply<KIND>(__uint_as_float(g.w), __uint_as_float(w.w), s1.w, s2.w, o));
  if (KIND != B2C_OPT_SGD || o.momentum != 0.f) *s1p = s1;
  if (KIND != B2C_OPT_SGD) *s2p = s2;
  return r;
}

// ---- W = 1: the whole step is local.  Streaming, HBM bound: SGD+momentum reads g, w, m and writes w, m
// (20 bytes per parameter), Adam reads g, w, m, v and writes w, m, v (28).  Grid fills the chip.
template <int KIND>
__global__ void __launch_bounds__(256) k_fused_step_local(const float* __restrict__ grad, float* __restrict__ param,
                                                          float* __restrict__ st1, float* __restrict__ st2, size_t count,
                                                          float scale, const __grid_constant__ B2cOptK o) {
  const size_t nfull = count / 4;
  const uint4* g4 = reinterpret_cast<const uint4*>(grad);
  uint4* w4 = reinterpret_cast<uint4*>(param);
  float4* s1 = reinterpret_cast<float4*>(st1);
  float4* s2 = reinterpret_cast<float4*>(st2);
  const size_t stride = (size_t)gridDim.x * blockDim.x;
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < nfull; i += stride) {
    uint4 g = g4[i];
    if (scale != 1.f) g = nvls_scale<MODE_F32>(g, scale);
    w4[i] = opt_vec<KIND>(g, w4[i], s1 + i, s2 + i, o);
  }
  if (blockIdx.x == 0 && threadIdx.x < (count & 3)) {   // ragged tail, element-wise
    const size_t e = nfull * 4 + threadIdx.x;
    float a = 0.f, b = 0.f;
    if (!(KIND == B2C_OPT_SGD && (o.momentum == 0.f || o.first))) a = st1[e];
    if (KIND != B2C_OPT_SGD) b = st2[e];
    param[e] = opt_apply<KIND>(__fmul_rn(grad[e], scale), param[e], a, b, o);
    if (KIND != B2C_OPT_SGD || o.momentum != 0.f) st1[e] = a;
    if (KIND != B2C_OPT_SGD) st2[e] = b;
  }
}

// ---- peer-to-peer (registered buckets, or a local group): zero-copy two-shot whose second shot carries weights.
// a.buf[q] = rank q's GRADIENT bucket, a.pbuf[q] = rank q's PARAMETER bucket (both mapped here); state is local.
struct B2cFusedArgs {
  B2cKArgs k;
  void* pbuf[B2C_MAX_WORLD];
  float* st1[B2C_MAX_WORLD];     // per (virtual) rank: state of its shard (real mode: only [rank])
  float* st2[B2C_MAX_WORLD];
  B2cOptK opt;
};

template <int KIND, int W>
__global__ void __launch_bounds__(512, 1) k_fused_step_zc(const __grid_constant__ B2cFusedArgs fa) {
  const B2cKArgs& a = fa.k;
  constexpr int U = (W >= 8) ? 1 : (W >= 4) ? 2 : 4;
  Ctx c = prologue(a);
  const size_t nvec = a.count / 4;            // whole vectors only: buckets are padded to 16 bytes by the reducer
  const size_t nvs = (nvec + W - 1) / W;
  channel_barrier(a, c, c.epoch + 1u);
  if (!c.dead) {
    const size_t s0 = (size_t)c.rank * nvs;
    const size_t sl = (s0 >= nvec) ? 0 : ((s0 + nvs <= nvec) ? nvs : nvec - s0);
    char* p[W];
#pragma unroll
    for (int q = 0; q < W; ++q) p[q] = static_cast<char*>(a.buf[q]) + s0 * 16;
    const char* wloc = static_cast<const char*>(fa.pbuf[c.rank]) + s0 * 16;
    float4* s1 = reinterpret_cast<float4*>(fa.st1[c.rank]);
    float4* s2 = reinterpret_cast<float4*>(fa.st2[c.rank]);
    B2C_FOR_OWNED(sl, U, {
      uint4 v[U][W];
      uint4 w[U];
      _Pragma("unroll") for (int u = 0; u < U; ++u)
        _Pragma("unroll") for (int q = 0; q < W; ++q)
          if (pred[u]) v[u][q] = ld_sys_v4(p[q] + idx[u] * 16);
      _Pragma("unroll") for (int u = 0; u < U; ++u) if (pred[u]) w[u] = ld_sys_v4(wloc + idx[u] * 16);
      _Pragma("unroll") for (int u = 0; u < U; ++u) if (pred[u]) {
        Acc<MODE_F32> acc; acc.init(v[u][0], a.scale);
        _Pragma("unroll") for (int q = 1; q < W; ++q) acc.add(v[u][q], a.scale);
        const uint4 res = opt_vec<KIND>(acc.finish(), w[u], s1 + idx[u], s2 + idx[u], fa.opt);
        _Pragma("unroll") for (int k = 0; k < W; ++k) {  // own copy first, then peers rotated
          int q = c.rank + k; if (q >= W) q -= W;
          st_v4(static_cast<char*>(fa.pbuf[q]) + (s0 + idx[u]) * 16, res);
        }
      }
    })
  }
  channel_barrier(a, c, c.epoch + 2u);
  epilogue(c);
}

// ---- through the switch: multimem.ld_reduce the gradient shard, update, multimem.st the weights; optionally the
// root's module buffers (C4, the per-forward broadcast of the reference path) ride in the same launch.
struct B2cNvlsFusedArgs {
  B2cNvlsArgs n;          // n.off = gradient bucket offset in the arena
  size_t param_off;       // parameter bucket offset in the arena (same layout as the gradient bucket)
  float* st1;
  float* st2;
  size_t bcast_off;       // C4 fold: [bcast_off, +bcast_vecs*16) of the arena is multicast from bcast_root
  size_t bcast_vecs;
  int bcast_root;
  B2cOptK opt;
};

template <int KIND, int NU>
__global__ void __launch_bounds__(512, 1) k_fused_step_nvls(const __grid_constant__ B2cNvlsFusedArgs fa) {
  const B2cNvlsArgs& a = fa.n;
  __shared__ uint32_t s_epoch;
  __shared__ int s_dead;
  const int b = blockIdx.x;
  char* mc_flag = a.mc + (size_t)b * B2C_MC_FLAG_STRIDE;
  const uint32_t* uc_flag = reinterpret_cast<const uint32_t*>(a.uc + (size_t)b * B2C_MC_FLAG_STRIDE);
  unsigned long long* ph = reinterpret_cast<unsigned long long*>(a.ctl) + B2C_CTL_PHASE_U64;
  const uint32_t W = (uint32_t)a.world;
  pdl_wait();
  if (threadIdx.x == 0) {
    if (b == 0) ph[0] = globaltimer_ns();
    const uint32_t e = ld_volatile_u32(a.ctl + B2C_CTL_MC_EPOCH + b);
    s_epoch = e;
    s_dead = mc_arrive_wait<true>(mc_flag, uc_flag, (e + 1u) * W, a.timeout_ns, a.status, b) ? 0 : 1;
    if (b == 0) ph[1] = globaltimer_ns();
  }
  __syncthreads();
  pdl_launch_dependents();
  const bool dead = s_dead != 0;
  if (!dead) {
    const size_t nvs = (a.nvec + W - 1) / W;
    const size_t s0 = (size_t)a.rank * nvs;
    const size_t sl = (s0 >= a.nvec) ? 0 : ((s0 + nvs <= a.nvec) ? nvs : a.nvec - s0);
    const char* gmc = a.mc + a.off + s0 * 16;
    char* wmc = a.mc + fa.param_off + s0 * 16;
    const char* wuc = a.uc + fa.param_off + s0 * 16;
    float4* s1 = reinterpret_cast<float4*>(fa.st1);
    float4* s2 = reinterpret_cast<float4*>(fa.st2);
    const size_t step = (size_t)gridDim.x * a.chunk_vecs;
    for (size_t c0 = (size_t)b * a.chunk_vecs; c0 < sl; c0 += step) {
      const size_t c1 = (c0 + a.chunk_vecs < sl) ? c0 + a.chunk_vecs : sl;
      for (size_t i = c0 + threadIdx.x; i < c1; i += (size_t)blockDim.x * NU) {
        uint4 g[NU], w[NU];
#pragma unroll
        for (int u = 0; u < NU; ++u) {
          const size_t j = i + (size_t)u * blockDim.x;
          if (j < c1) { g[u] = multimem_ld_reduce_f32(gmc + j * 16); w[u] = ld_sys_v4(wuc + j * 16); }
        }
#pragma unroll
        for (int u = 0; u < NU; ++u) {
          const size_t j = i + (size_t)u * blockDim.x;
          if (j < c1) multimem_st_v4(wmc + j * 16, opt_vec<KIND>(nvls_scale<MODE_F32>(g[u], a.scale), w[u], s1 + j, s2 + j, fa.opt));
        }
      }
    }
    if (fa.bcast_vecs && a.rank == fa.bcast_root) {   // C4: the root's buffer bytes, in place, to every rank
      const char* src = a.uc + fa.bcast_off;
      char* dst = a.mc + fa.bcast_off;
      for (size_t i = (size_t)b * blockDim.x + threadIdx.x; i < fa.bcast_vecs; i += (size_t)gridDim.x * blockDim.x)
        multimem_st_v4(dst + i * 16, ld_sys_v4(src + i * 16));
    }
  }
  __syncthreads();
  if (threadIdx.x == 0) {
    const uint32_t e = s_epoch;
    if (!dead) mc_arrive_wait<true>(mc_flag, uc_flag, (e + 2u) * W, a.timeout_ns, a.status, b);
    if (b == 0) {
      const unsigned long long t = globaltimer_ns();
      ph[2] = t; ph[3] = t;
      phase_commit(a.ctl);
    }
    a.ctl[B2C_CTL_MC_EPOCH + b] = e + 2u;
  }
}


// ================================================================= W = 1: in-place scale pass
// A single rank has nothing to exchange: out = in * scale (one rounding).  Plain HBM-bound
// streaming kernel — no flags, no counters, small register footprint so that many CTAs share an SM
// next to the backward kernels.  16-byte aligned buffers only (the generic one-shot kernel takes
// the rest); the ragged tail (< one vector) is handled by the last thread block's first warp.
template <int MODE>
__global__ void __launch_bounds__(256) k_scale_inplace(void* __restrict__ buf, size_t count, float scale) {
  constexpr int VW = Mode<MODE>::VW;
  const size_t nfull = count / VW;
  uint4* p = static_cast<uint4*>(buf);
  const size_t stride = (size_t)gridDim.x * blockDim.x;
  size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  for (; i + 3 * stride < nfull; i += 4 * stride) {
    uint4 v0 = p[i], v1 = p[i + stride], v2 = p[i + 2 * stride], v3 = p[i + 3 * stride];
    Acc<MODE> a0, a1, a2, a3;
    a0.init(v0, scale); a1.init(v1, scale); a2.init(v2, scale); a3.init(v3, scale);
    p[i] = a0.finish(); p[i + stride] = a1.finish(); p[i + 2 * stride] = a2.finish(); p[i + 3 * stride] = a3.finish();
  }
  for (; i < nfull; i += stride) {
    Acc<MODE> a; a.init(p[i], scale);
    p[i] = a.finish();
  }
  if (blockIdx.x == gridDim.x - 1) {
    const size_t e = nfull * VW + threadIdx.x;
    if (e < count) Elem<MODE>::st(buf, e, __fmul_rn(Elem<MODE>::ld(buf, e), scale));
  }
}
