// device_utils.cuh -- small device-side building blocks: scoped loads/stores, deterministic block
// reductions, the cross-rank reduction over peer memory, %globaltimer.
//
// The cross-rank reduction replaces the reference's dot_gpu + MPI_Allreduce pair (S4:771-786, S4:871,
// S4:892, S4:925): per-CTA partials -> last-arriving CTA sums them in a fixed tree -> one 32-byte
// "LL" record (payload and sequence number in the same 8-byte word, so no separate flag and no fence
// on the critical path) is stored into every rank's slot area over NVLink -> every CTA of every rank
// polls its own local slot area and adds the per-rank values in rank order.  All ranks therefore
// obtain bit-identical sums and take identical branches.
#pragma once
#include <cuda_runtime.h>

#include "../../include/poisson_b200.h"
#include "ctx.h"

namespace pb {

__device__ __forceinline__ unsigned long long globaltimer_ns() {
  unsigned long long t;
  asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t));
  return t;
}

__device__ __forceinline__ unsigned long long ld_relaxed_sys_u64(const unsigned long long* p) {
  unsigned long long v;
  asm volatile("ld.relaxed.sys.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ void st_release_sys_u64(unsigned long long* p, unsigned long long v) {
  asm volatile("st.release.sys.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}
__device__ __forceinline__ void st_relaxed_sys_u64(unsigned long long* p, unsigned long long v) {
  asm volatile("st.relaxed.sys.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}
__device__ __forceinline__ unsigned int ld_acquire_gpu_u32(const unsigned int* p) {
  unsigned int v;
  asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ int ld_relaxed_gpu_s32(const int* p) {
  int v;
  asm volatile("ld.relaxed.gpu.global.s32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ unsigned long long ld_acquire_sys_u64(const unsigned long long* p) {
  unsigned long long v;
  asm volatile("ld.acquire.sys.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ void red_add_relaxed_sys_u64(unsigned long long* p, unsigned long long v) {
  asm volatile("red.relaxed.sys.global.add.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}
__device__ __forceinline__ void fence_acq_rel_sys() { asm volatile("fence.acq_rel.sys;" ::: "memory"); }
__device__ __forceinline__ void fence_acq_rel_gpu() { asm volatile("fence.acq_rel.gpu;" ::: "memory"); }
// Orders earlier generic-proxy accesses (ours, or ones we acquired from other SMs / GPUs) before later
// async-proxy (TMA) accesses.
__device__ __forceinline__ void fence_proxy_async() { asm volatile("fence.proxy.async;" ::: "memory"); }

// Sum over the CTA with a fixed shape (lane tree, then warp 0 tree over the warp sums), so the result
// does not depend on scheduling.  `scratch` holds NV * 32 doubles.  Result valid in thread 0.
template <int NV>
__device__ __forceinline__ void block_sum(double (&v)[NV], double* scratch) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nwarps = (blockDim.x + 31) >> 5;
#pragma unroll
  for (int k = 0; k < NV; ++k) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v[k] += __shfl_down_sync(0xffffffffu, v[k], o);
  }
  __syncthreads();  // scratch may still be in use by a previous call
  if (lane == 0) {
#pragma unroll
    for (int k = 0; k < NV; ++k) scratch[k * 32 + warp] = v[k];
  }
  __syncthreads();
  if (warp == 0) {
#pragma unroll
    for (int k = 0; k < NV; ++k) {
      double x = lane < nwarps ? scratch[k * 32 + lane] : 0.0;
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) x += __shfl_down_sync(0xffffffffu, x, o);
      v[k] = x;
    }
  }
}

// Result of a cross-rank reduction, identical in every thread of every CTA of every rank.
struct RedResult {
  double v[kRedValues];
  bool ok;  // false: the watchdog fired
};

constexpr unsigned long long kDefaultWatchdogNs = 20ull * 1000ull * 1000ull * 1000ull;

// Cross-rank sum of NV (<= kRedValues) per-thread values.  Must be called by all threads of all CTAs
// [cta_lo, cta_lo + ncta) of a rank, the same number of times on every rank.  `seq` is the 1-based
// sequence number of this reduction (persisting across launches through Control::red_seq), `cta` the
// CTA index inside the rank.  `smem` provides 64 + 8 doubles of scratch.
// `sys_fences`: the reduction also publishes halo values that were written with plain stores into peer
// memory (system-scope release/acquire around the records).  Kernels that track halo completion with their
// own counters pass false and keep every fence of the reduction at GPU scope.
template <int NV>
__device__ __forceinline__ RedResult cross_rank_sum(const RankCtx& c, double (&val)[NV], unsigned int seq,
                                                    unsigned int arrive_target, int cta, int ncta,
                                                    double* smem, bool did_peer_stores, bool sys_fences) {
  static_assert(NV <= kRedValues, "too many values");
  __shared__ int s_last;
  block_sum<NV>(val, smem);
  if (threadIdx.x == 0) {
#pragma unroll
    for (int k = 0; k < NV; ++k) c.partials[cta * kPartialStride + k] = val[k];
    // release: partials (and any halo values this CTA stored into a neighbour GPU) before the arrival
    if (did_peer_stores && sys_fences)
      fence_acq_rel_sys();
    else
      fence_acq_rel_gpu();
    const unsigned int old = atomicAdd(c.arrive, 1u);
    s_last = (old + 1u == arrive_target) ? 1 : 0;
  }
  __syncthreads();
  if (s_last) {
    // aggregator: acquire everyone's partials, sum them in a fixed tree, publish to all ranks
    if (threadIdx.x == 0) fence_acq_rel_gpu();
    __syncthreads();
    double acc[NV];
#pragma unroll
    for (int k = 0; k < NV; ++k) acc[k] = 0.0;
    for (int i = threadIdx.x; i < ncta; i += blockDim.x) {
#pragma unroll
      for (int k = 0; k < NV; ++k) acc[k] += __ldcg(&c.partials[i * kPartialStride + k]);
    }
    block_sum<NV>(acc, smem);
    if (threadIdx.x == 0) {
#pragma unroll
      for (int k = 0; k < NV; ++k) smem[64 + k] = acc[k];
    }
    __syncthreads();
    if (threadIdx.x < c.nranks) {
      // lane rk publishes this rank's sums into rank rk's slot area (its own included)
      unsigned long long words[kSlotWords];
#pragma unroll
      for (int k = 0; k < kRedValues; ++k) {
        const unsigned long long bits = k < NV ? (unsigned long long)__double_as_longlong(smem[64 + k]) : 0ull;
        words[2 * k] = (bits & 0xffffffffull) | ((unsigned long long)seq << 32);
        words[2 * k + 1] = (bits >> 32) | ((unsigned long long)seq << 32);
      }
      const size_t off = ((size_t)(seq & 1u) * kMaxRanks + (size_t)c.rank) * kSlotWords;
      unsigned long long* dst = c.peer_slots[threadIdx.x] + off;
      if (sys_fences)
        fence_acq_rel_sys();  // release (cumulative over the halo stores acquired through the arrivals)
      else
        fence_acq_rel_gpu();
#pragma unroll
      for (int q = 0; q < kSlotWords; ++q) st_relaxed_sys_u64(dst + q, words[q]);
    }
    __syncthreads();
  }
  // everyone: poll the local slot area
  RedResult res;
  res.ok = true;
  if (threadIdx.x < 32) {
    const int lane = threadIdx.x;
    double mine[kRedValues];
#pragma unroll
    for (int k = 0; k < kRedValues; ++k) mine[k] = 0.0;
    bool ok = true;
    if (lane < c.nranks) {
      const unsigned long long* src = c.slots + ((size_t)(seq & 1u) * kMaxRanks + (size_t)lane) * kSlotWords;
      unsigned long long w[kSlotWords];
      unsigned long long t0 = 0;
      unsigned int spins = 0;
      const unsigned long long limit = c.ctl->watchdog_ns ? c.ctl->watchdog_ns : kDefaultWatchdogNs;
      while (true) {
        bool all = true;
#pragma unroll
        for (int q = 0; q < kSlotWords; ++q) {
          w[q] = ld_relaxed_sys_u64(src + q);
          all = all && ((unsigned int)(w[q] >> 32) == seq);
        }
        if (all) break;
        if ((++spins & 0x3ffu) == 0u) {
          const unsigned long long now = globaltimer_ns();
          if (t0 == 0) t0 = now;
          if (now - t0 > limit) {
            ok = false;
            break;
          }
        }
      }
#pragma unroll
      for (int k = 0; k < kRedValues; ++k) {
        const unsigned long long bits = (w[2 * k] & 0xffffffffull) | (w[2 * k + 1] << 32);
        mine[k] = __longlong_as_double((long long)bits);
      }
    }
    if (sys_fences)
      fence_acq_rel_sys();  // acquire: halo values stored by neighbours before their arrival
    else
      fence_acq_rel_gpu();
    ok = __all_sync(0xffffffffu, ok);
    // rank-ordered sum, identical everywhere
    double tot[kRedValues];
#pragma unroll
    for (int k = 0; k < kRedValues; ++k) {
      tot[k] = 0.0;
      for (int rk = 0; rk < c.nranks; ++rk) tot[k] += __shfl_sync(0xffffffffu, mine[k], rk);
    }
    if (lane == 0) {
#pragma unroll
      for (int k = 0; k < kRedValues; ++k) smem[64 + k] = tot[k];
      smem[64 + kRedValues] = ok ? 1.0 : 0.0;
    }
  }
  __syncthreads();
#pragma unroll
  for (int k = 0; k < kRedValues; ++k) res.v[k] = smem[64 + k];
  res.ok = smem[64 + kRedValues] != 0.0;
  __syncthreads();
  return res;
}

}  // namespace pb
