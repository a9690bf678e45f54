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
__device__ __forceinline__ unsigned long long ld_acquire_gpu_u64(const unsigned long long* p) {
  unsigned long long v;
  asm volatile("ld.acquire.gpu.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
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
// the same for global memory only: what every writing thread executes before the barrier that hands its stores to the
// thread issuing the TMA loads (measured: 44 cycles for a whole CTA)
__device__ __forceinline__ void fence_proxy_async_global() { asm volatile("fence.proxy.async.global;" ::: "memory"); }

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


// ---------------------------------------------------------------------------------------------------
// star_sum: the cross-rank reduction of the TMA kernel.  Same contract as cross_rank_sum (all threads of all CTAs
// of a rank call it, every rank the same number of times, bit-identical result everywhere), different mechanics:
// no atomic, no elected "last arriver", only warp 0 of a CTA takes part, and the only line that many SMs poll is
// polled with one lane and one load per value.  bench/experiments/hop_latency.cu measured the building blocks on a
// B200: a "store a 16-byte record, another SM polls it" hop costs 0.5 us on an idle line, a GPU-scope fence in front
// of the store 0.3 us more, 147 writers + one gathering warp + one result record 1.2 us per round (1.9 us with the
// two fences), records packed 16 bytes apart being the fastest layout; and a two-level tree in which warp 0 of every
// CTA polled 13 group records with 13 lanes was SLOWER (7.5 us) than cross_rank_sum (4.2 us), because ~150 warps
// spinning on the same four lines back up the two L2 slices that hold them.  Hence a star:
//
//   level 0  CTA sum: lane tree per warp -> shared memory -> one __syncthreads -> lane tree over the warp sums
//   level 1  lane 0 releases the CTA's global stores with one GPU-scope fence and writes an LL record (each 8-byte
//            word = 32 bits of payload | sequence number, so the record carries its own validity) into ll1[k][cta]
//   level 2  warp 0 of CTA 0 looks at all records of the rank at once (lane l copies records l, l+32, ... into shared
//            memory with cp.async, every copy of a look in flight together, so the fan-in costs one hop, not one per
//            record, and no registers) and adds them in a fixed order.  Several ranks: it stores the rank's record into every rank's slot area
//            over NVLink, polls its own slot area and adds the ranks' records in rank order.
//   level 3  the same warp fences and stores the result record; lane 0 of warp 0 of every other CTA polls it (147
//            such pollers on one line cost nothing measurable; private mailboxes written by CTA 0 were not faster),
//            then reads it once more as an acquire load.
//   The result goes through shared memory, second __syncthreads.
//
// The fences are the ones the data needs anyway: a reduction is also the grid-wide barrier between the phases (what
// the CTAs stored before it must be readable by every other CTA after it), and the chain member -> CTA 0 -> member is
// release / acquire at every link.  No buffer alternates except the slots: a CTA writes record s+1 only after it
// read result s, which CTA 0 wrote after consuming every record s, and CTA 0 writes result s+1 only after it has
// seen every record s+1.
//
// Replaces dot_gpu + MPI_Allreduce (S4:771-786, S4:871, S4:892, S4:925) like cross_rank_sum, which the plain-load
// persistent kernel keeps because its halo is published by system-scope fences inside the reduction.
__device__ __forceinline__ void ld_relaxed_gpu_v2_u64(const unsigned long long* p, unsigned long long& a,
                                                      unsigned long long& b) {
  asm volatile("ld.relaxed.gpu.global.v2.u64 {%0, %1}, [%2];" : "=l"(a), "=l"(b) : "l"(p) : "memory");
}
__device__ __forceinline__ void st_relaxed_gpu_v2_u64(unsigned long long* p, unsigned long long a, unsigned long long b) {
  asm volatile("st.relaxed.gpu.global.v2.u64 [%0], {%1, %2};" ::"l"(p), "l"(a), "l"(b) : "memory");
}
__device__ __forceinline__ void ld_relaxed_sys_v2_u64(const unsigned long long* p, unsigned long long& a,
                                                      unsigned long long& b) {
  asm volatile("ld.relaxed.sys.global.v2.u64 {%0, %1}, [%2];" : "=l"(a), "=l"(b) : "l"(p) : "memory");
}
__device__ __forceinline__ void st_relaxed_sys_v2_u64(unsigned long long* p, unsigned long long a, unsigned long long b) {
  asm volatile("st.relaxed.sys.global.v2.u64 [%0], {%1, %2};" ::"l"(p), "l"(a), "l"(b) : "memory");
}

__device__ __forceinline__ void ll_encode(double v, unsigned int seq, unsigned long long& w0, unsigned long long& w1) {
  const unsigned long long bits = (unsigned long long)__double_as_longlong(v);
  w0 = (bits & 0xffffffffull) | ((unsigned long long)seq << 32);
  w1 = (bits >> 32) | ((unsigned long long)seq << 32);
}
__device__ __forceinline__ double ll_decode(unsigned long long w0, unsigned long long w1) {
  return __longlong_as_double((long long)((w0 & 0xffffffffull) | (w1 << 32)));
}

// One look at NV LL values, `stride` u64 apart: true when every word carries `seq` (then `out` holds the values).
template <int NV, bool SYS>
__device__ __forceinline__ bool ll_peek(const unsigned long long* rec, size_t stride, unsigned int seq, double (&out)[NV]) {
  bool all = true;
#pragma unroll
  for (int k = 0; k < NV; ++k) {
    unsigned long long w0, w1;
    if (SYS)
      ld_relaxed_sys_v2_u64(rec + k * stride, w0, w1);
    else
      ld_relaxed_gpu_v2_u64(rec + k * stride, w0, w1);
    all = all && ((unsigned int)(w0 >> 32) == seq) && ((unsigned int)(w1 >> 32) == seq);
    out[k] = ll_decode(w0, w1);
  }
  return all;
}

// Poll such a record until it carries `seq` (bounded by `limit` ns).
template <int NV, bool SYS>
__device__ __forceinline__ bool ll_poll(const unsigned long long* rec, size_t stride, unsigned int seq,
                                        unsigned long long limit, double (&out)[NV]) {
  unsigned long long t0 = 0;
  unsigned int spins = 0;
  while (!ll_peek<NV, SYS>(rec, stride, seq, out)) {
    if ((++spins & 0x3ffu) == 0u) {
      const unsigned long long now = globaltimer_ns();
      if (t0 == 0) t0 = now;
      if (now - t0 > limit) return false;
    }
  }
  return true;
}

// fixed-shape sum over the 32 lanes, result in every lane (xor butterfly: partners hold a + b and b + a, the same bits)
__device__ __forceinline__ double warp_sum_all(double x) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) x += __shfl_xor_sync(0xffffffffu, x, o);
  return x;
}

// ---- clusters of two CTAs: the first level of the star --------------------------------------------------------------
// When the kernel is launched with thread-block clusters of two (148 = 2 x 74 CTAs pack perfectly), the odd CTA of a
// pair does not send a record through L2: lane 0 releases the CTA's stores with its fence and writes the CTA's sums
// straight into the even CTA's shared memory (distributed shared memory, st.async completing on the even CTA's
// mbarrier); the even CTA adds them to its own and sends ONE record for the pair.  Half as many records travel to
// CTA 0, and the odd CTA's fence runs beside the even one's instead of in front of a store CTA 0 waits for.
// Measured: the stand-alone model (bench/experiments/hop_latency.cu, 148 one-warp CTAs, fenced) gains 0.6-0.8 us per
// round, but inside the real kernels the reduction gets slower (3.52 against 3.31 us), so the launchers use clusters
// only on request (PB_CLUSTERS=1, see launch_cooperative_pairs).
struct PairState {
  unsigned long long bar;     // mbarrier of the even CTA: one arrival (its own) + the bytes of the partner's sums
  double vals[kRedValues];    // the odd CTA's sums
  unsigned int round;         // reductions so far in this launch (phase parity of `bar`)
};
__device__ __forceinline__ PairState* star_pair_state() {
  __shared__ __align__(16) PairState st;
  return &st;
}
__device__ __forceinline__ unsigned int pb_smem_u32(const void* p) { return (unsigned int)__cvta_generic_to_shared(p); }
__device__ __forceinline__ unsigned int mapa_shared_cluster(unsigned int addr, unsigned int cta_rank) {
  unsigned int r;
  asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(r) : "r"(addr), "r"(cta_rank));
  return r;
}
__device__ __forceinline__ void st_async_cluster_b64(unsigned int remote_addr, unsigned long long v, unsigned int remote_bar) {
  asm volatile("st.async.weak.shared::cluster.mbarrier::complete_tx::bytes.b64 [%0], %1, [%2];" ::"r"(remote_addr), "l"(v),
               "r"(remote_bar)
               : "memory");
}
// all threads of both CTAs of the cluster (kernel start and end only)
__device__ __forceinline__ void cluster_barrier_all() {
  asm volatile("barrier.cluster.arrive.release.aligned;\n\tbarrier.cluster.wait.acquire.aligned;" ::: "memory");
}
// Kernel start (all threads; `pair` is the kernel's launch mode): the even CTA's mbarrier must exist before its
// partner's first write.  Kernel end (all threads, on EVERY exit path): nobody leaves while its partner may still
// write into its shared memory.
__device__ __forceinline__ void star_pair_init(int pair) {
  if (!pair) return;
  PairState* ps = star_pair_state();
  if (threadIdx.x == 0) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(pb_smem_u32(&ps->bar)) : "memory");
    ps->round = 0;
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  cluster_barrier_all();
}
__device__ __forceinline__ void star_pair_exit(int pair) {
  if (pair) cluster_barrier_all();
}
// lane 0 of warp 0 of an even CTA: wait for the partner's sums of this round (bounded) and add them
template <int NV>
__device__ __forceinline__ bool star_pair_receive(double (&tot)[NV], unsigned long long limit) {
  PairState* ps = star_pair_state();
  const unsigned int bar = pb_smem_u32(&ps->bar), parity = ps->round & 1u;
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"((unsigned int)(NV * 8)) : "memory");
  unsigned long long t0 = 0;
  unsigned int spins = 0;
  bool ok = true;
  while (true) {
    unsigned int done;
    asm volatile(
        "{\n.reg .pred p;\n"
        "mbarrier.try_wait.parity.acquire.cluster.shared::cta.b64 p, [%1], %2;\n"
        "selp.u32 %0, 1, 0, p;\n}"
        : "=r"(done)
        : "r"(bar), "r"(parity)
        : "memory");
    if (done) break;
    if ((++spins & 0xffu) == 0u) {
      const unsigned long long now = globaltimer_ns();
      if (t0 == 0) t0 = now;
      if (now - t0 > limit) {
        ok = false;
        break;
      }
    }
  }
  ps->round += 1;
#pragma unroll
  for (int k = 0; k < NV; ++k) tot[k] += ps->vals[k];  // fixed order: even CTA + odd CTA
  return ok;
}
// lane 0 of warp 0 of an odd CTA: release, then hand the sums to the even CTA (cluster rank 0)
template <int NV>
__device__ __forceinline__ void star_pair_send(const double (&tot)[NV]) {
  PairState* ps = star_pair_state();
  fence_acq_rel_gpu();  // release: this CTA's stores of the phase (cumulative over the caller's barrier)
  const unsigned int rvals = mapa_shared_cluster(pb_smem_u32(&ps->vals[0]), 0), rbar = mapa_shared_cluster(pb_smem_u32(&ps->bar), 0);
#pragma unroll
  for (int k = 0; k < NV; ++k) st_async_cluster_b64(rvals + 8u * k, (unsigned long long)__double_as_longlong(tot[k]), rbar);
}

#define PB_EXCHANGE_INLINE __forceinline__

// Staging area of CTA 0's gather: the records travel global -> shared memory with cp.async (L2 only, no register
// per load in flight), kGatherSlots records per value.  One instance per kernel.
constexpr int kGatherSlots = 160;
__device__ __forceinline__ unsigned long long* star_gather_buf() {
  __shared__ __align__(16) unsigned long long buf[kRedValues * kGatherSlots * 2];
  return buf;
}
__device__ __forceinline__ void cp_async_cg_16(void* smem_dst, const void* gmem_src) {
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"((unsigned int)__cvta_generic_to_shared(smem_dst)),
               "l"(gmem_src)
               : "memory");
}
__device__ __forceinline__ void cp_async_wait_all() { asm volatile("cp.async.wait_all;" ::: "memory"); }

// Levels 1-3 of star_sum, executed by warp 0 of every CTA: tot_io holds the CTA's sums on entry and the global sums
// on return.  The gather must not cost the tile loops of the calling kernel any registers (a version that kept its 20
// loads of a look in registers made ptxas spill in the tile loops, inlined or not: phases +10 %), hence cp.async.
// TRACE (micro-benchmark only): thread 0 adds the clock64 distance between the numbered steps to trace[3..7].
template <int NV, bool TRACE>
__device__ PB_EXCHANGE_INLINE bool star_exchange(const RankCtx* cp, double* tot_io, unsigned long long* gbuf, unsigned int seq,
                                                 int cta, int ncta, int pair, unsigned long long limit, long long* trace) {
  const RankCtx& c = *cp;
  const int lane = threadIdx.x & 31;
  long long tr_prev = 0;
  if constexpr (TRACE) tr_prev = clock64();
#define PB_TRACE(i)                                   \
  if constexpr (TRACE) {                              \
    const long long now_ = clock64();                 \
    if (threadIdx.x == 0) trace[i] += now_ - tr_prev; \
    tr_prev = now_;                                   \
  }
  // CTA records, one 16-byte LL pair per value: value k of CTA i at ll1 + (k * kMaxCtas + i) * 2, so that a warp
  // reads 32 consecutive records with one fully coalesced 512-byte load.  The result record `res` (NV pairs) is the
  // one line every other CTA polls.
  unsigned long long* ll1 = reinterpret_cast<unsigned long long*>(reinterpret_cast<char*>(c.ctl) + kLl1Offset);
  unsigned long long* res = reinterpret_cast<unsigned long long*>(reinterpret_cast<char*>(c.ctl) + kResOffset);
  constexpr size_t kValStride = (size_t)kMaxCtas * 2;
  bool ok = true;
  double tot[NV];
#pragma unroll
  for (int k = 0; k < NV; ++k) tot[k] = tot_io[k];
  // clusters of two: the even CTA of a pair speaks for both (record index = pair index), the odd one has handed its
  // sums over already and only waits for the result
  const int rec = pair ? cta >> 1 : cta, nrec = pair ? ncta >> 1 : ncta;
  if (pair && !(cta & 1) && lane == 0) ok = star_pair_receive<NV>(tot, limit);
  if (cta != 0) {
    // ---- level 1: release + record; level 3: wait for the result ----
    if (lane == 0) {
      if (pair && (cta & 1)) {
        star_pair_send<NV>(tot);
      } else {
        fence_acq_rel_gpu();  // release: this CTA's (and its partner's) stores of the phase
#pragma unroll
        for (int k = 0; k < NV; ++k) {
          unsigned long long w0, w1;
          ll_encode(tot[k], ok ? seq : seq ^ 0x80000000u, w0, w1);
          st_relaxed_gpu_v2_u64(ll1 + k * kValStride + (size_t)rec * 2, w0, w1);
        }
      }
      PB_TRACE(3);  // release fence + record store
      if (c.nranks == 1) {
        ok = ll_poll<NV, false>(res, 2, seq, limit, tot);
        PB_TRACE(6);  // result poll
        // acquire: one more look at the record, this time as an acquire load, which synchronises with CTA 0's release
        // (fence + store) -- what every CTA released before its record is readable from here on
        (void)ld_acquire_gpu_u64(res);
        PB_TRACE(7);  // acquire
      }
    }
    if (c.nranks > 1) {
      // Several ranks: every CTA reads the ranks' records in its own slot area directly (lane rk polls rank rk's
      // record) and adds them in rank order: one hop less than going through CTA 0 again.  But not before CTA 0 says
      // that its gather is over (the result record, polled by one lane): nranks lanes per CTA spinning on the slot
      // lines while CTA 0 still gathers slowed that gather down (8 GPUs: 9.6 us per reduction).
      const size_t slot_off = ((size_t)(seq & 1u) * kMaxRanks) * kSlotWords;
      double r[NV];
#pragma unroll
      for (int k = 0; k < NV; ++k) r[k] = 0.0;
      if (lane == 0) {
        double dummy[1];
        ok = ll_poll<1, false>(res, 2, seq, limit, dummy) && ok;
        (void)ld_acquire_gpu_u64(res);  // acquire: CTA 0 stored it after its fence
      }
      __syncwarp();
      if (lane < c.nranks) ok = ll_poll<NV, true>(c.slots + slot_off + (size_t)lane * kSlotWords, 2, seq, limit, r) && ok;
      ok = __all_sync(0xffffffffu, ok);
#pragma unroll
      for (int k = 0; k < NV; ++k) {
        tot[k] = 0.0;
        for (int rk = 0; rk < c.nranks; ++rk) tot[k] += __shfl_sync(0xffffffffu, r[k], rk);
      }
    }
  } else {
    // ---- level 2: CTA 0 gathers the records of the rank (its own from registers; its own stores of the phase are
    //      released by the fence in front of the result store) ----
    {
      // Lane l owns records l, l + 32, ...: the first kGatherSlots records of every value are copied to shared memory
      // with cp.async, all copies of a look in flight together, and checked there; any further ones (more than 160
      // CTAs per rank) are looked at one after the other.
      double acc[NV];
      unsigned long long t0 = 0;
      unsigned int spins = 0;
      const int nfast = min(nrec, kGatherSlots);
      while (true) {
        bool all = true;
        for (int i = lane == 0 ? 32 : lane; i < nfast; i += 32) {
#pragma unroll
          for (int k = 0; k < NV; ++k) cp_async_cg_16(gbuf + ((size_t)k * kGatherSlots + i) * 2, ll1 + k * kValStride + (size_t)i * 2);
        }
        cp_async_wait_all();  // every lane reads back only what it copied itself
#pragma unroll
        for (int k = 0; k < NV; ++k) acc[k] = lane == 0 ? tot[k] : 0.0;  // CTA 0's own sum comes from its registers
        for (int i = lane == 0 ? 32 : lane; i < nfast; i += 32) {
#pragma unroll
          for (int k = 0; k < NV; ++k) {
            const ulonglong2 w = *reinterpret_cast<const ulonglong2*>(gbuf + ((size_t)k * kGatherSlots + i) * 2);
            all = all && ((unsigned int)(w.x >> 32) == seq) && ((unsigned int)(w.y >> 32) == seq);
            acc[k] += ll_decode(w.x, w.y);
          }
        }
        for (int i = lane + kGatherSlots; i < nrec; i += 32) {
          double v[NV];
          all = ll_peek<NV, false>(ll1 + (size_t)i * 2, kValStride, seq, v) && all;
#pragma unroll
          for (int k = 0; k < NV; ++k) acc[k] += v[k];
        }
        if (__all_sync(0xffffffffu, all)) break;
        if ((++spins & 0x3ffu) == 0u) {
          const unsigned long long now = globaltimer_ns();
          if (t0 == 0) t0 = now;
          if (__any_sync(0xffffffffu, now - t0 > limit)) {  // warp-uniform: the loop holds warp-wide votes
            ok = false;
            break;
          }
        }
      }
#pragma unroll
      for (int k = 0; k < NV; ++k) tot[k] = warp_sum_all(acc[k]);
      ok = __all_sync(0xffffffffu, ok);  // lane 0 may have timed out waiting for its partner CTA
    }
    PB_TRACE(4);  // gather of the CTA records
    if (c.nranks > 1) {
      // ---- across the ranks: one record per rank into everybody's slot area, rank-ordered sum ----
      // The fence acquires the CTAs' releases and passes them on with the record that lands in this rank's own slot
      // area, which is what the other CTAs of this rank poll.  (Other ranks never read this rank's arrays through the
      // reduction: halo rows travel as self-validating records.)
      const size_t slot_off = ((size_t)(seq & 1u) * kMaxRanks) * kSlotWords;
      fence_acq_rel_gpu();
      if (lane < c.nranks) {  // lane rk stores this rank's record into rank rk's slot area (its own included)
        unsigned long long* dst = c.peer_slots[lane] + slot_off + (size_t)c.rank * kSlotWords;
#pragma unroll
        for (int k = 0; k < NV; ++k) {
          unsigned long long w0, w1;
          ll_encode(tot[k], ok ? seq : seq ^ 0x80000000u, w0, w1);  // a gather that timed out publishes nothing valid
          st_relaxed_sys_v2_u64(dst + 2 * k, w0, w1);
        }
      }
      if (lane == 0) {  // "the gather of this rank is over": the other CTAs may start polling their slots
        unsigned long long w0, w1;
        ll_encode(0.0, seq, w0, w1);
        st_relaxed_gpu_v2_u64(res, w0, w1);
      }
      double r[NV];
#pragma unroll
      for (int k = 0; k < NV; ++k) r[k] = 0.0;
      if (lane < c.nranks) ok = ll_poll<NV, true>(c.slots + slot_off + (size_t)lane * kSlotWords, 2, seq, limit, r) && ok;
#pragma unroll
      for (int k = 0; k < NV; ++k) {
        tot[k] = 0.0;
        for (int rk = 0; rk < c.nranks; ++rk) tot[k] += __shfl_sync(0xffffffffu, r[k], rk);
      }
    }
    // ---- level 3: the result record (one rank only; with several ranks everybody reads the slots) ----
    if (c.nranks == 1) fence_acq_rel_gpu();  // acquire the CTAs' releases, pass them on with CTA 0's own stores
    ok = __all_sync(0xffffffffu, ok);
    if (c.nranks == 1 && lane < NV) {
      unsigned long long w0, w1;
      // a gather that timed out publishes a record nobody accepts: the other CTAs then time out by themselves
      ll_encode(lane == 0 ? tot[0] : tot[NV - 1], ok ? seq : seq ^ 0x80000000u, w0, w1);
      st_relaxed_gpu_v2_u64(res + 2 * lane, w0, w1);
    }
    PB_TRACE(5);  // cross-rank part, fence, result store
  }
#undef PB_TRACE
  if (lane == 0) {
#pragma unroll
    for (int k = 0; k < NV; ++k) tot_io[k] = tot[k];
  }
  return ok;
}

// `pair`: the kernel runs with clusters of two CTAs (star_pair_init was called); ncta is then even.
template <int NV, bool TRACE = false>
__device__ __forceinline__ RedResult star_sum(const RankCtx& c, double (&val)[NV], unsigned int seq, int cta, int ncta,
                                              unsigned long long limit, double* smem, long long* trace = nullptr, int pair = 0) {
  static_assert(NV <= kRedValues, "too many values");
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nwarps = (blockDim.x + 31) >> 5;  // nwarps <= 16
  long long tr_prev = 0;
  if constexpr (TRACE) tr_prev = clock64();
#define PB_TRACE(i)                                   \
  if constexpr (TRACE) {                              \
    const long long now_ = clock64();                 \
    if (threadIdx.x == 0) trace[i] += now_ - tr_prev; \
    tr_prev = now_;                                   \
  }
  // ---- level 0: CTA sum ----
  // Two values share one shuffle tree: after the first exchange the lower half-warp carries the first value, the
  // upper half-warp the second (6 fp64 shuffle steps instead of 10; the shape is fixed, so the bits are reproducible).
  if constexpr (NV == 2) {
    const bool lo = lane < 16;
    double x = (lo ? val[0] : val[1]) + __shfl_xor_sync(0xffffffffu, lo ? val[1] : val[0], 16);
#pragma unroll
    for (int o = 8; o > 0; o >>= 1) x += __shfl_xor_sync(0xffffffffu, x, o);
    if ((lane & 15) == 0) smem[(lane >> 4) * 32 + warp] = x;
  } else {
#pragma unroll
    for (int k = 0; k < NV; ++k) {
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) val[k] += __shfl_down_sync(0xffffffffu, val[k], o);
    }
    if (lane == 0) {
#pragma unroll
      for (int k = 0; k < NV; ++k) smem[k * 32 + warp] = val[k];
    }
  }
  // Every thread orders its generic-proxy global stores of the phase before later async-proxy (TMA) reads of them;
  // the barrier then hands them to lane 0 of warp 0, whose GPU-scope fence releases them to the other CTAs.
  fence_proxy_async_global();
  PB_TRACE(0);  // warp shuffles, smem write, proxy fence
  __syncthreads();
  PB_TRACE(1);  // barrier
  if (warp == 0) {
    double tot[NV];
    if constexpr (NV == 2) {  // at most 16 warps: the half-warps reduce one value each
      const int wl = lane & 15;
      double x = wl < nwarps ? smem[(lane >> 4) * 32 + wl] : 0.0;
#pragma unroll
      for (int o = 8; o > 0; o >>= 1) x += __shfl_xor_sync(0xffffffffu, x, o);
      tot[0] = __shfl_sync(0xffffffffu, x, 0), tot[1] = __shfl_sync(0xffffffffu, x, 16);
    } else {
#pragma unroll
      for (int k = 0; k < NV; ++k) tot[k] = warp_sum_all(lane < nwarps ? smem[k * 32 + lane] : 0.0);
    }
    PB_TRACE(2);  // lane tree over the warp sums
    double* io = smem + 72;  // 72..73: in/out of the exchange (generic pointer into shared memory)
    if (lane == 0) {
#pragma unroll
      for (int k = 0; k < NV; ++k) io[k] = tot[k];
    }
    __syncwarp();
    const bool ok = star_exchange<NV, TRACE>(&c, io, star_gather_buf(), seq, cta, ncta, pair, limit, trace);
    if (lane == 0) {
#pragma unroll
      for (int k = 0; k < NV; ++k) smem[64 + k] = io[k];
      smem[64 + kRedValues] = ok ? 1.0 : 0.0;
    }
  }
  __syncthreads();
  RedResult res;
#pragma unroll
  for (int k = 0; k < kRedValues; ++k) res.v[k] = k < NV ? smem[64 + k] : 0.0;
  res.ok = smem[64 + kRedValues] != 0.0;
#undef PB_TRACE
  return res;
}

}  // namespace pb
