// hop_latency.cu -- what one "store a record, somebody polls it" hop costs on a B200, alone and under contention.
//
// Stand-alone (no library): nvcc -O3 -std=c++17 -gencode arch=compute_100a,code=sm_100a hop_latency.cu -o hop_latency
// The in-kernel reduction of the solver (star_sum, csrc/cuda/device_utils.cuh) is a chain of such hops; this program
// measures the pieces so that the chain can be designed from numbers ("Cross-rank reduction"):
//   T1  ping-pong of one sequence-tagged 16-byte record between two CTAs on different SMs  -> ns per one-way hop
//   T2  one writer, N pollers on the SAME record, every poller acknowledges into a private slot which the writer
//       polls (fan-out through a shared line + fan-in of N records into one warp)           -> ns per round
//   T3  the same, but the writer stores N private mailboxes (one per poller) instead         -> ns per round
//   T4  T1 with a GPU-scope fence before every store (what a release costs on an idle SM)
//   T5  the star: N writers, warp 0 of CTA 0 gathers all records with every load of a look in flight, one result
//       record that everybody polls; record stride 16 / 32 / 128 bytes, with and without the two fences
//   T6  T5 with thread-block clusters of two CTAs: the odd CTA hands its value to the even one through distributed
//       shared memory (cluster barrier), so only half as many records travel through L2
#include <cooperative_groups.h>
#include <cuda_runtime.h>

#include <cstdio>
#include <cstdlib>
#include <vector>

namespace cg = cooperative_groups;

__device__ __forceinline__ unsigned long long gtime() {
  unsigned long long t;
  asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t));
  return t;
}
__device__ __forceinline__ void ld2(const unsigned long long* p, unsigned long long& a, unsigned long long& b) {
  asm volatile("ld.relaxed.gpu.global.v2.u64 {%0, %1}, [%2];" : "=l"(a), "=l"(b) : "l"(p) : "memory");
}
__device__ __forceinline__ void st2(unsigned long long* p, unsigned long long a, unsigned long long b) {
  asm volatile("st.relaxed.gpu.global.v2.u64 [%0], {%1, %2};" ::"l"(p), "l"(a), "l"(b) : "memory");
}
__device__ __forceinline__ void fence_gpu() { asm volatile("fence.acq_rel.gpu;" ::: "memory"); }

__device__ __forceinline__ void wait_rec(const unsigned long long* p, unsigned long long want) {
  unsigned long long a, b;
  do {
    ld2(p, a, b);
  } while (a != want || b != want);
}

constexpr int kLine = 16;  // u64 per 128-byte line

// T1 / T4.  rec[0] (line 0) and rec[kLine] (line 1).
__global__ void pingpong(unsigned long long* rec, int iters, int fenced, unsigned long long* out_ns) {
  if (threadIdx.x != 0 || blockIdx.x > 1) return;
  const int me = blockIdx.x;
  unsigned long long t0 = gtime();
  for (unsigned long long it = 1; it <= (unsigned long long)iters; ++it) {
    if (me == 0) {
      if (fenced) fence_gpu();
      st2(rec, it, it);
      wait_rec(rec + kLine, it);
    } else {
      wait_rec(rec, it);
      if (fenced) fence_gpu();
      st2(rec + kLine, it, it);
    }
  }
  if (me == 0) *out_ns = gtime() - t0;
}

// T2 / T3: CTA 0 is the writer, CTAs 1..N poll.  shared_line = 1: everybody polls rec[0]; 0: poller c polls its own
// mailbox mb[c * stride].  Acks: ack[c * kLine] (one line per poller), polled by warp 0 of CTA 0, lane-strided (one
// record after the other per lane: this fan-in is deliberately naive, T5 does it properly).
__global__ void fanout(unsigned long long* rec, unsigned long long* mb, unsigned long long* ack, int npoll, int iters,
                       int shared_line, int mb_stride, unsigned long long* out_ns) {
  const int cta = blockIdx.x, lane = threadIdx.x;
  if (cta > npoll) return;
  unsigned long long t0 = gtime();
  for (unsigned long long it = 1; it <= (unsigned long long)iters; ++it) {
    if (cta == 0) {
      if (shared_line) {
        if (lane == 0) st2(rec, it, it);
      } else {
        for (int c = 1 + lane; c <= npoll; c += 32) st2(mb + (size_t)c * mb_stride, it, it);
      }
      for (int c = 1 + lane; c <= npoll; c += 32) wait_rec(ack + (size_t)c * kLine, it);
      __syncwarp();
    } else if (lane == 0) {
      wait_rec(shared_line ? rec : mb + (size_t)cta * mb_stride, it);
      st2(ack + (size_t)cta * kLine, it, it);
    }
  }
  if (cta == 0 && lane == 0) *out_ns = gtime() - t0;
}

// T5: the star.
__global__ void star(unsigned long long* rec, unsigned long long* res, int n, int iters, int stride, int fenced,
                     unsigned long long* out_ns) {
  const int cta = blockIdx.x, lane = threadIdx.x;
  if (cta > n) return;
  unsigned long long t0 = gtime();
  for (unsigned long long it = 1; it <= (unsigned long long)iters; ++it) {
    if (cta == 0) {
      while (true) {
        unsigned long long a[5], b[5];
        bool all = true;
#pragma unroll
        for (int j = 0; j < 5; ++j) {
          const int i = 1 + lane + 32 * j;
          if (i <= n) ld2(rec + (size_t)i * stride, a[j], b[j]);
        }
#pragma unroll
        for (int j = 0; j < 5; ++j) {
          const int i = 1 + lane + 32 * j;
          if (i <= n) all = all && a[j] == it && b[j] == it;
        }
        if (__all_sync(0xffffffffu, all)) break;
      }
      if (fenced) fence_gpu();
      if (lane == 0) st2(res, it, it);
    } else if (lane == 0) {
      if (fenced) fence_gpu();
      st2(rec + (size_t)cta * stride, it, it);
      wait_rec(res, it);
    }
  }
  if (cta == 0 && lane == 0) *out_ns = gtime() - t0;
}

// T6: the star with clusters of two.  CTA pair (2q, 2q + 1): the odd CTA writes its value into the even CTA's shared
// memory, cluster barrier, the even CTA stores the pair's record (pair 0's even CTA is the gatherer and keeps its own
// in registers); everybody polls the result.  74 pairs -> 73 records through L2.
__global__ void __cluster_dims__(2, 1, 1) star_clusters(unsigned long long* rec, unsigned long long* res, int npairs, int iters,
                                                       int fenced, unsigned long long* out_ns) {
  __shared__ unsigned long long s_from_peer;
  cg::cluster_group cluster = cg::this_cluster();
  const int cta = blockIdx.x, lane = threadIdx.x, pair = cta >> 1;
  const bool even = (cta & 1) == 0;
  unsigned long long* peer_slot = cluster.map_shared_rank(&s_from_peer, 0);
  unsigned long long t0 = gtime();
  for (unsigned long long it = 1; it <= (unsigned long long)iters; ++it) {
    if (!even && lane == 0) {
      if (fenced) fence_gpu();  // the CTA's data must be released before anybody learns that the CTA is done
      *peer_slot = it;
    }
    cluster.sync();
    if (even) {
      if (pair == 0) {
        while (true) {
          unsigned long long a[3], b[3];
          bool all = s_from_peer == it;
#pragma unroll
          for (int j = 0; j < 3; ++j) {
            const int i = 1 + lane + 32 * j;
            if (i < npairs) ld2(rec + (size_t)i * 2, a[j], b[j]);
          }
#pragma unroll
          for (int j = 0; j < 3; ++j) {
            const int i = 1 + lane + 32 * j;
            if (i < npairs) all = all && a[j] == it && b[j] == it;
          }
          if (__all_sync(0xffffffffu, all)) break;
        }
        if (fenced) fence_gpu();
        if (lane == 0) st2(res, it, it);
      } else if (lane == 0) {
        if (fenced) fence_gpu();
        st2(rec + (size_t)pair * 2, it, it);  // the odd CTA's value arrived with the cluster barrier
      }
    }
    if (lane == 0 && !(even && pair == 0)) wait_rec(res, it);
    __syncwarp();
  }
  if (cta == 0 && lane == 0) *out_ns = gtime() - t0;
}

#define CK(x)                                                                      \
  do {                                                                             \
    cudaError_t e_ = (x);                                                          \
    if (e_ != cudaSuccess) {                                                       \
      std::printf("CUDA error %s at %s:%d\n", cudaGetErrorString(e_), __FILE__, __LINE__); \
      std::exit(1);                                                                \
    }                                                                              \
  } while (0)

int main() {
  int iters = 5000;
  unsigned long long *buf, *out;
  const size_t words = 4096 * kLine;
  CK(cudaMalloc(&buf, words * 8));
  CK(cudaMallocManaged(&out, 8));
  unsigned long long* rec = buf;
  unsigned long long* mb = buf + 64 * kLine;
  unsigned long long* ack = buf + 1024 * kLine;
  unsigned long long* res = buf + 3000 * kLine;
  auto reset = [&] { CK(cudaMemset(buf, 0, words * 8)); };
  for (int fenced = 0; fenced <= 1; ++fenced) {
    reset();
    void* args[] = {&rec, &iters, &fenced, &out};
    CK(cudaLaunchCooperativeKernel((void*)pingpong, dim3(2), dim3(32), args, 0, 0));
    CK(cudaDeviceSynchronize());
    std::printf("{\"test\": \"pingpong\", \"fence_before_store\": %d, \"ns_per_hop\": %.1f}\n", fenced,
                (double)*out / (2.0 * iters));
  }
  for (int shared_line = 1; shared_line >= 0; --shared_line) {
    for (int stride : {4, 16}) {  // mailboxes 32 B apart (4 per line) or one per line
      if (shared_line && stride != 4) continue;
      for (int npoll : {1, 12, 147}) {
        reset();
        void* args[] = {&rec, &mb, &ack, &npoll, &iters, &shared_line, &stride, &out};
        CK(cudaLaunchCooperativeKernel((void*)fanout, dim3(148), dim3(32), args, 0, 0));
        CK(cudaDeviceSynchronize());
        std::printf("{\"test\": \"fanout+naive_fanin\", \"pollers\": %d, \"broadcast\": \"%s\", \"ns_per_round\": %.1f}\n", npoll,
                    shared_line ? "one shared record" : (stride == 4 ? "private mailboxes, 4 per line" : "private mailboxes, 1 per line"),
                    (double)*out / iters);
      }
    }
  }
  for (int fenced = 0; fenced <= 1; ++fenced) {
    for (int stride : {2, 4, 16}) {
      for (int n : {12, 36, 147}) {
        reset();
        void* args[] = {&ack, &res, &n, &iters, &stride, &fenced, &out};
        CK(cudaLaunchCooperativeKernel((void*)star, dim3(148), dim3(32), args, 0, 0));
        CK(cudaDeviceSynchronize());
        std::printf("{\"test\": \"star\", \"writers\": %d, \"record_stride_bytes\": %d, \"fenced\": %d, \"ns_per_round\": %.1f}\n", n,
                    stride * 8, fenced, (double)*out / iters);
      }
    }
  }
  // T6: clusters of two, launched cooperatively (all 148 CTAs must be resident)
  for (int fenced = 0; fenced <= 1; ++fenced) {
    reset();
    int npairs = 74;
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3(148);
    cfg.blockDim = dim3(32);
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeCooperative;
    attr[0].val.cooperative = 1;
    cfg.attrs = attr;
    cfg.numAttrs = 1;
    cudaError_t e = cudaLaunchKernelEx(&cfg, star_clusters, ack, res, npairs, iters, fenced, out);
    if (e == cudaSuccess) e = cudaDeviceSynchronize();
    if (e != cudaSuccess) {
      std::printf("{\"test\": \"star with clusters of 2\", \"fenced\": %d, \"error\": \"%s\"}\n", fenced, cudaGetErrorString(e));
      cudaGetLastError();
      continue;
    }
    std::printf("{\"test\": \"star with clusters of 2 (DSMEM + cluster barrier, then 73 records)\", \"fenced\": %d, \"ns_per_round\": %.1f}\n",
                fenced, (double)*out / iters);
  }
  return 0;
}
