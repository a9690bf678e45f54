// apply_a_op.cu -- stand-alone operator application with fused inner product:  Ap = A p,  dot = (Ap, p).
//
// A library-level version of what the solver does inside its phase AB, for callers that hold their own
// arrays in the padded block layout (csrc/core/decomp.h): rows li = 0..nx+1, columns lj = 0..ny+1 stored at
// [li * pitch + lj + pb_col_offset()], ring cells filled by the caller (zero on the boundary of Omega).
// Replaces apply_A_kernel (S4:507-536) + dot_gpu(Ap, p) (S4:771-786, S4:865) of the reference: one tile
// per CTA writes Ap and one partial sum per tile into a scratch array, a second single-CTA kernel adds the
// partials in a fixed tree (deterministic) and scales by h1*h2.  Nothing is copied to the host.
#include <cuda_runtime.h>

#include "../api/error.h"
#include "launch.h"
#include "phases.cuh"

namespace pb {

constexpr int kOpThreads = 256;

__global__ void __launch_bounds__(kOpThreads)
apply_a_tile_kernel(Grid g, int nx, int ny, int pitch, const double* __restrict__ p, const double* __restrict__ a,
                    const double* __restrict__ b, double* __restrict__ Ap, double* __restrict__ tile_partials) {
  __shared__ double s_red[64];
  const int tiles_y = (ny + kTileY - 1) / kTileY;
  const int tx = blockIdx.x / tiles_y, ty = blockIdx.x - tx * tiles_y;
  const int li0 = 1 + tx * kTileX, lj0 = 1 + ty * kTileY;
  const int cx = min(kTileX, nx - li0 + 1), cy = min(kTileY, ny - lj0 + 1);
  const StencilConsts K = make_stencil_consts(g);
  double acc[1] = {0.0};
  const int col = threadIdx.x % kTileY;
  if (col < cy) {
    for (int row = threadIdx.x / kTileY; row < cx; row += kOpThreads / kTileY) {
      const size_t c = (size_t)(li0 + row) * pitch + (lj0 + col + kColOff);
      const double pc = p[c];
      const double v = stencil_value<kDivMarkstein>(K, pc, p[c - pitch], p[c + pitch], p[c - 1], p[c + 1], a[c],
                                                    a[c + pitch], b[c], b[c + 1]);
      Ap[c] = v;
      acc[0] = fma(v, pc, acc[0]);
    }
  }
  block_sum<1>(acc, s_red);
  if (threadIdx.x == 0) tile_partials[blockIdx.x] = acc[0];
}

__global__ void __launch_bounds__(1024)
apply_a_finalize_kernel(const double* __restrict__ tile_partials, int ntiles, double h1h2, double* __restrict__ dot_out) {
  __shared__ double s_red[64];
  double acc[1] = {0.0};
  for (int i = threadIdx.x; i < ntiles; i += blockDim.x) acc[0] += tile_partials[i];
  block_sum<1>(acc, s_red);
  if (threadIdx.x == 0) *dot_out = acc[0] * h1h2;
}

int apply_a_tiles(int nx, int ny) { return ((nx + kTileX - 1) / kTileX) * ((ny + kTileY - 1) / kTileY); }

// Shared launcher: `scratch` must hold apply_a_tiles(nx, ny) doubles.
cudaError_t launch_apply_a_dot(int M, int N, int nx, int ny, int pitch, const double* p, const double* a,
                               const double* b, double* Ap, double* dot_out, double* scratch, cudaStream_t stream) {
  const Grid g = make_grid(M, N);
  const int ntiles = apply_a_tiles(nx, ny);
  apply_a_tile_kernel<<<ntiles, kOpThreads, 0, stream>>>(g, nx, ny, pitch, p, a, b, Ap, scratch);
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) return e;
  apply_a_finalize_kernel<<<1, 1024, 0, stream>>>(scratch, ntiles, g.h1h2, dot_out);
  return cudaGetLastError();
}

}  // namespace pb

extern "C" int pb_apply_A_dot(int M, int N, int nx, int ny, int pitch, const void* p, const void* a, const void* b,
                              void* Ap, void* dot_out, void* cuda_stream) {
  try {
    if (M < 2 || N < 2 || nx < 1 || ny < 1 || pitch < ny + 2 + pb::kColOff)
      throw std::runtime_error("pb_apply_A_dot: invalid block shape");
    cudaStream_t stream = reinterpret_cast<cudaStream_t>(cuda_stream);
    double* scratch = nullptr;
    PB_CUDA_CHECK(cudaMallocAsync(&scratch, sizeof(double) * (size_t)pb::apply_a_tiles(nx, ny), stream));
    const cudaError_t e = pb::launch_apply_a_dot(M, N, nx, ny, pitch, static_cast<const double*>(p),
                                                 static_cast<const double*>(a), static_cast<const double*>(b),
                                                 static_cast<double*>(Ap), static_cast<double*>(dot_out), scratch, stream);
    PB_CUDA_CHECK(cudaFreeAsync(scratch, stream));
    PB_CUDA_CHECK(e);
    return 0;
  } catch (const std::exception& ex) {
    pb::set_last_error(ex.what());
    return -1;
  }
}

// ---------------------------------------------------------------------------------------------------
// pb_apply_A_dot_noalloc -- the same operator for callers that must not allocate.
//
// Some users of this library cannot let a call reach the device allocator: an application that has given
// every byte of device memory to its own pool (a further cudaMalloc fails or breaks its budget), a service
// on a latency-critical path that cannot wait on the driver's allocation lock, or a host program whose
// rules forbid any allocation once start-up is over.  pb_apply_A_dot takes its per-tile scratch from the
// stream's memory pool on every call; this variant keeps the scratch in the file-scope array below instead
// and never calls cudaMalloc, cudaFree, their async forms, or any other allocating routine.
//
// The array is part of the library image: every GPU on which this library's code gets loaded holds these
// 512 KiB from then until the process exits, whether or not this entry point is ever used.
//
// Limit: kNoAllocMaxTiles tiles of 32 x 64 nodes, i.e. ceil(nx/32) * ceil(ny/64) <= 65536 (for example
// 8192 x 16384 owned nodes).  Larger blocks are rejected with an error; nothing is launched.
//
// Set-up: `g_apply_a_scratch` is a device symbol.  In host code that name is only a stand-in, not a
// device address, so the first call made on a given device (with that device current) asks the runtime for
// the real address once, with cudaGetSymbolAddress, checks the result and caches it per device.  Make that
// first call during initialisation, before any stream capture is begun on the device.  The cached address is
// never refreshed: cudaDeviceReset destroys the context the array lives in and leaves the cache pointing at
// freed memory, so this entry point must not be used on a device after it has been reset.
//
// Sharing: all calls on a device use the one array, and it keeps whatever the previous call left in it.
// Nothing in it has to start at zero -- each call writes every per-tile slot it later reads, and reads
// no other slot -- so no clearing pass is needed.  Two calls whose kernels overlap on the GPU (different
// streams, or different host threads without ordering between them) would write the same slots and race:
// keep calls on one device ordered, e.g. by issuing them into one stream.
namespace pb {

constexpr int kNoAllocMaxTiles = 65536;
static __device__ __align__(256) double g_apply_a_scratch[kNoAllocMaxTiles];

constexpr int kNoAllocMaxDevices = 64;
static double* g_apply_a_scratch_addr[kNoAllocMaxDevices];  // device address of the array, per device ordinal

// Address of the scratch array on the current device; resolved on first use per device.
static cudaError_t noalloc_scratch_address(double** out) {
  int dev = -1;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess) return e;
  if (dev < 0 || dev >= kNoAllocMaxDevices) return cudaErrorInvalidDevice;
  double* addr = __atomic_load_n(&g_apply_a_scratch_addr[dev], __ATOMIC_ACQUIRE);
  if (!addr) {
    void* sym = nullptr;
    e = cudaGetSymbolAddress(&sym, g_apply_a_scratch);
    if (e != cudaSuccess) return e;
    addr = static_cast<double*>(sym);
    __atomic_store_n(&g_apply_a_scratch_addr[dev], addr, __ATOMIC_RELEASE);  // same value from every thread
  }
  *out = addr;
  return cudaSuccess;
}

}  // namespace pb

extern "C" int pb_apply_A_dot_noalloc(int M, int N, int nx, int ny, int pitch, const void* p, const void* a,
                                      const void* b, void* Ap, void* dot_out, void* cuda_stream) {
  try {
    if (M < 2 || N < 2 || nx < 1 || ny < 1 || pitch < ny + 2 + pb::kColOff)
      throw std::runtime_error("pb_apply_A_dot_noalloc: invalid block shape");
    const long long ntiles = (long long)((nx + pb::kTileX - 1) / pb::kTileX) * (long long)((ny + pb::kTileY - 1) / pb::kTileY);
    if (ntiles > pb::kNoAllocMaxTiles)
      throw std::runtime_error("pb_apply_A_dot_noalloc: block of " + std::to_string(nx) + " x " + std::to_string(ny) +
                               " nodes needs " + std::to_string(ntiles) + " tiles of scratch, the built-in array holds " +
                               std::to_string(pb::kNoAllocMaxTiles) + "; use pb_apply_A_dot for blocks this large");
    double* scratch = nullptr;
    PB_CUDA_CHECK(pb::noalloc_scratch_address(&scratch));
    PB_CUDA_CHECK(pb::launch_apply_a_dot(M, N, nx, ny, pitch, static_cast<const double*>(p), static_cast<const double*>(a),
                                         static_cast<const double*>(b), static_cast<double*>(Ap),
                                         static_cast<double*>(dot_out), scratch, reinterpret_cast<cudaStream_t>(cuda_stream)));
    return 0;
  } catch (const std::exception& ex) {
    pb::set_last_error(ex.what());
    return -1;
  }
}
