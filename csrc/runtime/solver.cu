// solver.cu -- see solver.h.
#include "solver.h"

#include <algorithm>
#include <cstddef>
#include <cstring>
#include <vector>

#include "../api/error.h"
#include "../cuda/launch.h"

namespace pb {

static_assert(offsetof(Control, max_iter) - offsetof(Control, delta) == offsetof(DeviceOptions, max_iter),
              "DeviceOptions must mirror the option block of Control");
static_assert(offsetof(Control, watchdog_ns) - offsetof(Control, delta) == offsetof(DeviceOptions, watchdog_ns),
              "DeviceOptions must mirror the option block of Control");
static_assert(sizeof(Control) <= kSlotsOffset, "Control overflows its page");
static_assert(2 * kMaxRanks * kSlotWords * 8 <= kArriveOffset - kSlotsOffset, "slot area overflows");
static_assert(kPartialsOffset + kMaxCtas * kPartialStride * 8 <= kCtaTimesOffset, "partials overflow");
static_assert(kCtaTimesOffset + kMaxCtas * kCtaTimeWords * 8 <= kLl1Offset, "per-CTA report area overflows");
static_assert(kLl1Offset + kRedValues * kMaxCtas * 16 <= kResOffset, "star_sum CTA records overflow");
static_assert(kResOffset + kRedValues * 16 <= kTileCodeOffset, "star_sum result record overflows");
static_assert(kTileCodeOffset + kTileCodeBytes <= kHeaderBytes, "tile codes overflow");
static_assert(kHeaderBytes % 4096 == 0, "arrays must start page-aligned (TMA base alignment)");

Solver::Solver(int M, int N, int Px, int Py, int first_rank, int nlocal, int device)
    : grid_(make_grid(M, N)), pg_{Px, Py}, first_rank_(first_rank), device_(device) {
  if (M < 2 || N < 2) throw std::runtime_error("Solver: M and N must be >= 2");
  if (Px < 1 || Py < 1 || Px * Py > kMaxRanks) throw std::runtime_error("Solver: process grid must have 1..16 ranks");
  if (Px > M - 1 || Py > N - 1) throw std::runtime_error("Solver: more blocks than interior nodes along an axis");
  if (first_rank < 0 || nlocal < 1 || first_rank + nlocal > Px * Py)
    throw std::runtime_error("Solver: local rank range outside the process grid");
  PB_CUDA_CHECK(cudaSetDevice(device_));
  PB_CUDA_CHECK(cudaDeviceGetAttribute(&sm_count_, cudaDevAttrMultiProcessorCount, device_));
  PB_CUDA_CHECK(cudaDeviceGetAttribute(&l2_bytes_, cudaDevAttrL2CacheSize, device_));
  peer_base_.assign(nranks(), nullptr);
  for (int l = 0; l < nlocal; ++l) {
    blocks_.push_back(make_local_block(M, N, pg_, first_rank + l));
    char* base = nullptr;
    const size_t bytes = arena_bytes(blocks_.back());
    PB_CUDA_CHECK(cudaMalloc(&base, bytes));
    PB_CUDA_CHECK(cudaMemset(base, 0, bytes));
    arenas_.push_back(base);
    peer_base_[first_rank + l] = base;
  }
  PB_CUDA_CHECK(cudaMalloc(&dctx_, sizeof(RankCtx) * nlocal));
  PB_CUDA_CHECK(cudaMalloc(&dmaps_, tensor_map_bytes() * kMapVariants * kMapsPerVariant * nlocal));
  PB_CUDA_CHECK(cudaMallocHost(&pinned_opts_, sizeof(DeviceOptions)));
  PB_CUDA_CHECK(cudaMallocHost(&pinned_ctl_, sizeof(Control) * nlocal));
  for (auto& e : ev_) PB_CUDA_CHECK(cudaEventCreate(&e));
  PB_CUDA_CHECK(cudaDeviceSynchronize());
}

Solver::~Solver() {
  cudaSetDevice(device_);
  cudaDeviceSynchronize();
  for (void* p : opened_ipc_) cudaIpcCloseMemHandle(p);
  for (char* a : arenas_) cudaFree(a);
  if (dctx_) cudaFree(dctx_);
  if (dmaps_) cudaFree(dmaps_);
  if (pinned_opts_) cudaFreeHost(pinned_opts_);
  if (pinned_ctl_) cudaFreeHost(pinned_ctl_);
  for (auto& e : ev_) cudaEventDestroy(e);
}

const void* Solver::host_maps(int l) const { return hmaps_ + (size_t)l * kMapVariants * kMapsPerVariant * tensor_map_bytes(); }

double* Solver::array(int l, int which) const {
  if (which < 0 || which > 7) throw std::runtime_error("Solver::array: bad index");
  // arena order: r, w, p0, p1, Ap, a, b, D  (r first so that peers find it at a fixed offset)
  static const int order[8] = {1, 0, 2, 3, 4, 5, 6, 7};  // which -> slot
  return reinterpret_cast<double*>(arenas_[l] + kHeaderBytes) + (size_t)order[which] * blocks_[l].elems();
}

void Solver::ipc_handle(int l, void* out64) const {
  cudaIpcMemHandle_t h;
  PB_CUDA_CHECK(cudaSetDevice(device_));
  PB_CUDA_CHECK(cudaIpcGetMemHandle(&h, arenas_[l]));
  static_assert(sizeof(h) == 64, "unexpected IPC handle size");
  std::memcpy(out64, &h, 64);
}

void Solver::open_peer(int rank, const void* handle64) {
  if (rank < 0 || rank >= nranks()) throw std::runtime_error("open_peer: bad rank");
  if (rank >= first_rank_ && rank < first_rank_ + nlocal()) return;  // local
  cudaIpcMemHandle_t h;
  std::memcpy(&h, handle64, 64);
  void* base = nullptr;
  PB_CUDA_CHECK(cudaSetDevice(device_));
  PB_CUDA_CHECK(cudaIpcOpenMemHandle(&base, h, cudaIpcMemLazyEnablePeerAccess));
  opened_ipc_.push_back(base);
  peer_base_[rank] = base;
}

void Solver::set_peer_base(int rank, void* base) {
  if (rank < 0 || rank >= nranks()) throw std::runtime_error("set_peer_base: bad rank");
  peer_base_[rank] = base;
}

void Solver::finalize() {
  PB_CUDA_CHECK(cudaSetDevice(device_));
  // Ranks whose arena is unknown are unreachable: no halo stores, no in-kernel reductions (the
  // library-collective baseline runs like this); the persistent method then refuses to start.
  fabric_complete_ = true;
  for (int rk = 0; rk < nranks(); ++rk)
    if (!peer_base_[rk]) fabric_complete_ = false;
  hctx_.clear();
  for (int l = 0; l < nlocal(); ++l) {
    const LocalBlock& b = blocks_[l];
    RankCtx c;
    std::memset(&c, 0, sizeof(c));
    c.g = grid_;
    c.i0 = b.i0, c.j0 = b.j0, c.nx = b.nx, c.ny = b.ny, c.pitch = b.pitch;
    c.rank = b.rank, c.nranks = nranks();
    c.w = array(l, 0), c.r = array(l, 1), c.p[0] = array(l, 2), c.p[1] = array(l, 3);
    c.Ap = array(l, 4), c.a = array(l, 5), c.b = array(l, 6), c.D = array(l, 7);
    c.cls = reinterpret_cast<unsigned char*>(arenas_[l] + kHeaderBytes + 8 * b.elems() * sizeof(double));
    for (int d = 0; d < 4; ++d) {
      c.nbr[d] = b.nbr[d];
      if (b.nbr[d] >= 0 && peer_base_[b.nbr[d]]) {
        const LocalBlock nb = make_local_block(grid_.M, grid_.N, pg_, b.nbr[d]);
        c.nbr_r[d] = reinterpret_cast<double*>(static_cast<char*>(peer_base_[b.nbr[d]]) + kHeaderBytes);
        c.nbr_pitch[d] = nb.pitch, c.nbr_nx[d] = nb.nx, c.nbr_ny[d] = nb.ny;
        c.nbr_halo_cnt[d] = reinterpret_cast<unsigned long long*>(static_cast<char*>(peer_base_[b.nbr[d]]) + kHaloCntOffset) + (d ^ 1);
        c.nbr_ll_cap[d] = ll_capacity(nb);
        c.nbr_ll[d] = reinterpret_cast<unsigned long long*>(static_cast<char*>(peer_base_[b.nbr[d]]) + halo_ll_offset(nb)) +
                      (size_t)(d ^ 1) * 2 * c.nbr_ll_cap[d] * 2;
      }
    }
    c.partials = reinterpret_cast<double*>(arenas_[l] + kPartialsOffset);
    c.arrive = reinterpret_cast<unsigned int*>(arenas_[l] + kArriveOffset);
    c.halo_cnt = reinterpret_cast<unsigned long long*>(arenas_[l] + kHaloCntOffset);
    c.halo_ll = reinterpret_cast<unsigned long long*>(arenas_[l] + halo_ll_offset(b));
    c.ll_cap = ll_capacity(b);
    c.slots = reinterpret_cast<unsigned long long*>(arenas_[l] + kSlotsOffset);
    for (int rk = 0; rk < nranks(); ++rk)
      if (peer_base_[rk])
        c.peer_slots[rk] = reinterpret_cast<unsigned long long*>(static_cast<char*>(peer_base_[rk]) + kSlotsOffset);
    c.ctl = control(l);
    c.sys_scope = nlocal() != nranks() ? 1 : 0;
    c.tiles_x = (b.nx + kTileX - 1) / kTileX;
    c.tiles_y = (b.ny + kTileY - 1) / kTileY;
    c.tmaps = dmaps_ + (size_t)l * kMapVariants * kMapsPerVariant * tensor_map_bytes();
    hctx_.push_back(c);
  }
  // tensor maps for both tile heights (host-encoded, copied once)
  hmaps_store_.assign(tensor_map_bytes() * kMapVariants * kMapsPerVariant * nlocal() + 64, 0);
  char* hm = hmaps_ = reinterpret_cast<char*>((reinterpret_cast<uintptr_t>(hmaps_store_.data()) + 63) & ~uintptr_t(63));
  for (int l = 0; l < nlocal(); ++l) {
    char* dst = hm + (size_t)l * kMapVariants * kMapsPerVariant * tensor_map_bytes();
    if (!build_tensor_maps(hctx_[l], 32, dst) || !build_tensor_maps(hctx_[l], 16, dst + kMapsPerVariant * tensor_map_bytes()) ||
        !build_tensor_maps(hctx_[l], 64, dst + 2 * kMapsPerVariant * tensor_map_bytes()))
      throw std::runtime_error("cuTensorMapEncodeTiled failed");
  }
  PB_CUDA_CHECK(cudaMemcpy(dmaps_, hm, tensor_map_bytes() * kMapVariants * kMapsPerVariant * nlocal(), cudaMemcpyHostToDevice));
  PB_CUDA_CHECK(cudaMemcpy(dctx_, hctx_.data(), sizeof(RankCtx) * nlocal(), cudaMemcpyHostToDevice));
  finalized_ = true;
}

}  // namespace pb
