// ba_stage.cu — HostStager: uploads from PAGEABLE host memory at link speed (ba_internal.h).
//
// The drop-in call hands the solver the reference's own containers (CSfM.cpp:311-343: std::vector<Point2d>,
// std::vector<int>, the map's point storage) — pageable memory.  cudaMemcpyAsync from such memory is staged by the driver
// at ~10 GB/s and blocks the calling thread (11 ms for the 100 MB of a Venice-sized call).  Here a few threads copy
// slices of the source into small pinned rings and hand every chunk to the copy engine as soon as it is staged.
#include <cstring>

#include "ba_internal.h"

namespace hsfm {

bool HostStager::pageable(const void* p) {
  if (!p) return false;
  cudaPointerAttributes a;
  if (cudaPointerGetAttributes(&a, p) != cudaSuccess) { cudaGetLastError(); return true; }  // old-style error for plain host memory
  return a.type == cudaMemoryTypeUnregistered;
}

int HostStager::init() {
  if (pinned) return 0;
  HSFM_CUDA(cudaMallocHost(&pinned, (size_t)LANES * CHUNKS * CHUNK));
  for (int l = 0; l < LANES; ++l)
    for (int c = 0; c < CHUNKS; ++c) {
      HSFM_CUDA(cudaEventCreateWithFlags(&ev[l][c], cudaEventDisableTiming));
      used[l][c] = false;
    }
  return 0;
}

void HostStager::destroy() {
  if (!pinned) return;
  for (int l = 0; l < LANES; ++l)
    for (int c = 0; c < CHUNKS; ++c)
      if (ev[l][c]) { cudaEventDestroy(ev[l][c]); ev[l][c] = nullptr; }
  cudaFreeHost(pinned);
  pinned = nullptr;
}

// One lane's ring: a slot is refilled only after the DMA that last read it has completed.
int HostStager::lane_copy(int lane, void* dst, const void* src, size_t bytes, cudaStream_t st) {
  char* ring = pinned + (size_t)lane * CHUNKS * CHUNK;
  int slot = 0;
  for (size_t off = 0; off < bytes; off += CHUNK, slot = (slot + 1) % CHUNKS) {
    const size_t len = std::min(CHUNK, bytes - off);
    char* buf = ring + (size_t)slot * CHUNK;
    if (used[lane][slot]) HSFM_CUDA(cudaEventSynchronize(ev[lane][slot]));
    std::memcpy(buf, (const char*)src + off, len);
    HSFM_CUDA(cudaMemcpyAsync((char*)dst + off, buf, len, cudaMemcpyHostToDevice, st));
    HSFM_CUDA(cudaEventRecord(ev[lane][slot], st));
    used[lane][slot] = true;
  }
  return 0;
}

namespace {
// slice `part` of `parts` of a job, on 256-byte boundaries
inline void slice(const HostStager::Job& j, int part, int parts, size_t* off, size_t* len) {
  const size_t per = ((j.bytes + parts - 1) / parts + 255) & ~(size_t)255;
  *off = std::min(j.bytes, per * (size_t)part);
  *len = std::min(j.bytes - *off, per);
}
}  // namespace

int HostStager::copy(const std::vector<Job>& jobs, cudaStream_t st, int device) {
  int rc = init();
  if (rc) return rc;
  int rcs[LANES] = {};
  auto work = [&](int lane) {
    if (lane) cudaSetDevice(device);
    for (const Job& j : jobs) {
      size_t off, len;
      slice(j, lane, LANES, &off, &len);
      if (len && !rcs[lane]) rcs[lane] = lane_copy(lane, (char*)j.dst + off, (const char*)j.src + off, len, st);
    }
  };
  std::thread th[LANES - 1];
  for (int l = 1; l < LANES; ++l) th[l - 1] = std::thread(work, l);
  work(0);
  for (int l = 1; l < LANES; ++l) th[l - 1].join();
  for (int l = 0; l < LANES; ++l)
    if (rcs[l]) return rcs[l];
  return 0;
}

void HostStager::start(const std::vector<Job>& jobs, cudaStream_t st, int device, std::vector<std::thread>& threads, int* rc) {
  *rc = init();
  if (*rc) return;
  // lanes 1 .. LANES-1 (the caller must not use copy() until it has joined the threads: the rings are per lane)
  for (int l = 1; l < LANES; ++l)
    threads.emplace_back([this, jobs, st, device, l, rc]() {
      cudaSetDevice(device);
      for (const Job& j : jobs) {
        size_t off, len;
        slice(j, l - 1, LANES - 1, &off, &len);
        if (len) {
          const int r = lane_copy(l, (char*)j.dst + off, (const char*)j.src + off, len, st);
          if (r) *rc = r;  // any failing lane will do
        }
      }
    });
}

}  // namespace hsfm
