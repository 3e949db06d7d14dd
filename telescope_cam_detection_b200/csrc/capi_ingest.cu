// Frame ring: the ingest end of the per-frame path.
//
// Replaces the copy chain of the reference's capture loop (/root/reference/src/stream_capture_gpu_ffmpeg.py:253-283):
//   stdout.read(frame_size)            a fresh bytes object per frame (allocation + one copy out of the pipe)
//   torch.from_numpy(..).to(device)    a pageable upload: the driver stages it through its own pinned buffer
//   img.clone() twice                  two device copies (latest_frame and the queue entry)
// with: the decoder's raw bgr24 bytes are read() straight into a PINNED ring slot (no intermediate object), ONE
// asynchronous upload per frame on the ring's copy stream, and consumers get the ring's device slot itself -- a slot is
// only reused n_slots frames later, which is what the two clones bought.  The fd is any byte stream of raw frames:
// ffmpeg's `-f rawvideo -pix_fmt bgr24 pipe:1` in production (the reference's command line), a test writer here.
#include <atomic>
#include <cerrno>
#include <cstring>
#include <unistd.h>
#include <vector>

#include "tcd_common.cuh"

using namespace tcd;

namespace tcd {
extern std::atomic<uint64_t> g_h2d_bytes;
}

struct tcd_frame_ring {
  int device = 0, n = 0;
  size_t frame_bytes = 0;
  uint8_t* host = nullptr;   // n pinned slots
  uint8_t* dev = nullptr;    // n device slots
  cudaStream_t copy = nullptr;
  std::vector<cudaEvent_t> ev;
  std::vector<char> in_flight;
};

extern "C" {

tcd_status tcd_frame_ring_create(int32_t n_slots, int64_t frame_bytes, int32_t device, tcd_frame_ring** out) {
  TCD_REQUIRE(out && n_slots >= 2 && n_slots <= 1024 && frame_bytes > 0, "frame_ring_create: bad arguments");
  TCD_CUDA_OK(cudaSetDevice(device));
  tcd_frame_ring* r = new tcd_frame_ring();
  r->device = device;
  r->n = n_slots;
  r->frame_bytes = (size_t)frame_bytes;
  // slots 256-byte aligned so that every device slot can be a TMA / vector-load base
  const size_t pitch = ((size_t)frame_bytes + 255) & ~(size_t)255;
  cudaError_t e = cudaHostAlloc((void**)&r->host, pitch * n_slots, cudaHostAllocDefault);
  if (e == cudaSuccess) e = cudaMalloc((void**)&r->dev, pitch * n_slots);
  if (e == cudaSuccess) e = cudaStreamCreateWithFlags(&r->copy, cudaStreamNonBlocking);
  r->ev.resize(n_slots);
  r->in_flight.assign(n_slots, 0);
  for (int i = 0; i < n_slots && e == cudaSuccess; ++i) e = cudaEventCreateWithFlags(&r->ev[i], cudaEventDisableTiming);
  if (e != cudaSuccess) {
    set_last_error("frame_ring_create: %s", cudaGetErrorString(e));
    if (r->host) cudaFreeHost(r->host);
    if (r->dev) cudaFree(r->dev);
    delete r;
    return TCD_ERR_CUDA;
  }
  *out = r;
  return TCD_OK;
}

void tcd_frame_ring_destroy(tcd_frame_ring* r) {
  if (!r) return;
  cudaSetDevice(r->device);
  cudaStreamSynchronize(r->copy);
  for (auto e : r->ev) cudaEventDestroy(e);
  cudaStreamDestroy(r->copy);
  cudaFreeHost(r->host);
  cudaFree(r->dev);
  delete r;
}

static inline size_t ring_pitch(const tcd_frame_ring* r) { return (r->frame_bytes + 255) & ~(size_t)255; }

uint8_t* tcd_frame_ring_host(tcd_frame_ring* r, int32_t slot) {
  return (r && slot >= 0 && slot < r->n) ? r->host + ring_pitch(r) * slot : nullptr;
}
const uint8_t* tcd_frame_ring_device(tcd_frame_ring* r, int32_t slot) {
  return (r && slot >= 0 && slot < r->n) ? r->dev + ring_pitch(r) * slot : nullptr;
}

/* Fills the slot's pinned buffer from `fd` (blocking reads until frame_bytes arrived).  *got = bytes read: less than
 * frame_bytes means end of stream (or a read error, TCD_ERR_ARG with errno in the message).  Called without the GIL
 * through ctypes, so a capture thread per camera costs the interpreter nothing while it waits for the decoder. */
tcd_status tcd_frame_ring_read_fd(tcd_frame_ring* r, int32_t slot, int32_t fd, int64_t* got) {
  TCD_REQUIRE(r && got && slot >= 0 && slot < r->n, "frame_ring_read_fd: bad arguments");
  if (r->in_flight[slot]) {   // the slot's previous upload must have left the pinned buffer
    TCD_CUDA_OK(cudaSetDevice(r->device));
    TCD_CUDA_OK(cudaEventSynchronize(r->ev[slot]));
    r->in_flight[slot] = 0;
  }
  uint8_t* dst = r->host + ring_pitch(r) * slot;
  size_t have = 0;
  while (have < r->frame_bytes) {
    const ssize_t k = read(fd, dst + have, r->frame_bytes - have);
    if (k == 0) break;
    if (k < 0) {
      if (errno == EINTR) continue;
      *got = (int64_t)have;
      set_last_error("frame_ring_read_fd: read(%d): %s", fd, strerror(errno));
      return TCD_ERR_INVALID;
    }
    have += (size_t)k;
  }
  *got = (int64_t)have;
  return TCD_OK;
}

/* Starts the slot's upload (pinned -> device slot, asynchronous on the ring's copy stream). */
tcd_status tcd_frame_ring_publish(tcd_frame_ring* r, int32_t slot) {
  TCD_REQUIRE(r && slot >= 0 && slot < r->n, "frame_ring_publish: bad slot");
  TCD_CUDA_OK(cudaSetDevice(r->device));
  const size_t off = ring_pitch(r) * slot;
  TCD_CUDA_OK(cudaMemcpyAsync(r->dev + off, r->host + off, r->frame_bytes, cudaMemcpyHostToDevice, r->copy));
  TCD_CUDA_OK(cudaEventRecord(r->ev[slot], r->copy));
  r->in_flight[slot] = 1;
  g_h2d_bytes.fetch_add((uint64_t)r->frame_bytes, std::memory_order_relaxed);
  return TCD_OK;
}

/* Orders `stream`'s later work after the slot's upload; stream = NULL waits on the host instead. */
tcd_status tcd_frame_ring_wait(tcd_frame_ring* r, int32_t slot, void* stream) {
  TCD_REQUIRE(r && slot >= 0 && slot < r->n, "frame_ring_wait: bad slot");
  TCD_CUDA_OK(cudaSetDevice(r->device));
  if (!r->in_flight[slot]) return TCD_OK;
  if (stream) TCD_CUDA_OK(cudaStreamWaitEvent((cudaStream_t)stream, r->ev[slot], 0));
  else TCD_CUDA_OK(cudaEventSynchronize(r->ev[slot]));
  return TCD_OK;
}

}  // extern "C"
