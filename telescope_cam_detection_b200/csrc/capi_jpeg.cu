// C ABI of the JPEG encoder (include/tcd.h, "JPEG encode").  The handle owns the constant tables, a workspace for
// max_batch frames of max_h x max_w and pinned staging for host frames / the returned files.
#include <cstring>
#include <mutex>
#include <vector>

#include "kernels.cuh"
#include "tcd_common.cuh"

using namespace tcd;

namespace {
const uint8_t kStdLuma[64] = {16, 11, 10, 16, 24, 40, 51, 61, 12, 12, 14, 19, 26, 58, 60, 55, 14, 13, 16, 24, 40, 57,
                              69, 56, 14, 17, 22, 29, 51, 87, 80, 62, 18, 22, 37, 56, 68, 109, 103, 77, 24, 35, 55, 64,
                              81, 104, 113, 92, 49, 64, 78, 87, 103, 121, 120, 101, 72, 92, 95, 98, 112, 100, 103, 99};
const uint8_t kStdChroma[64] = {17, 18, 24, 47, 99, 99, 99, 99, 18, 21, 26, 66, 99, 99, 99, 99, 24, 26, 56, 99, 99, 99,
                                99, 99, 47, 66, 99, 99, 99, 99, 99, 99, 99, 99, 99, 99, 99, 99, 99, 99, 99, 99, 99, 99,
                                99, 99, 99, 99, 99, 99, 99, 99, 99, 99, 99, 99, 99, 99, 99, 99, 99, 99, 99, 99};
const uint8_t kZigzag[64] = {0, 1, 8, 16, 9, 2, 3, 10, 17, 24, 32, 25, 18, 11, 4, 5, 12, 19, 26, 33, 40, 48,
                             41, 34, 27, 20, 13, 6, 7, 14, 21, 28, 35, 42, 49, 56, 57, 50, 43, 36, 29, 22, 15, 23,
                             30, 37, 44, 51, 58, 59, 52, 45, 38, 31, 39, 46, 53, 60, 61, 54, 47, 55, 62, 63};
// Annex K Huffman tables: (bits[16], values)
const uint8_t kDcLBits[16] = {0, 1, 5, 1, 1, 1, 1, 1, 1, 0, 0, 0, 0, 0, 0, 0};
const uint8_t kDcCBits[16] = {0, 3, 1, 1, 1, 1, 1, 1, 1, 1, 1, 0, 0, 0, 0, 0};
const uint8_t kDcVals[12] = {0, 1, 2, 3, 4, 5, 6, 7, 8, 9, 10, 11};
const uint8_t kAcLBits[16] = {0, 2, 1, 3, 3, 2, 4, 3, 5, 5, 4, 4, 0, 0, 1, 0x7d};
const uint8_t kAcLVals[162] = {
    0x01, 0x02, 0x03, 0x00, 0x04, 0x11, 0x05, 0x12, 0x21, 0x31, 0x41, 0x06, 0x13, 0x51, 0x61, 0x07, 0x22, 0x71, 0x14, 0x32, 0x81,
    0x91, 0xa1, 0x08, 0x23, 0x42, 0xb1, 0xc1, 0x15, 0x52, 0xd1, 0xf0, 0x24, 0x33, 0x62, 0x72, 0x82, 0x09, 0x0a, 0x16, 0x17, 0x18,
    0x19, 0x1a, 0x25, 0x26, 0x27, 0x28, 0x29, 0x2a, 0x34, 0x35, 0x36, 0x37, 0x38, 0x39, 0x3a, 0x43, 0x44, 0x45, 0x46, 0x47, 0x48,
    0x49, 0x4a, 0x53, 0x54, 0x55, 0x56, 0x57, 0x58, 0x59, 0x5a, 0x63, 0x64, 0x65, 0x66, 0x67, 0x68, 0x69, 0x6a, 0x73, 0x74, 0x75,
    0x76, 0x77, 0x78, 0x79, 0x7a, 0x83, 0x84, 0x85, 0x86, 0x87, 0x88, 0x89, 0x8a, 0x92, 0x93, 0x94, 0x95, 0x96, 0x97, 0x98, 0x99,
    0x9a, 0xa2, 0xa3, 0xa4, 0xa5, 0xa6, 0xa7, 0xa8, 0xa9, 0xaa, 0xb2, 0xb3, 0xb4, 0xb5, 0xb6, 0xb7, 0xb8, 0xb9, 0xba, 0xc2, 0xc3,
    0xc4, 0xc5, 0xc6, 0xc7, 0xc8, 0xc9, 0xca, 0xd2, 0xd3, 0xd4, 0xd5, 0xd6, 0xd7, 0xd8, 0xd9, 0xda, 0xe1, 0xe2, 0xe3, 0xe4, 0xe5,
    0xe6, 0xe7, 0xe8, 0xe9, 0xea, 0xf1, 0xf2, 0xf3, 0xf4, 0xf5, 0xf6, 0xf7, 0xf8, 0xf9, 0xfa};
const uint8_t kAcCBits[16] = {0, 2, 1, 2, 4, 4, 3, 4, 7, 5, 4, 4, 0, 1, 2, 0x77};
const uint8_t kAcCVals[162] = {
    0x00, 0x01, 0x02, 0x03, 0x11, 0x04, 0x05, 0x21, 0x31, 0x06, 0x12, 0x41, 0x51, 0x07, 0x61, 0x71, 0x13, 0x22, 0x32, 0x81, 0x08,
    0x14, 0x42, 0x91, 0xa1, 0xb1, 0xc1, 0x09, 0x23, 0x33, 0x52, 0xf0, 0x15, 0x62, 0x72, 0xd1, 0x0a, 0x16, 0x24, 0x34, 0xe1, 0x25,
    0xf1, 0x17, 0x18, 0x19, 0x1a, 0x26, 0x27, 0x28, 0x29, 0x2a, 0x35, 0x36, 0x37, 0x38, 0x39, 0x3a, 0x43, 0x44, 0x45, 0x46, 0x47,
    0x48, 0x49, 0x4a, 0x53, 0x54, 0x55, 0x56, 0x57, 0x58, 0x59, 0x5a, 0x63, 0x64, 0x65, 0x66, 0x67, 0x68, 0x69, 0x6a, 0x73, 0x74,
    0x75, 0x76, 0x77, 0x78, 0x79, 0x7a, 0x82, 0x83, 0x84, 0x85, 0x86, 0x87, 0x88, 0x89, 0x8a, 0x92, 0x93, 0x94, 0x95, 0x96, 0x97,
    0x98, 0x99, 0x9a, 0xa2, 0xa3, 0xa4, 0xa5, 0xa6, 0xa7, 0xa8, 0xa9, 0xaa, 0xb2, 0xb3, 0xb4, 0xb5, 0xb6, 0xb7, 0xb8, 0xb9, 0xba,
    0xc2, 0xc3, 0xc4, 0xc5, 0xc6, 0xc7, 0xc8, 0xc9, 0xca, 0xd2, 0xd3, 0xd4, 0xd5, 0xd6, 0xd7, 0xd8, 0xd9, 0xda, 0xe2, 0xe3, 0xe4,
    0xe5, 0xe6, 0xe7, 0xe8, 0xe9, 0xea, 0xf2, 0xf3, 0xf4, 0xf5, 0xf6, 0xf7, 0xf8, 0xf9, 0xfa};

void derive_codes(const uint8_t* bits, const uint8_t* vals, uint16_t* code, uint8_t* size) {
  unsigned int c = 0;
  int k = 0;
  for (int l = 1; l <= 16; ++l) {
    for (int i = 0; i < bits[l - 1]; ++i, ++k) {
      code[vals[k]] = (uint16_t)c++;
      size[vals[k]] = (uint8_t)l;
    }
    c <<= 1;
  }
}

void quality_tables(int quality, uint16_t out[128]) {
  quality = quality < 1 ? 1 : (quality > 100 ? 100 : quality);
  const int scale = quality < 50 ? 5000 / quality : 200 - quality * 2;
  for (int t = 0; t < 2; ++t)
    for (int i = 0; i < 64; ++i) {
      long v = ((long)(t ? kStdChroma[i] : kStdLuma[i]) * scale + 50) / 100;
      out[t * 64 + i] = (uint16_t)(v < 1 ? 1 : (v > 255 ? 255 : v));
    }
}

std::vector<uint8_t> file_header(int h, int w, const uint16_t q[128]) {
  std::vector<uint8_t> o;
  auto put = [&](std::initializer_list<int> b) { for (int v : b) o.push_back((uint8_t)v); };
  put({0xff, 0xd8, 0xff, 0xe0, 0, 16, 'J', 'F', 'I', 'F', 0, 1, 1, 0, 0, 1, 0, 1, 0, 0});
  for (int t = 0; t < 2; ++t) {
    put({0xff, 0xdb, 0, 67, t});
    for (int i = 0; i < 64; ++i) o.push_back((uint8_t)q[t * 64 + kZigzag[i]]);
  }
  put({0xff, 0xc0, 0, 17, 8, h >> 8, h & 255, w >> 8, w & 255, 3, 1, 0x22, 0, 2, 0x11, 1, 3, 0x11, 1});
  struct T { int id; const uint8_t* bits; const uint8_t* vals; int n; };
  const T tabs[4] = {{0x00, kDcLBits, kDcVals, 12}, {0x10, kAcLBits, kAcLVals, 162}, {0x01, kDcCBits, kDcVals, 12},
                     {0x11, kAcCBits, kAcCVals, 162}};
  for (const T& t : tabs) {
    const int len = 2 + 1 + 16 + t.n;
    put({0xff, 0xc4, len >> 8, len & 255, t.id});
    o.insert(o.end(), t.bits, t.bits + 16);
    o.insert(o.end(), t.vals, t.vals + t.n);
  }
  put({0xff, 0xda, 0, 12, 3, 1, 0x00, 2, 0x11, 3, 0x11, 0, 63, 0});
  return o;
}
}  // namespace

struct tcd_jpeg_encoder {
  int device = 0, max_h = 0, max_w = 0, max_batch = 0;
  cudaStream_t stream = nullptr;
  size_t frame_bytes = 0, coef_bytes = 0, cnt_bytes = 0, off_bytes = 0, word_bytes = 0, out_cap = 0, per_frame = 0;
  uint8_t* ws = nullptr;          // per frame: [staged frame][coef][bit_cnt][bit_off][words][out]
  uint16_t* qtab_dev = nullptr;   // [max_batch][128]
  unsigned long long* meta_dev = nullptr;   // [max_batch][4]
  unsigned long long* meta_host = nullptr;  // pinned
  JpegImg* desc_dev = nullptr;
  uint8_t* pinned = nullptr;      // staging for the returned files: max_batch x out_cap would be huge -> sized per call
  size_t pinned_bytes = 0;
  std::mutex mu;
};

extern "C" {

tcd_status tcd_jpeg_encoder_create(int32_t max_h, int32_t max_w, int32_t max_batch, int32_t device, tcd_jpeg_encoder** out) {
  TCD_REQUIRE(out && max_h >= 1 && max_w >= 1 && max_h <= 16384 && max_w <= 16384 && max_batch >= 1 && max_batch <= 256,
              "jpeg_encoder_create: bad arguments");
  TCD_CUDA_OK(cudaSetDevice(device));
  if (int rc = require_sm100()) return (tcd_status)rc;
  tcd_jpeg_encoder* e = new tcd_jpeg_encoder();
  e->device = device;
  e->max_h = max_h;
  e->max_w = max_w;
  e->max_batch = max_batch;
  const size_t mcus = (size_t)cdiv(max_w, 16) * cdiv(max_h, 16), blocks = mcus * 6;
  auto al = [](size_t v) { return (v + 255) & ~(size_t)255; };
  e->frame_bytes = al((size_t)max_h * max_w * 3);
  e->coef_bytes = al(blocks * 64 * sizeof(int16_t));
  e->cnt_bytes = al(blocks * sizeof(unsigned int));
  e->off_bytes = al(blocks * sizeof(unsigned long long));
  e->word_bytes = al(blocks * 208 + 64);           // worst case: 20 DC bits + 63 x 26 AC bits per block
  e->out_cap = al(1024 + 2 * (blocks * 208 + 64)); // every stream byte stuffed
  e->per_frame = e->frame_bytes + e->coef_bytes + e->cnt_bytes + e->off_bytes + e->word_bytes + e->out_cap;
  cudaError_t ce = cudaMalloc((void**)&e->ws, e->per_frame * max_batch);
  if (ce == cudaSuccess) ce = cudaMalloc((void**)&e->qtab_dev, sizeof(uint16_t) * 128 * max_batch);
  if (ce == cudaSuccess) ce = cudaMalloc((void**)&e->meta_dev, sizeof(unsigned long long) * 4 * max_batch);
  if (ce == cudaSuccess) ce = cudaHostAlloc((void**)&e->meta_host, sizeof(unsigned long long) * 4 * max_batch, cudaHostAllocDefault);
  if (ce == cudaSuccess) ce = cudaMalloc((void**)&e->desc_dev, sizeof(JpegImg) * max_batch);
  if (ce == cudaSuccess) ce = cudaStreamCreateWithFlags(&e->stream, cudaStreamNonBlocking);
  if (ce != cudaSuccess) {
    set_last_error("jpeg_encoder_create: %s", cudaGetErrorString(ce));
    cudaFree(e->ws);
    delete e;
    return TCD_ERR_CUDA;
  }
  uint8_t zz_pos[64];
  for (int i = 0; i < 64; ++i) zz_pos[kZigzag[i]] = (uint8_t)i;
  static uint16_t code[4][256];
  static uint8_t size[4][256];
  memset(code, 0, sizeof(code));
  memset(size, 0, sizeof(size));
  derive_codes(kDcLBits, kDcVals, code[0], size[0]);
  derive_codes(kAcLBits, kAcLVals, code[1], size[1]);
  derive_codes(kDcCBits, kDcVals, code[2], size[2]);
  derive_codes(kAcCBits, kAcCVals, code[3], size[3]);
  if (int rc = jpeg_set_tables(zz_pos, &code[0][0], &size[0][0])) {
    tcd_jpeg_encoder_destroy(e);
    return (tcd_status)rc;
  }
  *out = e;
  return TCD_OK;
}

void tcd_jpeg_encoder_destroy(tcd_jpeg_encoder* e) {
  if (!e) return;
  cudaSetDevice(e->device);
  if (e->stream) cudaStreamSynchronize(e->stream);
  cudaFree(e->ws);
  cudaFree(e->qtab_dev);
  cudaFree(e->meta_dev);
  cudaFreeHost(e->meta_host);
  cudaFree(e->desc_dev);
  cudaFreeHost(e->pinned);
  if (e->stream) cudaStreamDestroy(e->stream);
  delete e;
}

/* n BGR u8 frames (host or device pointers, row strides in bytes; NULL / 0 = tight) -> n baseline JPEG files
 * (YCbCr 4:2:0, standard Huffman tables, the IJG quality scale) in HOST buffers out[i] of out_cap[i] bytes;
 * out_size[i] = file length.  The files are byte-identical to cv2.imencode('.jpg', frame, [IMWRITE_JPEG_QUALITY, q]). */
tcd_status tcd_jpeg_encode(tcd_jpeg_encoder* e, const uint8_t* const* frames, const int32_t* h, const int32_t* w,
                           const int64_t* stride, int32_t n, int32_t frames_on_device, int32_t quality,
                           uint8_t* const* out, const int64_t* out_cap, int64_t* out_size) {
  TCD_REQUIRE(e && frames && h && w && out && out_cap && out_size, "jpeg_encode: null argument");
  TCD_REQUIRE(n >= 1 && n <= e->max_batch, "jpeg_encode: %d frames (the encoder was created for %d)", n, e->max_batch);
  std::lock_guard<std::mutex> lock(e->mu);
  TCD_CUDA_OK(cudaSetDevice(e->device));
  uint16_t q[128];
  quality_tables(quality, q);
  std::vector<JpegImg> d(n);
  std::vector<std::vector<uint8_t>> hdr(n);
  int max_mx = 0, max_my = 0;
  for (int i = 0; i < n; ++i) {
    TCD_REQUIRE(h[i] >= 1 && w[i] >= 1 && h[i] <= e->max_h && w[i] <= e->max_w, "jpeg_encode: frame %d is %dx%d (max %dx%d)", i,
                h[i], w[i], e->max_h, e->max_w);
    uint8_t* base = e->ws + e->per_frame * i;
    const size_t tight = (size_t)w[i] * 3;
    const size_t st = stride && stride[i] > 0 ? (size_t)stride[i] : tight;
    JpegImg& m = d[i];
    if (frames_on_device) {
      m.src = frames[i];
      m.stride = (long long)st;
    } else {
      TCD_CUDA_OK(cudaMemcpy2DAsync(base, tight, frames[i], st, tight, (size_t)h[i], cudaMemcpyHostToDevice, e->stream));
      m.src = base;
      m.stride = (long long)tight;
    }
    m.h = h[i];
    m.w = w[i];
    m.mcus_x = cdiv(w[i], 16);
    m.mcus_y = cdiv(h[i], 16);
    m.blocks_x = cdiv(w[i], 8);
    m.blocks_y = cdiv(h[i], 8);
    max_mx = std::max(max_mx, m.mcus_x);
    max_my = std::max(max_my, m.mcus_y);
    m.qtab = e->qtab_dev + 128 * i;
    uint8_t* p = base + e->frame_bytes;
    m.coef = (int16_t*)p; p += e->coef_bytes;
    m.bit_cnt = (unsigned int*)p; p += e->cnt_bytes;
    m.bit_off = (unsigned long long*)p; p += e->off_bytes;
    m.words = (unsigned int*)p; p += e->word_bytes;
    m.out = p;
    m.out_cap = (long long)e->out_cap;
    m.meta = e->meta_dev + 4 * i;
    hdr[i] = file_header(h[i], w[i], q);
    m.header_len = (int)hdr[i].size();
    const size_t blocks = (size_t)m.mcus_x * m.mcus_y * 6;
    TCD_CUDA_OK(cudaMemsetAsync(m.words, 0, blocks * 208 + 64, e->stream));
    TCD_CUDA_OK(cudaMemcpyAsync(e->qtab_dev + 128 * i, q, sizeof(q), cudaMemcpyHostToDevice, e->stream));
  }
  TCD_CUDA_OK(cudaMemcpyAsync(e->desc_dev, d.data(), sizeof(JpegImg) * n, cudaMemcpyHostToDevice, e->stream));
  if (int rc = jpeg_launch(e->desc_dev, n, max_mx, max_my, e->stream)) return (tcd_status)rc;
  TCD_CUDA_OK(cudaMemcpyAsync(e->meta_host, e->meta_dev, sizeof(unsigned long long) * 4 * n, cudaMemcpyDeviceToHost, e->stream));
  TCD_CUDA_OK(cudaStreamSynchronize(e->stream));   // sizes known; `q`, `d` and host frames are done with
  for (int i = 0; i < n; ++i) {
    TCD_REQUIRE(e->meta_host[4 * i + 2] == 0, "jpeg_encode: frame %d overflowed the workspace", i);
    const long long sz = (long long)e->meta_host[4 * i + 1];
    out_size[i] = sz;
    TCD_REQUIRE(sz <= out_cap[i], "jpeg_encode: frame %d needs %lld bytes, the buffer holds %lld", i, sz, (long long)out_cap[i]);
    memcpy(out[i], hdr[i].data(), hdr[i].size());
    TCD_CUDA_OK(cudaMemcpyAsync(out[i] + hdr[i].size(), d[i].out + hdr[i].size(), (size_t)sz - hdr[i].size(), cudaMemcpyDeviceToHost,
                                e->stream));
  }
  TCD_CUDA_OK(cudaStreamSynchronize(e->stream));
  return TCD_OK;
}

/* upper bound of a file for an h x w frame (what out_cap must be to never fail) */
int64_t tcd_jpeg_max_bytes(int32_t h, int32_t w) {
  const int64_t blocks = (int64_t)((w + 15) / 16) * ((h + 15) / 16) * 6;
  return 1024 + 2 * (blocks * 208 + 64);
}

}  // extern "C"
