// Baseline JPEG encoder on the GPU (the result path's frame encode).
//
// Replaces cv2.imencode('.jpg', frame, [IMWRITE_JPEG_QUALITY, q]) (/root/reference/src/web_server.py:923 -- one encode
// per camera per MJPEG tick -- and src/snapshot_saver.py:160, :358-404) for frames that are already in HBM: the file
// that comes back is BYTE-IDENTICAL to the one the reference's encoder (libjpeg-turbo inside cv2) writes -- same
// fixed-point colour conversion, 4:2:0 downsampling bias, integer "islow" DCT, quantisation, Annex K Huffman tables,
// edge rules (restated in oracle/jpeg.py, pinned on cv2 itself).
//   jpeg_dct_kernel    64x16 pixels (4 MCUs) per CTA: BGR tile -> smem, colour conversion + chroma downsampling in
//                      smem, 24 blocks x (row pass, column pass, quantise) -> int16 coefficients in zigzag order
//   jpeg_count_kernel  one thread per 8x8 block: the number of bits its Huffman code takes (DC prediction needs only
//                      the neighbour's DC, so blocks are independent)
//   jpeg_scan_kernel   one CTA per frame: exclusive prefix sum of the bit counts (<= 200 k blocks)
//   jpeg_emit_kernel   one thread per block: the code bits OR-ed into the zeroed bit stream at the block's offset
//   jpeg_stuff_kernel  0xFF -> 0xFF 0x00 byte stuffing (count, scan, scatter in one CTA-per-frame launch) + EOI
// HBM bytes per 1080p frame: 6.2 MB pixels read, 6.3 MB coefficients written and read twice, ~1 MB of stream.
#include <vector>

#include "kernels.cuh"
#include "tcd_common.cuh"

namespace tcd {

__constant__ uint8_t c_zigzag_pos[64];       // natural index -> position in zigzag order
__constant__ uint16_t c_huff_code[4][256];   // 0 DC luma, 1 AC luma, 2 DC chroma, 3 AC chroma
__constant__ uint8_t c_huff_size[4][256];

__device__ __forceinline__ int dsc(int x, int n) { return (x + (1 << (n - 1))) >> n; }

// one 8-point pass of the IJG "islow" forward DCT (jfdctint.c); FIRST: row pass (results scaled up by 2^2)
template <bool FIRST>
__device__ __forceinline__ void fdct8(int (&d)[8]) {
  constexpr int CB = 13, P1 = 2;
  constexpr int F0298 = 2446, F0390 = 3196, F0541 = 4433, F0765 = 6270, F0899 = 7373, F1175 = 9633, F1501 = 12299,
                F1847 = 15137, F1961 = 16069, F2053 = 16819, F2562 = 20995, F3072 = 25172;
  int t0 = d[0] + d[7], t7 = d[0] - d[7], t1 = d[1] + d[6], t6 = d[1] - d[6];
  int t2 = d[2] + d[5], t5 = d[2] - d[5], t3 = d[3] + d[4], t4 = d[3] - d[4];
  const int t10 = t0 + t3, t13 = t0 - t3, t11 = t1 + t2, t12 = t1 - t2;
  constexpr int n = FIRST ? CB - P1 : CB + P1;
  if (FIRST) {
    d[0] = (t10 + t11) << P1;
    d[4] = (t10 - t11) << P1;
  } else {
    d[0] = dsc(t10 + t11, P1);
    d[4] = dsc(t10 - t11, P1);
  }
  int z1 = (t12 + t13) * F0541;
  d[2] = dsc(z1 + t13 * F0765, n);
  d[6] = dsc(z1 - t12 * F1847, n);
  z1 = t4 + t7;
  int z2 = t5 + t6, z3 = t4 + t6, z4 = t5 + t7;
  const int z5 = (z3 + z4) * F1175;
  t4 *= F0298; t5 *= F2053; t6 *= F3072; t7 *= F1501;
  z1 *= -F0899; z2 *= -F2562; z3 *= -F1961; z4 *= -F0390;
  z3 += z5; z4 += z5;
  d[7] = dsc(t4 + z1 + z3, n);
  d[5] = dsc(t5 + z2 + z4, n);
  d[3] = dsc(t6 + z2 + z3, n);
  d[1] = dsc(t7 + z1 + z4, n);
}

constexpr int kJpgMcuPerCta = 4;

// grid (ceil(mcus_x / 4), mcus_y, n_img), 256 threads
__global__ void __launch_bounds__(256) jpeg_dct_kernel(const JpegImg* __restrict__ imgs) {
  const JpegImg im = imgs[blockIdx.z];
  const int mx0 = blockIdx.x * kJpgMcuPerCta, my = blockIdx.y;
  if (mx0 >= im.mcus_x || my >= im.mcus_y) return;
  __shared__ uint8_t s_bgr[16][64 * 3 + 4];
  __shared__ int16_t s_y[16][64 + 2];
  __shared__ int16_t s_c[2][8][32 + 2];
  __shared__ int s_blk[24][8][9];        // 24 blocks, row-pass results (padded against bank conflicts)
  const int tid = threadIdx.x;
  // ---- BGR tile with the luma edge rule (clamp to the last column / row)
  const int x0 = mx0 * 16, y0 = my * 16;
  for (int i = tid; i < 16 * 64; i += 256) {
    const int r = i >> 6, c = i & 63;
    const int sy = min(y0 + r, im.h - 1), sx = min(x0 + c, im.w - 1);
    const uint8_t* p = im.src + (long long)sy * im.stride + (long long)sx * 3;
    s_bgr[r][c * 3 + 0] = p[0];
    s_bgr[r][c * 3 + 1] = p[1];
    s_bgr[r][c * 3 + 2] = p[2];
  }
  __syncthreads();
  // ---- luma (level-shifted)
  for (int i = tid; i < 16 * 64; i += 256) {
    const int r = i >> 6, c = i & 63;
    const int b = s_bgr[r][c * 3], g = s_bgr[r][c * 3 + 1], rr = s_bgr[r][c * 3 + 2];
    s_y[r][c] = (int16_t)(((19595 * rr + 38470 * g + 7471 * b + 32768) >> 16) - 128);
  }
  // ---- chroma: 2x2 sums of per-pixel Cb / Cr; rows beyond the (evened) image height repeat the last chroma row
  const int he = (im.h + 1) & ~1;
  for (int i = tid; i < 8 * 32; i += 256) {
    const int cr_ = i >> 5, cc = i & 31;
    const int cyg = min(my * 8 + cr_, he / 2 - 1);      // global chroma row after the bottom rule
    const int r0 = 2 * cyg - y0;                        // tile row (the tile is already clamped to h - 1)
    int sb = 0, sr = 0;
#pragma unroll
    for (int dy = 0; dy < 2; ++dy)
#pragma unroll
      for (int dx = 0; dx < 2; ++dx) {
        const uint8_t* p = &s_bgr[r0 + dy][(2 * cc + dx) * 3];
        const int b = p[0], g = p[1], rr = p[2];
        sb += (-11059 * rr - 21709 * g + 32768 * b + (128 << 16) + 32767) >> 16;
        sr += (32768 * rr - 27439 * g - 5329 * b + (128 << 16) + 32767) >> 16;
      }
    const int bias = (cc & 1) ? 2 : 1;                  // the tile starts at an even chroma column
    s_c[0][cr_][cc] = (int16_t)(((sb + bias) >> 2) - 128);
    s_c[1][cr_][cc] = (int16_t)(((sr + bias) >> 2) - 128);
  }
  __syncthreads();
  // ---- 24 blocks: MCU m holds blocks m*6 + {Y00, Y01, Y10, Y11, Cb, Cr}
  if (tid < 192) {
    const int blk = tid >> 3, row = tid & 7, m = blk / 6, k = blk - m * 6;
    int d[8];
    if (k < 4) {
      const int by = (k >> 1) * 8 + row, bx = m * 16 + (k & 1) * 8;
#pragma unroll
      for (int j = 0; j < 8; ++j) d[j] = s_y[by][bx + j];
    } else {
#pragma unroll
      for (int j = 0; j < 8; ++j) d[j] = s_c[k - 4][row][m * 8 + j];
    }
    fdct8<true>(d);
#pragma unroll
    for (int j = 0; j < 8; ++j) s_blk[blk][row][j] = d[j];
  }
  __syncthreads();
  if (tid < 192) {
    const int blk = tid >> 3, col = tid & 7, m = blk / 6, k = blk - m * 6;
    if (mx0 + m >= im.mcus_x) return;
    int d[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) d[j] = s_blk[blk][j][col];
    fdct8<false>(d);
    const uint16_t* q = im.qtab + (k < 4 ? 0 : 64);
    int16_t* out = im.coef + ((size_t)(my * im.mcus_x + mx0 + m) * 6 + k) * 64;
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      const int nat = j * 8 + col;
      const int qv = (int)q[nat] << 3;
      const int a = abs(d[j]);
      const int r = (a + (qv >> 1)) / qv;
      out[c_zigzag_pos[nat]] = (int16_t)(d[j] < 0 ? -r : r);
    }
  }
}

// The DC a luma block is coded with: dummy blocks (beyond ceil(w/8) x ceil(h/8)) take the DC of the block coded just
// before them in their MCU.
__device__ __forceinline__ bool jpeg_is_dummy(const JpegImg& im, int mcu, int k) {
  if (k >= 4) return false;
  const int my = mcu / im.mcus_x, mx = mcu - my * im.mcus_x;
  return (2 * my + (k >> 1)) >= im.blocks_y || (2 * mx + (k & 1)) >= im.blocks_x;
}
__device__ __forceinline__ int jpeg_dc(const JpegImg& im, int mcu, int k) {
  while (k > 0 && k < 4 && jpeg_is_dummy(im, mcu, k)) --k;
  return im.coef[((size_t)mcu * 6 + k) * 64];
}
__device__ __forceinline__ int jpeg_pred(const JpegImg& im, int mcu, int k) {
  if (k >= 1 && k < 4) return jpeg_dc(im, mcu, k - 1);
  if (mcu == 0) return 0;
  return jpeg_dc(im, mcu - 1, k == 0 ? 3 : k);
}
__device__ __forceinline__ int jpeg_nbits(int v) { return 32 - __clz(abs(v)); }

// MODE 0: count bits; MODE 1: emit bits at bit_off[block]
template <int MODE>
__global__ void __launch_bounds__(128) jpeg_huff_kernel(const JpegImg* __restrict__ imgs) {
  const JpegImg im = imgs[blockIdx.y];
  const int b = blockIdx.x * 128 + threadIdx.x;
  const int nblk = im.mcus_x * im.mcus_y * 6;
  if (b >= nblk) return;
  const int mcu = b / 6, k = b - mcu * 6;
  const int tdc = k < 4 ? 0 : 2, tac = tdc + 1;
  const bool dummy = jpeg_is_dummy(im, mcu, k);
  const int16_t* cf = im.coef + (size_t)b * 64;
  unsigned long long acc = 0;   // MODE 1: pending bits, left-aligned count in nacc
  int nacc = 0;
  unsigned long long pos = MODE ? im.bit_off[b] : 0;
  unsigned int total = 0;
  auto put = [&](unsigned int code, int size) {
    if (MODE == 0) {
      total += size;
      return;
    }
    acc = (acc << size) | (code & ((1u << size) - 1u));
    nacc += size;
    // with 32 or more pending bits: fill the current word to its end (the stream position becomes word aligned), then
    // whole words; what stays pending is always less than a word
    while (nacc >= 32) {
      const int room = 32 - (int)(pos & 31);
      const unsigned int bits = (unsigned int)((acc >> (nacc - room)) & ((1ull << room) - 1ull));
      atomicOr(im.words + (pos >> 5), bits);
      pos += room;
      nacc -= room;
    }
  };
  const int diff = jpeg_dc(im, mcu, k) - jpeg_pred(im, mcu, k);
  int nb = jpeg_nbits(diff);
  put(c_huff_code[tdc][nb], c_huff_size[tdc][nb]);
  if (nb) put((unsigned int)(diff >= 0 ? diff : diff - 1), nb);
  if (dummy) {
    put(c_huff_code[tac][0], c_huff_size[tac][0]);      // EOB: every AC coefficient of a dummy block is zero
  } else {
    int run = 0;
    for (int i = 1; i < 64; ++i) {
      const int v = cf[i];
      if (v == 0) {
        ++run;
        continue;
      }
      while (run > 15) {
        put(c_huff_code[tac][0xF0], c_huff_size[tac][0xF0]);
        run -= 16;
      }
      nb = jpeg_nbits(v);
      const int sym = (run << 4) | nb;
      put(c_huff_code[tac][sym], c_huff_size[tac][sym]);
      put((unsigned int)(v >= 0 ? v : v - 1), nb);
      run = 0;
    }
    if (run > 0) put(c_huff_code[tac][0], c_huff_size[tac][0]);
  }
  if (MODE == 0) {
    im.bit_cnt[b] = total;
  } else {
    while (nacc > 0) {   // the tail: left-align the remaining bits inside their word(s)
      const int room = 32 - (int)(pos & 31);
      const int take = min(room, nacc);
      const unsigned int bits = (unsigned int)((acc >> (nacc - take)) & ((1ull << take) - 1ull));
      atomicOr(im.words + (pos >> 5), bits << (room - take));
      pos += take;
      nacc -= take;
    }
  }
}

// one CTA (1024 threads) per frame: exclusive scan of bit_cnt -> bit_off (64-bit), total -> meta[0]
__global__ void __launch_bounds__(1024) jpeg_scan_kernel(const JpegImg* __restrict__ imgs) {
  const JpegImg im = imgs[blockIdx.x];
  const int n = im.mcus_x * im.mcus_y * 6;
  __shared__ unsigned long long s_part[1024];
  const int tid = threadIdx.x;
  const int per = (n + 1023) / 1024;
  const int lo = min(n, tid * per), hi = min(n, lo + per);
  unsigned long long s = 0;
  for (int i = lo; i < hi; ++i) s += im.bit_cnt[i];
  s_part[tid] = s;
  __syncthreads();
  for (int o = 1; o < 1024; o <<= 1) {
    const unsigned long long v = tid >= o ? s_part[tid - o] : 0;
    __syncthreads();
    s_part[tid] += v;
    __syncthreads();
  }
  unsigned long long run = tid ? s_part[tid - 1] : 0;
  for (int i = lo; i < hi; ++i) {
    im.bit_off[i] = run;
    run += im.bit_cnt[i];
  }
  if (tid == 1023) im.meta[0] = s_part[1023];
}

// one CTA per frame: byte stuffing.  The bit stream is big-endian inside each 32-bit word.
__global__ void __launch_bounds__(1024) jpeg_stuff_kernel(const JpegImg* __restrict__ imgs) {
  const JpegImg im = imgs[blockIdx.x];
  const unsigned long long bits = im.meta[0];
  const long long nbytes = (long long)((bits + 7) >> 3);
  __shared__ long long s_part[1024];
  const int tid = threadIdx.x;
  const long long per = (nbytes + 1023) / 1024;
  const long long lo = min(nbytes, (long long)tid * per), hi = min(nbytes, lo + per);
  auto byte_at = [&](long long i) -> unsigned int {
    unsigned int v = (im.words[i >> 2] >> (24 - 8 * (int)(i & 3))) & 0xffu;
    if (i == nbytes - 1 && (bits & 7)) v |= (1u << (8 - (int)(bits & 7))) - 1u;   // pad the last byte with ones
    return v;
  };
  long long ff = 0;
  for (long long i = lo; i < hi; ++i) ff += byte_at(i) == 0xffu;
  s_part[tid] = ff;
  __syncthreads();
  for (int o = 1; o < 1024; o <<= 1) {
    const long long v = tid >= o ? s_part[tid - o] : 0;
    __syncthreads();
    s_part[tid] += v;
    __syncthreads();
  }
  long long w = lo + (tid ? s_part[tid - 1] : 0);
  uint8_t* out = im.out + im.header_len;
  const long long cap = im.out_cap - im.header_len - 2;
  for (long long i = lo; i < hi; ++i) {
    const unsigned int v = byte_at(i);
    if (w < cap) out[w] = (uint8_t)v;
    ++w;
    if (v == 0xffu) {
      if (w < cap) out[w] = 0;
      ++w;
    }
  }
  if (tid == 1023) {
    const long long total = nbytes + s_part[1023];
    if (total <= cap) {
      out[total] = 0xff;
      out[total + 1] = 0xd9;
    }
    im.meta[1] = (unsigned long long)(im.header_len + total + 2);
    im.meta[2] = total <= cap ? 0ull : 1ull;     // overflow flag
  }
}

int jpeg_set_tables(const uint8_t* zigzag_pos, const uint16_t* code, const uint8_t* size) {
  TCD_CUDA_OK(cudaMemcpyToSymbol(c_zigzag_pos, zigzag_pos, 64));
  TCD_CUDA_OK(cudaMemcpyToSymbol(c_huff_code, code, 4 * 256 * sizeof(uint16_t)));
  TCD_CUDA_OK(cudaMemcpyToSymbol(c_huff_size, size, 4 * 256));
  return TCD_OK;
}

int jpeg_launch(const JpegImg* imgs_dev, int n, int max_mcus_x, int max_mcus_y, cudaStream_t stream) {
  const int max_blocks = max_mcus_x * max_mcus_y * 6;
  jpeg_dct_kernel<<<dim3(cdiv(max_mcus_x, kJpgMcuPerCta), max_mcus_y, n), 256, 0, stream>>>(imgs_dev);
  jpeg_huff_kernel<0><<<dim3(cdiv(max_blocks, 128), n), 128, 0, stream>>>(imgs_dev);
  jpeg_scan_kernel<<<n, 1024, 0, stream>>>(imgs_dev);
  jpeg_huff_kernel<1><<<dim3(cdiv(max_blocks, 128), n), 128, 0, stream>>>(imgs_dev);
  jpeg_stuff_kernel<<<n, 1024, 0, stream>>>(imgs_dev);
  TCD_CUDA_OK(cudaGetLastError());
  return TCD_OK;
}

}  // namespace tcd
