// CUDA-core kernels: the thin first convolution (K=25, too thin for tensor cores), the fused
// classifier head, pooling backward, and direct/tiled fallbacks used for unusual shapes and as the
// on-device oracle of the tcgen05 kernels.  Replaces the TensorFlow kernels the reference calls at
// horovod/tensorflow_mnist.py:42-73 and horovod/tensorflow_mnist_gpu.py:47-88.
#include "../include/b200dp.h"
#include "common.cuh"

namespace {

__device__ __forceinline__ float bf(const bf16* p) { return __bfloat162float(*p); }

// ------------------------------------------------------------------------------------------------
// conv1: 5x5 SAME, Cin=1 -> C1, + bias + ReLU + 2x2 maxpool, one image per block.
// Lane = output channel (weights live in registers), input tile + halo in shared memory.
// ------------------------------------------------------------------------------------------------
// bump_step == 2: the control block the step counter lives in follows the engine's layout — hyper-parameters (lr, beta1,
// beta2, eps, eps_mode as doubles) at +16 bytes, the running powers beta1^t / beta2^t (doubles, TensorFlow's beta-power
// accumulators, horovod/tensorflow_mnist.py:130) at +96, and the Adam coefficients of THIS step (5 floats: beta1, beta2, eps,
// step size, v scale) at +128.  Updating the powers by one multiplication here, once per step, takes the two fp64 pow()
// calls out of every optimizer kernel's prologue (2-3 us of a single thread in front of each of them).
__device__ __forceinline__ void adam_coef_update(int* step_ptr) {
  const double* hyper = reinterpret_cast<const double*>(reinterpret_cast<char*>(step_ptr) + 16);
  double* pw = reinterpret_cast<double*>(reinterpret_cast<char*>(step_ptr) + 96);
  float* cf = reinterpret_cast<float*>(reinterpret_cast<char*>(step_ptr) + 128);
  const double lr = hyper[0], b1 = hyper[1], b2 = hyper[2], eps = hyper[3];
  const double b1p = pw[0] * b1, b2p = pw[1] * b2;
  pw[0] = b1p; pw[1] = b2p;
  const double bc1 = 1.0 - b1p, bc2 = 1.0 - b2p;
  cf[0] = (float)b1; cf[1] = (float)b2; cf[2] = (float)eps;
  if (hyper[4] == 0.0) { cf[3] = (float)(lr * sqrt(bc2) / bc1); cf[4] = 1.f; }
  else { cf[3] = (float)(lr / bc1); cf[4] = (float)(1.0 / sqrt(bc2)); }
}

__global__ void __launch_bounds__(256) conv1_pool_fwd_kernel(const float* __restrict__ x, const bf16* __restrict__ w,
                                                             const bf16* __restrict__ bias, bf16* __restrict__ out,
                                                             uint8_t* __restrict__ argmax, int C1, int* step_ptr,
                                                             int bump_step, float* loss_ring, int* correct_ring, int ring) {
  __shared__ float img[32][33];
  const int b = blockIdx.x;
  pdl_wait();
  pdl_launch_dependents();
  if (bump_step && b == 0 && blockIdx.y == 0 && threadIdx.x == 0 && step_ptr) {
    const int s = *step_ptr + 1;          // the step that starts now
    *step_ptr = s;
    if (ring > 0) { loss_ring[s % ring] = 0.f; correct_ring[s % ring] = 0; }
    if (bump_step == 2) adam_coef_update(step_ptr);
  }
  for (int i = threadIdx.x; i < 32 * 32; i += blockDim.x) {
    const int r = i >> 5, c = i & 31;
    const int yy = r - 2, xx = c - 2;
    img[r][c] = (yy >= 0 && yy < 28 && xx >= 0 && xx < 28) ? x[(size_t)b * 784 + yy * 28 + xx] : 0.f;
  }
  __syncthreads();
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nwarp = blockDim.x >> 5;
  for (int c0 = 0; c0 < C1; c0 += 32) {
    const int c = c0 + lane;
    float wr[25];
#pragma unroll
    for (int t = 0; t < 25; ++t) wr[t] = c < C1 ? bf(w + t * C1 + c) : 0.f;
    const float bb = c < C1 ? bf(bias + c) : 0.f;
    for (int p = blockIdx.y * 49 + warp; p < (blockIdx.y + 1) * 49; p += nwarp) {
      const int ph = p / 14, pw = p % 14;
      // the four conv outputs under one pooling window share a 6x6 input patch: read it once into registers
      float win[6][6];
#pragma unroll
      for (int r = 0; r < 6; ++r)
#pragma unroll
        for (int cc = 0; cc < 6; ++cc) win[r][cc] = img[2 * ph + r][2 * pw + cc];
      float best = -1.f;
      int bi = 0;
#pragma unroll
      for (int q = 0; q < 4; ++q) {
        float acc = bb;
#pragma unroll
        for (int kh = 0; kh < 5; ++kh)
#pragma unroll
          for (int kw = 0; kw < 5; ++kw) acc = fmaf(win[(q >> 1) + kh][(q & 1) + kw], wr[kh * 5 + kw], acc);
        acc = fmaxf(acc, 0.f);
        if (acc > best) { best = acc; bi = q; }
      }
      if (c < C1) {
        const size_t o = ((size_t)b * 196 + p) * C1 + c;
        out[o] = __float2bfloat16(best);
        argmax[o] = (uint8_t)bi;
      }
    }
  }
}

// conv1 weight gradient straight from the pooled-activation gradient: the pooling window passes the
// gradient to its argmax position only, so
//   dW[tap,c] = sum_{b,ph,pw} x[b, 2ph+i*+kh-2, 2pw+j*+kw-2] * da1[b,ph,pw,c] * (a1 > 0).
__global__ void __launch_bounds__(256) conv1_wgrad_kernel(const float* __restrict__ x, const bf16* __restrict__ da1,
                                                          const bf16* __restrict__ a1, const uint8_t* __restrict__ argmax,
                                                          float* __restrict__ dw, float* __restrict__ dbias, int B, int C1) {
  __shared__ float img[32][33];
  __shared__ float red[8][26][32];
  pdl_wait();
  pdl_launch_dependents();
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nwarp = blockDim.x >> 5;
  for (int c0 = 0; c0 < C1; c0 += 32) {
    const int c = c0 + lane;
    float acc[26];
#pragma unroll
    for (int t = 0; t < 26; ++t) acc[t] = 0.f;
    for (int b = blockIdx.x; b < B; b += gridDim.x) {
      __syncthreads();
      for (int i = threadIdx.x; i < 32 * 32; i += blockDim.x) {
        const int r = i >> 5, cc = i & 31;
        const int yy = r - 2, xx = cc - 2;
        img[r][cc] = (yy >= 0 && yy < 28 && xx >= 0 && xx < 28) ? x[(size_t)b * 784 + yy * 28 + xx] : 0.f;
      }
      __syncthreads();
      if (c < C1) {
        // all global loads of this warp's (at most 7) pixels are issued before the first FMA
        float gv[7];
        int qv[7];
#pragma unroll
        for (int i = 0; i < 7; ++i) {
          const int p = blockIdx.y * 49 + warp + i * 8;
          gv[i] = 0.f; qv[i] = 0;
          if (warp + i * 8 < 49) {
            const size_t o = ((size_t)b * 196 + p) * C1 + c;
            gv[i] = bf(a1 + o) > 0.f ? bf(da1 + o) : 0.f;
            qv[i] = argmax[o];
          }
        }
#pragma unroll
        for (int i = 0; i < 7; ++i) {
          if (warp + i * 8 >= 49) continue;
          const int p = blockIdx.y * 49 + warp + i * 8;
          const float g = gv[i];
          const int q = qv[i];
          const int y = 2 * (p / 14) + (q >> 1), xq = 2 * (p % 14) + (q & 1);
#pragma unroll
          for (int kh = 0; kh < 5; ++kh)
#pragma unroll
            for (int kw = 0; kw < 5; ++kw) acc[kh * 5 + kw] = fmaf(img[y + kh][xq + kw], g, acc[kh * 5 + kw]);
          acc[25] += g;
        }
      }
    }
    __syncthreads();
#pragma unroll
    for (int t = 0; t < 26; ++t) red[warp][t][lane] = acc[t];
    __syncthreads();
    for (int i = threadIdx.x; i < 26 * 32; i += blockDim.x) {
      const int t = i >> 5, l = i & 31;
      float s = 0.f;
      for (int wv = 0; wv < nwarp; ++wv) s += red[wv][t][l];
      if (c0 + l < C1) {
        if (t < 25) atomicAdd(dw + t * C1 + c0 + l, s);
        else atomicAdd(dbias + c0 + l, s);
      }
    }
  }
}

// ------------------------------------------------------------------------------------------------
// keras128 first layer: KxK VALID, Cin=1 -> C1, + bias + ReLU (no pooling), NHWC bf16 out.
// Lane = output channel (taps in registers), the image in shared memory; blockIdx.y splits the pixels.
// Keras: Conv2D(32, (3,3), activation='relu') — the shape BASELINE.json's prose names.
// ------------------------------------------------------------------------------------------------
template <int K>
__global__ void __launch_bounds__(256) conv1_valid_fwd_kernel(const float* __restrict__ x, const bf16* __restrict__ w,
                                                              const bf16* __restrict__ bias, bf16* __restrict__ out, int C1,
                                                              int* step_ptr, int bump_step, float* loss_ring,
                                                              int* correct_ring, int ring) {
  constexpr int HO = 28 - K + 1, NPIX = HO * HO;
  __shared__ float img[28][29];
  const int b = blockIdx.x;
  pdl_wait();
  pdl_launch_dependents();
  if (bump_step && b == 0 && blockIdx.y == 0 && threadIdx.x == 0 && step_ptr) {
    const int s = *step_ptr + 1;
    *step_ptr = s;
    if (ring > 0) { loss_ring[s % ring] = 0.f; correct_ring[s % ring] = 0; }
    if (bump_step == 2) adam_coef_update(step_ptr);
  }
  for (int i = threadIdx.x; i < 784; i += blockDim.x) img[i / 28][i % 28] = x[(size_t)b * 784 + i];
  __syncthreads();
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nwarp = blockDim.x >> 5;
  const int per = (NPIX + gridDim.y - 1) / gridDim.y;
  const int p0 = blockIdx.y * per, p1 = min(NPIX, p0 + per);
  for (int c0 = 0; c0 < C1; c0 += 32) {
    const int c = c0 + lane;
    float wr[K * K];
#pragma unroll
    for (int t = 0; t < K * K; ++t) wr[t] = c < C1 ? bf(w + t * C1 + c) : 0.f;
    const float bb = c < C1 ? bf(bias + c) : 0.f;
    for (int p = p0 + warp; p < p1; p += nwarp) {
      const int y = p / HO, xq = p % HO;
      float acc = bb;
#pragma unroll
      for (int kh = 0; kh < K; ++kh)
#pragma unroll
        for (int kw = 0; kw < K; ++kw) acc = fmaf(img[y + kh][xq + kw], wr[kh * K + kw], acc);
      if (c < C1) out[((size_t)b * NPIX + p) * C1 + c] = __float2bfloat16(fmaxf(acc, 0.f));
    }
  }
}

// dW[tap,c] = sum_{b,pix} x[b, pix+tap] * da1[b,pix,c] * (a1 > 0);  dbias[c] = the same sum without x.
template <int K>
__global__ void __launch_bounds__(256) conv1_valid_wgrad_kernel(const float* __restrict__ x, const bf16* __restrict__ da1,
                                                                const bf16* __restrict__ a1, float* __restrict__ dw,
                                                                float* __restrict__ dbias, int B, int C1) {
  constexpr int HO = 28 - K + 1, NPIX = HO * HO, NT = K * K + 1;
  __shared__ float img[28][29];
  __shared__ float red[8][NT][32];
  pdl_wait();
  pdl_launch_dependents();
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nwarp = blockDim.x >> 5;
  const int per = (NPIX + gridDim.y - 1) / gridDim.y;
  const int p0 = blockIdx.y * per, p1 = min(NPIX, p0 + per);
  for (int c0 = 0; c0 < C1; c0 += 32) {
    const int c = c0 + lane;
    float acc[NT];
#pragma unroll
    for (int t = 0; t < NT; ++t) acc[t] = 0.f;
    for (int b = blockIdx.x; b < B; b += gridDim.x) {
      __syncthreads();
      for (int i = threadIdx.x; i < 784; i += blockDim.x) img[i / 28][i % 28] = x[(size_t)b * 784 + i];
      __syncthreads();
      if (c < C1) {
        for (int pb = p0 + warp; pb < p1; pb += 4 * nwarp) {
          // four pixels per trip: their loads are in flight together
          float gv[4];
#pragma unroll
          for (int i = 0; i < 4; ++i) {
            const int p = pb + i * nwarp;
            gv[i] = 0.f;
            if (p < p1) {
              const size_t o = ((size_t)b * NPIX + p) * C1 + c;
              gv[i] = bf(a1 + o) > 0.f ? bf(da1 + o) : 0.f;
            }
          }
#pragma unroll
          for (int i = 0; i < 4; ++i) {
            const int p = pb + i * nwarp;
            if (p >= p1) continue;
            const int y = p / HO, xq = p % HO;
            const float g = gv[i];
#pragma unroll
            for (int kh = 0; kh < K; ++kh)
#pragma unroll
              for (int kw = 0; kw < K; ++kw) acc[kh * K + kw] = fmaf(img[y + kh][xq + kw], g, acc[kh * K + kw]);
            acc[K * K] += g;
          }
        }
      }
    }
    __syncthreads();
#pragma unroll
    for (int t = 0; t < NT; ++t) red[warp][t][lane] = acc[t];
    __syncthreads();
    for (int i = threadIdx.x; i < NT * 32; i += blockDim.x) {
      const int t = i >> 5, l = i & 31;
      float s = 0.f;
      for (int wv = 0; wv < nwarp; ++wv) s += red[wv][t][l];
      if (c0 + l < C1) {
        if (t < K * K) atomicAdd(dw + t * C1 + c0 + l, s);
        else atomicAdd(dbias + c0 + l, s);
      }
    }
  }
}

// 3x3 variants with eight channels per lane: 4 lanes cover a 32-channel chunk of one pixel, a warp works on 8 consecutive
// pixels, so activations move as 16-byte accesses (512 contiguous bytes per warp) instead of 2-byte ones.  C1 % 8 == 0.
__device__ __forceinline__ void unpack_bf16x8(const uint4 v, float* f) {
  const uint32_t w[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
  for (int j = 0; j < 4; ++j) {
    f[2 * j] = __uint_as_float(w[j] << 16);
    f[2 * j + 1] = __uint_as_float(w[j] & 0xFFFF0000u);
  }
}

__global__ void __launch_bounds__(256) conv1_valid3_fwd_v8_kernel(const float* __restrict__ x, const bf16* __restrict__ w,
                                                                  const bf16* __restrict__ bias, bf16* __restrict__ out, int C1,
                                                                  int* step_ptr, int bump_step, float* loss_ring,
                                                                  int* correct_ring, int ring) {
  constexpr int K = 3, HO = 26, NPIX = HO * HO;
  __shared__ float img[28][29];
  const int b = blockIdx.x;
  pdl_wait();
  pdl_launch_dependents();
  if (bump_step && b == 0 && blockIdx.y == 0 && threadIdx.x == 0 && step_ptr) {
    const int s = *step_ptr + 1;
    *step_ptr = s;
    if (ring > 0) { loss_ring[s % ring] = 0.f; correct_ring[s % ring] = 0; }
    if (bump_step == 2) adam_coef_update(step_ptr);
  }
  for (int i = threadIdx.x; i < 784; i += blockDim.x) img[i / 28][i % 28] = x[(size_t)b * 784 + i];
  __syncthreads();
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nwarp = blockDim.x >> 5;
  const int sub = lane >> 2, cq = lane & 3;
  const int per = (NPIX + gridDim.y - 1) / gridDim.y;
  const int p0 = blockIdx.y * per, p1 = min(NPIX, p0 + per);
  for (int c0 = 0; c0 < C1; c0 += 32) {
    const int c = c0 + cq * 8;
    if (c >= C1) continue;
    float wr[K * K][8], bb[8];
#pragma unroll
    for (int t = 0; t < K * K; ++t) unpack_bf16x8(__ldg(reinterpret_cast<const uint4*>(w + t * C1 + c)), wr[t]);
    unpack_bf16x8(__ldg(reinterpret_cast<const uint4*>(bias + c)), bb);
    for (int p = p0 + warp * 8 + sub; p < p1; p += nwarp * 8) {
      const int y = p / HO, xq = p % HO;
      float acc[8];
#pragma unroll
      for (int e = 0; e < 8; ++e) acc[e] = bb[e];
#pragma unroll
      for (int kh = 0; kh < K; ++kh)
#pragma unroll
        for (int kw = 0; kw < K; ++kw) {
          const float v = img[y + kh][xq + kw];
#pragma unroll
          for (int e = 0; e < 8; ++e) acc[e] = fmaf(v, wr[kh * K + kw][e], acc[e]);
        }
      const uint2 lo = pack_bf16x4(make_float4(fmaxf(acc[0], 0.f), fmaxf(acc[1], 0.f), fmaxf(acc[2], 0.f), fmaxf(acc[3], 0.f)));
      const uint2 hi = pack_bf16x4(make_float4(fmaxf(acc[4], 0.f), fmaxf(acc[5], 0.f), fmaxf(acc[6], 0.f), fmaxf(acc[7], 0.f)));
      *reinterpret_cast<uint4*>(out + ((size_t)b * NPIX + p) * C1 + c) = make_uint4(lo.x, lo.y, hi.x, hi.y);
    }
  }
}

__global__ void __launch_bounds__(256) conv1_valid3_wgrad_v8_kernel(const float* __restrict__ x, const bf16* __restrict__ da1,
                                                                    const bf16* __restrict__ a1, float* __restrict__ dw,
                                                                    float* __restrict__ dbias, int B, int C1) {
  constexpr int K = 3, HO = 26, NPIX = HO * HO, NT = K * K + 1;
  __shared__ float img[28][29];
  __shared__ float red[8][NT][32];
  pdl_wait();
  pdl_launch_dependents();
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nwarp = blockDim.x >> 5;
  const int sub = lane >> 2, cq = lane & 3;
  const int per = (NPIX + gridDim.y - 1) / gridDim.y;
  const int p0 = blockIdx.y * per, p1 = min(NPIX, p0 + per);
  for (int c0 = 0; c0 < C1; c0 += 32) {
    const int c = c0 + cq * 8;
    const bool cvalid = c < C1;
    float acc[NT][8];
#pragma unroll
    for (int t = 0; t < NT; ++t)
#pragma unroll
      for (int e = 0; e < 8; ++e) acc[t][e] = 0.f;
    for (int b = blockIdx.x; b < B; b += gridDim.x) {
      __syncthreads();
      for (int i = threadIdx.x; i < 784; i += blockDim.x) img[i / 28][i % 28] = x[(size_t)b * 784 + i];
      __syncthreads();
      if (cvalid) {
        for (int pb = p0 + warp * 8 + sub; pb < p1; pb += 2 * nwarp * 8) {
          // two pixels per trip: four 16-byte loads in flight per lane
          uint4 dv[2], av[2];
#pragma unroll
          for (int i = 0; i < 2; ++i) {
            const int p = pb + i * nwarp * 8;
            if (p < p1) {
              const size_t o = ((size_t)b * NPIX + p) * C1 + c;
              dv[i] = __ldg(reinterpret_cast<const uint4*>(da1 + o));
              av[i] = __ldg(reinterpret_cast<const uint4*>(a1 + o));
            }
          }
#pragma unroll
          for (int i = 0; i < 2; ++i) {
            const int p = pb + i * nwarp * 8;
            if (p >= p1) continue;
            float g[8], a[8];
            unpack_bf16x8(dv[i], g);
            unpack_bf16x8(av[i], a);
#pragma unroll
            for (int e = 0; e < 8; ++e) g[e] = a[e] > 0.f ? g[e] : 0.f;
            const int y = p / HO, xq = p % HO;
#pragma unroll
            for (int kh = 0; kh < K; ++kh)
#pragma unroll
              for (int kw = 0; kw < K; ++kw) {
                const float v = img[y + kh][xq + kw];
#pragma unroll
                for (int e = 0; e < 8; ++e) acc[kh * K + kw][e] = fmaf(v, g[e], acc[kh * K + kw][e]);
              }
#pragma unroll
            for (int e = 0; e < 8; ++e) acc[K * K][e] += g[e];
          }
        }
      }
    }
    // fold the 8 pixel slots of the warp (lane bits 2..4), then the warps, then one atomic per (tap, channel) per block
#pragma unroll
    for (int t = 0; t < NT; ++t)
#pragma unroll
      for (int e = 0; e < 8; ++e) {
        float v = acc[t][e];
        v += __shfl_xor_sync(0xffffffffu, v, 4);
        v += __shfl_xor_sync(0xffffffffu, v, 8);
        v += __shfl_xor_sync(0xffffffffu, v, 16);
        acc[t][e] = v;
      }
    __syncthreads();
    if (sub == 0) {
#pragma unroll
      for (int t = 0; t < NT; ++t)
#pragma unroll
        for (int e = 0; e < 8; ++e) red[warp][t][cq * 8 + e] = acc[t][e];
    }
    __syncthreads();
    for (int i = threadIdx.x; i < NT * 32; i += blockDim.x) {
      const int t = i >> 5, l = i & 31;
      float sum = 0.f;
      for (int wv = 0; wv < nwarp; ++wv) sum += red[wv][t][l];
      if (c0 + l < C1) {
        if (t < K * K) atomicAdd(dw + t * C1 + c0 + l, sum);
        else atomicAdd(dbias + c0 + l, sum);
      }
    }
  }
}

// Inverted dropout over a bf16 tensor (in place allowed): element i keeps its value * 1/(1-rate) when Philox lane i%4 of
// counter (i/4, 0, step, layer) is >= rate * 2^32 — the same stream b200dp/utils/philox.py draws on the host.
__global__ void __launch_bounds__(256) dropout_bf16_kernel(const bf16* __restrict__ in, bf16* __restrict__ out, long long n4,
                                                           float rate, uint64_t seed, const int* step_ptr, int layer_id) {
  pdl_wait();
  pdl_launch_dependents();
  const uint32_t step = step_ptr ? (uint32_t)*step_ptr : 0u;
  const uint32_t thresh = (uint32_t)fminf(rate * 4294967296.f, 4294967295.f);
  const float sc = 1.f / (1.f - rate);
  for (long long q = (long long)blockIdx.x * blockDim.x + threadIdx.x; q < n4; q += (long long)gridDim.x * blockDim.x) {
    const uint4 rnd = dropout_rand4((uint32_t)q, step, (uint32_t)layer_id, seed);
    const uint2 raw = *reinterpret_cast<const uint2*>(in + 4 * q);
    const __nv_bfloat162 lo = *reinterpret_cast<const __nv_bfloat162*>(&raw.x), hi = *reinterpret_cast<const __nv_bfloat162*>(&raw.y);
    float4 v = make_float4(__low2float(lo), __high2float(lo), __low2float(hi), __high2float(hi));
    v.x = rnd.x >= thresh ? v.x * sc : 0.f;
    v.y = rnd.y >= thresh ? v.y * sc : 0.f;
    v.z = rnd.z >= thresh ? v.z * sc : 0.f;
    v.w = rnd.w >= thresh ? v.w * sc : 0.f;
    *reinterpret_cast<uint2*>(out + 4 * q) = pack_bf16x4(v);
  }
}

// ------------------------------------------------------------------------------------------------
// Classifier head, one batch row per block:
//   h = dropout(relu(z + b1));  logits = h W2 + b2;  loss/acc;  dlogits = (softmax - onehot)/B;
//   dz = (dlogits W2^T) * relu'/dropout mask.   z is cleared afterwards (split-K accumulator).
// ------------------------------------------------------------------------------------------------
template <int NC_MAX>
__global__ void __launch_bounds__(256) head_fwd_bwd_kernel(float* __restrict__ z, const bf16* __restrict__ b1,
                                                           const bf16* __restrict__ w2, const bf16* __restrict__ b2,
                                                           const int* __restrict__ labels, bf16* __restrict__ h_out,
                                                           bf16* __restrict__ dz_out, float* __restrict__ dlogits_out,
                                                           float* __restrict__ logits_out, float* __restrict__ loss_sum,
                                                           int* __restrict__ correct, int B, int D, int NC, float drop_rate,
                                                           int training, uint64_t seed, const int* step_ptr, int layer_id,
                                                           int zero_z, int ring, int stage_w2) {
  extern __shared__ float hs[];            // D floats: bf16-rounded h, then (optionally) a copy of W2
  __shared__ float part[8][NC_MAX];
  __shared__ float dl[NC_MAX];
  const int b = blockIdx.x;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5, nwarp = blockDim.x >> 5;
  // W2 ([D,NC] bf16, 20 KB for the MNIST head) is read twice per row with a stride the coalescer dislikes:
  // stage it once per block with 16-byte loads and serve both passes from shared memory.  The weights were written by
  // the previous step's update, long complete (every kernel of the chain passes its own pdl_wait before it lets its
  // successor start), so the staging runs BEFORE pdl_wait, underneath the tail of the dense GEMM that produces z.
  const bf16* w2p = w2;
  if (stage_w2) {
    bf16* w2s = reinterpret_cast<bf16*>(hs + D);
    const int n16 = D * NC * 2 / 16;
    const uint4* src = reinterpret_cast<const uint4*>(w2);
    uint4* dst = reinterpret_cast<uint4*>(w2s);
    for (int i = threadIdx.x; i < n16; i += blockDim.x) dst[i] = __ldg(src + i);
    w2p = w2s;
    __syncthreads();
  }
  pdl_wait();
  pdl_launch_dependents();
  const bool drop = training && drop_rate > 0.f;
  const uint32_t step = step_ptr ? (uint32_t)*step_ptr : 0u;
  const uint32_t thresh = (uint32_t)fminf(drop_rate * 4294967296.f, 4294967295.f);
  const float keep_scale = drop ? 1.f / (1.f - drop_rate) : 1.f;
  float acc[NC_MAX];
#pragma unroll
  for (int k = 0; k < NC_MAX; ++k) acc[k] = 0.f;
  float* zr = z + (size_t)b * D;
  for (int j4 = tid * 4; j4 < D; j4 += blockDim.x * 4) {
    float4 zv = *reinterpret_cast<float4*>(zr + j4);
    if (zero_z) *reinterpret_cast<float4*>(zr + j4) = make_float4(0.f, 0.f, 0.f, 0.f);
    uint4 rnd = make_uint4(~0u, ~0u, ~0u, ~0u);
    if (drop) rnd = dropout_rand4((uint32_t)(((size_t)b * D + j4) >> 2), step, (uint32_t)layer_id, seed);
    const float zz[4] = {zv.x, zv.y, zv.z, zv.w};
    const uint32_t rr[4] = {rnd.x, rnd.y, rnd.z, rnd.w};
#pragma unroll
    for (int e = 0; e < 4; ++e) {
      const int j = j4 + e;
      float hv = fmaxf(zz[e] + bf(b1 + j), 0.f);
      hv = (!drop || rr[e] >= thresh) ? hv * keep_scale : 0.f;
      const bf16 hb = __float2bfloat16(hv);
      h_out[(size_t)b * D + j] = hb;
      hv = __bfloat162float(hb);
      hs[j] = hv;
      if (hv != 0.f) {
#pragma unroll
        for (int k = 0; k < NC_MAX; ++k)
          if (k < NC) acc[k] = fmaf(hv, bf(w2p + (size_t)j * NC + k), acc[k]);
      }
    }
  }
#pragma unroll
  for (int k = 0; k < NC_MAX; ++k) {
    const float s = warp_sum(acc[k]);
    if (lane == 0) part[warp][k] = s;
  }
  __syncthreads();
  if (tid == 0) {
    float lg[NC_MAX], mx = -1e30f;
    int am = 0;
    for (int k = 0; k < NC; ++k) {
      float s = bf(b2 + k);
      for (int wv = 0; wv < nwarp; ++wv) s += part[wv][k];
      lg[k] = s;
      if (logits_out) logits_out[(size_t)b * NC + k] = s;
      if (s > mx) { mx = s; am = k; }
    }
    float se = 0.f;
    for (int k = 0; k < NC; ++k) se += expf(lg[k] - mx);
    const int y = labels[b];
    const float lse = mx + logf(se);
    const float invB = 1.f / (float)B;
    const int slot = ring > 0 ? (int)(step % (uint32_t)ring) : 0;   // ring > 0: per-step mean loss, else running sum
    atomicAdd(loss_sum + slot, ring > 0 ? (lse - lg[y]) * invB : lse - lg[y]);
    if (am == y) atomicAdd(correct + slot, 1);
    for (int k = 0; k < NC; ++k) {
      const float d = (expf(lg[k] - lse) - (k == y ? 1.f : 0.f)) * invB;
      dl[k] = d;
      if (dlogits_out) dlogits_out[(size_t)b * NC + k] = d;
    }
  }
  if (!training || dz_out == nullptr) return;
  __syncthreads();
  float dlr[NC_MAX];
#pragma unroll
  for (int k = 0; k < NC_MAX; ++k) dlr[k] = k < NC ? dl[k] : 0.f;
  for (int j = tid; j < D; j += blockDim.x) {
    float g = 0.f;
    if (hs[j] > 0.f) {
#pragma unroll
      for (int k = 0; k < NC_MAX; ++k)
        if (k < NC) g = fmaf(dlr[k], bf(w2p + (size_t)j * NC + k), g);
      g *= keep_scale;
    }
    dz_out[(size_t)b * D + j] = __float2bfloat16(g);
  }
}

// dW2[j,k] = sum_b h[b,j] dlogits[b,k];  db1[j] = sum_b dz[b,j];  db2[k] = sum_b dlogits[b,k]
template <int NC_MAX>
__global__ void __launch_bounds__(256) head_wgrad_kernel(const bf16* __restrict__ h, const bf16* __restrict__ dz,
                                                         const float* __restrict__ dlogits, float* __restrict__ dw2,
                                                         float* __restrict__ db2, float* __restrict__ db1, int B, int D, int NC) {
  __shared__ float red[32][8][NC_MAX + 1];
  extern __shared__ float dls[];                                // B x NC dlogits, staged once per block
  pdl_wait();
  pdl_launch_dependents();
  for (int i = threadIdx.x; i < B * NC; i += blockDim.x) dls[i] = dlogits[i];
  __syncthreads();
  const int jj = threadIdx.x & 7, bl = threadIdx.x >> 3;       // 8 columns x 32 batch lanes
  const int j = blockIdx.x * 8 + jj;
  float acc[NC_MAX + 1];
#pragma unroll
  for (int k = 0; k <= NC_MAX; ++k) acc[k] = 0.f;
  if (j < D) {
#pragma unroll 4
    for (int b = bl; b < B; b += 32) {
      const float hv = bf(h + (size_t)b * D + j);
      acc[NC_MAX] += bf(dz + (size_t)b * D + j);
#pragma unroll
      for (int k = 0; k < NC_MAX; ++k)
        if (k < NC) acc[k] = fmaf(hv, dls[b * NC + k], acc[k]);
    }
  }
#pragma unroll
  for (int k = 0; k <= NC_MAX; ++k) red[bl][jj][k] = acc[k];
  __syncthreads();
  for (int i = threadIdx.x; i < 8 * (NC_MAX + 1); i += blockDim.x) {
    const int c = i / (NC_MAX + 1), k = i % (NC_MAX + 1);
    const int col = blockIdx.x * 8 + c;
    if (col >= D) continue;
    float s = 0.f;
    for (int r = 0; r < 32; ++r) s += red[r][c][k];
    if (k == NC_MAX) db1[col] = s;
    else if (k < NC) dw2[(size_t)col * NC + k] = s;
  }
  if (blockIdx.x == 0 && (int)threadIdx.x < NC) {
    float s = 0.f;
    for (int b = 0; b < B; ++b) s += dls[b * NC + threadIdx.x];
    db2[threadIdx.x] = s;
  }
}

// ------------------------------------------------------------------------------------------------
// grid-stride with a stride that is a multiple of C: every thread stays on one channel, so the conv bias
// gradient (sum of dy over batch and space) accumulates in a register and costs one atomic per thread group.
__global__ void __launch_bounds__(256) unpool_relu_bwd_kernel(const bf16* __restrict__ dpool, const bf16* __restrict__ pooled,
                                                              const uint8_t* __restrict__ argmax, bf16* __restrict__ dy,
                                                              float* __restrict__ dbias, int HP, int WP, int C, long long total,
                                                              float gscale) {
  __shared__ float red[256];
  pdl_wait();
  pdl_launch_dependents();
  float bsum = 0.f;
  const long long stride = (long long)gridDim.x * blockDim.x;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += stride) {
    const int c = (int)(i % C);
    long long p = i / C;
    const int pw = (int)(p % WP); p /= WP;
    const int ph = (int)(p % HP);
    const long long b = p / HP;
    const float g = bf(pooled + i) > 0.f ? bf(dpool + i) * gscale : 0.f;
    bsum += g;
    const int q = argmax[i];
    const int W = 2 * WP;
    const size_t base = (((size_t)b * 2 * HP + 2 * ph) * W + 2 * pw) * C + c;
    const bf16 zero = __float2bfloat16(0.f), gv = __float2bfloat16(g);
    dy[base] = q == 0 ? gv : zero;
    dy[base + C] = q == 1 ? gv : zero;
    dy[base + (size_t)W * C] = q == 2 ? gv : zero;
    dy[base + (size_t)W * C + C] = q == 3 ? gv : zero;
  }
  if (dbias == nullptr) return;
  // blockDim (256) and the stride are multiples of C when C divides 256; otherwise fall back to per-thread atomics
  if (256 % C == 0) {
    red[threadIdx.x] = bsum;
    __syncthreads();
    if ((int)threadIdx.x < C) {
      float s = 0.f;
      for (int t = threadIdx.x; t < 256; t += C) s += red[t];
      atomicAdd(dbias + threadIdx.x, s);
    }
  } else if (stride % C == 0) {
    const long long first = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (first < total) atomicAdd(dbias + (int)(first % C), bsum);
  }
}

// Eight channels per thread (16-byte loads of dpool/pooled, 8-byte argmax, four 16-byte stores): the scalar kernel above
// moves 2 bytes per memory instruction and is latency-bound.  Needs C % 8 == 0 and (for dbias) 256 % (C/8) == 0, which
// keeps every thread on one channel group for the whole grid-stride loop.
__global__ void __launch_bounds__(256) unpool_relu_bwd_v8_kernel(const bf16* __restrict__ dpool, const bf16* __restrict__ pooled,
                                                                 const uint8_t* __restrict__ argmax, bf16* __restrict__ dy,
                                                                 float* __restrict__ dbias, int HP, int WP, int C,
                                                                 long long ngroups, float gscale) {
  __shared__ float red[256][9];
  pdl_wait();
  pdl_launch_dependents();
  const int CG = C >> 3;
  float bs[8];
#pragma unroll
  for (int e = 0; e < 8; ++e) bs[e] = 0.f;
  const long long stride = (long long)gridDim.x * blockDim.x;
  const int W = 2 * WP;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < ngroups; i += stride) {
    const int cg = (int)(i % CG);
    long long p = i / CG;
    const int pw = (int)(p % WP); p /= WP;
    const int ph = (int)(p % HP);
    const long long b = p / HP;
    const uint4 dp = __ldg(reinterpret_cast<const uint4*>(dpool) + i);
    const uint4 po = __ldg(reinterpret_cast<const uint4*>(pooled) + i);
    const uint2 am = __ldg(reinterpret_cast<const uint2*>(argmax) + i);
    const uint32_t dpw[4] = {dp.x, dp.y, dp.z, dp.w}, pow_[4] = {po.x, po.y, po.z, po.w};
    uint32_t gb[8];           // gradient as bf16 bits, per channel
    uint32_t qq[8];
#pragma unroll
    for (int e = 0; e < 8; ++e) {
      const uint32_t dbits = (dpw[e >> 1] >> (16 * (e & 1))) & 0xFFFFu, pbits = (pow_[e >> 1] >> (16 * (e & 1))) & 0xFFFFu;
      const float pv = __uint_as_float(pbits << 16), dv = __uint_as_float(dbits << 16);
      const float g = pv > 0.f ? dv * gscale : 0.f;
      bs[e] += g;
      gb[e] = (uint32_t)__bfloat16_as_ushort(__float2bfloat16(g));
      qq[e] = ((e < 4 ? am.x : am.y) >> (8 * (e & 3))) & 0xFFu;
    }
    bf16* base = dy + ((((size_t)b * 2 * HP + 2 * ph) * W + 2 * pw) * C + cg * 8);
#pragma unroll
    for (int qd = 0; qd < 4; ++qd) {
      uint32_t o[4];
#pragma unroll
      for (int j = 0; j < 4; ++j)
        o[j] = (qq[2 * j] == (uint32_t)qd ? gb[2 * j] : 0u) | ((qq[2 * j + 1] == (uint32_t)qd ? gb[2 * j + 1] : 0u) << 16);
      *reinterpret_cast<uint4*>(base + (size_t)(qd >> 1) * W * C + (qd & 1) * C) = make_uint4(o[0], o[1], o[2], o[3]);
    }
  }
  if (dbias == nullptr) return;
#pragma unroll
  for (int e = 0; e < 8; ++e) red[threadIdx.x][e] = bs[e];
  __syncthreads();
  if ((int)threadIdx.x < C) {
    const int cg = threadIdx.x >> 3, e = threadIdx.x & 7;
    float sum = 0.f;
    for (int t = cg; t < 256; t += CG) sum += red[t][e];
    atomicAdd(dbias + threadIdx.x, sum);
  }
}

__global__ void maxpool_fwd_kernel(const bf16* __restrict__ in, bf16* __restrict__ out, uint8_t* __restrict__ argmax,
                                   int H, int W, int C, long long total) {
  const int HP = H / 2, WP = W / 2;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
    const int c = (int)(i % C);
    long long p = i / C;
    const int pw = (int)(p % WP); p /= WP;
    const int ph = (int)(p % HP);
    const long long b = p / HP;
    const size_t base = (((size_t)b * H + 2 * ph) * W + 2 * pw) * C + c;
    const float v0 = bf(in + base), v1 = bf(in + base + C), v2 = bf(in + base + (size_t)W * C), v3 = bf(in + base + (size_t)W * C + C);
    float best = v0; int q = 0;
    if (v1 > best) { best = v1; q = 1; }
    if (v2 > best) { best = v2; q = 2; }
    if (v3 > best) { best = v3; q = 3; }
    out[i] = __float2bfloat16(best);
    if (argmax) argmax[i] = (uint8_t)q;
  }
}

// ------------------------------------------------------------------------------------------------
// Direct convolutions (fallback / oracle).  Output spatial size Ho = H + 2*pad - k + 1.
// ------------------------------------------------------------------------------------------------
__global__ void conv_direct_fwd_kernel(const bf16* __restrict__ in, const bf16* __restrict__ w, const bf16* __restrict__ bias,
                                       bf16* __restrict__ out, int H, int W, int Ci, int Co, int k, int pad, int relu,
                                       long long total) {
  const int Ho = H + 2 * pad - k + 1, Wo = W + 2 * pad - k + 1;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
    const int co = (int)(i % Co);
    long long p = i / Co;
    const int wo = (int)(p % Wo); p /= Wo;
    const int ho = (int)(p % Ho);
    const long long b = p / Ho;
    float acc = bias ? bf(bias + co) : 0.f;
    for (int kh = 0; kh < k; ++kh) {
      const int hi = ho + kh - pad;
      if (hi < 0 || hi >= H) continue;
      for (int kw = 0; kw < k; ++kw) {
        const int wi = wo + kw - pad;
        if (wi < 0 || wi >= W) continue;
        const bf16* ip = in + (((size_t)b * H + hi) * W + wi) * Ci;
        const bf16* wp = w + ((size_t)(kh * k + kw) * Ci) * Co + co;
        for (int ci = 0; ci < Ci; ++ci) acc = fmaf(bf(ip + ci), bf(wp + (size_t)ci * Co), acc);
      }
    }
    if (relu) acc = fmaxf(acc, 0.f);
    out[i] = __float2bfloat16(acc);
  }
}

__global__ void conv_direct_dgrad_kernel(const bf16* __restrict__ dout, const bf16* __restrict__ w, bf16* __restrict__ din,
                                         int H, int W, int Ci, int Co, int k, int pad, long long total) {
  const int Ho = H + 2 * pad - k + 1, Wo = W + 2 * pad - k + 1;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
    const int ci = (int)(i % Ci);
    long long p = i / Ci;
    const int wi = (int)(p % W); p /= W;
    const int hi = (int)(p % H);
    const long long b = p / H;
    float acc = 0.f;
    for (int kh = 0; kh < k; ++kh) {
      const int ho = hi - kh + pad;
      if (ho < 0 || ho >= Ho) continue;
      for (int kw = 0; kw < k; ++kw) {
        const int wo = wi - kw + pad;
        if (wo < 0 || wo >= Wo) continue;
        const bf16* dp = dout + (((size_t)b * Ho + ho) * Wo + wo) * Co;
        const bf16* wp = w + ((size_t)(kh * k + kw) * Ci + ci) * Co;
        for (int co = 0; co < Co; ++co) acc = fmaf(bf(dp + co), bf(wp + co), acc);
      }
    }
    din[i] = __float2bfloat16(acc);
  }
}

// one thread per (tap, ci, co); blockIdx.y splits the batch; fp32 atomics into dw / dbias
__global__ void conv_direct_wgrad_kernel(const bf16* __restrict__ in, const bf16* __restrict__ dout, float* __restrict__ dw,
                                         float* __restrict__ dbias, int B, int H, int W, int Ci, int Co, int k, int pad) {
  const int Ho = H + 2 * pad - k + 1, Wo = W + 2 * pad - k + 1;
  const long long total = (long long)k * k * Ci * Co;
  const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= total) return;
  const int co = (int)(i % Co);
  long long p = i / Co;
  const int ci = (int)(p % Ci); p /= Ci;
  const int kw = (int)(p % k), kh = (int)(p / k);
  const int bchunk = (B + gridDim.y - 1) / gridDim.y;
  const int b0 = blockIdx.y * bchunk, b1 = min(B, b0 + bchunk);
  float acc = 0.f, bacc = 0.f;
  const bool do_bias = dbias != nullptr && kh == 0 && kw == 0 && ci == 0;
  for (int b = b0; b < b1; ++b)
    for (int ho = 0; ho < Ho; ++ho) {
      const int hi = ho + kh - pad;
      const bool hin = hi >= 0 && hi < H;
      if (!hin && !do_bias) continue;
      for (int wo = 0; wo < Wo; ++wo) {
        const float d = bf(dout + (((size_t)b * Ho + ho) * Wo + wo) * Co + co);
        if (do_bias) bacc += d;
        const int wi = wo + kw - pad;
        if (hin && wi >= 0 && wi < W) acc = fmaf(bf(in + (((size_t)b * H + hi) * W + wi) * Ci + ci), d, acc);
      }
    }
  atomicAdd(dw + i, acc);
  if (do_bias) atomicAdd(dbias + co, bacc);
}

// ------------------------------------------------------------------------------------------------
// Tiled CUDA-core GEMM, 64x64 tile, 16-deep k slices, 4x4 register micro-tile.
// ------------------------------------------------------------------------------------------------
template <bool TA, bool TB>
__global__ void __launch_bounds__(256) gemm_simple_kernel(const bf16* __restrict__ A, const bf16* __restrict__ Bm, void* __restrict__ C,
                                                          int M, int N, int K, int accumulate, int out_bf16) {
  __shared__ float As[16][65], Bs[16][65];
  const int tx = threadIdx.x & 15, ty = threadIdx.x >> 4;
  const int m0 = blockIdx.y * 64, n0 = blockIdx.x * 64;
  float acc[4][4] = {};
  for (int k0 = 0; k0 < K; k0 += 16) {
    for (int i = threadIdx.x; i < 64 * 16; i += 256) {
      int mm, kk;
      if (TA) { mm = i & 63; kk = i >> 6; } else { kk = i & 15; mm = i >> 4; }
      const int gm = m0 + mm, gk = k0 + kk;
      float v = 0.f;
      if (gm < M && gk < K) v = bf(TA ? A + (size_t)gk * M + gm : A + (size_t)gm * K + gk);
      As[kk][mm] = v;
    }
    for (int i = threadIdx.x; i < 64 * 16; i += 256) {
      int nn, kk;
      if (TB) { kk = i & 15; nn = i >> 4; } else { nn = i & 63; kk = i >> 6; }
      const int gn = n0 + nn, gk = k0 + kk;
      float v = 0.f;
      if (gn < N && gk < K) v = bf(TB ? Bm + (size_t)gn * K + gk : Bm + (size_t)gk * N + gn);
      Bs[kk][nn] = v;
    }
    __syncthreads();
#pragma unroll
    for (int kk = 0; kk < 16; ++kk) {
      float a[4], b[4];
#pragma unroll
      for (int i = 0; i < 4; ++i) { a[i] = As[kk][ty * 4 + i]; b[i] = Bs[kk][tx * 4 + i]; }
#pragma unroll
      for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j) acc[i][j] = fmaf(a[i], b[j], acc[i][j]);
    }
    __syncthreads();
  }
  for (int i = 0; i < 4; ++i)
    for (int j = 0; j < 4; ++j) {
      const int gm = m0 + ty * 4 + i, gn = n0 + tx * 4 + j;
      if (gm >= M || gn >= N) continue;
      const size_t o = (size_t)gm * N + gn;
      if (out_bf16) {
        bf16* c = (bf16*)C;
        c[o] = __float2bfloat16(acc[i][j] + (accumulate ? __bfloat162float(c[o]) : 0.f));
      } else {
        float* c = (float*)C;
        c[o] = acc[i][j] + (accumulate ? c[o] : 0.f);
      }
    }
}

__global__ void cast_f32_bf16_kernel(const float* __restrict__ s, bf16* __restrict__ d, long long n) {
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x)
    d[i] = __float2bfloat16(s[i]);
}

__global__ void bias_relu_bf16_kernel(const float* __restrict__ z, const bf16* __restrict__ bias, bf16* __restrict__ out,
                                      int cols, int relu, long long total) {
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
    float v = z[i] + (bias ? bf(bias + (i % cols)) : 0.f);
    if (relu) v = fmaxf(v, 0.f);
    out[i] = __float2bfloat16(v);
  }
}

__global__ void adam_local_kernel(float* __restrict__ p, float* __restrict__ m, float* __restrict__ v, const float* __restrict__ g,
                                  bf16* __restrict__ w16, long long n, float b1, float b2, float eps, float step_size,
                                  float vscale, float gscale) {
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x) {
    const float gg = g[i] * gscale;
    const float mm = b1 * m[i] + (1.f - b1) * gg;
    const float vv = b2 * v[i] + (1.f - b2) * gg * gg;
    const float pp = p[i] - step_size * mm / (sqrtf(vv) * vscale + eps);
    m[i] = mm; v[i] = vv; p[i] = pp;
    if (w16) w16[i] = __float2bfloat16(pp);
  }
}

__global__ void l2_flush_kernel(float4* p, size_t n) {
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x)
    p[i] = make_float4(1.f, 2.f, 3.f, (float)i);
}

inline int grid_for(long long total, int threads, int cap = 148 * 16) {
  long long g = (total + threads - 1) / threads;
  if (g < 1) g = 1;
  return (int)(g > cap ? cap : g);
}

}  // namespace

extern "C" {

int b200_conv1_pool_fwd(const float* x, const void* w, const void* bias, void* out, uint8_t* argmax, int B, int C1,
                        int* step_ptr, int bump_step, float* loss_ring, int* correct_ring, int ring, void* stream) {
  B200_REQUIRE(B > 0 && C1 > 0, "conv1: bad shape B=%d C1=%d", B, C1);
  B200_CHECK(launch_pdl(conv1_pool_fwd_kernel, dim3(B, 4), dim3(256), 0, (cudaStream_t)stream, x, (const bf16*)w, (const bf16*)bias,
                        (bf16*)out, argmax, C1, step_ptr, bump_step, loss_ring, correct_ring, ring));
  return B200_OK;
}

int b200_conv1_wgrad(const float* x, const void* da1, const void* a1, const uint8_t* argmax, float* dw, float* dbias,
                     int B, int C1, void* stream) {
  B200_REQUIRE(B > 0 && C1 > 0, "conv1 wgrad: bad shape");
  const int grid = B < 74 ? B : 74;          // several images per block: one atomic flush per block, not per image
  B200_CHECK(launch_pdl(conv1_wgrad_kernel, dim3(grid, 4), dim3(256), 0, (cudaStream_t)stream, x, (const bf16*)da1, (const bf16*)a1,
                        (const uint8_t*)argmax, dw, dbias, B, C1));
  return B200_OK;
}

int b200_head_fwd_bwd(float* z, const void* b1, const void* w2, const void* b2, const int* labels, void* h, void* dz,
                      float* dlogits, float* logits_out, float* loss_sum, int* correct, int B, int D, int NC,
                      float drop_rate, int training, uint64_t seed, const int* step_ptr, int layer_id, int zero_z,
                      int ring, void* stream) {
  B200_REQUIRE(NC >= 1 && NC <= 16, "head: NC=%d outside [1,16]", NC);
  B200_REQUIRE(D % 4 == 0 && D > 0 && (size_t)D * 4 <= 200 * 1024, "head: D=%d must be a multiple of 4 and fit shared memory", D);
  B200_REQUIRE(drop_rate >= 0.f && drop_rate < 1.f, "head: drop_rate %f", drop_rate);
  const size_t w2_bytes = (size_t)D * NC * 2;
  const int stage_w2 = (w2_bytes % 16 == 0 && ((uintptr_t)w2 & 15) == 0 && w2_bytes <= 64 * 1024) ? 1 : 0;
  const size_t smem = (size_t)D * sizeof(float) + (stage_w2 ? w2_bytes : 0);
  static bool attr_set = false;
  if (!attr_set) {
    B200_CHECK(cudaFuncSetAttribute(head_fwd_bwd_kernel<16>, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
    attr_set = true;
  }
  B200_CHECK(launch_pdl(head_fwd_bwd_kernel<16>, dim3(B), dim3(256), smem, (cudaStream_t)stream, z, (const bf16*)b1, (const bf16*)w2,
                        (const bf16*)b2, labels, (bf16*)h, (bf16*)dz, dlogits, logits_out, loss_sum, correct, B, D, NC, drop_rate,
                        training, seed, step_ptr, layer_id, zero_z, ring, stage_w2));
  return B200_OK;
}

int b200_head_wgrad(const void* h, const void* dz, const float* dlogits, float* dw2, float* db2, float* db1, int B, int D,
                    int NC, void* stream) {
  B200_REQUIRE(NC >= 1 && NC <= 16, "head wgrad: NC=%d outside [1,16]", NC);
  B200_REQUIRE((size_t)B * NC * 4 <= 160 * 1024, "head wgrad: batch %d too large for the staged dlogits", B);
  static bool wg_attr = false;
  if (!wg_attr) {
    B200_CHECK(cudaFuncSetAttribute(head_wgrad_kernel<16>, cudaFuncAttributeMaxDynamicSharedMemorySize, 160 * 1024));
    wg_attr = true;
  }
  B200_CHECK(launch_pdl(head_wgrad_kernel<16>, dim3(cdiv(D, 8)), dim3(256), (size_t)B * NC * sizeof(float), (cudaStream_t)stream,
                        (const bf16*)h, (const bf16*)dz, dlogits, dw2, db2, db1, B, D, NC));
  return B200_OK;
}

// `pooled` may be the tensor AFTER an inverted dropout of keep-scale `gscale` (keras128): a dropped or ReLU-clamped element
// is 0 there and passes no gradient, a kept one passes dpool * gscale.
int b200_unpool_relu_bwd_scaled(const void* dpool, const void* pooled, const uint8_t* argmax, void* dy, float* dbias, int B,
                                int HP, int WP, int C, float gscale, void* stream) {
  const long long total = (long long)B * HP * WP * C;
  const bool aligned = (((uintptr_t)dpool | (uintptr_t)pooled | (uintptr_t)dy) & 15) == 0 && ((uintptr_t)argmax & 7) == 0;
  if (C % 8 == 0 && C <= 256 && 256 % (C / 8) == 0 && aligned) {
    const long long ngroups = total / 8;
    B200_CHECK(launch_pdl(unpool_relu_bwd_v8_kernel, dim3(grid_for(ngroups, 256, 148 * 8)), dim3(256), 0, (cudaStream_t)stream,
                          (const bf16*)dpool, (const bf16*)pooled, argmax, (bf16*)dy, dbias, HP, WP, C, ngroups, gscale));
    return B200_OK;
  }
  int grid = grid_for(total, 256, 148 * 4);
  if (dbias && 256 % C != 0) {
    // keep the stride a multiple of C so a thread never changes channel
    while (grid > 1 && (256LL * grid) % C != 0) --grid;
    B200_REQUIRE((256LL * grid) % C == 0, "unpool: cannot build a channel-stable stride for C=%d", C);
  }
  B200_CHECK(launch_pdl(unpool_relu_bwd_kernel, dim3(grid), dim3(256), 0, (cudaStream_t)stream, (const bf16*)dpool, (const bf16*)pooled,
                        argmax, (bf16*)dy, dbias, HP, WP, C, total, gscale));
  return B200_OK;
}

int b200_unpool_relu_bwd(const void* dpool, const void* pooled, const uint8_t* argmax, void* dy, float* dbias, int B, int HP,
                         int WP, int C, void* stream) {
  return b200_unpool_relu_bwd_scaled(dpool, pooled, argmax, dy, dbias, B, HP, WP, C, 1.f, stream);
}

int b200_conv1_valid_fwd(const float* x, const void* w, const void* bias, void* out, int B, int C1, int k, int* step_ptr,
                         int bump_step, float* loss_ring, int* correct_ring, int ring, void* stream) {
  B200_REQUIRE(B > 0 && C1 > 0, "conv1_valid: bad shape B=%d C1=%d", B, C1);
  B200_REQUIRE(k == 3 || k == 5, "conv1_valid: kernel size %d not instantiated (3 or 5)", k);
  const bool v8 = k == 3 && C1 % 8 == 0 && (((uintptr_t)w | (uintptr_t)bias | (uintptr_t)out) & 15) == 0;
  auto kern = v8 ? conv1_valid3_fwd_v8_kernel : (k == 3 ? conv1_valid_fwd_kernel<3> : conv1_valid_fwd_kernel<5>);
  B200_CHECK(launch_pdl(kern, dim3(B, 4), dim3(256), 0, (cudaStream_t)stream, x, (const bf16*)w, (const bf16*)bias, (bf16*)out, C1,
                        step_ptr, bump_step, loss_ring, correct_ring, ring));
  return B200_OK;
}

int b200_conv1_valid_wgrad(const float* x, const void* da1, const void* a1, float* dw, float* dbias, int B, int C1, int k,
                           void* stream) {
  B200_REQUIRE(B > 0 && C1 > 0, "conv1_valid wgrad: bad shape");
  B200_REQUIRE(k == 3 || k == 5, "conv1_valid wgrad: kernel size %d not instantiated (3 or 5)", k);
  const bool v8 = k == 3 && C1 % 8 == 0 && (((uintptr_t)da1 | (uintptr_t)a1) & 15) == 0;
  auto kern = v8 ? conv1_valid3_wgrad_v8_kernel : (k == 3 ? conv1_valid_wgrad_kernel<3> : conv1_valid_wgrad_kernel<5>);
  const int grid = B < 74 ? B : 74;
  B200_CHECK(launch_pdl(kern, dim3(grid, 4), dim3(256), 0, (cudaStream_t)stream, x, (const bf16*)da1, (const bf16*)a1, dw, dbias, B, C1));
  return B200_OK;
}

int b200_dropout_bf16(const void* in, void* out, long long n, float rate, uint64_t seed, const int* step_ptr, int layer_id,
                      void* stream) {
  B200_REQUIRE(n > 0 && n % 4 == 0, "dropout: element count %lld must be a positive multiple of 4", n);
  B200_REQUIRE(rate >= 0.f && rate < 1.f, "dropout: rate %f outside [0,1)", rate);
  B200_REQUIRE((n >> 2) <= 0xFFFFFFFFLL, "dropout: tensor too large for a 32-bit Philox counter");
  B200_CHECK(launch_pdl(dropout_bf16_kernel, dim3(grid_for(n / 4, 256, 148 * 8)), dim3(256), 0, (cudaStream_t)stream, (const bf16*)in,
                        (bf16*)out, n / 4, rate, seed, step_ptr, layer_id));
  return B200_OK;
}

int b200_maxpool_fwd(const void* in, void* out, uint8_t* argmax, int B, int H, int W, int C, void* stream) {
  B200_REQUIRE(H % 2 == 0 && W % 2 == 0, "maxpool: odd spatial size %dx%d", H, W);
  const long long total = (long long)B * (H / 2) * (W / 2) * C;
  maxpool_fwd_kernel<<<grid_for(total, 256), 256, 0, (cudaStream_t)stream>>>((const bf16*)in, (bf16*)out, argmax, H, W, C, total);
  B200_LAUNCH_CHECK();
  return B200_OK;
}

int b200_conv_direct_fwd(const void* in, const void* w, const void* bias, void* out, int B, int H, int W, int Ci, int Co,
                         int k, int pad, int relu, void* stream) {
  const int Ho = H + 2 * pad - k + 1, Wo = W + 2 * pad - k + 1;
  B200_REQUIRE(Ho > 0 && Wo > 0, "conv: empty output");
  const long long total = (long long)B * Ho * Wo * Co;
  conv_direct_fwd_kernel<<<grid_for(total, 256, 1 << 20), 256, 0, (cudaStream_t)stream>>>(
      (const bf16*)in, (const bf16*)w, (const bf16*)bias, (bf16*)out, H, W, Ci, Co, k, pad, relu, total);
  B200_LAUNCH_CHECK();
  return B200_OK;
}

int b200_conv_direct_dgrad(const void* dout, const void* w, void* din, int B, int H, int W, int Ci, int Co, int k, int pad,
                           void* stream) {
  const long long total = (long long)B * H * W * Ci;
  conv_direct_dgrad_kernel<<<grid_for(total, 256, 1 << 20), 256, 0, (cudaStream_t)stream>>>((const bf16*)dout, (const bf16*)w,
                                                                                           (bf16*)din, H, W, Ci, Co, k, pad, total);
  B200_LAUNCH_CHECK();
  return B200_OK;
}

int b200_conv_direct_wgrad(const void* in, const void* dout, float* dw, float* dbias, int B, int H, int W, int Ci, int Co,
                           int k, int pad, void* stream) {
  const long long total = (long long)k * k * Ci * Co;
  const int gx = cdiv(total, 128);
  int gy = (148 * 8 + gx - 1) / gx;
  if (gy > B) gy = B;
  if (gy < 1) gy = 1;
  conv_direct_wgrad_kernel<<<dim3(gx, gy), 128, 0, (cudaStream_t)stream>>>((const bf16*)in, (const bf16*)dout, dw, dbias, B, H, W,
                                                                          Ci, Co, k, pad);
  B200_LAUNCH_CHECK();
  return B200_OK;
}

int b200_gemm_simple(const void* A, const void* Bm, void* C, int M, int N, int K, int transA, int transB, int accumulate,
                     int out_bf16, void* stream) {
  B200_REQUIRE(M > 0 && N > 0 && K > 0, "gemm: bad shape %dx%dx%d", M, N, K);
  dim3 grid(cdiv(N, 64), cdiv(M, 64));
  cudaStream_t st = (cudaStream_t)stream;
  if (!transA && !transB) gemm_simple_kernel<false, false><<<grid, 256, 0, st>>>((const bf16*)A, (const bf16*)Bm, C, M, N, K, accumulate, out_bf16);
  if (transA && !transB) gemm_simple_kernel<true, false><<<grid, 256, 0, st>>>((const bf16*)A, (const bf16*)Bm, C, M, N, K, accumulate, out_bf16);
  if (!transA && transB) gemm_simple_kernel<false, true><<<grid, 256, 0, st>>>((const bf16*)A, (const bf16*)Bm, C, M, N, K, accumulate, out_bf16);
  if (transA && transB) gemm_simple_kernel<true, true><<<grid, 256, 0, st>>>((const bf16*)A, (const bf16*)Bm, C, M, N, K, accumulate, out_bf16);
  B200_LAUNCH_CHECK();
  return B200_OK;
}

int b200_cast_f32_to_bf16(const float* src, void* dst, long long n, void* stream) {
  cast_f32_bf16_kernel<<<grid_for(n, 256), 256, 0, (cudaStream_t)stream>>>(src, (bf16*)dst, n);
  B200_LAUNCH_CHECK();
  return B200_OK;
}

int b200_bias_relu_bf16(const float* z, const void* bias, void* out, int rows, int cols, int relu, void* stream) {
  const long long total = (long long)rows * cols;
  bias_relu_bf16_kernel<<<grid_for(total, 256), 256, 0, (cudaStream_t)stream>>>(z, (const bf16*)bias, (bf16*)out, cols, relu, total);
  B200_LAUNCH_CHECK();
  return B200_OK;
}

int b200_adam_local(float* master, float* m, float* v, const float* grad, void* w16, long long n, int step, double lr,
                    double b1, double b2, double eps, int eps_mode, float grad_scale, void* stream) {
  B200_REQUIRE(step >= 1, "adam: step must be >= 1");
  const double bc1 = 1.0 - pow(b1, (double)step), bc2 = 1.0 - pow(b2, (double)step);
  const float step_size = (float)(eps_mode == 0 ? lr * sqrt(bc2) / bc1 : lr / bc1);
  const float vscale = (float)(eps_mode == 0 ? 1.0 : 1.0 / sqrt(bc2));
  adam_local_kernel<<<grid_for(n, 256), 256, 0, (cudaStream_t)stream>>>(master, m, v, grad, (bf16*)w16, n, (float)b1, (float)b2,
                                                                        (float)eps, step_size, vscale, grad_scale);
  B200_LAUNCH_CHECK();
  return B200_OK;
}

int b200_l2_flush(void* scratch, size_t bytes, void* stream) {
  l2_flush_kernel<<<148 * 4, 512, 0, (cudaStream_t)stream>>>((float4*)scratch, bytes / 16);
  B200_LAUNCH_CHECK();
  return B200_OK;
}

}  // extern "C"
