// Implicit-GEMM convolution on tcgen05 for sm_100a (NHWC bf16, stride 1, any padding 0..k-1: SAME and VALID).
//
// No im2col buffer exists anywhere: the filter taps are applied by the TMA unit — for tap (kh,kw) the
// producer loads the SAME 4-D box {channels, 16, 8, 1} of the activation tensor shifted by (kh-pad, kw-pad),
// and out-of-bounds rows/columns (the SAME halo, and the part of a 16-wide tile past the image edge) arrive as zeros.
// The source (TMA) tensor and the produced tensor may differ in size: a VALID forward reads HxW and writes
// (H-k+1)x(W-k+1); its dgrad reads the small tensor and writes the big one ("full" correlation through OOB zeros).
// An M tile is an 8x16 pixel patch of one image (row r = 16*hh + ww), so the 2x2 max-pool partners of a pixel
// sit in the same warp's TMEM lanes (lane^1, lane^16, lane^17) and pooling is three shuffles in the epilogue.
//
//   forward : D[pix, co] = sum_{tap,ci} in[pix + tap, ci] * W[tap, ci, co]      A K-major, B MN-major
//   dgrad   : D[pix, ci] = sum_{tap,co} dy[pix - tap, co] * W[tap, ci, co]      A K-major, B K-major
//   wgrad   : D[(tap,ci), co] = sum_pix in[pix + tap, ci] * dy[pix, co]         A MN-major (taps stacked in M), B MN-major
//
// Replaces cuDNN's conv2d fwd/dgrad/wgrad behind layers.conv2d (horovod/tensorflow_mnist.py:49-60;
// horovod/tensorflow_mnist_gpu.py:47-56) — SURVEY §2.5a K5, K6, K13.
#include <stdlib.h>

#include "../include/b200dp.h"
#include "tc_sm100.cuh"

namespace {

using namespace tc;

constexpr int TILE_H = 8, TILE_W = 16, TILE_PIX = 128;

struct ConvParams {
  int B, Ho, Wo, Ci, Co, k, pad;      // Ho x Wo: spatial size of the tensor this launch PRODUCES
  int mode;          // 0 forward, 1 dgrad
  int KC;            // reduction channels per k-block (32 or 64)
  int BN;            // output channels per CTA
  int kc_blocks;     // reduction-channel blocks per tap
  int relu, pool;
  const bf16* bias;
  bf16* out;
  uint8_t* argmax;
  int tiles_h, tiles_w;       // ceil(Ho / 8), ceil(Wo / 16)
  uint32_t tmem_cols;
  int nstage;
};

__global__ void __launch_bounds__(192, 1) conv_tc_kernel(const __grid_constant__ CUtensorMap tmA,
                                                         const __grid_constant__ CUtensorMap tmB, const ConvParams p) {
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  uint8_t* smem = (uint8_t*)(((uintptr_t)smem_raw + 1023) & ~(uintptr_t)1023);
  const int a_bytes = TILE_PIX * p.KC * 2;
  const int b_bytes = p.BN * p.KC * 2;
  const int stage_bytes = a_bytes + ((b_bytes + 1023) & ~1023);
  uint64_t* full = (uint64_t*)(smem + p.nstage * stage_bytes);
  uint64_t* empty = full + p.nstage;
  uint64_t* tmem_full = empty + p.nstage;
  uint32_t* tmem_slot = (uint32_t*)(tmem_full + 1);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int n0 = blockIdx.x * p.BN;
  const int tiles_hw = p.tiles_h * p.tiles_w;
  const int img = blockIdx.y / tiles_hw, tile = blockIdx.y % tiles_hw;
  const int h0 = (tile / p.tiles_w) * TILE_H, w0 = (tile % p.tiles_w) * TILE_W;
  const int nkb = p.k * p.k * p.kc_blocks;
  const uint32_t swz = p.KC == 64 ? SW_128B : SW_64B;
  const int row_bytes = p.KC * 2;

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tmA);
    tma_prefetch_desc(&tmB);
    for (int s = 0; s < p.nstage; ++s) { mbar_init(full + s, 1); mbar_init(empty + s, 1); }
    mbar_init(tmem_full, 1);
    fence_barrier_init();
  }
  if (warp == 1) tmem_alloc(tmem_slot, p.tmem_cols);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  pdl_wait();          // everything above overlapped the previous kernel; its outputs are visible from here on
  pdl_launch_dependents();

  if (warp == 0) {
    const bool leader = elect_one();
    int stage = 0;
    uint32_t phase = 0;
    int tap = 0, cb = 0, kh = 0, kw = 0;
    for (int kb = 0; kb < nkb; ++kb) {
      const int dh = p.mode == 0 ? kh - p.pad : p.pad - kh;
      const int dw = p.mode == 0 ? kw - p.pad : p.pad - kw;
      mbar_wait(empty + stage, phase ^ 1, 11);
      if (leader) {
        uint8_t* sA = smem + stage * stage_bytes;
        uint8_t* sB = sA + a_bytes;
        mbar_arrive_expect_tx(full + stage, (uint32_t)(a_bytes + b_bytes));
        tma_load_4d(sA, &tmA, full + stage, cb * p.KC, w0 + dw, h0 + dh, img);
        if (p.mode == 0) {
          // rows = (tap, ci) block, cols = co: MN-major B, one 64-column box per 64 output channels
          for (int j = 0; j < p.BN / 64; ++j)
            tma_load_2d(sB + j * (p.KC * 128), &tmB, full + stage, n0 + 64 * j, tap * p.Ci + cb * p.KC);
        } else {
          // rows = ci (this CTA's N), cols = co block: K-major B
          tma_load_2d(sB, &tmB, full + stage, cb * p.KC, tap * p.Ci + n0);
        }
      }
      __syncwarp();
      if (++cb == p.kc_blocks) { cb = 0; ++tap; if (++kw == p.k) { kw = 0; ++kh; } }
      if (++stage == p.nstage) { stage = 0; phase ^= 1; }
    }
  } else if (warp == 1) {
    const bool leader = elect_one();
    const uint32_t idesc = make_idesc_bf16(128, p.BN, 0, p.mode == 0 ? 1 : 0);
    const uint64_t a_hi = make_smem_desc(0, 16, 8 * row_bytes, swz);
    const uint64_t b_hi = p.mode == 0 ? make_smem_desc(0, p.KC * 128, 1024, SW_128B) : make_smem_desc(0, 16, 8 * row_bytes, swz);
    const uint32_t b_step = (p.mode == 0 ? 2048u : 32u) >> 4;
    const uint32_t s0 = (smem_u32(smem) & 0x3FFFFu) >> 4, stage_step = (uint32_t)stage_bytes >> 4;
    const int ksteps = p.KC / 16;
    int stage = 0;
    uint32_t phase = 0;
    for (int kb = 0; kb < nkb; ++kb) {
      mbar_wait(full + stage, phase, 12);
      tc_fence_after();
      if (leader) {
        const uint32_t a_lo = s0 + stage * stage_step, b_lo = a_lo + ((uint32_t)a_bytes >> 4);
        for (int ks = 0; ks < ksteps; ++ks)
          umma_bf16(tmem_base, a_hi + (a_lo + ks * 2u), b_hi + (b_lo + ks * b_step), idesc, (kb > 0 || ks > 0) ? 1u : 0u);
        umma_commit(empty + stage);
        if (kb == nkb - 1) umma_commit(tmem_full);
      }
      __syncwarp();
      if (++stage == p.nstage) { stage = 0; phase ^= 1; }
    }
  } else {
    mbar_wait(tmem_full, 0, 13);
    tc_fence_after();
    const int q = warp & 3;
    const int hh = 2 * q + (lane >> 4), ww = w0 + (lane & 15);
    const int h = h0 + hh;
    const uint32_t taddr = tmem_base + ((uint32_t)(q * 32) << 16);
    const int Cout = p.mode == 0 ? p.Co : p.Ci;
    for (int c = 0; c < p.BN; c += 16) {
      uint32_t r[16];
      tmem_ld_32x16(taddr + c, r);
      tmem_ld_wait();
      float v[16];
#pragma unroll
      for (int j = 0; j < 16; ++j) {
        v[j] = __uint_as_float(r[j]);
        if (p.bias) v[j] += __bfloat162float(p.bias[n0 + c + j]);
        if (p.relu) v[j] = fmaxf(v[j], 0.f);
      }
      if (p.pool) {
        uint32_t packed[8], am[4];
        pool2x2_bf16x16<1, 16>(v, packed, am);
        if (lane < 16 && !(lane & 1) && ww < p.Wo && h < p.Ho) {
          const size_t o = (((size_t)img * (p.Ho / 2) + h / 2) * (p.Wo / 2) + ww / 2) * Cout + n0 + c;
          *reinterpret_cast<uint4*>(p.out + o) = make_uint4(packed[0], packed[1], packed[2], packed[3]);
          *reinterpret_cast<uint4*>(p.out + o + 8) = make_uint4(packed[4], packed[5], packed[6], packed[7]);
          if (p.argmax) *reinterpret_cast<uint4*>(p.argmax + o) = make_uint4(am[0], am[1], am[2], am[3]);
        }
      } else if (ww < p.Wo && h < p.Ho) {
        const size_t o = (((size_t)img * p.Ho + h) * p.Wo + ww) * Cout + n0 + c;
        uint2 lo = pack_bf16x4(make_float4(v[0], v[1], v[2], v[3])), l2 = pack_bf16x4(make_float4(v[4], v[5], v[6], v[7]));
        uint2 hi = pack_bf16x4(make_float4(v[8], v[9], v[10], v[11])), h2 = pack_bf16x4(make_float4(v[12], v[13], v[14], v[15]));
        *reinterpret_cast<uint4*>(p.out + o) = make_uint4(lo.x, lo.y, l2.x, l2.y);
        *reinterpret_cast<uint4*>(p.out + o + 8) = make_uint4(hi.x, hi.y, h2.x, h2.y);
      }
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 1) tmem_dealloc(tmem_base, p.tmem_cols);
}

// ------------------------------------------------------------------------------------------------
// wgrad: M tile = G stacked (tap, channel-block) pairs, G*CB = 128; K = pixels.
// ------------------------------------------------------------------------------------------------
struct WgradParams {
  int B, Ci, Co, k, pad;
  int CB;            // channels per A box: 32 (64-byte rows) or 64 (128-byte rows)
  int G;             // pairs per M tile = 128 / CB
  int cblocks;       // Ci / CB
  int npairs;        // k*k*cblocks
  int BN;
  int tiles_h, tiles_w, ntiles, tiles_per_split;   // pixel tiles run over the dy tensor (Ho x Wo)
  float* dw;
  uint32_t tmem_cols;
  int nstage;
};

__global__ void __launch_bounds__(192, 1) conv_wgrad_tc_kernel(const __grid_constant__ CUtensorMap tmIn,
                                                               const __grid_constant__ CUtensorMap tmDy, const WgradParams p) {
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  uint8_t* smem = (uint8_t*)(((uintptr_t)smem_raw + 1023) & ~(uintptr_t)1023);
  const int box_bytes = TILE_PIX * p.CB * 2;          // one (tap, channel-block) box
  const int a_bytes = p.G * box_bytes;                // 32 KB
  const int b_bytes = p.BN * TILE_PIX * 2;
  const int stage_bytes = a_bytes + b_bytes;
  uint64_t* full = (uint64_t*)(smem + p.nstage * stage_bytes);
  uint64_t* empty = full + p.nstage;
  uint64_t* tmem_full = empty + p.nstage;
  uint32_t* tmem_slot = (uint32_t*)(tmem_full + 1);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int mt = blockIdx.x, n0 = blockIdx.y * p.BN;
  const int t0 = blockIdx.z * p.tiles_per_split, t1 = min(p.ntiles, t0 + p.tiles_per_split);
  if (t0 >= t1) return;
  const int pair0 = mt * p.G;
  const int npair_here = min(p.G, p.npairs - pair0);

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tmIn);
    tma_prefetch_desc(&tmDy);
    for (int s = 0; s < p.nstage; ++s) { mbar_init(full + s, 1); mbar_init(empty + s, 1); }
    mbar_init(tmem_full, 1);
    fence_barrier_init();
  }
  if (warp == 1) tmem_alloc(tmem_slot, p.tmem_cols);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  pdl_wait();          // everything above overlapped the previous kernel; its outputs are visible from here on
  pdl_launch_dependents();

  if (warp == 0) {
    const bool leader = elect_one();
    int stage = 0;
    uint32_t phase = 0;
    for (int t = t0; t < t1; ++t) {
      const int tiles_hw = p.tiles_h * p.tiles_w;
      const int img = t / tiles_hw, tile = t % tiles_hw;
      const int h0 = (tile / p.tiles_w) * TILE_H, w0 = (tile % p.tiles_w) * TILE_W;
      mbar_wait(empty + stage, phase ^ 1, 21);
      if (leader) {
        uint8_t* sA = smem + stage * stage_bytes;
        uint8_t* sB = sA + a_bytes;
        mbar_arrive_expect_tx(full + stage, (uint32_t)(npair_here * box_bytes + b_bytes));
        for (int j = 0; j < npair_here; ++j) {
          const int pr = pair0 + j;
          const int tap = pr / p.cblocks, cb = pr % p.cblocks;
          const int kh = tap / p.k, kw = tap % p.k;
          tma_load_4d(sA + j * box_bytes, &tmIn, full + stage, cb * p.CB, w0 + kw - p.pad, h0 + kh - p.pad, img);
        }
        for (int j = 0; j < p.BN / 64; ++j) tma_load_4d(sB + j * (TILE_PIX * 128), &tmDy, full + stage, n0 + 64 * j, w0, h0, img);
      }
      __syncwarp();
      if (++stage == p.nstage) { stage = 0; phase ^= 1; }
    }
  } else if (warp == 1) {
    const bool leader = elect_one();
    const uint32_t idesc = make_idesc_bf16(128, p.BN, 1, 1);
    const uint32_t a_swz = p.CB == 64 ? SW_128B : SW_64B;
    const int a_row = p.CB * 2;
    const uint64_t a_hi = make_smem_desc(0, box_bytes, 8 * a_row, a_swz);
    const uint64_t b_hi = make_smem_desc(0, TILE_PIX * 128, 1024, SW_128B);
    const uint32_t a_step = (uint32_t)(16 * a_row) >> 4, b_step = 2048u >> 4;
    const uint32_t s0 = (smem_u32(smem) & 0x3FFFFu) >> 4, stage_step = (uint32_t)stage_bytes >> 4;
    int stage = 0;
    uint32_t phase = 0;
    for (int t = t0; t < t1; ++t) {
      mbar_wait(full + stage, phase, 22);
      tc_fence_after();
      if (leader) {
        const uint32_t a_lo = s0 + stage * stage_step, b_lo = a_lo + ((uint32_t)a_bytes >> 4);
#pragma unroll
        for (int ks = 0; ks < TILE_PIX / 16; ++ks)      // 16 pixels (K) per instruction = two 8-row swizzle atoms
          umma_bf16(tmem_base, a_hi + (a_lo + ks * a_step), b_hi + (b_lo + ks * b_step), idesc, (t > t0 || ks > 0) ? 1u : 0u);
        umma_commit(empty + stage);
        if (t == t1 - 1) umma_commit(tmem_full);
      }
      __syncwarp();
      if (++stage == p.nstage) { stage = 0; phase ^= 1; }
    }
  } else {
    mbar_wait(tmem_full, 0, 23);
    tc_fence_after();
    const int q = warp & 3;
    const int row = q * 32 + lane;
    const int j = row / p.CB, ch = row % p.CB;
    const uint32_t taddr = tmem_base + ((uint32_t)(q * 32) << 16);
    const bool valid = j < npair_here;
    const int pr = pair0 + (valid ? j : 0);
    const int tap = pr / p.cblocks, cb = pr % p.cblocks;
    float* dst_row = p.dw + ((size_t)tap * p.Ci + cb * p.CB + ch) * p.Co + n0;
    for (int c = 0; c < p.BN; c += 16) {
      uint32_t r[16];
      tmem_ld_32x16(taddr + c, r);
      tmem_ld_wait();
      if (valid) {
#pragma unroll
        for (int e = 0; e < 16; e += 4)
          red_add_f4(dst_row + c + e, make_float4(__uint_as_float(r[e]), __uint_as_float(r[e + 1]), __uint_as_float(r[e + 2]),
                                                  __uint_as_float(r[e + 3])));
      }
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 1) tmem_dealloc(tmem_base, p.tmem_cols);
}

// dbias[c] += sum over rows of dy[rows, C]
__global__ void colsum_bf16_kernel(const bf16* __restrict__ dy, float* __restrict__ dbias, long long rows, int C) {
  const long long per = (rows + gridDim.x - 1) / gridDim.x;
  const long long r0 = blockIdx.x * per, r1 = min(rows, r0 + per);
  for (int c = threadIdx.x; c < C; c += blockDim.x) {
    float s = 0.f;
    for (long long r = r0; r < r1; ++r) s += __bfloat162float(dy[r * C + c]);
    atomicAdd(dbias + c, s);
  }
}

int act_tmap(CUtensorMap* tm, const void* base, int B, int H, int W, int C, int box_c) {
  const uint64_t dims[4] = {(uint64_t)C, (uint64_t)W, (uint64_t)H, (uint64_t)B};
  const uint64_t str[3] = {(uint64_t)C * 2, (uint64_t)W * C * 2, (uint64_t)H * W * C * 2};
  const uint32_t box[4] = {(uint32_t)box_c, TILE_W, TILE_H, 1};
  return tc::make_tmap_bf16(tm, base, 4, dims, str, box, box_c * 2);
}

int set_smem(const void* fn) {
  B200_CHECK(cudaFuncSetAttribute(fn, cudaFuncAttributeMaxDynamicSharedMemorySize, 220 * 1024));
  return B200_OK;
}

// H x W is the spatial size of the forward INPUT; the forward output (= dgrad input) is Hf x Wf = H+2*pad-k+1.
int launch_conv(int mode, const void* act, const void* w, const void* bias, void* out, uint8_t* argmax, int B, int H, int W,
                int Ci, int Co, int k, int pad, int relu, int pool, cudaStream_t st) {
  const int Cred = mode == 0 ? Ci : Co;      // reduction channels
  const int Cout = mode == 0 ? Co : Ci;      // channels produced
  B200_REQUIRE(k >= 1 && k <= 7 && pad >= 0 && pad < k, "conv_tc: kernel size %d (<= 7) / padding %d (< k) unsupported", k, pad);
  const int Hf = H + 2 * pad - k + 1, Wf = W + 2 * pad - k + 1;
  B200_REQUIRE(W >= 1 && H >= 1 && Hf >= 1 && Wf >= 1, "conv_tc: empty tensor (%dx%d -> %dx%d)", H, W, Hf, Wf);
  B200_REQUIRE(Cred % 32 == 0, "conv_tc: reduction channels %d must be a multiple of 32", Cred);
  const int Hs = mode == 0 ? H : Hf, Ws = mode == 0 ? W : Wf;       // tensor the TMA reads
  const int Ho = mode == 0 ? Hf : H, Wo = mode == 0 ? Wf : W;       // tensor this launch writes
  B200_REQUIRE(!pool || (Ho % 2 == 0 && Wo % 2 == 0), "conv_tc: pooling needs an even output size, got %dx%d", Ho, Wo);
  ConvParams p;
  p.B = B; p.Ho = Ho; p.Wo = Wo; p.Ci = Ci; p.Co = Co; p.k = k; p.pad = pad; p.mode = mode;
  p.KC = Cred % 64 == 0 ? 64 : 32;
  p.kc_blocks = Cred / p.KC;
  if (mode == 0) {
    B200_REQUIRE(Cout % 64 == 0, "conv_tc fwd: output channels %d must be a multiple of 64", Cout);
    p.BN = Cout % 128 == 0 ? 128 : 64;
  } else {
    B200_REQUIRE(Cout % 32 == 0, "conv_tc dgrad: input channels %d must be a multiple of 32", Cout);
    p.BN = Cout % 128 == 0 ? 128 : (Cout % 64 == 0 ? 64 : 32);
  }
  p.relu = relu; p.pool = pool; p.bias = (const bf16*)bias; p.out = (bf16*)out; p.argmax = argmax;
  p.tiles_h = (Ho + TILE_H - 1) / TILE_H;
  p.tiles_w = (Wo + TILE_W - 1) / TILE_W;
  uint32_t cols = 32;
  while ((int)cols < p.BN) cols <<= 1;
  p.tmem_cols = cols;
  const int a_bytes = TILE_PIX * p.KC * 2, b_bytes = p.BN * p.KC * 2;
  const int stage_bytes = a_bytes + ((b_bytes + 1023) & ~1023);
  p.nstage = 6;
  while (p.nstage > 2 && (size_t)p.nstage * stage_bytes > 190 * 1024) --p.nstage;
  const size_t smem = (size_t)p.nstage * stage_bytes + 1024 + 256;

  CUtensorMap tmA, tmB;
  int rc;
  if ((rc = act_tmap(&tmA, act, B, Hs, Ws, Cred, p.KC))) return rc;
  {
    const uint64_t dims[2] = {(uint64_t)Co, (uint64_t)k * k * Ci}, str[1] = {(uint64_t)Co * 2};
    if (mode == 0) {
      const uint32_t box[2] = {64, (uint32_t)p.KC};
      if ((rc = tc::make_tmap_bf16(&tmB, w, 2, dims, str, box, 128))) return rc;
    } else {
      const uint32_t box[2] = {(uint32_t)p.KC, (uint32_t)p.BN};
      if ((rc = tc::make_tmap_bf16(&tmB, w, 2, dims, str, box, p.KC * 2))) return rc;
    }
  }
  static bool attr = false;
  if (!attr) { if ((rc = set_smem((const void*)conv_tc_kernel))) return rc; attr = true; }
  dim3 grid(Cout / p.BN, B * p.tiles_h * p.tiles_w);
  B200_CHECK(launch_pdl(conv_tc_kernel, grid, dim3(192), smem, st, tmA, tmB, p));
  return B200_OK;
}

}  // namespace

int b200_conv_halo_launch(int mode, const void* act, const void* w, const void* bias, void* out, uint8_t* argmax, int B, int H,
                          int W, int Ci, int Co, int k, int relu, int pool, cudaStream_t st);

int b200_conv_res_launch(int mode, const void* act, const void* w, const void* bias, void* out, uint8_t* argmax, int B, int H,
                         int W, int Ci, int Co, int k, int pad, int relu, int pool, cudaStream_t st);

// B200DP_CONV_RES=0 keeps larger / VALID images on the shifted-TMA-box kernel instead of the smem-resident one
static bool res_enabled() {
  const char* e = getenv("B200DP_CONV_RES");
  return !(e && e[0] == '0');
}

static bool halo_enabled() {
  const char* e = getenv("B200DP_CONV_HALO");
  return !(e && e[0] == '0');
}

extern "C" {

int b200_conv_tc_fwd_pad(const void* in, const void* w, const void* bias, void* out, uint8_t* argmax, int B, int H, int W,
                         int Ci, int Co, int k, int pad, int relu, int pool, void* stream) {
  if (halo_enabled() && pad == k / 2 && (k & 1)) {
    const int rc = b200_conv_halo_launch(0, in, w, bias, out, argmax, B, H, W, Ci, Co, k, relu, pool, (cudaStream_t)stream);
    if (rc != B200_ERR_UNSUPPORTED) return rc;
  }
  if (res_enabled()) {
    const int rc = b200_conv_res_launch(0, in, w, bias, out, argmax, B, H, W, Ci, Co, k, pad, relu, pool, (cudaStream_t)stream);
    if (rc != B200_ERR_UNSUPPORTED) return rc;
  }
  return launch_conv(0, in, w, bias, out, argmax, B, H, W, Ci, Co, k, pad, relu, pool, (cudaStream_t)stream);
}

int b200_conv_tc_fwd(const void* in, const void* w, const void* bias, void* out, uint8_t* argmax, int B, int H, int W, int Ci,
                     int Co, int k, int relu, int pool, void* stream) {
  B200_REQUIRE(k % 2 == 1, "conv_tc: SAME padding needs an odd kernel size, got %d", k);
  return b200_conv_tc_fwd_pad(in, w, bias, out, argmax, B, H, W, Ci, Co, k, k / 2, relu, pool, stream);
}

int b200_conv_tc_dgrad_pad(const void* dout, const void* w, void* din, int B, int H, int W, int Ci, int Co, int k, int pad,
                           void* stream) {
  if (halo_enabled() && pad == k / 2 && (k & 1)) {
    const int rc = b200_conv_halo_launch(1, dout, w, nullptr, din, nullptr, B, H, W, Ci, Co, k, 0, 0, (cudaStream_t)stream);
    if (rc != B200_ERR_UNSUPPORTED) return rc;
  }
  if (res_enabled()) {
    const int rc = b200_conv_res_launch(1, dout, w, nullptr, din, nullptr, B, H, W, Ci, Co, k, pad, 0, 0, (cudaStream_t)stream);
    if (rc != B200_ERR_UNSUPPORTED) return rc;
  }
  return launch_conv(1, dout, w, nullptr, din, nullptr, B, H, W, Ci, Co, k, pad, 0, 0, (cudaStream_t)stream);
}

int b200_conv_tc_dgrad(const void* dout, const void* w, void* din, const void* mask_src, int B, int H, int W, int Ci, int Co,
                       int k, void* stream) {
  (void)mask_src;
  B200_REQUIRE(k % 2 == 1, "conv_tc: SAME padding needs an odd kernel size, got %d", k);
  return b200_conv_tc_dgrad_pad(dout, w, din, B, H, W, Ci, Co, k, k / 2, stream);
}

int b200_conv_tc_wgrad_pad(const void* in, const void* dout, float* dw, float* dbias, int B, int H, int W, int Ci, int Co,
                           int k, int pad, void* stream) {
  B200_REQUIRE(k >= 1 && k <= 7 && pad >= 0 && pad < k, "conv_tc wgrad: kernel size %d / padding %d unsupported", k, pad);
  const int Ho = H + 2 * pad - k + 1, Wo = W + 2 * pad - k + 1;
  B200_REQUIRE(Ho >= 1 && Wo >= 1, "conv_tc wgrad: empty output");
  B200_REQUIRE(Ci % 32 == 0 && (Ci == 32 || Ci % 64 == 0), "conv_tc wgrad: Ci=%d must be 32 or a multiple of 64", Ci);
  B200_REQUIRE(Co % 64 == 0, "conv_tc wgrad: Co=%d must be a multiple of 64", Co);
  cudaStream_t st = (cudaStream_t)stream;
  WgradParams p;
  p.B = B; p.Ci = Ci; p.Co = Co; p.k = k; p.pad = pad;
  p.CB = Ci == 32 ? 32 : 64;
  p.G = 128 / p.CB;
  p.cblocks = Ci / p.CB;
  p.npairs = k * k * p.cblocks;
  p.BN = Co % 128 == 0 ? 128 : 64;
  p.tiles_h = (Ho + TILE_H - 1) / TILE_H;
  p.tiles_w = (Wo + TILE_W - 1) / TILE_W;
  p.ntiles = B * p.tiles_h * p.tiles_w;
  p.dw = dw;
  uint32_t cols = 32;
  while ((int)cols < p.BN) cols <<= 1;
  p.tmem_cols = cols;
  const int mtiles = (p.npairs + p.G - 1) / p.G, ntile_n = Co / p.BN;
  static int target_ctas = 0;
  if (!target_ctas) {
    const char* e = getenv("B200DP_WGRAD_CTAS");      // tuning knob: CTAs the split over pixel tiles aims for
    target_ctas = e ? atoi(e) : 148;               // one wave: 12.8 us vs 16.3 us at two CTAs per SM (two uneven waves) on the MNIST conv2
    if (target_ctas < 1) target_ctas = 148;
  }
  int split = target_ctas / (mtiles * ntile_n);
  if (split < 1) split = 1;
  if (split > p.ntiles) split = p.ntiles;
  p.tiles_per_split = (p.ntiles + split - 1) / split;
  split = (p.ntiles + p.tiles_per_split - 1) / p.tiles_per_split;
  const int stage_bytes = 128 * 128 * 2 + p.BN * TILE_PIX * 2;
  p.nstage = 4;
  while (p.nstage > 2 && (size_t)p.nstage * stage_bytes > 190 * 1024) --p.nstage;
  const size_t smem = (size_t)p.nstage * stage_bytes + 1024 + 256;
  CUtensorMap tmIn, tmDy;
  int rc;
  if ((rc = act_tmap(&tmIn, in, B, H, W, Ci, p.CB))) return rc;
  if ((rc = act_tmap(&tmDy, dout, B, Ho, Wo, Co, 64))) return rc;
  static bool attr = false;
  if (!attr) { if ((rc = set_smem((const void*)conv_wgrad_tc_kernel))) return rc; attr = true; }
  B200_CHECK(launch_pdl(conv_wgrad_tc_kernel, dim3(mtiles, ntile_n, split), dim3(192), smem, st, tmIn, tmDy, p));
  if (dbias) {
    colsum_bf16_kernel<<<148, Co < 256 ? (Co + 31) / 32 * 32 : 256, 0, st>>>((const bf16*)dout, dbias, (long long)B * Ho * Wo, Co);
    B200_LAUNCH_CHECK();
  }
  return B200_OK;
}

int b200_conv_tc_wgrad(const void* in, const void* dout, float* dw, float* dbias, int B, int H, int W, int Ci, int Co, int k,
                       void* stream) {
  B200_REQUIRE(k % 2 == 1, "conv_tc: SAME padding needs an odd kernel size, got %d", k);
  return b200_conv_tc_wgrad_pad(in, dout, dw, dbias, B, H, W, Ci, Co, k, k / 2, stream);
}

}  // extern "C"
