// Persistent implicit-GEMM convolution with shared-memory-resident operands (fwd and dgrad) for sm_100a.
//
// conv_tc.cu applies filter taps by re-loading a shifted TMA box per tap: simple and general, but every
// activation byte crosses L2->SM k*k times.  When the whole filter bank fits in shared memory (the MNIST CNN's
// second convolution: 25 taps x 32 x 64 bf16 = 100 KB) this kernel instead
//   * loads the filter bank ONCE per CTA (TMA, 128B-swizzled tiles, one mbarrier per tap so the first MMAs can
//     start after 4 KB), and keeps it for every image the CTA processes (persistent loop, grid = min(#SM, B));
//   * loads each image ONCE (two loader warps, coalesced 16-byte loads) into the UMMA *no-swizzle* canonical
//     layout [channel-chunk of 8][pixel row-major, 16 wide][8 channels]: consecutive pixels of a chunk are 16 B
//     apart, so an 8-pixel core matrix is 128 contiguous bytes (SBO = 128 B) and the two 8-channel chunks of a
//     K=16 step are a chunk stride apart (LBO).  In that layout a filter tap is nothing but a different
//     16-byte-aligned START ADDRESS in the A descriptor: tap (kh,kw) starts (kh*16 + kw - pad) pixels later.
//     The halo is zeroed once per CTA and never written again: columns W..15 of each 16-wide row double as the
//     right padding of that row and the left padding of the next.
// Accumulators are double-buffered in TMEM (image parity x 2 row-tiles) so the pooling epilogue of one tile
// overlaps the MMAs of the next.  Same math and same epilogue as conv_tc.cu.
//
// Replaces cuDNN conv2d fwd/dgrad behind layers.conv2d (horovod/tensorflow_mnist.py:56-60) — SURVEY K5/K6/K13.
#include <stdlib.h>

#include "../include/b200dp.h"
#include "tc_sm100.cuh"

int b200_stable_weights();     // runtime.cu

namespace {

using namespace tc;

constexpr int ROWPIX = 16;           // pixels per stored image row (W real + zero columns)
constexpr int MAXTAPS = 49;

struct HaloParams {
  int B, H, W, Ci, Co, k, pad;
  int mode;            // 0 forward, 1 dgrad
  int Cred, N;         // reduction channels, produced channels
  int rows;            // stored rows per image buffer = 16 + 2*pad
  int chunk_bytes;     // rows * 16 pixels * 16 B
  int img_bytes;       // (Cred/8) * chunk_bytes
  int wtile_bytes;     // one tap's weight tile
  int relu, pool;
  const bf16* act;     // input activation [B,H,W,Cred]
  const bf16* bias;
  bf16* out;
  uint8_t* argmax;
  uint32_t tmem_cols;
  int dbg;             // experiment switches: 1 skip MMAs, 2 skip image loads, 4 skip weight loads
  int early_w;         // the filter bank was NOT written by the kernel launched just before: load it ahead of the dependency wait
};

__device__ __forceinline__ void halo_epilogue(const HaloParams& p, const float* sbias, uint32_t taddr, int img, int h0, int lane, int q) {
  const int hh = 2 * q + (lane >> 4), ww = lane & 15;
  const int h = h0 + hh;
  for (int c = 0; c < p.N; c += 16) {
    uint32_t r[16];
    tmem_ld_32x16(taddr + c, r);
    tmem_ld_wait();
    float v[16];
#pragma unroll
    for (int j = 0; j < 16; ++j) {
      v[j] = __uint_as_float(r[j]);
      v[j] += sbias[c + j];
      if (p.relu) v[j] = fmaxf(v[j], 0.f);
    }
    if (p.pool) {
      uint32_t packed[8], am[4];
      pool2x2_bf16x16<1, 16>(v, packed, am);
      if (lane < 16 && !(lane & 1) && ww < p.W && h < p.H) {
        const size_t o = (((size_t)img * (p.H / 2) + h / 2) * (p.W / 2) + ww / 2) * p.N + c;
        *reinterpret_cast<uint4*>(p.out + o) = make_uint4(packed[0], packed[1], packed[2], packed[3]);
        *reinterpret_cast<uint4*>(p.out + o + 8) = make_uint4(packed[4], packed[5], packed[6], packed[7]);
        if (p.argmax) *reinterpret_cast<uint4*>(p.argmax + o) = make_uint4(am[0], am[1], am[2], am[3]);
      }
    } else if (ww < p.W && h < p.H) {
      const size_t o = (((size_t)img * p.H + h) * p.W + ww) * p.N + c;
      uint2 lo = pack_bf16x4(make_float4(v[0], v[1], v[2], v[3])), l2 = pack_bf16x4(make_float4(v[4], v[5], v[6], v[7]));
      uint2 hi = pack_bf16x4(make_float4(v[8], v[9], v[10], v[11])), h2 = pack_bf16x4(make_float4(v[12], v[13], v[14], v[15]));
      *reinterpret_cast<uint4*>(p.out + o) = make_uint4(lo.x, lo.y, l2.x, l2.y);
      *reinterpret_cast<uint4*>(p.out + o + 8) = make_uint4(hi.x, hi.y, h2.x, h2.y);
    }
  }
}

template <int KK, int KSTEPS, int MODE>
__global__ void __launch_bounds__(320, 1) conv_halo_kernel(const __grid_constant__ CUtensorMap tmW, const HaloParams p) {
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  uint8_t* smem = (uint8_t*)(((uintptr_t)smem_raw + 1023) & ~(uintptr_t)1023);
  constexpr int ntaps = KK * KK;
  uint8_t* sW = smem;                                             // ntaps tiles, 1024-aligned each
  uint8_t* sImg0 = sW + ntaps * p.wtile_bytes + 128;              // 128 B of zeros precede every image buffer
  const int img_stride = p.img_bytes + 128;
  uint64_t* wfull = (uint64_t*)(sImg0 + 2 * img_stride);
  uint64_t* afull = wfull + ntaps;
  uint64_t* aempty = afull + 2;
  uint64_t* tfull = aempty + 2;
  uint64_t* tempty = tfull + 4;
  uint32_t* tmem_slot = (uint32_t*)(tempty + 4);
  float* sbias = (float*)(tmem_slot + 4);                      // p.N floats (zeros when the conv has no bias)

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int nimg = (p.B - (int)blockIdx.x + (int)gridDim.x - 1) / (int)gridDim.x;    // images this CTA handles

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tmW);
    for (int t = 0; t < ntaps; ++t) mbar_init(wfull + t, 1);
    for (int i = 0; i < 2; ++i) { mbar_init(afull + i, 128); mbar_init(aempty + i, 1); }
    for (int i = 0; i < 4; ++i) { mbar_init(tfull + i, 1); mbar_init(tempty + i, 4); }
    fence_barrier_init();
  }
  // Zero both image buffers and their guard bytes once: the loader warps only ever write real pixels, so the halo
  // (rows -pad..-1 and H.., columns W..15) and the guards stay zero for every image this CTA processes.
  {
    uint4* z = reinterpret_cast<uint4*>(sImg0 - 128);
    const int n16 = (2 * img_stride + 128) / 16;
    for (int i = threadIdx.x; i < n16; i += blockDim.x) z[i] = make_uint4(0u, 0u, 0u, 0u);
  }
  fence_proxy_async();
  if (warp == 1) tmem_alloc(tmem_slot, p.tmem_cols);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  auto load_filters = [&]() {
    // filter bank: once per CTA, one barrier per tap so the first MMAs start after 4 KB
    for (int t = 0; t < ntaps; ++t) {
      if (p.dbg & 4) { mbar_arrive(wfull + t); continue; }
      mbar_arrive_expect_tx(wfull + t, (uint32_t)p.wtile_bytes);
      tma_load_2d(sW + t * p.wtile_bytes, &tmW, wfull + t, 0, t * p.Ci);     // [Ci rows][Co cols]
    }
  };
  // Inside a training step the weights were written by the PREVIOUS step's update, complete long before this kernel's
  // predecessor even started (every kernel of the chain passes its own dependency wait before it releases its
  // successor), so the 100 KB filter bank streams in underneath the predecessor's tail instead of after it.
  if (p.early_w && warp == 0 && elect_one()) load_filters();
  pdl_wait();          // everything above overlapped the previous kernel; its outputs are visible from here on
  pdl_launch_dependents();
  // (the bias, a weight written by the optimizer kernel, is staged by the epilogue warps alone — see below — so the TMA,
  // MMA and loader warps do not wait for that global round trip)

  if (warp == 0) {
    const bool leader = elect_one();
    if (leader && !p.early_w) load_filters();
    __syncwarp();
  } else if (warp == 1) {
    // Uniform loops, fully unrolled over taps and K steps; one elected lane issues.  An A descriptor is a constant
    // high word plus (image base + tap pixel offset + chunk offset) in the 14-bit start-address field.
    const bool leader = elect_one();
    const uint32_t idesc = make_idesc_bf16(128, p.N, 0, MODE == 0 ? 1 : 0);
    const uint64_t a_hi = make_smem_desc(0, (uint32_t)p.chunk_bytes, 128, SW_NONE);
    const uint64_t b_hi = MODE == 0 ? make_smem_desc(0, 0, 1024, SW_128B) : make_smem_desc(0, 16, 1024, SW_128B);
    constexpr uint32_t b_step = (MODE == 0 ? 2048u : 32u) >> 4;
    const uint32_t a_step = (uint32_t)(2 * p.chunk_bytes) >> 4;
    const uint32_t w_lo = (smem_u32(sW) & 0x3FFFFu) >> 4, w_step = (uint32_t)p.wtile_bytes >> 4;
    constexpr int PAD = KK / 2;
    for (int it = 0; it < nimg; ++it) {
      const int buf = it & 1;
      mbar_wait(afull + buf, (uint32_t)(it >> 1) & 1u, 32);
      const uint32_t i_lo = (smem_u32(sImg0 + buf * img_stride) & 0x3FFFFu) >> 4;
#pragma unroll
      for (int tile = 0; tile < 2; ++tile) {
        const int acc = buf * 2 + tile;
        mbar_wait(tempty + acc, ((uint32_t)(it >> 1) & 1u) ^ 1u, 33);
        tc_fence_after();
        const uint32_t d_tmem = tmem_base + (uint32_t)(acc * p.N);
#pragma unroll
        for (int t = 0; t < ntaps; ++t) {
          if (it == 0 && tile == 0) { mbar_wait(wfull + t, 0, 34); tc_fence_after(); }
          const int kh = t / KK, kw = t % KK;
          // first input pixel (row-major, 16 per row; buffer row 0 = image row -pad) that feeds output pixel (8*tile, 0)
          const int pix = MODE == 0 ? (tile * 8 + kh) * ROWPIX + (kw - PAD) : (tile * 8 + 2 * PAD - kh) * ROWPIX + (PAD - kw);
          if (leader && !((p.dbg & 1) && t > 0)) {
#pragma unroll
            for (int ks = 0; ks < KSTEPS; ++ks)
              umma_bf16(d_tmem, a_hi + (uint64_t)(i_lo + (uint32_t)pix + ks * a_step), b_hi + (uint64_t)(w_lo + t * w_step + ks * b_step),
                        idesc, (t > 0 || ks > 0) ? 1u : 0u);
          }
        }
        if (leader) umma_commit(tfull + acc);
        __syncwarp();
      }
      if (leader) umma_commit(aempty + buf);          // both tiles' MMAs have consumed the image buffer
      __syncwarp();
    }
  } else if (warp >= 6) {
    // ===== image loaders (4 warps): coalesced 16-byte global loads, scattered into the chunked no-swizzle layout
    // [channel chunk][row][16 px][8 ch].  (A 5-D TMA box can produce this layout too, but with a 16-byte inner
    // dimension it issues one request per pixel-chunk and was the bottleneck of the kernel.)
    const int lt = threadIdx.x - 192;
    const int CK = p.Cred / 8;
    const int pieces = p.H * p.W * CK;
    constexpr int UNR = 13;              // 128 threads x 13 x 16 B covers the 25 KB dgrad image in one round trip
    for (int it = 0; it < nimg; ++it) {
      const int buf = it & 1;
      mbar_wait(aempty + buf, (((uint32_t)(it >> 1)) & 1u) ^ 1u, 31);
      const int img = blockIdx.x + it * gridDim.x;
      const uint4* src = reinterpret_cast<const uint4*>(p.act + (size_t)img * p.H * p.W * p.Cred);
      uint8_t* dst = sImg0 + buf * img_stride;
      if (!(p.dbg & 2)) {
        for (int base = 0; base < pieces; base += 128 * UNR) {
          uint4 v[UNR];
#pragma unroll
          for (int u = 0; u < UNR; ++u) {
            const int idx = base + u * 128 + lt;
            if (idx < pieces) v[u] = __ldg(src + idx);
          }
#pragma unroll
          for (int u = 0; u < UNR; ++u) {
            const int idx = base + u * 128 + lt;
            if (idx < pieces) {
              const int pixel = idx / CK, chunk = idx - pixel * CK;
              const int h = pixel / p.W, w = pixel - h * p.W;
              *reinterpret_cast<uint4*>(dst + chunk * p.chunk_bytes + ((h + p.pad) * ROWPIX + w) * 16) = v[u];
            }
          }
        }
      }
      fence_proxy_async();                 // generic-proxy stores -> visible to the tensor core's async proxy
      mbar_arrive(afull + buf);
    }
  } else {
    const int q = warp & 3;
    for (int i = threadIdx.x - 64; i < p.N; i += 128) sbias[i] = p.bias ? __bfloat162float(p.bias[i]) : 0.f;
    asm volatile("bar.sync 1, 128;" ::: "memory");          // warps 2-5 only
    for (int it = 0; it < nimg; ++it) {
      const int img = blockIdx.x + it * gridDim.x;
      for (int tile = 0; tile < 2; ++tile) {
        const int acc = (it & 1) * 2 + tile;
        mbar_wait(tfull + acc, (uint32_t)(it >> 1) & 1u, 35);
        tc_fence_after();
        if (!(p.dbg & 8)) halo_epilogue(p, sbias, tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)(acc * p.N), img, tile * 8, lane, q);
        tc_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive(tempty + acc);
      }
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 1) tmem_dealloc(tmem_base, p.tmem_cols);
}

}  // namespace

// Returns B200_ERR_UNSUPPORTED (without setting an error) when the shape does not fit this specialisation.
int b200_conv_halo_launch(int mode, const void* act, const void* w, const void* bias, void* out, uint8_t* argmax, int B, int H,
                          int W, int Ci, int Co, int k, int relu, int pool, cudaStream_t st) {
  const int pad = k / 2;
  const int Cred = mode == 0 ? Ci : Co, N = mode == 0 ? Co : Ci;
  const int ntaps = k * k;
  const long wbytes = (long)ntaps * Ci * Co * 2;
  if (k % 2 == 0 || ntaps > MAXTAPS || H > 16 || W + pad > 16 || Cred % 16 != 0 || Cred > 128) return B200_ERR_UNSUPPORTED;
  if (mode == 0 && N != 64) return B200_ERR_UNSUPPORTED;                       // MN-major B tile: one 128-byte row of co
  if (mode == 1 && (Cred != 64 || N % 16 != 0 || N > 256)) return B200_ERR_UNSUPPORTED;   // K-major B tile: 128-byte rows of co
  if (pool && (H % 2 || W % 2)) return B200_ERR_UNSUPPORTED;
  HaloParams p;
  p.B = B; p.H = H; p.W = W; p.Ci = Ci; p.Co = Co; p.k = k; p.pad = pad; p.mode = mode;
  p.Cred = Cred; p.N = N;
  p.rows = 16 + 2 * pad;
  p.chunk_bytes = p.rows * ROWPIX * 16;
  p.img_bytes = (Cred / 8) * p.chunk_bytes;
  p.wtile_bytes = Ci * Co * 2;
  if (p.wtile_bytes % 1024 != 0) return B200_ERR_UNSUPPORTED;
  p.relu = relu; p.pool = pool; p.bias = (const bf16*)bias; p.out = (bf16*)out; p.argmax = argmax;
  uint32_t cols = 32;
  while ((int)cols < 4 * N) cols <<= 1;
  if (cols > 512) return B200_ERR_UNSUPPORTED;
  p.tmem_cols = cols;
  { const char* e = getenv("B200DP_HALO_DBG"); p.dbg = e ? atoi(e) : 0; }
  p.early_w = b200_stable_weights();
  const size_t smem = 1024 + (size_t)wbytes + 128 + 2 * ((size_t)p.img_bytes + 128) + (ntaps + 12) * 8 + 64 + (size_t)N * 4;
  if (smem > 220 * 1024) return B200_ERR_UNSUPPORTED;

  CUtensorMap tmW;
  int rc;
  p.act = (const bf16*)act;
  {
    const uint64_t dims[2] = {(uint64_t)Co, (uint64_t)ntaps * Ci}, str[1] = {(uint64_t)Co * 2};
    const uint32_t box[2] = {(uint32_t)Co, (uint32_t)Ci};
    if ((rc = tc::make_tmap_bf16(&tmW, w, 2, dims, str, box, 128))) return rc;
  }
  const int ksteps = Cred / 16;
  void (*kern)(const CUtensorMap, const HaloParams) = nullptr;
  if (k == 5 && mode == 0 && ksteps == 2) kern = conv_halo_kernel<5, 2, 0>;
  else if (k == 5 && mode == 1 && ksteps == 4) kern = conv_halo_kernel<5, 4, 1>;
  else if (k == 3 && mode == 0 && ksteps == 2) kern = conv_halo_kernel<3, 2, 0>;
  else if (k == 3 && mode == 1 && ksteps == 4) kern = conv_halo_kernel<3, 4, 1>;
  else return B200_ERR_UNSUPPORTED;
  static const void* configured[8] = {};
  bool seen = false;
  for (const void* c : configured) seen |= (c == (const void*)kern);
  if (!seen) {
    B200_CHECK(cudaFuncSetAttribute((const void*)kern, cudaFuncAttributeMaxDynamicSharedMemorySize, 220 * 1024));
    for (const void*& c : configured)
      if (c == nullptr) { c = (const void*)kern; break; }
  }
  int grid = B < 148 ? B : 148;
  B200_CHECK(launch_pdl(kern, dim3(grid), dim3(320), smem, st, tmW, p));
  return B200_OK;
}
