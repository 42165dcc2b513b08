// tcgen05 GEMM for sm_100a: D[M,N] = A.B, bf16 operands staged by TMA into 128B-swizzled shared
// memory, fp32 accumulators in TMEM, warp-specialised (TMA producer / single-thread MMA issuer /
// 4 epilogue warps), mbarrier pipeline, optional split-K with vector red.add.
//
// Replaces the cuBLAS/Eigen matmuls behind layers.dense (horovod/tensorflow_mnist.py:65-70;
// horovod/tensorflow_mnist_gpu.py:62-66) — forward, dgrad and wgrad all map onto this kernel
// through the K-major / MN-major operand switches, so weights and activations are never transposed
// in memory.
#include <stdlib.h>

#include <mutex>

#include "../include/b200dp.h"
#include "tc_sm100.cuh"

const float* b200_comm_coef_cache(b200_comm_t h, int rank);   // comm.cu

namespace tc {

EncodeTiledFn get_encode_tiled() {
  static EncodeTiledFn fn = nullptr;
  static std::once_flag once;
  std::call_once(once, [] {
    void* p = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) == cudaSuccess &&
        q == cudaDriverEntryPointSuccess)
      fn = (EncodeTiledFn)p;
  });
  return fn;
}

int make_tmap_bf16(CUtensorMap* out, const void* base, int rank, const uint64_t* dims, const uint64_t* strides_bytes,
                   const uint32_t* box, uint32_t swizzle_bytes) {
  EncodeTiledFn enc = get_encode_tiled();
  B200_REQUIRE(enc != nullptr, "cuTensorMapEncodeTiled is not available from this driver");
  B200_REQUIRE(((uintptr_t)base & 15) == 0, "TMA base pointer %p is not 16-byte aligned", base);
  cuuint64_t gdim[5], gstr[4];
  cuuint32_t bx[5], es[5];
  for (int i = 0; i < rank; ++i) { gdim[i] = dims[i]; bx[i] = box[i]; es[i] = 1; }
  for (int i = 0; i + 1 < rank; ++i) {
    B200_REQUIRE(strides_bytes[i] % 16 == 0, "TMA stride %llu of dim %d is not a multiple of 16 bytes", (unsigned long long)strides_bytes[i], i + 1);
    gstr[i] = strides_bytes[i];
  }
  CUtensorMapSwizzle sw = swizzle_bytes == 128 ? CU_TENSOR_MAP_SWIZZLE_128B
                          : swizzle_bytes == 64 ? CU_TENSOR_MAP_SWIZZLE_64B
                          : swizzle_bytes == 32 ? CU_TENSOR_MAP_SWIZZLE_32B
                                                : CU_TENSOR_MAP_SWIZZLE_NONE;
  CUresult r = enc(out, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, (cuuint32_t)rank, const_cast<void*>(base), gdim, gstr, bx, es,
                   CU_TENSOR_MAP_INTERLEAVE_NONE, sw, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  B200_REQUIRE(r == CUDA_SUCCESS, "cuTensorMapEncodeTiled failed with CUresult %d (rank %d, dims %llu x %llu, box %u x %u)", (int)r,
               rank, (unsigned long long)dims[0], (unsigned long long)(rank > 1 ? dims[1] : 1), box[0], rank > 1 ? box[1] : 1);
  return B200_OK;
}

}  // namespace tc

namespace {

using namespace tc;

constexpr int BM = 128, BK = 64, MAX_STAGE = 4;
constexpr int A_TILE_BYTES = BM * BK * 2;     // 16 KB

struct GemmParams {
  void* C;
  int M, N, K, ldc;
  int a_mn, b_mn, out_mode, bn;
  int kb_total, kb_per_split;
  uint32_t tmem_cols;
  int nstage;          // pipeline depth: min(4, k-blocks per CTA) — a short K keeps shared memory small (more CTAs per SM)
  // out_mode 3 (GEMM -> reduce-scatter): row block r*rows_per_owner.. belongs to rank r; the tile is red.add'ed straight
  // into that rank's buffer over NVLink, so the reduce-scatter of the gradient allreduce happens tile by tile here
  float* peerC[B200_MAX_RANKS];
  int rows_per_owner;
  int bulk_rs;         // out_mode 3: 1 = bulk reduce-stores from a shared-memory staging row, 0 = per-thread 16-byte red.add
  // out_mode 4 (wgrad GEMM -> Adam): the tile is the cross-rank gradient SUM of rows this rank owns; the epilogue applies
  // the 1/size scale and the Adam update to the fp32 master weights / moments at flat index param_base + row*ldc + col and
  // stores the refreshed bf16 weights into every rank's working copy.  The gradient itself is never written anywhere.
  float* master;
  float* adam_m;
  float* adam_v;
  bf16* w16[B200_MAX_RANKS];
  int nranks;
  long long param_base;
  const int* step_ptr;
  const double* hyper;   // lr, beta1, beta2, eps, eps_mode (0: TF placement, 1: torch placement)
  float gscale;
  int tma_push;          // 1: the bf16 tile leaves through ONE TMA tensor store per destination rank (needs 16-byte aligned rows)
  const float* coef;     // this step's Adam coefficients (5 floats) when the engine caches them; else computed from step/hyper
};

// out_mode 4: one tensor map per destination rank over the bf16 rows this rank owns (stores clip at the slice's end)
struct TmapPack {
  CUtensorMap m[B200_MAX_RANKS];
};

struct AdamCoefG {
  float b1, b2, eps, step_size, vscale;
};
// The sweep runs on 12 warps per SM, so it is bound by its dependent instruction chain, not by HBM, unless the chain is
// short: sqrt and reciprocal are the SFU approximations (2 ulp; IEEE sqrtf and '/' expand to ~60 instructions with
// fix-up branches per element and made this epilogue 3x slower than the memory system allows).  The owner rank computes
// each weight once and every replica receives that value, so replicas stay bitwise identical whatever the rounding.
__device__ __forceinline__ void adam1(float g, float& p, float& m, float& v, const AdamCoefG& c) {
  m = fmaf(c.b1, m, (1.f - c.b1) * g);
  v = fmaf(c.b2, v, (1.f - c.b2) * g * g);
  float s, r;
  asm("sqrt.approx.ftz.f32 %0, %1;" : "=f"(s) : "f"(v));
  asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(fmaf(s, c.vscale, c.eps)));
  p = fmaf(-c.step_size * m, r, p);
}

__global__ void __launch_bounds__(192, 2) gemm_tc_kernel(const __grid_constant__ CUtensorMap tmA,
                                                         const __grid_constant__ CUtensorMap tmB,
                                                         const __grid_constant__ TmapPack tmW, const GemmParams p) {
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  uint8_t* smem = (uint8_t*)(((uintptr_t)smem_raw + 1023) & ~(uintptr_t)1023);
  const int stage_bytes = A_TILE_BYTES + p.bn * BK * 2;
  const int NSTAGE = p.nstage;
  uint64_t* full = (uint64_t*)(smem + NSTAGE * stage_bytes);
  uint64_t* empty = full + NSTAGE;
  uint64_t* tmem_full = empty + NSTAGE;
  uint32_t* tmem_slot = (uint32_t*)(tmem_full + 1);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int n0 = blockIdx.x * p.bn, m0 = blockIdx.y * BM;
  const int kb0 = blockIdx.z * p.kb_per_split;
  const int kb1 = min(p.kb_total, kb0 + p.kb_per_split);
  if (kb0 >= kb1) return;

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tmA);
    tma_prefetch_desc(&tmB);
    for (int s = 0; s < NSTAGE; ++s) { mbar_init(full + s, 1); mbar_init(empty + s, 1); }
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
    // ===== TMA producer: the whole warp walks the loop (uniform control flow), one elected lane issues =====
    const bool leader = elect_one();
    int stage = 0;
    uint32_t phase = 0;
    for (int kb = kb0; kb < kb1; ++kb) {
      mbar_wait(empty + stage, phase ^ 1, 1);
      if (leader) {
        uint8_t* sA = smem + stage * stage_bytes;
        uint8_t* sB = sA + A_TILE_BYTES;
        mbar_arrive_expect_tx(full + stage, (uint32_t)stage_bytes);
        const int k0 = kb * BK;
        if (!p.a_mn) {
          tma_load_2d(sA, &tmA, full + stage, k0, m0);
        } else {
          tma_load_2d(sA, &tmA, full + stage, m0, k0);
          tma_load_2d(sA + 8192, &tmA, full + stage, m0 + 64, k0);
        }
        if (!p.b_mn) {
          tma_load_2d(sB, &tmB, full + stage, k0, n0);
        } else {
          for (int j = 0; j < p.bn / 64; ++j) tma_load_2d(sB + j * 8192, &tmB, full + stage, n0 + 64 * j, k0);
        }
      }
      __syncwarp();
      if (++stage == NSTAGE) { stage = 0; phase ^= 1; }
    }
  } else if (warp == 1) {
    // ===== MMA issuer: uniform loop, one elected lane drives the tensor core. Descriptors are a constant high
    // word plus a running start-address field, so an issue costs a handful of (uniform-datapath) instructions.
    const bool leader = elect_one();
    const uint32_t idesc = make_idesc_bf16(BM, p.bn, p.a_mn, p.b_mn);
    const uint64_t a_hi = p.a_mn ? make_smem_desc(0, 8192, 1024, SW_128B) : make_smem_desc(0, 16, 1024, SW_128B);
    const uint64_t b_hi = p.b_mn ? make_smem_desc(0, 8192, 1024, SW_128B) : make_smem_desc(0, 16, 1024, SW_128B);
    const uint32_t a_step = (p.a_mn ? 2048u : 32u) >> 4, b_step = (p.b_mn ? 2048u : 32u) >> 4;
    const uint32_t s0 = (smem_u32(smem) & 0x3FFFFu) >> 4, stage_step = (uint32_t)stage_bytes >> 4;
    int stage = 0;
    uint32_t phase = 0;
    for (int kb = kb0; kb < kb1; ++kb) {
      mbar_wait(full + stage, phase, 2);
      tc_fence_after();
      if (leader) {
        const uint32_t a_lo = s0 + stage * stage_step, b_lo = a_lo + (A_TILE_BYTES >> 4);
#pragma unroll
        for (int k = 0; k < BK / 16; ++k)
          umma_bf16(tmem_base, a_hi + (a_lo + k * a_step), b_hi + (b_lo + k * b_step), idesc, (kb > kb0 || k > 0) ? 1u : 0u);
        umma_commit(empty + stage);                 // frees the smem slot when these MMAs retire
        if (kb == kb1 - 1) umma_commit(tmem_full);  // accumulator complete
      }
      __syncwarp();
      if (++stage == NSTAGE) { stage = 0; phase ^= 1; }
    }
  } else {
    // ===== epilogue: TMEM -> registers -> global (thread = one output row, 16 columns per step) =====
    if (p.out_mode != 4) {    // mode 4 computes its Adam coefficients first and waits for the accumulator afterwards
      mbar_wait(tmem_full, 0, 3);
      tc_fence_after();
    }
    const int q = warp & 3;                         // TMEM lane quarter this warp may touch
    const int row = m0 + q * 32 + lane;
    const uint32_t taddr = tmem_base + ((uint32_t)(q * 32) << 16);
    const int out_mode = p.out_mode, N = p.N, bn = p.bn;
    const bool row_ok = row < p.M;
    bf16* const c16 = (bf16*)p.C + (size_t)row * p.ldc;
    float* const c32 = (out_mode == 3 ? p.peerC[min(row / p.rows_per_owner, B200_MAX_RANKS - 1)] : (float*)p.C) + (size_t)row * p.ldc;
    if (out_mode == 3 && p.bulk_rs) {
      // GEMM -> reduce-scatter, bulk form: each thread parks 64 columns of its row (256 contiguous bytes) in shared
      // memory and hands them to the TMA unit as ONE reduce-store (cp.reduce.async.bulk ... add.f32) aimed at the
      // owner rank's buffer — peer memory over NVLink, or local.  256-byte packets instead of 16-byte atomics.
      float* stg = reinterpret_cast<float*>(smem + NSTAGE * stage_bytes + 256) + (size_t)(q * 32 + lane) * 64;
      const uint32_t stg_u = smem_u32(stg);
      for (int c = 0; c < bn; c += 64) {
#pragma unroll
        for (int hq = 0; hq < 4; ++hq) {
          uint32_t rr[16];
          if (c + hq * 16 < bn) {
            tmem_ld_32x16(taddr + c + hq * 16, rr);
            tmem_ld_wait();
#pragma unroll
            for (int j = 0; j < 16; j += 4)
              *reinterpret_cast<uint4*>(stg + hq * 16 + j) = make_uint4(rr[j], rr[j + 1], rr[j + 2], rr[j + 3]);
          }
        }
        fence_proxy_async();
        const int col = n0 + c;
        const int ncol = min(64, min(bn - c, N - col));
        if (row_ok && ncol > 0) {
          if (ncol % 4 == 0) {
            asm volatile("cp.reduce.async.bulk.global.shared::cta.bulk_group.add.f32 [%0], [%1], %2;" ::"l"(c32 + col), "r"(stg_u),
                         "r"(ncol * 4)
                         : "memory");
          } else {
            for (int j = 0; j < ncol; ++j) atomicAdd_system(c32 + col + j, stg[j]);
          }
        }
        asm volatile("cp.async.bulk.commit_group;" ::: "memory");
        asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory");      // staging row may be overwritten now
      }
      asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");
    } else if (out_mode == 4) {
      // wgrad GEMM -> Adam, part 1 (the four TMEM warps): Adam coefficients (fp64 bias corrections from the device step
      // counter, overlapped with the main loop), then the whole accumulator tile goes TMEM -> shared memory.  The
      // pipeline stages are free by then (every MMA has retired), so the tile reuses them: row pitch bn floats, 16-byte
      // columns XOR-swizzled with the row so that this row-per-thread store and the coalesced sweep below are both
      // conflict-free.  Part 2 (all six warps) follows the role branches.
      if (warp == 2 && lane == 0 && p.coef != nullptr) {
        AdamCoefG* co = reinterpret_cast<AdamCoefG*>(smem + NSTAGE * stage_bytes + 128);
        co->b1 = p.coef[0]; co->b2 = p.coef[1]; co->eps = p.coef[2]; co->step_size = p.coef[3]; co->vscale = p.coef[4];
      } else if (warp == 2 && lane == 0) {
        AdamCoefG* co = reinterpret_cast<AdamCoefG*>(smem + NSTAGE * stage_bytes + 128);
        const int t = *p.step_ptr;
        const double lr = p.hyper[0], b1 = p.hyper[1], b2 = p.hyper[2], eps = p.hyper[3];
        const double bc1 = 1.0 - pow(b1, (double)t), bc2 = 1.0 - pow(b2, (double)t);
        co->b1 = (float)b1; co->b2 = (float)b2; co->eps = (float)eps;
        if (p.hyper[4] == 0.0) { co->step_size = (float)(lr * sqrt(bc2) / bc1); co->vscale = 1.f; }
        else { co->step_size = (float)(lr / bc1); co->vscale = (float)(1.0 / sqrt(bc2)); }
      }
      mbar_wait(tmem_full, 0, 3);
      tc_fence_after();
      float* tile = reinterpret_cast<float*>(smem);
      const int rl = q * 32 + lane;
      const int f4_per_row = bn >> 2;
      for (int c0 = 0; c0 < bn; c0 += 16) {
        uint32_t rr[16];
        tmem_ld_32x16(taddr + c0, rr);
        tmem_ld_wait();
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          const int c4 = (c0 >> 2) + j;
          *reinterpret_cast<uint4*>(tile + (size_t)rl * bn + ((c4 ^ (rl & 7)) << 2)) =
              make_uint4(rr[4 * j], rr[4 * j + 1], rr[4 * j + 2], rr[4 * j + 3]);
        }
      }
      (void)f4_per_row;
    } else {
    uint32_t r[2][16];
    tmem_ld_32x16(taddr, r[0]);
#pragma unroll 1
    for (int c = 0; c < bn; c += 32) {
#pragma unroll
      for (int half = 0; half < 2; ++half) {
        const int cc = c + half * 16;
        if (cc >= bn) break;
        tmem_ld_wait();                                              // r[half] is ready
        if (cc + 16 < bn) tmem_ld_32x16(taddr + cc + 16, r[half ^ 1]);  // next chunk in flight while this one is stored
        if (!row_ok) continue;
        const int col = n0 + cc;
        const int valid = N - col;
        const uint32_t* v = r[half];
        if (valid >= 16) {
          if (out_mode == 1) {
            uint2 t0 = pack_bf16x4(make_float4(__uint_as_float(v[0]), __uint_as_float(v[1]), __uint_as_float(v[2]), __uint_as_float(v[3])));
            uint2 t1 = pack_bf16x4(make_float4(__uint_as_float(v[4]), __uint_as_float(v[5]), __uint_as_float(v[6]), __uint_as_float(v[7])));
            uint2 t2 = pack_bf16x4(make_float4(__uint_as_float(v[8]), __uint_as_float(v[9]), __uint_as_float(v[10]), __uint_as_float(v[11])));
            uint2 t3 = pack_bf16x4(make_float4(__uint_as_float(v[12]), __uint_as_float(v[13]), __uint_as_float(v[14]), __uint_as_float(v[15])));
            *reinterpret_cast<uint4*>(c16 + col) = make_uint4(t0.x, t0.y, t1.x, t1.y);
            *reinterpret_cast<uint4*>(c16 + col + 8) = make_uint4(t2.x, t2.y, t3.x, t3.y);
          } else {
#pragma unroll
            for (int j = 0; j < 16; j += 4) {
              const float4 f = make_float4(__uint_as_float(v[j]), __uint_as_float(v[j + 1]), __uint_as_float(v[j + 2]), __uint_as_float(v[j + 3]));
              if (out_mode == 2) red_add_f4(c32 + col + j, f);
              else if (out_mode == 3) red_add_sys_f4(c32 + col + j, f);
              else st_f4(c32 + col + j, f);
            }
          }
        } else if (valid > 0) {
#pragma unroll
          for (int j = 0; j < 16; ++j) {
            if (j < valid) {
              const float f = __uint_as_float(v[j]);
              if (out_mode == 1) c16[col + j] = __float2bfloat16(f);
              else if (out_mode == 2) atomicAdd(c32 + col + j, f);
              else if (out_mode == 3) atomicAdd_system(c32 + col + j, f);
              else c32[col + j] = f;
            }
          }
        }
      }
    }
    }
  }
  if (p.out_mode == 4) {
    // Part 2: all 192 threads sweep the tile coalesced (consecutive threads = consecutive 16-byte columns of a parameter
    // row) through master/m/v with 18 independent 16-byte loads in flight per thread — an HBM-bound pass needs ~64 KB
    // in flight per SM — apply Adam and store fp32 state + the bf16 weights of every rank.
    __syncthreads();
    const AdamCoefG co = *reinterpret_cast<const AdamCoefG*>(smem + NSTAGE * stage_bytes + 128);
    const float* tile = reinterpret_cast<const float*>(smem);
    bf16* wt = reinterpret_cast<bf16*>(smem + NSTAGE * stage_bytes + 256);     // [128][bn] bf16, row-major: the TMA store's source
    const int bn = p.bn, f4_per_row = bn >> 2, total = BM * f4_per_row;
    const int nr = p.nranks;
    const float gs = p.gscale;
    constexpr int U = 6;
#pragma unroll 1
    for (int f0 = threadIdx.x; f0 < total; f0 += U * 192) {
      float4 pp[U], mm[U], vv[U];
      long long idx[U];
      bool ok[U];
#pragma unroll
      for (int u = 0; u < U; ++u) {
        const int f = f0 + u * 192;
        const int rl = f / f4_per_row, c4 = f - rl * f4_per_row;
        const int grow = m0 + rl, col = n0 + c4 * 4;
        ok[u] = f < total && grow < p.M && col < p.N;
        idx[u] = p.param_base + (long long)grow * p.ldc + col;
        if (ok[u]) { pp[u] = ld_f4(p.master + idx[u]); mm[u] = ld_f4(p.adam_m + idx[u]); vv[u] = ld_f4(p.adam_v + idx[u]); }
      }
#pragma unroll
      for (int u = 0; u < U; ++u) {
        if (!ok[u]) continue;
        const int f = f0 + u * 192;
        const int rl = f / f4_per_row, c4 = f - rl * f4_per_row;
        const float4 g = *reinterpret_cast<const float4*>(tile + (size_t)rl * bn + ((c4 ^ (rl & 7)) << 2));
        adam1(g.x * gs, pp[u].x, mm[u].x, vv[u].x, co); adam1(g.y * gs, pp[u].y, mm[u].y, vv[u].y, co);
        adam1(g.z * gs, pp[u].z, mm[u].z, vv[u].z, co); adam1(g.w * gs, pp[u].w, mm[u].w, vv[u].w, co);
        st_f4(p.master + idx[u], pp[u]); st_f4(p.adam_m + idx[u], mm[u]); st_f4(p.adam_v + idx[u], vv[u]);
        const uint2 w = pack_bf16x4(pp[u]);
        if (p.tma_push) {
          *reinterpret_cast<uint2*>(wt + (size_t)rl * bn + c4 * 4) = w;
        } else {
#pragma unroll 1
          for (int r = 0; r < nr; ++r) *reinterpret_cast<uint2*>(p.w16[r] + idx[u]) = w;
        }
      }
    }
    if (p.tma_push) {
      // The refreshed bf16 tile goes to every rank (this one included) as ONE TMA tensor store each: the copy engine
      // streams the 128 rows over NVLink while the SM is already done.  Per-thread 8-byte stores to 8 peers held this
      // kernel at 27-29 us on 2 and 8 GPUs (10-12 us of it is the sweep): remote stores back up in the store queue.
      fence_proxy_async();                 // generic-proxy writes of the tile -> visible to the async proxy
      __syncthreads();
      if (threadIdx.x == 0) {
        for (int r = 0; r < nr; ++r)
          asm volatile("cp.async.bulk.tensor.2d.global.shared::cta.bulk_group [%0, {%1, %2}], [%3];" ::"l"(&tmW.m[r]), "r"(n0), "r"(m0),
                       "r"(smem_u32(wt))
                       : "memory");
        asm volatile("cp.async.bulk.commit_group;" ::: "memory");
        asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");     // stores performed before the grid completes
        if (nr > 1 && p.tma_push == 2) __threadfence_system();
      }
    } else if (nr > 1) {
      __threadfence_system();   // weights pushed into the peers are visible before this grid completes
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 1) tmem_dealloc(tmem_base, p.tmem_cols);
}

}  // namespace

struct AdamEpilogue {
  float* master; float* m; float* v;
  bf16* w16[B200_MAX_RANKS];
  int nranks;
  long long param_base;
  const int* step_ptr;
  const double* hyper;
  float gscale;
  const float* coef;
};

static int gemm_tc_launch(const void* A, const void* B, void* C, float* const* peerC, int rows_per_owner, int M, int N, int K, int lda,
                          int ldb, int ldc, int a_mn_major, int b_mn_major, int out_mode, int split_k, int block_n, void* stream,
                          const AdamEpilogue* adam = nullptr, const TmapPack* wmaps = nullptr) {
  B200_REQUIRE(M > 0 && N > 0 && K > 0, "gemm_tc: bad shape %dx%dx%d", M, N, K);
  B200_REQUIRE(out_mode >= 0 && out_mode <= 4, "gemm_tc: out_mode %d", out_mode);
  B200_REQUIRE((out_mode == 4) == (adam != nullptr), "gemm_tc: out_mode 4 needs the Adam epilogue arguments");
  B200_REQUIRE(split_k >= 1, "gemm_tc: split_k %d", split_k);
  B200_REQUIRE(split_k == 1 || out_mode >= 2, "gemm_tc: split_k > 1 needs an accumulating out_mode");
  B200_REQUIRE(block_n >= 16 && block_n <= 256 && block_n % 16 == 0, "gemm_tc: block_n %d must be a multiple of 16 in [16,256]", block_n);
  B200_REQUIRE(!b_mn_major || block_n % 64 == 0, "gemm_tc: MN-major B needs block_n %% 64 == 0 (got %d)", block_n);
  B200_REQUIRE(lda % 8 == 0 && ldb % 8 == 0, "gemm_tc: lda/ldb must be multiples of 8 elements (16 bytes)");
  B200_REQUIRE(ldc % 4 == 0 && ((uintptr_t)C & 15) == 0, "gemm_tc: C must be 16-byte aligned with ldc %% 4 == 0");
  B200_REQUIRE(out_mode != 4 || (N % 4 == 0 && split_k == 1), "gemm_tc: the Adam epilogue needs N %% 4 == 0 and no split-K");

  CUtensorMap tmA, tmB;
  int rc;
  if (!a_mn_major) {   // stored [M,K]
    const uint64_t dims[2] = {(uint64_t)K, (uint64_t)M}, str[1] = {(uint64_t)lda * 2};
    const uint32_t box[2] = {64, 128};
    if ((rc = tc::make_tmap_bf16(&tmA, A, 2, dims, str, box, 128))) return rc;
  } else {             // stored [K,M]
    const uint64_t dims[2] = {(uint64_t)M, (uint64_t)K}, str[1] = {(uint64_t)lda * 2};
    const uint32_t box[2] = {64, 64};
    if ((rc = tc::make_tmap_bf16(&tmA, A, 2, dims, str, box, 128))) return rc;
  }
  if (!b_mn_major) {   // stored [N,K]
    const uint64_t dims[2] = {(uint64_t)K, (uint64_t)N}, str[1] = {(uint64_t)ldb * 2};
    const uint32_t box[2] = {64, (uint32_t)block_n};
    if ((rc = tc::make_tmap_bf16(&tmB, B, 2, dims, str, box, 128))) return rc;
  } else {             // stored [K,N]
    const uint64_t dims[2] = {(uint64_t)N, (uint64_t)K}, str[1] = {(uint64_t)ldb * 2};
    const uint32_t box[2] = {64, 64};
    if ((rc = tc::make_tmap_bf16(&tmB, B, 2, dims, str, box, 128))) return rc;
  }
  GemmParams p;
  memset(&p, 0, sizeof p);
  p.rows_per_owner = rows_per_owner > 0 ? rows_per_owner : 1;
  if (peerC)
    for (int r = 0; r < B200_MAX_RANKS; ++r) p.peerC[r] = peerC[r];
  if (adam) {
    p.master = adam->master; p.adam_m = adam->m; p.adam_v = adam->v;
    for (int r = 0; r < B200_MAX_RANKS; ++r) p.w16[r] = adam->w16[r];
    p.nranks = adam->nranks; p.param_base = adam->param_base; p.step_ptr = adam->step_ptr; p.hyper = adam->hyper;
    p.gscale = adam->gscale;
    {
      // wait_group 0 above has already waited for the stores to be performed; the extra system fence (B200DP_PUSH_FENCE=1)
      // measured as a third of this kernel's stall samples (MEMBAR.SC.SYS + ERRBAR) and is off by default
      const char* e = getenv("B200DP_PUSH_FENCE");
      p.tma_push = wmaps ? ((e && e[0] == '1') ? 2 : 1) : 0;
    }
    p.coef = adam->coef;
  }
  p.C = C; p.M = M; p.N = N; p.K = K; p.ldc = ldc;
  p.a_mn = a_mn_major; p.b_mn = b_mn_major; p.out_mode = out_mode; p.bn = block_n;
  p.kb_total = (K + BK - 1) / BK;
  if (split_k > p.kb_total) split_k = p.kb_total;
  p.kb_per_split = (p.kb_total + split_k - 1) / split_k;
  split_k = (p.kb_total + p.kb_per_split - 1) / p.kb_per_split;
  uint32_t cols = 32;
  while ((int)cols < block_n) cols <<= 1;
  p.tmem_cols = cols;
  p.nstage = p.kb_per_split < MAX_STAGE ? p.kb_per_split : MAX_STAGE;
  if (p.nstage < 2) p.nstage = 2;
  if (out_mode == 4) {
    // the accumulator tile (128 x bn fp32) is parked in the stage buffers: two CTAs per SM matter more than pipeline depth
    // for this memory-bound epilogue, so at most 3 stages, and at least enough bytes for the tile
    if (p.nstage > 3) p.nstage = 3;
    while ((size_t)p.nstage * (A_TILE_BYTES + block_n * BK * 2) < (size_t)BM * block_n * 4) ++p.nstage;
  }
  {
    const char* e = getenv("B200DP_GEMM_RS_BULK");
    p.bulk_rs = (out_mode == 3 && !(e && e[0] == '0')) ? 1 : 0;
  }
  const size_t smem = (size_t)p.nstage * (A_TILE_BYTES + block_n * BK * 2) + 1024 /*align slack*/ + 256 /*barriers*/ +
                      (p.bulk_rs ? 128 * 64 * sizeof(float) : 0) + (p.tma_push ? (size_t)BM * block_n * 2 : 0);
  static size_t smem_set = 0;
  if (smem > smem_set) {
    B200_CHECK(cudaFuncSetAttribute(gemm_tc_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
    smem_set = 200 * 1024;
  }
  B200_REQUIRE(smem <= 200 * 1024, "gemm_tc: %zu bytes of shared memory needed", smem);
  dim3 grid((N + block_n - 1) / block_n, (M + BM - 1) / BM, split_k);
  static const TmapPack no_maps = {};
  B200_CHECK(launch_pdl(gemm_tc_kernel, grid, dim3(192), smem, (cudaStream_t)stream, tmA, tmB, wmaps ? *wmaps : no_maps, p));
  return B200_OK;
}

extern "C" int b200_gemm_tc(const void* A, const void* B, void* C, int M, int N, int K, int lda, int ldb, int ldc,
                            int a_mn_major, int b_mn_major, int out_mode, int split_k, int block_n, void* stream) {
  B200_REQUIRE(out_mode >= 0 && out_mode <= 2, "gemm_tc: out_mode %d", out_mode);
  return gemm_tc_launch(A, B, C, nullptr, 0, M, N, K, lda, ldb, ldc, a_mn_major, b_mn_major, out_mode, split_k, block_n, stream);
}

// GEMM fused with the reduce-scatter of a gradient allreduce: D = A.B is never stored locally; every 128xBN tile is
// added (red.add, system scope) into the buffer of the rank that owns those rows, over NVLink, as soon as the tile
// leaves TMEM.  c_seg/c_off name the [M,N] fp32 accumulation buffer inside a symmetric segment (it must be zero
// when the step starts); rank r owns rows [r*M/n, (r+1)*M/n).  M must be divisible by the world size.
extern "C" int b200_gemm_tc_reduce_scatter(b200_comm_t comm, const void* A, const void* B, int c_seg, size_t c_off, int nranks,
                                           int M, int N, int K, int lda, int ldb, int ldc, int a_mn_major, int b_mn_major,
                                           int block_n, void* stream) {
  B200_REQUIRE(nranks >= 1 && nranks <= B200_MAX_RANKS && M % nranks == 0, "gemm_tc_reduce_scatter: M=%d not divisible by %d ranks", M, nranks);
  B200_REQUIRE(c_off % 16 == 0, "gemm_tc_reduce_scatter: offset not 16-byte aligned");
  float* peers[B200_MAX_RANKS] = {};
  for (int r = 0; r < nranks; ++r) {
    char* base = (char*)b200_comm_seg_ptr(comm, c_seg, r);
    B200_REQUIRE(base != nullptr, "gemm_tc_reduce_scatter: segment %d is not mapped for rank %d", c_seg, r);
    peers[r] = (float*)(base + c_off);
  }
  for (int r = nranks; r < B200_MAX_RANKS; ++r) peers[r] = peers[nranks - 1];
  return gemm_tc_launch(A, B, peers[0], peers, M / nranks, M, N, K, lda, ldb, ldc, a_mn_major, b_mn_major, 3, 1, block_n, stream);
}

// Dense-layer weight gradient fused with its optimizer step — and, across ranks, with the gradient "allreduce" itself:
// A and B hold the activations / upstream gradients of the GLOBAL batch (every rank's rows, gathered by
// b200_allgather_acts), so D = A^T.B restricted to the rows this rank owns IS the cross-rank gradient sum of those rows,
// accumulated in fp32 in TMEM in a fixed order.  The epilogue scales it (1/size), applies Adam to the owner's fp32
// master weights / moments and stores the new bf16 weights into every rank's working copy over NVLink.
//   A: stored [K, M] (activations of the owned input-feature slice, M contiguous), B: stored [K, N] (dz), K = nranks*batch.
//   state_seg holds master | m | v at the given byte offsets; element (row, col) of D is parameter
//   param_start + (row0 + row) * N + col.  w16_seg/w16_off: bf16 working weights (peer-mapped: all ranks are written).
//   as_rank: -1 = this communicator's rank; an emulated communicator (all ranks on one GPU) names the rank to run as.
//   The kernel waits on nothing: order it after the gather (and after every reader of the old weights) with streams.
extern "C" int b200_gemm_tc_wgrad_adam(b200_comm_t comm, const void* A, const void* B, int M, int N, int K, int lda, int ldb,
                                       int state_seg, size_t master_off, size_t m_off, size_t v_off, int w16_seg, size_t w16_off,
                                       int ctrl_seg, size_t step_off, size_t hyper_off, long long param_start, int row0,
                                       int as_rank, float gscale, int block_n, void* stream) {
  const int nranks = b200_comm_nranks(comm);
  if (as_rank < 0) as_rank = b200_comm_rank(comm);      // emulated communicators name the rank to run as
  B200_REQUIRE(as_rank >= 0 && as_rank < nranks, "gemm_tc_wgrad_adam: rank %d outside world of %d", as_rank, nranks);
  B200_REQUIRE(param_start % 4 == 0 && N % 4 == 0, "gemm_tc_wgrad_adam: parameter range not 16-byte aligned");
  AdamEpilogue ep;
  memset(&ep, 0, sizeof ep);
  char* st = (char*)b200_comm_seg_ptr(comm, state_seg, as_rank);
  char* ct = (char*)b200_comm_seg_ptr(comm, ctrl_seg, as_rank);
  B200_REQUIRE(st && ct, "gemm_tc_wgrad_adam: state/ctrl segment not mapped");
  ep.master = (float*)(st + master_off); ep.m = (float*)(st + m_off); ep.v = (float*)(st + v_off);
  for (int r = 0; r < nranks; ++r) {
    char* w = (char*)b200_comm_seg_ptr(comm, w16_seg, r);
    B200_REQUIRE(w != nullptr, "gemm_tc_wgrad_adam: weight segment %d is not mapped for rank %d", w16_seg, r);
    ep.w16[r] = (bf16*)(w + w16_off);
  }
  ep.nranks = nranks;
  ep.param_base = param_start + (long long)row0 * N;
  ep.step_ptr = (const int*)(ct + step_off);
  ep.hyper = (const double*)(ct + hyper_off);
  ep.gscale = gscale;
  ep.coef = b200_comm_coef_cache(comm, as_rank);
  // TMA tensor stores need 16-byte aligned rows: parameter index of the slice's first element a multiple of 8, N too
  static int tma_ok = -1;
  if (tma_ok < 0) {
    const char* e = getenv("B200DP_TMA_PUSH");
    tma_ok = (e && e[0] == '0') ? 0 : 1;
  }
  TmapPack maps;
  const bool push = tma_ok && nranks > 1 && ep.param_base % 8 == 0 && N % 8 == 0;   // alone, direct stores are 1.3 us faster
  if (push) {
    memset(&maps, 0, sizeof maps);
    for (int r = 0; r < nranks; ++r) {
      const uint64_t dims[2] = {(uint64_t)N, (uint64_t)M}, str[1] = {(uint64_t)N * 2};
      const uint32_t box[2] = {(uint32_t)block_n, 128};
      int rc = tc::make_tmap_bf16(&maps.m[r], ep.w16[r] + ep.param_base, 2, dims, str, box, 0);
      if (rc) return rc;
    }
  }
  return gemm_tc_launch(A, B, ep.master, nullptr, 0, M, N, K, lda, ldb, N, 1, 1, 4, 1, block_n, stream, &ep, push ? &maps : nullptr);
}
