// Peer-memory collectives over NVLink 5 / NVSwitch for sm_100a.
//
// What this replaces: Horovod's background thread, negotiation,
// fusion buffer and its MPI/NCCL allreduce/broadcast ops, pulled in by the reference at
// horovod/Dockerfile:64-65 and driven from horovod/tensorflow_mnist.py:133-143.
//
// Design: every rank owns "symmetric segments" (same size, same layout on every rank) whose
// base pointers are mapped into every peer (CUDA IPC).  A collective is ONE kernel that
//   1. crosses a flag barrier held in peer memory (st.release.sys / ld.acquire.sys, epochs kept
//      in device memory so a CUDA-graph replay needs no host-side state),
//   2. reads the peers' gradient buckets directly over NVLink, sums them in fixed rank order
//      (bitwise-identical result on every rank),
//   3. applies the 1/size scale and the Adam update to the fp32 master weights and writes the
//      bf16 working copy — one-shot: every rank updates everything; two-shot: rank r reduces and
//      updates slice r only and pushes the updated bf16 slice into every peer.
// No NCCL, no separate scale/cast/optimizer kernels.
//
// "Emulated" communicators place all N ranks' segments on ONE GPU and run the N ranks as
// blockIdx.y slices of one cooperative launch; this is how the protocol is tested on a single
// B200 (ranks that wait on one another must never be separate launches on one GPU).
#include <cooperative_groups.h>
#include <stdarg.h>
#include <stdlib.h>
#include <string.h>

#include <mutex>
#include <string>
#include <vector>

#include "../include/b200dp.h"
#include "common.cuh"
#include "tc_sm100.cuh"

namespace {

constexpr int MAXR = B200_MAX_RANKS;
constexpr int MAXCH = B200_MAX_CHANNELS;
constexpr int MAXB = B200_MAX_COMM_BLOCKS;
constexpr size_t PAD_FLAGS_BYTES = (size_t)MAXCH * MAXB * MAXR * sizeof(uint32_t);
constexpr size_t PAD_EPOCH_BYTES = (size_t)MAXCH * MAXB * sizeof(uint32_t);
constexpr size_t PAD_BYTES = PAD_FLAGS_BYTES + PAD_EPOCH_BYTES;

thread_local std::string g_err;

struct Seg {
  void* base[MAXR] = {};
  size_t bytes = 0;
  bool opened[MAXR] = {};
  bool vmm = false;           // VMM/multicast-backed: mappings are released with the process, not with cudaFree
};

struct Comm {
  int rank = 0, n = 1, device = 0;
  bool emulated = false;
  std::vector<Seg> segs;
  unsigned long long timeout_cycles = 40ull * 1000 * 1000 * 1000;   // ~20 s at 2 GHz
  uint32_t skew_ns = 0;
  int poison = 0;             // debug: consumed gradient buffers are filled with NaN instead of zeros (b200_comm_set_debug_poison)
  void* emu_args = nullptr;   // device scratch for per-rank argument blocks (emulated only)
  // Step-result mirror (b200_comm_set_step_mirror): the fused kernel of `mirror_channel` — the last kernel of a training
  // step — copies that step's loss / correct-count ring slot into pinned host memory, so the host reads a step's result
  // without a device-to-host copy operation in the stream between two steps.
  // Push inbox (b200_comm_set_inbox): one-shot launches on `inbox_channel` do not pull the peers' gradients after a
  // barrier (two NVLink round trips) — every rank stores its slice into each peer's inbox and then raises the flag, so
  // the data rides in front of the flag and the wait costs one one-way latency.
  const float* coef_cache[MAXR] = {};   // per rank (emulated communicators hold all of them); null: kernels compute the coefficients
  int inbox_channel = -1, inbox_seg = -1;
  size_t inbox_bytes_per_parity = 0;
  int mirror_channel = -1, mirror_ring = 0;
  const float* mirror_loss = nullptr;
  const int* mirror_correct = nullptr;
  float* mirror_host_loss = nullptr;
  int* mirror_host_correct = nullptr;
};

struct PadPtrs {
  uint32_t* flags[MAXR];   // every rank's flag array (peer-mapped)
  uint32_t* epochs;        // this rank's epoch counters
  uint32_t skew_ns;        // debug: pseudo-random delay (< skew_ns) before every barrier arrival, to provoke races
  uint32_t trace;          // debug (B200DP_COMM_TRACE=1): block 0 prints the phase times of every 256th launch
};

// -------------------------------------------------------------------------------------------
// Flag barrier between the same-index block of every rank.
// -------------------------------------------------------------------------------------------
__device__ __forceinline__ void rank_barrier(const PadPtrs& pad, int channel, int bid, int rank, int n,
                                             uint32_t epoch, unsigned long long timeout) {
  __syncthreads();
  if (pad.skew_ns) {       // race-hunting mode: ranks/blocks reach the barrier at scattered times
    uint32_t h = (uint32_t)rank * 0x9E3779B9u ^ (uint32_t)bid * 0x85EBCA6Bu ^ epoch * 0xC2B2AE35u;
    h ^= h >> 15; h *= 0x2C1B3C6Du; h ^= h >> 12;
    __nanosleep(h % pad.skew_ns);
    __syncthreads();
  }
  if ((int)threadIdx.x < n) {
    const int peer = threadIdx.x;
    const size_t slot = ((size_t)channel * MAXB + bid) * MAXR;
    st_release_sys(pad.flags[peer] + slot + rank, epoch);
    const uint32_t* mine = pad.flags[rank] + slot + peer;
    const long long t0 = clock64();
    while ((int32_t)(ld_acquire_sys(mine) - epoch) < 0) {
      if ((unsigned long long)(clock64() - t0) > timeout) {
        printf("b200dp: collective timeout: rank %d channel %d block %d never saw epoch %u from rank %d (has %u)\n",
               rank, channel, bid, epoch, peer, ld_acquire_sys(mine));
        __trap();
      }
    }
  }
  __syncthreads();
}

__device__ __forceinline__ void bulk_g2s(void* dst_smem, const void* src, uint32_t bytes, uint64_t* bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(tc::smem_u32(dst_smem)),
               "l"(src), "r"(bytes), "r"(tc::smem_u32(bar))
               : "memory");
}
__device__ __forceinline__ void bulk_s2g(void* dst, const void* src_smem, uint32_t bytes) {
  asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(dst), "r"(tc::smem_u32(src_smem)), "r"(bytes) : "memory");
}

struct Range {
  long long start;    // elements
  long long numel;    // multiple of 4
  int algo;           // 0 one-shot, 1 two-shot, 3 two-shot on a gradient the producer already reduce-scattered
  int zero_other;     // clear the other-parity gradient buffer over this range
};

struct FusedArgs {
  int rank, n, channel;
  int reduce;                 // 0: gradients are already final locally (single rank / after Adasum)
  PadPtrs pad;
  const float* grad[MAXR];    // current-parity gradient buffers of every rank
  float* grad_other;          // local other-parity buffer
  float* master;              // local fp32 weights / Adam moments
  float* m;
  float* v;
  bf16* w16[MAXR];            // bf16 working weights of every rank
  const int* step_ptr;        // device: 1-based step being applied
  const double* hyper;        // device: lr, beta1, beta2, eps, eps_mode
  float scale;                // 1/size for Average, 1 for Sum
  int nranges;
  Range ranges[B200_MAX_RANGES];
  unsigned long long timeout;
  const float* mir_loss;      // step-result mirror (all null when unused): device rings -> mapped pinned host rings
  const int* mir_correct;
  float* mir_host_loss;
  int* mir_host_correct;
  int mir_ring;
  int poison;                 // debug: "clear the other-parity buffer" writes NaN (the engine zeroes it again before use)
  const float* coef_cache;    // Adam coefficients of this step, precomputed by the step's first kernel (else null: computed here)
  int bulk;                   // WIDE kernel, one-shot ranges only: peers' slices arrive by TMA bulk copies into shared memory
  int push;                   // WIDE kernel, one-shot ranges only: slices are pushed into the peers' inboxes ahead of the flag
  float* inbox[MAXR];         // every rank's inbox for this parity: [source rank][compact range index]
  long long inbox_stride;     // floats per source rank (sum of the range lengths)
};

struct AdamCoef {
  float b1, b2, eps, step_size, vscale;
};

__device__ __forceinline__ float adam_elem(float g, float& p, float& m, float& v, const AdamCoef& c) {
  m = c.b1 * m + (1.f - c.b1) * g;
  v = c.b2 * v + (1.f - c.b2) * g * g;
  p -= c.step_size * m / (sqrtf(v) * c.vscale + c.eps);
  return p;
}

// Peer loads are issued four at a time before their adds (4 independent 16-byte NVLink reads in flight per thread
// without blowing the register budget); the sum runs in fixed rank order so every rank computes the same bits.
__device__ __forceinline__ float4 reduce_ptrs(const float* const* ptrs, int n, long long i) {
  float4 s = make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
  for (int r0 = 0; r0 < MAXR; r0 += 4) {
    if (r0 < n) {
      float4 t[4];
#pragma unroll
      for (int k = 0; k < 4; ++k)
        if (r0 + k < n) t[k] = ld_f4(ptrs[r0 + k] + i);
#pragma unroll
      for (int k = 0; k < 4; ++k)
        if (r0 + k < n) {
          if (r0 + k == 0) s = t[k];
          else { s.x += t[k].x; s.y += t[k].y; s.z += t[k].z; s.w += t[k].w; }
        }
    }
  }
  return s;
}

// Latency form for the small buckets: all peers' loads are in flight at once (one NVLink round trip instead of two at
// N=8); the sum still runs in rank order.  Needs ~32 more registers than the 64 the 512-thread kernel may use.
__device__ __forceinline__ float4 reduce_ptrs_wide(const float* const* ptrs, int n, long long i) {
  float4 t[MAXR];
#pragma unroll
  for (int k = 0; k < MAXR; ++k)
    if (k < n) t[k] = ld_f4(ptrs[k] + i);
  float4 s = t[0];
#pragma unroll
  for (int k = 1; k < MAXR; ++k)
    if (k < n) { s.x += t[k].x; s.y += t[k].y; s.z += t[k].z; s.w += t[k].w; }
  return s;
}

template <bool WIDE>
__device__ __forceinline__ float4 reduce_peers(const FusedArgs& a, long long i) {
  float4 s = WIDE ? reduce_ptrs_wide(a.grad, a.reduce ? a.n : 1, i) : reduce_ptrs(a.grad, a.reduce ? a.n : 1, i);
  s.x *= a.scale; s.y *= a.scale; s.z *= a.scale; s.w *= a.scale;
  return s;
}

// WIDE: blocks of at most 256 threads (the small, latency-bound buckets); otherwise 512 threads x 2 blocks per SM.
template <bool EMU, bool WIDE>
__global__ void __launch_bounds__(WIDE ? 256 : 512, 2) fused_allreduce_adam_kernel(const FusedArgs areal, const FusedArgs* aemu) {
  const FusedArgs& a = EMU ? aemu[blockIdx.y] : areal;
  const int bid = blockIdx.x;
  const long long tid = (long long)bid * blockDim.x + threadIdx.x;
  const long long nthr = (long long)gridDim.x * blockDim.x;
  uint32_t epoch = a.pad.epochs[a.channel * MAXB + bid];
  const bool multi = a.n > 1 && a.reduce;

  __shared__ AdamCoef coef;
  if (threadIdx.x == 0 && a.coef_cache != nullptr) {
    coef.b1 = a.coef_cache[0]; coef.b2 = a.coef_cache[1]; coef.eps = a.coef_cache[2];
    coef.step_size = a.coef_cache[3]; coef.vscale = a.coef_cache[4];
  } else if (threadIdx.x == 0) {
    const int t = *a.step_ptr;
    const double lr = a.hyper[0], b1 = a.hyper[1], b2 = a.hyper[2], eps = a.hyper[3];
    const double bc1 = 1.0 - pow(b1, (double)t), bc2 = 1.0 - pow(b2, (double)t);
    coef.b1 = (float)b1;
    coef.b2 = (float)b2;
    coef.eps = (float)eps;
    if (a.hyper[4] == 0.0) {            // TF placement: lr_t * m / (sqrt(v) + eps)
      coef.step_size = (float)(lr * sqrt(bc2) / bc1);
      coef.vscale = 1.f;
    } else {                            // torch placement: lr/bc1 * m / (sqrt(v)/sqrt(bc2) + eps)
      coef.step_size = (float)(lr / bc1);
      coef.vscale = (float)(1.0 / sqrt(bc2));
    }
  }
  if (WIDE && a.push && multi) {
    // ---- push-based one-shot ------------------------------------------------------------------------------------
    // 1. store this rank's slice of every range into each peer's inbox (row = this rank), 2. fence + flag exchange
    // (the barrier doubles as "data has arrived"), 3. sum own gradient + the N-1 inbox rows in rank order, Adam.
    const int me = a.rank;
    const bool tr = a.pad.trace && bid == 0 && threadIdx.x == 0 && (epoch & 255u) == 0;
    unsigned long long t0 = 0, t1 = 0, t2 = 0, t3 = 0;
    if (tr) asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t0));
    long long cbase = 0;
    for (int ri = 0; ri < a.nranges; ++ri) {
      const Range rg = a.ranges[ri];
      const long long nq = rg.numel / 4, per = (nq + gridDim.x - 1) / gridDim.x;
      const long long q0 = min(nq, (long long)bid * per), q1 = min(nq, q0 + per);
      for (long long q = q0 + threadIdx.x; q < q1; q += blockDim.x) {
        const float4 g = ld_f4(a.grad[me] + rg.start + q * 4);
        // the fp32 state of these parameters was last touched a whole step ago (150 MB of traffic since): start
        // bringing it into L2 now, underneath the flag exchange
        asm volatile("prefetch.global.L2 [%0];" ::"l"(a.master + rg.start + q * 4));
        asm volatile("prefetch.global.L2 [%0];" ::"l"(a.m + rg.start + q * 4));
        asm volatile("prefetch.global.L2 [%0];" ::"l"(a.v + rg.start + q * 4));
        float* dst = nullptr;
#pragma unroll 1
        for (int r = 0; r < a.n; ++r)
          if (r != me) { dst = a.inbox[r] + (long long)me * a.inbox_stride + cbase + q * 4; st_f4(dst, g); }
      }
      cbase += rg.numel;
    }
    if (tr) asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t1));
    // no per-thread fence: the barrier's __syncthreads orders every thread's inbox stores before the flag threads, and
    // their st.release.sys is cumulative over everything that happened before it
    rank_barrier(a.pad, a.channel, bid, a.rank, a.n, ++epoch, a.timeout);
    if (tr) asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t2));
    const AdamCoef c = coef;
    const float clr = a.poison ? __int_as_float(0x7fc00000) : 0.f;
    cbase = 0;
    for (int ri = 0; ri < a.nranges; ++ri) {
      const Range rg = a.ranges[ri];
      const long long nq = rg.numel / 4, per = (nq + gridDim.x - 1) / gridDim.x;
      const long long q0 = min(nq, (long long)bid * per), q1 = min(nq, q0 + per);
      for (long long q = q0 + threadIdx.x; q < q1; q += blockDim.x) {
        const long long i = rg.start + q * 4;
        float4 p = ld_f4(a.master + i), m = ld_f4(a.m + i), v = ld_f4(a.v + i);
        float4 t[MAXR];
#pragma unroll
        for (int r = 0; r < MAXR; ++r)
          if (r < a.n) t[r] = r == me ? ld_f4(a.grad[me] + i) : ld_f4(a.inbox[me] + (long long)r * a.inbox_stride + cbase + q * 4);
        float4 g = t[0];
#pragma unroll
        for (int r = 1; r < MAXR; ++r)
          if (r < a.n) { g.x += t[r].x; g.y += t[r].y; g.z += t[r].z; g.w += t[r].w; }
        g.x *= a.scale; g.y *= a.scale; g.z *= a.scale; g.w *= a.scale;
        adam_elem(g.x, p.x, m.x, v.x, c); adam_elem(g.y, p.y, m.y, v.y, c);
        adam_elem(g.z, p.z, m.z, v.z, c); adam_elem(g.w, p.w, m.w, v.w, c);
        st_f4(a.master + i, p); st_f4(a.m + i, m); st_f4(a.v + i, v);
        *reinterpret_cast<uint2*>(a.w16[me] + i) = pack_bf16x4(p);
        if (rg.zero_other) st_f4(a.grad_other + i, make_float4(clr, clr, clr, clr));
      }
      cbase += rg.numel;
    }
    if (threadIdx.x == 0) a.pad.epochs[a.channel * MAXB + bid] = epoch;
    if (a.mir_host_loss != nullptr && bid == 0 && threadIdx.x == 0) {
      const int slot = (int)((uint32_t)*a.step_ptr % (uint32_t)a.mir_ring);
      a.mir_host_loss[slot] = a.mir_loss[slot];
      a.mir_host_correct[slot] = a.mir_correct[slot];
    }
    if (tr) {
      asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t3));
      printf("b200dp trace: rank %d ch %d epoch %u: push %llu ns, barrier %llu ns, reduce+adam %llu ns\n", a.rank, a.channel, epoch,
             t1 - t0, t2 - t1, t3 - t2);
    }
    return;
  }
  if (WIDE && a.bulk && multi && threadIdx.x == 0) {
    extern __shared__ __align__(128) uint8_t fsm0[];
    tc::mbar_init(reinterpret_cast<uint64_t*>(fsm0), 1);
    tc::fence_barrier_init();
  }
  if (multi) rank_barrier(a.pad, a.channel, bid, a.rank, a.n, ++epoch, a.timeout);
  else __syncthreads();
  const AdamCoef c = coef;
  // After the barrier above every peer has finished reading the other-parity buffer, so it may be cleared for the next
  // accumulation.  Poison mode writes NaN instead: a reader that is late after all (a protocol bug) then poisons its
  // weights and the loss instead of quietly summing zeros; the engine zeroes the buffer itself before it is used again.
  const float clr = a.poison ? __int_as_float(0x7fc00000) : 0.f;

  if (WIDE && a.bulk && multi) {
    // One-shot over the copy engine: every block owns a contiguous run of 16-byte columns of each range; one thread
    // pulls that run from all N ranks into shared memory with TMA bulk copies (one NVLink round trip for everything,
    // no long-latency loads parked in the SM's load queue in front of a co-resident kernel's), then the block sums
    // the N copies in rank order and applies Adam.
    extern __shared__ __align__(128) uint8_t fsm[];
    uint64_t* bar = reinterpret_cast<uint64_t*>(fsm);
    float4* buf = reinterpret_cast<float4*>(fsm + 128);
    // (the barrier was initialised before the rank barrier, see below)
    if (threadIdx.x == 0) {
      uint32_t total = 0;
      for (int ri = 0; ri < a.nranges; ++ri) {
        const long long nq = a.ranges[ri].numel / 4, per = (nq + gridDim.x - 1) / gridDim.x;
        const long long q0 = min(nq, (long long)bid * per), q1 = min(nq, q0 + per);
        total += (uint32_t)((q1 - q0) * 16 * a.n);
      }
      if (total) {
        tc::mbar_arrive_expect_tx(bar, total);
        float4* dst = buf;
        for (int ri = 0; ri < a.nranges; ++ri) {
          const long long nq = a.ranges[ri].numel / 4, per = (nq + gridDim.x - 1) / gridDim.x;
          const long long q0 = min(nq, (long long)bid * per), q1 = min(nq, q0 + per);
          if (q1 > q0)
            for (int r = 0; r < a.n; ++r) {
              bulk_g2s(dst, a.grad[r] + a.ranges[ri].start + q0 * 4, (uint32_t)((q1 - q0) * 16), bar);
              dst += q1 - q0;
            }
        }
      } else {
        tc::mbar_arrive(bar);
      }
    }
    tc::mbar_wait(bar, 0, 41);
    const float4* src = buf;
    for (int ri = 0; ri < a.nranges; ++ri) {
      const Range rg = a.ranges[ri];
      const long long nq = rg.numel / 4, per = (nq + gridDim.x - 1) / gridDim.x;
      const long long q0 = min(nq, (long long)bid * per), q1 = min(nq, q0 + per);
      const int cnt = (int)(q1 - q0);
      for (int j = threadIdx.x; j < cnt; j += blockDim.x) {
        const long long i = rg.start + (q0 + j) * 4;
        float4 p = ld_f4(a.master + i), m = ld_f4(a.m + i), v = ld_f4(a.v + i);
        float4 g = src[j];
        for (int r = 1; r < a.n; ++r) {
          const float4 t = src[(size_t)r * cnt + j];
          g.x += t.x; g.y += t.y; g.z += t.z; g.w += t.w;
        }
        g.x *= a.scale; g.y *= a.scale; g.z *= a.scale; g.w *= a.scale;
        adam_elem(g.x, p.x, m.x, v.x, c); adam_elem(g.y, p.y, m.y, v.y, c);
        adam_elem(g.z, p.z, m.z, v.z, c); adam_elem(g.w, p.w, m.w, v.w, c);
        st_f4(a.master + i, p); st_f4(a.m + i, m); st_f4(a.v + i, v);
        *reinterpret_cast<uint2*>(a.w16[a.rank] + i) = pack_bf16x4(p);
        if (rg.zero_other) st_f4(a.grad_other + i, make_float4(clr, clr, clr, clr));
      }
      src += (size_t)a.n * cnt;
    }
    if (threadIdx.x == 0) a.pad.epochs[a.channel * MAXB + bid] = epoch;
    if (a.mir_host_loss != nullptr && bid == 0 && threadIdx.x == 0) {
      const int slot = (int)((uint32_t)*a.step_ptr % (uint32_t)a.mir_ring);
      a.mir_host_loss[slot] = a.mir_loss[slot];
      a.mir_host_correct[slot] = a.mir_correct[slot];
    }
    return;
  }

  bool any_twoshot = false;
  for (int ri = 0; ri < a.nranges; ++ri) {
    const Range rg = a.ranges[ri];
    if (rg.algo == 0 || !multi) {
      // one-shot: every rank reduces and updates the whole range; all loads of an element (N peer gradients,
      // master, m, v) are issued before the first use
      const long long nq = rg.numel / 4;
      for (long long q = tid; q < nq; q += nthr) {
        const long long i = rg.start + q * 4;
        float4 p = ld_f4(a.master + i), m = ld_f4(a.m + i), v = ld_f4(a.v + i);
        const float4 g = reduce_peers<WIDE>(a, i);
        adam_elem(g.x, p.x, m.x, v.x, c); adam_elem(g.y, p.y, m.y, v.y, c);
        adam_elem(g.z, p.z, m.z, v.z, c); adam_elem(g.w, p.w, m.w, v.w, c);
        st_f4(a.master + i, p); st_f4(a.m + i, m); st_f4(a.v + i, v);
        *reinterpret_cast<uint2*>(a.w16[a.rank] + i) = pack_bf16x4(p);
        if (rg.zero_other) st_f4(a.grad_other + i, make_float4(clr, clr, clr, clr));
      }
    } else {
      // two-shot: reduce + update the owned slice, push the bf16 result to every rank
      any_twoshot = true;
      long long chunk = (rg.numel + a.n - 1) / a.n;
      chunk = (chunk + 3) / 4 * 4;
      const long long lo = min((long long)a.rank * chunk, rg.numel);
      const long long hi = min(lo + chunk, rg.numel);
      const bool pre_reduced = rg.algo == 3;   // the producing GEMM already reduce-scattered into the owner (red.add over NVLink)
      for (long long q = tid; q < (hi - lo) / 4; q += nthr) {
        const long long i = rg.start + lo + q * 4;
        float4 g;
        if (pre_reduced) {
          g = ld_f4(a.grad[a.rank] + i);
          g.x *= a.scale; g.y *= a.scale; g.z *= a.scale; g.w *= a.scale;
        } else {
          g = reduce_peers<WIDE>(a, i);
        }
        float4 p = ld_f4(a.master + i), m = ld_f4(a.m + i), v = ld_f4(a.v + i);
        adam_elem(g.x, p.x, m.x, v.x, c); adam_elem(g.y, p.y, m.y, v.y, c);
        adam_elem(g.z, p.z, m.z, v.z, c); adam_elem(g.w, p.w, m.w, v.w, c);
        st_f4(a.master + i, p); st_f4(a.m + i, m); st_f4(a.v + i, v);
        const uint2 w = pack_bf16x4(p);
#pragma unroll 1
        for (int r = 0; r < a.n; ++r) *reinterpret_cast<uint2*>(a.w16[r] + i) = w;
      }
      if (rg.zero_other)
        for (long long q = tid; q < rg.numel / 4; q += nthr)
          st_f4(a.grad_other + rg.start + q * 4, make_float4(clr, clr, clr, clr));
    }
  }
  if (any_twoshot) {
    __threadfence_system();
    rank_barrier(a.pad, a.channel, bid, a.rank, a.n, ++epoch, a.timeout);
  }
  if (threadIdx.x == 0) a.pad.epochs[a.channel * MAXB + bid] = epoch;
  if (a.mir_host_loss != nullptr && bid == 0 && threadIdx.x == 0) {
    // the head kernel finished this step's slot long ago (stream order); posted stores over PCIe, visible to the host
    // once this kernel has completed
    const int slot = (int)((uint32_t)*a.step_ptr % (uint32_t)a.mir_ring);
    a.mir_host_loss[slot] = a.mir_loss[slot];
    a.mir_host_correct[slot] = a.mir_correct[slot];
  }
}

// -------------------------------------------------------------------------------------------
// Plain allreduce / broadcast / all-gather on staging buffers (public hvd.allreduce/broadcast).
// -------------------------------------------------------------------------------------------
struct PlainArgs {
  int rank, n, channel, algo, root;
  PadPtrs pad;
  const float* src[MAXR];
  float* dst[MAXR];
  long long numel;            // multiple of 4
  float scale;
  unsigned long long timeout;
};

template <bool EMU>
__global__ void __launch_bounds__(512) allreduce_plain_kernel(const PlainArgs areal, const PlainArgs* aemu) {
  const PlainArgs& a = EMU ? aemu[blockIdx.y] : areal;
  const int bid = blockIdx.x;
  const long long tid = (long long)bid * blockDim.x + threadIdx.x, nthr = (long long)gridDim.x * blockDim.x;
  uint32_t epoch = a.pad.epochs[a.channel * MAXB + bid];
  rank_barrier(a.pad, a.channel, bid, a.rank, a.n, ++epoch, a.timeout);
  if (a.algo == 0) {
    for (long long q = tid; q < a.numel / 4; q += nthr) {
      float4 s = reduce_ptrs(a.src, a.n, q * 4);
      s.x *= a.scale; s.y *= a.scale; s.z *= a.scale; s.w *= a.scale;
      st_f4(a.dst[a.rank] + q * 4, s);
    }
  } else {
    long long chunk = (a.numel + a.n - 1) / a.n;
    chunk = (chunk + 3) / 4 * 4;
    const long long lo = min((long long)a.rank * chunk, a.numel), hi = min(lo + chunk, a.numel);
    for (long long q = tid; q < (hi - lo) / 4; q += nthr) {
      const long long i = lo + q * 4;
      float4 s = reduce_ptrs(a.src, a.n, i);
      s.x *= a.scale; s.y *= a.scale; s.z *= a.scale; s.w *= a.scale;
      for (int r = 0; r < a.n; ++r) st_f4(a.dst[r] + i, s);
    }
    __threadfence_system();
    rank_barrier(a.pad, a.channel, bid, a.rank, a.n, ++epoch, a.timeout);
  }
  if (threadIdx.x == 0) a.pad.epochs[a.channel * MAXB + bid] = epoch;
}

// mode 0: broadcast src[root] -> dst[rank];  mode 1: in-place all-gather of rank-owned slices of dst
template <bool EMU>
__global__ void __launch_bounds__(512) copy_collective_kernel(const PlainArgs areal, const PlainArgs* aemu, int mode) {
  const PlainArgs& a = EMU ? aemu[blockIdx.y] : areal;
  const int bid = blockIdx.x;
  const long long tid = (long long)bid * blockDim.x + threadIdx.x, nthr = (long long)gridDim.x * blockDim.x;
  uint32_t epoch = a.pad.epochs[a.channel * MAXB + bid];
  rank_barrier(a.pad, a.channel, bid, a.rank, a.n, ++epoch, a.timeout);
  if (mode == 0) {
    // Large messages: scatter + all-gather, so the root's outbound link carries ~2S instead of (N-1)S.
    // Phase 1: rank r pulls slice r from the root (the root copies everything locally).
    long long chunk = (a.numel + a.n - 1) / a.n;
    chunk = (chunk + 3) / 4 * 4;
    const bool staged = a.n > 2 && a.numel * 4 >= (1 << 20);
    if (a.rank == a.root) {
      if (a.src[a.root] != a.dst[a.rank])
        for (long long q = tid; q < a.numel / 4; q += nthr) st_f4(a.dst[a.rank] + q * 4, ld_f4(a.src[a.root] + q * 4));
    } else if (!staged) {
      for (long long q = tid; q < a.numel / 4; q += nthr) st_f4(a.dst[a.rank] + q * 4, ld_f4(a.src[a.root] + q * 4));
    } else {
      const long long lo = min((long long)a.rank * chunk, a.numel), hi = min(lo + chunk, a.numel);
      for (long long q = tid; q < (hi - lo) / 4; q += nthr) st_f4(a.dst[a.rank] + lo + q * 4, ld_f4(a.src[a.root] + lo + q * 4));
    }
    if (staged) {
      __threadfence_system();
      rank_barrier(a.pad, a.channel, bid, a.rank, a.n, ++epoch, a.timeout);
      // Phase 2: pull every other slice from the rank that owns it (the root's own slice comes from the root).
      if (a.rank != a.root) {
        for (int r = 0; r < a.n; ++r) {
          if (r == a.rank) continue;
          const long long lo = min((long long)r * chunk, a.numel), hi = min(lo + chunk, a.numel);
          const float* from = r == a.root ? a.src[a.root] : a.dst[r];
          for (long long q = tid; q < (hi - lo) / 4; q += nthr) st_f4(a.dst[a.rank] + lo + q * 4, ld_f4(from + lo + q * 4));
        }
      }
    }
  } else {
    long long chunk = (a.numel + a.n - 1) / a.n;
    chunk = (chunk + 3) / 4 * 4;
    for (int r = 0; r < a.n; ++r) {
      if (r == a.rank) continue;
      const long long lo = min((long long)r * chunk, a.numel), hi = min(lo + chunk, a.numel);
      for (long long q = tid; q < (hi - lo) / 4; q += nthr) {
        const long long i = lo + q * 4;
        st_f4(a.dst[a.rank] + i, ld_f4(a.dst[r] + i));
      }
    }
  }
  rank_barrier(a.pad, a.channel, bid, a.rank, a.n, ++epoch, a.timeout);
  if (threadIdx.x == 0) a.pad.epochs[a.channel * MAXB + bid] = epoch;
}

// -------------------------------------------------------------------------------------------
// All-gather of the dense layer's backward operands (the "allreduce" of the dense weight gradient, turned inside
// out): rank r collects every rank's dz [B, D] and the column slice r of every rank's flattened activations a2 [B, F]
// into local [n*B, D] / [n*B, F/n] buffers.  The wgrad GEMM that follows (b200_gemm_tc_wgrad_adam) contracts over the
// GLOBAL batch, so the cross-rank gradient sum of the rows rank r owns appears in its accumulators without the
// 12.8 MB gradient ever crossing NVLink: 2.5 MB of activations do instead (N=8).  One barrier: a rank arrives when its
// own dz/a2 are final AND it no longer reads the old dense weights (the caller orders the launch after its dgrad).
// -------------------------------------------------------------------------------------------
struct GatherArgs {
  int rank, n, channel;
  PadPtrs pad;
  const bf16* act[MAXR];      // every rank's a2 [B, F]
  const bf16* dz[MAXR];       // every rank's dz [B, D]
  bf16* act_out;              // local [n*B, F/n]
  bf16* dz_out;               // local [n*B, D]
  int B, F, D;
  unsigned long long timeout;
};

__device__ __forceinline__ uint4 ld_u4(const void* p) {
  uint4 v;
  asm volatile("ld.global.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(p));
  return v;
}

template <bool EMU>
__global__ void __launch_bounds__(128, 8) allgather_acts_kernel(const GatherArgs areal, const GatherArgs* aemu) {
  const GatherArgs& a = EMU ? aemu[blockIdx.y] : areal;
  const int bid = blockIdx.x;
  uint32_t epoch = a.pad.epochs[a.channel * MAXB + bid];
  rank_barrier(a.pad, a.channel, bid, a.rank, a.n, ++epoch, a.timeout);
  if (threadIdx.x == 0) a.pad.epochs[a.channel * MAXB + bid] = epoch;
  const int Fs = a.F / a.n;                       // owned feature slice (multiple of 8 elements)
  const int dz_v = a.B * a.D / 8;                 // 16-byte vectors of one rank's dz
  const int fs_v = Fs / 8;                        // 16-byte vectors per row of the slice
  const int act_v = a.B * fs_v;
  const int per_rank = dz_v + act_v;
  const long long total = (long long)per_rank * a.n;
  const long long tid = (long long)bid * blockDim.x + threadIdx.x, nthr = (long long)gridDim.x * blockDim.x;
  // A peer load takes ~3 us over NVLink, so the copy is as fast as the bytes it keeps in flight: eight independent
  // 16-byte loads per thread, 128-thread blocks of <= 64 registers (8 K registers: they fit on an SM beside a
  // convolution CTA), one block per SM -> 2.4 MB in flight, the 2.9 MB gather of the 8-rank MNIST step in ~two round trips.
  constexpr int UG = 8;
  for (long long i0 = tid; i0 < total; i0 += UG * nthr) {
    uint4 v[UG];
#pragma unroll
    for (int u = 0; u < UG; ++u) {
      const long long i = i0 + u * nthr;
      if (i < total) {
        const int pr = (int)(i / per_rank);
        const int p = (pr + a.rank) % a.n;        // start with the own rank, then walk the peers: spreads the link load
        const int j = (int)(i % per_rank);
        if (j < dz_v) {
          v[u] = ld_u4(reinterpret_cast<const uint4*>(a.dz[p]) + j);
        } else {
          const int k = j - dz_v, b = k / fs_v, c = k % fs_v;
          v[u] = ld_u4(reinterpret_cast<const uint4*>(a.act[p] + (size_t)b * a.F + (size_t)a.rank * Fs) + c);
        }
      }
    }
    // destinations are recomputed after the loads were issued: eight live 64-bit pointers would push the kernel past
    // the 64 registers that let its blocks sit beside a convolution CTA
#pragma unroll
    for (int u = 0; u < UG; ++u) {
      const long long i = i0 + u * nthr;
      if (i < total) {
        const int pr = (int)(i / per_rank);
        const int p = (pr + a.rank) % a.n;
        const int j = (int)(i % per_rank);
        uint4* dst = j < dz_v ? reinterpret_cast<uint4*>(a.dz_out) + (size_t)p * dz_v + j
                              : reinterpret_cast<uint4*>(a.act_out) + (size_t)p * act_v + (j - dz_v);
        *dst = v[u];
      }
    }
  }
}

// The same gather on the copy engine of the SM (TMA bulk copies, no multicast): one thread per block moves 16 KB pieces
// peer -> shared memory -> local memory through a ring of four slots with mbarrier completion.  Why not plain loads: a
// peer load stays ~3 us in the SM's load queue, and the loads of whatever convolution CTA shares that SM queue up
// behind it (measured: the conv2 dgrad beside the load-based gather went from 14 to 29 us).  Bulk copies bypass that
// queue.  The ring is two slots (32 KB) on purpose: with 64 KB the blocks could not share an SM with a resident
// convolution CTA (186 KB), took whole SMs away from the 128-CTA dgrad that starts right behind them, and pushed 20 of its
// CTAs into a second wave (measured at N=8: dgrad 13 -> 26 us).  One block per SM x 32 KB = 4.7 MB in flight.
constexpr int GSLOT = 16384, GSLOTS = 2;     // 32 KB per block: small enough to sit on an SM BESIDE a 186 KB convolution CTA

template <bool EMU>
__global__ void __launch_bounds__(32) allgather_acts_bulk_kernel(const GatherArgs areal, const GatherArgs* aemu) {
  extern __shared__ __align__(128) uint8_t gsm[];
  const GatherArgs& a = EMU ? aemu[blockIdx.y] : areal;
  const int bid = blockIdx.x;
  uint32_t epoch = a.pad.epochs[a.channel * MAXB + bid];
  rank_barrier(a.pad, a.channel, bid, a.rank, a.n, ++epoch, a.timeout);
  if (threadIdx.x != 0) return;
  a.pad.epochs[a.channel * MAXB + bid] = epoch;
  uint64_t* full = reinterpret_cast<uint64_t*>(gsm + GSLOTS * GSLOT);
  for (int s = 0; s < GSLOTS; ++s) tc::mbar_init(full + s, 1);
  tc::fence_barrier_init();
  const int Fs = a.F / a.n;
  const int row_b = Fs * 2;                                   // bytes of one row of the owned slice (multiple of 16)
  const long long dz_b = (long long)a.B * a.D * 2;            // bytes of one rank's dz (multiple of 16)
  const int n_dz = (int)((dz_b + GSLOT - 1) / GSLOT);
  const int rg = row_b <= GSLOT ? GSLOT / row_b : 0;          // rows per piece; 0: a row is split into column segments
  const int segs = rg ? 1 : (row_b + GSLOT - 1) / GSLOT;
  const int n_act = rg ? (a.B + rg - 1) / rg : a.B * segs;
  const int per_peer = n_dz + n_act;
  const int total = per_peer * a.n;
  const int cnt = bid < total ? (total - bid + (int)gridDim.x - 1) / (int)gridDim.x : 0;

  // piece k of this block -> loads into slot k % GSLOTS; returns the bytes and where they go
  auto issue = [&](int k, bf16** dst_out, uint32_t* bytes_out, bool do_load) {
    const int i = bid + k * (int)gridDim.x;
    const int pr = i / per_peer, j = i - pr * per_peer;
    const int p = (pr + a.rank) % a.n;                        // own rank first, then the peers in ring order
    uint8_t* slot = gsm + (k % GSLOTS) * GSLOT;
    uint64_t* bar = full + (k % GSLOTS);
    if (j < n_dz) {
      const long long off = (long long)j * GSLOT;
      const uint32_t bytes = (uint32_t)min((long long)GSLOT, dz_b - off);
      if (do_load) { tc::mbar_arrive_expect_tx(bar, bytes); bulk_g2s(slot, (const uint8_t*)a.dz[p] + off, bytes, bar); }
      *dst_out = reinterpret_cast<bf16*>((uint8_t*)a.dz_out + (size_t)p * dz_b + off);
      *bytes_out = bytes;
    } else if (rg) {
      const int b0 = (j - n_dz) * rg, nb = min(rg, a.B - b0);
      if (do_load) {
        tc::mbar_arrive_expect_tx(bar, (uint32_t)(nb * row_b));
        for (int r = 0; r < nb; ++r)
          bulk_g2s(slot + r * row_b, a.act[p] + (size_t)(b0 + r) * a.F + (size_t)a.rank * Fs, (uint32_t)row_b, bar);
      }
      *dst_out = a.act_out + ((size_t)p * a.B + b0) * Fs;
      *bytes_out = (uint32_t)(nb * row_b);
    } else {
      const int q = j - n_dz, b = q / segs, sg = q - b * segs;
      const int c0 = sg * GSLOT;
      const uint32_t bytes = (uint32_t)min(GSLOT, row_b - c0);
      if (do_load) {
        tc::mbar_arrive_expect_tx(bar, bytes);
        bulk_g2s(slot, (const uint8_t*)(a.act[p] + (size_t)b * a.F + (size_t)a.rank * Fs) + c0, bytes, bar);
      }
      *dst_out = reinterpret_cast<bf16*>((uint8_t*)(a.act_out + ((size_t)p * a.B + b) * Fs) + c0);
      *bytes_out = bytes;
    }
  };
  bf16* dst;
  uint32_t bytes;
  for (int k = 0; k < min(cnt, GSLOTS); ++k) issue(k, &dst, &bytes, true);
  for (int k = 0; k < cnt; ++k) {
    tc::mbar_wait(full + (k % GSLOTS), (uint32_t)((k / GSLOTS) & 1), 40);
    issue(k, &dst, &bytes, false);
    bulk_s2g(dst, gsm + (k % GSLOTS) * GSLOT, bytes);
    asm volatile("cp.async.bulk.commit_group;" ::: "memory");
    if (k + GSLOTS < cnt) {
      asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory");     // the slot has been read out: refill it
      issue(k + GSLOTS, &dst, &bytes, true);
    }
  }
  asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");              // all stores have landed before the grid completes
}

// -------------------------------------------------------------------------------------------
// Adasum: log2(N) levels; at level L rank r pairs with r ^ (1<<L).
//   a' = (1 - a.b/(2|a|^2)) a + (1 - a.b/(2|b|^2)) b      per tensor segment
// "a" is always the lower rank's vector so both partners compute bit-identical results.
// Kernel 1 writes per-block partial dot products (fixed block->element mapping), kernel 2 sums
// the partials in block order and combines — deterministic, no atomics.
// -------------------------------------------------------------------------------------------
struct AdasumArgs {
  int rank, n, channel, level;
  PadPtrs pad;
  const float* cur[MAXR];     // level-input vectors of every rank
  float* next;                // local level-output vector
  double* partials;           // local scratch [gridDim.x][nseg][3]
  int nseg;
  long long seg_start[B200_MAX_SEGMENTS];
  long long seg_numel[B200_MAX_SEGMENTS];
  unsigned long long timeout;
};

template <bool EMU>
__global__ void __launch_bounds__(256) adasum_dots_kernel(const AdasumArgs areal, const AdasumArgs* aemu) {
  const AdasumArgs& a = EMU ? aemu[blockIdx.y] : areal;
  const int bid = blockIdx.x;
  uint32_t epoch = a.pad.epochs[a.channel * MAXB + bid];
  rank_barrier(a.pad, a.channel, bid, a.rank, a.n, ++epoch, a.timeout);
  if (threadIdx.x == 0) a.pad.epochs[a.channel * MAXB + bid] = epoch;
  const int partner = a.rank ^ (1 << a.level);
  const float* va = a.cur[min(a.rank, partner)];
  const float* vb = a.cur[max(a.rank, partner)];
  __shared__ double red[3][8];
  for (int s = 0; s < a.nseg; ++s) {
    double dab = 0, daa = 0, dbb = 0;
    const float* pa = va + a.seg_start[s];
    const float* pb = vb + a.seg_start[s];
    for (long long i = (long long)bid * blockDim.x + threadIdx.x; i < a.seg_numel[s]; i += (long long)gridDim.x * blockDim.x) {
      const double x = pa[i], y = pb[i];
      dab += x * y; daa += x * x; dbb += y * y;
    }
    for (int o = 16; o > 0; o >>= 1) {
      dab += __shfl_xor_sync(0xffffffffu, dab, o);
      daa += __shfl_xor_sync(0xffffffffu, daa, o);
      dbb += __shfl_xor_sync(0xffffffffu, dbb, o);
    }
    const int w = threadIdx.x >> 5, l = threadIdx.x & 31;
    if (l == 0) { red[0][w] = dab; red[1][w] = daa; red[2][w] = dbb; }
    __syncthreads();
    if (threadIdx.x < 3) {
      double t = 0;
      for (int k = 0; k < (int)(blockDim.x >> 5); ++k) t += red[threadIdx.x][k];
      a.partials[((size_t)bid * a.nseg + s) * 3 + threadIdx.x] = t;
    }
    __syncthreads();
  }
}

template <bool EMU>
__global__ void __launch_bounds__(256) adasum_combine_kernel(const AdasumArgs areal, const AdasumArgs* aemu, int nblk_dots) {
  const AdasumArgs& a = EMU ? aemu[blockIdx.y] : areal;
  const int partner = a.rank ^ (1 << a.level);
  const float* va = a.cur[min(a.rank, partner)];
  const float* vb = a.cur[max(a.rank, partner)];
  __shared__ float ca_s[B200_MAX_SEGMENTS], cb_s[B200_MAX_SEGMENTS];
  if ((int)threadIdx.x < a.nseg) {
    double dab = 0, daa = 0, dbb = 0;
    for (int b = 0; b < nblk_dots; ++b) {
      const double* p = a.partials + ((size_t)b * a.nseg + threadIdx.x) * 3;
      dab += p[0]; daa += p[1]; dbb += p[2];
    }
    ca_s[threadIdx.x] = (float)(1.0 - (daa > 0 ? dab / (2.0 * daa) : 0.0));
    cb_s[threadIdx.x] = (float)(1.0 - (dbb > 0 ? dab / (2.0 * dbb) : 0.0));
  }
  __syncthreads();
  for (int s = 0; s < a.nseg; ++s) {
    const float ca = ca_s[s], cb = cb_s[s];
    const long long st = a.seg_start[s];
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < a.seg_numel[s]; i += (long long)gridDim.x * blockDim.x)
      a.next[st + i] = ca * va[st + i] + cb * vb[st + i];
  }
}

// -------------------------------------------------------------------------------------------
int check_seg(Comm* c, int seg, size_t off, size_t bytes, const char* what) {
  B200_REQUIRE(seg >= 0 && seg < (int)c->segs.size(), "%s: segment %d does not exist", what, seg);
  B200_REQUIRE(off + bytes <= c->segs[seg].bytes, "%s: [%zu, %zu) outside segment %d of %zu bytes", what, off, off + bytes, seg, c->segs[seg].bytes);
  B200_REQUIRE(off % 16 == 0, "%s: offset %zu not 16-byte aligned", what, off);
  return B200_OK;
}

inline char* seg_ptr(Comm* c, int seg, int peer, size_t off) { return (char*)c->segs[seg].base[peer] + off; }

void fill_pad(Comm* c, int as_rank, PadPtrs* p) {
  for (int r = 0; r < c->n; ++r) p->flags[r] = (uint32_t*)seg_ptr(c, 0, r, 0);
  p->epochs = (uint32_t*)seg_ptr(c, 0, as_rank, PAD_FLAGS_BYTES);
  p->skew_ns = c->skew_ns;
  static int tr = -1;
  if (tr < 0) { const char* e = getenv("B200DP_COMM_TRACE"); tr = (e && e[0] == '1') ? 1 : 0; }
  p->trace = (uint32_t)tr;
}

template <typename Args, typename KReal, typename KEmu, typename... Extra>
int launch_ranked(Comm* c, Args* per_rank, KReal kreal, KEmu kemu, int nblocks, int nthreads, cudaStream_t st, Extra... extra) {
  if (!c->emulated) {
    kreal<<<nblocks, nthreads, 0, st>>>(per_rank[0], nullptr, extra...);
    B200_LAUNCH_CHECK();
    return B200_OK;
  }
  B200_REQUIRE(sizeof(Args) * c->n <= 65536, "argument block too large");
  B200_CHECK(cudaMemcpyAsync(c->emu_args, per_rank, sizeof(Args) * c->n, cudaMemcpyHostToDevice, st));
  const Args* dev = (const Args*)c->emu_args;
  Args dummy = per_rank[0];
  void* params[] = {(void*)&dummy, (void*)&dev, (void*)&extra...};
  B200_CHECK(cudaLaunchCooperativeKernel((void*)kemu, dim3(nblocks, c->n), dim3(nthreads), params, 0, st));
  // per_rank is host stack memory: the async copy above must finish before we return
  B200_CHECK(cudaStreamSynchronize(st));
  return B200_OK;
}

template <typename Args, typename KReal, typename KEmu>
int launch_ranked_smem(Comm* c, Args* per_rank, KReal kreal, KEmu kemu, int nblocks, int nthreads, size_t smem, cudaStream_t st) {
  if (!c->emulated) {
    kreal<<<nblocks, nthreads, smem, st>>>(per_rank[0], nullptr);
    B200_LAUNCH_CHECK();
    return B200_OK;
  }
  B200_REQUIRE(sizeof(Args) * c->n <= 65536, "argument block too large");
  B200_CHECK(cudaMemcpyAsync(c->emu_args, per_rank, sizeof(Args) * c->n, cudaMemcpyHostToDevice, st));
  const Args* dev = (const Args*)c->emu_args;
  Args dummy = per_rank[0];
  void* params[] = {(void*)&dummy, (void*)&dev};
  B200_CHECK(cudaLaunchCooperativeKernel((void*)kemu, dim3(nblocks, c->n), dim3(nthreads), params, smem, st));
  B200_CHECK(cudaStreamSynchronize(st));
  return B200_OK;
}

}  // namespace

void b200_set_error(const char* fmt, ...) {
  char buf[1024];
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(buf, sizeof buf, fmt, ap);
  va_end(ap);
  g_err = buf;
}

void b200_forget_mcsegs(void* comm);   // defined with the NVLS code below

const float* b200_comm_coef_cache(b200_comm_t h, int rank) {
  Comm* c = (Comm*)h;
  return (rank >= 0 && rank < MAXR) ? c->coef_cache[rank] : nullptr;
}

extern "C" {

const char* b200_last_error(void) { return g_err.c_str(); }

size_t b200_comm_pad_bytes(void) { return PAD_BYTES; }

int b200_comm_create(int rank, int nranks, int emulated, b200_comm_t* out) {
  B200_REQUIRE(nranks >= 1 && nranks <= MAXR, "nranks %d outside [1, %d]", nranks, MAXR);
  B200_REQUIRE(rank >= 0 && rank < nranks, "rank %d outside world of %d", rank, nranks);
  Comm* c = new Comm;
  c->rank = emulated ? 0 : rank;
  c->n = nranks;
  c->emulated = emulated != 0;
  B200_CHECK(cudaGetDevice(&c->device));
  if (c->emulated) B200_CHECK(cudaMalloc(&c->emu_args, 65536));
  *out = c;
  return B200_OK;
}

int b200_comm_rank(b200_comm_t h) { return ((Comm*)h)->rank; }
int b200_comm_nranks(b200_comm_t h) { return ((Comm*)h)->n; }
int b200_comm_is_emulated(b200_comm_t h) { return ((Comm*)h)->emulated ? 1 : 0; }

// The fused kernels read the step's Adam coefficients (5 floats) from (ctrl_seg, off) instead of computing them; the
// engine's first kernel of a step writes them (b200_conv1_pool_fwd with bump_step = 2).  ctrl_seg < 0 switches it off.
int b200_comm_set_coef_cache(b200_comm_t h, int ctrl_seg, size_t off) {
  Comm* c = (Comm*)h;
  for (int r = 0; r < MAXR; ++r) c->coef_cache[r] = nullptr;
  if (ctrl_seg < 0) return B200_OK;
  B200_REQUIRE(ctrl_seg < (int)c->segs.size() && off + 20 <= c->segs[ctrl_seg].bytes && off % 4 == 0, "set_coef_cache: bad segment/offset");
  for (int r = 0; r < c->n; ++r)
    if (c->emulated || r == c->rank) c->coef_cache[r] = (const float*)((char*)c->segs[ctrl_seg].base[r] + off);
  return B200_OK;
}

// seg: a symmetric segment of 2 * bytes_per_parity bytes (zero or garbage, it is only read after being written);
// one-shot launches on `channel` whose ranges fit (nranks * sum(numel) * 4 <= bytes_per_parity) then push instead of pull.
int b200_comm_set_inbox(b200_comm_t h, int channel, int seg, size_t bytes_per_parity) {
  Comm* c = (Comm*)h;
  if (channel < 0) { c->inbox_channel = -1; return B200_OK; }
  B200_REQUIRE(channel < MAXCH && seg >= 0 && seg < (int)c->segs.size() && 2 * bytes_per_parity <= c->segs[seg].bytes &&
                   bytes_per_parity % 16 == 0, "set_inbox: bad channel/segment/size");
  c->inbox_channel = channel; c->inbox_seg = seg; c->inbox_bytes_per_parity = bytes_per_parity;
  return B200_OK;
}

// loss/correct rings at (ctrl_seg, loss_off / correct_off), `ring` slots each; host_loss/host_correct: pinned host memory
// the device can address (cudaHostAlloc / torch pin_memory under UVA), `ring` slots each.  channel < 0 switches it off.
int b200_comm_set_step_mirror(b200_comm_t h, int channel, int ctrl_seg, size_t loss_off, size_t correct_off, int ring,
                              float* host_loss, int* host_correct) {
  Comm* c = (Comm*)h;
  if (channel < 0) { c->mirror_channel = -1; return B200_OK; }
  B200_REQUIRE(channel < MAXCH && ring > 0 && host_loss && host_correct, "set_step_mirror: bad arguments");
  B200_REQUIRE(ctrl_seg >= 0 && ctrl_seg < (int)c->segs.size() && loss_off + (size_t)ring * 4 <= c->segs[ctrl_seg].bytes &&
                   correct_off + (size_t)ring * 4 <= c->segs[ctrl_seg].bytes, "set_step_mirror: rings outside segment %d", ctrl_seg);
  const int me = c->emulated ? 0 : c->rank;
  c->mirror_loss = (const float*)((char*)c->segs[ctrl_seg].base[me] + loss_off);
  c->mirror_correct = (const int*)((char*)c->segs[ctrl_seg].base[me] + correct_off);
  c->mirror_host_loss = host_loss; c->mirror_host_correct = host_correct; c->mirror_ring = ring;
  c->mirror_channel = channel;
  return B200_OK;
}

int b200_comm_set_timeout_ms(b200_comm_t h, double ms) {
  Comm* c = (Comm*)h;
  int khz = 0;
  B200_CHECK(cudaDeviceGetAttribute(&khz, cudaDevAttrClockRate, c->device));
  c->timeout_cycles = (unsigned long long)(ms * (double)khz);
  return B200_OK;
}

// Debug aid: gradient buffers a fused kernel has consumed are filled with NaN instead of zeros.
int b200_comm_set_debug_poison(b200_comm_t h, int on) {
  ((Comm*)h)->poison = on ? 1 : 0;
  return B200_OK;
}

// Debug aid: every barrier arrival is preceded by a pseudo-random sleep below max_ns nanoseconds.
int b200_comm_set_debug_skew_ns(b200_comm_t h, unsigned max_ns) {
  ((Comm*)h)->skew_ns = max_ns;
  return B200_OK;
}

// Allocates the local part of a new symmetric segment (zero-filled). Returns its id and, for real
// communicators, the 64-byte CUDA IPC handle to hand to the peers. Segment 0 must be the signal
// pad (b200_comm_pad_bytes()).
int b200_comm_seg_alloc(b200_comm_t h, size_t bytes, int* seg_id, void* ipc_handle_out) {
  Comm* c = (Comm*)h;
  B200_REQUIRE(bytes > 0 && bytes % 256 == 0, "segment size %zu must be a positive multiple of 256", bytes);
  B200_REQUIRE(!c->segs.empty() || bytes >= PAD_BYTES, "segment 0 is the signal pad and needs >= %zu bytes", PAD_BYTES);
  Seg s;
  s.bytes = bytes;
  if (c->emulated) {
    for (int r = 0; r < c->n; ++r) {
      B200_CHECK(cudaMalloc(&s.base[r], bytes));
      B200_CHECK(cudaMemset(s.base[r], 0, bytes));
    }
  } else {
    B200_CHECK(cudaMalloc(&s.base[c->rank], bytes));
    B200_CHECK(cudaMemset(s.base[c->rank], 0, bytes));
    if (ipc_handle_out && c->n > 1) {
      cudaIpcMemHandle_t hd;
      B200_CHECK(cudaIpcGetMemHandle(&hd, s.base[c->rank]));
      memcpy(ipc_handle_out, &hd, sizeof hd);
    }
  }
  B200_CHECK(cudaDeviceSynchronize());
  c->segs.push_back(s);
  *seg_id = (int)c->segs.size() - 1;
  return B200_OK;
}

// all_handles: nranks x 64 bytes, rank order (gathered by the caller over its control plane).
int b200_comm_seg_open(b200_comm_t h, int seg_id, const void* all_handles) {
  Comm* c = (Comm*)h;
  B200_REQUIRE(seg_id >= 0 && seg_id < (int)c->segs.size(), "segment %d does not exist", seg_id);
  if (c->emulated || c->n == 1) return B200_OK;
  Seg& s = c->segs[seg_id];
  for (int r = 0; r < c->n; ++r) {
    if (r == c->rank || s.opened[r]) continue;
    cudaIpcMemHandle_t hd;
    memcpy(&hd, (const char*)all_handles + (size_t)r * sizeof hd, sizeof hd);
    B200_CHECK(cudaIpcOpenMemHandle(&s.base[r], hd, cudaIpcMemLazyEnablePeerAccess));
    s.opened[r] = true;
  }
  return B200_OK;
}

void* b200_comm_seg_ptr(b200_comm_t h, int seg_id, int peer) {
  Comm* c = (Comm*)h;
  if (seg_id < 0 || seg_id >= (int)c->segs.size() || peer < 0 || peer >= c->n) return nullptr;
  return c->segs[seg_id].base[peer];
}

int b200_comm_destroy(b200_comm_t h) {
  Comm* c = (Comm*)h;
  if (!c) return B200_OK;
  cudaDeviceSynchronize();
  for (Seg& s : c->segs)
    for (int r = 0; r < c->n; ++r) {
      if (!s.base[r] || s.vmm) continue;
      if (c->emulated || r == c->rank) cudaFree(s.base[r]);
      else if (s.opened[r]) cudaIpcCloseMemHandle(s.base[r]);
    }
  if (c->emu_args) cudaFree(c->emu_args);
  b200_forget_mcsegs(c);
  delete c;
  return B200_OK;
}

// The fused step: allreduce(grad) -> scale -> Adam(master, m, v) -> bf16 weights, per range.
// All buffers live in symmetric segments and are addressed as (segment, byte offset).
// In an emulated communicator the call runs every rank (each on its own copy of the segments).
int b200_fused_allreduce_adam(b200_comm_t h, int channel, int nblocks, int nthreads, int reduce,
                              int grad_seg, size_t grad_cur_off, size_t grad_other_off,
                              int state_seg, size_t master_off, size_t m_off, size_t v_off,
                              int w16_seg, size_t w16_off, int ctrl_seg, size_t step_off, size_t hyper_off,
                              float scale, int nranges, const long long* range_start, const long long* range_numel,
                              const int* range_algo, const int* range_zero_other, void* stream) {
  Comm* c = (Comm*)h;
  B200_REQUIRE(channel >= 0 && channel < MAXCH, "channel %d outside [0, %d)", channel, MAXCH);
  B200_REQUIRE(nblocks >= 1 && nblocks <= MAXB, "nblocks %d outside [1, %d]", nblocks, MAXB);
  B200_REQUIRE(nthreads >= 32 && nthreads <= 512 && nthreads % 32 == 0, "nthreads %d invalid", nthreads);
  B200_REQUIRE(nranges >= 1 && nranges <= B200_MAX_RANGES, "nranges %d outside [1, %d]", nranges, B200_MAX_RANGES);
  long long max_end = 0;
  for (int i = 0; i < nranges; ++i) {
    B200_REQUIRE(range_start[i] % 4 == 0 && range_numel[i] % 4 == 0 && range_numel[i] >= 0, "range %d not 4-element aligned", i);
    max_end = max_end > range_start[i] + range_numel[i] ? max_end : range_start[i] + range_numel[i];
  }
  int rc;
  if ((rc = check_seg(c, grad_seg, grad_cur_off, max_end * 4, "grad"))) return rc;
  if ((rc = check_seg(c, grad_seg, grad_other_off, max_end * 4, "grad_other"))) return rc;
  if ((rc = check_seg(c, state_seg, master_off, max_end * 4, "master"))) return rc;
  if ((rc = check_seg(c, state_seg, m_off, max_end * 4, "adam_m"))) return rc;
  if ((rc = check_seg(c, state_seg, v_off, max_end * 4, "adam_v"))) return rc;
  if ((rc = check_seg(c, w16_seg, w16_off, max_end * 2, "w16"))) return rc;
  if ((rc = check_seg(c, ctrl_seg, step_off, 16, "step"))) return rc;
  if ((rc = check_seg(c, ctrl_seg, hyper_off, 48, "hyper"))) return rc;

  FusedArgs args[MAXR];
  const int nr = c->emulated ? c->n : 1;
  for (int k = 0; k < nr; ++k) {
    const int me = c->emulated ? k : c->rank;
    FusedArgs& a = args[k];
    memset(&a, 0, sizeof a);
    a.rank = me; a.n = c->n; a.channel = channel; a.reduce = reduce;
    fill_pad(c, me, &a.pad);
    for (int r = 0; r < c->n; ++r) {
      a.grad[r] = (const float*)seg_ptr(c, grad_seg, r, grad_cur_off);
      a.w16[r] = (bf16*)seg_ptr(c, w16_seg, r, w16_off);
    }
    if (!reduce) a.grad[0] = (const float*)seg_ptr(c, grad_seg, me, grad_cur_off);
    a.grad_other = (float*)seg_ptr(c, grad_seg, me, grad_other_off);
    a.master = (float*)seg_ptr(c, state_seg, me, master_off);
    a.m = (float*)seg_ptr(c, state_seg, me, m_off);
    a.v = (float*)seg_ptr(c, state_seg, me, v_off);
    a.step_ptr = (const int*)seg_ptr(c, ctrl_seg, me, step_off);
    a.hyper = (const double*)seg_ptr(c, ctrl_seg, me, hyper_off);
    a.scale = scale;
    a.nranges = nranges;
    for (int i = 0; i < nranges; ++i) a.ranges[i] = Range{range_start[i], range_numel[i], range_algo[i], range_zero_other[i]};
    a.timeout = c->timeout_cycles;
    a.poison = c->poison;
    a.coef_cache = c->coef_cache[me];
    if (channel == c->mirror_channel && !c->emulated) {
      a.mir_loss = c->mirror_loss; a.mir_correct = c->mirror_correct;
      a.mir_host_loss = c->mirror_host_loss; a.mir_host_correct = c->mirror_host_correct; a.mir_ring = c->mirror_ring;
    }
  }
  if (nthreads <= 256) {
    // small buckets: TMA bulk prefetch of the peers' slices when every range is one-shot and the slices fit shared memory
    static int bulk_ok = -1;
    if (bulk_ok < 0) {
      const char* e = getenv("B200DP_ONESHOT_BULK");
      bulk_ok = (e && e[0] == '0') ? 0 : 1;
      if (bulk_ok) {
        B200_CHECK(cudaFuncSetAttribute(fused_allreduce_adam_kernel<false, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, 160 * 1024));
        B200_CHECK(cudaFuncSetAttribute(fused_allreduce_adam_kernel<true, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, 160 * 1024));
      }
    }
    size_t smem = 0;
    bool bulk = bulk_ok && reduce && c->n > 1;
    for (int i = 0; i < nranges && bulk; ++i) {
      if (range_algo[i] != 0) bulk = false;
      smem += (size_t)((range_numel[i] / 4 + nblocks - 1) / nblocks) * 16 * c->n;
    }
    if (bulk && smem + 128 > 160 * 1024) bulk = false;
    // push inbox registered for this channel: data travels in front of the flag (one one-way latency)
    bool push = false;
    if (channel == c->inbox_channel && reduce && c->n > 1) {
      long long ctot = 0;
      push = true;
      for (int i = 0; i < nranges; ++i) {
        if (range_algo[i] != 0) push = false;
        ctot += range_numel[i];
      }
      if ((size_t)ctot * 4 * c->n > c->inbox_bytes_per_parity) push = false;
      if (push) {
        const size_t poff = grad_cur_off > grad_other_off ? c->inbox_bytes_per_parity : 0;
        for (int k = 0; k < nr; ++k) {
          args[k].push = 1;
          args[k].inbox_stride = ctot;
          for (int r = 0; r < c->n; ++r) args[k].inbox[r] = (float*)seg_ptr(c, c->inbox_seg, r, poff);
        }
        bulk = false;
      }
    }
    smem = bulk ? smem + 128 : 0;
    for (int k = 0; k < nr; ++k) args[k].bulk = bulk ? 1 : 0;
    return launch_ranked_smem(c, args, fused_allreduce_adam_kernel<false, true>, fused_allreduce_adam_kernel<true, true>, nblocks,
                              nthreads, smem, (cudaStream_t)stream);
  }
  return launch_ranked(c, args, fused_allreduce_adam_kernel<false, false>, fused_allreduce_adam_kernel<true, false>, nblocks, nthreads,
                       (cudaStream_t)stream);
}

static int plain_args(Comm* c, PlainArgs* args, int channel, int algo, int root, int src_seg, size_t src_off,
                      int dst_seg, size_t dst_off, long long numel, float scale) {
  B200_REQUIRE(channel >= 0 && channel < MAXCH, "channel %d outside [0, %d)", channel, MAXCH);
  B200_REQUIRE(numel >= 0 && numel % 4 == 0, "numel %lld must be a multiple of 4", numel);
  B200_REQUIRE(root >= 0 && root < c->n, "root %d outside world of %d", root, c->n);
  int rc;
  if ((rc = check_seg(c, src_seg, src_off, numel * 4, "src"))) return rc;
  if ((rc = check_seg(c, dst_seg, dst_off, numel * 4, "dst"))) return rc;
  const int nr = c->emulated ? c->n : 1;
  for (int k = 0; k < nr; ++k) {
    const int me = c->emulated ? k : c->rank;
    PlainArgs& a = args[k];
    memset(&a, 0, sizeof a);
    a.rank = me; a.n = c->n; a.channel = channel; a.algo = algo; a.root = root;
    fill_pad(c, me, &a.pad);
    for (int r = 0; r < c->n; ++r) {
      a.src[r] = (const float*)seg_ptr(c, src_seg, r, src_off);
      a.dst[r] = (float*)seg_ptr(c, dst_seg, r, dst_off);
    }
    a.numel = numel;
    a.scale = scale;
    a.timeout = c->timeout_cycles;
  }
  return B200_OK;
}

// dst = scale * sum_r src_r  (fp32). algo 0 one-shot (dst local), 1 two-shot (dst written on every rank).
int b200_allreduce(b200_comm_t h, int channel, int nblocks, int nthreads, int algo, int src_seg, size_t src_off,
                   int dst_seg, size_t dst_off, long long numel, float scale, void* stream) {
  Comm* c = (Comm*)h;
  B200_REQUIRE(nblocks >= 1 && nblocks <= MAXB && nthreads >= 32 && nthreads <= 512, "bad launch shape");
  PlainArgs args[MAXR];
  int rc = plain_args(c, args, channel, algo, 0, src_seg, src_off, dst_seg, dst_off, numel, scale);
  if (rc) return rc;
  return launch_ranked(c, args, allreduce_plain_kernel<false>, allreduce_plain_kernel<true>, nblocks, nthreads, (cudaStream_t)stream);
}

// Every rank copies root's src into its own dst (peer reads over NVLink).
int b200_broadcast(b200_comm_t h, int channel, int nblocks, int nthreads, int root, int src_seg, size_t src_off,
                   int dst_seg, size_t dst_off, long long numel, void* stream) {
  Comm* c = (Comm*)h;
  B200_REQUIRE(nblocks >= 1 && nblocks <= MAXB && nthreads >= 32 && nthreads <= 512, "bad launch shape");
  PlainArgs args[MAXR];
  int rc = plain_args(c, args, channel, 0, root, src_seg, src_off, dst_seg, dst_off, numel, 1.f);
  if (rc) return rc;
  return launch_ranked(c, args, copy_collective_kernel<false>, copy_collective_kernel<true>, nblocks, nthreads, (cudaStream_t)stream, 0);
}

// In-place all-gather: rank r owns slice r (same slicing as the two-shot path) of the buffer.
int b200_allgather_slices(b200_comm_t h, int channel, int nblocks, int nthreads, int seg, size_t off, long long numel, void* stream) {
  Comm* c = (Comm*)h;
  B200_REQUIRE(nblocks >= 1 && nblocks <= MAXB && nthreads >= 32 && nthreads <= 512, "bad launch shape");
  PlainArgs args[MAXR];
  int rc = plain_args(c, args, channel, 0, 0, seg, off, seg, off, numel, 1.f);
  if (rc) return rc;
  return launch_ranked(c, args, copy_collective_kernel<false>, copy_collective_kernel<true>, nblocks, nthreads, (cudaStream_t)stream, 1);
}

// One-sided pull: copy [off, off+bytes) of `peer`'s copy of a segment into the same place of the local copy
// (rank 0 uses it to collect the owner-sharded Adam state when it checkpoints alone).
int b200_comm_pull_from_peer(b200_comm_t h, int seg, size_t off, size_t bytes, int peer, void* stream) {
  Comm* c = (Comm*)h;
  int rc;
  if ((rc = check_seg(c, seg, off, bytes, "pull_from_peer"))) return rc;
  B200_REQUIRE(peer >= 0 && peer < c->n, "peer %d outside world of %d", peer, c->n);
  const int me = c->emulated ? 0 : c->rank;
  if (peer == me || bytes == 0) return B200_OK;
  B200_CHECK(cudaMemcpyAsync(seg_ptr(c, seg, me, off), seg_ptr(c, seg, peer, off), bytes, cudaMemcpyDefault, (cudaStream_t)stream));
  return B200_OK;
}

// act/dz: (segment, offset) of every rank's a2 [B,F] / dz [B,D] (bf16); act_out/dz_out: (segment, offset) of the local
// gathered [n*B, F/n] / [n*B, D] buffers.  F must be divisible by 8*nranks, B*D by 8.
int b200_allgather_acts(b200_comm_t h, int channel, int nblocks, int act_seg, size_t act_off, int dz_seg, size_t dz_off,
                        int out_seg, size_t act_out_off, size_t dz_out_off, int B, int F, int D, void* stream) {
  Comm* c = (Comm*)h;
  B200_REQUIRE(channel >= 0 && channel < MAXCH, "channel %d outside [0, %d)", channel, MAXCH);
  B200_REQUIRE(nblocks >= 1 && nblocks <= MAXB, "nblocks %d outside [1, %d]", nblocks, MAXB);
  B200_REQUIRE(B > 0 && F % (8 * c->n) == 0 && (B * D) % 8 == 0, "allgather_acts: F=%d must be a multiple of 8*%d, B*D of 8", F, c->n);
  int rc;
  if ((rc = check_seg(c, act_seg, act_off, (size_t)B * F * 2, "act"))) return rc;
  if ((rc = check_seg(c, dz_seg, dz_off, (size_t)B * D * 2, "dz"))) return rc;
  if ((rc = check_seg(c, out_seg, act_out_off, (size_t)B * F * 2, "act_out"))) return rc;
  if ((rc = check_seg(c, out_seg, dz_out_off, (size_t)c->n * B * D * 2, "dz_out"))) return rc;
  GatherArgs args[MAXR];
  const int nr = c->emulated ? c->n : 1;
  for (int k = 0; k < nr; ++k) {
    const int me = c->emulated ? k : c->rank;
    GatherArgs& a = args[k];
    memset(&a, 0, sizeof a);
    a.rank = me; a.n = c->n; a.channel = channel;
    fill_pad(c, me, &a.pad);
    for (int r = 0; r < c->n; ++r) {
      a.act[r] = (const bf16*)seg_ptr(c, act_seg, r, act_off);
      a.dz[r] = (const bf16*)seg_ptr(c, dz_seg, r, dz_off);
    }
    a.act_out = (bf16*)seg_ptr(c, out_seg, me, act_out_off);
    a.dz_out = (bf16*)seg_ptr(c, out_seg, me, dz_out_off);
    a.B = B; a.F = F; a.D = D;
    a.timeout = c->timeout_cycles;
  }
  static int bulk = -1;
  if (bulk < 0) {
    const char* e = getenv("B200DP_GATHER_BULK");
    bulk = (e && e[0] == '0') ? 0 : 1;
    if (bulk) {
      const size_t sm = (size_t)GSLOTS * GSLOT + 64;
      B200_CHECK(cudaFuncSetAttribute(allgather_acts_bulk_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sm));
      B200_CHECK(cudaFuncSetAttribute(allgather_acts_bulk_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sm));
    }
  }
  if (!bulk)
    return launch_ranked(c, args, allgather_acts_kernel<false>, allgather_acts_kernel<true>, nblocks, 128, (cudaStream_t)stream);
  return launch_ranked_smem(c, args, allgather_acts_bulk_kernel<false>, allgather_acts_bulk_kernel<true>, nblocks < 148 ? nblocks : 148, 32,
                            (size_t)GSLOTS * GSLOT + 64, (cudaStream_t)stream);
}

size_t b200_adasum_scratch_bytes(int nblocks, int nseg) { return (size_t)nblocks * nseg * 3 * sizeof(double); }

// One Adasum level: next = combine(cur[lower], cur[higher]) for partner = rank ^ (1 << level).
// cur/next are (segment, offset) fp32 vectors; scratch is local memory of
// b200_adasum_scratch_bytes(nblocks, nseg) bytes per rank, addressed the same way.
int b200_adasum_level(b200_comm_t h, int channel, int nblocks, int level, int cur_seg, size_t cur_off, int next_seg,
                      size_t next_off, int scratch_seg, size_t scratch_off, int nseg, const long long* seg_start,
                      const long long* seg_numel, void* stream) {
  Comm* c = (Comm*)h;
  B200_REQUIRE(channel >= 0 && channel < MAXCH, "channel %d outside [0, %d)", channel, MAXCH);
  B200_REQUIRE(nblocks >= 1 && nblocks <= MAXB, "nblocks %d outside [1, %d]", nblocks, MAXB);
  B200_REQUIRE((c->n & (c->n - 1)) == 0, "Adasum needs a power-of-two world size, got %d", c->n);
  B200_REQUIRE(level >= 0 && (1 << level) < c->n, "level %d invalid for %d ranks", level, c->n);
  B200_REQUIRE(nseg >= 1 && nseg <= B200_MAX_SEGMENTS, "nseg %d outside [1, %d]", nseg, B200_MAX_SEGMENTS);
  long long max_end = 0;
  for (int i = 0; i < nseg; ++i) max_end = max_end > seg_start[i] + seg_numel[i] ? max_end : seg_start[i] + seg_numel[i];
  int rc;
  if ((rc = check_seg(c, cur_seg, cur_off, max_end * 4, "adasum cur"))) return rc;
  if ((rc = check_seg(c, next_seg, next_off, max_end * 4, "adasum next"))) return rc;
  if ((rc = check_seg(c, scratch_seg, scratch_off, b200_adasum_scratch_bytes(nblocks, nseg), "adasum scratch"))) return rc;
  AdasumArgs args[MAXR];
  const int nr = c->emulated ? c->n : 1;
  for (int k = 0; k < nr; ++k) {
    const int me = c->emulated ? k : c->rank;
    AdasumArgs& a = args[k];
    memset(&a, 0, sizeof a);
    a.rank = me; a.n = c->n; a.channel = channel; a.level = level;
    fill_pad(c, me, &a.pad);
    for (int r = 0; r < c->n; ++r) a.cur[r] = (const float*)seg_ptr(c, cur_seg, r, cur_off);
    a.next = (float*)seg_ptr(c, next_seg, me, next_off);
    a.partials = (double*)seg_ptr(c, scratch_seg, me, scratch_off);
    a.nseg = nseg;
    for (int i = 0; i < nseg; ++i) { a.seg_start[i] = seg_start[i]; a.seg_numel[i] = seg_numel[i]; }
    a.timeout = c->timeout_cycles;
  }
  cudaStream_t st = (cudaStream_t)stream;
  rc = launch_ranked(c, args, adasum_dots_kernel<false>, adasum_dots_kernel<true>, nblocks, 256, st);
  if (rc) return rc;
  if (!c->emulated) {
    adasum_combine_kernel<false><<<nblocks * 2, 256, 0, st>>>(args[0], nullptr, nblocks);
    B200_LAUNCH_CHECK();
    return B200_OK;
  }
  // emulated: ranks do not wait on each other inside the combine kernel, so a plain launch is fine
  adasum_combine_kernel<true><<<dim3(nblocks * 2, c->n), 256, 0, st>>>(args[0], (const AdasumArgs*)c->emu_args, nblocks);
  B200_LAUNCH_CHECK();
  return B200_OK;
}

}  // extern "C"

// =====================================================================================================
// b200_adasum_pair_static — the one entry point of this library that allocates nothing.
//
// It computes the same pairwise Adasum combination as one level of b200_adasum_level
//     out = (1 - a.b / (2|a|^2)) a + (1 - a.b / (2|b|^2)) b        per tensor segment
// for two fp32 vectors the caller already holds (plain device pointers, or peer pointers it mapped itself).
// The per-block partial dot products, which b200_adasum_level keeps in caller-provided segment memory, live
// here in a fixed array compiled into the library image.
//
// Who needs this: code that runs INSIDE somebody else's memory regime and therefore must not, or cannot, call
// cudaMalloc — a gradient hook plugged into a framework whose caching allocator already owns the whole
// device (a fresh cudaMalloc fails there, or succeeds only by fragmenting the pool), a serving process that
// carved up its memory budget at start-up, or a latency-sensitive loop that cannot afford the device-wide
// synchronisation cudaMalloc/cudaFree imply.  Such callers cannot create a communicator (its segments are
// cudaMalloc'ed) just to obtain a few kilobytes of scratch.
//
// Cost and limits: the scratch array is B200_ADASUM_STATIC_MAX_BLOCKS x B200_MAX_SEGMENTS x 3 doubles = 48 KiB
// of device memory on EVERY GPU that loads this library, held until the process exits, used or not.
// Requests beyond those maxima are rejected with an error, never truncated.
//
// Sharing: all calls on a device use the same array.  It keeps whatever the previous call left in it; that is
// harmless because the first kernel overwrites every slot the second kernel reads (nothing has to start at
// zero), but it also means two calls whose kernels overlap on the GPU (different streams, or different host
// threads without ordering) race on the scratch and produce garbage.  Serialise them on one stream.
//
// Set-up: the array's name is only a host-side handle; the device address is looked up with
// cudaGetSymbolAddress once per device, on the first call made while that device is current, and cached.
// Make that first call during start-up, before any stream capture is begun.
// =====================================================================================================
namespace {

constexpr int STATIC_MAX_BLOCKS = B200_ADASUM_STATIC_MAX_BLOCKS;
constexpr int STATIC_SCRATCH_DOUBLES = STATIC_MAX_BLOCKS * B200_MAX_SEGMENTS * 3;

static __device__ __align__(256) double g_adasum_static_scratch[STATIC_SCRATCH_DOUBLES];

struct PairArgs {
  const float* a;
  const float* b;
  float* out;
  double* partials;           // device address of g_adasum_static_scratch
  int nseg;
  long long seg_start[B200_MAX_SEGMENTS];
  long long seg_numel[B200_MAX_SEGMENTS];
};

// Same block/thread -> element mapping and the same reduction order as adasum_dots_kernel, so the result
// matches b200_adasum_level bit for bit; no barrier, because both vectors are already complete.
__global__ void __launch_bounds__(256) adasum_pair_dots_kernel(const PairArgs p) {
  const int bid = blockIdx.x;
  __shared__ double red[3][8];
  for (int s = 0; s < p.nseg; ++s) {
    double dab = 0, daa = 0, dbb = 0;
    const float* pa = p.a + p.seg_start[s];
    const float* pb = p.b + p.seg_start[s];
    for (long long i = (long long)bid * blockDim.x + threadIdx.x; i < p.seg_numel[s]; i += (long long)gridDim.x * blockDim.x) {
      const double x = pa[i], y = pb[i];
      dab += x * y; daa += x * x; dbb += y * y;
    }
    for (int o = 16; o > 0; o >>= 1) {
      dab += __shfl_xor_sync(0xffffffffu, dab, o);
      daa += __shfl_xor_sync(0xffffffffu, daa, o);
      dbb += __shfl_xor_sync(0xffffffffu, dbb, o);
    }
    const int w = threadIdx.x >> 5, l = threadIdx.x & 31;
    if (l == 0) { red[0][w] = dab; red[1][w] = daa; red[2][w] = dbb; }
    __syncthreads();
    if (threadIdx.x < 3) {
      double t = 0;
      for (int k = 0; k < (int)(blockDim.x >> 5); ++k) t += red[threadIdx.x][k];
      p.partials[((size_t)bid * p.nseg + s) * 3 + threadIdx.x] = t;     // every slot read later is written here
    }
    __syncthreads();
  }
}

__global__ void __launch_bounds__(256) adasum_pair_combine_kernel(const PairArgs p, int nblk_dots) {
  __shared__ float ca_s[B200_MAX_SEGMENTS], cb_s[B200_MAX_SEGMENTS];
  if ((int)threadIdx.x < p.nseg) {
    double dab = 0, daa = 0, dbb = 0;
    for (int b = 0; b < nblk_dots; ++b) {
      const double* q = p.partials + ((size_t)b * p.nseg + threadIdx.x) * 3;
      dab += q[0]; daa += q[1]; dbb += q[2];
    }
    ca_s[threadIdx.x] = (float)(1.0 - (daa > 0 ? dab / (2.0 * daa) : 0.0));
    cb_s[threadIdx.x] = (float)(1.0 - (dbb > 0 ? dab / (2.0 * dbb) : 0.0));
  }
  __syncthreads();
  for (int s = 0; s < p.nseg; ++s) {
    const float ca = ca_s[s], cb = cb_s[s];
    const long long st = p.seg_start[s];
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < p.seg_numel[s]; i += (long long)gridDim.x * blockDim.x)
      p.out[st + i] = ca * p.a[st + i] + cb * p.b[st + i];
  }
}

constexpr int STATIC_MAX_DEVICES = 64;
double* g_static_scratch_addr[STATIC_MAX_DEVICES] = {};
std::mutex g_static_scratch_mu;

// Device address of the static scratch on the CURRENT device (looked up once per device).
int static_scratch_for_current_device(double** out) {
  int dev = -1;
  B200_CHECK(cudaGetDevice(&dev));
  B200_REQUIRE(dev >= 0 && dev < STATIC_MAX_DEVICES, "adasum_pair_static: device ordinal %d not supported (max %d)", dev, STATIC_MAX_DEVICES);
  std::lock_guard<std::mutex> lk(g_static_scratch_mu);
  if (g_static_scratch_addr[dev] == nullptr) {
    double* ptr = nullptr;
    // the symbol name is a host-side stand-in; only this call yields an address a kernel may dereference
    B200_CHECK(cudaGetSymbolAddress((void**)&ptr, g_adasum_static_scratch));
    B200_REQUIRE(ptr != nullptr, "adasum_pair_static: cudaGetSymbolAddress returned a null address on device %d", dev);
    g_static_scratch_addr[dev] = ptr;
  }
  *out = g_static_scratch_addr[dev];
  return B200_OK;
}

}  // namespace

extern "C" int b200_adasum_pair_static(const float* a, const float* b, float* out, int nblocks, int nseg,
                                       const long long* seg_start, const long long* seg_numel, void* stream) {
  B200_REQUIRE(a != nullptr && b != nullptr && out != nullptr, "adasum_pair_static: null vector pointer");
  B200_REQUIRE(nblocks >= 1 && nblocks <= STATIC_MAX_BLOCKS,
               "adasum_pair_static: nblocks %d outside [1, %d] — the built-in scratch holds partials for at most %d blocks", nblocks,
               STATIC_MAX_BLOCKS, STATIC_MAX_BLOCKS);
  B200_REQUIRE(nseg >= 1 && nseg <= B200_MAX_SEGMENTS,
               "adasum_pair_static: nseg %d outside [1, %d] — the built-in scratch holds at most %d tensor segments per call", nseg,
               B200_MAX_SEGMENTS, B200_MAX_SEGMENTS);
  B200_REQUIRE(seg_start != nullptr && seg_numel != nullptr, "adasum_pair_static: null segment table");
  PairArgs p;
  memset(&p, 0, sizeof p);
  p.a = a; p.b = b; p.out = out; p.nseg = nseg;
  for (int i = 0; i < nseg; ++i) {
    B200_REQUIRE(seg_start[i] >= 0 && seg_numel[i] >= 0, "adasum_pair_static: segment %d has a negative start or length", i);
    p.seg_start[i] = seg_start[i];
    p.seg_numel[i] = seg_numel[i];
  }
  int rc = static_scratch_for_current_device(&p.partials);
  if (rc) return rc;
  cudaStream_t st = (cudaStream_t)stream;
  // A call is two kernels that hand the partial sums to each other through the shared scratch.  Host threads calling
  // into ONE stream must not interleave them (dots A, dots B, combine A would combine with B's sums), so the pair is
  // enqueued under a lock; the stream then keeps whole calls in order.  Calls on different streams can still overlap on
  // the GPU — that remains the caller's to avoid.
  static std::mutex launch_mu;
  std::lock_guard<std::mutex> lk(launch_mu);
  adasum_pair_dots_kernel<<<nblocks, 256, 0, st>>>(p);
  B200_LAUNCH_CHECK();
  adasum_pair_combine_kernel<<<nblocks * 2, 256, 0, st>>>(p, nblocks);
  B200_LAUNCH_CHECK();
  return B200_OK;
}

// =====================================================================================================
// NVLS: symmetric segments backed by the CUDA VMM API and bound to an NVSwitch multicast object, so a kernel
// can reduce IN the switch (multimem.ld_reduce) and broadcast THROUGH it (multimem.st).  SURVEY §5.8 / W23.
// A multicast segment is also peer-mapped like an IPC segment, so every P2P kernel above works on it unchanged.
//   alloc  : cuMemCreate(POSIX fd) + local map; rank 0 also cuMulticastCreate.   -> fds go to the peers (SCM_RIGHTS,
//            done by the Python control plane)
//   open   : import peers' memory + the multicast object, cuMulticastAddDevice
//   (control-plane barrier: every device must be added before any memory is bound)
//   bind   : cuMulticastBindMem(local memory) + map the multicast object
// =====================================================================================================
#include <cuda.h>
#include <unistd.h>

namespace {

struct DriverApi {
  CUresult (*MemCreate)(CUmemGenericAllocationHandle*, size_t, const CUmemAllocationProp*, unsigned long long);
  CUresult (*MemRelease)(CUmemGenericAllocationHandle);
  CUresult (*MemAddressReserve)(CUdeviceptr*, size_t, size_t, CUdeviceptr, unsigned long long);
  CUresult (*MemAddressFree)(CUdeviceptr, size_t);
  CUresult (*MemMap)(CUdeviceptr, size_t, size_t, CUmemGenericAllocationHandle, unsigned long long);
  CUresult (*MemUnmap)(CUdeviceptr, size_t);
  CUresult (*MemSetAccess)(CUdeviceptr, size_t, const CUmemAccessDesc*, size_t);
  CUresult (*MemExportToShareableHandle)(void*, CUmemGenericAllocationHandle, CUmemAllocationHandleType, unsigned long long);
  CUresult (*MemImportFromShareableHandle)(CUmemGenericAllocationHandle*, void*, CUmemAllocationHandleType);
  CUresult (*MemGetAllocationGranularity)(size_t*, const CUmemAllocationProp*, CUmemAllocationGranularity_flags);
  CUresult (*MulticastCreate)(CUmemGenericAllocationHandle*, const CUmulticastObjectProp*);
  CUresult (*MulticastAddDevice)(CUmemGenericAllocationHandle, CUdevice);
  CUresult (*MulticastBindMem)(CUmemGenericAllocationHandle, size_t, CUmemGenericAllocationHandle, size_t, size_t, unsigned long long);
  CUresult (*MulticastGetGranularity)(size_t*, const CUmulticastObjectProp*, CUmulticastGranularity_flags);
  CUresult (*DeviceGet)(CUdevice*, int);
  CUresult (*DeviceGetAttribute)(int*, CUdevice_attribute, CUdevice);
  bool ok = false;
};

template <typename T>
bool drv_sym(const char* name, T* out) {
  void* p = nullptr;
  cudaDriverEntryPointQueryResult q;
  if (cudaGetDriverEntryPoint(name, &p, cudaEnableDefault, &q) != cudaSuccess || q != cudaDriverEntryPointSuccess || !p) return false;
  *out = (T)p;
  return true;
}

DriverApi* drv() {
  static DriverApi api;
  static std::once_flag once;
  std::call_once(once, [] {
    bool ok = true;
    ok &= drv_sym("cuMemCreate", &api.MemCreate);
    ok &= drv_sym("cuMemRelease", &api.MemRelease);
    ok &= drv_sym("cuMemAddressReserve", &api.MemAddressReserve);
    ok &= drv_sym("cuMemAddressFree", &api.MemAddressFree);
    ok &= drv_sym("cuMemMap", &api.MemMap);
    ok &= drv_sym("cuMemUnmap", &api.MemUnmap);
    ok &= drv_sym("cuMemSetAccess", &api.MemSetAccess);
    ok &= drv_sym("cuMemExportToShareableHandle", &api.MemExportToShareableHandle);
    ok &= drv_sym("cuMemImportFromShareableHandle", &api.MemImportFromShareableHandle);
    ok &= drv_sym("cuMemGetAllocationGranularity", &api.MemGetAllocationGranularity);
    ok &= drv_sym("cuMulticastCreate", &api.MulticastCreate);
    ok &= drv_sym("cuMulticastAddDevice", &api.MulticastAddDevice);
    ok &= drv_sym("cuMulticastBindMem", &api.MulticastBindMem);
    ok &= drv_sym("cuMulticastGetGranularity", &api.MulticastGetGranularity);
    ok &= drv_sym("cuDeviceGet", &api.DeviceGet);
    ok &= drv_sym("cuDeviceGetAttribute", &api.DeviceGetAttribute);
    api.ok = ok;
  });
  return &api;
}

#define B200_DRV(expr)                                                                                  \
  do {                                                                                                  \
    CUresult _r = (expr);                                                                               \
    if (_r != CUDA_SUCCESS) {                                                                           \
      b200_set_error("%s:%d: %s -> CUresult %d", __FILE__, __LINE__, #expr, (int)_r);                   \
      return B200_ERR_CUDA;                                                                             \
    }                                                                                                   \
  } while (0)

struct McSeg {
  int seg_id = -1;
  size_t size = 0;
  CUmemGenericAllocationHandle mem[MAXR] = {};
  CUmemGenericAllocationHandle mc = 0;
  void* mc_base = nullptr;
};
std::vector<std::pair<Comm*, McSeg>> g_mcsegs;
std::mutex g_mc_mu;

}  // namespace
void b200_forget_mcsegs(void* comm) {
  std::lock_guard<std::mutex> lk(g_mc_mu);
  for (size_t i = g_mcsegs.size(); i-- > 0;)
    if ((void*)g_mcsegs[i].first == comm) g_mcsegs.erase(g_mcsegs.begin() + i);
}
namespace {

McSeg* find_mc(Comm* c, int seg) {
  for (auto& e : g_mcsegs)
    if (e.first == c && e.second.seg_id == seg) return &e.second;
  return nullptr;
}

int map_rw(DriverApi* d, CUmemGenericAllocationHandle hnd, size_t size, size_t gran, int dev, void** out) {
  CUdeviceptr va = 0;
  B200_DRV(d->MemAddressReserve(&va, size, gran, 0, 0));
  B200_DRV(d->MemMap(va, size, 0, hnd, 0));
  CUmemAccessDesc acc;
  memset(&acc, 0, sizeof acc);
  acc.location.type = CU_MEM_LOCATION_TYPE_DEVICE;
  acc.location.id = dev;
  acc.flags = CU_MEM_ACCESS_FLAGS_PROT_READWRITE;
  B200_DRV(d->MemSetAccess(va, size, &acc, 1));
  *out = (void*)va;
  return B200_OK;
}

// ---- kernels -------------------------------------------------------------------------------------
__device__ __forceinline__ float4 multimem_ld_reduce_f4(const float* mc) {
  float4 v;
  asm volatile("multimem.ld_reduce.relaxed.sys.global.add.v4.f32 {%0,%1,%2,%3}, [%4];"
               : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "l"(mc) : "memory");
  return v;
}
__device__ __forceinline__ void multimem_st_f4(float* mc, float4 v) {
  asm volatile("multimem.st.relaxed.sys.global.v4.f32 [%0], {%1,%2,%3,%4};" ::"l"(mc), "f"(v.x), "f"(v.y), "f"(v.z), "f"(v.w) : "memory");
}
__device__ __forceinline__ void multimem_st_b64(void* mc, uint2 v) {
  const unsigned long long x = ((unsigned long long)v.y << 32) | v.x;
  asm volatile("multimem.st.relaxed.sys.global.b64 [%0], %1;" ::"l"(mc), "l"(x) : "memory");
}

struct NvlsArgs {
  int rank, n, channel, mode;      // mode 0: plain allreduce (dst = scale*sum), 1: fused Adam
  PadPtrs pad;
  const float* src_mc;             // multicast view of the gradient / source buffer
  float* dst_mc;                   // mode 0: multicast view of the result buffer
  bf16* w16_mc;                    // mode 1: multicast view of the bf16 weights
  float* grad_other;               // mode 1: local other-parity gradient buffer (cleared)
  float* master; float* m; float* v;
  const int* step_ptr; const double* hyper;
  float scale;
  long long start, numel;
  int zero_other;
  unsigned long long timeout;
};

// Rank r owns slice r: the switch pulls that slice from every GPU and returns one sum (multimem.ld_reduce);
// the owner applies scale (+Adam) and the switch replicates the store to every GPU (multimem.st).
__global__ void __launch_bounds__(512) nvls_allreduce_kernel(const NvlsArgs a) {
  const int bid = blockIdx.x;
  const long long tid = (long long)bid * blockDim.x + threadIdx.x, nthr = (long long)gridDim.x * blockDim.x;
  uint32_t epoch = a.pad.epochs[a.channel * MAXB + bid];
  __shared__ AdamCoef coef;
  if (a.mode == 1 && threadIdx.x == 0) {
    const int t = *a.step_ptr;
    const double lr = a.hyper[0], b1 = a.hyper[1], b2 = a.hyper[2], eps = a.hyper[3];
    const double bc1 = 1.0 - pow(b1, (double)t), bc2 = 1.0 - pow(b2, (double)t);
    coef.b1 = (float)b1; coef.b2 = (float)b2; coef.eps = (float)eps;
    if (a.hyper[4] == 0.0) { coef.step_size = (float)(lr * sqrt(bc2) / bc1); coef.vscale = 1.f; }
    else { coef.step_size = (float)(lr / bc1); coef.vscale = (float)(1.0 / sqrt(bc2)); }
  }
  rank_barrier(a.pad, a.channel, bid, a.rank, a.n, ++epoch, a.timeout);
  const AdamCoef c = coef;
  long long chunk = (a.numel + a.n - 1) / a.n;
  chunk = (chunk + 3) / 4 * 4;
  const long long lo = min((long long)a.rank * chunk, a.numel), hi = min(lo + chunk, a.numel);
  // four in-switch reductions in flight per thread: the round trip through the NVSwitch is several microseconds
  const long long nq = (hi - lo) / 4;
  for (long long q0 = tid; q0 < nq; q0 += 4 * nthr) {
    float4 g[4];
#pragma unroll
    for (int u = 0; u < 4; ++u) {
      const long long q = q0 + u * nthr;
      if (q < nq) g[u] = multimem_ld_reduce_f4(a.src_mc + a.start + lo + q * 4);
    }
#pragma unroll
    for (int u = 0; u < 4; ++u) {
      const long long q = q0 + u * nthr;
      if (q >= nq) continue;
      const long long i = a.start + lo + q * 4;
      float4 gg = g[u];
      gg.x *= a.scale; gg.y *= a.scale; gg.z *= a.scale; gg.w *= a.scale;
      if (a.mode == 0) {
        multimem_st_f4(a.dst_mc + i, gg);
      } else {
        float4 p = ld_f4(a.master + i), m = ld_f4(a.m + i), v = ld_f4(a.v + i);
        adam_elem(gg.x, p.x, m.x, v.x, c); adam_elem(gg.y, p.y, m.y, v.y, c);
        adam_elem(gg.z, p.z, m.z, v.z, c); adam_elem(gg.w, p.w, m.w, v.w, c);
        st_f4(a.master + i, p); st_f4(a.m + i, m); st_f4(a.v + i, v);
        multimem_st_b64(a.w16_mc + i, pack_bf16x4(p));
      }
    }
  }
  if (a.mode == 1 && a.zero_other)
    for (long long q = tid; q < a.numel / 4; q += nthr) st_f4(a.grad_other + a.start + q * 4, make_float4(0.f, 0.f, 0.f, 0.f));
  __threadfence_system();
  rank_barrier(a.pad, a.channel, bid, a.rank, a.n, ++epoch, a.timeout);
  if (threadIdx.x == 0) a.pad.epochs[a.channel * MAXB + bid] = epoch;
}

struct NvlsBcastArgs {
  int rank, n, channel, root;
  PadPtrs pad;
  const float* src;       // root's local source
  float* dst_mc;          // multicast view of the destination
  long long numel;
  unsigned long long timeout;
};

// Broadcast through the switch: the root stores once to the multicast address, the NVSwitch replicates the
// store into every GPU's copy of the segment.  The root's link carries S bytes, not (N-1)*S.
__global__ void __launch_bounds__(512) nvls_broadcast_kernel(const NvlsBcastArgs a) {
  const int bid = blockIdx.x;
  const long long tid = (long long)bid * blockDim.x + threadIdx.x, nthr = (long long)gridDim.x * blockDim.x;
  uint32_t epoch = a.pad.epochs[a.channel * MAXB + bid];
  rank_barrier(a.pad, a.channel, bid, a.rank, a.n, ++epoch, a.timeout);     // destination free, source ready
  if (a.rank == a.root) {
    const long long nq = a.numel / 4;
    for (long long q0 = tid; q0 < nq; q0 += 4 * nthr) {
      float4 v[4];
#pragma unroll
      for (int u = 0; u < 4; ++u)
        if (q0 + u * nthr < nq) v[u] = ld_f4(a.src + (q0 + u * nthr) * 4);
#pragma unroll
      for (int u = 0; u < 4; ++u)
        if (q0 + u * nthr < nq) multimem_st_f4(a.dst_mc + (q0 + u * nthr) * 4, v[u]);
    }
    __threadfence_system();
  }
  rank_barrier(a.pad, a.channel, bid, a.rank, a.n, ++epoch, a.timeout);     // data has landed everywhere
  if (threadIdx.x == 0) a.pad.epochs[a.channel * MAXB + bid] = epoch;
}

}  // namespace

extern "C" {

// 1 when this device can join an NVSwitch multicast team (NVLS), else 0.
int b200_nvls_supported(void) {
  DriverApi* d = drv();
  if (!d->ok) return 0;
  int dev = 0, v = 0;
  if (cudaGetDevice(&dev) != cudaSuccess) return 0;
  cudaFree(0);
  CUdevice cd;
  if (d->DeviceGet(&cd, dev) != CUDA_SUCCESS) return 0;
  if (d->DeviceGetAttribute(&v, CU_DEVICE_ATTRIBUTE_MULTICAST_SUPPORTED, cd) != CUDA_SUCCESS) return 0;
  return v != 0;
}

// Step 1 (every rank): allocate the local part. mem_fd_out receives a POSIX fd of the local memory; on rank 0
// mc_fd_out receives the fd of the freshly created multicast object (-1 elsewhere). The caller ships the fds.
int b200_comm_mcseg_alloc(b200_comm_t h, size_t bytes, int* seg_id, int* mem_fd_out, int* mc_fd_out, size_t* actual_bytes) {
  Comm* c = (Comm*)h;
  DriverApi* d = drv();
  B200_REQUIRE(d->ok, "driver VMM/multicast entry points are not available");
  B200_REQUIRE(!c->emulated && c->n > 1, "multicast segments need a real multi-GPU communicator");
  B200_REQUIRE(!c->segs.empty(), "allocate the signal pad (segment 0) first");
  B200_CHECK(cudaFree(0));
  CUmulticastObjectProp mp;
  memset(&mp, 0, sizeof mp);
  mp.numDevices = (unsigned)c->n;
  mp.handleTypes = CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR;
  mp.size = bytes;
  size_t gran = 0;
  B200_DRV(d->MulticastGetGranularity(&gran, &mp, CU_MULTICAST_GRANULARITY_RECOMMENDED));
  CUmemAllocationProp ap;
  memset(&ap, 0, sizeof ap);
  ap.type = CU_MEM_ALLOCATION_TYPE_PINNED;
  ap.location.type = CU_MEM_LOCATION_TYPE_DEVICE;
  ap.location.id = c->device;
  ap.requestedHandleTypes = CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR;
  size_t mgran = 0;
  B200_DRV(d->MemGetAllocationGranularity(&mgran, &ap, CU_MEM_ALLOC_GRANULARITY_RECOMMENDED));
  if (mgran > gran) gran = mgran;
  const size_t size = (bytes + gran - 1) / gran * gran;
  mp.size = size;
  McSeg ms;
  ms.size = size;
  B200_DRV(d->MemCreate(&ms.mem[c->rank], size, &ap, 0));
  Seg s;
  s.bytes = size;
  s.vmm = true;
  int rc = map_rw(d, ms.mem[c->rank], size, gran, c->device, &s.base[c->rank]);
  if (rc) return rc;
  B200_CHECK(cudaMemset(s.base[c->rank], 0, size));
  B200_CHECK(cudaDeviceSynchronize());
  int fd = -1;
  B200_DRV(d->MemExportToShareableHandle(&fd, ms.mem[c->rank], CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR, 0));
  *mem_fd_out = fd;
  *mc_fd_out = -1;
  if (c->rank == 0) {
    B200_DRV(d->MulticastCreate(&ms.mc, &mp));
    int mfd = -1;
    B200_DRV(d->MemExportToShareableHandle(&mfd, ms.mc, CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR, 0));
    *mc_fd_out = mfd;
  }
  c->segs.push_back(s);
  ms.seg_id = (int)c->segs.size() - 1;
  *seg_id = ms.seg_id;
  if (actual_bytes) *actual_bytes = size;
  std::lock_guard<std::mutex> lk(g_mc_mu);
  g_mcsegs.emplace_back(c, ms);
  return B200_OK;
}

// Step 2: mem_fds[r] = fd of rank r's memory as received in THIS process (own entry ignored); mc_fd = fd of the
// multicast object as received here (ignored on rank 0). Imports, peer-maps, and adds this device to the team.
int b200_comm_mcseg_open(b200_comm_t h, int seg_id, const int* mem_fds, int mc_fd) {
  Comm* c = (Comm*)h;
  DriverApi* d = drv();
  McSeg* ms = find_mc(c, seg_id);
  B200_REQUIRE(ms != nullptr, "segment %d is not a multicast segment", seg_id);
  Seg& s = c->segs[seg_id];
  CUmemAllocationProp ap;
  memset(&ap, 0, sizeof ap);
  ap.type = CU_MEM_ALLOCATION_TYPE_PINNED;
  ap.location.type = CU_MEM_LOCATION_TYPE_DEVICE;
  ap.location.id = c->device;
  size_t gran = 0;
  B200_DRV(d->MemGetAllocationGranularity(&gran, &ap, CU_MEM_ALLOC_GRANULARITY_RECOMMENDED));
  for (int r = 0; r < c->n; ++r) {
    if (r == c->rank) continue;
    B200_DRV(d->MemImportFromShareableHandle(&ms->mem[r], (void*)(uintptr_t)mem_fds[r], CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR));
    int rc = map_rw(d, ms->mem[r], ms->size, gran, c->device, &s.base[r]);
    if (rc) return rc;
    close(mem_fds[r]);
  }
  if (c->rank != 0) {
    B200_DRV(d->MemImportFromShareableHandle(&ms->mc, (void*)(uintptr_t)mc_fd, CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR));
    close(mc_fd);
  }
  CUdevice cd;
  B200_DRV(d->DeviceGet(&cd, c->device));
  B200_DRV(d->MulticastAddDevice(ms->mc, cd));
  return B200_OK;
}

// Step 3 (after a control-plane barrier: every rank has added its device): bind the local memory and map the
// multicast object.
int b200_comm_mcseg_bind(b200_comm_t h, int seg_id) {
  Comm* c = (Comm*)h;
  DriverApi* d = drv();
  McSeg* ms = find_mc(c, seg_id);
  B200_REQUIRE(ms != nullptr, "segment %d is not a multicast segment", seg_id);
  B200_DRV(d->MulticastBindMem(ms->mc, 0, ms->mem[c->rank], 0, ms->size, 0));
  CUmulticastObjectProp mp;
  memset(&mp, 0, sizeof mp);
  mp.numDevices = (unsigned)c->n;
  mp.handleTypes = CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR;
  mp.size = ms->size;
  size_t gran = 0;
  B200_DRV(d->MulticastGetGranularity(&gran, &mp, CU_MULTICAST_GRANULARITY_RECOMMENDED));
  return map_rw(d, ms->mc, ms->size, gran, c->device, &ms->mc_base);
}

void* b200_comm_seg_mc_ptr(b200_comm_t h, int seg_id) {
  McSeg* ms = find_mc((Comm*)h, seg_id);
  return ms ? ms->mc_base : nullptr;
}

// dst = scale * sum over ranks of src, reduced in the NVSwitch. src/dst live in multicast segments.
int b200_allreduce_nvls(b200_comm_t h, int channel, int nblocks, int nthreads, int src_seg, size_t src_off, int dst_seg,
                        size_t dst_off, long long numel, float scale, void* stream) {
  Comm* c = (Comm*)h;
  McSeg *ss = find_mc(c, src_seg), *ds = find_mc(c, dst_seg);
  B200_REQUIRE(ss && ds && ss->mc_base && ds->mc_base, "allreduce_nvls: source and destination must be bound multicast segments");
  B200_REQUIRE(channel >= 0 && channel < MAXCH && nblocks >= 1 && nblocks <= MAXB && nthreads >= 32 && nthreads <= 512, "bad launch shape");
  B200_REQUIRE(numel % 4 == 0, "numel %lld must be a multiple of 4", numel);
  int rc;
  if ((rc = check_seg(c, src_seg, src_off, numel * 4, "nvls src"))) return rc;
  if ((rc = check_seg(c, dst_seg, dst_off, numel * 4, "nvls dst"))) return rc;
  NvlsArgs a;
  memset(&a, 0, sizeof a);
  a.rank = c->rank; a.n = c->n; a.channel = channel; a.mode = 0;
  fill_pad(c, c->rank, &a.pad);
  a.src_mc = (const float*)((char*)ss->mc_base + src_off);
  a.dst_mc = (float*)((char*)ds->mc_base + dst_off);
  a.scale = scale; a.start = 0; a.numel = numel; a.timeout = c->timeout_cycles;
  nvls_allreduce_kernel<<<nblocks, nthreads, 0, (cudaStream_t)stream>>>(a);
  B200_LAUNCH_CHECK();
  return B200_OK;
}

// Broadcast root's src into every rank's dst with one multicast store stream (dst must be a bound multicast segment).
int b200_broadcast_nvls(b200_comm_t h, int channel, int nblocks, int nthreads, int root, int src_seg, size_t src_off,
                        int dst_seg, size_t dst_off, long long numel, void* stream) {
  Comm* c = (Comm*)h;
  McSeg* ds = find_mc(c, dst_seg);
  B200_REQUIRE(ds && ds->mc_base, "broadcast_nvls: destination must be a bound multicast segment");
  B200_REQUIRE(channel >= 0 && channel < MAXCH && nblocks >= 1 && nblocks <= MAXB && nthreads >= 32 && nthreads <= 512, "bad launch shape");
  B200_REQUIRE(numel % 4 == 0 && root >= 0 && root < c->n, "broadcast_nvls: bad numel/root");
  int rc;
  if ((rc = check_seg(c, src_seg, src_off, numel * 4, "nvls bcast src"))) return rc;
  if ((rc = check_seg(c, dst_seg, dst_off, numel * 4, "nvls bcast dst"))) return rc;
  NvlsBcastArgs a;
  memset(&a, 0, sizeof a);
  a.rank = c->rank; a.n = c->n; a.channel = channel; a.root = root;
  fill_pad(c, c->rank, &a.pad);
  a.src = (const float*)seg_ptr(c, src_seg, c->rank, src_off);
  a.dst_mc = (float*)((char*)ds->mc_base + dst_off);
  a.numel = numel; a.timeout = c->timeout_cycles;
  nvls_broadcast_kernel<<<nblocks, nthreads, 0, (cudaStream_t)stream>>>(a);
  B200_LAUNCH_CHECK();
  return B200_OK;
}

// NVLS form of b200_fused_allreduce_adam for one range: in-switch reduction of the gradient slice, Adam on the
// owner, multicast store of the bf16 weights. grad and w16 must live in bound multicast segments.
int b200_fused_allreduce_adam_nvls(b200_comm_t h, int channel, int nblocks, int nthreads, int grad_seg, size_t grad_cur_off,
                                   size_t grad_other_off, int state_seg, size_t master_off, size_t m_off, size_t v_off,
                                   int w16_seg, size_t w16_off, int ctrl_seg, size_t step_off, size_t hyper_off, float scale,
                                   long long start, long long numel, int zero_other, void* stream) {
  Comm* c = (Comm*)h;
  McSeg *gs = find_mc(c, grad_seg), *ws = find_mc(c, w16_seg);
  B200_REQUIRE(gs && ws && gs->mc_base && ws->mc_base, "fused nvls: gradient and bf16-weight segments must be bound multicast segments");
  B200_REQUIRE(channel >= 0 && channel < MAXCH && nblocks >= 1 && nblocks <= MAXB && nthreads >= 32 && nthreads <= 512, "bad launch shape");
  B200_REQUIRE(start % 4 == 0 && numel % 4 == 0, "range not 4-element aligned");
  const long long end = start + numel;
  int rc;
  if ((rc = check_seg(c, grad_seg, grad_cur_off, end * 4, "grad"))) return rc;
  if ((rc = check_seg(c, grad_seg, grad_other_off, end * 4, "grad_other"))) return rc;
  if ((rc = check_seg(c, state_seg, master_off, end * 4, "master"))) return rc;
  if ((rc = check_seg(c, state_seg, m_off, end * 4, "adam_m"))) return rc;
  if ((rc = check_seg(c, state_seg, v_off, end * 4, "adam_v"))) return rc;
  if ((rc = check_seg(c, w16_seg, w16_off, end * 2, "w16"))) return rc;
  NvlsArgs a;
  memset(&a, 0, sizeof a);
  a.rank = c->rank; a.n = c->n; a.channel = channel; a.mode = 1;
  fill_pad(c, c->rank, &a.pad);
  a.src_mc = (const float*)((char*)gs->mc_base + grad_cur_off);
  a.w16_mc = (bf16*)((char*)ws->mc_base + w16_off);
  a.grad_other = (float*)seg_ptr(c, grad_seg, c->rank, grad_other_off);
  a.master = (float*)seg_ptr(c, state_seg, c->rank, master_off);
  a.m = (float*)seg_ptr(c, state_seg, c->rank, m_off);
  a.v = (float*)seg_ptr(c, state_seg, c->rank, v_off);
  a.step_ptr = (const int*)seg_ptr(c, ctrl_seg, c->rank, step_off);
  a.hyper = (const double*)seg_ptr(c, ctrl_seg, c->rank, hyper_off);
  a.scale = scale; a.start = start; a.numel = numel; a.zero_other = zero_other; a.timeout = c->timeout_cycles;
  nvls_allreduce_kernel<<<nblocks, nthreads, 0, (cudaStream_t)stream>>>(a);
  B200_LAUNCH_CHECK();
  return B200_OK;
}

}  // extern "C"
