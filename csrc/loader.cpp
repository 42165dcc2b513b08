// Native minibatch loader: a background thread assembles shuffled batches into caller-provided (pinned) slots.
//
// The reference feeds its training loop from an infinite Python generator: every pass draws a fresh permutation of
// the whole dataset, yields consecutive full batches and drops the tail (horovod/tensorflow_mnist.py:76-85); Keras does
// the same inside fit (horovod/tensorflow_mnist_gpu.py:173).  At 0.13 ms per training step the row gather of a batch
// (128 x 3136 bytes from random rows) is a visible slice of the host's per-step budget, so it moves here: one producer
// thread per loader keeps a ring of `nslots` batches ahead of the consumer, copying rows with memcpy into buffers the
// caller owns (pinned host memory on a GPU box, so the H2D copy that follows is a straight DMA).
//
// Determinism: the permutation of epoch e is a Fisher-Yates shuffle driven by splitmix64 seeded with (seed, e); the
// Python twin in b200dp/data/loader.py reproduces it bit for bit, which is how the tests check the native path.
#include <stdarg.h>
#include <stdint.h>
#include <string.h>

#include <condition_variable>
#include <mutex>
#include <thread>
#include <vector>

#include "../include/b200dp.h"

void b200_set_error(const char* fmt, ...);

namespace {

inline uint64_t splitmix64(uint64_t& s) {
  uint64_t z = (s += 0x9E3779B97F4A7C15ull);
  z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
  z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
  return z ^ (z >> 31);
}

struct Loader {
  const uint8_t* x;        // [n, row_bytes]
  const int32_t* y;        // [n]
  long long n;
  size_t row_bytes;
  int batch, nslots, shuffle;
  uint64_t seed;
  std::vector<uint8_t*> sx;
  std::vector<int32_t*> sy;
  // ring state: slot i holds batch number `filled[i]` once ready
  std::vector<long long> ready_batch;     // batch index stored in the slot, -1 if free
  long long next_produce = 0, next_consume = 0;
  std::mutex mu;
  std::condition_variable cv_free, cv_ready;
  bool stop = false;
  std::thread worker;
  std::vector<long long> perm;
  long long perm_epoch = -1;

  long long batches_per_epoch() const { return n / batch; }

  void make_perm(long long epoch) {
    perm.resize((size_t)n);
    for (long long i = 0; i < n; ++i) perm[(size_t)i] = i;
    if (shuffle) {
      uint64_t s = seed ^ (0xD1B54A32D192ED03ull * (uint64_t)(epoch + 1));
      for (long long i = n - 1; i > 0; --i) {
        const long long j = (long long)(splitmix64(s) % (uint64_t)(i + 1));
        const long long t = perm[(size_t)i];
        perm[(size_t)i] = perm[(size_t)j];
        perm[(size_t)j] = t;
      }
    }
    perm_epoch = epoch;
  }

  void fill(int slot, long long b) {
    const long long bpe = batches_per_epoch();
    const long long epoch = b / bpe, k = b % bpe;
    if (epoch != perm_epoch) make_perm(epoch);
    uint8_t* dx = sx[(size_t)slot];
    int32_t* dy = sy[(size_t)slot];
    for (int i = 0; i < batch; ++i) {
      const long long r = perm[(size_t)(k * batch + i)];
      memcpy(dx + (size_t)i * row_bytes, x + (size_t)r * row_bytes, row_bytes);
      dy[i] = y[r];
    }
  }

  void run() {
    for (;;) {
      long long b;
      int slot;
      {
        std::unique_lock<std::mutex> lk(mu);
        cv_free.wait(lk, [&] { return stop || ready_batch[(size_t)(next_produce % nslots)] < 0; });
        if (stop) return;
        b = next_produce;
        slot = (int)(b % nslots);
      }
      fill(slot, b);
      {
        std::lock_guard<std::mutex> lk(mu);
        ready_batch[(size_t)slot] = b;
        ++next_produce;
      }
      cv_ready.notify_all();
    }
  }
};

}  // namespace

extern "C" {

int b200_loader_create(const void* x, const int* y, long long n, size_t row_bytes, int batch, int nslots, int shuffle,
                       uint64_t seed, void* const* slot_x, int* const* slot_y, b200_loader_t* out) {
  if (!x || !y || !slot_x || !slot_y || !out) { b200_set_error("loader: null argument"); return 2; }
  if (n <= 0 || batch <= 0 || batch > n || row_bytes == 0 || nslots < 2 || nslots > 64) {
    b200_set_error("loader: bad shape (n=%lld, batch=%d, row_bytes=%zu, nslots=%d; need 1 <= batch <= n, 2 <= nslots <= 64)", n, batch,
                   row_bytes, nslots);
    return 2;
  }
  Loader* L = new Loader();
  L->x = (const uint8_t*)x; L->y = y; L->n = n; L->row_bytes = row_bytes; L->batch = batch; L->nslots = nslots;
  L->shuffle = shuffle; L->seed = seed;
  for (int i = 0; i < nslots; ++i) {
    if (!slot_x[i] || !slot_y[i]) { delete L; b200_set_error("loader: slot %d is null", i); return 2; }
    L->sx.push_back((uint8_t*)slot_x[i]);
    L->sy.push_back((int32_t*)slot_y[i]);
  }
  L->ready_batch.assign((size_t)nslots, -1);
  L->worker = std::thread([L] { L->run(); });
  *out = L;
  return 0;
}

// Blocks until the next batch (in order) is ready; *slot is the ring slot that holds it, *batch_index its running number.
int b200_loader_next(b200_loader_t h, int* slot, long long* batch_index) {
  Loader* L = (Loader*)h;
  if (!L || !slot) { b200_set_error("loader: null handle"); return 2; }
  std::unique_lock<std::mutex> lk(L->mu);
  const long long b = L->next_consume;
  const int s = (int)(b % L->nslots);
  L->cv_ready.wait(lk, [&] { return L->stop || L->ready_batch[(size_t)s] == b; });
  if (L->stop) { b200_set_error("loader: stopped"); return 2; }
  ++L->next_consume;
  *slot = s;
  if (batch_index) *batch_index = b;
  return 0;
}

// Hands a slot back to the producer once the consumer no longer reads it (after the H2D copy of that batch finished).
int b200_loader_release(b200_loader_t h, int slot) {
  Loader* L = (Loader*)h;
  if (!L || slot < 0 || slot >= L->nslots) { b200_set_error("loader: bad slot %d", slot); return 2; }
  {
    std::lock_guard<std::mutex> lk(L->mu);
    L->ready_batch[(size_t)slot] = -1;
  }
  L->cv_free.notify_all();
  return 0;
}

long long b200_loader_batches_per_epoch(b200_loader_t h) {
  Loader* L = (Loader*)h;
  return L ? L->batches_per_epoch() : -1;
}

int b200_loader_destroy(b200_loader_t h) {
  Loader* L = (Loader*)h;
  if (!L) return 0;
  {
    std::lock_guard<std::mutex> lk(L->mu);
    L->stop = true;
  }
  L->cv_free.notify_all();
  L->cv_ready.notify_all();
  if (L->worker.joinable()) L->worker.join();
  delete L;
  return 0;
}

}  // extern "C"
