// Native micro-batch producer: worker threads cut (input, target) token windows out of one flat token array and write
// them -- already sliced for this context-parallel rank (contiguous or zig-zag) -- into a ring of caller-owned (pinned)
// host buffers, ahead of the training loop.
//
// Reference counterpart: scaletorch/data/dataloader.py:107-292 (torch DataLoader + DistributedSampler + Python collate
// with a one-batch software prefetch) over the "concat_chunk" dataset of scaletorch/data/dataset.py:63-86 (all tokens
// concatenated, cut into chunks of S + 1).  Here the same sampling semantics live in ~200 lines of C++ so that the
// host side of a step is: wait for a slot, cudaMemcpyAsync from pinned memory, release the slot.
//
// Semantics (mirrored by the pure-Python oracle in scaletorch_b200/data/native_loader.py):
//   sample w            = tokens[w * (S + 1) : (w + 1) * (S + 1)]                 (num_samples = n_tokens / (S + 1))
//   epoch permutation   = Fisher-Yates driven by splitmix64(seed + epoch)          (identity when shuffle == 0)
//   rank's sample list  = perm[data_rank], perm[data_rank + data_world], ...        (tail dropped to a multiple of mbs)
//   batch b             = samples [b * mbs, (b + 1) * mbs) of that list
//   CP slice            = positions [cp_rank * S/cp, +S/cp) or, zig-zag, chunks cp_rank and 2cp-1-cp_rank of 2cp
//   slot layout         = int64 [2][mbs][S/cp]: inputs, then next-token targets
//
// The library allocates no buffers: the ring belongs to the caller.  All entry points are thread-safe with respect to
// the worker threads; one consumer thread is assumed.
#include <atomic>
#include <condition_variable>
#include <cstdint>
#include <mutex>
#include <numeric>
#include <thread>
#include <vector>

namespace {

inline uint64_t splitmix64(uint64_t& s) {
  uint64_t z = (s += 0x9E3779B97F4A7C15ull);
  z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
  z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
  return z ^ (z >> 31);
}

struct Loader {
  const int32_t* tokens;
  long n_tokens;
  int S, mbs, cp_rank, cp_size, zigzag, data_rank, data_world, shuffle, n_slots;
  uint64_t seed;
  int64_t* ring;
  long num_samples, per_rank, batches_per_epoch;
  int s_local;

  std::mutex mu;
  std::condition_variable cv;
  std::vector<int> state;          // per slot: 0 free, 1 being filled, 2 ready, 3 held by the consumer
  std::vector<long> slot_batch;    // batch index a slot holds / is being filled with
  long next_claim = 0;             // next batch index a worker may claim
  long next_serve = 0;             // next batch index the consumer gets
  bool stop = false;
  std::vector<std::thread> workers;
  long cached_epoch = -1;
  std::vector<long> perm;          // permutation of the cached epoch (guarded by perm_mu)
  std::mutex perm_mu;

  long slot_elems() const { return 2L * mbs * s_local; }

  void epoch_perm(long epoch, std::vector<long>& out) {
    std::lock_guard<std::mutex> g(perm_mu);
    if (epoch != cached_epoch) {
      perm.resize(num_samples);
      std::iota(perm.begin(), perm.end(), 0L);
      if (shuffle) {
        uint64_t s = seed + uint64_t(epoch) * 0x632BE59BD9B4E019ull;
        for (long i = num_samples - 1; i > 0; --i) {
          const long j = long(splitmix64(s) % uint64_t(i + 1));
          std::swap(perm[i], perm[j]);
        }
      }
      cached_epoch = epoch;
    }
    out = perm;
  }

  // copy the CP slice of sample `w` (inputs: offset 0, targets: offset 1) to dst[s_local]
  void copy_slice(long w, int shift, int64_t* dst) const {
    const int32_t* src = tokens + w * long(S + 1) + shift;
    if (!zigzag) {
      const int lo = cp_rank * s_local;
      for (int i = 0; i < s_local; ++i) dst[i] = src[lo + i];
    } else {
      const int c = s_local / 2;
      const int a = cp_rank * c, b = (2 * cp_size - 1 - cp_rank) * c;
      for (int i = 0; i < c; ++i) dst[i] = src[a + i];
      for (int i = 0; i < c; ++i) dst[c + i] = src[b + i];
    }
  }

  void fill(long batch, int slot) {
    const long epoch = batch / batches_per_epoch, in_epoch = batch % batches_per_epoch;
    std::vector<long> p;
    epoch_perm(epoch, p);
    int64_t* base = ring + long(slot) * slot_elems();
    for (int m = 0; m < mbs; ++m) {
      const long k = in_epoch * mbs + m;                    // k-th sample of this rank in this epoch
      const long w = p[data_rank + k * data_world];
      copy_slice(w, 0, base + long(m) * s_local);
      copy_slice(w, 1, base + long(mbs) * s_local + long(m) * s_local);
    }
  }

  void work() {
    for (;;) {
      long batch;
      int slot;
      {
        std::unique_lock<std::mutex> lk(mu);
        cv.wait(lk, [&] { return stop || state[next_claim % n_slots] == 0; });
        if (stop) return;
        batch = next_claim++;
        slot = int(batch % n_slots);
        state[slot] = 1;
        slot_batch[slot] = batch;
      }
      fill(batch, slot);
      {
        std::lock_guard<std::mutex> g(mu);
        state[slot] = 2;
      }
      cv.notify_all();
    }
  }
};

}  // namespace

extern "C" {

// Returns an opaque handle, or nullptr when the arguments are inconsistent (sequence not divisible by the CP layout,
// fewer samples than one batch per rank, ...).  `ring` must hold n_slots * 2 * mbs * (S / cp_size) int64 values.
void* stb_loader_create(const int32_t* tokens, long n_tokens, int seq_len, int mbs, int cp_rank, int cp_size,
                        int zigzag, int data_rank, int data_world, int shuffle, uint64_t seed, int n_slots,
                        int64_t* ring, int n_threads) {
  if (!tokens || !ring || seq_len <= 0 || mbs <= 0 || cp_size <= 0 || n_slots < 2 || n_threads < 1) return nullptr;
  if (cp_rank < 0 || cp_rank >= cp_size || data_rank < 0 || data_rank >= data_world) return nullptr;
  if (seq_len % (zigzag ? 2 * cp_size : cp_size) != 0) return nullptr;
  auto* L = new Loader();
  L->tokens = tokens; L->n_tokens = n_tokens; L->S = seq_len; L->mbs = mbs; L->cp_rank = cp_rank; L->cp_size = cp_size;
  L->zigzag = zigzag; L->data_rank = data_rank; L->data_world = data_world; L->shuffle = shuffle; L->seed = seed;
  L->n_slots = n_slots; L->ring = ring;
  L->s_local = seq_len / cp_size;
  L->num_samples = n_tokens / (seq_len + 1);
  L->per_rank = L->num_samples / data_world;
  L->batches_per_epoch = L->per_rank / mbs;
  if (L->batches_per_epoch < 1) {
    delete L;
    return nullptr;
  }
  L->state.assign(n_slots, 0);
  L->slot_batch.assign(n_slots, -1);
  for (int i = 0; i < n_threads; ++i) L->workers.emplace_back([L] { L->work(); });
  return L;
}

long stb_loader_batches_per_epoch(void* h) { return static_cast<Loader*>(h)->batches_per_epoch; }
long stb_loader_num_samples(void* h) { return static_cast<Loader*>(h)->num_samples; }

// Blocks until the next batch (in order) is ready; returns its slot index and marks it held.  *batch_out = its index.
int stb_loader_next(void* h, long* batch_out) {
  auto* L = static_cast<Loader*>(h);
  std::unique_lock<std::mutex> lk(L->mu);
  const long want = L->next_serve;
  const int slot = int(want % L->n_slots);
  L->cv.wait(lk, [&] { return L->state[slot] == 2 && L->slot_batch[slot] == want; });
  L->state[slot] = 3;
  L->next_serve++;
  if (batch_out) *batch_out = want;
  return slot;
}

// The consumer is done with the slot (its H2D copy has completed): workers may refill it.
int stb_loader_release(void* h, int slot) {
  auto* L = static_cast<Loader*>(h);
  {
    std::lock_guard<std::mutex> g(L->mu);
    if (slot < 0 || slot >= L->n_slots || L->state[slot] != 3) return 1;
    L->state[slot] = 0;
  }
  L->cv.notify_all();
  return 0;
}

void stb_loader_destroy(void* h) {
  auto* L = static_cast<Loader*>(h);
  if (!L) return;
  {
    std::lock_guard<std::mutex> g(L->mu);
    L->stop = true;
  }
  L->cv.notify_all();
  for (auto& t : L->workers) t.join();
  delete L;
}

}  // extern "C"
