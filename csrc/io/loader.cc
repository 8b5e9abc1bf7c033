// Multi-thread batch reader (the README's "ReadByQueue" variant, /root/reference/README.md:5-7; the single-thread
// form is mnist.train.next_batch(128), distribute_training.py:224).
//
// The data set stays where Python put it; worker threads gather rows into a ring of caller-provided batch
// buffers (pinned host memory on a GPU box, so the trainer's cudaMemcpyAsync reads them directly) without
// holding the GIL. The ORDER of examples is not decided here: Python pushes one index permutation per epoch
// (the same numpy generator DataSet.next_batch uses), and batch k is positions [k*B, (k+1)*B) of the
// concatenated stream of permutations — which is exactly next_batch's "tail of one epoch + head of the next"
// roll-over, so the native reader and the Python loader produce identical batches.
#include <atomic>
#include <condition_variable>
#include <cstring>
#include <deque>
#include <mutex>
#include <thread>
#include <vector>

#include "coding.h"

namespace {

struct Epoch {
  int64_t base;  // stream position of its first index
  std::vector<int64_t> perm;
};

struct Loader {
  const float* images = nullptr;
  const int64_t* labels = nullptr;
  int64_t n = 0, feat = 0;
  int batch = 0, ring = 0;
  std::vector<float*> x_slots;
  std::vector<int64_t*> y_slots;

  std::mutex mu;
  std::condition_variable cv_work, cv_ready;
  std::deque<Epoch> epochs;   // index stream, oldest first
  int64_t stream_end = 0;     // positions [0, stream_end) have been pushed
  int64_t next_claim = 0;     // next batch number a worker may take
  int64_t released = 0;       // batches [0, released) have been handed back by the consumer
  int64_t consumed = 0;       // batches [0, consumed) have been returned by next()
  std::vector<int64_t> ready;  // ready[slot] = batch number + 1 currently held by the slot (0: none)
  bool stop = false;
  std::vector<std::thread> threads;

  // copies the indices of batch k into idx; caller holds mu and has checked that the stream covers it
  void indices_of(int64_t k, std::vector<int64_t>* idx) {
    int64_t pos = k * batch;
    idx->resize(batch);
    size_t e = 0;
    for (int i = 0; i < batch; ++i, ++pos) {
      while (pos >= epochs[e].base + static_cast<int64_t>(epochs[e].perm.size())) ++e;
      (*idx)[i] = epochs[e].perm[pos - epochs[e].base];
    }
  }

  void drop_finished_epochs() {  // epochs entirely below every batch still to be claimed
    const int64_t low = next_claim * batch;
    while (epochs.size() > 1 && epochs.front().base + static_cast<int64_t>(epochs.front().perm.size()) <= low)
      epochs.pop_front();
  }

  void worker() {
    std::vector<int64_t> idx;
    for (;;) {
      int64_t k;
      {
        std::unique_lock<std::mutex> lk(mu);
        cv_work.wait(lk, [&] {
          return stop || ((next_claim + 1) * batch <= stream_end && next_claim < released + ring);
        });
        if (stop) return;
        k = next_claim++;
        indices_of(k, &idx);
        drop_finished_epochs();
      }
      const int slot = static_cast<int>(k % ring);
      float* x = x_slots[slot];
      int64_t* y = y_slots[slot];
      for (int i = 0; i < batch; ++i) {
        memcpy(x + static_cast<int64_t>(i) * feat, images + idx[i] * feat, sizeof(float) * feat);
        y[i] = labels[idx[i]];
      }
      {
        std::lock_guard<std::mutex> lk(mu);
        ready[slot] = k + 1;
      }
      cv_ready.notify_all();
    }
  }
};

}  // namespace

extern "C" {

void* ttdio_loader_create(const float* images, const int64_t* labels, int64_t n, int64_t feat, int32_t batch,
                          int32_t ring, int32_t threads, float* const* x_slots, int64_t* const* y_slots) {
  if (!images || !labels || n < 1 || feat < 1 || batch < 1 || ring < 2 || threads < 1 || !x_slots || !y_slots) {
    ttdio::set_error("ttdio_loader_create: bad argument");
    return nullptr;
  }
  Loader* L = new Loader();
  L->images = images; L->labels = labels; L->n = n; L->feat = feat; L->batch = batch; L->ring = ring;
  L->x_slots.assign(x_slots, x_slots + ring);
  L->y_slots.assign(y_slots, y_slots + ring);
  L->ready.assign(ring, 0);
  for (int t = 0; t < threads; ++t) L->threads.emplace_back([L] { L->worker(); });
  return L;
}

/* Appends one epoch's permutation of [0, n) to the index stream. */
int ttdio_loader_push_epoch(void* h, const int64_t* perm, int64_t count) {
  Loader* L = static_cast<Loader*>(h);
  if (!L || !perm || count < 1) {
    ttdio::set_error("ttdio_loader_push_epoch: bad argument");
    return 1;
  }
  for (int64_t i = 0; i < count; ++i)
    if (perm[i] < 0 || perm[i] >= L->n) {
      ttdio::set_error("ttdio_loader_push_epoch: index out of range");
      return 1;
    }
  {
    std::lock_guard<std::mutex> lk(L->mu);
    Epoch e;
    e.base = L->stream_end;
    e.perm.assign(perm, perm + count);
    L->stream_end += count;
    L->epochs.push_back(std::move(e));
  }
  L->cv_work.notify_all();
  return 0;
}

/* Stream positions pushed but not yet consumed by next(): lets the caller keep an epoch or two ahead. */
int64_t ttdio_loader_backlog(void* h) {
  Loader* L = static_cast<Loader*>(h);
  std::lock_guard<std::mutex> lk(L->mu);
  return L->stream_end - L->consumed * L->batch;
}

/* Blocks until the next batch (in order) is assembled; returns its ring slot, -1 on timeout, -2 when closed. */
int ttdio_loader_next(void* h, int32_t timeout_ms) {
  Loader* L = static_cast<Loader*>(h);
  std::unique_lock<std::mutex> lk(L->mu);
  const int64_t k = L->consumed;
  const int slot = static_cast<int>(k % L->ring);
  const bool ok = L->cv_ready.wait_for(lk, std::chrono::milliseconds(timeout_ms),
                                       [&] { return L->stop || L->ready[slot] == k + 1; });
  if (L->stop) return -2;
  if (!ok) return -1;
  L->consumed = k + 1;
  return slot;
}

/* The consumer is done with every batch up to and including number `upto` (0-based): their slots may be refilled. */
int ttdio_loader_release(void* h, int64_t upto) {
  Loader* L = static_cast<Loader*>(h);
  {
    std::lock_guard<std::mutex> lk(L->mu);
    if (upto + 1 > L->consumed) upto = L->consumed - 1;
    if (upto + 1 > L->released) L->released = upto + 1;
  }
  L->cv_work.notify_all();
  return 0;
}

void ttdio_loader_destroy(void* h) {
  Loader* L = static_cast<Loader*>(h);
  if (!L) return;
  {
    std::lock_guard<std::mutex> lk(L->mu);
    L->stop = true;
  }
  L->cv_work.notify_all();
  L->cv_ready.notify_all();
  for (auto& t : L->threads) t.join();
  delete L;
}

}  // extern "C"
