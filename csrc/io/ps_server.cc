// Native parameter-server task (the `--job_name=ps` process of /root/reference/distribute_training.py:179-181, whose
// work in the reference is done by TensorFlow's C++ gRPC worker service + executor).
//
// One ps task = one listening TCP endpoint at its own entry of --ps_hosts (:25-27, :173-177). Workers connect to
// every ps and speak a request/reply protocol: a 32-byte header {op, a, b, c} (int64, little endian) followed by
// an optional float32 payload whose length is implied by the op and the shard size. The ps holds its shard of the
// variables (round-robin placement, global_step with ps 0) and runs what TF placed on the ps device:
//
//   PULL         -> {global_step} + weights                                    (M1)
//   PUSH_ASYNC   ApplyGradientDescent at once, unlocked semantics, ps-0 bumps global_step   (:150-152, M2)
//   PUSH_SYNC    ConditionalAccumulator: gradients tagged with a stale local_step are dropped; the n-th fresh one
//                triggers mean -> apply -> ONE global_step increment -> tokens                (:144-148, M3/M5)
//   TOKEN        FIFOQueue dequeue: blocks until a token is available                        (M4)
//   INIT / READY chief initialises or restores, the others poll until it has                 (:209-215, M7)
//   GET_STATE    weights + EMA shadows for the Saver                                         (M8)
//   DONE         a worker leaves; when all have left, serve() returns (server.join() never does)
//
// Threads: an accept loop plus one thread per connection; the shard, the accumulator and the token queue are
// guarded by one mutex (the apply is a few hundred kilobytes), a condition variable wakes token waiters.
// A worker that dies simply loses its connection; the ps keeps its state, so a restarted worker reconnects and
// carries on (what MonitoredTrainingSession's recovery relies on, :208).
#include <arpa/inet.h>
#include <errno.h>
#include <netdb.h>
#include <netinet/in.h>
#include <netinet/tcp.h>
#include <poll.h>
#include <sys/socket.h>
#include <unistd.h>

#include <algorithm>
#include <atomic>
#include <chrono>
#include <cmath>
#include <condition_variable>
#include <cstring>
#include <deque>
#include <mutex>
#include <set>
#include <string>
#include <thread>
#include <vector>

#include "coding.h"

namespace {

enum Op : int64_t { OP_READY = 0, OP_INIT, OP_PULL, OP_PUSH_ASYNC, OP_PUSH_SYNC, OP_TOKEN, OP_GET_STATE, OP_DONE,
                    OP_SYNC_INIT, OP_HELLO };

struct PsConfig {  // mirrors ttdps_config in ttd_io.h
  int64_t numel;
  int32_t n_workers, n_agg, tokens_per_step, owns_global_step;
  int32_t reg_kind;  // 0 none, 1 L1, 2 L2
  int32_t use_ema;
  float reg_scale, ema_decay;
  double idle_timeout_secs;  // > 0: serve() gives up after this long without any request (once initialised)
};

bool read_full(int fd, void* buf, size_t n) {
  char* p = static_cast<char*>(buf);
  while (n > 0) {
    const ssize_t r = ::recv(fd, p, n, 0);
    if (r == 0) return false;
    if (r < 0) {
      if (errno == EINTR) continue;
      return false;
    }
    p += r;
    n -= static_cast<size_t>(r);
  }
  return true;
}

bool write_full(int fd, const void* buf, size_t n) {
  const char* p = static_cast<const char*>(buf);
  while (n > 0) {
    const ssize_t r = ::send(fd, p, n, MSG_NOSIGNAL);
    if (r < 0) {
      if (errno == EINTR) continue;
      return false;
    }
    p += r;
    n -= static_cast<size_t>(r);
  }
  return true;
}

struct Server {
  PsConfig cfg;
  std::vector<float> w, ema, acc, reg_mask;
  int listen_fd = -1;
  int port = 0;

  std::mutex mu;
  std::condition_variable cv;
  bool initialized = false;
  int64_t global_step = 0;
  int64_t acc_step = 0;
  int32_t acc_count = 0;
  std::deque<int64_t> tokens;
  int64_t release_epoch = 0;  // bumped whenever a worker leaves: blocked token requests give up
  std::set<int64_t> done_workers;
  int64_t applied = 0, dropped_stale = 0, pulls = 0;
  std::chrono::steady_clock::time_point last_msg;
  bool stopping = false;

  std::vector<int> conn_fds;  // open connections (their threads are detached and counted in active_conns)
  int active_conns = 0;

  void apply(const float* grad, float scale, double lr, int64_t gs_after) {
    const int64_t n = cfg.numel;
    const float lrf = static_cast<float>(lr);
    float* W = w.data();
    const float* M = reg_mask.data();
    if (cfg.reg_kind == 2) {
      for (int64_t i = 0; i < n; ++i) W[i] -= lrf * (grad[i] * scale + cfg.reg_scale * W[i] * M[i]);
    } else if (cfg.reg_kind == 1) {
      for (int64_t i = 0; i < n; ++i) {
        const float s = W[i] > 0.f ? 1.f : (W[i] < 0.f ? -1.f : 0.f);
        W[i] -= lrf * (grad[i] * scale + cfg.reg_scale * s * M[i]);
      }
    } else {
      for (int64_t i = 0; i < n; ++i) W[i] -= lrf * (grad[i] * scale);
    }
    if (cfg.use_ema) {  // ExponentialMovingAverage(decay, num_updates=global_step): min(decay, (1+n)/(10+n))
      const double nn = static_cast<double>(gs_after);
      const double d = std::min(static_cast<double>(cfg.ema_decay), (1.0 + nn) / (10.0 + nn));
      const float k = static_cast<float>(1.0 - d);
      float* E = ema.data();
      for (int64_t i = 0; i < n; ++i) E[i] -= k * (E[i] - W[i]);
    }
    ++applied;
  }

  // one request; returns false when the connection should be closed
  bool handle(int fd, int64_t* worker_id, std::vector<float>* scratch) {
    int64_t hdr[4];
    if (!read_full(fd, hdr, sizeof(hdr))) return false;
    const int64_t op = hdr[0], a = hdr[1], b = hdr[2];
    const int64_t n = cfg.numel;
    int64_t reply[4] = {op, 0, 0, 0};
    auto touch = [&] { last_msg = std::chrono::steady_clock::now(); };
    switch (op) {
      case OP_HELLO: {
        *worker_id = a;
        std::lock_guard<std::mutex> lk(mu);
        touch();
        reply[1] = 1;
        return write_full(fd, reply, sizeof(reply));
      }
      case OP_READY: {
        std::lock_guard<std::mutex> lk(mu);
        touch();
        reply[1] = initialized ? 1 : 0;
        reply[2] = global_step;
        return write_full(fd, reply, sizeof(reply));
      }
      case OP_INIT: {
        scratch->resize(static_cast<size_t>(2 * n));
        if (!read_full(fd, scratch->data(), sizeof(float) * 2 * n)) return false;
        std::lock_guard<std::mutex> lk(mu);
        touch();
        std::copy(scratch->begin(), scratch->begin() + n, w.begin());
        std::copy(scratch->begin() + n, scratch->end(), ema.begin());
        global_step = acc_step = a;
        std::fill(acc.begin(), acc.end(), 0.f);
        acc_count = 0;
        tokens.clear();
        initialized = true;
        reply[1] = 1;
        return write_full(fd, reply, sizeof(reply));
      }
      case OP_SYNC_INIT: {
        {
          std::lock_guard<std::mutex> lk(mu);
          touch();
          if (cfg.owns_global_step)
            for (int64_t i = 0; i < a && i < 4096; ++i) tokens.push_back(global_step);
        }
        cv.notify_all();
        reply[1] = 1;
        return write_full(fd, reply, sizeof(reply));
      }
      case OP_PULL: {
        scratch->resize(static_cast<size_t>(n));
        {
          std::lock_guard<std::mutex> lk(mu);
          touch();
          ++pulls;
          reply[1] = global_step;
          std::copy(w.begin(), w.end(), scratch->begin());
        }
        return write_full(fd, reply, sizeof(reply)) && write_full(fd, scratch->data(), sizeof(float) * n);
      }
      case OP_PUSH_ASYNC: {
        scratch->resize(static_cast<size_t>(n));
        if (!read_full(fd, scratch->data(), sizeof(float) * n)) return false;
        double lr;
        memcpy(&lr, &a, sizeof(lr));
        std::lock_guard<std::mutex> lk(mu);
        touch();
        if (cfg.owns_global_step) ++global_step;
        apply(scratch->data(), 1.0f, lr, b + 1);
        reply[1] = global_step;
        return write_full(fd, reply, sizeof(reply));
      }
      case OP_PUSH_SYNC: {
        scratch->resize(static_cast<size_t>(n));
        if (!read_full(fd, scratch->data(), sizeof(float) * n)) return false;
        double lr;
        memcpy(&lr, &a, sizeof(lr));
        bool notify = false;
        {
          std::lock_guard<std::mutex> lk(mu);
          touch();
          if (b < acc_step) {
            ++dropped_stale;  // ConditionalAccumulator drops stale gradients silently
          } else {
            const float* g = scratch->data();
            float* A = acc.data();
            for (int64_t i = 0; i < n; ++i) A[i] += g[i];
            if (++acc_count >= cfg.n_agg) {  // take_grad(n): mean of n fresh gradients
              const int64_t new_gs = acc_step + 1;
              apply(A, 1.0f / static_cast<float>(acc_count), lr, new_gs);
              std::fill(acc.begin(), acc.end(), 0.f);
              acc_count = 0;
              acc_step = new_gs;
              if (cfg.owns_global_step) {
                global_step = new_gs;
                for (int i = 0; i < cfg.tokens_per_step; ++i) tokens.push_back(new_gs);
                notify = true;
              }
            }
          }
          reply[1] = global_step;
        }
        if (notify) cv.notify_all();
        return write_full(fd, reply, sizeof(reply));
      }
      case OP_TOKEN: {
        std::unique_lock<std::mutex> lk(mu);
        touch();
        const int64_t epoch0 = release_epoch;
        // A worker that already sent DONE contributes no more gradients, so with n_agg == n_workers no further
        // aggregate can complete: a token request that arrives AFTER that DONE (a restarted worker, uneven stop
        // timing) must not wait for one. It gets the current global_step at once and reaches its stop check.
        const bool nobody_left = !done_workers.empty() && tokens.empty();
        cv.wait(lk, [&] { return nobody_left || stopping || !tokens.empty() || release_epoch != epoch0; });
        if (!tokens.empty()) {
          reply[1] = tokens.front();
          tokens.pop_front();
        } else {
          reply[1] = global_step;  // nobody will complete another aggregate: let the worker reach its stop check
        }
        reply[2] = global_step;
        lk.unlock();
        return write_full(fd, reply, sizeof(reply));
      }
      case OP_GET_STATE: {
        scratch->resize(static_cast<size_t>(2 * n));
        {
          std::lock_guard<std::mutex> lk(mu);
          touch();
          reply[1] = global_step;
          std::copy(w.begin(), w.end(), scratch->begin());
          std::copy(ema.begin(), ema.end(), scratch->begin() + n);
        }
        return write_full(fd, reply, sizeof(reply)) && write_full(fd, scratch->data(), sizeof(float) * 2 * n);
      }
      case OP_DONE: {
        {
          std::lock_guard<std::mutex> lk(mu);
          touch();
          done_workers.insert(*worker_id);
          ++release_epoch;
        }
        cv.notify_all();
        reply[1] = 1;
        write_full(fd, reply, sizeof(reply));
        return false;
      }
      default:
        return false;
    }
  }

  void connection(int fd) {
    int one = 1;
    setsockopt(fd, IPPROTO_TCP, TCP_NODELAY, &one, sizeof(one));
    int64_t worker_id = -1 - fd;  // until HELLO names it
    std::vector<float> scratch;
    while (handle(fd, &worker_id, &scratch)) {
    }
    ::close(fd);
    {
      std::lock_guard<std::mutex> lk(mu);
      conn_fds.erase(std::remove(conn_fds.begin(), conn_fds.end(), fd), conn_fds.end());
      --active_conns;  // last touch of the server by this thread: destroy() may free it once this reaches zero
      cv.notify_all();
    }
  }
};

}  // namespace

extern "C" {

void* ttdps_server_create(const char* bind_host, int32_t port, const void* config, const float* reg_mask) {
  const PsConfig* c = static_cast<const PsConfig*>(config);
  if (!c || c->numel < 0 || c->n_workers < 1 || c->n_agg < 1) {
    ttdio::set_error("ttdps_server_create: bad configuration");
    return nullptr;
  }
  Server* S = new Server();
  S->cfg = *c;
  S->w.assign(static_cast<size_t>(c->numel), 0.f);
  S->ema.assign(static_cast<size_t>(c->numel), 0.f);
  S->acc.assign(static_cast<size_t>(c->numel), 0.f);
  S->reg_mask.assign(static_cast<size_t>(c->numel), 0.f);
  if (reg_mask) std::copy(reg_mask, reg_mask + c->numel, S->reg_mask.begin());
  S->last_msg = std::chrono::steady_clock::now();
  const int fd = ::socket(AF_INET, SOCK_STREAM, 0);
  if (fd < 0) {
    ttdio::set_error(std::string("ttdps_server_create: socket: ") + strerror(errno));
    delete S;
    return nullptr;
  }
  int one = 1;
  setsockopt(fd, SOL_SOCKET, SO_REUSEADDR, &one, sizeof(one));
  sockaddr_in addr;
  memset(&addr, 0, sizeof(addr));
  addr.sin_family = AF_INET;
  addr.sin_port = htons(static_cast<uint16_t>(port));
  addr.sin_addr.s_addr = htonl(INADDR_ANY);
  if (bind_host && *bind_host && strcmp(bind_host, "0.0.0.0") != 0) {
    addrinfo hints, *res = nullptr;
    memset(&hints, 0, sizeof(hints));
    hints.ai_family = AF_INET;
    if (getaddrinfo(bind_host, nullptr, &hints, &res) == 0 && res) {
      addr.sin_addr = reinterpret_cast<sockaddr_in*>(res->ai_addr)->sin_addr;
      freeaddrinfo(res);
    }  // an unresolvable name falls back to all interfaces
  }
  if (::bind(fd, reinterpret_cast<sockaddr*>(&addr), sizeof(addr)) != 0 || ::listen(fd, 64) != 0) {
    ttdio::set_error(std::string("ttdps_server_create: bind/listen on port ") + std::to_string(port) + ": " +
                     strerror(errno));
    ::close(fd);
    delete S;
    return nullptr;
  }
  socklen_t len = sizeof(addr);
  getsockname(fd, reinterpret_cast<sockaddr*>(&addr), &len);
  S->port = ntohs(addr.sin_port);
  S->listen_fd = fd;
  return S;
}

int32_t ttdps_server_port(void* h) { return static_cast<Server*>(h)->port; }

/* Accepts and serves until every worker has sent DONE (returns 0) or the idle limit passes (returns 1). */
int ttdps_server_serve(void* h) {
  Server* S = static_cast<Server*>(h);
  int rc = 0;
  for (;;) {
    {
      std::lock_guard<std::mutex> lk(S->mu);
      if (static_cast<int>(S->done_workers.size()) >= S->cfg.n_workers) break;
      if (S->cfg.idle_timeout_secs > 0 && S->initialized) {
        const double idle =
            std::chrono::duration<double>(std::chrono::steady_clock::now() - S->last_msg).count();
        if (idle > S->cfg.idle_timeout_secs) {
          rc = 1;
          break;
        }
      }
    }
    pollfd p;
    p.fd = S->listen_fd;
    p.events = POLLIN;
    p.revents = 0;
    const int r = ::poll(&p, 1, 100);
    if (r > 0 && (p.revents & POLLIN)) {
      const int cfd = ::accept(S->listen_fd, nullptr, nullptr);
      if (cfd >= 0) {
        std::lock_guard<std::mutex> lk(S->mu);
        S->conn_fds.push_back(cfd);
        ++S->active_conns;
        std::thread([S, cfd] { S->connection(cfd); }).detach();  // a reconnecting worker does not pile up joinable threads
      }
    }
  }
  {
    std::lock_guard<std::mutex> lk(S->mu);
    S->stopping = true;
    ++S->release_epoch;
  }
  S->cv.notify_all();
  return rc;
}

/* out[0..5] = applied, dropped_stale, pulls, global_step, initialized, workers done */
void ttdps_server_stats(void* h, int64_t* out) {
  Server* S = static_cast<Server*>(h);
  std::lock_guard<std::mutex> lk(S->mu);
  out[0] = S->applied;
  out[1] = S->dropped_stale;
  out[2] = S->pulls;
  out[3] = S->global_step;
  out[4] = S->initialized ? 1 : 0;
  out[5] = static_cast<int64_t>(S->done_workers.size());
}

void ttdps_server_destroy(void* h) {
  Server* S = static_cast<Server*>(h);
  if (!S) return;
  {
    std::lock_guard<std::mutex> lk(S->mu);
    S->stopping = true;
    ++S->release_epoch;
    for (int fd : S->conn_fds) ::shutdown(fd, SHUT_RDWR);  // unblocks threads sitting in recv
  }
  S->cv.notify_all();
  if (S->listen_fd >= 0) ::close(S->listen_fd);
  {
    std::unique_lock<std::mutex> lk(S->mu);
    S->cv.wait(lk, [&] { return S->active_conns == 0; });
  }
  delete S;
}

}  // extern "C"
