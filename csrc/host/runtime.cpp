// Host runtime services: message pump, versioned key/value mailbox, locality-aware delay
// scheduler.
//
// Reference behaviour (re-implemented natively):
//  * GraphManager/kernel/DrMessagePump.h:22-295 — actor-style runtime: messages are delivered to
//    a listener under the listener's own lock by a pool of worker threads; delayed messages sit
//    in a time-ordered multimap.
//  * ProcessService/ProcessService.cs:674-746 — per-process versioned key/value mailbox:
//    SetValue bumps the version, BlockOnStatus(key, lastSeenVersion, timeout) returns at once if a
//    newer version exists, else waits (the GM <-> worker control plane).
//  * LocalScheduler/LocalScheduler.cs:52-53, 132-268 and Queues.cs:249-453 — delay scheduling:
//    a process is offered to the queues of the computers it has affinity for, after 1 s to their
//    racks, after 2 s to the whole cluster; hard constraints never relax.
#include "../../include/dryad_host.h"

#include <string.h>

#include <chrono>
#include <condition_variable>
#include <deque>
#include <map>
#include <mutex>
#include <string>
#include <thread>
#include <vector>

namespace {

using Clock = std::chrono::steady_clock;

// ---- message pump -------------------------------------------------------------------------------
typedef void (*pump_handler)(int64_t listener, int64_t message, void* user);

struct Msg { int64_t listener, message; };

struct Pump {
  std::mutex mu;
  std::condition_variable cv;
  std::deque<Msg> q;
  std::multimap<Clock::time_point, Msg> delayed;
  std::map<int64_t, std::mutex*> listener_locks;     // delivery happens under the listener's lock
  std::vector<std::thread> workers;
  pump_handler handler = nullptr;
  void* user = nullptr;
  bool stop = false;
  int64_t delivered = 0;
  int64_t notifiers = 0;          // stream-completion callbacks created and not yet fired (guarded by mu)
};

void pump_worker(Pump* p) {
  for (;;) {
    Msg m;
    {
      std::unique_lock<std::mutex> l(p->mu);
      for (;;) {
        auto now = Clock::now();
        while (!p->delayed.empty() && p->delayed.begin()->first <= now) {
          p->q.push_back(p->delayed.begin()->second);
          p->delayed.erase(p->delayed.begin());
        }
        if (!p->q.empty() || p->stop) break;
        if (p->delayed.empty()) p->cv.wait(l);
        else {
          // copy the deadline: wait_until keeps a reference to its argument while the mutex is
          // released, and another worker may erase that multimap node meanwhile (found by ASan)
          const Clock::time_point deadline = p->delayed.begin()->first;
          p->cv.wait_until(l, deadline);
        }
      }
      if (p->q.empty() && p->stop) return;
      m = p->q.front(); p->q.pop_front();
    }
    std::mutex* lk;
    {
      std::lock_guard<std::mutex> l(p->mu);
      auto& slot = p->listener_locks[m.listener];
      if (!slot) slot = new std::mutex();
      lk = slot;
    }
    {
      std::lock_guard<std::mutex> l(*lk);
      if (p->handler) p->handler(m.listener, m.message, p->user);
    }
    { std::lock_guard<std::mutex> l(p->mu); p->delivered++; }
    p->cv.notify_all();
  }
}

// ---- mailbox ------------------------------------------------------------------------------------------
struct Value { uint64_t version = 0; std::string status; std::string data; };
struct Mailbox {
  std::mutex mu;
  std::condition_variable cv;
  std::map<std::string, Value> kv;
};

// ---- delay scheduler ---------------------------------------------------------------------------------
struct Proc { int64_t id; std::vector<int> computers; bool hard; Clock::time_point submitted; bool taken = false; };
struct Sched {
  std::mutex mu;
  std::vector<int> rack_of;             // computer -> rack
  std::deque<Proc> procs;               // FIFO of waiting processes
  double rack_delay = 1.0, cluster_delay = 2.0;
};

}  // namespace

extern "C" {

// ---- pump ----------------------------------------------------------------------------------------------
void* pump_create(int threads, pump_handler handler, void* user) {
  Pump* p = new Pump();
  p->handler = handler; p->user = user;
  for (int i = 0; i < (threads > 0 ? threads : 4); ++i) p->workers.emplace_back(pump_worker, p);
  return p;
}
void pump_post(void* h, int64_t listener, int64_t message, double delay_seconds) {
  Pump* p = (Pump*)h;
  {
    std::lock_guard<std::mutex> l(p->mu);
    Msg m{listener, message};
    if (delay_seconds > 0)
      p->delayed.emplace(Clock::now() + std::chrono::duration_cast<Clock::duration>(std::chrono::duration<double>(delay_seconds)), m);
    else p->q.push_back(m);
  }
  p->cv.notify_all();
}
// Stream-completion notifier: `pump_notifier_fn()` is a `void(void*)` suitable for
// cudaLaunchHostFunc; when the stream reaches it, the message is posted to the pump and delivered
// under the listener's lock like any other — a vertex's GPU work reports its own completion to the
// graph-manager event loop without any thread blocking in a stream synchronize
// (GraphManager/kernel/DrMessagePump.h:81-87 delivers process/vertex state changes the same way).
struct Notifier { Pump* p; int64_t listener, message; };
void* pump_make_notifier(void* h, int64_t listener, int64_t message) {
  Pump* p = (Pump*)h;
  { std::lock_guard<std::mutex> l(p->mu); ++p->notifiers; }
  return new Notifier{p, listener, message};
}
// Runs on a CUDA driver thread when the stream reaches the callback.  pump_destroy waits until every
// outstanding notifier has fired, so `n->p` is alive here even when the job is torn down with GPU
// work still in flight.
static void pump_notifier_fire(void* arg) {
  Notifier* n = (Notifier*)arg;
  pump_post(n->p, n->listener, n->message, 0.0);
  { std::lock_guard<std::mutex> l(n->p->mu); --n->p->notifiers; }
  n->p->cv.notify_all();
  delete n;
}
void* pump_notifier_fn() { return (void*)&pump_notifier_fire; }
// Wait until `count` messages have been delivered in total (or timeout). Returns delivered count.
int64_t pump_wait_delivered(void* h, int64_t count, double timeout_seconds) {
  Pump* p = (Pump*)h;
  std::unique_lock<std::mutex> l(p->mu);
  p->cv.wait_for(l, std::chrono::duration<double>(timeout_seconds), [&] { return p->delivered >= count; });
  return p->delivered;
}
void pump_destroy(void* h) {
  Pump* p = (Pump*)h;
  {
    // callbacks queued on CUDA streams still point at this pump: let them fire first (bounded wait:
    // a stream that never reaches its callback must not hang the teardown; such a notifier is leaked
    // together with the pump rather than left to post into freed memory)
    std::unique_lock<std::mutex> l(p->mu);
    if (!p->cv.wait_for(l, std::chrono::seconds(30), [&] { return p->notifiers == 0; })) {
      p->stop = true;
      l.unlock();
      p->cv.notify_all();
      for (auto& t : p->workers) t.join();
      return;                                            // pump intentionally not freed
    }
    p->stop = true;
  }
  p->cv.notify_all();
  for (auto& t : p->workers) t.join();
  for (auto& kv : p->listener_locks) delete kv.second;
  delete p;
}

// ---- mailbox -------------------------------------------------------------------------------------------
void* mbox_create(void) { return new Mailbox(); }
void mbox_destroy(void* h) { delete (Mailbox*)h; }
uint64_t mbox_set(void* h, const char* key, const char* short_status, const void* data, int64_t n) {
  Mailbox* m = (Mailbox*)h;
  uint64_t v;
  {
    std::lock_guard<std::mutex> l(m->mu);
    Value& val = m->kv[key];
    val.version++;
    val.status = short_status ? short_status : "";
    val.data.assign((const char*)data, (size_t)(n > 0 ? n : 0));
    v = val.version;
  }
  m->cv.notify_all();
  return v;
}
// Returns the current version (> last_seen) and copies up to cap bytes of the value, or 0 on timeout.
uint64_t mbox_block_on_status(void* h, const char* key, uint64_t last_seen, double timeout_seconds,
                              char* status_out, int32_t status_cap, void* data_out, int64_t cap, int64_t* n_out) {
  Mailbox* m = (Mailbox*)h;
  std::unique_lock<std::mutex> l(m->mu);
  bool ok = m->cv.wait_for(l, std::chrono::duration<double>(timeout_seconds), [&] {
    auto it = m->kv.find(key);
    return it != m->kv.end() && it->second.version > last_seen;
  });
  if (!ok) return 0;
  Value& val = m->kv[key];
  if (status_out && status_cap > 0) snprintf(status_out, (size_t)status_cap, "%s", val.status.c_str());
  int64_t n = (int64_t)val.data.size();
  if (n_out) *n_out = n;
  if (data_out && cap > 0) memcpy(data_out, val.data.data(), (size_t)(n < cap ? n : cap));
  return val.version;
}

// ---- delay scheduler ----------------------------------------------------------------------------------
void* ls_create(int ncomputers, const int32_t* rack_of, double rack_delay, double cluster_delay) {
  Sched* s = new Sched();
  s->rack_of.assign(rack_of, rack_of + ncomputers);
  s->rack_delay = rack_delay; s->cluster_delay = cluster_delay;
  return s;
}
void ls_destroy(void* h) { delete (Sched*)h; }
void ls_submit(void* h, int64_t proc_id, const int32_t* computers, int n, int hard) {
  Sched* s = (Sched*)h;
  std::lock_guard<std::mutex> l(s->mu);
  Proc p; p.id = proc_id; p.computers.assign(computers, computers + n); p.hard = hard != 0; p.submitted = Clock::now();
  s->procs.push_back(p);
}
// An idle computer asks for work.  `age_bias` is added to every process's waiting time (tests use it
// instead of sleeping).  Returns a process id or -1.
int64_t ls_computer_idle(void* h, int computer, double age_bias) {
  Sched* s = (Sched*)h;
  std::lock_guard<std::mutex> l(s->mu);
  auto now = Clock::now();
  int rack = computer >= 0 && computer < (int)s->rack_of.size() ? s->rack_of[computer] : -1;
  // pass 1: processes with affinity for this computer; pass 2: rack-level (aged); pass 3: cluster-level
  for (int pass = 0; pass < 3; ++pass) {
    for (auto it = s->procs.begin(); it != s->procs.end(); ++it) {
      double age = std::chrono::duration<double>(now - it->submitted).count() + age_bias;
      bool match = false;
      if (pass == 0) {
        for (int c : it->computers) if (c == computer) match = true;
        if (it->computers.empty()) match = true;                 // no preference
      } else if (it->hard) {
        match = false;                                           // hard constraints never relax
      } else if (pass == 1 && age >= s->rack_delay) {
        for (int c : it->computers) if (c >= 0 && c < (int)s->rack_of.size() && s->rack_of[c] == rack) match = true;
      } else if (pass == 2 && age >= s->cluster_delay) {
        match = true;
      }
      if (match) { int64_t id = it->id; s->procs.erase(it); return id; }
    }
  }
  return -1;
}
int ls_waiting(void* h) { Sched* s = (Sched*)h; std::lock_guard<std::mutex> l(s->mu); return (int)s->procs.size(); }

}  // extern "C"
