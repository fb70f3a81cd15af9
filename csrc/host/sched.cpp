// Graph manager core: stage DAG -> vertex graph, versioned execution attempts, failure
// accounting, upstream invalidation, speculative duplicates, stage statistics (straggler model),
// dynamic managers (aggregation tree, distributor sizing, broadcast tree).
//
// Reference behaviour (re-implemented natively, as the reference does in C++/CLI):
//  * GraphManager/vertex/DrGraph.cpp:86-482        job state machine, failure budget (abort after
//    maxActiveFailureCount = 6, GraphManager/graph/DrGraphParameters.cpp:52), read-error ->
//    re-run the producer (:392-415), shutdown when all vertices completed (:355-367).
//  * GraphManager/vertex/DrVertex.cpp:356-1171     versioned attempts: pending / running records /
//    completed record; first completed version wins, others are cancelled (:820-826).
//  * GraphManager/vertex/DrCohort.cpp:852-929      EnsurePendingVersion policy: at most 3
//    concurrent running versions, duplicates only on request.
//  * DryadLinqGraphManager/GraphBuilder.cs:420-504 ConnectPointwise (1:N, N:1, N:N) and
//    ConnectCrossProduct.
//  * GraphManager/stagemanager/DrStageStatistics.cpp:23-652  time = startup + multiplier*bytes
//    fitted by robust regression over random subsamples; outlier threshold; first estimate at 50 %
//    completion; stages of <= 10 vertices fall back to the default threshold (:93-111).
//  * GraphManager/stagemanager/DrDefaultManager.cpp:664-714  CheckForDuplicates.
//  * GraphManager/stagemanager/DrDynamicAggregateManager.h:174-179, DrDynamicDistributor.h:64,
//    DrDynamicBroadcast.cpp:125-130  dynamic graph rewrites.
//  * GraphManager/vertex/DrClique.h:28-64, DrCohort.cpp:852-929  start cliques / gangs: vertices
//    joined by FIFO or pipe edges start together once every member's external (file) inputs exist,
//    and fail together (a streamed edge keeps nothing to re-read).
//  * GraphManager/stagemanager/DrPipelineSplitManager.h:22-83  a dynamic split of a stage is
//    followed 1:1 by the pointwise stages downstream (semi variant: side inputs are re-broadcast).
//  * GraphManager/stagemanager/DrDynamicRangeDistributor.h:22-78  consumer count of a range
//    distributor from sample volume / sampling rate.
#include "../../include/dryad_host.h"

#include <math.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include <algorithm>
#include <map>
#include <mutex>
#include <random>
#include <string>
#include <vector>

namespace {

enum VState { V_WAITING = 0, V_READY = 1, V_RUNNING = 2, V_COMPLETED = 3, V_FAILED = 4 };
enum JState { J_NOT_STARTED = 0, J_RUNNING = 1, J_STOPPING = 2, J_SUCCEEDED = 3, J_FAILED = 4 };

struct Attempt { int version; double start; };

struct Vertex {
  int id, stage, part;
  int state = V_WAITING;
  int next_version = 0;
  int completed_version = -1;
  int failures = 0;
  int pending_inputs = 0;       // outstanding EXTERNAL (file) inputs; streamed edges are not counted
  int clique = -1;              // start clique (gang) id, -1 = starts alone
  bool removed = false;         // replaced by a dynamic resize of its stage
  bool duplicate_requested = false;
  std::vector<Attempt> running;
  std::vector<int> inputs;      // producer vertex ids (with multiplicity)
  std::vector<uint8_t> input_fifo;
  std::vector<int> consumers;
  std::vector<uint8_t> consumer_fifo;
  double seconds = 0, bytes_in = 0;
};

struct Sample { double bytes, seconds; };

struct Stage {
  std::string name;
  int nparts;
  std::vector<int> vids;        // vertex ids of the live partitions
  std::vector<Sample> samples;
  double startup = 0, per_byte = 0, sigma = 0, inlier_max = 0;
  bool have_model = false;
  int completed = 0;
};

struct Conn { int src, dst, kind, channel; };

struct GM {
  std::mutex mu;
  std::vector<Stage> stages;
  std::vector<Vertex> verts;
  std::vector<Conn> conns;
  std::vector<std::vector<int>> cliques;
  int live = 0;                 // vertices that are not `removed`
  int jstate = J_NOT_STARTED;
  int max_failures = 6;
  int max_running_versions = 3;
  bool speculative = true;
  double default_outlier_s = 600.0, min_outlier_s = 10.0, nonparam_fraction = 0.5;
  int completed = 0;
  int total_failures = 0, duplicates_started = 0, invalidations = 0, gang_restarts = 0;
  std::string error;
};

void make_ready_if_possible(GM* g, Vertex& v) {
  if (v.removed) return;
  if (v.clique < 0) {
    if (v.state == V_WAITING && v.pending_inputs == 0) v.state = V_READY;
    return;
  }
  // DrStartClique::StartVersionIfReady: the whole clique becomes runnable at once, and only when
  // no member still waits for an external input
  const std::vector<int>& m = g->cliques[v.clique];
  for (int id : m) {
    const Vertex& o = g->verts[id];
    if (o.state == V_COMPLETED) continue;
    if (o.state != V_WAITING || o.pending_inputs != 0) return;
  }
  for (int id : m) if (g->verts[id].state == V_WAITING) g->verts[id].state = V_READY;
}

// Take back a completed vertex's output (it has to run again); consumers that have not completed
// wait for it once more.
void revoke(GM* g, Vertex& up) {
  if (up.state != V_COMPLETED) return;
  up.state = V_WAITING; up.completed_version = -1; g->completed--; g->stages[up.stage].completed--;
  g->invalidations++;
  for (size_t i = 0; i < up.consumers.size(); ++i) {
    if (up.consumer_fifo[i]) continue;
    Vertex& cv = g->verts[up.consumers[i]];
    if (cv.state != V_COMPLETED) { cv.pending_inputs++; if (cv.state == V_READY || (cv.state == V_RUNNING && cv.running.empty())) cv.state = V_WAITING; }
  }
}

int add_edges(GM* g, const Conn& c) {
  Stage& S = g->stages[c.src]; Stage& Dd = g->stages[c.dst];
  const bool fifo = c.channel != 0;
  auto edge = [&](int sp, int dp) {
    Vertex& sv = g->verts[S.vids[sp]]; Vertex& dv = g->verts[Dd.vids[dp]];
    dv.inputs.push_back(sv.id); dv.input_fifo.push_back(fifo); sv.consumers.push_back(dv.id); sv.consumer_fifo.push_back(fifo);
    if (!fifo && sv.state != V_COMPLETED) dv.pending_inputs++;
  };
  if (c.kind == 1) { for (int a = 0; a < S.nparts; ++a) for (int b = 0; b < Dd.nparts; ++b) edge(a, b); }
  else if (S.nparts == Dd.nparts) { for (int i = 0; i < S.nparts; ++i) edge(i, i); }
  else if (S.nparts == 1) { for (int b = 0; b < Dd.nparts; ++b) edge(0, b); }
  else if (Dd.nparts == 1) { for (int a = 0; a < S.nparts; ++a) edge(a, 0); }
  else if (S.nparts % Dd.nparts == 0) { int f = S.nparts / Dd.nparts; for (int a = 0; a < S.nparts; ++a) edge(a, a / f); }
  else if (Dd.nparts % S.nparts == 0) { int f = Dd.nparts / S.nparts; for (int b = 0; b < Dd.nparts; ++b) edge(b / f, b); }
  else return -2;               // widths that do not divide cannot be connected pointwise
  return 0;
}

void new_vertices(GM* g, int stage, int nparts) {
  Stage& s = g->stages[stage];
  s.vids.clear(); s.nparts = nparts;
  for (int p = 0; p < nparts; ++p) {
    Vertex v; v.id = (int)g->verts.size(); v.stage = stage; v.part = p;
    g->verts.push_back(v); s.vids.push_back(v.id); g->live++;
  }
}

// Connected components over streamed edges = start cliques.
void build_cliques(GM* g) {
  std::vector<int> parent(g->verts.size());
  for (size_t i = 0; i < parent.size(); ++i) parent[i] = (int)i;
  auto find = [&](int x) { while (parent[x] != x) { parent[x] = parent[parent[x]]; x = parent[x]; } return x; };
  for (auto& v : g->verts) {
    if (v.removed) continue;
    for (size_t i = 0; i < v.inputs.size(); ++i) if (v.input_fifo[i] && !g->verts[v.inputs[i]].removed) parent[find(v.id)] = find(v.inputs[i]);
  }
  std::map<int, std::vector<int>> comp;
  for (auto& v : g->verts) if (!v.removed) comp[find(v.id)].push_back(v.id);
  g->cliques.clear();
  for (auto& v : g->verts) v.clique = -1;
  for (auto& kv : comp) {
    if (kv.second.size() < 2) continue;
    for (int id : kv.second) g->verts[id].clique = (int)g->cliques.size();
    g->cliques.push_back(kv.second);
  }
}

// Robust fit time = a + b*bytes: least squares over random subsamples, keep the fit with the
// smallest median absolute residual; sigma from the inlier residuals.
void refit(Stage& s) {
  size_t n = s.samples.size();
  if (n < 4) { s.have_model = false; return; }
  std::mt19937 rng(12345u + (unsigned)n);
  double best_med = 1e300, ba = 0, bb = 0;
  const int trials = 10;
  for (int t = 0; t < trials; ++t) {
    size_t m = std::max<size_t>(3, (size_t)(n * 0.5));
    std::vector<size_t> idx(n);
    for (size_t i = 0; i < n; ++i) idx[i] = i;
    std::shuffle(idx.begin(), idx.end(), rng);
    double sx = 0, sy = 0, sxx = 0, sxy = 0;
    for (size_t i = 0; i < m; ++i) {
      const Sample& p = s.samples[idx[i]];
      sx += p.bytes; sy += p.seconds; sxx += p.bytes * p.bytes; sxy += p.bytes * p.seconds;
    }
    double den = m * sxx - sx * sx, a, b;
    if (fabs(den) < 1e-9 * (1.0 + sxx)) { b = 0; a = sy / m; }        // ill-conditioned: constant model
    else { b = (m * sxy - sx * sy) / den; a = (sy - b * sx) / m; }
    if (b < 0) { b = 0; a = sy / m; }
    std::vector<double> res(n);
    for (size_t i = 0; i < n; ++i) res[i] = fabs(s.samples[i].seconds - (a + b * s.samples[i].bytes));
    std::nth_element(res.begin(), res.begin() + n / 2, res.end());
    if (res[n / 2] < best_med) { best_med = res[n / 2]; ba = a; bb = b; }
  }
  // sigma over the inliers (residual <= 3 * 1.4826 * MAD)
  double cut = 3.0 * 1.4826 * best_med + 1e-9, ss = 0, imax = 0; int cnt = 0;
  for (auto& p : s.samples) {
    double r = p.seconds - (ba + bb * p.bytes);
    if (fabs(r) <= cut) { ss += r * r; ++cnt; imax = std::max(imax, p.seconds); }
  }
  s.inlier_max = imax;
  s.startup = ba; s.per_byte = bb; s.sigma = cnt > 1 ? sqrt(ss / (cnt - 1)) : best_med;
  s.have_model = true;
}

double outlier_threshold(const GM* g, const Stage& s, double bytes) {
  if (s.nparts <= 10 || !s.have_model || s.completed * 2 < s.nparts)
    return g->default_outlier_s;
  double est = s.startup + s.per_byte * bytes + 3.0 * s.sigma;
  // non-parametric guard: never below the slowest INLIER seen so far times a margin
  est = std::max(est, s.inlier_max * (1.0 + g->nonparam_fraction));
  return std::max(g->min_outlier_s, std::min(est, g->default_outlier_s));
}

void check_done(GM* g) {
  if (g->jstate == J_RUNNING && g->completed == g->live) g->jstate = J_SUCCEEDED;
}

}  // namespace

extern "C" {

void* gm_create(int max_failures, int speculative) {
  GM* g = new GM();
  if (max_failures > 0) g->max_failures = max_failures;
  g->speculative = speculative != 0;
  return g;
}
void gm_destroy(void* h) { delete (GM*)h; }

void gm_set_outlier_params(void* h, double default_s, double min_s, double nonparam_fraction) {
  GM* g = (GM*)h;
  std::lock_guard<std::mutex> l(g->mu);
  g->default_outlier_s = default_s; g->min_outlier_s = min_s; g->nonparam_fraction = nonparam_fraction;
}

int gm_add_stage(void* h, const char* name, int nparts) {
  GM* g = (GM*)h;
  std::lock_guard<std::mutex> l(g->mu);
  Stage s; s.name = name ? name : ""; s.nparts = nparts;
  g->stages.push_back(s);
  new_vertices(g, (int)g->stages.size() - 1, nparts);
  return (int)g->stages.size() - 1;
}

// kind 0 = pointwise (N:N, 1:N broadcast, N:1 gather, k*N:N / N:k*N grouped), 1 = cross product
// (every src -> every dst).  channel 0 = file (the producer's output persists), 1 = FIFO / pipe
// (streamed: producer and consumer belong to one start clique).
int gm_connect_ex(void* h, int src_stage, int dst_stage, int kind, int channel) {
  GM* g = (GM*)h;
  std::lock_guard<std::mutex> l(g->mu);
  if (src_stage < 0 || dst_stage < 0 || src_stage >= (int)g->stages.size() || dst_stage >= (int)g->stages.size()) return -1;
  Conn c{src_stage, dst_stage, kind, channel};
  int rc = add_edges(g, c);
  if (rc == 0) g->conns.push_back(c);
  return rc;
}
int gm_connect(void* h, int src_stage, int dst_stage, int kind) { return gm_connect_ex(h, src_stage, dst_stage, kind, 0); }

int gm_start(void* h) {
  GM* g = (GM*)h;
  std::lock_guard<std::mutex> l(g->mu);
  if (g->jstate != J_NOT_STARTED) return -1;
  g->jstate = g->live == 0 ? J_SUCCEEDED : J_RUNNING;
  build_cliques(g);
  for (auto& v : g->verts) make_ready_if_possible(g, v);
  return 0;
}

// Dynamic resize of a stage that has not started (DrDynamicDistributionManager changes the width
// of a consumer stage once its producers' sizes are known).  With `follow` != 0 the new width is
// propagated to the pointwise stages downstream that had the same width (DrPipelineSplitManager;
// stages fed by other, narrower inputs keep those as re-broadcast side inputs = the semi variant).
// Returns the number of stages resized, or -1 if a vertex of an affected stage already started.
int gm_resize_stage(void* h, int stage, int new_parts, int follow) {
  GM* g = (GM*)h;
  std::lock_guard<std::mutex> l(g->mu);
  if (stage < 0 || stage >= (int)g->stages.size() || new_parts < 1) return -1;
  std::vector<int> todo{stage};
  const int old_parts = g->stages[stage].nparts;
  if (follow)
    for (size_t k = 0; k < todo.size(); ++k)
      for (auto& c : g->conns)
        if (c.src == todo[k] && c.kind == 0 && g->stages[c.dst].nparts == old_parts &&
            std::find(todo.begin(), todo.end(), c.dst) == todo.end()) todo.push_back(c.dst);
  for (int st : todo)
    for (int id : g->stages[st].vids) { const Vertex& v = g->verts[id]; if (v.state != V_WAITING && v.state != V_READY) return -1; }
  auto affected = [&](int st) { return std::find(todo.begin(), todo.end(), st) != todo.end(); };
  // drop every edge that touches a resized stage, retire its vertices, create the new ones
  for (auto& v : g->verts) {
    if (v.removed) continue;
    const bool mine = affected(v.stage);
    std::vector<int> in2; std::vector<uint8_t> inf2; int pend = 0;
    for (size_t i = 0; i < v.inputs.size(); ++i) {
      const Vertex& up = g->verts[v.inputs[i]];
      if (mine || affected(up.stage)) continue;
      in2.push_back(v.inputs[i]); inf2.push_back(v.input_fifo[i]);
      if (!v.input_fifo[i] && up.state != V_COMPLETED) ++pend;
    }
    std::vector<int> c2; std::vector<uint8_t> cf2;
    for (size_t i = 0; i < v.consumers.size(); ++i) {
      if (mine || affected(g->verts[v.consumers[i]].stage)) continue;
      c2.push_back(v.consumers[i]); cf2.push_back(v.consumer_fifo[i]);
    }
    v.inputs.swap(in2); v.input_fifo.swap(inf2); v.consumers.swap(c2); v.consumer_fifo.swap(cf2);
    if (v.state == V_WAITING || v.state == V_READY) { v.pending_inputs = pend; v.state = V_WAITING; }
  }
  for (int st : todo) {
    for (int id : g->stages[st].vids) { g->verts[id].removed = true; g->live--; }
    new_vertices(g, st, new_parts);
  }
  int bad = 0;
  for (auto& c : g->conns) if ((affected(c.src) || affected(c.dst)) && add_edges(g, c) != 0) bad = 1;
  if (bad) { g->jstate = J_FAILED; g->error = "dynamic resize produced widths that cannot be connected pointwise"; return -2; }
  if (g->jstate == J_RUNNING) {
    build_cliques(g);
    for (auto& v : g->verts) make_ready_if_possible(g, v);
  }
  return (int)todo.size();
}

// Graph rewrite WHILE THE JOB RUNS (DrDynamicAggregateManager: an internal vertex takes over a group
// of one consumer's inputs; DrDynamicBroadcastManager: a copy vertex stands between a producer and a
// group of its consumers).  Adds ONE vertex to `stage` (a stage made with 0 partitions for this
// purpose) whose inputs are `inputs`; in every listed consumer, the file edges from those inputs are
// replaced by one edge from the new vertex (appended at the end of the consumer's input list).
// Consumers must not have started.  Returns the new vertex id, or <0 and leaves the graph untouched.
int gm_splice(void* h, int stage, const int* inputs, int n_in, const int* consumers, int n_cons) {
  GM* g = (GM*)h;
  std::lock_guard<std::mutex> l(g->mu);
  if (stage < 0 || stage >= (int)g->stages.size() || n_in < 1 || n_cons < 1) return -1;
  if (g->jstate != J_RUNNING && g->jstate != J_NOT_STARTED) return -1;
  const int nv = (int)g->verts.size();
  for (int i = 0; i < n_in; ++i) {
    if (inputs[i] < 0 || inputs[i] >= nv || g->verts[inputs[i]].removed) return -2;
  }
  for (int c = 0; c < n_cons; ++c) {
    if (consumers[c] < 0 || consumers[c] >= nv) return -3;
    const Vertex& cv = g->verts[consumers[c]];
    if (cv.removed || cv.clique >= 0 || (cv.state != V_WAITING && cv.state != V_READY) || !cv.running.empty()) return -4;
    std::vector<int> left = cv.inputs;                    // every input must be an edge of every consumer
    for (int i = 0; i < n_in; ++i) {
      bool found = false;
      for (size_t k = 0; k < left.size(); ++k)
        if (left[k] == inputs[i] && !cv.input_fifo[k]) { left[k] = -1; found = true; break; }
      if (!found) return -5;
    }
  }
  {
    Vertex v; v.id = nv; v.stage = stage; v.part = g->stages[stage].nparts;
    g->verts.push_back(v);
  }
  g->stages[stage].vids.push_back(nv); g->stages[stage].nparts++; g->live++;
  for (int i = 0; i < n_in; ++i) {
    Vertex& up = g->verts[inputs[i]];
    Vertex& me = g->verts[nv];
    me.inputs.push_back(up.id); me.input_fifo.push_back(0);
    if (up.state != V_COMPLETED) me.pending_inputs++;
    up.consumers.push_back(nv); up.consumer_fifo.push_back(0);
  }
  for (int c = 0; c < n_cons; ++c) {
    Vertex& cv = g->verts[consumers[c]];
    for (int i = 0; i < n_in; ++i) {
      Vertex& up = g->verts[inputs[i]];
      for (size_t k = 0; k < cv.inputs.size(); ++k)
        if (cv.inputs[k] == up.id && !cv.input_fifo[k]) {
          cv.inputs.erase(cv.inputs.begin() + k); cv.input_fifo.erase(cv.input_fifo.begin() + k);
          if (up.state != V_COMPLETED && cv.pending_inputs > 0) cv.pending_inputs--;
          break;
        }
      for (size_t k = 0; k < up.consumers.size(); ++k)
        if (up.consumers[k] == cv.id && !up.consumer_fifo[k]) {
          up.consumers.erase(up.consumers.begin() + k); up.consumer_fifo.erase(up.consumer_fifo.begin() + k);
          break;
        }
    }
    cv.inputs.push_back(nv); cv.input_fifo.push_back(0); cv.pending_inputs++;
    if (cv.state == V_READY) cv.state = V_WAITING;
    g->verts[nv].consumers.push_back(cv.id); g->verts[nv].consumer_fifo.push_back(0);
  }
  if (g->jstate == J_RUNNING) make_ready_if_possible(g, g->verts[nv]);
  return nv;
}

// Producer vertex ids of a vertex, in input order (what `bad_input` of gm_report_failed indexes).
int gm_vertex_inputs(void* h, int vertex, int* out, int cap) {
  GM* g = (GM*)h; std::lock_guard<std::mutex> l(g->mu);
  if (vertex < 0 || vertex >= (int)g->verts.size()) return -1;
  const Vertex& v = g->verts[vertex];
  for (int i = 0; i < (int)v.inputs.size() && i < cap; ++i) out[i] = v.inputs[i];
  return (int)v.inputs.size();
}

int gm_stage_vertices(void* h, int stage, int* out, int cap) {
  GM* g = (GM*)h; std::lock_guard<std::mutex> l(g->mu);
  if (stage < 0 || stage >= (int)g->stages.size()) return -1;
  const Stage& s = g->stages[stage];
  for (int i = 0; i < (int)s.vids.size() && i < cap; ++i) out[i] = s.vids[i];
  return (int)s.vids.size();
}

// Start clique of a vertex: returns its size (1 = starts alone) and writes up to cap member ids.
int gm_vertex_clique(void* h, int vertex, int* out, int cap) {
  GM* g = (GM*)h; std::lock_guard<std::mutex> l(g->mu);
  if (vertex < 0 || vertex >= (int)g->verts.size()) return -1;
  const Vertex& v = g->verts[vertex];
  if (v.clique < 0) { if (cap > 0) out[0] = vertex; return 1; }
  const std::vector<int>& m = g->cliques[v.clique];
  for (int i = 0; i < (int)m.size() && i < cap; ++i) out[i] = m[i];
  return (int)m.size();
}

// Next vertex attempt to launch: a READY vertex, or a RUNNING vertex with a duplicate request.
// Returns 1 and fills (vertex, version), 0 when nothing is runnable now, <0 when the job is over.
int gm_next_ready(void* h, double now, int* vertex, int* version) {
  GM* g = (GM*)h;
  std::lock_guard<std::mutex> l(g->mu);
  if (g->jstate != J_RUNNING) return -1;
  for (auto& v : g->verts) {
    if (v.removed) continue;
    bool dup = v.state == V_RUNNING && v.duplicate_requested && (int)v.running.size() < g->max_running_versions;
    if (v.state == V_READY || dup) {
      Attempt a; a.version = v.next_version++; a.start = now;
      v.running.push_back(a);
      if (dup) { v.duplicate_requested = false; g->duplicates_started++; }
      v.state = V_RUNNING;
      *vertex = v.id; *version = a.version;
      return 1;
    }
  }
  return 0;
}

// Returns 1 if this version won (its outputs are the vertex's outputs), 0 if it lost the race.
int gm_report_completed(void* h, int vertex, int version, double seconds, double bytes_in) {
  GM* g = (GM*)h;
  std::lock_guard<std::mutex> l(g->mu);
  if (vertex < 0 || vertex >= (int)g->verts.size()) return -1;
  Vertex& v = g->verts[vertex];
  auto it = std::find_if(v.running.begin(), v.running.end(), [&](const Attempt& a) { return a.version == version; });
  if (it == v.running.end()) return 0;             // cancelled earlier
  v.running.erase(it);
  if (v.state == V_COMPLETED) return 0;            // another version already won
  v.state = V_COMPLETED; v.completed_version = version; v.seconds = seconds; v.bytes_in = bytes_in;
  v.running.clear();                               // gang cancel of the other versions
  v.duplicate_requested = false;
  Stage& s = g->stages[v.stage];
  s.samples.push_back({bytes_in, seconds});
  s.completed++;
  if (s.completed * 2 >= s.nparts && (s.completed == (s.nparts + 1) / 2 || (s.completed * 20) % std::max(1, s.nparts) == 0 || s.completed == s.nparts))
    refit(s);
  g->completed++;
  for (size_t i = 0; i < v.consumers.size(); ++i) {
    Vertex& cv = g->verts[v.consumers[i]];
    if (!v.consumer_fifo[i] && cv.pending_inputs > 0) cv.pending_inputs--;
    make_ready_if_possible(g, cv);
  }
  check_done(g);
  return 1;
}

// bad_input >= 0: the attempt failed READING input #bad_input -> the producer's output is revoked
// and the producer is re-executed; otherwise the failure counts against this vertex.
// Returns 0 = will retry, 1 = job failed (budget exhausted), 2 = upstream invalidated.
int gm_report_failed(void* h, int vertex, int version, int bad_input, const char* message) {
  GM* g = (GM*)h;
  std::lock_guard<std::mutex> l(g->mu);
  if (vertex < 0 || vertex >= (int)g->verts.size()) return -1;
  Vertex& v = g->verts[vertex];
  auto it = std::find_if(v.running.begin(), v.running.end(), [&](const Attempt& a) { return a.version == version; });
  if (it != v.running.end()) v.running.erase(it);
  if (v.state == V_COMPLETED) return 0;
  g->total_failures++;
  if (bad_input >= 0 && bad_input < (int)v.inputs.size()) {
    Vertex& up = g->verts[v.inputs[bad_input]];
    up.failures++;
    if (up.failures >= g->max_failures) {
      g->jstate = J_FAILED; g->error = std::string("input unreadable too many times: ") + (message ? message : "");
      return 1;
    }
    if (up.state == V_COMPLETED) { revoke(g, up); make_ready_if_possible(g, up); }
    if (v.running.empty()) v.state = V_WAITING;
    make_ready_if_possible(g, v);
    return 2;
  }
  v.failures++;
  if (v.failures >= g->max_failures) {
    g->jstate = J_FAILED;
    g->error = std::string("vertex ") + g->stages[v.stage].name + "[" + std::to_string(v.part) + "] failed " +
               std::to_string(v.failures) + " times: " + (message ? message : "");
    return 1;
  }
  if (v.clique >= 0) {
    // gang semantics: a streamed edge keeps nothing to re-read, so every member that has not
    // finished restarts, and finished members whose streamed consumer has not finished run again
    g->gang_restarts++;
    for (int id : g->cliques[v.clique]) {
      Vertex& o = g->verts[id];
      if (o.state == V_COMPLETED) {
        bool needed = false;
        for (size_t i = 0; i < o.consumers.size(); ++i)
          if (o.consumer_fifo[i] && g->verts[o.consumers[i]].state != V_COMPLETED) needed = true;
        if (needed) revoke(g, o);
      } else { o.running.clear(); o.duplicate_requested = false; o.state = V_WAITING; }
    }
    make_ready_if_possible(g, v);
    return 0;
  }
  if (v.running.empty()) { v.state = V_WAITING; make_ready_if_possible(g, v); }
  return 0;
}

// Straggler check (1 s timer in the reference): request a duplicate for every vertex whose oldest
// running attempt exceeds its stage's outlier threshold.  Returns the number of requests made.
int gm_check_duplicates(void* h, double now) {
  GM* g = (GM*)h;
  std::lock_guard<std::mutex> l(g->mu);
  if (!g->speculative || g->jstate != J_RUNNING) return 0;
  int n = 0;
  for (auto& v : g->verts) {
    if (v.removed || v.clique >= 0) continue;        // gang members are never duplicated alone
    if (v.state != V_RUNNING || v.running.empty() || v.duplicate_requested) continue;
    if ((int)v.running.size() >= g->max_running_versions) continue;
    double thr = outlier_threshold(g, g->stages[v.stage], v.bytes_in);
    double oldest = v.running.front().start;
    if (now - oldest > thr) { v.duplicate_requested = true; ++n; }
  }
  return n;
}

int gm_job_state(void* h) { GM* g = (GM*)h; std::lock_guard<std::mutex> l(g->mu); return g->jstate; }
int gm_vertex_count(void* h) { GM* g = (GM*)h; std::lock_guard<std::mutex> l(g->mu); return (int)g->verts.size(); }
int gm_live_vertex_count(void* h) { GM* g = (GM*)h; std::lock_guard<std::mutex> l(g->mu); return g->live; }
int gm_gang_restarts(void* h) { GM* g = (GM*)h; std::lock_guard<std::mutex> l(g->mu); return g->gang_restarts; }
int gm_vertex_info(void* h, int vertex, int* stage, int* part, int* state, int* completed_version, int* failures) {
  GM* g = (GM*)h; std::lock_guard<std::mutex> l(g->mu);
  if (vertex < 0 || vertex >= (int)g->verts.size()) return -1;
  Vertex& v = g->verts[vertex];
  *stage = v.stage; *part = v.part; *state = v.removed ? 5 : v.state; *completed_version = v.completed_version; *failures = v.failures;
  return 0;
}
int gm_counters(void* h, int* completed, int* failures, int* duplicates, int* invalidations) {
  GM* g = (GM*)h; std::lock_guard<std::mutex> l(g->mu);
  *completed = g->completed; *failures = g->total_failures; *duplicates = g->duplicates_started; *invalidations = g->invalidations;
  return 0;
}
int gm_error(void* h, char* buf, int cap) {
  GM* g = (GM*)h; std::lock_guard<std::mutex> l(g->mu);
  snprintf(buf, (size_t)cap, "%s", g->error.c_str());
  return (int)g->error.size();
}
int gm_stage_model(void* h, int stage, double* startup, double* per_byte, double* sigma, double* threshold_for_bytes, double bytes) {
  GM* g = (GM*)h; std::lock_guard<std::mutex> l(g->mu);
  if (stage < 0 || stage >= (int)g->stages.size()) return -1;
  Stage& s = g->stages[stage];
  *startup = s.startup; *per_byte = s.per_byte; *sigma = s.sigma;
  *threshold_for_bytes = outlier_threshold(g, s, bytes);
  return s.have_model ? 1 : 0;
}

// ---- dynamic managers ---------------------------------------------------------------------------
// Aggregation tree (DrDynamicAggregateManager): group completed producer outputs first by machine,
// then by pod, then overall, bounded by input count and bytes.  Writes group ids per input;
// returns the number of groups.  Inputs larger than skip_bytes stay alone.
int dyn_aggregate_groups(int n, const int64_t* bytes, const int32_t* machine, const int32_t* pod,
                         int max_inputs, int64_t max_bytes, int64_t skip_bytes, int32_t* group_out) {
  std::vector<int> order(n);
  for (int i = 0; i < n; ++i) order[i] = i;
  std::stable_sort(order.begin(), order.end(), [&](int a, int b) {
    if (pod[a] != pod[b]) return pod[a] < pod[b];
    return machine[a] < machine[b];
  });
  int groups = 0, cur = -1, cnt = 0; int64_t acc = 0; int cur_pod = -1;
  for (int k = 0; k < n; ++k) {
    int i = order[k];
    if (bytes[i] > skip_bytes) { group_out[i] = groups++; cur = -1; continue; }
    bool fits = cur >= 0 && cnt < max_inputs && acc + bytes[i] <= max_bytes && pod[i] == cur_pod;
    if (!fits) { cur = groups++; cnt = 0; acc = 0; cur_pod = pod[i]; }
    group_out[i] = cur; cnt++; acc += bytes[i];
  }
  return groups;
}

// Distributor sizing (DrDynamicDistributionManager): consumers = ceil(total bytes / data per vertex).
int dyn_consumer_count(int64_t total_bytes, int64_t data_per_vertex, int min_count, int max_count) {
  if (data_per_vertex <= 0) data_per_vertex = 1ll << 30;
  int64_t c = (total_bytes + data_per_vertex - 1) / data_per_vertex;
  if (c < min_count) c = min_count;
  if (c > max_count) c = max_count;
  return (int)c;
}

// Range distributor sizing (DrDynamicRangeDistributor): the producers are SAMPLERS, so the data
// volume is estimated as sample bytes / sampling rate; the resulting consumer count is also handed
// to the bucketizer as its partition-count argument.
int dyn_range_consumer_count(int64_t sample_bytes, double sampling_rate, int64_t data_per_vertex,
                             int min_count, int max_count) {
  if (sampling_rate <= 0) sampling_rate = 0.001;
  double est = (double)sample_bytes / sampling_rate;
  if (est > 9.0e18) est = 9.0e18;
  return dyn_consumer_count((int64_t)est, data_per_vertex, min_count, max_count);
}

// Broadcast tree (DrDynamicBroadcast.cpp:125-130): >= 5 consumers get floor(sqrt(n)) copy vertices;
// copy_of[i] = which copy feeds consumer i.  Returns the number of copies (0 = direct).
int dyn_broadcast_tree(int consumers, int32_t* copy_of) {
  if (consumers < 5) { for (int i = 0; i < consumers; ++i) copy_of[i] = -1; return 0; }
  int copies = (int)floor(sqrt((double)consumers));
  for (int i = 0; i < consumers; ++i) copy_of[i] = i % copies;
  return copies;
}

}  // extern "C"
