// Vertex command / status wire format and resource-affinity merging.
//
// Reference behaviour (re-implemented natively):
//  * DryadVertex/VertexHost/system/common/include/dvertexcommand.h:48-243 and
//    GraphManager/vertex/DrVertexCommand.cpp — DVertexCommandBlock (Start / ReOpenChannels /
//    Terminate, vertex id + version, argument list, input and output channel descriptions with
//    URI and expected length) and DVertexStatus (state, per-channel processed bytes and error
//    code), serialised with the tagged type-length-value property format of
//    GraphManager/gang/DrProperty.cpp (16-bit tag ids).
//  * GraphManager/kernel/DrResources.cpp — DrAffinityMerger: affinity weights (bytes) accumulate
//    per locality up the Core/Socket/Computer/Rack/Cluster hierarchy; the best level is the lowest
//    one holding at least a threshold fraction of the total weight.
#include "../../include/dryad_host.h"

#include <stdio.h>
#include <string.h>

#include <map>
#include <string>
#include <vector>

namespace {

enum Tag : uint16_t {
  T_BEGIN = 0x1000, T_END = 0x1001, T_COMMAND = 0x1010, T_VERTEX_ID = 0x1011, T_VERSION = 0x1012,
  T_ARG = 0x1020, T_INPUT_URI = 0x1030, T_INPUT_LEN = 0x1031, T_OUTPUT_URI = 0x1040, T_OUTPUT_HINT = 0x1041,
  T_STATE = 0x1050, T_CH_BYTES = 0x1051, T_CH_ERROR = 0x1052, T_EXIT_CODE = 0x1053, T_MESSAGE = 0x1054,
};

struct Block {
  int32_t command = 0, vertex_id = 0, version = 0, state = 0, exit_code = 0;
  std::vector<std::string> args, in_uri, out_uri;
  std::vector<int64_t> in_len, out_hint, ch_bytes;
  std::vector<int32_t> ch_err;
  std::string message;
  std::string wire;
};

void put_tlv(std::string& w, uint16_t tag, const void* p, uint32_t n) {
  w.append((const char*)&tag, 2); w.append((const char*)&n, 4); w.append((const char*)p, n);
}
void put_i64(std::string& w, uint16_t tag, int64_t v) { put_tlv(w, tag, &v, 8); }
void put_str(std::string& w, uint16_t tag, const std::string& s) { put_tlv(w, tag, s.data(), (uint32_t)s.size()); }

}  // namespace

extern "C" {

void* vc_create(int command, int vertex_id, int version) { Block* b = new Block(); b->command = command; b->vertex_id = vertex_id; b->version = version; return b; }
void vc_destroy(void* h) { delete (Block*)h; }
void vc_add_arg(void* h, const char* a) { ((Block*)h)->args.push_back(a ? a : ""); }
void vc_add_input(void* h, const char* uri, int64_t length) { Block* b = (Block*)h; b->in_uri.push_back(uri ? uri : ""); b->in_len.push_back(length); }
void vc_add_output(void* h, const char* uri, int64_t size_hint) { Block* b = (Block*)h; b->out_uri.push_back(uri ? uri : ""); b->out_hint.push_back(size_hint); }
void vc_set_status(void* h, int state, int exit_code, const char* message) { Block* b = (Block*)h; b->state = state; b->exit_code = exit_code; b->message = message ? message : ""; }
void vc_add_channel_status(void* h, int64_t bytes, int error) { Block* b = (Block*)h; b->ch_bytes.push_back(bytes); b->ch_err.push_back(error); }

// Serialise; returns the wire size, copies up to cap bytes.
int64_t vc_serialize(void* h, void* out, int64_t cap) {
  Block* b = (Block*)h;
  std::string& w = b->wire; w.clear();
  put_i64(w, T_BEGIN, 1);
  put_i64(w, T_COMMAND, b->command); put_i64(w, T_VERTEX_ID, b->vertex_id); put_i64(w, T_VERSION, b->version);
  for (auto& a : b->args) put_str(w, T_ARG, a);
  for (size_t i = 0; i < b->in_uri.size(); ++i) { put_str(w, T_INPUT_URI, b->in_uri[i]); put_i64(w, T_INPUT_LEN, b->in_len[i]); }
  for (size_t i = 0; i < b->out_uri.size(); ++i) { put_str(w, T_OUTPUT_URI, b->out_uri[i]); put_i64(w, T_OUTPUT_HINT, b->out_hint[i]); }
  put_i64(w, T_STATE, b->state); put_i64(w, T_EXIT_CODE, b->exit_code);
  if (!b->message.empty()) put_str(w, T_MESSAGE, b->message);
  for (size_t i = 0; i < b->ch_bytes.size(); ++i) { put_i64(w, T_CH_BYTES, b->ch_bytes[i]); put_i64(w, T_CH_ERROR, b->ch_err[i]); }
  put_i64(w, T_END, 1);
  if (out && cap > 0) memcpy(out, w.data(), (size_t)((int64_t)w.size() < cap ? (int64_t)w.size() : cap));
  return (int64_t)w.size();
}

// Parse; returns a new block or NULL on a malformed buffer (unknown tags are skipped).
void* vc_parse(const void* data, int64_t n) {
  const char* p = (const char*)data; const char* e = p + n;
  Block* b = new Block(); bool began = false, ended = false;
  while (p + 6 <= e) {
    uint16_t tag; uint32_t len; memcpy(&tag, p, 2); memcpy(&len, p + 2, 4); p += 6;
    if (p + len > e) { delete b; return nullptr; }
    int64_t v = 0; if (len == 8) memcpy(&v, p, 8);
    std::string s(p, len);
    switch (tag) {
      case T_BEGIN: began = true; break;
      case T_END: ended = true; break;
      case T_COMMAND: b->command = (int32_t)v; break;
      case T_VERTEX_ID: b->vertex_id = (int32_t)v; break;
      case T_VERSION: b->version = (int32_t)v; break;
      case T_ARG: b->args.push_back(s); break;
      case T_INPUT_URI: b->in_uri.push_back(s); break;
      case T_INPUT_LEN: b->in_len.push_back(v); break;
      case T_OUTPUT_URI: b->out_uri.push_back(s); break;
      case T_OUTPUT_HINT: b->out_hint.push_back(v); break;
      case T_STATE: b->state = (int32_t)v; break;
      case T_EXIT_CODE: b->exit_code = (int32_t)v; break;
      case T_MESSAGE: b->message = s; break;
      case T_CH_BYTES: b->ch_bytes.push_back(v); break;
      case T_CH_ERROR: b->ch_err.push_back((int32_t)v); break;
      default: break;
    }
    p += len;
  }
  if (!began || !ended || p != e || b->in_uri.size() != b->in_len.size() || b->out_uri.size() != b->out_hint.size()) { delete b; return nullptr; }
  return b;
}

int vc_get_header(void* h, int* command, int* vertex_id, int* version, int* state, int* exit_code) {
  Block* b = (Block*)h; *command = b->command; *vertex_id = b->vertex_id; *version = b->version; *state = b->state; *exit_code = b->exit_code; return 0;
}
int vc_count(void* h, int what) {   // 0 args, 1 inputs, 2 outputs, 3 channel statuses
  Block* b = (Block*)h;
  return what == 0 ? (int)b->args.size() : what == 1 ? (int)b->in_uri.size() : what == 2 ? (int)b->out_uri.size() : (int)b->ch_bytes.size();
}
int64_t vc_get_item(void* h, int what, int i, char* buf, int cap) {   // returns the number (length / hint / bytes / error) and copies the string
  Block* b = (Block*)h; const std::string* s = nullptr; int64_t num = 0;
  if (what == 0 && i < (int)b->args.size()) s = &b->args[i];
  else if (what == 1 && i < (int)b->in_uri.size()) { s = &b->in_uri[i]; num = b->in_len[i]; }
  else if (what == 2 && i < (int)b->out_uri.size()) { s = &b->out_uri[i]; num = b->out_hint[i]; }
  else if (what == 3 && i < (int)b->ch_bytes.size()) { num = b->ch_bytes[i]; if (buf && cap > 0) snprintf(buf, (size_t)cap, "%d", b->ch_err[i]); return num; }
  else if (what == 4) s = &b->message;
  if (s && buf && cap > 0) snprintf(buf, (size_t)cap, "%s", s->c_str());
  return num;
}

// Affinity merge: inputs = (computer, weight) pairs; rack_of maps computer -> rack.  Chooses the
// preferred level: 0 = a computer, 1 = a rack, 2 = cluster; *choice = computer or rack id.
int affinity_merge(int n, const int32_t* computer, const int64_t* weight, const int32_t* rack_of, int ncomputers,
                   double threshold, int32_t* choice) {
  std::map<int, int64_t> by_c, by_r; int64_t total = 0;
  for (int i = 0; i < n; ++i) {
    by_c[computer[i]] += weight[i]; total += weight[i];
    int r = computer[i] >= 0 && computer[i] < ncomputers ? rack_of[computer[i]] : -1;
    by_r[r] += weight[i];
  }
  if (total <= 0) { *choice = -1; return 2; }
  int best = -1; int64_t bw = -1;
  for (auto& kv : by_c) if (kv.second > bw) { bw = kv.second; best = kv.first; }
  if ((double)bw >= threshold * (double)total) { *choice = best; return 0; }
  best = -1; bw = -1;
  for (auto& kv : by_r) if (kv.second > bw) { bw = kv.second; best = kv.first; }
  if ((double)bw >= threshold * (double)total) { *choice = best; return 1; }
  *choice = -1; return 2;
}

}  // extern "C"
