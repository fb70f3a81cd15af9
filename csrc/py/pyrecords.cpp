// CPython extension: the tagged record encoding of dryad_b200/channel/records.py at C speed.
//
// Reference roles: the generated per-type serializers and the binary writer / reader the vertex runtime
// uses for every record that crosses a channel (LinqToDryad/DryadLinqCodeGen.cs:496-1479,
// DryadLinqBinaryWriter.cs:336-546, DryadLinqBinaryReader.cs).  The wire format is the one defined in
// channel/records.py (byte-exact: same tags, compact ints, strings as char count + byte count + UTF-8);
// this module handles the types that make up almost all records — None, bool, int64, float, str, bytes,
// tuple, list, dict, Record, Pair, LineRecord — and hands anything else (big integers, Grouping,
// ForkTuple, GUID, DateTime, Decimal, custom serializers, subclasses) to the Python implementation, value
// by value, so the two always produce and accept the same bytes.
//
//   configure(Record, Pair, LineRecord, encode_fallback, decode_fallback)
//   encode_one(obj) -> bytes
//   encode_many(sequence) -> (data: bytes, offsets: bytes)     offsets = int64[n + 1], record i = data[o[i]:o[i+1]]
//   decode_many(buffer) -> list                               whole records, back to back
#define PY_SSIZE_T_CLEAN
#include <Python.h>

#include <stdint.h>
#include <string.h>

#include <string>
#include <vector>

namespace {

enum Tag { T_NONE = 0, T_FALSE, T_TRUE, T_INT, T_FLOAT, T_STR, T_BYTES, T_TUPLE, T_LIST, T_RECORD, T_PAIR, T_GROUP, T_LINE,
           T_BIGINT, T_DICT, T_FORK };

PyObject* g_record = nullptr;      // types (borrowed for comparison, owned references kept)
PyObject* g_pair = nullptr;
PyObject* g_line = nullptr;
PyObject* g_enc_fallback = nullptr;
PyObject* g_dec_fallback = nullptr;

inline bool put_compact(std::string& out, Py_ssize_t v) {
  if (v < 0 || v >= ((Py_ssize_t)1 << 31)) { PyErr_SetString(PyExc_ValueError, "compact int out of range"); return false; }
  if (v < 0x80) { out.push_back((char)v); return true; }
  const char b[4] = {(char)((v >> 24) | 0x80), (char)((v >> 16) & 0xFF), (char)((v >> 8) & 0xFF), (char)(v & 0xFF)};
  out.append(b, 4);
  return true;
}

inline bool put_string(std::string& out, PyObject* s) {
  Py_ssize_t nb = 0;
  const char* p = PyUnicode_AsUTF8AndSize(s, &nb);
  if (!p) return false;
  if (!put_compact(out, PyUnicode_GET_LENGTH(s)) || !put_compact(out, nb)) return false;
  out.append(p, (size_t)nb);
  return true;
}

bool fallback_encode(PyObject* obj, std::string& out) {
  if (!g_enc_fallback) { PyErr_SetString(PyExc_RuntimeError, "record codec is not configured"); return false; }
  PyObject* b = PyObject_CallOneArg(g_enc_fallback, obj);
  if (!b) return false;
  char* p = nullptr;
  Py_ssize_t n = 0;
  if (PyBytes_AsStringAndSize(b, &p, &n) < 0) { Py_DECREF(b); return false; }
  out.append(p, (size_t)n);
  Py_DECREF(b);
  return true;
}

bool encode(PyObject* obj, std::string& out) {
  if (obj == Py_None) { out.push_back((char)T_NONE); return true; }
  if (obj == Py_True) { out.push_back((char)T_TRUE); return true; }
  if (obj == Py_False) { out.push_back((char)T_FALSE); return true; }
  PyTypeObject* tp = Py_TYPE(obj);
  if (tp == &PyLong_Type) {
    int overflow = 0;
    const long long v = PyLong_AsLongLongAndOverflow(obj, &overflow);
    if (overflow) return fallback_encode(obj, out);            // big integer: text form
    if (v == -1 && PyErr_Occurred()) return false;
    out.push_back((char)T_INT);
    out.append((const char*)&v, 8);                             // little-endian hosts only (x86-64 / aarch64)
    return true;
  }
  if (tp == &PyFloat_Type) {
    const double d = PyFloat_AS_DOUBLE(obj);
    out.push_back((char)T_FLOAT);
    out.append((const char*)&d, 8);
    return true;
  }
  if (tp == &PyUnicode_Type) { out.push_back((char)T_STR); return put_string(out, obj); }
  if (g_line && (PyObject*)tp == g_line) { out.push_back((char)T_LINE); return put_string(out, obj); }
  if (tp == &PyBytes_Type) {
    out.push_back((char)T_BYTES);
    if (!put_compact(out, PyBytes_GET_SIZE(obj))) return false;
    out.append(PyBytes_AS_STRING(obj), (size_t)PyBytes_GET_SIZE(obj));
    return true;
  }
  const bool is_tuple = tp == &PyTuple_Type, is_list = tp == &PyList_Type;
  if (is_tuple || is_list || (g_pair && (PyObject*)tp == g_pair)) {
    if (Py_EnterRecursiveCall(" while encoding a record")) return false;
    bool ok = true;
    if (!is_tuple && !is_list) {
      out.push_back((char)T_PAIR);
      ok = PyTuple_GET_SIZE(obj) == 2 && encode(PyTuple_GET_ITEM(obj, 0), out) && encode(PyTuple_GET_ITEM(obj, 1), out);
      if (!ok && !PyErr_Occurred()) PyErr_SetString(PyExc_ValueError, "a Pair has two fields");
    } else {
      const Py_ssize_t n = is_tuple ? PyTuple_GET_SIZE(obj) : PyList_GET_SIZE(obj);
      out.push_back((char)(is_tuple ? T_TUPLE : T_LIST));
      ok = put_compact(out, n);
      for (Py_ssize_t i = 0; ok && i < n; ++i) {
        if (!is_tuple && PyList_GET_SIZE(obj) != n) { PyErr_SetString(PyExc_RuntimeError, "list changed size while it was encoded"); ok = false; break; }
        ok = encode(is_tuple ? PyTuple_GET_ITEM(obj, i) : PyList_GET_ITEM(obj, i), out);
      }
    }
    Py_LeaveRecursiveCall();
    return ok;
  }
  const bool is_record = g_record && (PyObject*)tp == g_record;
  if (is_record || tp == &PyDict_Type) {
    if (is_record) {                                            // field names must be plain strings
      PyObject *k, *v;
      Py_ssize_t pos = 0;
      while (PyDict_Next(obj, &pos, &k, &v))
        if (Py_TYPE(k) != &PyUnicode_Type) return fallback_encode(obj, out);
    }
    if (Py_EnterRecursiveCall(" while encoding a record")) return false;
    out.push_back((char)(is_record ? T_RECORD : T_DICT));
    bool ok = put_compact(out, PyDict_GET_SIZE(obj));
    PyObject *k, *v;
    Py_ssize_t pos = 0;
    while (ok && PyDict_Next(obj, &pos, &k, &v)) ok = (is_record ? put_string(out, k) : encode(k, out)) && encode(v, out);
    Py_LeaveRecursiveCall();
    return ok;
  }
  return fallback_encode(obj, out);
}

// ---- decoding --------------------------------------------------------------------------------------------------
struct Reader { const unsigned char* p; Py_ssize_t n; Py_ssize_t pos; PyObject* owner; };

inline bool need(Reader& r, Py_ssize_t k) {
  if (k < 0 || r.pos + k > r.n) { PyErr_SetString(PyExc_ValueError, "truncated record stream"); return false; }
  return true;
}

inline bool get_compact(Reader& r, Py_ssize_t* v) {
  if (!need(r, 1)) return false;
  const unsigned b0 = r.p[r.pos];
  if (b0 < 0x80) { *v = b0; r.pos += 1; return true; }
  if (!need(r, 4)) return false;
  *v = ((Py_ssize_t)(b0 & 0x7F) << 24) | ((Py_ssize_t)r.p[r.pos + 1] << 16) | ((Py_ssize_t)r.p[r.pos + 2] << 8) | r.p[r.pos + 3];
  r.pos += 4;
  return true;
}

inline PyObject* get_string(Reader& r) {
  Py_ssize_t nc, nb;
  if (!get_compact(r, &nc) || !get_compact(r, &nb) || !need(r, nb)) return nullptr;
  PyObject* s = PyUnicode_DecodeUTF8((const char*)r.p + r.pos, nb, "strict");
  r.pos += nb;
  return s;
}

PyObject* decode(Reader& r);

PyObject* fallback_decode(Reader& r, Py_ssize_t at) {
  if (!g_dec_fallback) { PyErr_SetString(PyExc_RuntimeError, "record codec is not configured"); return nullptr; }
  PyObject* res = PyObject_CallFunction(g_dec_fallback, "On", r.owner, at);
  if (!res) return nullptr;
  if (!PyTuple_Check(res) || PyTuple_GET_SIZE(res) != 2) { Py_DECREF(res); PyErr_SetString(PyExc_TypeError, "decode fallback must return (object, position)"); return nullptr; }
  const Py_ssize_t np = PyLong_AsSsize_t(PyTuple_GET_ITEM(res, 1));
  if (np < 0 && PyErr_Occurred()) { Py_DECREF(res); return nullptr; }
  if (np <= at || np > r.n) { Py_DECREF(res); PyErr_SetString(PyExc_ValueError, "decode fallback returned a bad position"); return nullptr; }
  PyObject* obj = PyTuple_GET_ITEM(res, 0);
  Py_INCREF(obj);
  Py_DECREF(res);
  r.pos = np;
  return obj;
}

PyObject* decode(Reader& r) {
  if (!need(r, 1)) return nullptr;
  const Py_ssize_t at = r.pos;
  const unsigned t = r.p[r.pos++];
  switch (t) {
    case T_NONE: Py_RETURN_NONE;
    case T_TRUE: Py_RETURN_TRUE;
    case T_FALSE: Py_RETURN_FALSE;
    case T_INT: {
      if (!need(r, 8)) return nullptr;
      long long v; memcpy(&v, r.p + r.pos, 8); r.pos += 8;
      return PyLong_FromLongLong(v);
    }
    case T_FLOAT: {
      if (!need(r, 8)) return nullptr;
      double d; memcpy(&d, r.p + r.pos, 8); r.pos += 8;
      return PyFloat_FromDouble(d);
    }
    case T_STR: return get_string(r);
    case T_LINE: {
      if (!g_line) { r.pos = at; return fallback_decode(r, at); }
      PyObject* s = get_string(r);
      if (!s) return nullptr;
      PyObject* l = PyObject_CallOneArg(g_line, s);
      Py_DECREF(s);
      return l;
    }
    case T_BYTES: {
      Py_ssize_t n;
      if (!get_compact(r, &n) || !need(r, n)) return nullptr;
      PyObject* b = PyBytes_FromStringAndSize((const char*)r.p + r.pos, n);
      r.pos += n;
      return b;
    }
    case T_TUPLE: case T_LIST: {
      Py_ssize_t n;
      if (!get_compact(r, &n)) return nullptr;
      if (n > r.n - r.pos) { PyErr_SetString(PyExc_ValueError, "truncated record stream"); return nullptr; }   // every item takes >= 1 byte
      if (Py_EnterRecursiveCall(" while decoding a record")) return nullptr;
      PyObject* seq = t == T_TUPLE ? PyTuple_New(n) : PyList_New(n);
      for (Py_ssize_t i = 0; seq && i < n; ++i) {
        PyObject* v = decode(r);
        if (!v) { Py_CLEAR(seq); break; }
        if (t == T_TUPLE) PyTuple_SET_ITEM(seq, i, v); else PyList_SET_ITEM(seq, i, v);
      }
      Py_LeaveRecursiveCall();
      return seq;
    }
    case T_PAIR: {
      if (!g_pair) { r.pos = at; return fallback_decode(r, at); }
      if (Py_EnterRecursiveCall(" while decoding a record")) return nullptr;
      PyObject* a = decode(r);
      PyObject* b = a ? decode(r) : nullptr;
      Py_LeaveRecursiveCall();
      PyObject* res = (a && b) ? PyObject_CallFunctionObjArgs(g_pair, a, b, nullptr) : nullptr;
      Py_XDECREF(a); Py_XDECREF(b);
      return res;
    }
    case T_RECORD: case T_DICT: {
      if (t == T_RECORD && !g_record) { r.pos = at; return fallback_decode(r, at); }
      Py_ssize_t n;
      if (!get_compact(r, &n)) return nullptr;
      if (Py_EnterRecursiveCall(" while decoding a record")) return nullptr;
      PyObject* d = t == T_RECORD ? PyObject_CallNoArgs(g_record) : PyDict_New();
      for (Py_ssize_t i = 0; d && i < n; ++i) {
        PyObject* k = t == T_RECORD ? get_string(r) : decode(r);
        PyObject* v = k ? decode(r) : nullptr;
        if (!k || !v || PyDict_SetItem(d, k, v) < 0) { Py_XDECREF(k); Py_XDECREF(v); Py_CLEAR(d); break; }
        Py_DECREF(k); Py_DECREF(v);
      }
      Py_LeaveRecursiveCall();
      return d;
    }
    default:
      r.pos = at;
      return fallback_decode(r, at);
  }
}

// ---- module functions ------------------------------------------------------------------------------------------
PyObject* py_configure(PyObject*, PyObject* args) {
  PyObject *rec, *pair, *line, *enc, *dec;
  if (!PyArg_ParseTuple(args, "OOOOO", &rec, &pair, &line, &enc, &dec)) return nullptr;
  PyObject** slots[5] = {&g_record, &g_pair, &g_line, &g_enc_fallback, &g_dec_fallback};
  PyObject* vals[5] = {rec, pair, line, enc, dec};
  for (int i = 0; i < 5; ++i) {
    PyObject* v = vals[i] == Py_None ? nullptr : vals[i];
    Py_XINCREF(v);
    Py_XSETREF(*slots[i], v);
  }
  Py_RETURN_NONE;
}

PyObject* py_encode_one(PyObject*, PyObject* obj) {
  std::string out;
  out.reserve(64);
  if (!encode(obj, out)) return nullptr;
  return PyBytes_FromStringAndSize(out.data(), (Py_ssize_t)out.size());
}

PyObject* py_encode_many(PyObject*, PyObject* seq) {
  PyObject* it = PyObject_GetIter(seq);
  if (!it) return nullptr;
  std::string out;
  std::vector<int64_t> offs;
  out.reserve(1 << 16);
  offs.push_back(0);
  PyObject* item;
  while ((item = PyIter_Next(it))) {
    const bool ok = encode(item, out);
    Py_DECREF(item);
    if (!ok) { Py_DECREF(it); return nullptr; }
    offs.push_back((int64_t)out.size());
  }
  Py_DECREF(it);
  if (PyErr_Occurred()) return nullptr;
  PyObject* data = PyBytes_FromStringAndSize(out.data(), (Py_ssize_t)out.size());
  PyObject* o = PyBytes_FromStringAndSize((const char*)offs.data(), (Py_ssize_t)(offs.size() * sizeof(int64_t)));
  if (!data || !o) { Py_XDECREF(data); Py_XDECREF(o); return nullptr; }
  PyObject* res = PyTuple_Pack(2, data, o);
  Py_DECREF(data); Py_DECREF(o);
  return res;
}

PyObject* py_decode_many(PyObject*, PyObject* buf) {
  Py_buffer view;
  if (PyObject_GetBuffer(buf, &view, PyBUF_SIMPLE) < 0) return nullptr;
  Reader r{(const unsigned char*)view.buf, view.len, 0, buf};
  PyObject* out = PyList_New(0);
  while (out && r.pos < r.n) {
    PyObject* v = decode(r);
    if (!v || PyList_Append(out, v) < 0) { Py_XDECREF(v); Py_CLEAR(out); break; }
    Py_DECREF(v);
  }
  PyBuffer_Release(&view);
  return out;
}

PyMethodDef methods[] = {
  {"configure", py_configure, METH_VARARGS, "configure(Record, Pair, LineRecord, encode_fallback, decode_fallback)"},
  {"encode_one", py_encode_one, METH_O, "encode_one(obj) -> bytes"},
  {"encode_many", py_encode_many, METH_O, "encode_many(seq) -> (data, int64 offsets as bytes)"},
  {"decode_many", py_decode_many, METH_O, "decode_many(buffer) -> list of records"},
  {nullptr, nullptr, 0, nullptr}};

PyModuleDef moduledef = {PyModuleDef_HEAD_INIT, "_dryad_records", "native record codec", -1, methods, nullptr, nullptr, nullptr, nullptr};

}  // namespace

PyMODINIT_FUNC PyInit__dryad_records(void) { return PyModule_Create(&moduledef); }
