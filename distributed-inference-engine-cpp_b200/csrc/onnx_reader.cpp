// See onnx_reader.h.  Wire-format decoder: varint (0), fixed64 (1), length-delimited (2),
// fixed32 (5).  Groups (3/4) do not occur in ONNX files and are rejected.
#include "onnx_reader.h"

#include <fcntl.h>
#include <sys/mman.h>
#include <sys/stat.h>
#include <unistd.h>

#include <cstring>
#include <stdexcept>

namespace rn {
namespace {

struct Span {
  const uint8_t* p;
  const uint8_t* end;
  bool empty() const { return p >= end; }
};

[[noreturn]] void bad(const char* what) { throw std::runtime_error(std::string("onnx: malformed file: ") + what); }

uint64_t varint(Span& s) {
  uint64_t v = 0;
  for (int shift = 0; shift < 70; shift += 7) {
    if (s.p >= s.end) bad("truncated varint");
    uint8_t b = *s.p++;
    v |= uint64_t(b & 0x7F) << shift;
    if (!(b & 0x80)) return v;
  }
  bad("varint too long");
}

struct Field {
  uint32_t num;
  uint32_t wire;
  uint64_t v;   // wire 0: value; wire 1/5: raw little-endian bits
  Span sub;     // wire 2
};

bool next(Span& s, Field& f) {
  if (s.empty()) return false;
  uint64_t key = varint(s);
  f.num = uint32_t(key >> 3);
  f.wire = uint32_t(key & 7);
  switch (f.wire) {
    case 0: f.v = varint(s); break;
    case 1:
      if (s.end - s.p < 8) bad("truncated fixed64");
      std::memcpy(&f.v, s.p, 8); s.p += 8; break;
    case 2: {
      uint64_t n = varint(s);
      if (uint64_t(s.end - s.p) < n) bad("truncated length-delimited field");
      f.sub = Span{s.p, s.p + n};
      s.p += n;
      break;
    }
    case 5: {
      if (s.end - s.p < 4) bad("truncated fixed32");
      uint32_t x; std::memcpy(&x, s.p, 4); f.v = x; s.p += 4; break;
    }
    default: bad("unsupported wire type");
  }
  return true;
}

std::string str(const Span& s) { return std::string(reinterpret_cast<const char*>(s.p), size_t(s.end - s.p)); }

float bits_to_float(uint64_t v) {
  uint32_t x = uint32_t(v);
  float f;
  std::memcpy(&f, &x, 4);
  return f;
}

// repeated int64: either one varint per field occurrence (wire 0) or a packed run (wire 2)
void add_ints(const Field& f, std::vector<int64_t>& out) {
  if (f.wire == 0) {
    out.push_back(int64_t(f.v));
  } else if (f.wire == 2) {
    Span s = f.sub;
    while (!s.empty()) out.push_back(int64_t(varint(s)));
  } else {
    bad("int field with unexpected wire type");
  }
}

void add_floats(const Field& f, std::vector<float>& out) {
  if (f.wire == 5) {
    out.push_back(bits_to_float(f.v));
  } else if (f.wire == 2) {
    size_t n = size_t(f.sub.end - f.sub.p);
    if (n % 4) bad("packed float run not a multiple of 4 bytes");
    size_t at = out.size();
    out.resize(at + n / 4);
    std::memcpy(out.data() + at, f.sub.p, n);
  } else {
    bad("float field with unexpected wire type");
  }
}

OnnxTensor parse_tensor(Span s) {
  OnnxTensor t;
  Span raw{nullptr, nullptr};
  Field f;
  while (next(s, f)) {
    switch (f.num) {
      case 1: add_ints(f, t.dims); break;
      case 2: t.data_type = int(f.v); break;
      case 4: add_floats(f, t.f32); break;
      case 7: add_ints(f, t.i64); break;
      case 8: t.name = str(f.sub); break;
      case 9: raw = f.sub; break;
      case 13: case 14:
        if (f.num == 14 && f.v == 1) throw std::runtime_error("onnx: external tensor data is not supported");
        break;
      default: break;
    }
  }
  int64_t n = t.numel();
  if (t.data_type == 1) {
    if (raw.p) {
      if (int64_t(raw.end - raw.p) != n * 4) bad("raw_data size does not match dims (float)");
      t.f32.resize(size_t(n));
      std::memcpy(t.f32.data(), raw.p, size_t(n) * 4);
    } else if (int64_t(t.f32.size()) != n) {
      bad("float_data size does not match dims");
    }
  } else if (t.data_type == 7) {
    if (raw.p) {
      if (int64_t(raw.end - raw.p) != n * 8) bad("raw_data size does not match dims (int64)");
      t.i64.resize(size_t(n));
      std::memcpy(t.i64.data(), raw.p, size_t(n) * 8);
    } else if (int64_t(t.i64.size()) != n) {
      bad("int64_data size does not match dims");
    }
  } else {
    throw std::runtime_error("onnx: tensor '" + t.name + "' has unsupported data_type " + std::to_string(t.data_type));
  }
  return t;
}

OnnxAttr parse_attr(Span s, std::string& name) {
  OnnxAttr a;
  Field f;
  while (next(s, f)) {
    switch (f.num) {
      case 1: name = str(f.sub); break;
      case 2: a.f = bits_to_float(f.v); a.has_f = true; break;
      case 3: a.i = int64_t(f.v); a.has_i = true; break;
      case 4: a.s = str(f.sub); break;
      case 7: add_floats(f, a.floats); break;
      case 8: add_ints(f, a.ints); break;
      case 20: a.type = int(f.v); break;
      default: break;  // tensors / graphs / doc strings: not needed on this path
    }
  }
  return a;
}

OnnxNode parse_node(Span s) {
  OnnxNode n;
  Field f;
  while (next(s, f)) {
    switch (f.num) {
      case 1: n.inputs.push_back(str(f.sub)); break;
      case 2: n.outputs.push_back(str(f.sub)); break;
      case 3: n.name = str(f.sub); break;
      case 4: n.op = str(f.sub); break;
      case 5: {
        std::string k;
        OnnxAttr a = parse_attr(f.sub, k);
        n.attrs[k] = std::move(a);
        break;
      }
      default: break;
    }
  }
  return n;
}

OnnxValueInfo parse_value_info(Span s) {
  OnnxValueInfo vi;
  Field f;
  while (next(s, f)) {
    if (f.num == 1) vi.name = str(f.sub);
    if (f.num == 2 && f.wire == 2) {                 // TypeProto
      Span ty = f.sub; Field g;
      while (next(ty, g)) {
        if (g.num != 1 || g.wire != 2) continue;     // tensor_type
        Span tt = g.sub; Field h;
        while (next(tt, h)) {
          if (h.num != 2 || h.wire != 2) continue;   // shape
          Span sh = h.sub; Field d;
          while (next(sh, d)) {
            if (d.num != 1 || d.wire != 2) continue; // dim
            int64_t val = -1;                        // dim_param or nothing -> dynamic
            Span dm = d.sub; Field e;
            while (next(dm, e))
              if (e.num == 1 && e.wire == 0) val = int64_t(e.v);
            vi.dims.push_back(val);
          }
        }
      }
    }
  }
  return vi;
}

struct Mapped {
  void* base = MAP_FAILED;
  size_t len = 0;
  ~Mapped() { if (base != MAP_FAILED) munmap(base, len); }
};

}  // namespace

int64_t OnnxNode::attr_i(const std::string& k, int64_t dflt) const {
  auto it = attrs.find(k);
  return (it == attrs.end() || !it->second.has_i) ? dflt : it->second.i;
}
float OnnxNode::attr_f(const std::string& k, float dflt) const {
  auto it = attrs.find(k);
  return (it == attrs.end() || !it->second.has_f) ? dflt : it->second.f;
}
std::vector<int64_t> OnnxNode::attr_ints(const std::string& k, std::vector<int64_t> dflt) const {
  auto it = attrs.find(k);
  return (it == attrs.end() || it->second.ints.empty()) ? dflt : it->second.ints;
}

OnnxModel load_onnx(const std::string& path) {
  int fd = ::open(path.c_str(), O_RDONLY);
  if (fd < 0) throw std::runtime_error("onnx: cannot open model file '" + path + "'");
  struct stat st;
  if (fstat(fd, &st) != 0 || st.st_size == 0) {
    ::close(fd);
    throw std::runtime_error("onnx: model file '" + path + "' is empty or unreadable");
  }
  Mapped m;
  m.len = size_t(st.st_size);
  m.base = mmap(nullptr, m.len, PROT_READ, MAP_PRIVATE, fd, 0);
  ::close(fd);
  if (m.base == MAP_FAILED) throw std::runtime_error("onnx: mmap failed for '" + path + "'");

  OnnxModel model;
  Span top{static_cast<const uint8_t*>(m.base), static_cast<const uint8_t*>(m.base) + m.len};
  Span graph{nullptr, nullptr};
  Field f;
  while (next(top, f)) {
    if (f.num == 1 && f.wire == 0) model.ir_version = int64_t(f.v);
    if (f.num == 7 && f.wire == 2) graph = f.sub;
    if (f.num == 8 && f.wire == 2) {
      Span os = f.sub; Field g;
      std::string domain; int64_t ver = 0;
      while (next(os, g)) {
        if (g.num == 1) domain = str(g.sub);
        if (g.num == 2) ver = int64_t(g.v);
      }
      if (domain.empty() || domain == "ai.onnx") model.opset = std::max(model.opset, ver);
    }
  }
  if (!graph.p) throw std::runtime_error("onnx: '" + path + "' has no graph");

  std::vector<OnnxValueInfo> all_inputs;
  while (next(graph, f)) {
    if (f.wire != 2) continue;
    switch (f.num) {
      case 1: model.nodes.push_back(parse_node(f.sub)); break;
      case 5: {
        OnnxTensor t = parse_tensor(f.sub);
        std::string k = t.name;
        model.initializers.emplace(std::move(k), std::move(t));
        break;
      }
      case 11: all_inputs.push_back(parse_value_info(f.sub)); break;
      case 12: model.outputs.push_back(parse_value_info(f.sub)); break;
      default: break;
    }
  }
  // IR < 4 lists initializers among the graph inputs; ONNX Runtime's session inputs exclude them.
  for (auto& vi : all_inputs)
    if (!model.initializers.count(vi.name)) model.inputs.push_back(std::move(vi));
  if (model.inputs.empty()) throw std::runtime_error("onnx: model has no (non-initializer) input");
  if (model.outputs.empty()) throw std::runtime_error("onnx: model has no output");
  return model;
}

}  // namespace rn
