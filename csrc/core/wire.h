// Hand-written protobuf wire codec for the model-carrying messages.
//
// The build image has no C++ protobuf, so the native controller core reads and writes the
// `metisfl.Model` / `FederatedModel` / `TensorSpec` messages (reference schema
// metisfl/proto/model.proto:42-110) directly in wire format.  Parsing is zero-copy: tensor
// payloads stay `string_view`s into the stored request bytes, which is what lets aggregation read
// learner models without the reference's deserialise-copy (proto_tensor_serde.h:14-24).
#pragma once
#include <cstdint>
#include <cstring>
#include <string>
#include <string_view>
#include <vector>

namespace mfl::wire {

enum WireType { kVarint = 0, kFixed64 = 1, kLen = 2, kFixed32 = 5 };

struct Reader {
  const uint8_t* p;
  const uint8_t* end;
  explicit Reader(std::string_view s)
      : p(reinterpret_cast<const uint8_t*>(s.data())), end(p + s.size()) {}
  bool done() const { return p >= end; }
  bool varint(uint64_t& v) {
    v = 0;
    for (int shift = 0; shift < 64 && p < end; shift += 7) {
      const uint8_t b = *p++;
      v |= (uint64_t)(b & 0x7f) << shift;
      if (!(b & 0x80)) return true;
    }
    return false;
  }
  bool tag(uint32_t& field, int& wt) {
    uint64_t t;
    if (!varint(t)) return false;
    field = (uint32_t)(t >> 3);
    wt = (int)(t & 7);
    return field != 0;
  }
  bool bytes(std::string_view& out) {
    uint64_t n;
    if (!varint(n) || n > (uint64_t)(end - p)) return false;
    out = std::string_view(reinterpret_cast<const char*>(p), (size_t)n);
    p += n;
    return true;
  }
  bool skip(int wt) {
    uint64_t v;
    std::string_view s;
    switch (wt) {
      case kVarint: return varint(v);
      case kFixed64: if (end - p < 8) return false; p += 8; return true;
      case kLen: return bytes(s);
      case kFixed32: if (end - p < 4) return false; p += 4; return true;
    }
    return false;
  }
};

inline void put_varint(std::string& out, uint64_t v) {
  while (v >= 0x80) {
    out.push_back((char)(v | 0x80));
    v >>= 7;
  }
  out.push_back((char)v);
}
inline size_t varint_size(uint64_t v) {
  size_t n = 1;
  while (v >= 0x80) { v >>= 7; ++n; }
  return n;
}
inline void put_tag(std::string& out, uint32_t field, int wt) { put_varint(out, ((uint64_t)field << 3) | wt); }
inline void put_uint(std::string& out, uint32_t field, uint64_t v) {
  if (v == 0) return;  // proto3 default elision
  put_tag(out, field, kVarint);
  put_varint(out, v);
}
inline void put_bytes(std::string& out, uint32_t field, std::string_view s) {
  put_tag(out, field, kLen);
  put_varint(out, s.size());
  out.append(s.data(), s.size());
}

struct DeviceSlot {  // extension field TensorSpec.device_slot = 15
  bool present = false;
  uint32_t device_rank = 0, slot = 0;
  uint64_t offset_bytes = 0, version = 0;
};

struct TensorView {
  uint32_t length = 0;
  std::vector<int64_t> dims;
  int dtype = 0;        // metisfl.DType.Type
  int byte_order = 0;   // metisfl.DType.ByteOrder
  bool fortran = false;
  std::string_view value;
  DeviceSlot slot;
};

struct VariableView {
  std::string_view name;
  bool trainable = false;
  bool ciphertext = false;
  TensorView t;
};

bool ParseTensorSpec(std::string_view bytes, TensorView& t);
bool ParseModel(std::string_view bytes, std::vector<VariableView>& vars);
// Appends one `Model.variables` entry (field 1) to `out`; `value` replaces v.t.value when given.
void AppendVariable(std::string& out, const VariableView& v, const std::string_view* value = nullptr);
bool ParseFederatedModel(std::string_view bytes, uint32_t& contributors, uint32_t& iteration, std::string_view& model);
std::string BuildFederatedModel(uint32_t contributors, uint32_t iteration, std::string_view model);

int DTypeSize(int dtype);

}  // namespace mfl::wire
