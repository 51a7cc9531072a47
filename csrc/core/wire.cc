#include "core/wire.h"

namespace mfl::wire {

int DTypeSize(int dt) {
  static const int kSize[12] = {1, 2, 4, 8, 1, 2, 4, 8, 4, 8, 2, 2};
  return (dt >= 0 && dt < 12) ? kSize[dt] : 0;
}

static bool ParseDType(std::string_view b, TensorView& t) {
  Reader r(b);
  uint32_t f; int wt; uint64_t v;
  while (!r.done()) {
    if (!r.tag(f, wt)) return false;
    if (wt == kVarint && f >= 1 && f <= 3) {
      if (!r.varint(v)) return false;
      if (f == 1) t.dtype = (int)v;
      else if (f == 2) t.byte_order = (int)v;
      else t.fortran = v != 0;
    } else if (!r.skip(wt)) return false;
  }
  return true;
}

static bool ParseSlot(std::string_view b, DeviceSlot& s) {
  Reader r(b);
  uint32_t f; int wt; uint64_t v;
  s.present = true;
  while (!r.done()) {
    if (!r.tag(f, wt)) return false;
    if (wt == kVarint && f >= 1 && f <= 4) {
      if (!r.varint(v)) return false;
      if (f == 1) s.device_rank = (uint32_t)v;
      else if (f == 2) s.slot = (uint32_t)v;
      else if (f == 3) s.offset_bytes = v;
      else s.version = v;
    } else if (!r.skip(wt)) return false;
  }
  return true;
}

bool ParseTensorSpec(std::string_view bytes, TensorView& t) {
  Reader r(bytes);
  uint32_t f; int wt; uint64_t v; std::string_view s;
  while (!r.done()) {
    if (!r.tag(f, wt)) return false;
    if (f == 1 && wt == kVarint) {
      if (!r.varint(v)) return false;
      t.length = (uint32_t)v;
    } else if (f == 2 && wt == kLen) {  // packed repeated int64
      if (!r.bytes(s)) return false;
      Reader pr(s);
      while (!pr.done()) {
        if (!pr.varint(v)) return false;
        t.dims.push_back((int64_t)v);
      }
    } else if (f == 2 && wt == kVarint) {  // unpacked form
      if (!r.varint(v)) return false;
      t.dims.push_back((int64_t)v);
    } else if (f == 3 && wt == kLen) {
      if (!r.bytes(s) || !ParseDType(s, t)) return false;
    } else if (f == 4 && wt == kLen) {
      if (!r.bytes(t.value)) return false;
    } else if (f == 15 && wt == kLen) {
      if (!r.bytes(s) || !ParseSlot(s, t.slot)) return false;
    } else if (!r.skip(wt)) return false;
  }
  return true;
}

static bool ParseVariable(std::string_view b, VariableView& var) {
  Reader r(b);
  uint32_t f; int wt; uint64_t v; std::string_view s, inner;
  while (!r.done()) {
    if (!r.tag(f, wt)) return false;
    if (f == 1 && wt == kLen) {
      if (!r.bytes(var.name)) return false;
    } else if (f == 2 && wt == kVarint) {
      if (!r.varint(v)) return false;
      var.trainable = v != 0;
    } else if ((f == 3 || f == 4) && wt == kLen) {  // Plaintext/CiphertextTensor{tensor_spec = 1}
      if (!r.bytes(s)) return false;
      var.ciphertext = f == 4;
      Reader tr(s);
      uint32_t tf; int twt;
      while (!tr.done()) {
        if (!tr.tag(tf, twt)) return false;
        if (tf == 1 && twt == kLen) {
          if (!tr.bytes(inner) || !ParseTensorSpec(inner, var.t)) return false;
        } else if (!tr.skip(twt)) return false;
      }
    } else if (!r.skip(wt)) return false;
  }
  return true;
}

bool ParseModel(std::string_view bytes, std::vector<VariableView>& vars) {
  Reader r(bytes);
  uint32_t f; int wt; std::string_view s;
  while (!r.done()) {
    if (!r.tag(f, wt)) return false;
    if (f == 1 && wt == kLen) {
      if (!r.bytes(s)) return false;
      vars.emplace_back();
      if (!ParseVariable(s, vars.back())) return false;
    } else if (!r.skip(wt)) return false;
  }
  return true;
}

void AppendVariable(std::string& out, const VariableView& v, const std::string_view* value) {
  const std::string_view val = value ? *value : v.t.value;
  // DType
  std::string dt;
  put_uint(dt, 1, (uint64_t)v.t.dtype);
  put_uint(dt, 2, (uint64_t)v.t.byte_order);
  put_uint(dt, 3, v.t.fortran ? 1 : 0);
  // TensorSpec header (everything but the payload, which is appended in place)
  std::string ts;
  put_uint(ts, 1, v.t.length);
  if (!v.t.dims.empty()) {
    std::string packed;
    for (int64_t d : v.t.dims) put_varint(packed, (uint64_t)d);
    put_bytes(ts, 2, packed);
  }
  put_bytes(ts, 3, dt);
  std::string slot;
  if (v.t.slot.present) {
    std::string sb;
    put_uint(sb, 1, v.t.slot.device_rank);
    put_uint(sb, 2, v.t.slot.slot);
    put_uint(sb, 3, v.t.slot.offset_bytes);
    put_uint(sb, 4, v.t.slot.version);
    put_bytes(slot, 15, sb);
  }
  const size_t val_field = val.empty() ? 0 : 1 + varint_size(val.size()) + val.size();
  const size_t ts_len = ts.size() + val_field + slot.size();
  const size_t tensor_len = 1 + varint_size(ts_len) + ts_len;  // {Plain,Cipher}textTensor{tensor_spec=1}
  std::string head;
  if (!v.name.empty()) put_bytes(head, 1, v.name);
  put_uint(head, 2, v.trainable ? 1 : 0);
  const size_t var_len = head.size() + 1 + varint_size(tensor_len) + tensor_len;
  out.reserve(out.size() + var_len + 16);
  put_tag(out, 1, kLen);
  put_varint(out, var_len);
  out.append(head);
  put_tag(out, v.ciphertext ? 4 : 3, kLen);
  put_varint(out, tensor_len);
  put_tag(out, 1, kLen);
  put_varint(out, ts_len);
  out.append(ts);
  if (!val.empty()) put_bytes(out, 4, val);
  out.append(slot);
}

bool ParseFederatedModel(std::string_view bytes, uint32_t& contributors, uint32_t& iteration, std::string_view& model) {
  Reader r(bytes);
  uint32_t f; int wt; uint64_t v;
  contributors = iteration = 0;
  model = std::string_view();
  while (!r.done()) {
    if (!r.tag(f, wt)) return false;
    if (f == 1 && wt == kVarint) { if (!r.varint(v)) return false; contributors = (uint32_t)v; }
    else if (f == 2 && wt == kVarint) { if (!r.varint(v)) return false; iteration = (uint32_t)v; }
    else if (f == 3 && wt == kLen) { if (!r.bytes(model)) return false; }
    else if (!r.skip(wt)) return false;
  }
  return true;
}

std::string BuildFederatedModel(uint32_t contributors, uint32_t iteration, std::string_view model) {
  std::string out;
  out.reserve(model.size() + 24);
  put_uint(out, 1, contributors);
  put_uint(out, 2, iteration);
  put_bytes(out, 3, model);
  return out;
}

}  // namespace mfl::wire
