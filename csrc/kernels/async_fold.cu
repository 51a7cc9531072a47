// K3 fused over NVLink: one asynchronous community update in ONE launch on ONE GPU.
//
// What it replaces (behaviour, not code): the reference's rolling update, which for every finishing learner
// deserialises and re-serialises whole protobuf tensors on the controller's CPU
//   S += w_new x_new - w_old x_old ; C = T(S / z)     federated_rolling_average_base.cc:175-293,
//                                                     federated_recency.cc:27-94 (FedRec)
// and, for the FedAsync extension,   C = (1 - a_t) C + a_t x   (a_t = alpha * s(staleness)),
// followed by a RunTask that ships C to the learner again (controller.cc:696-759).
//
// B200 design.  Asynchronous rules fold one learner at a time, so a collective launch on all GPUs would make seven
// GPUs wait for one.  Instead ONE rank - the executor - owns the fp32 rolling state (S for FedRec, an fp32 master
// of C for FedAsync; both rules are  state = decay * state + w_new x_new - w_old x_old,  C = cast(state * post)).
// Its kernel reads the finisher's slot (and, for FedRec, the slot holding that learner's previous contribution)
// once over NVLink with peer loads, updates the state in local HBM, and pushes the new community model into EVERY
// GPU's community buffer: one multimem.st through the NVSwitch when the buffer has a multicast address, peer stores
// otherwise.  No other GPU launches anything.  Ordering is by monotonic counters in symmetric memory:
//   * `wait_flag >= wait_value`  - the owner's stream released "task n of learner g is published" (flag_signal);
//   * `version_flags[r] = new_version` - the last CTA releases the new version to every rank once all stores are
//     fenced; learners pick a community buffer by version (snapshot_install below).
// The community model is kept in n_comm >= 3 rotating buffers (version v lives in buffer v % n_comm), so a learner
// that is still copying version v is not overwritten by the fold that produces v + 1 or v + 2; the copy re-reads
// the version at the end and reports a torn read if more than n_comm - 2 folds went by.
#include <algorithm>

#include "common.cuh"

namespace mfl {

constexpr int kFoldUnit = 8;

struct FoldArgs {
  const char* x_new;
  const char* x_old;
  float w_new, w_old, decay, post;
  float* state;
  uint64_t state_off[MFL_MAX_SEGMENTS];  // element offset of each segment's fp32 state
  char* out[MFL_MAX_PEERS];
  char* mc_out;
  int world;
  mfl_segment seg[MFL_MAX_SEGMENTS];
  int n_seg;
  unsigned long long* version_flags[MFL_MAX_PEERS];
  unsigned long long new_version;
  const unsigned long long* wait_flag;
  unsigned long long wait_value;
  unsigned int* done_counter;
};

__device__ __forceinline__ void multimem_st16(void* p, const uint4& v) {
  asm volatile("multimem.st.relaxed.sys.global.v4.f32 [%0], {%1,%2,%3,%4};"
               :: "l"(p), "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w) : "memory");
}

// 16 bytes to every rank's community buffer at byte offset `off`
__device__ __forceinline__ void push16(const FoldArgs& a, size_t off, const uint4& v) {
  if (a.mc_out) {
    multimem_st16(a.mc_out + off, v);
  } else {
    for (int t = 0; t < a.world; ++t) st_stream(a.out[t] + off, v);
  }
}

template <int DT>
__device__ __forceinline__ void fold_segment(const FoldArgs& a, const mfl_segment& sg, float* st) {
  constexpr int kVec = (DT == MFL_FLOAT32) ? 2 : 1;  // 16-byte vectors per 8-element unit
  const size_t units = sg.numel / kFoldUnit;
  const size_t stride = (size_t)gridDim.x * blockDim.x;
  const uint4* xn = reinterpret_cast<const uint4*>(a.x_new + sg.offset_bytes);
  const uint4* xo = a.x_old ? reinterpret_cast<const uint4*>(a.x_old + sg.offset_bytes) : nullptr;
  for (size_t u = (size_t)blockIdx.x * blockDim.x + threadIdx.x; u < units; u += stride) {
    // issue every load of the unit before the first use: the slot loads cross NVLink (~2 us)
    uint4 rn[kVec], ro[kVec];
#pragma unroll
    for (int k = 0; k < kVec; ++k) rn[k] = ld_peer(xn + kVec * u + k);
    if (xo) {
#pragma unroll
      for (int k = 0; k < kVec; ++k) ro[k] = ld_peer(xo + kVec * u + k);
    }
    const uint4 s0 = ld_peer(reinterpret_cast<const uint4*>(st) + 2 * u);
    const uint4 s1 = ld_peer(reinterpret_cast<const uint4*>(st) + 2 * u + 1);
    float s[kFoldUnit], x[kFoldUnit];
    VecTraits<MFL_FLOAT32>::unpack(s0, s);
    VecTraits<MFL_FLOAT32>::unpack(s1, s + 4);
#pragma unroll
    for (int k = 0; k < kVec; ++k) VecTraits<DT>::unpack(rn[k], x + k * VecTraits<DT>::N);
#pragma unroll
    for (int i = 0; i < kFoldUnit; ++i) s[i] = fmaf(a.w_new, x[i], a.decay * s[i]);
    if (xo) {
#pragma unroll
      for (int k = 0; k < kVec; ++k) VecTraits<DT>::unpack(ro[k], x + k * VecTraits<DT>::N);
#pragma unroll
      for (int i = 0; i < kFoldUnit; ++i) s[i] = fmaf(-a.w_old, x[i], s[i]);
    }
    st_stream(reinterpret_cast<uint4*>(st) + 2 * u, VecTraits<MFL_FLOAT32>::pack(s));
    st_stream(reinterpret_cast<uint4*>(st) + 2 * u + 1, VecTraits<MFL_FLOAT32>::pack(s + 4));
#pragma unroll
    for (int i = 0; i < kFoldUnit; ++i) s[i] *= a.post;
#pragma unroll
    for (int k = 0; k < kVec; ++k)
      push16(a, sg.offset_bytes + (kVec * u + k) * sizeof(uint4), VecTraits<DT>::pack(s + k * VecTraits<DT>::N));
  }
  // < 8 trailing elements
  if (blockIdx.x == 0 && threadIdx.x < sg.numel % kFoldUnit) {
    const size_t i = units * kFoldUnit + threadIdx.x;
    float s = a.decay * st[i] + a.w_new * load_as_float(a.x_new + sg.offset_bytes, DT, i);
    if (a.x_old) s -= a.w_old * load_as_float(a.x_old + sg.offset_bytes, DT, i);
    st[i] = s;
    for (int t = 0; t < a.world; ++t) store_from_float(a.out[t] + sg.offset_bytes, DT, i, s * a.post);
  }
}

// integer bookkeeping variables (BatchNorm's num_batches_tracked ...): the newest contribution wins
__device__ __forceinline__ void fold_segment_newest(const FoldArgs& a, const mfl_segment& sg) {
  const size_t nbytes = sg.numel * dtype_size(sg.dtype);
  const size_t vecs = nbytes / sizeof(uint4);
  const size_t stride = (size_t)gridDim.x * blockDim.x;
  const uint4* xn = reinterpret_cast<const uint4*>(a.x_new + sg.offset_bytes);
  for (size_t v = (size_t)blockIdx.x * blockDim.x + threadIdx.x; v < vecs; v += stride)
    push16(a, sg.offset_bytes + v * sizeof(uint4), ld_peer(xn + v));
  if (blockIdx.x == 0 && threadIdx.x < nbytes % sizeof(uint4)) {
    const size_t b = vecs * sizeof(uint4) + threadIdx.x;
    const char c = a.x_new[sg.offset_bytes + b];
    for (int t = 0; t < a.world; ++t) a.out[t][sg.offset_bytes + b] = c;
  }
}

__global__ void __launch_bounds__(256) async_fold_kernel(const __grid_constant__ FoldArgs a) {
  if (a.wait_flag) {  // the finisher's slot is complete once its owner released task number `wait_value`
    if (threadIdx.x == 0) counter_wait_ge(a.wait_flag, a.wait_value);
    __syncthreads();
  }
  for (int s = 0; s < a.n_seg; ++s) {
    const mfl_segment& sg = a.seg[s];
    float* st = a.state + a.state_off[s];
    if (sg.exact || !(sg.dtype == MFL_BFLOAT16 || sg.dtype == MFL_FLOAT32 || sg.dtype == MFL_FLOAT16)) {
      fold_segment_newest(a, sg);
    } else if (sg.dtype == MFL_BFLOAT16) {
      fold_segment<MFL_BFLOAT16>(a, sg, st);
    } else if (sg.dtype == MFL_FLOAT32) {
      fold_segment<MFL_FLOAT32>(a, sg, st);
    } else {
      fold_segment<MFL_FLOAT16>(a, sg, st);
    }
  }
  // publish: every CTA fences its stores, the last one to arrive releases the version to all ranks
  if (!a.done_counter) return;  // local use (mfl_rolling_update): stream order is all the ordering there is
  __syncthreads();
  if (threadIdx.x == 0) {
    __threadfence_system();
    const unsigned int prev = atomicAdd(a.done_counter, 1u);
    if (prev == gridDim.x - 1) {
      atomicExch(a.done_counter, 0u);  // self-resetting: the next fold on this stream starts from zero
      __threadfence_system();
      for (int t = 0; t < a.world; ++t) counter_publish(a.version_flags[t], a.new_version);
    }
  }
}

__global__ void flag_signal_kernel(unsigned long long* flag, unsigned long long value) {
  __threadfence_system();  // everything earlier on this stream is visible before the counter moves
  counter_publish(flag, value);
}

__global__ void flag_wait_kernel(const unsigned long long* flag, unsigned long long value) {
  counter_wait_ge(flag, value);
}

// status[0] = version this install reads, status[1] = number of CTAs that saw a torn read
__global__ void version_snapshot_kernel(const unsigned long long* version_flag, unsigned long long* status) {
  status[0] = counter_read(version_flag);
}

struct InstallArgs {
  const char* comm[8];
  int n_comm;
  const unsigned long long* version_flag;
  unsigned long long* status;
  char* dst;
  size_t nbytes;
  float* master;       // optional fp32 copy of segment 0
  mfl_segment seg0;
};

__global__ void __launch_bounds__(512) snapshot_install_kernel(const __grid_constant__ InstallArgs a) {
  const unsigned long long v = a.status[0];
  const char* src = a.comm[v % (unsigned long long)a.n_comm];
  const size_t vecs = a.nbytes / sizeof(uint4);
  const size_t stride = (size_t)gridDim.x * blockDim.x;
  const size_t lane = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  for (size_t i = lane; i < vecs; i += stride)
    st_stream(reinterpret_cast<uint4*>(a.dst) + i, ld_peer(reinterpret_cast<const uint4*>(src) + i));
  if (a.master) {
    const mfl_segment& sg = a.seg0;
    const char* c = src + sg.offset_bytes;
    const size_t units = sg.numel / kFoldUnit;
    for (size_t u = lane; u < units; u += stride) {
      float f[kFoldUnit];
      if (sg.dtype == MFL_BFLOAT16) VecTraits<MFL_BFLOAT16>::unpack(ld_peer(reinterpret_cast<const uint4*>(c) + u), f);
      else if (sg.dtype == MFL_FLOAT16) VecTraits<MFL_FLOAT16>::unpack(ld_peer(reinterpret_cast<const uint4*>(c) + u), f);
      else {
        VecTraits<MFL_FLOAT32>::unpack(ld_peer(reinterpret_cast<const uint4*>(c) + 2 * u), f);
        VecTraits<MFL_FLOAT32>::unpack(ld_peer(reinterpret_cast<const uint4*>(c) + 2 * u + 1), f + 4);
      }
      st_stream(reinterpret_cast<uint4*>(a.master) + 2 * u, VecTraits<MFL_FLOAT32>::pack(f));
      st_stream(reinterpret_cast<uint4*>(a.master) + 2 * u + 1, VecTraits<MFL_FLOAT32>::pack(f + 4));
    }
    if (blockIdx.x == 0 && threadIdx.x < sg.numel % kFoldUnit) {
      const size_t i = units * kFoldUnit + threadIdx.x;
      a.master[i] = load_as_float(c, sg.dtype, i);
    }
  }
  // the buffer we copied is rewritten by the fold that produces version v + n_comm, which cannot start before
  // version v + n_comm - 1 is out: seeing anything newer than v + n_comm - 2 here means the copy may be torn
  __syncthreads();
  if (threadIdx.x == 0) {
    const unsigned long long now = counter_read(a.version_flag);
    if (now > v + (unsigned long long)(a.n_comm - 2)) atomicAdd(a.status + 1, 1ull);
  }
}

static bool aligned16(const void* p) { return (reinterpret_cast<uintptr_t>(p) & 15) == 0; }

}  // namespace mfl

using namespace mfl;

MFL_SPIN_TIMEOUT_READER(mfl_spin_timeouts_async)

extern "C" int mfl_async_fold(const void* x_new, double w_new, const void* x_old, double w_old, double decay,
                              double post_scale, float* state, const uint64_t* state_elem_offsets,
                              const mfl_segment* segments, int n_segments, void* const* out_bufs, void* mc_out,
                              int world, unsigned long long* const* version_flags, unsigned long long new_version,
                              const unsigned long long* wait_flag, unsigned long long wait_value,
                              unsigned int* done_counter, int n_blocks, void* stream) {
  MFL_ARG(x_new && state && segments && out_bufs, "async_fold: null argument");
  MFL_ARG((version_flags != nullptr) == (done_counter != nullptr), "async_fold: version flags and the completion "
          "counter come together (both, or neither for a purely local update)");
  MFL_ARG(n_segments >= 1 && n_segments <= MFL_MAX_SEGMENTS, "async_fold: n_segments=%d", n_segments);
  MFL_ARG(world >= 1 && world <= MFL_MAX_PEERS, "async_fold: world=%d", world);
  MFL_ARG(n_blocks >= 1 && n_blocks <= 8 * sm_count(), "async_fold: n_blocks=%d", n_blocks);
  FoldArgs a;
  a.x_new = (const char*)x_new;
  a.x_old = (const char*)x_old;
  a.w_new = (float)w_new; a.w_old = (float)w_old; a.decay = (float)decay; a.post = (float)post_scale;
  a.state = state;
  a.mc_out = (char*)mc_out;
  a.world = world;
  a.n_seg = n_segments;
  MFL_ARG(aligned16(x_new) && aligned16(x_old) && aligned16(state) && aligned16(mc_out), "async_fold: misaligned buffer");
  for (int s = 0; s < n_segments; ++s) {
    a.seg[s] = segments[s];
    a.state_off[s] = state_elem_offsets[s];
    MFL_ARG(segments[s].offset_bytes % 16 == 0 && dtype_size(segments[s].dtype) && state_elem_offsets[s] % 4 == 0,
            "async_fold: segment %d misaligned or bad dtype", s);
  }
  for (int t = 0; t < world; ++t) {
    a.out[t] = (char*)out_bufs[t];
    a.version_flags[t] = version_flags ? version_flags[t] : nullptr;
    MFL_ARG(out_bufs[t] && aligned16(out_bufs[t]) && (!version_flags || version_flags[t]), "async_fold: bad output %d", t);
  }
  a.new_version = new_version;
  a.wait_flag = wait_flag;
  a.wait_value = wait_value;
  a.done_counter = done_counter;
  async_fold_kernel<<<n_blocks, 256, 0, (cudaStream_t)stream>>>(a);
  return check(cudaGetLastError(), "async_fold_kernel launch");
}

extern "C" int mfl_flag_signal(unsigned long long* flag, unsigned long long value, void* stream) {
  MFL_ARG(flag, "flag_signal: null flag");
  flag_signal_kernel<<<1, 1, 0, (cudaStream_t)stream>>>(flag, value);
  return check(cudaGetLastError(), "flag_signal_kernel launch");
}

extern "C" int mfl_flag_wait(const unsigned long long* flag, unsigned long long value, void* stream) {
  MFL_ARG(flag, "flag_wait: null flag");
  flag_wait_kernel<<<1, 1, 0, (cudaStream_t)stream>>>(flag, value);
  return check(cudaGetLastError(), "flag_wait_kernel launch");
}

extern "C" int mfl_snapshot_install(const void* const* community_bufs, int n_comm,
                                    const unsigned long long* version_flag, unsigned long long* status, void* dst,
                                    size_t nbytes, float* master_f32, const mfl_segment* segment0, int n_blocks,
                                    void* stream) {
  MFL_ARG(community_bufs && n_comm >= 1 && n_comm <= 8 && version_flag && status && dst, "snapshot_install: bad arguments");
  MFL_ARG(nbytes % 16 == 0 && aligned16(dst), "snapshot_install: size and destination must be 16-byte multiples");
  MFL_ARG(!master_f32 || (segment0 && aligned16(master_f32)), "snapshot_install: master needs segment 0");
  InstallArgs a;
  for (int i = 0; i < n_comm; ++i) {
    a.comm[i] = (const char*)community_bufs[i];
    MFL_ARG(community_bufs[i] && aligned16(community_bufs[i]), "snapshot_install: bad community buffer %d", i);
  }
  a.n_comm = n_comm;
  a.version_flag = version_flag;
  a.status = status;
  a.dst = (char*)dst;
  a.nbytes = nbytes;
  a.master = master_f32;
  if (segment0) a.seg0 = *segment0;
  if (n_blocks < 1) n_blocks = 2 * sm_count();
  cudaStream_t st = (cudaStream_t)stream;
  version_snapshot_kernel<<<1, 1, 0, st>>>(version_flag, status);
  snapshot_install_kernel<<<n_blocks, 512, 0, st>>>(a);
  return check(cudaGetLastError(), "snapshot_install_kernel launch");
}
