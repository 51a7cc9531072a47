// K2: fused optimizer step + publish.
//
// Behavioural spec (reference, CPU/TF): the five optimizers selectable from OptimizerConfig
// (model.proto:116-152; keras_model_ops.py:245-283) and FedProx's proximal update
// `var -= lr*grad + lr*mu*(var - vstar)` (optimizers/fed_prox.py:50-60); the upload of the trained
// model that follows a task (learner.py:197-206 -> MarkTaskCompleted).
//
// B200 design: one persistent, 16-byte vectorised launch over the whole flat parameter arena
// (no per-tensor launches).  fp32 master weights / optimizer state stay in HBM; the same pass
// writes the low-precision working copy the next forward reads and - on the last local step -
// the (optionally pre-scaled) copy in the peer-visible publish slot, so "upload" costs one extra
// store per element instead of a serialise + RPC.  Gradients are optionally zeroed in the same
// pass so no memset kernel is needed before the next backward.
#include <algorithm>

#include "common.cuh"

namespace mfl {

struct OptArgs {
  float* master;
  void* model;
  const void* grad;
  float* s1;
  float* s2;
  const float* anchor;
  void* publish;
  size_t numel;
  mfl_optim_cfg cfg;
  float bc1, bc2;  // Adam bias corrections 1-beta^t
  int zero_grad;
};

__device__ __forceinline__ float sgn(float x) { return (x > 0.f) - (x < 0.f); }

template <int KIND>
__device__ __forceinline__ float update_one(const OptArgs& a, float p, float g, float& m, float& v, float anchor) {
  const mfl_optim_cfg& c = a.cfg;
  g *= c.grad_scale;
  if constexpr (KIND == 0) {  // vanilla SGD with L1/L2 regularisation
    g += c.l2 * p + c.l1 * sgn(p);
    return p - c.lr * g;
  } else if constexpr (KIND == 1) {  // momentum SGD (Keras form: m = mu*m - lr*g ; p += m)
    m = c.momentum * m - c.lr * g;
    return p + m;
  } else if constexpr (KIND == 2) {  // FedProx
    return p - c.lr * (g + c.mu * (p - anchor));
  } else {  // Adam / AdamW
    if constexpr (KIND == 4) p *= (1.f - c.lr * c.weight_decay);
    m = c.beta1 * m + (1.f - c.beta1) * g;
    v = c.beta2 * v + (1.f - c.beta2) * g * g;
    const float mhat = m / a.bc1, vhat = v / a.bc2;
    return p - c.lr * mhat / (sqrtf(vhat) + c.eps);
  }
}

template <int DT> __device__ __forceinline__ void load8(const void* base, size_t u, float* f) {
  if constexpr (DT == MFL_FLOAT32) {
    const uint4* p = reinterpret_cast<const uint4*>(base) + 2 * u;
    uint4 a = *p, b = *(p + 1);
    VecTraits<MFL_FLOAT32>::unpack(a, f);
    VecTraits<MFL_FLOAT32>::unpack(b, f + 4);
  } else {
    VecTraits<DT>::unpack(*(reinterpret_cast<const uint4*>(base) + u), f);
  }
}
template <int DT> __device__ __forceinline__ void store8(void* base, size_t u, const float* f) {
  if constexpr (DT == MFL_FLOAT32) {
    uint4* p = reinterpret_cast<uint4*>(base) + 2 * u;
    *p = VecTraits<MFL_FLOAT32>::pack(f);
    *(p + 1) = VecTraits<MFL_FLOAT32>::pack(f + 4);
  } else {
    *(reinterpret_cast<uint4*>(base) + u) = VecTraits<DT>::pack(f);
  }
}
template <int DT> __device__ __forceinline__ void zero8(void* base, size_t u) {
  const uint4 z = make_uint4(0, 0, 0, 0);
  if constexpr (DT == MFL_FLOAT32) {
    uint4* p = reinterpret_cast<uint4*>(base) + 2 * u;
    *p = z; *(p + 1) = z;
  } else {
    *(reinterpret_cast<uint4*>(base) + u) = z;
  }
}

template <int KIND, int MODEL_DT, int GRAD_DT, int PUB_DT>
__global__ void __launch_bounds__(256) optimizer_step_kernel(const __grid_constant__ OptArgs a) {
  const size_t units = a.numel / 8;
  const size_t stride = (size_t)gridDim.x * blockDim.x;
  const float ps = a.cfg.publish_scale;
  for (size_t u = (size_t)blockIdx.x * blockDim.x + threadIdx.x; u < units; u += stride) {
    float p[8], g[8], m[8], v[8], an[8];
    load8<MFL_FLOAT32>(a.master, u, p);
    load8<GRAD_DT>(a.grad, u, g);
    if constexpr (KIND == 1 || KIND >= 3) load8<MFL_FLOAT32>(a.s1, u, m);
    if constexpr (KIND >= 3) load8<MFL_FLOAT32>(a.s2, u, v);
    if constexpr (KIND == 2) load8<MFL_FLOAT32>(a.anchor, u, an);
#pragma unroll
    for (int i = 0; i < 8; ++i) p[i] = update_one<KIND>(a, p[i], g[i], m[i], v[i], an[i]);
    store8<MFL_FLOAT32>(a.master, u, p);
    if constexpr (KIND == 1 || KIND >= 3) store8<MFL_FLOAT32>(a.s1, u, m);
    if constexpr (KIND >= 3) store8<MFL_FLOAT32>(a.s2, u, v);
    if (a.model) store8<MODEL_DT>(a.model, u, p);
    if (a.zero_grad) zero8<GRAD_DT>(const_cast<void*>(a.grad), u);
    if (a.publish) {
      if (ps != 1.f) {
#pragma unroll
        for (int i = 0; i < 8; ++i) p[i] *= ps;
      }
      // the slot is read by peer GPUs: streaming store, no L1 allocation
      if constexpr (PUB_DT == MFL_FLOAT32) {
        uint4* q = reinterpret_cast<uint4*>(a.publish) + 2 * u;
        st_stream(q, VecTraits<MFL_FLOAT32>::pack(p));
        st_stream(q + 1, VecTraits<MFL_FLOAT32>::pack(p + 4));
      } else {
        st_stream(reinterpret_cast<uint4*>(a.publish) + u, VecTraits<PUB_DT>::pack(p));
      }
    }
  }
  if (blockIdx.x == 0 && threadIdx.x < a.numel % 8) {  // scalar tail
    const size_t i = units * 8 + threadIdx.x;
    float p = a.master[i], g = load_as_float(a.grad, GRAD_DT, i), m = 0.f, v = 0.f, an = 0.f;
    if constexpr (KIND == 1 || KIND >= 3) m = a.s1[i];
    if constexpr (KIND >= 3) v = a.s2[i];
    if constexpr (KIND == 2) an = a.anchor[i];
    p = update_one<KIND>(a, p, g, m, v, an);
    a.master[i] = p;
    if constexpr (KIND == 1 || KIND >= 3) a.s1[i] = m;
    if constexpr (KIND >= 3) a.s2[i] = v;
    if (a.model) store_from_float(a.model, MODEL_DT, i, p);
    if (a.zero_grad) store_from_float(const_cast<void*>(a.grad), GRAD_DT, i, 0.f);
    if (a.publish) store_from_float(a.publish, PUB_DT, i, p * ps);
  }
}

// Instantiated combinations: grads share the working copy's dtype (they are autograd's grads of
// it); the publish slot is either that dtype or fp32 (pre-scaled publishing keeps full precision).
template <int KIND, int DT>
static void launch_pub(int pub_dt, const OptArgs& a, int blocks, cudaStream_t st) {
  if (pub_dt == MFL_FLOAT32) optimizer_step_kernel<KIND, DT, DT, MFL_FLOAT32><<<blocks, 256, 0, st>>>(a);
  else optimizer_step_kernel<KIND, DT, DT, DT><<<blocks, 256, 0, st>>>(a);
}
template <int KIND>
static void launch_model(int model_dt, int pub_dt, const OptArgs& a, int blocks, cudaStream_t st) {
  switch (model_dt) {
    case MFL_FLOAT32: launch_pub<KIND, MFL_FLOAT32>(pub_dt, a, blocks, st); break;
    case MFL_FLOAT16: launch_pub<KIND, MFL_FLOAT16>(pub_dt, a, blocks, st); break;
    default: launch_pub<KIND, MFL_BFLOAT16>(pub_dt, a, blocks, st); break;
  }
}

static bool f3(int dt) { return dt == MFL_FLOAT32 || dt == MFL_BFLOAT16 || dt == MFL_FLOAT16; }
static bool al16(const void* p) { return (reinterpret_cast<uintptr_t>(p) & 15) == 0; }

}  // namespace mfl

using namespace mfl;

extern "C" int mfl_optimizer_step(float* master, void* model, int model_dtype, const void* grad, int grad_dtype,
                                  float* state1, float* state2, const float* anchor, void* publish,
                                  int publish_dtype, size_t numel, const mfl_optim_cfg* cfg, int zero_grad,
                                  void* stream) {
  MFL_ARG(master && grad && cfg, "optimizer_step: master, grad and cfg are required");
  MFL_ARG(cfg->kind >= 0 && cfg->kind <= 4, "optimizer_step: unknown optimizer kind %d", cfg->kind);
  MFL_ARG(f3(grad_dtype) && (!model || f3(model_dtype)) && (!publish || f3(publish_dtype)),
          "optimizer_step: dtypes must be FLOAT32/FLOAT16/BFLOAT16");
  MFL_ARG((cfg->kind != 1 && cfg->kind < 3) || state1, "optimizer_step: state1 (momentum / m) is required");
  MFL_ARG(cfg->kind < 3 || state2, "optimizer_step: state2 (v) is required for Adam/AdamW");
  MFL_ARG(cfg->kind != 2 || anchor, "optimizer_step: FedProx needs the anchor (community) weights");
  MFL_ARG(al16(master) && al16(model) && al16(grad) && al16(state1) && al16(state2) && al16(anchor) && al16(publish),
          "optimizer_step: all arenas must be 16-byte aligned");
  if (numel == 0) return 0;
  OptArgs a;
  a.master = master; a.model = model; a.grad = grad; a.s1 = state1; a.s2 = state2; a.anchor = anchor;
  a.publish = publish; a.numel = numel; a.cfg = *cfg; a.zero_grad = zero_grad;
  const int t = cfg->step > 0 ? cfg->step : 1;
  a.bc1 = 1.f - powf(cfg->beta1, (float)t);
  a.bc2 = 1.f - powf(cfg->beta2, (float)t);
  const int md = grad_dtype, pd = publish ? publish_dtype : md;
  MFL_ARG(!model || model_dtype == grad_dtype, "optimizer_step: model and grad dtypes must match (%d vs %d)",
          model_dtype, grad_dtype);
  MFL_ARG(pd == md || pd == MFL_FLOAT32, "optimizer_step: publish dtype must be the model dtype or FLOAT32");
  const int blocks = (int)std::max<size_t>(1, std::min<size_t>((numel / 8 + 255) / 256, (size_t)sm_count() * 4));
  cudaStream_t st = (cudaStream_t)stream;
  switch (cfg->kind) {
    case 0: launch_model<0>(md, pd, a, blocks, st); break;
    case 1: launch_model<1>(md, pd, a, blocks, st); break;
    case 2: launch_model<2>(md, pd, a, blocks, st); break;
    case 3: launch_model<3>(md, pd, a, blocks, st); break;
    default: launch_model<4>(md, pd, a, blocks, st); break;
  }
  return check(cudaGetLastError(), "optimizer_step_kernel launch");
}
