// autograd_bn.cpp — the SyncBatchNorm autograd node as a C++ torch::autograd::Function.
//
// Host plumbing only: it does what syncbn.py's _FusedSyncBN does (layout checks, the two small
// allocations, the sequence number, saving tensors for the backward) and then calls the same
// two C-ABI entry points, b200dp_bn_forward / b200dp_bn_backward (include/b200dp.h), which it
// finds with dlsym in the libb200dp.so handed to bind().  There is no arithmetic here and no
// other path to the GPU.  Why it exists: a python autograd.Function costs ~25 us per layer and
// direction (and its backward needs the GIL on the autograd thread); a ResNet-18 on CIFAR
// images runs a BN layer in about that time, so the python node, not the GPU, set the step
// time of the small BASELINE configs (TORCH/nn/modules/_functions.py:7-209 is the node this
// replaces in the reference stack; it is python there too, with five ATen calls and two
// collectives per layer).
#include <dlfcn.h>
#include <torch/extension.h>
#include <c10/cuda/CUDAGuard.h>
#include <c10/cuda/CUDAStream.h>

#include <atomic>
#include <map>
#include <mutex>

namespace {

using at::Tensor;
using torch::autograd::AutogradContext;
using torch::autograd::variable_list;

typedef int (*fwd_fn)(const void*, const void*, void*, int, int64_t, int, void*, float*, const float*,
                      const float*, float, float, float*, float*, float*, int, void*, uint32_t, void*,
                      uint32_t*, int64_t*);
typedef int (*bwd_fn)(const void*, const void*, const void*, void*, void*, int, int64_t, int, const float*,
                      const float*, const float*, int, void*, float*, float*, float*, void*, uint32_t,
                      void*);
typedef uint32_t (*seq_fn)(void*);
typedef size_t (*scratch_fn)(int);
typedef const char* (*errstr_fn)(int);

fwd_fn g_fwd = nullptr;
bwd_fn g_bwd = nullptr;
seq_fn g_seq = nullptr;
scratch_fn g_scratch_bytes = nullptr;
errstr_fn g_errstr = nullptr;
std::atomic<long long> g_fwd_calls{0}, g_bwd_calls{0};

std::mutex g_scratch_mu;
// zero-filled reduction scratch per (device, stream); never destroyed: a static destructor would hand
// CUDA memory back after torch's allocator has gone at interpreter exit
auto* g_scratch = new std::map<std::pair<int, void*>, Tensor>();

void* bn_scratch(const at::Device& dev, void* stream) {
  std::lock_guard<std::mutex> lk(g_scratch_mu);
  auto key = std::make_pair((int)dev.index(), stream);
  auto it = g_scratch->find(key);
  if (it == g_scratch->end()) {
    const int64_t n = (int64_t)g_scratch_bytes(0);
    it = g_scratch->emplace(key, at::zeros({n}, at::TensorOptions().dtype(at::kByte).device(dev))).first;
  }
  return it->second.data_ptr();
}

void check(int rc, const char* what) {
  if (rc == 0) return;
  TORCH_CHECK(false, what, " failed with code ", rc, ": ", g_errstr ? g_errstr(rc) : "?");
}

int dtype_code(const Tensor& t) {
  if (t.scalar_type() == at::kFloat) return 0;
  if (t.scalar_type() == at::kBFloat16) return 1;
  TORCH_CHECK(false, "libb200dp handles float32 and bfloat16 tensors, not ", t.scalar_type());
}

// x laid out so that its memory is [M, C] row-major (channels last)
Tensor as_rows(const Tensor& x) {
  switch (x.dim()) {
    case 4: return x.is_contiguous(at::MemoryFormat::ChannelsLast) ? x : x.contiguous(at::MemoryFormat::ChannelsLast);
    case 2: return x.contiguous();
    case 5: return x.contiguous(at::MemoryFormat::ChannelsLast3d);
    case 3: return x.unsqueeze(-1).contiguous(at::MemoryFormat::ChannelsLast).squeeze(-1);
    default: TORCH_CHECK_VALUE(false, "expected at least 2D input (got ", x.dim(), "D input)");
  }
}

Tensor f32_view(const c10::optional<Tensor>& t) {
  if (!t.has_value() || !t->defined()) return Tensor();
  if (t->scalar_type() == at::kFloat && t->is_contiguous()) return t->detach();
  return t->detach().to(at::kFloat).contiguous();
}

inline void* ptr(const Tensor& t) { return t.defined() ? t.data_ptr() : nullptr; }
inline void* ptr(const c10::optional<Tensor>& t) { return (t.has_value() && t->defined()) ? t->data_ptr() : nullptr; }

class FusedSyncBN : public torch::autograd::Function<FusedSyncBN> {
 public:
  // running_mean / running_var travel as addresses: they are updated in place by the kernel and
  // never take part in autograd (the python node passes data_ptr() too).
  static Tensor forward(AutogradContext* ctx, const Tensor& x_in, const c10::optional<Tensor>& residual_in,
                        const c10::optional<Tensor>& weight, const c10::optional<Tensor>& bias,
                        int64_t running_mean, int64_t running_var, double eps, double momentum, bool relu,
                        int64_t comm, int64_t world, bool mask_words, int64_t batches_tracked) {
    TORCH_CHECK(g_fwd != nullptr, "_b200dp_autograd.bind() has not been called");
    Tensor x = as_rows(x_in);
    const int64_t C = x.size(1);
    const int64_t M = C ? x.numel() / C : 0;
    TORCH_CHECK_VALUE(!(M == 1 && world < 2), "Expected more than 1 value per channel when training, got input size ", M);
    const int dt = dtype_code(x);
    const bool has_res = residual_in.has_value() && residual_in->defined();
    Tensor residual;
    if (has_res) {
      TORCH_CHECK_VALUE(residual_in->sizes() == x.sizes() && residual_in->scalar_type() == x.scalar_type(),
                        "residual must have the shape and dtype of the input");
      residual = as_rows(*residual_in);
    }
    Tensor w32 = f32_view(weight), b32 = f32_view(bias);
    c10::cuda::CUDAGuard guard(x.device());
    void* stream = (void*)c10::cuda::getCurrentCUDAStream(x.device().index()).stream();
    // one fp32 allocation: [stats_local (2C+1) | pad | save = mean, invstd, count (2C+1)]
    const int64_t off = (2 * C + 1 + 3) / 4 * 4;
    Tensor aux = at::empty({off + 2 * C + 1}, x.options().dtype(at::kFloat));
    Tensor y = at::empty_like(x);
    float* aux_p = aux.data_ptr<float>();
    float* sv = aux_p + off;
    Tensor maskw;
    if (mask_words && relu && has_res && dt == 1 && at::GradMode::is_enabled())
      maskw = at::empty({x.numel() / 8}, x.options().dtype(at::kInt));
    void* h = world > 1 ? (void*)comm : nullptr;
    const uint32_t seq = world > 1 ? g_seq(h) : 0u;
    g_fwd_calls.fetch_add(1, std::memory_order_relaxed);
    check(g_fwd(x.data_ptr(), ptr(residual), y.data_ptr(), dt, M, (int)C, bn_scratch(x.device(), stream), aux_p,
                (const float*)ptr(w32), (const float*)ptr(b32), (float)eps, (float)momentum, (float*)running_mean,
                (float*)running_var, sv, relu ? 1 : 0, h, seq, stream, (uint32_t*)ptr(maskw),
                (int64_t*)batches_tracked),
          "b200dp_bn_forward");
    // ReLU mask in the backward: recomputed from x when nothing was added, else from y (or the mask words)
    const int relu_mode = !relu ? 0 : (!has_res ? 1 : (maskw.defined() ? 3 : 2));
    ctx->saved_data["M"] = M;
    ctx->saved_data["C"] = C;
    ctx->saved_data["dt"] = (int64_t)dt;
    ctx->saved_data["off"] = off;
    ctx->saved_data["relu_mode"] = (int64_t)relu_mode;
    ctx->saved_data["has_res"] = has_res;
    ctx->saved_data["comm"] = comm;
    ctx->saved_data["world"] = world;
    const bool has_w = weight.has_value() && weight->defined(), has_b = bias.has_value() && bias->defined();
    ctx->saved_data["has_w"] = has_w;
    ctx->saved_data["has_b"] = has_b;
    ctx->saved_data["w_dtype"] = (int64_t)(has_w ? weight->scalar_type() : at::kFloat);
    ctx->saved_data["b_dtype"] = (int64_t)(has_b ? bias->scalar_type() : at::kFloat);
    ctx->save_for_backward({x, relu_mode == 2 ? y : maskw, w32, b32, aux});
    return y;
  }

  static variable_list backward(AutogradContext* ctx, variable_list grads) {
    auto saved = ctx->get_saved_variables();
    const Tensor &x = saved[0], &y = saved[1], &w32 = saved[2], &b32 = saved[3], &aux = saved[4];
    const int64_t M = ctx->saved_data["M"].toInt(), C = ctx->saved_data["C"].toInt();
    const int dt = (int)ctx->saved_data["dt"].toInt();
    const int64_t off = ctx->saved_data["off"].toInt();
    const int relu_mode = (int)ctx->saved_data["relu_mode"].toInt();
    const bool has_res = ctx->saved_data["has_res"].toBool();
    const bool has_w = ctx->saved_data["has_w"].toBool(), has_b = ctx->saved_data["has_b"].toBool();
    const int64_t world = ctx->saved_data["world"].toInt();
    // autograd edges exist for the tensor arguments that were passed, in argument order
    size_t edge = 1;
    const size_t e_res = has_res ? edge++ : 0, e_w = has_w ? edge++ : 0, e_b = has_b ? edge++ : 0;
    Tensor dy = grads[0];
    if (dy.scalar_type() != x.scalar_type()) dy = dy.to(x.scalar_type());
    dy = as_rows(dy);
    c10::cuda::CUDAGuard guard(x.device());
    void* stream = (void*)c10::cuda::getCurrentCUDAStream(x.device().index()).stream();
    // one fp32 allocation: [sums (2C) | grad_weight (C) | grad_bias (C)]
    Tensor out = at::empty({4 * C}, x.options().dtype(at::kFloat));
    Tensor dx = at::empty_like(x);
    Tensor dres;
    if (has_res && ctx->needs_input_grad(e_res)) dres = at::empty_like(x);
    void* h = world > 1 ? (void*)ctx->saved_data["comm"].toInt() : nullptr;
    const uint32_t seq = world > 1 ? g_seq(h) : 0u;
    float* out_p = out.data_ptr<float>();
    const float* sv = aux.data_ptr<float>() + off;
    g_bwd_calls.fetch_add(1, std::memory_order_relaxed);
    check(g_bwd(dy.data_ptr(), x.data_ptr(), ptr(y), dx.data_ptr(), ptr(dres), dt, M, (int)C, sv,
                (const float*)ptr(w32), (const float*)ptr(b32), relu_mode, bn_scratch(x.device(), stream), out_p,
                has_w ? out_p + 2 * C : nullptr, has_b ? out_p + 3 * C : nullptr, h, seq, stream),
          "b200dp_bn_backward");
    Tensor gw, gb;
    if (has_w && ctx->needs_input_grad(e_w)) {
      gw = out.narrow(0, 2 * C, C);
      const auto wt = (at::ScalarType)ctx->saved_data["w_dtype"].toInt();
      if (wt != at::kFloat) gw = gw.to(wt);
    }
    if (has_b && ctx->needs_input_grad(e_b)) {
      gb = out.narrow(0, 3 * C, C);
      const auto bt = (at::ScalarType)ctx->saved_data["b_dtype"].toInt();
      if (bt != at::kFloat) gb = gb.to(bt);
    }
    return {dx, dres, gw, gb, Tensor(), Tensor(), Tensor(), Tensor(), Tensor(), Tensor(), Tensor(), Tensor(), Tensor()};
  }
};

Tensor fused_syncbn(const Tensor& x, const c10::optional<Tensor>& residual, const c10::optional<Tensor>& weight,
                    const c10::optional<Tensor>& bias, int64_t running_mean, int64_t running_var, double eps,
                    double momentum, bool relu, int64_t comm, int64_t world, bool mask_words,
                    int64_t batches_tracked) {
  // batches_tracked: address of the layer's int64 num_batches_tracked on the device (0: the caller counts)
  return FusedSyncBN::apply(x, residual, weight, bias, running_mean, running_var, eps, momentum, relu, comm, world,
                            mask_words, batches_tracked);
}

void bind(const std::string& lib_path) {
  void* lib = dlopen(lib_path.c_str(), RTLD_NOW | RTLD_GLOBAL);
  TORCH_CHECK(lib != nullptr, "cannot load ", lib_path, ": ", dlerror());
  g_fwd = (fwd_fn)dlsym(lib, "b200dp_bn_forward");
  g_bwd = (bwd_fn)dlsym(lib, "b200dp_bn_backward");
  g_seq = (seq_fn)dlsym(lib, "b200dp_comm_next_seq");
  g_scratch_bytes = (scratch_fn)dlsym(lib, "b200dp_bn_scratch_bytes");
  g_errstr = (errstr_fn)dlsym(lib, "b200dp_error_string");
  TORCH_CHECK(g_fwd && g_bwd && g_seq && g_scratch_bytes && g_errstr, lib_path, " does not export the SyncBN entry points");
}

std::pair<long long, long long> counters() { return {g_fwd_calls.load(), g_bwd_calls.load()}; }

}  // namespace

PYBIND11_MODULE(TORCH_EXTENSION_NAME, m) {
  m.def("bind", &bind, "dlopen libb200dp.so and resolve the SyncBN entry points");
  m.def("fused_syncbn", &fused_syncbn, "y = act(bn(x) [+ residual]) with batch statistics shared across ranks");
  m.def("counters", &counters, "(forward calls, backward calls) made through this module");
}
