// Optional C++ DDP communication hook (opt-in: B2C_CPP_HOOK=1).
//
// The default hook is ddp_hook.allreduce_hook (Python, on the autograd thread).  For host-bound tiny
// models its interpreter time is visible (profiles/README.md, config #2); this translation unit
// registers the same logic as a c10d::CommHookInterface directly on the Reducer, so a bucket goes
// Reducer -> runHook -> b2c_allreduce with no Python frame in between.
//
// libb2c.so stays torch-free: this file is the ONLY place that sees torch headers, it reaches the
// library through the function pointers it is given (the C-ABI of include/b2c.h), and it is built
// separately (ray_lightning_accelerators_b200/_b2c_torch_hook*.so) by __graft_entry__.build().
//
// Buckets must already be reducible without per-bucket set-up: they live in the multicast arena
// (DDP constructed inside Comm.torch_pool(), W >= 4) or go through the staged kernels.
#include <torch/extension.h>
#include <torch/csrc/distributed/c10d/comm.hpp>
#include <torch/csrc/distributed/c10d/reducer.hpp>
#include <ATen/cuda/CUDAEvent.h>
#include <c10/cuda/CUDAGuard.h>
#include <c10/cuda/CUDAStream.h>

#include <deque>

namespace {

using allreduce_fn = int (*)(void* comm, void* buf, size_t count, int dtype, int wire_dtype, float scale, int algo,
                             void* stream);
using error_fn = const char* (*)();

struct HookState {
  allreduce_fn allreduce = nullptr;
  error_fn last_error = nullptr;
  void* comm = nullptr;
  float scale = 1.f;
  int algo = 0;
  size_t inline_bytes = 1 << 20;
  bool side_used = false;
  int64_t buckets = 0;
  c10::optional<c10::cuda::CUDAStream> side;
  // optional per-bucket device timing (bench.py's in-step roofline): event pairs around every call
  struct Timed {
    at::cuda::CUDAEvent start{cudaEventDefault};
    at::cuda::CUDAEvent stop{cudaEventDefault};
    int64_t nbytes = 0;
  };
  bool timing = false;
  std::deque<Timed> timed;
};

class B2cCommHook : public c10d::CommHookInterface {
 public:
  explicit B2cCommHook(std::shared_ptr<HookState> st) : st_(std::move(st)) {}

  c10::intrusive_ptr<c10::ivalue::Future> runHook(c10d::GradBucket& bucket) override {
    at::Tensor& buf = bucket.getBufferRef();
    int dtype = -1;
    if (buf.scalar_type() == at::kFloat) dtype = 0;          // B2C_F32
    else if (buf.scalar_type() == at::kBFloat16) dtype = 1;  // B2C_BF16
    if (dtype < 0 || !buf.is_contiguous()) {
      error_ = dtype < 0 ? std::string("b2c hook: gradients must be float32 or bfloat16, got ") + c10::toString(buf.scalar_type())
                         : std::string("b2c hook: bucket buffer must be contiguous");
      auto bad = c10::make_intrusive<c10::ivalue::Future>(c10::TensorType::get());
      finish(bad, buf);
      return bad;
    }
    const size_t nbytes = buf.numel() * buf.element_size();
    st_->buckets++;
    if (!buf.is_cuda()) {  // plumbing tests on a CPU box: the callee decides what a NULL stream means
      call(buf, dtype, nullptr);
      auto fut = c10::make_intrusive<c10::ivalue::Future>(c10::TensorType::get());
      finish(fut, buf);
      return fut;
    }
    const auto dev = buf.device();
    auto cur = c10::cuda::getCurrentCUDAStream(dev.index());
    std::vector<c10::Device> devices{dev};
    if (nbytes <= st_->inline_bytes) {
      // latency-bound bucket: straight onto the backward stream (keep the comm stream-ordered first)
      if (st_->side_used) {
        at::cuda::CUDAEvent ev;
        ev.record(*st_->side);
        ev.block(cur);
      }
      timed_call(buf, dtype, cur, nbytes);
      auto fut = c10::make_intrusive<c10::ivalue::Future>(c10::TensorType::get(), devices);
      finish(fut, buf);
      return fut;
    }
    if (!st_->side) st_->side = c10::cuda::getStreamFromPool(/*isHighPriority=*/true, dev.index());
    st_->side_used = true;
    at::cuda::CUDAEvent ready;
    ready.record(cur);
    ready.block(*st_->side);
    c10::cuda::CUDAStreamGuard guard(*st_->side);  // the Future records its event on the current stream
    timed_call(buf, dtype, *st_->side, nbytes);
    auto fut = c10::make_intrusive<c10::ivalue::Future>(c10::TensorType::get(), devices);
    finish(fut, buf);
    return fut;
  }

  at::Tensor parseHookResult(const c10::IValue& result) override {
    TORCH_INTERNAL_ASSERT(result.isTensor(), "b2c hook: expected a tensor result");
    return result.toTensor();
  }

 private:
  // Never throws: an exception escaping runHook unwinds through the autograd engine's hook machinery and
  // was seen to crash the process; a failed call instead completes the Future with the error, which the
  // Reducer re-raises (as a Python RuntimeError) when it waits for the bucket.
  void call(at::Tensor& buf, int dtype, void* stream) {
    const int rc = st_->allreduce(st_->comm, buf.data_ptr(), (size_t)buf.numel(), dtype, dtype, st_->scale, st_->algo,
                                  stream);
    if (rc != 0 && error_.empty())
      error_ = "b2c_allreduce failed (" + std::to_string(rc) + "): " + (st_->last_error ? st_->last_error() : "?");
  }
  void timed_call(at::Tensor& buf, int dtype, const c10::cuda::CUDAStream& stream, size_t nbytes) {
    if (!st_->timing) {
      call(buf, dtype, stream.stream());
      return;
    }
    st_->timed.emplace_back();
    auto& t = st_->timed.back();
    t.nbytes = (int64_t)nbytes;
    t.start.record(stream);
    call(buf, dtype, stream.stream());
    t.stop.record(stream);
  }
  void finish(const c10::intrusive_ptr<c10::ivalue::Future>& fut, at::Tensor& buf) {
    if (error_.empty()) {
      fut->markCompleted(c10::IValue(buf));
    } else {
      fut->setError(std::make_exception_ptr(std::runtime_error(error_)));
      error_.clear();
    }
  }
  std::string error_;
  std::shared_ptr<HookState> st_;
};

std::vector<std::shared_ptr<HookState>>& registry() {
  static std::vector<std::shared_ptr<HookState>> r;
  return r;
}

// fn / err are the addresses of b2c_allreduce / b2c_last_error (ctypes.cast(..., c_void_p).value),
// comm the b2c_comm*; returns an id for buckets_seen().
int64_t register_hook(const std::shared_ptr<c10d::Reducer>& reducer, uint64_t fn, uint64_t err, uint64_t comm,
                      double scale, int64_t algo, int64_t inline_bytes) {
  TORCH_CHECK(reducer, "b2c hook: reducer is null");
  TORCH_CHECK(fn != 0, "b2c hook: allreduce function pointer is null");
  auto st = std::make_shared<HookState>();
  st->allreduce = reinterpret_cast<allreduce_fn>(fn);
  st->last_error = reinterpret_cast<error_fn>(err);
  st->comm = reinterpret_cast<void*>(comm);
  st->scale = (float)scale;
  st->algo = (int)algo;
  st->inline_bytes = (size_t)inline_bytes;
  registry().push_back(st);
  reducer->register_comm_hook(std::make_unique<B2cCommHook>(st));
  return (int64_t)registry().size() - 1;
}

void enable_timing(int64_t id, bool on) {
  TORCH_CHECK(id >= 0 && (size_t)id < registry().size(), "bad hook id");
  registry()[id]->timing = on;
  if (!on) registry()[id]->timed.clear();
}

// (microseconds, bytes) of every call since timing was enabled; waits for the events, then forgets them
std::vector<std::pair<double, int64_t>> collect_timing(int64_t id) {
  TORCH_CHECK(id >= 0 && (size_t)id < registry().size(), "bad hook id");
  auto& st = *registry()[id];
  std::vector<std::pair<double, int64_t>> out;
  for (auto& t : st.timed) {
    t.stop.synchronize();
    out.emplace_back((double)t.start.elapsed_time(t.stop) * 1e3, t.nbytes);
  }
  st.timed.clear();
  return out;
}

int64_t buckets_seen(int64_t id) {
  TORCH_CHECK(id >= 0 && (size_t)id < registry().size(), "bad hook id");
  return registry()[id]->buckets;
}

}  // namespace

PYBIND11_MODULE(TORCH_EXTENSION_NAME, m) {
  m.def("register_hook", &register_hook, "register the b2c comm hook on a c10d Reducer");
  m.def("buckets_seen", &buckets_seen);
  m.def("enable_timing", &enable_timing);
  m.def("collect_timing", &collect_timing);
}
