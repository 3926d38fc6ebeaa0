// Launch-program executor: the native replacement for what the reference leaves to the Python interpreter,
// autograd's engine threads and DataParallel's parallel_apply threads (utils/train_utils.py:59-70,138;
// torch/nn/parallel/parallel_apply.py).
//
// A training step of the UNet is a FIXED sequence of kernel launches (forward, loss, hand-written backward,
// collectives, optimizer).  Python builds that sequence once as a `dpt_program` (a flat array of op
// records pointing at pre-built plans and static buffers); running a step is then one C call that walks the
// array and launches on up to two streams (compute + communication) with events between them.  No Python,
// no GIL, no allocation in the step; the whole program is CUDA-graph capturable.
//
// `dpt_run_parallel` runs one program per device from persistent C++ worker threads: the single-process
// multi-GPU (DataParallel) engine drives N GPUs from one process without serialising on the GIL.
#include <atomic>
#include <condition_variable>
#include <memory>
#include <mutex>
#include <thread>
#include <vector>

#include "../../include/dpt_b200.h"
#include "../common/host.h"

namespace dpt {
extern std::atomic<long long> g_launch_count;
}

struct dpt_program {
  std::vector<dpt_op> ops;
  std::vector<cudaEvent_t> events;
  int device = 0;
  ~dpt_program() {
    for (cudaEvent_t e : events)
      if (e) cudaEventDestroy(e);
  }
};

using namespace dpt;

extern "C" int dpt_program_create(int device, int n_events, dpt_program** out) {
  if (!out) return fail("program_create: null out");
  std::unique_ptr<dpt_program> p(new dpt_program());
  p->device = device;
  int cur = 0;
  DPT_CUDA_OK(cudaGetDevice(&cur));
  DPT_CUDA_OK(cudaSetDevice(device));
  for (int i = 0; i < n_events; ++i) {
    cudaEvent_t e = nullptr;
    cudaError_t err = cudaEventCreateWithFlags(&e, cudaEventDisableTiming);
    if (err != cudaSuccess) {
      cudaSetDevice(cur);
      return fail("cudaEventCreate failed: %s", cudaGetErrorString(err));
    }
    p->events.push_back(e);
  }
  DPT_CUDA_OK(cudaSetDevice(cur));
  *out = p.release();
  return 0;
}

extern "C" void dpt_program_destroy(dpt_program* p) { delete p; }

extern "C" int dpt_program_add(dpt_program* p, const dpt_op* op) {
  if (!p || !op) return fail("program_add: null argument");
  if (op->kind < 0 || op->kind >= DPT_OP__COUNT) return fail("program_add: bad op kind %d", op->kind);
  if ((op->kind == DPT_OP_EVENT_RECORD || op->kind == DPT_OP_STREAM_WAIT) &&
      (op->i[0] < 0 || op->i[0] >= static_cast<long long>(p->events.size())))
    return fail("program_add: event index %lld out of range", op->i[0]);
  p->ops.push_back(*op);
  return static_cast<int>(p->ops.size()) - 1;
}

// Make `stream` (any stream of the program's device, e.g. the caller's copy stream) wait for the program's event
// `event` as last recorded by a DPT_OP_EVENT_RECORD op: input staging waits for "the step has finished reading its
// input buffer" without a host synchronisation.
extern "C" int dpt_program_stream_wait(const dpt_program* p, int event, void* stream) {
  if (!p || event < 0 || event >= static_cast<int>(p->events.size())) return fail("program_stream_wait: bad program / event");
  DPT_CUDA_OK(cudaStreamWaitEvent(static_cast<cudaStream_t>(stream), p->events[event], 0));
  return 0;
}

extern "C" int dpt_program_size(const dpt_program* p) { return p ? static_cast<int>(p->ops.size()) : 0; }

static int run_op(const dpt_program* prog, const dpt_op& o, cudaStream_t s) {
  void* st = s;
  switch (o.kind) {
    case DPT_OP_CONV:
      return dpt_conv_plan_launch(static_cast<const dpt_conv_plan*>(o.p[0]), st);
    case DPT_OP_WGRAD:
      return dpt_wgrad_plan_launch(static_cast<const dpt_wgrad_plan*>(o.p[0]), st);
    case DPT_OP_PACK_INPUT:
      return dpt_pack_input_ex(o.p[0], (int)o.i[4], o.i[4] ? o.f[0] : 1.f, o.p[1], (int)o.i[0], (int)o.i[1], (int)o.i[2],
                               (int)o.i[3], st);
    case DPT_OP_MAXPOOL:
      return dpt_maxpool2x2(o.p[0], o.p[1], (int)o.i[0], (int)o.i[1], (int)o.i[2], (int)o.i[3], st);
    case DPT_OP_POOL_BWD:
      return dpt_pool_bwd_fused(o.p[0], o.p[1], o.p[2], o.p[3], (int)o.i[0], (int)o.i[1], (int)o.i[2], (int)o.i[3],
                                (int)o.i[4], st);
    case DPT_OP_CHANNEL_SUM:
      return dpt_channel_sum_ex(o.p[0], static_cast<float*>(o.p[1]), o.i[0], (int)o.i[1], o.f[0], (int)o.i[3], o.p[2],
                                static_cast<size_t>(o.i[2]), st);
    case DPT_OP_PREP_WEIGHTS:
      return dpt_prep_weights(static_cast<const dpt_prep_entry*>(o.p[0]), (int)o.i[0], st);
    case DPT_OP_HEAD_FWD:
      return dpt_head_sigmoid_fwd(o.p[0], static_cast<const float*>(o.p[1]), static_cast<const float*>(o.p[2]), o.p[3],
                                  (int)o.i[1], static_cast<float*>(o.p[4]), static_cast<float*>(o.p[5]), o.i[0],
                                  (int)o.i[2], st);
    case DPT_OP_HEAD_FINALIZE: {
      // p[0..7] = partial arrays, i[0] = n_parts, i[1] = blocks per part, i[2] = log slot, f[0] = total pixels
      // (as float: exact below 2^24; larger counts are passed in i[3]), f[1] = dice weight, p[8] = state, p[9] = log
      const float* parts[8];
      int nb[8];
      const int n = (int)o.i[0];
      for (int k = 0; k < n && k < 8; ++k) {
        parts[k] = static_cast<const float*>(o.p[k]);
        nb[k] = (int)o.i[1];
      }
      return dpt_head_finalize(parts, nb, n, static_cast<double>(o.i[3]), o.f[1], static_cast<float*>(o.p[8]),
                               static_cast<float*>(o.p[9]), (int)o.i[2], st);
    }
    case DPT_OP_HEAD_BWD:
      return dpt_head_sigmoid_bwd_ex(o.p[0], static_cast<const float*>(o.p[1]), static_cast<const float*>(o.p[2]), o.p[3],
                                     (int)o.i[1], static_cast<const float*>(o.p[4]), o.p[5], static_cast<float*>(o.p[6]),
                                     static_cast<float*>(o.p[7]), static_cast<float*>(o.p[8]), o.i[0], (int)o.i[2],
                                     o.f[0], o.f[1], (int)o.i[3], (int)o.i[4], st);
    case DPT_OP_OPT_TICK:
      return dpt_opt_tick(static_cast<float*>(o.p[0]), st);
    case DPT_OP_ADAM:
      return dpt_adam_step(static_cast<float*>(o.p[0]), static_cast<float*>(o.p[1]), static_cast<float*>(o.p[2]),
                           static_cast<float*>(o.p[3]), o.i[0], static_cast<const float*>(o.p[4]), o.f[0], o.f[1],
                           o.f[2], (int)o.i[1], st);
    case DPT_OP_RMSPROP:
      return dpt_rmsprop_step(static_cast<float*>(o.p[0]), static_cast<float*>(o.p[1]), static_cast<float*>(o.p[2]),
                              static_cast<float*>(o.p[3]), o.i[0], static_cast<const float*>(o.p[4]), o.f[3], o.f[0],
                              o.f[1], o.f[4], o.f[2], (int)o.i[1], st);
    case DPT_OP_FUSED_AR:
      return dpt_fused_allreduce_opt(static_cast<const dpt_team*>(o.p[0]), static_cast<const dpt_fused_ar_desc*>(o.p[1]),
                                     st);
    case DPT_OP_BCAST:
      return dpt_team_bcast(static_cast<const dpt_team*>(o.p[0]), static_cast<void* const*>(o.p[1]), o.i[0], (int)o.i[1],
                            (int)o.i[2], st);
    case DPT_OP_BCAST_MULTIMEM:
      return dpt_team_bcast_multimem(static_cast<const dpt_team*>(o.p[0]), o.p[1], o.p[2], o.i[0], (int)o.i[1], (int)o.i[2], st);
    case DPT_OP_ALLREDUCE_SUM:
      return dpt_team_allreduce_sum(static_cast<const dpt_team*>(o.p[0]), static_cast<void* const*>(o.p[1]), o.i[0],
                                    o.f[0], (int)o.i[1], st);
    case DPT_OP_MEMSET:
      DPT_CUDA_OK(cudaMemsetAsync(o.p[0], (int)o.i[1], static_cast<size_t>(o.i[0]), s));
      return 0;
    case DPT_OP_MEMCPY:
      DPT_CUDA_OK(cudaMemcpyAsync(o.p[0], o.p[1], static_cast<size_t>(o.i[0]), cudaMemcpyDefault, s));
      return 0;
    case DPT_OP_BN_FWD_TRAIN:
      return dpt_bn_forward_train(o.p[0], o.p[1], static_cast<const float*>(o.p[2]), static_cast<const float*>(o.p[3]),
                                  static_cast<float*>(o.p[4]), static_cast<float*>(o.p[5]), static_cast<float*>(o.p[6]),
                                  static_cast<float*>(o.p[7]), o.i[0], (int)o.i[1], o.f[0], o.f[1], (int)o.i[2], o.p[8],
                                  static_cast<size_t>(o.i[3]), st);
    case DPT_OP_BN_FWD_EVAL:
      return dpt_bn_forward_eval(o.p[0], o.p[1], static_cast<const float*>(o.p[2]), static_cast<const float*>(o.p[3]),
                                 static_cast<const float*>(o.p[4]), static_cast<const float*>(o.p[5]),
                                 static_cast<float*>(o.p[6]), o.i[0], (int)o.i[1], o.f[0], (int)o.i[2], st);
    case DPT_OP_BN_BWD:
      return dpt_bn_backward(o.p[0], o.p[1], o.p[2], static_cast<const float*>(o.p[3]), static_cast<float*>(o.p[4]),
                             static_cast<float*>(o.p[5]), o.i[0], (int)o.i[1], (int)o.i[2], o.p[6],
                             static_cast<size_t>(o.i[3]), st);
    case DPT_OP_BILINEAR_FWD:
      return dpt_bilinear_up2x_fwd(o.p[0], o.p[1], (int)o.i[0], (int)o.i[1], (int)o.i[2], (int)o.i[3], st);
    case DPT_OP_BILINEAR_BWD:
      return dpt_bilinear_up2x_bwd(o.p[0], o.p[1], (int)o.i[0], (int)o.i[1], (int)o.i[2], (int)o.i[3], st);
    case DPT_OP_HEAD_SOFTMAX_FWD:
      return dpt_head_softmax_fwd(o.p[0], static_cast<const float*>(o.p[1]), static_cast<const float*>(o.p[2]), o.p[3],
                                  (int)o.i[2], static_cast<float*>(o.p[4]), static_cast<float*>(o.p[5]),
                                  static_cast<float*>(o.p[6]), o.i[0], o.i[1], (int)o.i[3], (int)o.i[4], st);
    case DPT_OP_HEAD_SOFTMAX_BWD:
      return dpt_head_softmax_bwd(o.p[0], static_cast<const float*>(o.p[1]), static_cast<const float*>(o.p[2]), o.p[3],
                                  (int)o.i[1], static_cast<const float*>(o.p[4]), o.p[5], static_cast<float*>(o.p[6]),
                                  static_cast<float*>(o.p[7]), o.i[0], (int)o.i[2], (int)o.i[3], o.f[0], (int)o.i[4], o.p[8],
                                  static_cast<size_t>(o.i[5]), st);
    case DPT_OP_HEAD_SOFTMAX_FINALIZE: {
      const float* parts[8];
      const int n = (int)o.i[0];
      for (int k = 0; k < n && k < 8; ++k) parts[k] = static_cast<const float*>(o.p[k]);
      return dpt_head_softmax_finalize(parts, n, static_cast<double>(o.i[1]), static_cast<float*>(o.p[8]), st);
    }
    case DPT_OP_BN_FINALIZE:
      return dpt_bn_finalize(static_cast<const float*>(o.p[0]), (int)o.i[0], (int)o.i[1], static_cast<const float*>(o.p[1]),
                             static_cast<const float*>(o.p[2]), static_cast<float*>(o.p[3]), static_cast<float*>(o.p[4]),
                             static_cast<float*>(o.p[5]), o.i[2], (int)o.i[3], o.f[0], o.f[1], st);
    case DPT_OP_BN_APPLY:
      return dpt_bn_apply(o.p[0], o.p[1], o.p[2], static_cast<const float*>(o.p[3]), (int)o.i[0], (int)o.i[1], (int)o.i[2],
                          (int)o.i[3], (int)o.i[4], st);
    case DPT_OP_BN_EVAL_PREPARE:
      return dpt_bn_eval_prepare(static_cast<const float*>(o.p[0]), static_cast<const float*>(o.p[1]),
                                 static_cast<const float*>(o.p[2]), static_cast<const float*>(o.p[3]),
                                 static_cast<float*>(o.p[4]), (int)o.i[0], o.f[0], st);
    case DPT_OP_EVENT_RECORD:
      DPT_CUDA_OK(cudaEventRecord(prog->events[o.i[0]], s));
      return 0;
    case DPT_OP_STREAM_WAIT:
      DPT_CUDA_OK(cudaStreamWaitEvent(s, prog->events[o.i[0]], 0));
      return 0;
    default:
      return fail("program: unknown op kind %d", o.kind);
  }
}

extern "C" int dpt_program_run(const dpt_program* p, int first, int last, void* stream_main, void* stream_aux) {
  if (!p) return fail("program_run: null program");
  const int n = static_cast<int>(p->ops.size());
  if (last < 0 || last > n) last = n;
  if (first < 0) first = 0;
  // The program's buffers, tensor maps and events belong to p->device: launch there whatever device the caller has
  // current (fit(device='cuda:1') while device 0 is current), and put the caller's device back afterwards.
  int cur = p->device;
  DPT_CUDA_OK(cudaGetDevice(&cur));
  if (cur != p->device) DPT_CUDA_OK(cudaSetDevice(p->device));
  int rc = 0;
  for (int k = first; k < last; ++k) {
    const dpt_op& o = p->ops[k];
    cudaStream_t s = static_cast<cudaStream_t>(o.stream == 1 ? stream_aux : stream_main);
    if (run_op(p, o, s) != 0) {
      std::string prev = last_error_ref();
      rc = fail("program op %d (kind %d) failed: %s", k, o.kind, prev.c_str());
      break;
    }
  }
  if (cur != p->device) cudaSetDevice(cur);
  return rc;
}

// ------------------------------------------------------------------------------------------------------
// Persistent per-device worker threads (DataParallel engine).
// ------------------------------------------------------------------------------------------------------
namespace {

struct Worker {
  std::thread th;
  std::mutex mu;
  std::condition_variable cv;
  bool has_job = false, done = true, quit = false;
  const dpt_program* prog = nullptr;
  int first = 0, last = -1;
  void* s_main = nullptr;
  void* s_aux = nullptr;
  int rc = 0;
  std::string err;
  int device = 0;

  void loop() {
    cudaSetDevice(device);
    for (;;) {
      std::unique_lock<std::mutex> lk(mu);
      cv.wait(lk, [&] { return has_job || quit; });
      if (quit) return;
      has_job = false;
      lk.unlock();
      int r = dpt_program_run(prog, first, last, s_main, s_aux);
      std::string e = r ? last_error_ref() : std::string();
      lk.lock();
      rc = r;
      err = e;
      done = true;
      cv.notify_all();
    }
  }
};

struct Pool {
  std::vector<std::unique_ptr<Worker>> workers;  // index = device
  std::mutex mu;
  Worker* get(int device) {
    std::lock_guard<std::mutex> g(mu);
    if (static_cast<int>(workers.size()) <= device) workers.resize(device + 1);
    if (!workers[device]) {
      workers[device].reset(new Worker());
      Worker* w = workers[device].get();
      w->device = device;
      w->th = std::thread([w] { w->loop(); });
    }
    return workers[device].get();
  }
  ~Pool() {
    for (auto& w : workers)
      if (w) {
        {
          std::lock_guard<std::mutex> g(w->mu);
          w->quit = true;
        }
        w->cv.notify_all();
        if (w->th.joinable()) w->th.join();
      }
  }
};

Pool& pool() {
  static Pool p;
  return p;
}

}  // namespace

extern "C" int dpt_run_parallel(int n, dpt_program* const* programs, const int* first, const int* last,
                                void* const* streams_main, void* const* streams_aux) {
  if (n <= 0) return 0;
  std::vector<Worker*> ws(n);
  for (int k = 0; k < n; ++k) {
    Worker* w = pool().get(programs[k]->device);
    ws[k] = w;
    std::lock_guard<std::mutex> g(w->mu);
    w->prog = programs[k];
    w->first = first ? first[k] : 0;
    w->last = last ? last[k] : -1;
    w->s_main = streams_main[k];
    w->s_aux = streams_aux ? streams_aux[k] : streams_main[k];
    w->done = false;
    w->has_job = true;
    w->cv.notify_all();
  }
  int rc = 0;
  for (int k = 0; k < n; ++k) {
    Worker* w = ws[k];
    std::unique_lock<std::mutex> lk(w->mu);
    w->cv.wait(lk, [&] { return w->done; });
    if (w->rc != 0 && rc == 0) {
      rc = w->rc;
      fail("device %d: %s", w->device, w->err.c_str());
    }
  }
  return rc;
}
