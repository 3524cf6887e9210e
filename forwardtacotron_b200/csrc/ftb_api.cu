// C-ABI entry points of libftb200.so that own state: handle lifetime and plan tables.
// (Stateless kernels export their own extern "C" wrappers next to the kernel.)
#include "ftb_internal.h"

using namespace ftb;

namespace {
// Plans allocate, set function attributes and launch on the handle's device: make it current for the call and put
// the caller's device back afterwards (a process that drives several GPUs keeps one handle per device).
struct DeviceScope {
  int prev = -1;
  bool switched = false;
  explicit DeviceScope(int device) {
    if (cudaGetDevice(&prev) == cudaSuccess && prev != device) switched = cudaSetDevice(device) == cudaSuccess;
  }
  ~DeviceScope() { if (switched) cudaSetDevice(prev); }
};
}  // namespace

extern "C" int ftb_version(void) { return FTB_VERSION; }

extern "C" const char* ftb_error_string(int code) {
  switch (code) {
    case 0: return "ok";
    case FTB_E_ARG: return "invalid argument (null pointer / bad size / bad enum)";
    case FTB_E_ALIGN: return "pointer or leading dimension not 16-byte aligned";
    case FTB_E_SHAPE: return "shape not supported by the sm_100a kernels";
    case FTB_E_PLAN: return "unknown plan id";
    case FTB_E_DRIVER: return "cuTensorMapEncodeTiled unavailable (no CUDA driver)";
    case FTB_E_RESIDENT: return "persistent recurrence grid does not fit on the device";
    case FTB_E_LIMIT: return "size exceeds the fixed capacity of the no-allocation entry point (FTB_LR_STATIC_MAX_*)";
    default: break;
  }
  if (code > 0) return cudaGetErrorString(static_cast<cudaError_t>(code));
  return "unknown ftb error";
}

extern "C" int ftb_create(int device, ftb_handle* out) {
  if (!out) return FTB_E_ARG;
  cudaError_t ce = cudaSetDevice(device);
  if (ce != cudaSuccess) return static_cast<int>(ce);
  cudaDeviceProp prop;
  ce = cudaGetDeviceProperties(&prop, device);
  if (ce != cudaSuccess) return static_cast<int>(ce);
  if (prop.major != 10) return FTB_E_SHAPE;          // sm_100a only: no fallback path
  ftb_context* c = new ftb_context();
  c->device = device;
  c->num_sms = prop.multiProcessorCount;
  ce = cudaMalloc(&c->err_word, sizeof(int));
  if (ce == cudaSuccess) ce = cudaMemset(c->err_word, 0, sizeof(int));
  if (ce != cudaSuccess) { cudaFree(c->err_word); delete c; return static_cast<int>(ce); }
  *out = c;
  return 0;
}

extern "C" int ftb_destroy(ftb_handle h) {
  if (!h) return FTB_E_ARG;
  for (auto& kv : h->gemm_plans) gemm_plan_free(kv.second);
  for (auto& kv : h->rnn_plans) rnn_plan_free(kv.second);
  for (auto& kv : h->tail_plans) tail_plan_free(kv.second);
  for (auto& kv : h->graphs) cudaGraphExecDestroy(kv.second);
  cudaFree(h->err_word);
  delete h;
  return 0;
}

extern "C" int64_t ftb_launch_count(ftb_handle h) { return h ? h->launches : -1; }

extern "C" int ftb_graph_capture_begin(ftb_handle h, void* stream) {
  if (!h || !stream) return FTB_E_ARG;             // the legacy default stream cannot be captured
  return static_cast<int>(cudaStreamBeginCapture(static_cast<cudaStream_t>(stream), cudaStreamCaptureModeThreadLocal));
}

extern "C" int ftb_graph_capture_end(ftb_handle h, void* stream, int* graph_out) {
  if (!h || !stream || !graph_out) return FTB_E_ARG;
  cudaGraph_t graph = nullptr;
  cudaError_t ce = cudaStreamEndCapture(static_cast<cudaStream_t>(stream), &graph);
  if (ce != cudaSuccess) { cudaGetLastError(); return static_cast<int>(ce); }
  cudaGraphExec_t exec = nullptr;
  ce = cudaGraphInstantiate(&exec, graph, 0);
  cudaGraphDestroy(graph);
  if (ce != cudaSuccess) { cudaGetLastError(); return static_cast<int>(ce); }
  const int id = h->next_plan++;
  h->graphs[id] = exec;
  *graph_out = id;
  return 0;
}

extern "C" int ftb_graph_launch(ftb_handle h, int graph, void* stream) {
  if (!h) return FTB_E_ARG;
  auto it = h->graphs.find(graph);
  if (it == h->graphs.end()) return FTB_E_PLAN;
  return static_cast<int>(cudaGraphLaunch(it->second, static_cast<cudaStream_t>(stream)));
}

extern "C" int ftb_graph_destroy(ftb_handle h, int graph) {
  if (!h) return FTB_E_ARG;
  auto it = h->graphs.find(graph);
  if (it == h->graphs.end()) return FTB_E_PLAN;
  cudaGraphExecDestroy(it->second);
  h->graphs.erase(it);
  return 0;
}

extern "C" int ftb_gemm_plan_create(ftb_handle h, const ftb_gemm_problem* problems, int nproblems, int* plan_out) {
  if (!h || !plan_out) return FTB_E_ARG;
  DeviceScope scope(h->device);
  GemmPlan plan;
  const int rc = gemm_plan_build(problems, nproblems, &plan);
  if (rc) return rc;
  const int id = h->next_plan++;
  h->gemm_plans[id] = plan;
  *plan_out = id;
  return 0;
}

extern "C" int ftb_gemm_plan_run(ftb_handle h, int plan, int impl, void* stream) {
  if (!h) return FTB_E_ARG;
  DeviceScope scope(h->device);
  auto it = h->gemm_plans.find(plan);
  if (it == h->gemm_plans.end()) return FTB_E_PLAN;
  h->launches += 1;
  return gemm_plan_launch(it->second, impl, h->num_sms, static_cast<cudaStream_t>(stream));
}

extern "C" int ftb_gemm_plan_info(ftb_handle h, int plan, int* cta_pairs, int* tiles) {
  if (!h) return FTB_E_ARG;
  auto it = h->gemm_plans.find(plan);
  if (it == h->gemm_plans.end()) return FTB_E_PLAN;
  if (cta_pairs) *cta_pairs = it->second.pair ? 1 : 0;
  if (tiles) *tiles = it->second.total_tiles;
  return 0;
}

extern "C" int ftb_gemm_plan_destroy(ftb_handle h, int plan) {
  if (!h) return FTB_E_ARG;
  DeviceScope scope(h->device);
  auto it = h->gemm_plans.find(plan);
  if (it == h->gemm_plans.end()) return FTB_E_PLAN;
  gemm_plan_free(it->second);
  h->gemm_plans.erase(it);
  return 0;
}

extern "C" int ftb_gemm_plan_set_geometry(ftb_handle h, int plan, int seq_len, int seq_stride, void* stream) {
  if (!h) return FTB_E_ARG;
  DeviceScope scope(h->device);
  auto it = h->gemm_plans.find(plan);
  if (it == h->gemm_plans.end()) return FTB_E_PLAN;
  return gemm_plan_set_geometry(it->second, seq_len, seq_stride, static_cast<cudaStream_t>(stream));
}

extern "C" int ftb_rnn_plan_create(ftb_handle h, const ftb_rnn_problem* problems, int nproblems, int* plan_out) {
  if (!h || !plan_out) return FTB_E_ARG;
  DeviceScope scope(h->device);
  RnnPlan plan;
  const int rc = rnn_plan_build(problems, nproblems, h->num_sms, &plan);
  if (rc) { rnn_plan_free(plan); return rc; }      // releases whatever the build had created before it failed
  const int id = h->next_plan++;
  h->rnn_plans[id] = plan;
  *plan_out = id;
  return 0;
}

extern "C" int ftb_rnn_plan_run(ftb_handle h, int plan, void* stream) {
  if (!h) return FTB_E_ARG;
  DeviceScope scope(h->device);
  auto it = h->rnn_plans.find(plan);
  if (it == h->rnn_plans.end()) return FTB_E_PLAN;
  h->launches += static_cast<int64_t>(it->second.launches.size() + it->second.hx.size());   // + sentinel pre-fills (L2 hand-off)
  return rnn_plan_launch(it->second, static_cast<cudaStream_t>(stream));
}

extern "C" int ftb_rnn_plan_set_geometry(ftb_handle h, int plan, int seq_len, int seq_stride, void* stream) {
  if (!h) return FTB_E_ARG;
  DeviceScope scope(h->device);
  auto it = h->rnn_plans.find(plan);
  if (it == h->rnn_plans.end()) return FTB_E_PLAN;
  return rnn_plan_set_geometry(it->second, seq_len, seq_stride, static_cast<cudaStream_t>(stream));
}

extern "C" int ftb_rnn_plan_info(ftb_handle h, int plan, int* handoff, int* kernel_launches, int* lanes) {
  if (!h) return FTB_E_ARG;
  auto it = h->rnn_plans.find(plan);
  if (it == h->rnn_plans.end()) return FTB_E_PLAN;
  if (handoff) *handoff = it->second.cluster ? 1 : 0;
  if (kernel_launches) *kernel_launches = static_cast<int>(it->second.launches.size());
  if (lanes) *lanes = it->second.lanes;
  return 0;
}

extern "C" int ftb_rnn_plan_destroy(ftb_handle h, int plan) {
  if (!h) return FTB_E_ARG;
  DeviceScope scope(h->device);
  auto it = h->rnn_plans.find(plan);
  if (it == h->rnn_plans.end()) return FTB_E_PLAN;
  rnn_plan_free(it->second);
  h->rnn_plans.erase(it);
  return 0;
}

extern "C" int ftb_cbhg_tail_plan_create(ftb_handle h, const ftb_cbhg_tail_problem* problem, int* plan_out) {
  if (!h || !plan_out) return FTB_E_ARG;
  DeviceScope scope(h->device);
  TailPlan plan;
  const int rc = tail_plan_build(problem, h->num_sms, &plan);
  if (rc) { tail_plan_free(plan); return rc; }
  const int id = h->next_plan++;
  h->tail_plans[id] = plan;
  *plan_out = id;
  return 0;
}

extern "C" int ftb_cbhg_tail_plan_run(ftb_handle h, int plan, void* stream) {
  if (!h) return FTB_E_ARG;
  DeviceScope scope(h->device);
  auto it = h->tail_plans.find(plan);
  if (it == h->tail_plans.end()) return FTB_E_PLAN;
  h->launches += 1;
  return tail_plan_launch(it->second, h->num_sms, static_cast<cudaStream_t>(stream));
}

extern "C" int ftb_cbhg_tail_plan_set_geometry(ftb_handle h, int plan, int seq_len, int seq_stride, void* stream) {
  if (!h) return FTB_E_ARG;
  DeviceScope scope(h->device);
  auto it = h->tail_plans.find(plan);
  if (it == h->tail_plans.end()) return FTB_E_PLAN;
  return tail_plan_set_geometry(it->second, seq_len, seq_stride, static_cast<cudaStream_t>(stream));
}

extern "C" int ftb_cbhg_tail_plan_destroy(ftb_handle h, int plan) {
  if (!h) return FTB_E_ARG;
  DeviceScope scope(h->device);
  auto it = h->tail_plans.find(plan);
  if (it == h->tail_plans.end()) return FTB_E_PLAN;
  tail_plan_free(it->second);
  h->tail_plans.erase(it);
  return 0;
}
