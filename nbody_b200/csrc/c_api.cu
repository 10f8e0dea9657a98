// c_api.cu -- the extern "C" boundary declared in include/nbody_b200.h: context, error plumbing,
// ParticleDataManager equivalents, and the direct / integrator / host-buffer / multi-GPU entry
// points.  Barnes-Hut, spatial hash, sort and bbox entry points live next to their kernels.
#include <atomic>
#include <cstdarg>
#include <cstdlib>
#include <cstring>
#include <map>
#include <new>
#include <string>

#include "common.cuh"

namespace nb {

static thread_local char g_err[512] = "";

void set_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof g_err, fmt, ap);
  va_end(ap);
}

int cuda_fail(cudaError_t e, const char* what, const char* file, int line) {
  set_error("CUDA Error: %s (%s) at %s:%d", cudaGetErrorString(e), what, file, line);
  cudaGetLastError();  // clear sticky-free errors
  return (int)e;
}

// Every reallocation of a live scratch buffer invalidates the pointers baked into captured graphs: a
// process-wide generation counter, compared by nb_graph_launch against the value at capture time.
static std::atomic<uint64_t> g_scratch_generation{0};
uint64_t scratch_generation() { return g_scratch_generation.load(std::memory_order_relaxed); }
static std::atomic<size_t> g_scratch_bytes{0};  // live DevBuf bytes of this process (nb_scratch_bytes)

int DevBuf::ensure(size_t bytes) {
  if (bytes <= cap) return NB_OK;
  if (p) {
    g_scratch_generation.fetch_add(1, std::memory_order_relaxed);
    cudaError_t e = cudaFree(p);
    g_scratch_bytes.fetch_sub(cap, std::memory_order_relaxed);
    p = nullptr;
    cap = 0;
    if (e != cudaSuccess) return cuda_fail(e, "cudaFree", __FILE__, __LINE__);
  }
  size_t want = bytes + bytes / 8 + 256;
  cudaError_t e = cudaMalloc(&p, want);
  if (e != cudaSuccess) {
    p = nullptr;
    set_error("Insufficient GPU memory - Required: %zu bytes", want);
    cudaGetLastError();
    return NB_ERR_RESOURCE;
  }
  cap = want;
  g_scratch_bytes.fetch_add(cap, std::memory_order_relaxed);
  return NB_OK;
}

void DevBuf::release() {
  if (p) {
    g_scratch_generation.fetch_add(1, std::memory_order_relaxed);
    g_scratch_bytes.fetch_sub(cap, std::memory_order_relaxed);
    cudaFree(p);
  }
  p = nullptr;
  cap = 0;
}

static cudaEvent_t prof_get_event(nb_context* ctx) {
  cudaEvent_t e = nullptr;
  if (!ctx->prof_pool.empty()) {
    e = ctx->prof_pool.back();
    ctx->prof_pool.pop_back();
  } else {
    cudaEventCreate(&e);
  }
  return e;
}
ProfScope::ProfScope(nb_context* c, cudaStream_t s) : ctx(c), st(s) {
  if (!ctx->prof || ctx->prof_ev.size() >= 65536) return;
  e0 = prof_get_event(ctx);
  e1 = prof_get_event(ctx);
  cudaEventRecord(e0, st);
}
ProfScope::~ProfScope() {
  if (!e0) return;
  cudaEventRecord(e1, st);
  ctx->prof_ev.emplace_back(e0, e1);
}
StageScope::StageScope(nb_context* c, cudaStream_t s, const char* n) : ctx(c), st(s), name(n) {
  if (!ctx->prof_stages || ctx->stage_ev.size() >= 262144) return;
  e0 = prof_get_event(ctx);
  e1 = prof_get_event(ctx);
  cudaEventRecord(e0, st);
}
StageScope::~StageScope() {
  if (!e0) return;
  cudaEventRecord(e1, st);
  ctx->stage_ev.push_back({name, e0, e1});
}

// implemented in the kernel files
int direct_run(nb_context* ctx, const nb_particles* d, float dt, float G, float eps2, bool step, cudaStream_t st);
int launch_store_acc(nb_context* ctx, const nb_particles* d, cudaStream_t st);
int launch_drift(nb_context* ctx, const nb_particles* d, float dt, cudaStream_t st);
int launch_kick(nb_context* ctx, const nb_particles* d, float dt, cudaStream_t st);
int kinetic_energy(nb_context* ctx, const nb_particles* d, double* ke, cudaStream_t st);
int potential_energy(nb_context* ctx, const nb_particles* d, float G, float eps, double* pe, cudaStream_t st);
int moments(nb_context* ctx, const nb_particles* d, double* out11, cudaStream_t st);
void bh_destroy(nb_context* ctx);
int bh_poll_error(nb_context* ctx);
int peer_poll_error(nb_context* ctx);
int sort_poll_error(nb_context* ctx);
int sort_init_device(nb_context* ctx);
int bh_init_device(nb_context* ctx);
void sort_destroy(nb_context* ctx);
void sh_destroy(nb_context* ctx);
void render_destroy(nb_context* ctx);

static int check_ctx(nb_context* ctx) {
  if (!ctx) {
    set_error("null context");
    return NB_ERR_INVALID;
  }
  cudaError_t e = cudaSetDevice(ctx->device);
  if (e != cudaSuccess) return cuda_fail(e, "cudaSetDevice", __FILE__, __LINE__);
  return NB_OK;
}

static int check_particles(const nb_particles* d, bool need_vel) {
  if (!d) {
    set_error("null particle data");
    return NB_ERR_INVALID;
  }
  if (d->count == 0) return NB_OK;
  if (!d->pos_x || !d->pos_y || !d->pos_z || !d->mass || !d->acc_x || !d->acc_y || !d->acc_z) {
    set_error("particle data has null position/mass/acceleration arrays");
    return NB_ERR_INVALID;
  }
  if (need_vel && (!d->vel_x || !d->vel_y || !d->vel_z || !d->acc_old_x || !d->acc_old_y || !d->acc_old_z)) {
    set_error("particle data has null velocity/acc_old arrays");
    return NB_ERR_INVALID;
  }
  return NB_OK;
}

}  // namespace nb

using namespace nb;

#define NB_ENTER(ctx)                       \
  do {                                      \
    if (int rc__ = check_ctx(ctx)) return rc__; \
  } while (0)

extern "C" {

int nb_version(void) { return 100; }  // 0.1.0

const char* nb_last_error(void) { return g_err; }

int nb_device_info(int device, int* sm_count, int* cc_major, int* cc_minor, size_t* total_mem, int* clock_khz) {
  int ndev = 0;
  cudaError_t e = cudaGetDeviceCount(&ndev);
  if (e != cudaSuccess || ndev == 0) {
    cudaGetLastError();
    set_error("no CUDA device available (%s); this library has no CPU fallback",
              e != cudaSuccess ? cudaGetErrorString(e) : "device count 0");
    return NB_ERR_NO_DEVICE;
  }
  cudaDeviceProp pr;
  NB_CUDA(cudaGetDeviceProperties(&pr, device));
  int clk = 0;
  cudaDeviceGetAttribute(&clk, cudaDevAttrClockRate, device);
  if (sm_count) *sm_count = pr.multiProcessorCount;
  if (cc_major) *cc_major = pr.major;
  if (cc_minor) *cc_minor = pr.minor;
  if (total_mem) *total_mem = pr.totalGlobalMem;
  if (clock_khz) *clock_khz = clk;
  return NB_OK;
}

int nb_context_create(int device, size_t max_bodies_hint, nb_context** out) {
  if (!out) {
    set_error("null out pointer");
    return NB_ERR_INVALID;
  }
  *out = nullptr;
  int sms = 0, maj = 0, min = 0;
  if (int rc = nb_device_info(device, &sms, &maj, &min, nullptr, nullptr)) return rc;
  if (maj != 10) {
    set_error("device %d is sm_%d%d; this library ships sm_100a code only and has no fallback", device, maj, min);
    return NB_ERR_NO_DEVICE;
  }
  NB_CUDA(cudaSetDevice(device));
  if (const char* e = getenv("NB_L2_FETCH_GRANULARITY")) {  // experiment knob: 32 / 64 / 128 bytes fetched per L2 miss
    const int v = atoi(e);
    if (v == 32 || v == 64 || v == 128) cudaDeviceSetLimit(cudaLimitMaxL2FetchGranularity, (size_t)v);
  }
  nb_context* ctx = new (std::nothrow) nb_context();
  if (!ctx) {
    set_error("out of host memory");
    return NB_ERR_RESOURCE;
  }
  ctx->device = device;
  ctx->sm_count = sms;
  ctx->cc_major = maj;
  ctx->cc_minor = min;
  (void)max_bodies_hint;
  // chunk counters of the direct kernel's two passes ({next chunk, CTAs that left} each): zeroed once here,
  // re-armed by the kernel itself, so the first launch may already sit inside a stream capture
  if (ctx->direct_work.ensure(4 * sizeof(unsigned)) != NB_OK ||
      cudaMemset(ctx->direct_work.p, 0, 4 * sizeof(unsigned)) != cudaSuccess) {
    set_error("cannot allocate the work counters: %s", cudaGetErrorString(cudaGetLastError()));
    ctx->direct_work.release();
    delete ctx;
    return NB_ERR_RESOURCE;
  }
  int rc_init = sort_init_device(ctx);  // per-device kernel attributes, set once here (not lazily from a sort)
  if (!rc_init) rc_init = bh_init_device(ctx);
  if (int rc = rc_init) {
    ctx->direct_work.release();
    delete ctx;
    return rc;
  }
  *out = ctx;
  return NB_OK;
}

int nb_context_destroy(nb_context* ctx) {
  if (!ctx) return NB_OK;
  cudaSetDevice(ctx->device);
  cudaDeviceSynchronize();
  nb_mgpu_peer_close(ctx);
  nb_mgpu_bh_close(ctx);
  render_destroy(ctx);
  bh_destroy(ctx);
  sh_destroy(ctx);
  sort_destroy(ctx);
  ctx->src.release();
  ctx->partial.release();
  ctx->direct_work.release();
  ctx->red.release();
  ctx->sort_keys2.release();
  ctx->sort_vals2.release();
  ctx->sort_hist.release();
  ctx->bbox_dev.release();
  if (ctx->h_red) cudaFreeHost(ctx->h_red);
  if (ctx->h_bbox) cudaFreeHost(ctx->h_bbox);
  if (ctx->stage_cap) nb_particles_free_device(&ctx->stage);
  if (ctx->own_stream) cudaStreamDestroy(ctx->own_stream);
  for (auto& pr : ctx->prof_ev) {
    cudaEventDestroy(pr.first);
    cudaEventDestroy(pr.second);
  }
  for (auto e : ctx->prof_pool) cudaEventDestroy(e);
  delete ctx;
  return NB_OK;
}

int nb_context_sm_count(const nb_context* ctx) { return ctx ? ctx->sm_count : 0; }
uint64_t nb_launch_count(const nb_context* ctx) { return ctx ? ctx->launches : 0; }

int nb_profile_enable(nb_context* ctx, int on) {
  NB_ENTER(ctx);
  ctx->prof = on != 0;
  return NB_OK;
}

int nb_profile_read(nb_context* ctx, double* kernel_ms, uint64_t* kernel_launches) {
  NB_ENTER(ctx);
  double total = 0.0;
  for (auto& pr : ctx->prof_ev) {
    NB_CUDA(cudaEventSynchronize(pr.second));
    float ms = 0.f;
    NB_CUDA(cudaEventElapsedTime(&ms, pr.first, pr.second));
    total += ms;
    ctx->prof_pool.push_back(pr.first);
    ctx->prof_pool.push_back(pr.second);
  }
  if (kernel_ms) *kernel_ms = total;
  if (kernel_launches) *kernel_launches = ctx->prof_ev.size();
  ctx->prof_ev.clear();
  return NB_OK;
}

int nb_profile_stages_enable(nb_context* ctx, int on) {
  NB_ENTER(ctx);
  ctx->prof_stages = on != 0;
  return NB_OK;
}

int nb_profile_stages_read(nb_context* ctx, char* buf, size_t cap) {
  NB_ENTER(ctx);
  if (!buf || cap == 0) {
    set_error("stage profile: null buffer");
    return NB_ERR_INVALID;
  }
  std::vector<std::string> order;
  std::map<std::string, std::pair<double, uint64_t>> sum;
  for (auto& r : ctx->stage_ev) {
    NB_CUDA(cudaEventSynchronize(r.e1));
    float ms = 0.f;
    NB_CUDA(cudaEventElapsedTime(&ms, r.e0, r.e1));
    auto it = sum.find(r.name);
    if (it == sum.end()) {
      order.push_back(r.name);
      sum[r.name] = {ms, 1};
    } else {
      it->second.first += ms;
      it->second.second += 1;
    }
    ctx->prof_pool.push_back(r.e0);
    ctx->prof_pool.push_back(r.e1);
  }
  ctx->stage_ev.clear();
  std::string out;
  char line[160];
  for (auto& nme : order) {
    snprintf(line, sizeof line, "%s %.6f %llu\n", nme.c_str(), sum[nme].first, (unsigned long long)sum[nme].second);
    out += line;
  }
  if (out.size() + 1 > cap) {
    set_error("stage profile: buffer too small (%zu bytes needed)", out.size() + 1);
    return NB_ERR_INVALID;
  }
  memcpy(buf, out.c_str(), out.size() + 1);
  return NB_OK;
}

size_t nb_scratch_bytes(const nb_context* ctx) {
  (void)ctx;
  return g_scratch_bytes.load(std::memory_order_relaxed);
}

uint64_t nb_scratch_generation(const nb_context* ctx) {
  (void)ctx;
  return scratch_generation();
}

int nb_synchronize(nb_context* ctx, nb_stream stream) {
  NB_ENTER(ctx);
  NB_CUDA(cudaStreamSynchronize((cudaStream_t)stream));
  // device-side failures of work that has now finished (eager or replayed from a graph): Barnes-Hut traversal
  // stack overflow, radix-sort look-back timeout, multi-GPU exchange timeout
  if (int rc = bh_poll_error(ctx)) return rc;
  if (int rc = sort_poll_error(ctx)) return rc;
  return peer_poll_error(ctx);
}

// ---- CUDA graphs: the steady-state step as one replayable launch -----------------------------------
// A step at small N is launch-bound (direct: 3 kernels, Barnes-Hut: ~20 kernels and memsets for a few
// hundred microseconds of work).  The host layer runs one step eagerly (all scratch sized), captures
// the next on a side stream and replays the instantiated graph from then on; any parameter baked into
// the launches (dt, G, eps, theta, body count, arrays) invalidates it.  Thin wrappers so that host code
// needs no CUDA headers.  Capturable: direct / Barnes-Hut forces and steps, the integrator kernels.
// Not capturable (they synchronise): spatial-hash build, energies, the renderer, *_export, *_stats.
struct nb_graph {
  cudaGraph_t graph = nullptr;
  cudaGraphExec_t exec = nullptr;
  int device = 0;
  nb_context* ctx = nullptr;
  uint64_t generation = 0;  // scratch generation the captured launches' pointers belong to
};

// raw device buffers for host code without CUDA headers (the renderer's framebuffer)
int nb_device_alloc(nb_context* ctx, size_t bytes, void** out) {
  NB_ENTER(ctx);
  if (!out) {
    set_error("device alloc: null argument");
    return NB_ERR_INVALID;
  }
  *out = nullptr;
  if (bytes == 0) return NB_OK;
  cudaError_t e = cudaMalloc(out, bytes);
  if (e != cudaSuccess) {
    cudaGetLastError();
    set_error("device alloc of %zu bytes failed: %s", bytes, cudaGetErrorString(e));
    return NB_ERR_RESOURCE;
  }
  return NB_OK;
}

int nb_device_free(nb_context* ctx, void* p) {
  NB_ENTER(ctx);
  if (p) NB_CUDA(cudaFree(p));
  return NB_OK;
}

int nb_device_download(nb_context* ctx, void* host_dst, const void* device_src, size_t bytes, nb_stream s) {
  NB_ENTER(ctx);
  if (bytes && (!host_dst || !device_src)) {
    set_error("device download: null argument");
    return NB_ERR_INVALID;
  }
  NB_CUDA(cudaMemcpyAsync(host_dst, device_src, bytes, cudaMemcpyDeviceToHost, (cudaStream_t)s));
  NB_CUDA(cudaStreamSynchronize((cudaStream_t)s));
  return NB_OK;
}

int nb_device_upload(nb_context* ctx, void* device_dst, const void* host_src, size_t bytes, nb_stream s) {
  NB_ENTER(ctx);
  if (bytes && (!host_src || !device_dst)) {
    set_error("device upload: null argument");
    return NB_ERR_INVALID;
  }
  NB_CUDA(cudaMemcpyAsync(device_dst, host_src, bytes, cudaMemcpyHostToDevice, (cudaStream_t)s));
  NB_CUDA(cudaStreamSynchronize((cudaStream_t)s));  // the host buffer may be pageable and reused at once
  return NB_OK;
}

int nb_stream_create(nb_context* ctx, nb_stream* out) {
  NB_ENTER(ctx);
  if (!out) {
    set_error("stream create: null argument");
    return NB_ERR_INVALID;
  }
  cudaStream_t st = nullptr;
  NB_CUDA(cudaStreamCreateWithFlags(&st, cudaStreamNonBlocking));
  *out = (nb_stream)st;
  return NB_OK;
}

int nb_stream_destroy(nb_context* ctx, nb_stream s) {
  NB_ENTER(ctx);
  if (s) NB_CUDA(cudaStreamDestroy((cudaStream_t)s));
  return NB_OK;
}

int nb_graph_begin_capture(nb_context* ctx, nb_stream s) {
  NB_ENTER(ctx);
  if (!s) {
    set_error("graph capture needs a non-default stream (nb_stream_create)");
    return NB_ERR_INVALID;
  }
  if (ctx->prof || ctx->prof_stages) {
    set_error("graph capture with nb_profile_enable / nb_profile_stages_enable on is not supported");
    return NB_ERR_INVALID;
  }
  ctx->capture_generation = scratch_generation();
  NB_CUDA(cudaStreamBeginCapture((cudaStream_t)s, cudaStreamCaptureModeThreadLocal));
  return NB_OK;
}

int nb_graph_end_capture(nb_context* ctx, nb_stream s, nb_graph** out) {
  NB_ENTER(ctx);
  if (!s || !out) {
    set_error("graph end capture: null argument");
    return NB_ERR_INVALID;
  }
  cudaGraph_t g = nullptr;
  NB_CUDA(cudaStreamEndCapture((cudaStream_t)s, &g));
  if (scratch_generation() != ctx->capture_generation) {
    // a scratch buffer grew while the step was being captured: earlier nodes hold the freed pointer
    cudaGraphDestroy(g);
    set_error("graph capture: scratch was reallocated during capture (run one eager step first)");
    return NB_ERR_INVALID;
  }
  cudaGraphExec_t e = nullptr;
  cudaError_t err = cudaGraphInstantiate(&e, g, 0);
  if (err != cudaSuccess) {
    cudaGraphDestroy(g);
    return cuda_fail(err, "cudaGraphInstantiate", __FILE__, __LINE__);
  }
  nb_graph* h = new (std::nothrow) nb_graph();
  if (!h) {
    cudaGraphExecDestroy(e);
    cudaGraphDestroy(g);
    set_error("out of host memory");
    return NB_ERR_RESOURCE;
  }
  h->graph = g;
  h->exec = e;
  h->device = ctx->device;
  h->ctx = ctx;
  h->generation = ctx->capture_generation;
  *out = h;
  return NB_OK;
}

int nb_graph_launch(nb_graph* g, nb_stream s) {
  if (!g || !g->exec) {
    set_error("graph launch: null graph");
    return NB_ERR_INVALID;
  }
  if (g->generation != scratch_generation()) {
    // context scratch (sort buffers, partials, tree arrays ...) was reallocated by another user of the context
    // after the capture: the graph's kernels would read and write freed memory
    set_error("graph launch: stale graph (context scratch was reallocated since capture); re-capture");
    return NB_ERR_STALE_GRAPH;
  }
  NB_CUDA(cudaSetDevice(g->device));
  // a replay runs no host-side checks of its own: report what earlier replays flagged (host-mapped flags)
  if (int rc = bh_poll_error(g->ctx)) return rc;
  if (int rc = sort_poll_error(g->ctx)) return rc;
  if (int rc = peer_poll_error(g->ctx)) return rc;
  NB_CUDA(cudaGraphLaunch(g->exec, (cudaStream_t)s));
  return NB_OK;
}

int nb_graph_is_stale(const nb_graph* g) { return (!g || g->generation != scratch_generation()) ? 1 : 0; }

int nb_graph_node_count(const nb_graph* g, size_t* nodes) {
  if (!g || !g->graph || !nodes) {
    set_error("graph node count: null argument");
    return NB_ERR_INVALID;
  }
  NB_CUDA(cudaGraphGetNodes(g->graph, nullptr, nodes));
  return NB_OK;
}

int nb_graph_destroy(nb_graph* g) {
  if (!g) return NB_OK;
  cudaSetDevice(g->device);
  if (g->exec) cudaGraphExecDestroy(g->exec);
  if (g->graph) cudaGraphDestroy(g->graph);
  delete g;
  return NB_OK;
}

// ---- ParticleDataManager ---------------------------------------------------------------------
static float** field_ptrs(nb_particles* d, float** out[13]) {
  float** f[13] = {&d->pos_x, &d->pos_y, &d->pos_z, &d->vel_x, &d->vel_y, &d->vel_z, &d->acc_x,
                   &d->acc_y, &d->acc_z, &d->acc_old_x, &d->acc_old_y, &d->acc_old_z, &d->mass};
  for (int k = 0; k < 13; k++) out[k] = f[k];
  return nullptr;
}

int nb_particles_alloc_device(nb_particles* d, size_t count) {
  if (!d) {
    set_error("null particle data");
    return NB_ERR_INVALID;
  }
  float** f[13];
  field_ptrs(d, f);
  for (int k = 0; k < 13; k++) *f[k] = nullptr;
  d->count = count;
  const size_t bytes = (count ? count : 1) * sizeof(float);
  for (int k = 0; k < 13; k++) {
    cudaError_t e = cudaMalloc((void**)f[k], bytes);
    if (e != cudaSuccess) {
      cudaGetLastError();
      for (int j = 0; j < k; j++) cudaFree(*f[j]);
      for (int j = 0; j < 13; j++) *f[j] = nullptr;
      d->count = 0;
      set_error("Insufficient GPU memory - Required: %zu bytes", bytes * 13);
      return NB_ERR_RESOURCE;
    }
  }
  for (int k = 6; k < 12; k++) {  // acc, acc_old (particle_init.cu:161-167)
    cudaError_t e = cudaMemset(*f[k], 0, bytes);
    if (e != cudaSuccess) {
      for (int j = 0; j < 13; j++) {
        cudaFree(*f[j]);
        *f[j] = nullptr;
      }
      d->count = 0;
      return cuda_fail(e, "cudaMemset", __FILE__, __LINE__);
    }
  }
  return NB_OK;
}

int nb_particles_free_device(nb_particles* d) {
  if (!d) return NB_OK;
  float** f[13];
  field_ptrs(d, f);
  for (int k = 0; k < 13; k++) {
    if (*f[k]) cudaFree(*f[k]);
    *f[k] = nullptr;
  }
  d->count = 0;
  return NB_OK;
}

static int copy_fields(nb_particles* dst, const nb_particles* src, const int* which, int nwhich, size_t n,
                       cudaMemcpyKind kind, cudaStream_t st) {
  float** fd[13];
  float** fs[13];
  field_ptrs(dst, fd);
  field_ptrs(const_cast<nb_particles*>(src), fs);
  for (int k = 0; k < nwhich; k++) {
    const int w = which[k];
    if (!*fd[w] || !*fs[w]) continue;
    NB_CUDA(cudaMemcpyAsync(*fd[w], *fs[w], n * sizeof(float), kind, st));
  }
  return NB_OK;
}

int nb_particles_upload(nb_context* ctx, nb_particles* d, const nb_particles* h, nb_stream s) {
  NB_ENTER(ctx);
  if (!d || !h || h->count > d->count) {
    set_error("upload: bad particle data (host count exceeds device count)");
    return NB_ERR_INVALID;
  }
  static const int which[10] = {0, 1, 2, 3, 4, 5, 6, 7, 8, 12};  // particle_init.cu:257-269
  return copy_fields(d, h, which, 10, h->count, cudaMemcpyHostToDevice, (cudaStream_t)s);
}

int nb_particles_download(nb_context* ctx, nb_particles* h, const nb_particles* d, nb_stream s) {
  NB_ENTER(ctx);
  if (!d || !h || d->count > h->count) {
    set_error("download: bad particle data (device count exceeds host count)");
    return NB_ERR_INVALID;
  }
  static const int which[10] = {0, 1, 2, 3, 4, 5, 6, 7, 8, 12};  // particle_init.cu:271-283
  if (int rc = copy_fields(h, d, which, 10, d->count, cudaMemcpyDeviceToHost, (cudaStream_t)s)) return rc;
  NB_CUDA(cudaStreamSynchronize((cudaStream_t)s));
  return NB_OK;
}

// ---- direct / integrator ---------------------------------------------------------------------
int nb_direct_forces(nb_context* ctx, const nb_particles* d, float G, float eps2, nb_stream s) {
  NB_ENTER(ctx);
  if (int rc = check_particles(d, false)) return rc;
  return direct_run(ctx, d, 0.f, G, eps2, false, (cudaStream_t)s);
}

int nb_store_accelerations(nb_context* ctx, const nb_particles* d, nb_stream s) {
  NB_ENTER(ctx);
  if (int rc = check_particles(d, true)) return rc;
  return launch_store_acc(ctx, d, (cudaStream_t)s);
}
int nb_update_positions(nb_context* ctx, const nb_particles* d, float dt, nb_stream s) {
  NB_ENTER(ctx);
  if (int rc = check_particles(d, true)) return rc;
  return launch_drift(ctx, d, dt, (cudaStream_t)s);
}
int nb_update_velocities(nb_context* ctx, const nb_particles* d, float dt, nb_stream s) {
  NB_ENTER(ctx);
  if (int rc = check_particles(d, true)) return rc;
  return launch_kick(ctx, d, dt, (cudaStream_t)s);
}

int nb_step_direct(nb_context* ctx, const nb_particles* d, float dt, float G, float eps2, nb_stream s) {
  NB_ENTER(ctx);
  if (int rc = check_particles(d, true)) return rc;
  return direct_run(ctx, d, dt, G, eps2, true, (cudaStream_t)s);
}

int nb_kinetic_energy(nb_context* ctx, const nb_particles* d, double* ke, nb_stream s) {
  NB_ENTER(ctx);
  if (!ke || !d || (d->count && (!d->vel_x || !d->vel_y || !d->vel_z || !d->mass))) {
    set_error("kinetic energy: null argument");
    return NB_ERR_INVALID;
  }
  return kinetic_energy(ctx, d, ke, (cudaStream_t)s);
}

int nb_potential_energy(nb_context* ctx, const nb_particles* d, float G, float eps, double* pe, nb_stream s) {
  NB_ENTER(ctx);
  if (!pe || !d || (d->count && (!d->pos_x || !d->pos_y || !d->pos_z || !d->mass))) {
    set_error("potential energy: null argument");
    return NB_ERR_INVALID;
  }
  return potential_energy(ctx, d, G, eps, pe, (cudaStream_t)s);
}

int nb_diagnostics(nb_context* ctx, const nb_particles* d, float G, float eps, int with_potential,
                   nb_diagnostics_t* out, nb_stream s) {
  NB_ENTER(ctx);
  if (!out) {
    set_error("diagnostics: null output");
    return NB_ERR_INVALID;
  }
  if (!d || (d->count && (!d->pos_x || !d->pos_y || !d->pos_z || !d->vel_x || !d->vel_y || !d->vel_z || !d->mass))) {
    set_error("diagnostics: null position/velocity/mass array");
    return NB_ERR_INVALID;
  }
  double m[11];
  if (int rc = moments(ctx, d, m, (cudaStream_t)s)) return rc;
  out->total_mass = m[0];
  for (int a = 0; a < 3; a++) {
    out->center_of_mass[a] = (m[0] != 0.0) ? m[1 + a] / m[0] : 0.0;
    out->momentum[a] = m[4 + a];
    out->angular_momentum[a] = m[7 + a];
  }
  out->kinetic = m[10];
  out->potential = 0.0;
  out->virial_ratio = 0.0;
  out->has_potential = 0;
  if (with_potential) {
    if (int rc = potential_energy(ctx, d, G, eps, &out->potential, (cudaStream_t)s)) return rc;
    out->has_potential = 1;
    if (out->potential != 0.0) out->virial_ratio = 2.0 * out->kinetic / (out->potential < 0 ? -out->potential : out->potential);
  }
  out->total_energy = out->kinetic + out->potential;
  return NB_OK;
}

// ---- host-buffer entry points ------------------------------------------------------------------
static int ensure_stage(nb_context* ctx, size_t n) {
  if (!ctx->own_stream) NB_CUDA(cudaStreamCreateWithFlags(&ctx->own_stream, cudaStreamNonBlocking));
  if (n <= ctx->stage_cap) {
    ctx->stage.count = n;
    return NB_OK;
  }
  if (ctx->stage_cap) nb_particles_free_device(&ctx->stage);
  ctx->stage_cap = 0;
  if (int rc = nb_particles_alloc_device(&ctx->stage, n)) return rc;
  ctx->stage_cap = n;
  return NB_OK;
}

int nb_host_direct_forces(nb_context* ctx, const nb_particles* h, float G, float eps2) {
  NB_ENTER(ctx);
  if (int rc = check_particles(h, false)) return rc;
  const size_t n = h->count;
  if (int rc = ensure_stage(ctx, n)) return rc;
  cudaStream_t st = ctx->own_stream;
  static const int in[4] = {0, 1, 2, 12};
  static const int out[3] = {6, 7, 8};
  if (int rc = copy_fields(&ctx->stage, h, in, 4, n, cudaMemcpyHostToDevice, st)) return rc;
  if (int rc = direct_run(ctx, &ctx->stage, 0.f, G, eps2, false, st)) return rc;
  if (int rc = copy_fields(const_cast<nb_particles*>(h), &ctx->stage, out, 3, n, cudaMemcpyDeviceToHost, st)) return rc;
  NB_CUDA(cudaStreamSynchronize(st));
  return NB_OK;
}

int nb_host_step_direct(nb_context* ctx, nb_particles* h, float dt, float G, float eps2, int steps,
                        int recompute_initial_forces) {
  NB_ENTER(ctx);
  if (int rc = check_particles(h, true)) return rc;
  const size_t n = h->count;
  if (int rc = ensure_stage(ctx, n)) return rc;
  cudaStream_t st = ctx->own_stream;
  static const int in[10] = {0, 1, 2, 3, 4, 5, 6, 7, 8, 12};
  static const int out[12] = {0, 1, 2, 3, 4, 5, 6, 7, 8, 9, 10, 11};
  if (int rc = copy_fields(&ctx->stage, h, in, 10, n, cudaMemcpyHostToDevice, st)) return rc;
  if (recompute_initial_forces)  // ParticleSystem::setState recomputes a(0): particle_system.cpp:283
    if (int rc = direct_run(ctx, &ctx->stage, 0.f, G, eps2, false, st)) return rc;
  for (int k = 0; k < steps; k++)
    if (int rc = direct_run(ctx, &ctx->stage, dt, G, eps2, true, st)) return rc;
  if (int rc = copy_fields(h, &ctx->stage, out, 12, n, cudaMemcpyDeviceToHost, st)) return rc;
  NB_CUDA(cudaStreamSynchronize(st));
  return NB_OK;
}

}  // extern "C"
