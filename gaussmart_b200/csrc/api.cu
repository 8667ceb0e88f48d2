// C ABI entry points (include/gaussmart_b200.h): workspace layouts, forward, backward,
// markVisible.  Host-side orchestration that replaces CudaRasterizer::Rasterizer::{forward,
// backward,markVisible} (rasterizer_impl.cu:141-153,198-458).
#include "common.cuh"
#include <cstdarg>
#include <cstdio>
#include <cstring>
#include <mutex>
#include <vector>

namespace gsm {

static thread_local char g_error[512] = "";

void set_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_error, sizeof(g_error), fmt, ap);
  va_end(ap);
}

int cuda_fail(cudaError_t e, const char* what) {
  set_error("CUDA error %d (%s) at %s", (int)e, cudaGetErrorString(e), what);
  return (int)e;
}

// ---- optional per-stage device timing (gsm_stage_timing_*): cudaEvent pairs recorded on the
// launching stream around every stage, resolved lazily when read.  Off by default.
static const char* const kStageNames[GSM_NUM_STAGES] = {
    "preprocess_fwd", "tile_scan", "scatter", "tile_sort", "sh_grad_expand", "render_fwd",
    "grad_zero", "render_bwd", "preprocess_bwd", "knn", "photometric_loss", "adam", "surface_regularizer", "peer_allreduce"};

struct StageTimer {
  std::mutex mu;
  bool on = false;
  struct Pending { int stage; cudaEvent_t a, b; };
  std::vector<cudaEvent_t> pool;
  std::vector<Pending> pending;
  double sum_ms[GSM_NUM_STAGES] = {};
  long long calls[GSM_NUM_STAGES] = {};
  cudaEvent_t get() {
    if (!pool.empty()) { cudaEvent_t e = pool.back(); pool.pop_back(); return e; }
    cudaEvent_t e = nullptr;
    cudaEventCreate(&e);
    return e;
  }
  void resolve() {
    for (auto& p : pending) {
      float ms = 0.f;
      if (cudaEventSynchronize(p.b) == cudaSuccess && cudaEventElapsedTime(&ms, p.a, p.b) == cudaSuccess) {
        sum_ms[p.stage] += ms;
        calls[p.stage]++;
      }
      pool.push_back(p.a);
      pool.push_back(p.b);
    }
    pending.clear();
  }
};
static StageTimer g_timer;

StageMark::StageMark(int stage, cudaStream_t stream) : stage_(stage), stream_(stream), a_(nullptr) {
  if (!g_timer.on) return;
  std::lock_guard<std::mutex> lk(g_timer.mu);
  a_ = g_timer.get();
  if (a_) cudaEventRecord(a_, stream_);
}
StageMark::~StageMark() {
  if (!a_) return;
  std::lock_guard<std::mutex> lk(g_timer.mu);
  cudaEvent_t b = g_timer.get();
  if (b) {
    cudaEventRecord(b, stream_);
    g_timer.pending.push_back({stage_, a_, b});
    if (g_timer.pending.size() > 8192) g_timer.resolve();
  }
}

GeometryLayout geometry_layout(int P) {
  GeometryLayout L{};
  const size_t n = (size_t)(P > 0 ? P : 0);
  size_t off = 0;
  auto take = [&](size_t bytes) { size_t o = off; off = align_up(off + bytes, 128); return o; };
  L.records = take(n * GSM_SURFEL_RECORD_BYTES);
  L.bbox = take(n * 8);
  L.cull = take(n * 16);
  L.clamp = take(n);
  L.rects = take(n * 16);
  L.flags = take(16);
  L.total = off + 128;
  return L;
}

ImageLayout image_layout(int width, int height) {
  ImageLayout L{};
  const size_t N = (size_t)width * height;
  const size_t Nt = (size_t)((width + TILE_X - 1) / TILE_X) * ((height + TILE_Y - 1) / TILE_Y);
  size_t off = 0;
  auto take = [&](size_t bytes) { size_t o = off; off = align_up(off + bytes, 128); return o; };
  L.final_T = take(N * 3 * 4);
  L.n_contrib = take(N * 2 * 4);
  L.ranges = take(Nt * 8);
  L.tile_count = take(Nt * 4 * TILE_SUBS);
  L.tile_cursor = take(Nt * 4 * TILE_SUBS);
  L.num_rendered = take(16);
  L.total = off + 128;
  return L;
}

BinningLayout binning_layout(size_t capacity, int width, int height) {
  (void)width; (void)height;
  BinningLayout L{};
  size_t off = 0;
  auto take = [&](size_t bytes) { size_t o = off; off = align_up(off + bytes, 128); return o; };
  L.point_list = take(capacity * 4);
  L.strip_mask = take(capacity + 4);
  L.pix_mask = take(capacity * 4 * STRIPS_PER_TILE);
  L.persistent = off;
  L.keys_sorted = take(capacity * 8);
  L.pairs = take(capacity * 8);
  L.total = off + 128;
  return L;
}

}  // namespace gsm

using namespace gsm;

extern "C" const char* gsm_last_error(void) { return g_error; }
extern "C" int gsm_abi_version(void) { return GSM_ABI_VERSION; }
extern "C" int gsm_stage_count(void) { return GSM_NUM_STAGES; }
extern "C" const char* gsm_stage_name(int i) { return (i >= 0 && i < GSM_NUM_STAGES) ? kStageNames[i] : ""; }
extern "C" int gsm_stage_timing_enable(int on) {
  std::lock_guard<std::mutex> lk(g_timer.mu);
  const int was = g_timer.on ? 1 : 0;
  g_timer.on = on != 0;
  return was;
}
extern "C" int gsm_stage_timing_read(double* sum_ms, long long* calls, int reset) {
  std::lock_guard<std::mutex> lk(g_timer.mu);
  g_timer.resolve();
  for (int i = 0; i < GSM_NUM_STAGES; i++) {
    if (sum_ms) sum_ms[i] = g_timer.sum_ms[i];
    if (calls) calls[i] = g_timer.calls[i];
    if (reset) { g_timer.sum_ms[i] = 0.0; g_timer.calls[i] = 0; }
  }
  return GSM_OK;
}
extern "C" size_t gsm_geometry_bytes(int P) { return geometry_layout(P).total; }
extern "C" size_t gsm_image_bytes(int width, int height) { return image_layout(width, height).total; }
extern "C" size_t gsm_binning_bytes(size_t capacity, int width, int height) {
  return binning_layout(capacity, width, height).total;
}
extern "C" size_t gsm_binning_persistent_bytes(size_t capacity) { return binning_layout(capacity, 0, 0).persistent + 128; }
extern "C" size_t gsm_binning_sorted_keys_offset(size_t capacity) { return binning_layout(capacity, 0, 0).keys_sorted; }
extern "C" size_t gsm_image_ranges_offset(int width, int height) { return image_layout(width, height).ranges; }
extern "C" size_t gsm_image_ncontrib_offset(int width, int height) { return image_layout(width, height).n_contrib; }
extern "C" size_t gsm_grad_workspace_bytes(int P) {
  return align_up((size_t)(P > 0 ? P : 0) * GSM_GRAD_RECORD_FLOATS * 4, 128) + 128;
}

// ---- status slots of the asynchronous forward ---------------------------------------------------------------
namespace {
struct StatusSlot {
  ForwardStatus* host;  // pinned
  cudaEvent_t ready;    // recorded right after the status copy, i.e. behind the scan, not behind the whole forward
  bool armed;
};
}  // namespace

extern "C" int gsm_status_create(void** out) {
  if (out == nullptr) { set_error("gsm_status_create: null output"); return GSM_ERR_INVALID_ARGUMENT; }
  StatusSlot* s = new StatusSlot{nullptr, nullptr, false};
  cudaError_t e = cudaHostAlloc((void**)&s->host, sizeof(ForwardStatus), cudaHostAllocDefault);
  if (e == cudaSuccess) e = cudaEventCreateWithFlags(&s->ready, cudaEventDisableTiming);
  if (e != cudaSuccess) {
    if (s->host) cudaFreeHost(s->host);
    delete s;
    return cuda_fail(e, "gsm_status_create");
  }
  memset(s->host, 0, sizeof(ForwardStatus));
  *out = s;
  return GSM_OK;
}
extern "C" int gsm_status_destroy(void* status) {
  StatusSlot* s = (StatusSlot*)status;
  if (s == nullptr) return GSM_OK;
  cudaEventDestroy(s->ready);
  cudaFreeHost(s->host);
  delete s;
  return GSM_OK;
}
extern "C" int gsm_status_query(void* status) {
  StatusSlot* s = (StatusSlot*)status;
  if (s == nullptr || !s->armed) return 1;
  const cudaError_t e = cudaEventQuery(s->ready);
  if (e == cudaSuccess) return 1;
  if (e == cudaErrorNotReady) return 0;
  return -cuda_fail(e, "gsm_status_query");
}
extern "C" int gsm_status_wait(void* status, long long* num_rendered, int* overflow, int* prefiltered_violation) {
  StatusSlot* s = (StatusSlot*)status;
  if (s == nullptr) { set_error("gsm_status_wait: null status"); return GSM_ERR_INVALID_ARGUMENT; }
  if (s->armed) GSM_CUDA(cudaEventSynchronize(s->ready));
  if (num_rendered) *num_rendered = (long long)s->host->num_rendered;
  if (overflow) *overflow = (int)s->host->overflow;
  if (prefiltered_violation) *prefiltered_violation = (int)s->host->prefiltered_violation;
  return GSM_OK;
}

static char* aligned_base(const void* p) {
  return reinterpret_cast<char*>(align_up(reinterpret_cast<size_t>(p), 128));
}

extern "C" int gsm_forward(const gsm_forward_args* a, void* stream_) {
  cudaStream_t stream = (cudaStream_t)stream_;
  if (a == nullptr) { set_error("gsm_forward: null args"); return GSM_ERR_INVALID_ARGUMENT; }
  const int P = a->P, W = a->width, H = a->height;
  const bool debug = (a->debug & 1) != 0;
  if (P < 0 || W <= 0 || H <= 0) { set_error("gsm_forward: bad sizes P=%d W=%d H=%d", P, W, H); return GSM_ERR_INVALID_ARGUMENT; }
  if (a->num_rendered_host == nullptr) { set_error("gsm_forward: num_rendered_host is null"); return GSM_ERR_INVALID_ARGUMENT; }
  if ((a->shs == nullptr) == (a->colors_precomp == nullptr) && P > 0) {
    set_error("gsm_forward: exactly one of shs / colors_precomp must be given"); return GSM_ERR_INVALID_ARGUMENT;
  }
  if (P > 0 && a->transMat_precomp == nullptr && (a->scales == nullptr || a->rotations == nullptr)) {
    set_error("gsm_forward: scales+rotations or transMat_precomp must be given"); return GSM_ERR_INVALID_ARGUMENT;
  }
  if (a->shs != nullptr && (a->D < 0 || a->D > 3 || (a->D + 1) * (a->D + 1) > a->M)) {
    set_error("gsm_forward: SH degree %d needs %d coefficients, M=%d", a->D, (a->D + 1) * (a->D + 1), a->M);
    return GSM_ERR_INVALID_ARGUMENT;
  }
  const int grid_x = (W + TILE_X - 1) / TILE_X, grid_y = (H + TILE_Y - 1) / TILE_Y;
  const int num_tiles = grid_x * grid_y;
  *a->num_rendered_host = 0;

  const GeometryLayout G = geometry_layout(P);
  const ImageLayout I = image_layout(W, H);
  if (a->geometry_bytes < G.total || a->image_bytes < I.total) {
    set_error("gsm_forward: geometry/image workspace too small"); return GSM_ERR_WORKSPACE_TOO_SMALL;
  }
  char* geom = aligned_base(a->geometry);
  char* img = aligned_base(a->image);
  float4* records = (float4*)(geom + G.records);
  uint32_t* flags = (uint32_t*)(geom + G.flags);
  uint2* ranges = (uint2*)(img + I.ranges);
  uint32_t* tile_count = (uint32_t*)(img + I.tile_count);
  uint32_t* tile_cursor = (uint32_t*)(img + I.tile_cursor);
  ForwardStatus* status_dev = (ForwardStatus*)(img + I.num_rendered);
  StatusSlot* slot = (StatusSlot*)a->status;
  const BinningLayout B = binning_layout(a->binning_capacity, W, H);
  if (slot != nullptr && P > 0 && (a->binning == nullptr || a->binning_bytes < B.total)) {
    set_error("gsm_forward: the asynchronous mode needs the binning workspace up front"); return GSM_ERR_WORKSPACE_TOO_SMALL;
  }

  int R = 0;
  if (P > 0) {
    GSM_CUDA(cudaMemsetAsync(flags, 0, 16, stream));
    GSM_CUDA(cudaMemsetAsync(tile_count, 0, sizeof(uint32_t) * (size_t)num_tiles * TILE_SUBS, stream));
    FwdPreprocessParams pp{};
    pp.P = P; pp.D = a->D; pp.M = a->M; pp.W = W; pp.H = H; pp.grid_x = grid_x; pp.grid_y = grid_y;
    pp.scale_modifier = a->scale_modifier; pp.prefiltered = a->prefiltered;
    pp.means3D = a->means3D; pp.scales = a->scales; pp.rotations = a->rotations; pp.opacities = a->opacities;
    pp.shs = a->shs; pp.transMat_precomp = a->transMat_precomp; pp.colors_precomp = a->colors_precomp;
    pp.shs_rest = a->shs ? a->shs_rest : nullptr; pp.raw = a->raw_params ? 1 : 0;
    pp.viewmatrix = a->viewmatrix; pp.projmatrix = a->projmatrix; pp.campos = a->campos;
    pp.radii = a->radii; pp.records = records; pp.bbox = (short4*)(geom + G.bbox); pp.cull = (float4*)(geom + G.cull); pp.clamp = (uint8_t*)(geom + G.clamp); pp.rects = (uint4*)(geom + G.rects); pp.tile_count = tile_count; pp.flags = flags;
    { StageMark m(GSM_STAGE_PREPROCESS_FWD, stream); launch_preprocess_fwd(pp, stream); }
    GSM_STAGE("forward preprocess");
    {
      StageMark m(GSM_STAGE_SCAN, stream);
      launch_tile_scan(tile_count, tile_cursor, ranges, num_tiles, (unsigned long long)a->binning_capacity, flags,
                       status_dev, stream);
    }
    GSM_STAGE("tile-count scan");
    if (slot != nullptr) {
      // asynchronous mode: R travels to the pinned slot behind the scan; nothing below depends on it on the host
      GSM_CUDA(cudaMemcpyAsync(slot->host, status_dev, sizeof(ForwardStatus), cudaMemcpyDeviceToHost, stream));
      GSM_CUDA(cudaEventRecord(slot->ready, stream));
      slot->armed = true;
      R = -1;
    } else {
      // S1: the one host<-device dependency of the reference's forward (rasterizer_impl.cu:282)
      ForwardStatus host_status{};
      GSM_CUDA(cudaMemcpyAsync(&host_status, status_dev, sizeof(ForwardStatus), cudaMemcpyDeviceToHost, stream));
      GSM_CUDA(cudaStreamSynchronize(stream));
      if (host_status.prefiltered_violation != 0) {
        set_error("Point is filtered although prefiltered is set. This shouldn't happen!");
        return GSM_ERR_INVALID_ARGUMENT;
      }
      if (host_status.num_rendered > 0x7fffffffull) { set_error("gsm_forward: too many instances"); return GSM_ERR_INVALID_ARGUMENT; }
      R = (int)host_status.num_rendered;
    }
  } else {
    GSM_CUDA(cudaMemsetAsync(ranges, 0, sizeof(uint2) * (size_t)num_tiles, stream));
    GSM_CUDA(cudaMemsetAsync(status_dev, 0, sizeof(ForwardStatus), stream));
    if (slot != nullptr) {
      GSM_CUDA(cudaMemcpyAsync(slot->host, status_dev, sizeof(ForwardStatus), cudaMemcpyDeviceToHost, stream));
      GSM_CUDA(cudaEventRecord(slot->ready, stream));
      slot->armed = true;
    }
  }
  *a->num_rendered_host = R;
  if (R >= 0 && (size_t)R > a->binning_capacity) {
    set_error("gsm_forward: binning capacity %zu < num_rendered %d", a->binning_capacity, R);
    return GSM_ERR_BINNING_OVERFLOW;
  }
  char* bin = nullptr;
  if (R != 0) {
    if (a->binning == nullptr || a->binning_bytes < B.total) {
      set_error("gsm_forward: binning workspace too small"); return GSM_ERR_WORKSPACE_TOO_SMALL;
    }
    bin = aligned_base(a->binning);
    uint2* pairs = (uint2*)(bin + B.pairs);
    // block masks ride in the top byte of the pairs' values when the surfel index fits 24 bits (debug bit 3 forces the
    // other path, in which the sort gathers box and slabs per instance, so that tests can cover it at small P)
    const int packed = (P <= (1 << 24) && !(a->debug & 8)) ? 1 : 0;
    {
      StageMark m(GSM_STAGE_DUPLICATE, stream);
      launch_scatter(P, (const uint4*)(geom + G.rects), a->radii, tile_cursor, pairs, grid_x, (const short4*)(geom + G.bbox),
                     (const float4*)(geom + G.cull), (a->debug & 2) ? 1 : 0, packed, status_dev, stream);
    }
    GSM_STAGE("scatter");
    {
      StageMark m(GSM_STAGE_SORT, stream);
      launch_tile_sort(ranges, pairs, (uint64_t*)(bin + B.keys_sorted), (uint32_t*)(bin + B.point_list),
                       (uint8_t*)(bin + B.strip_mask), (const short4*)(geom + G.bbox), (const float4*)(geom + G.cull), grid_x, num_tiles,
                       (a->debug & 2) ? 1 : 0, packed, status_dev, stream);
    }
    GSM_STAGE("tile sort");
  }
  if (P > 0) {  // the reference returns zero images without launching when P == 0 (rasterize_points.cu:99)
    FwdRenderParams rp{};
    rp.W = W; rp.H = H; rp.grid_x = grid_x; rp.grid_y = grid_y;
    rp.ranges = ranges;
    rp.point_list = bin ? (const uint32_t*)(bin + B.point_list) : nullptr;
    rp.records = records; rp.background = a->background;
    rp.strip_mask = bin ? (uint8_t*)(bin + B.strip_mask) : nullptr;
    // debug bit 2: inference render, no backward will follow -> the forward skips the hand-over (no per-entry pixel sets)
    rp.pix_mask = (bin && !(a->debug & 4)) ? (uint32_t*)(bin + B.pix_mask) : nullptr; rp.pix_plane = a->binning_capacity;
    rp.bbox = (const short4*)(geom + G.bbox); rp.no_cull = (a->debug & 2) ? 1 : 0;
    rp.final_T = (float*)(img + I.final_T); rp.n_contrib = (uint32_t*)(img + I.n_contrib);
    rp.out_color = a->out_color; rp.out_others = a->out_others;
    rp.stats = (unsigned long long*)a->work_counters;
    rp.status = status_dev;
    { StageMark m(GSM_STAGE_RENDER_FWD, stream); launch_render_fwd(rp, stream); }
    GSM_STAGE("forward render");
  } else {
    GSM_CUDA(cudaMemsetAsync(a->out_color, 0, sizeof(float) * 3 * (size_t)W * H, stream));
    GSM_CUDA(cudaMemsetAsync(a->out_others, 0, sizeof(float) * 7 * (size_t)W * H, stream));
  }
  return GSM_OK;
}

extern "C" int gsm_backward(const gsm_backward_args* a, void* stream_) {
  cudaStream_t stream = (cudaStream_t)stream_;
  if (a == nullptr) { set_error("gsm_backward: null args"); return GSM_ERR_INVALID_ARGUMENT; }
  const int P = a->P, W = a->width, H = a->height;
  const bool debug = (a->debug & 1) != 0;
  if (P < 0 || W <= 0 || H <= 0) { set_error("gsm_backward: bad sizes"); return GSM_ERR_INVALID_ARGUMENT; }
  if (P == 0) return GSM_OK;
  if (a->grad_workspace == nullptr || a->grad_workspace_bytes < gsm_grad_workspace_bytes(P)) {
    set_error("gsm_backward: grad workspace too small"); return GSM_ERR_WORKSPACE_TOO_SMALL;
  }
  const int grid_x = (W + TILE_X - 1) / TILE_X, grid_y = (H + TILE_Y - 1) / TILE_Y;
  const GeometryLayout G = geometry_layout(P);
  const ImageLayout I = image_layout(W, H);
  const char* geom = aligned_base(a->geometry);
  const char* img = aligned_base(a->image);
  const char* bin = a->binning ? aligned_base(a->binning) : nullptr;
  float* grad_records = (float*)aligned_base(a->grad_workspace);
  {
    StageMark m(GSM_STAGE_GRAD_ZERO, stream);
    GSM_CUDA(cudaMemsetAsync(grad_records, 0, (size_t)P * GSM_GRAD_RECORD_FLOATS * 4, stream));
  }

  BwdRenderParams rp{};
  rp.W = W; rp.H = H; rp.grid_x = grid_x; rp.grid_y = grid_y;
  rp.ranges = (const uint2*)(img + I.ranges);
  rp.point_list = (const uint32_t*)bin;  // point_list sits at offset 0 of the binning workspace
  rp.records = (const float4*)(geom + G.records);
  const BinningLayout B = binning_layout(a->binning_capacity, W, H);
  rp.strip_mask = bin ? (const uint8_t*)(bin + B.strip_mask) : nullptr;
  rp.pix_mask = bin ? (const uint32_t*)(bin + B.pix_mask) : nullptr; rp.pix_plane = a->binning_capacity;
  rp.bbox = (const short4*)(geom + G.bbox); rp.no_cull = (a->debug & 2) ? 1 : 0;
  rp.background = a->background;
  rp.final_T = (const float*)(img + I.final_T);
  rp.n_contrib = (const uint32_t*)(img + I.n_contrib);
  rp.dL_dpix = a->dL_dout_color; rp.dL_dothers = a->dL_dout_others;
  rp.grad_records = grad_records;
  { StageMark m(GSM_STAGE_RENDER_BWD, stream); launch_render_bwd(rp, stream); }
  GSM_STAGE("backward render");

  // W,H re-derived in fp32 exactly as backward.cu:613-614 with focal from rasterizer_impl.cu:387-388
  const float focal_y = H / (2.0f * a->tan_fovy);
  const float focal_x = W / (2.0f * a->tan_fovx);
  volatile float wx = focal_x * a->tan_fovx;
  volatile float wy = focal_y * a->tan_fovy;
  BwdPreprocessParams pp{};
  pp.P = P; pp.D = a->D; pp.M = a->M; pp.accumulate = a->accumulate;
  pp.W = (int)(wx * 2); pp.H = (int)(wy * 2);
  pp.means3D = a->means3D; pp.scales = a->scales; pp.rotations = a->rotations; pp.shs = a->shs;
  pp.transMat_precomp = a->transMat_precomp; pp.colors_precomp = a->colors_precomp;
  pp.viewmatrix = a->viewmatrix; pp.projmatrix = a->projmatrix; pp.campos = a->campos;
  pp.radii = a->radii; pp.records = rp.records; pp.clamp = (const uint8_t*)(geom + G.clamp); pp.grad_records = grad_records;
  pp.dL_dmeans2D = a->dL_dmeans2D; pp.dL_dcolors = a->dL_dcolors; pp.dL_dopacity = a->dL_dopacity;
  pp.dL_dmeans3D = a->dL_dmeans3D; pp.dL_dtransMat = a->dL_dtransMat; pp.dL_dsh = a->dL_dsh;
  pp.dL_dscales = a->dL_dscales; pp.dL_drotations = a->dL_drotations; pp.dL_drgb = a->dL_drgb;
  if (a->n_rgb_peers < 0 || a->n_rgb_peers > GSM_MAX_PEERS) { set_error("gsm_backward: n_rgb_peers out of range"); return GSM_ERR_INVALID_ARGUMENT; }
  pp.n_rgb_peers = a->shs ? a->n_rgb_peers : 0;
  for (int k = 0; k < pp.n_rgb_peers; k++) {
    if (a->dL_drgb_peers[k] == nullptr || (reinterpret_cast<size_t>(a->dL_drgb_peers[k]) & 15) != 0) {
      set_error("gsm_backward: dL_drgb_peers[%d] must be a 16-byte aligned device pointer", k); return GSM_ERR_INVALID_ARGUMENT;
    }
    pp.rgb_peers[k] = a->dL_drgb_peers[k];
  }
  pp.stat_norm = a->densify_grad_norm; pp.stat_vis = a->densify_visible; pp.max_radii = a->max_radii;
  pp.shs_rest = a->shs ? a->shs_rest : nullptr; pp.dL_dsh_rest = a->dL_dsh_rest; pp.opacities = a->opacities; pp.raw = a->raw_params ? 1 : 0;
  if (pp.raw && pp.opacities == nullptr) { set_error("gsm_backward: raw_params needs the raw opacities"); return GSM_ERR_INVALID_ARGUMENT; }
  if (pp.shs_rest != nullptr && pp.dL_dsh != nullptr && pp.dL_dsh_rest == nullptr && a->M > 1 && a->dL_drgb == nullptr && a->n_rgb_peers == 0) {
    set_error("gsm_backward: shs_rest given without dL_dsh_rest"); return GSM_ERR_INVALID_ARGUMENT;
  }
  { StageMark m(GSM_STAGE_PREPROCESS_BWD, stream); launch_preprocess_bwd(pp, stream); }
  GSM_STAGE("backward preprocess");
  return GSM_OK;
}

extern "C" int gsm_sh_grad_expand_strided(int P, int D, int M, int n_views, const float* means3D, const float* campos_all,
                                          const float* dL_drgb_all, size_t view_stride, float* dL_dsh, int accumulate,
                                          void* stream_) {
  cudaStream_t stream = (cudaStream_t)stream_;
  if (P < 0 || n_views < 0 || D < 0 || D > 3 || M < 0 || M > 16 || (D + 1) * (D + 1) > M || view_stride < (size_t)3 * (P > 0 ? P : 0)) {
    set_error("gsm_sh_grad_expand: bad sizes P=%d D=%d M=%d n_views=%d", P, D, M, n_views);
    return GSM_ERR_INVALID_ARGUMENT;
  }
  if (P == 0) return GSM_OK;
  StageMark m(GSM_STAGE_RANGES, stream);
  launch_sh_grad_expand(P, D, M, n_views, means3D, campos_all, dL_drgb_all, view_stride, dL_dsh, accumulate, stream);
  GSM_CUDA(cudaGetLastError());
  return GSM_OK;
}
extern "C" int gsm_sh_grad_expand(int P, int D, int M, int n_views, const float* means3D, const float* campos_all,
                                  const float* dL_drgb_all, float* dL_dsh, int accumulate, void* stream_) {
  return gsm_sh_grad_expand_strided(P, D, M, n_views, means3D, campos_all, dL_drgb_all, (size_t)3 * (P > 0 ? P : 0), dL_dsh,
                                    accumulate, stream_);
}

// ---- measured SIMT FP32 peak (bench aid): register-resident FMA chains, 64 warps per SM ------------------------
namespace gsm {
__global__ void __launch_bounds__(512) fp32_peak_kernel(float* __restrict__ out, int iters, float b, float c) {
  float a[8];
#pragma unroll
  for (int k = 0; k < 8; k++) a[k] = (float)(threadIdx.x + k);
  for (int i = 0; i < iters; i++) {
#pragma unroll
    for (int k = 0; k < 8; k++) a[k] = __fmaf_rn(a[k], b, c);   // 8 independent chains per thread
  }
  float s = 0.f;
#pragma unroll
  for (int k = 0; k < 8; k++) s += a[k];
  out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}
}  // namespace gsm

extern "C" int gsm_fp32_peak_probe(int iters, float* out, size_t out_floats, double* flops_out, void* stream_) {
  cudaStream_t stream = (cudaStream_t)stream_;
  int dev = 0, sms = 0;
  GSM_CUDA(cudaGetDevice(&dev));
  GSM_CUDA(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
  const int grid = sms * 4, block = 512;
  if (iters <= 0 || out == nullptr || out_floats < (size_t)grid * block) {
    set_error("gsm_fp32_peak_probe: needs iters > 0 and out[%d]", grid * block); return GSM_ERR_INVALID_ARGUMENT;
  }
  fp32_peak_kernel<<<grid, block, 0, stream>>>(out, iters, 1.0000001f, 1e-9f);
  GSM_CUDA(cudaGetLastError());
  if (flops_out) *flops_out = (double)grid * block * (double)iters * 8.0 * 2.0;
  return GSM_OK;
}

extern "C" int gsm_mark_visible(int P, const float* means3D, const float* viewmatrix,
                                const float* projmatrix, uint8_t* present, void* stream_) {
  (void)projmatrix;  // the reference's NDC test is commented out (auxiliary.h:200)
  cudaStream_t stream = (cudaStream_t)stream_;
  if (P < 0) { set_error("gsm_mark_visible: negative P"); return GSM_ERR_INVALID_ARGUMENT; }
  launch_mark_visible(P, means3D, viewmatrix, present, stream);
  GSM_CUDA(cudaGetLastError());
  return GSM_OK;
}
