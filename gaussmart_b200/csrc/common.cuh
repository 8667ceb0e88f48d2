// Internal definitions shared by the sm_100a kernels of gaussmart_b200.
// Not part of the C ABI (see include/gaussmart_b200.h).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stddef.h>
#include <mutex>
#include "../../include/gaussmart_b200.h"

namespace gsm {

// Tile geometry is fixed by the reference (cuda_rasterizer/config.h:15-17).
constexpr int TILE_X = 16;
constexpr int TILE_Y = 16;
constexpr int TILE_PIX = TILE_X * TILE_Y;
// A tile is composited by 8 independent warps, each owning one BLOCK_W x BLOCK_H = 32-pixel block
// ("strip" in the names below for historical reasons).  8x4 blocks meet ~20 % fewer (block, surfel)
// pairs than 16x2 strips for the few-pixel-wide splats that dominate (measured).
#ifndef GSM_BLOCK_W
#define GSM_BLOCK_W 8
#endif
constexpr int BLOCK_W = GSM_BLOCK_W;
constexpr int BLOCK_H = 32 / BLOCK_W;
constexpr int BLOCKS_X = TILE_X / BLOCK_W;          // blocks across a tile
constexpr int BLOCKS_Y = TILE_Y / BLOCK_H;          // blocks down a tile
constexpr int STRIPS_PER_TILE = BLOCKS_X * BLOCKS_Y;  // = 8: one mask bit per block, bit = by * BLOCKS_X + bx
static_assert(BLOCK_W * BLOCK_H == 32 && STRIPS_PER_TILE == 8, "one warp per block, 8 blocks per tile");
// pixel of lane `lane` of block `b` inside its tile
__host__ __device__ inline int block_px(int b, int lane) { return (b % BLOCKS_X) * BLOCK_W + (lane % BLOCK_W); }
__host__ __device__ inline int block_py(int b, int lane) { return (b / BLOCKS_X) * BLOCK_H + (lane / BLOCK_W); }
// mask of the blocks of a tile met by the inclusive pixel rectangle [c0,c1] x [r0,r1] (tile-local, clamped, non-empty)
__host__ __device__ inline uint32_t block_mask(int c0, int c1, int r0, int r1) {
  const uint32_t cm = ((2u << (c1 / BLOCK_W)) - 1u) & ~((1u << (c0 / BLOCK_W)) - 1u);
  uint32_t m = 0;
  for (int by = r0 / BLOCK_H; by <= r1 / BLOCK_H; by++) m |= cm << (by * BLOCKS_X);
  return m;
}
// Sub-cursors per tile for counting / appending instances (binning.cu), power of two.
constexpr int TILE_SUBS = 8;

// Compile-time constants of the reference (auxiliary.h:37-40).
constexpr float NEAR_N = 0.2f;
constexpr float FAR_N = 100.0f;
constexpr float FILTER_INV_SQUARE = 2.0f;
// (float)(3.0f * 0.707106f) as folded by the reference build: 0x4007C3AC
#define GSM_MIN_RADIUS_BITS 0x4007C3ACu
// far/(far-near) folded in fp32 by the reference build: 0x3F8041AB
#define GSM_FAR_RATIO_BITS 0x3F8041ABu

constexpr float ALPHA_MAX = 0.99f;
constexpr float ALPHA_MIN = 1.0f / 255.0f;
constexpr float T_EPS = 0.0001f;

// Packed per-surfel record (GSM_SURFEL_RECORD_BYTES = 80): five float4.
//   q0 = (Tu.x, Tu.y, Tu.z, centre.x)     Tu,Tv,Tw: rows of the splat->pixel homography
//   q1 = (Tv.x, Tv.y, Tv.z, centre.y)     centre: AABB centre used by the low-pass filter
//   q2 = (Tw.x, Tw.y, Tw.z, opacity)
//   q3 = (n.x,  n.y,  n.z,  view depth)   n: view-space normal, flipped to face the camera
//   q4 = (r,    g,    b,    clamp bits)   clamp bits: int bit c set when channel c was clamped at 0
constexpr int REC_F4 = 5;

// Conservative contribution box (short4 per surfel: xmin, ymin, xmax, ymax in integer pixel
// coordinates, inclusive).  A pixel outside it cannot pass the reference's alpha >= 1/255 test
// (forward.cu:385-392): alpha = opacity*exp(-rho/2) >= 1/255 needs rho = min(rho3d, rho2d) <=
// 2 ln(255 opacity) (<= 11.083 at opacity 1; the per-surfel bound is used, with a 3% + 0.2 margin
// on rho), so the box encloses the perspective-correct image of the disc u^2+v^2 <=
// CULL_RHO (the reference's own AABB construction, forward.cu:119-145, at a larger cutoff) united
// with the low-pass disc 2|d|^2 <= CULL_RHO, plus slack; surfels whose disc is not wholly in front
// of the camera plane get the unbounded box.  Compositing kernels use it only to SKIP work.
constexpr float CULL_RHO = 11.5f;
constexpr float CULL_SLACK_PX = 0.75f;
// Diagonal slabs ("cull record", one float4 per surfel): (lo, hi) of x+y and (lo, hi) of x-y over the same
// conservative contribution region as the box (the conic rho3d <= c^2 is an ellipse whose extent along a
// direction n is sqrt(n^T S n); united with the low-pass disc; same stretch and slack as the box).  Box and
// slabs bound the region by an octagon.  The tile sort drops from an instance's block mask the blocks that
// miss a slab; like the box this only ever SKIPS work.  Unbounded slabs (+-inf) where the box is unbounded.

// Gradient accumulation record (GSM_GRAD_RECORD_FLOATS = 20 floats, 80 B):
//   [0..2] dTu  [3..5] dTv  [6..8] dTw  [9..10] dcentre  [11..13] drgb  [14..16] dnormal
//   [17] dopacity  [18..19] pad
constexpr int GREC = GSM_GRAD_RECORD_FLOATS;
constexpr int GREC_USED = 18;

__host__ __device__ inline size_t align_up(size_t x, size_t a) { return (x + a - 1) / a * a; }

// ---- workspace layouts ------------------------------------------------------------------
struct GeometryLayout {
  size_t records;        // float4[P*5]
  size_t bbox;           // short4[P]  conservative integer pixel box of possible contribution
  size_t cull;           // float4[P]  diagonal slabs of the contribution region ("cull record")
  size_t clamp;          // u8[P]      bit c set when colour channel c was clamped at 0 (read by the backward preprocess)
  size_t rects;          // uint4[P]   (minx | miny<<16, maxx | maxy<<16, depth bits, 0): tile rect for the scatter
  size_t flags;          // u32[4]: [0] prefiltered-violation flag
  size_t total;
};
GeometryLayout geometry_layout(int P);

// Device-side status of a forward (image workspace; copied to the pinned status slot in the asynchronous mode).
// overflow != 0: R exceeded the binning capacity -> every kernel behind the scan treats the frame as EMPTY.
struct ForwardStatus {
  unsigned long long num_rendered;
  uint32_t overflow;
  uint32_t prefiltered_violation;
};

struct ImageLayout {
  size_t final_T;    // f32[3*N]  (T, M1, M2)
  size_t n_contrib;  // u32[2*N]  (last contributor, median contributor)
  size_t ranges;     // uint2[Nt]
  size_t tile_count;   // u32[Nt][TILE_SUBS]  instances per tile and sub-cursor (counted by the forward preprocess)
  size_t tile_cursor;  // u32[Nt][TILE_SUBS]  append cursors of the scatter
  size_t num_rendered; // ForwardStatus (R first)
  size_t total;
};
ImageLayout image_layout(int width, int height);

struct BinningLayout {
  size_t point_list;       // u32[cap]   depth-sorted surfel ids per tile   (kept for backward)
  size_t strip_mask;       // u8[cap]    per instance: which 32-pixel blocks of its tile it can touch (kept)
  size_t pix_mask;         // u32[8][cap] per block plane and instance: the pixels of the block that blended the entry
                           //            (written by the forward compositing, kept)
  size_t persistent;       // bytes from the start that must survive until backward
  size_t keys_sorted;      // u64[cap]   (tile << 32 | depth bits), the reference's sorted keys
  size_t pairs;            // uint2[cap] (depth bits, surfel) per tile segment, unsorted
  size_t total;
};
BinningLayout binning_layout(size_t capacity, int width, int height);

// ---- error plumbing -----------------------------------------------------------------------
void set_error(const char* fmt, ...);
int cuda_fail(cudaError_t e, const char* what);

#define GSM_CUDA(expr)                                   \
  do {                                                   \
    cudaError_t _e = (expr);                             \
    if (_e != cudaSuccess) return gsm::cuda_fail(_e, #expr); \
  } while (0)

#define GSM_STAGE(stage_name)                                                   \
  do {                                                                          \
    cudaError_t _e = cudaGetLastError();                                        \
    if (_e == cudaSuccess && debug) _e = cudaStreamSynchronize(stream);         \
    if (_e != cudaSuccess) return gsm::cuda_fail(_e, stage_name);               \
  } while (0)

// ---- per-device one-time configuration -----------------------------------------------------
// cudaFuncSetAttribute and __constant__ uploads are PER DEVICE; a per-process flag would leave the second device of a
// process unconfigured.  ensure(want, f) runs f() on the current device unless that device already reached `want`.
class PerDeviceLevel {
 public:
  template <class F>
  void ensure(size_t want, F&& f) {
    int d = 0;
    cudaGetDevice(&d);
    std::lock_guard<std::mutex> lk(mu_);
    if (d < 0 || d >= kMax) { f(); return; }
    if (level_[d] < want) { f(); level_[d] = want; }
  }
 private:
  static constexpr int kMax = 64;
  std::mutex mu_;
  size_t level_[kMax] = {};
};

// ---- optional per-stage timing (api.cu) --------------------------------------------------
enum {
  GSM_STAGE_PREPROCESS_FWD = 0, GSM_STAGE_SCAN, GSM_STAGE_DUPLICATE, GSM_STAGE_SORT, GSM_STAGE_RANGES,
  GSM_STAGE_RENDER_FWD, GSM_STAGE_GRAD_ZERO, GSM_STAGE_RENDER_BWD, GSM_STAGE_PREPROCESS_BWD, GSM_STAGE_KNN,
  GSM_STAGE_LOSS, GSM_STAGE_ADAM, GSM_STAGE_REGULARIZER, GSM_STAGE_PEER_ALLREDUCE
};
class StageMark {  // RAII: records a cudaEvent pair around a stage when timing is enabled
 public:
  StageMark(int stage, cudaStream_t stream);
  ~StageMark();
 private:
  int stage_;
  cudaStream_t stream_;
  cudaEvent_t a_;
};

// ---- raw-parameter activations (gaussian_model.py:104-113), evaluated as torch's CUDA kernels evaluate them ----------
struct ActivatedRotation { float4 q; float norm; };  // norm = max(||raw||, eps)
__device__ __forceinline__ float2 activate_scaling(float2 s) { return make_float2(expf(s.x), expf(s.y)); }  // torch.exp
__device__ __forceinline__ float activate_opacity(float x) {  // torch.sigmoid: 1 / (1 + exp(-x))
  return __fdiv_rn(1.0f, __fadd_rn(1.0f, expf(-x)));
}
__device__ __forceinline__ ActivatedRotation activate_rotation(float4 r) {
  // F.normalize(dim=-1): x / max(||x||_2, 1e-12); torch's row reduction over 4 elements adds the rounded squares as a
  // shuffle tree, (x0^2 + x2^2) + (x1^2 + x3^2) -- found by tools/activation_probe.py, bit-identical on 1 M rows
  const float n2 = __fadd_rn(__fadd_rn(__fmul_rn(r.x, r.x), __fmul_rn(r.z, r.z)), __fadd_rn(__fmul_rn(r.y, r.y), __fmul_rn(r.w, r.w)));
  const float n = fmaxf(__fsqrt_rn(n2), 1e-12f);
  ActivatedRotation a;
  a.q = make_float4(__fdiv_rn(r.x, n), __fdiv_rn(r.y, n), __fdiv_rn(r.z, n), __fdiv_rn(r.w, n));
  a.norm = n;
  return a;
}

// ---- TMA bulk copies of per-surfel rows (cp.async.bulk, 1-D) -----------------------------------------------------
// SH rows (12 M bytes per surfel, a multiple of 16 when M % 4 == 0) move between HBM and shared memory one ROW per
// instruction: loads complete on a warp's mbarrier, stores / reduce-adds belong to the issuing thread's bulk group.
// Rows sit SH_BULK_STRIDE floats apart in shared memory: 16-byte aligned for the copy engine, and conflict-free for
// per-lane 128-bit accesses (52 floats = 13 x 16 B, odd).
constexpr int SH_BULK_STRIDE = 52;  // floats per shared-memory row in bulk mode (M <= 16)
#ifdef __CUDACC__
// Transpose of the 32 x 32 bit matrix held one row per lane: on return, bit k of lane i is what bit i of lane k was.
// Five block-swap stages (16, 8, 4, 2, 1) of one shuffle each: ~30 warp-instructions, where the loop
// `for j: mine |= ((row[j] >> lane) & 1) << j` over shared memory costs ~130.  All 32 lanes must call it.
__device__ __forceinline__ uint32_t warp_transpose_bits(uint32_t x, int lane) {
#pragma unroll
  for (int k = 0; k < 5; k++) {
    const int j = 16 >> k;
    const uint32_t m = k == 0 ? 0x0000ffffu : k == 1 ? 0x00ff00ffu : k == 2 ? 0x0f0f0f0fu : k == 3 ? 0x33333333u : 0x55555555u;
    const uint32_t y = __shfl_xor_sync(0xffffffffu, x, j);
    x = (lane & j) ? (((y >> j) & m) | (x & ~m)) : ((x & m) | ((y & m) << j));
  }
  return x;
}

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint64_t* bar, int count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count) : "memory");
  asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void bulk_load(void* dst_smem, const void* src_gmem, uint32_t bytes, uint64_t* bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
               ::"r"(smem_u32(dst_smem)), "l"(src_gmem), "r"(bytes), "r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
  asm volatile(
      "{\n.reg .pred p;\nWAIT_%=:\n"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
      "@p bra DONE_%=;\nbra WAIT_%=;\nDONE_%=:\n}"
      ::"r"(smem_u32(bar)), "r"(parity) : "memory");
}

__device__ __forceinline__ void fence_proxy_async() {  // this thread's generic-proxy shared-memory writes -> visible to the copy engine
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
}
__device__ __forceinline__ void bulk_store(void* dst_gmem, const void* src_smem, uint32_t bytes) {
  asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(dst_gmem), "r"(smem_u32(src_smem)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void bulk_reduce_add_f32(void* dst_gmem, const void* src_smem, uint32_t bytes) {
  asm volatile("cp.reduce.async.bulk.global.shared::cta.bulk_group.add.f32 [%0], [%1], %2;" ::"l"(dst_gmem), "r"(smem_u32(src_smem)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void bulk_commit_and_wait_read() {  // the source rows may be reused / the CTA may exit afterwards
  asm volatile("cp.async.bulk.commit_group;" ::: "memory");
  asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory");
}
#endif

// ---- stage launchers (one translation unit each) ----------------------------------------
struct FwdPreprocessParams {
  int P, D, M, W, H;
  int grid_x, grid_y;
  float scale_modifier;
  int prefiltered;
  const float* means3D;
  const float* scales;
  const float* rotations;
  const float* opacities;
  const float* shs;
  const float* shs_rest;  // optional: shs = features_dc [P,1,3], shs_rest = features_rest [P,M-1,3]
  int raw;                // scales / rotations / opacities are the raw parameters: activated in-kernel
  const float* transMat_precomp;
  const float* colors_precomp;
  const float* viewmatrix;
  const float* projmatrix;
  const float* campos;
  int* radii;
  float4* records;
  short4* bbox;
  float4* cull;
  uint8_t* clamp;
  uint4* rects;
  uint32_t* tile_count;  // u32[Nt], zeroed by the caller: += 1 per (surfel, tile) instance
  uint32_t* flags;
};
void launch_preprocess_fwd(const FwdPreprocessParams& p, cudaStream_t stream);
void launch_mark_visible(int P, const float* means3D, const float* viewmatrix, uint8_t* present,
                         cudaStream_t stream);

void launch_tile_scan(const uint32_t* tile_count, uint32_t* tile_cursor, uint2* ranges, int num_tiles,
                      unsigned long long capacity, const uint32_t* flags, ForwardStatus* status, cudaStream_t stream);
void launch_scatter(int P, const uint4* rects, const int* radii, uint32_t* tile_cursor, uint2* pairs,
                    int grid_x, const short4* bbox, const float4* cull, int no_cull, int packed,
                    const ForwardStatus* status, cudaStream_t stream);
// packed: the scatter ran with P <= 2^24 and left the block masks in the top byte of the pairs' values
void launch_tile_sort(const uint2* ranges, uint2* pairs, uint64_t* keys_sorted, uint32_t* point_list,
                      uint8_t* strip_mask, const short4* bbox, const float4* cull, int grid_x, int num_tiles, int no_cull,
                      int packed, const ForwardStatus* status, cudaStream_t stream);

struct FwdRenderParams {
  int W, H, grid_x, grid_y;
  const uint2* ranges;
  const uint32_t* point_list;
  const float4* records;
  uint8_t* strip_mask;        // read, and refined: bits of (block, entry) pairs without a blended pixel are cleared
  uint32_t* pix_mask;         // out: [8][pix_plane] pixels of block b that blended list entry i, at [b * pix_plane + i]
  size_t pix_plane;
  const short4* bbox;         // per-surfel conservative contribution boxes
  int no_cull;                // evaluate every (pixel, entry) pair (debug bit 1)
  const float* background;
  float* final_T;
  uint32_t* n_contrib;
  float* out_color;
  float* out_others;
  unsigned long long* stats;  // optional work counters (NULL in production)
  const ForwardStatus* status;  // overflow => the frame is composited as empty
};
void launch_render_fwd(const FwdRenderParams& p, cudaStream_t stream);

struct BwdRenderParams {
  int W, H, grid_x, grid_y;
  const uint2* ranges;
  const uint32_t* point_list;
  const float4* records;
  const uint8_t* strip_mask;
  const uint32_t* pix_mask;   // as left by the forward compositing
  size_t pix_plane;
  const short4* bbox;
  int no_cull;
  const float* background;
  const float* final_T;
  const uint32_t* n_contrib;
  const float* dL_dpix;
  const float* dL_dothers;
  float* grad_records;  // [P*GREC], zero-initialised
  // (after an overflowed asynchronous forward n_contrib is 0 everywhere: the replay finds nothing to do)
};
void launch_render_bwd(const BwdRenderParams& p, cudaStream_t stream);

struct BwdPreprocessParams {
  int P, D, M, W, H;  // W,H: the sizes re-derived in fp32 as backward.cu:613-614 does
  int accumulate;     // outputs (except dL_dmeans2D) += instead of =
  const float* means3D;
  const float* scales;
  const float* rotations;
  const float* shs;
  const float* transMat_precomp;
  const float* colors_precomp;
  const float* viewmatrix;
  const float* projmatrix;
  const float* campos;
  const int* radii;
  const float4* records;
  const uint8_t* clamp;
  const float* grad_records;
  float* dL_dmeans2D;
  float* dL_dcolors;
  float* dL_dopacity;
  float* dL_dmeans3D;
  float* dL_dtransMat;
  float* dL_dsh;
  float* dL_dscales;
  float* dL_drotations;
  float* dL_drgb;  // optional [P,3]: factored mode, written INSTEAD of dL_dsh (see sh_grad_expand_kernel)
  float* rgb_peers[GSM_MAX_PEERS];  // factored mode over NVLink: the [P,3] slot of THIS view in every rank's receive buffer
  int n_rgb_peers;                  // (peer memory, 16-byte aligned); the rows are pushed there instead of dL_drgb
  float* stat_norm;  // optional [P]: += ||dL_dmeans2D.xy|| of visible surfels
  float* stat_vis;   // optional [P]: += 1 for visible surfels
  int* max_radii;    // optional [P]: max= radii
  const float* shs_rest;  // optional: split SH storage (see FwdPreprocessParams)
  float* dL_dsh_rest;     // with shs_rest: dL_dsh = d/d features_dc [P,1,3], dL_dsh_rest = d/d features_rest [P,M-1,3]
  const float* opacities; // raw mode only: the raw opacities (sigmoid' needs them)
  int raw;                // gradients are taken w.r.t. the RAW scales / rotations / opacities
};
void launch_preprocess_bwd(const BwdPreprocessParams& p, cudaStream_t stream);
void launch_sh_grad_expand(int P, int D, int M, int n_views, const float* means3D, const float* campos_all,
                           const float* drgb_all, size_t view_stride, float* dL_dsh, int accumulate, cudaStream_t stream);

}  // namespace gsm
