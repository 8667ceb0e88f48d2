// The all-reduce of the view-parallel exchange as one hand-written kernel over NVLink peer memory (new work: the
// reference is single-GPU).  Two-shot: with n ranks, rank r owns shard r of the buffer -- it loads that
// shard from EVERY rank's copy (peer loads over NVLink / NVSwitch), reduces in the fixed order rank 0..n-1, and stores
// the result into EVERY rank's copy (peer stores).  Each rank moves (n-1)/n of the buffer in and out, the bandwidth-
// optimal amount, in one launch; the sum is bitwise identical on all ranks.  The buffer is n_sum fp32 to be summed
// (surfel gradients 40 B/surfel + the two densification sums) followed by n_max int32 to be maxed (max_radii), so the
// three NCCL launches of round 1 (all-reduce SUM, all-reduce SUM, all-reduce MAX) are one kernel.
// The caller brackets it with cross-rank barriers (symmetric-memory signal pads) on the stream.
#include "common.cuh"
#include <cstdlib>

namespace gsm {

struct PeerBufs { void* p[GSM_MAX_PEERS]; };

__device__ __forceinline__ float4 ld_peer(const float4* a) {  // peer data is written by other GPUs: no read-only cache
  float4 v;
  asm volatile("ld.global.relaxed.sys.v4.f32 {%0,%1,%2,%3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "l"(a) : "memory");
  return v;
}

// SYS: system-scope relaxed loads (never served from a stale L1 line); otherwise plain loads -- safe here because every
// byte this kernel reads was written before the barrier that precedes its launch, and L1 starts a kernel invalidated
// 256 threads, ONE CTA per SM: the kernel is bound by NVLink (~0.7 ms for 260 MB at 8 ranks, NCCL's rate), not by
// the SMs, and it runs NEXT TO the local SH expand -- a fat launch (2 x 512 threads per SM) took the register file away
// from the expand, which then ran at less than half its speed.
constexpr int PEER_THREADS = 256;
template <int NR, bool SYS>
__global__ void __launch_bounds__(PEER_THREADS) peer_allreduce_kernel(const PeerBufs bufs, int rank, size_t sum_vec, size_t max_vec) {
  // vectors of 16 bytes; [0, sum_vec) fp32 sums, [sum_vec, sum_vec + max_vec) int32 maxima
  const size_t total = sum_vec + max_vec;
  const size_t per = (total + NR - 1) / NR;
  const size_t begin = per * (size_t)rank, end = min(begin + per, total);
  const size_t stride = (size_t)gridDim.x * blockDim.x;
  constexpr int U = NR <= 2 ? 8 : (NR <= 4 ? 4 : 3);   // U x NR 128-bit peer loads in flight per thread
  for (size_t v0 = begin + (size_t)blockIdx.x * blockDim.x + threadIdx.x; v0 < end; v0 += stride * U) {
    float4 x[U][NR];
#pragma unroll
    for (int u = 0; u < U; u++) {
      const size_t v = min(v0 + stride * u, end - 1);
#pragma unroll
      for (int k = 0; k < NR; k++)
        x[u][k] = SYS ? ld_peer(reinterpret_cast<const float4*>(bufs.p[k]) + v) : reinterpret_cast<const float4*>(bufs.p[k])[v];
    }
#pragma unroll
    for (int u = 0; u < U; u++) {
      const size_t v = v0 + stride * u;
      if (v >= end) continue;
      float4 r = x[u][0];
      if (v < sum_vec) {
#pragma unroll
        for (int k = 1; k < NR; k++) { r.x += x[u][k].x; r.y += x[u][k].y; r.z += x[u][k].z; r.w += x[u][k].w; }
      } else {
#pragma unroll
        for (int k = 1; k < NR; k++) {
          r.x = __int_as_float(max(__float_as_int(r.x), __float_as_int(x[u][k].x)));
          r.y = __int_as_float(max(__float_as_int(r.y), __float_as_int(x[u][k].y)));
          r.z = __int_as_float(max(__float_as_int(r.z), __float_as_int(x[u][k].z)));
          r.w = __int_as_float(max(__float_as_int(r.w), __float_as_int(x[u][k].w)));
        }
      }
#pragma unroll
      for (int k = 0; k < NR; k++) reinterpret_cast<float4*>(bufs.p[k])[v] = r;
    }
  }
}

// ---- the same all-reduce with WRITES only on the wire ("push" form) --------------------------------------------
// Measured on 8 B200s: a kernel that LOADS from peers moves ~350 GB/s per direction and, while it runs, starves every
// other kernel's memory traffic (the SH expand next to it made 10 % progress); posted peer STORES fused into a
// streaming kernel are nearly free (K6 pushes 420 MB of dL/dRGB rows for +0.19 ms).  So:
//   1. scatter: every rank pushes shard s of its partial buffer into rank s's staging area, slot [own rank];
//   2. (barrier)  sum + broadcast: rank s sums the n staged copies of its shard -- LOCAL loads, fixed order -- and
//      pushes the result into every rank's buffer.
template <int NR>
__global__ void __launch_bounds__(PEER_THREADS) peer_scatter_kernel(const float4* __restrict__ local, const PeerBufs staging,
                                                                     int rank, size_t total, size_t per) {
  const size_t stride = (size_t)gridDim.x * blockDim.x;
  for (size_t v = (size_t)blockIdx.x * blockDim.x + threadIdx.x; v < total; v += stride) {
    const size_t s = v / per;   // owner of this vector
    reinterpret_cast<float4*>(staging.p[s])[(size_t)rank * per + (v - s * per)] = __ldcs(local + v);
  }
}

template <int NR>
__global__ void __launch_bounds__(PEER_THREADS) peer_sum_broadcast_kernel(const float4* __restrict__ staged, const PeerBufs flats,
                                                                           int rank, size_t sum_vec, size_t max_vec, size_t per) {
  const size_t total = sum_vec + max_vec;
  const size_t begin = per * (size_t)rank, n = begin < total ? min(per, total - begin) : 0;
  const size_t stride = (size_t)gridDim.x * blockDim.x;
  for (size_t o = (size_t)blockIdx.x * blockDim.x + threadIdx.x; o < n; o += stride) {
    float4 x[NR];
#pragma unroll
    for (int k = 0; k < NR; k++) x[k] = staged[(size_t)k * per + o];   // written by rank k before the barrier: plain loads
    float4 r = x[0];
    if (begin + o < sum_vec) {
#pragma unroll
      for (int k = 1; k < NR; k++) { r.x += x[k].x; r.y += x[k].y; r.z += x[k].z; r.w += x[k].w; }
    } else {
#pragma unroll
      for (int k = 1; k < NR; k++) {
        r.x = __int_as_float(max(__float_as_int(r.x), __float_as_int(x[k].x)));
        r.y = __int_as_float(max(__float_as_int(r.y), __float_as_int(x[k].y)));
        r.z = __int_as_float(max(__float_as_int(r.z), __float_as_int(x[k].z)));
        r.w = __int_as_float(max(__float_as_int(r.w), __float_as_int(x[k].w)));
      }
    }
#pragma unroll
    for (int k = 0; k < NR; k++) reinterpret_cast<float4*>(flats.p[k])[begin + o] = r;
  }
}

// The same reduction through the NVSwitch's multicast address space (NVLS): multimem.ld_reduce has the SWITCH fetch
// the element from every rank's copy and return their sum (max) in one response, multimem.st has it write the result
// into every rank's copy -- one request per 16 bytes and direction instead of n_ranks, the additions done in the fabric.
__global__ void __launch_bounds__(PEER_THREADS) peer_allreduce_multimem_kernel(char* mc, int n_ranks, int rank, size_t sum_vec, size_t max_vec) {
  const size_t total = sum_vec + max_vec;
  const size_t per = (total + n_ranks - 1) / n_ranks;
  const size_t begin = per * (size_t)rank, end = min(begin + per, total);
  const size_t stride = (size_t)gridDim.x * blockDim.x;
  constexpr int U = 8;   // requests in flight per thread: a round trip through the switch is several microseconds
  for (size_t v0 = begin + (size_t)blockIdx.x * blockDim.x + threadIdx.x; v0 < end; v0 += stride * U) {
    float4 r[U];
#pragma unroll
    for (int u = 0; u < U; u++) {
      const size_t v = v0 + stride * u;
      char* a = mc + v * 16;
      if (v >= end) continue;
      if (v < sum_vec) {
        asm volatile("multimem.ld_reduce.relaxed.sys.global.add.v4.f32 {%0,%1,%2,%3}, [%4];"
                     : "=f"(r[u].x), "=f"(r[u].y), "=f"(r[u].z), "=f"(r[u].w) : "l"(a) : "memory");
      } else {
        int q[4];
#pragma unroll
        for (int k = 0; k < 4; k++) asm volatile("multimem.ld_reduce.relaxed.sys.global.max.s32 %0, [%1];" : "=r"(q[k]) : "l"(a + 4 * k) : "memory");
        r[u] = make_float4(__int_as_float(q[0]), __int_as_float(q[1]), __int_as_float(q[2]), __int_as_float(q[3]));
      }
    }
#pragma unroll
    for (int u = 0; u < U; u++) {
      const size_t v = v0 + stride * u;
      if (v >= end) continue;
      asm volatile("multimem.st.relaxed.sys.global.v4.f32 [%0], {%1,%2,%3,%4};" ::"l"(mc + v * 16), "f"(r[u].x), "f"(r[u].y), "f"(r[u].z), "f"(r[u].w) : "memory");
    }
  }
}

}  // namespace gsm

using namespace gsm;

extern "C" int gsm_peer_allreduce_multimem(void* multicast_buf, int n_ranks, int rank, size_t n_sum, size_t n_max, void* stream_) {
  cudaStream_t stream = (cudaStream_t)stream_;
  if (multicast_buf == nullptr || n_ranks < 1 || rank < 0 || rank >= n_ranks || (n_sum & 3) || (n_max & 3) ||
      (reinterpret_cast<size_t>(multicast_buf) & 15) != 0) {
    set_error("gsm_peer_allreduce_multimem: bad arguments"); return GSM_ERR_INVALID_ARGUMENT;
  }
  const size_t sum_vec = n_sum / 4, max_vec = n_max / 4;
  if (sum_vec + max_vec == 0 || n_ranks == 1) return GSM_OK;
  int dev = 0, sms = 0;
  GSM_CUDA(cudaGetDevice(&dev));
  GSM_CUDA(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
  const size_t per = (sum_vec + max_vec + n_ranks - 1) / n_ranks;
  const int grid = (int)min((size_t)sms, (per + PEER_THREADS - 1) / PEER_THREADS);
  StageMark m(GSM_STAGE_PEER_ALLREDUCE, stream);
  peer_allreduce_multimem_kernel<<<grid, PEER_THREADS, 0, stream>>>((char*)multicast_buf, n_ranks, rank, sum_vec, max_vec);
  GSM_CUDA(cudaGetLastError());
  return GSM_OK;
}

extern "C" int gsm_peer_allreduce(void* const* bufs, int n_ranks, int rank, size_t n_sum, size_t n_max, void* stream_) {
  cudaStream_t stream = (cudaStream_t)stream_;
  if (bufs == nullptr || n_ranks < 1 || n_ranks > GSM_MAX_PEERS || rank < 0 || rank >= n_ranks || (n_sum & 3) || (n_max & 3)) {
    set_error("gsm_peer_allreduce: bad arguments (1..%d ranks, counts multiples of 4)", GSM_MAX_PEERS); return GSM_ERR_INVALID_ARGUMENT;
  }
  PeerBufs pb{};
  for (int k = 0; k < n_ranks; k++) {
    if (bufs[k] == nullptr || (reinterpret_cast<size_t>(bufs[k]) & 15) != 0) { set_error("gsm_peer_allreduce: buffer %d null or misaligned", k); return GSM_ERR_INVALID_ARGUMENT; }
    pb.p[k] = bufs[k];
  }
  const size_t sum_vec = n_sum / 4, max_vec = n_max / 4;
  if (sum_vec + max_vec == 0 || n_ranks == 1) return GSM_OK;
  int dev = 0, sms = 0;
  GSM_CUDA(cudaGetDevice(&dev));
  GSM_CUDA(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
  const size_t per = (sum_vec + max_vec + n_ranks - 1) / n_ranks;
  const int grid = (int)min((size_t)sms * 2, (per + 511) / 512);
  StageMark m(GSM_STAGE_PEER_ALLREDUCE, stream);
  static const bool sys_loads = [] { const char* e = getenv("GSM_PEER_LD"); return e != nullptr && e[0] == 's'; }();
#define GSM_PEER_CASE(N) \
  case N: if (sys_loads) peer_allreduce_kernel<N, true><<<grid, PEER_THREADS, 0, stream>>>(pb, rank, sum_vec, max_vec); \
          else peer_allreduce_kernel<N, false><<<grid, PEER_THREADS, 0, stream>>>(pb, rank, sum_vec, max_vec); break;
  switch (n_ranks) {
    GSM_PEER_CASE(2) GSM_PEER_CASE(3) GSM_PEER_CASE(4) GSM_PEER_CASE(5) GSM_PEER_CASE(6) GSM_PEER_CASE(7)
    default: if (sys_loads) peer_allreduce_kernel<8, true><<<grid, PEER_THREADS, 0, stream>>>(pb, rank, sum_vec, max_vec);
             else peer_allreduce_kernel<8, false><<<grid, PEER_THREADS, 0, stream>>>(pb, rank, sum_vec, max_vec); break;
  }
#undef GSM_PEER_CASE
  GSM_CUDA(cudaGetLastError());
  return GSM_OK;
}

static int check_peers(void* const* bufs, int n_ranks, int rank, size_t n_sum, size_t n_max, const char* what) {
  if (bufs == nullptr || n_ranks < 1 || n_ranks > GSM_MAX_PEERS || rank < 0 || rank >= n_ranks || (n_sum & 3) || (n_max & 3)) {
    set_error("%s: bad arguments (1..%d ranks, counts multiples of 4)", what, GSM_MAX_PEERS); return GSM_ERR_INVALID_ARGUMENT;
  }
  for (int k = 0; k < n_ranks; k++)
    if (bufs[k] == nullptr || (reinterpret_cast<size_t>(bufs[k]) & 15) != 0) { set_error("%s: buffer %d null or misaligned", what, k); return GSM_ERR_INVALID_ARGUMENT; }
  return GSM_OK;
}

extern "C" size_t gsm_peer_staging_floats(int n_ranks, size_t n_sum, size_t n_max) {
  const size_t total = (n_sum + n_max) / 4;
  const size_t per = (total + (size_t)(n_ranks > 0 ? n_ranks : 1) - 1) / (size_t)(n_ranks > 0 ? n_ranks : 1);
  return per * 4 * (size_t)(n_ranks > 0 ? n_ranks : 1);
}

extern "C" int gsm_peer_scatter(const void* local_buf, void* const* staging, int n_ranks, int rank, size_t n_sum, size_t n_max,
                                void* stream_) {
  cudaStream_t stream = (cudaStream_t)stream_;
  if (int rc = check_peers(staging, n_ranks, rank, n_sum, n_max, "gsm_peer_scatter")) return rc;
  if (local_buf == nullptr || (reinterpret_cast<size_t>(local_buf) & 15) != 0) { set_error("gsm_peer_scatter: local buffer"); return GSM_ERR_INVALID_ARGUMENT; }
  const size_t total = (n_sum + n_max) / 4;
  if (total == 0) return GSM_OK;
  const size_t per = (total + n_ranks - 1) / n_ranks;
  PeerBufs pb{};
  for (int k = 0; k < n_ranks; k++) pb.p[k] = staging[k];
  int dev = 0, sms = 0;
  GSM_CUDA(cudaGetDevice(&dev));
  GSM_CUDA(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
  const int grid = (int)min((size_t)sms * 4, (total + PEER_THREADS - 1) / PEER_THREADS);
  StageMark m(GSM_STAGE_PEER_ALLREDUCE, stream);
  peer_scatter_kernel<8><<<grid, PEER_THREADS, 0, stream>>>((const float4*)local_buf, pb, rank, total, per);
  GSM_CUDA(cudaGetLastError());
  return GSM_OK;
}

extern "C" int gsm_peer_sum_broadcast(const void* my_staging, void* const* bufs, int n_ranks, int rank, size_t n_sum, size_t n_max,
                                      void* stream_) {
  cudaStream_t stream = (cudaStream_t)stream_;
  if (int rc = check_peers(bufs, n_ranks, rank, n_sum, n_max, "gsm_peer_sum_broadcast")) return rc;
  if (my_staging == nullptr || (reinterpret_cast<size_t>(my_staging) & 15) != 0) { set_error("gsm_peer_sum_broadcast: staging buffer"); return GSM_ERR_INVALID_ARGUMENT; }
  const size_t sum_vec = n_sum / 4, max_vec = n_max / 4, total = sum_vec + max_vec;
  if (total == 0) return GSM_OK;
  const size_t per = (total + n_ranks - 1) / n_ranks;
  PeerBufs pb{};
  for (int k = 0; k < n_ranks; k++) pb.p[k] = bufs[k];
  int dev = 0, sms = 0;
  GSM_CUDA(cudaGetDevice(&dev));
  GSM_CUDA(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
  const int grid = (int)min((size_t)sms * 2, (per + PEER_THREADS - 1) / PEER_THREADS);
  StageMark m(GSM_STAGE_PEER_ALLREDUCE, stream);
  const float4* st = (const float4*)my_staging;
#define GSM_SB_CASE(N) case N: peer_sum_broadcast_kernel<N><<<grid, PEER_THREADS, 0, stream>>>(st, pb, rank, sum_vec, max_vec, per); break;
  switch (n_ranks) {
    GSM_SB_CASE(1) GSM_SB_CASE(2) GSM_SB_CASE(3) GSM_SB_CASE(4) GSM_SB_CASE(5) GSM_SB_CASE(6) GSM_SB_CASE(7)
    default: peer_sum_broadcast_kernel<8><<<grid, PEER_THREADS, 0, stream>>>(st, pb, rank, sum_vec, max_vec, per); break;
  }
#undef GSM_SB_CASE
  GSM_CUDA(cudaGetLastError());
  return GSM_OK;
}
