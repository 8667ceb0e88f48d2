// The all-reduce of the view-parallel exchange as one hand-written kernel over NVLink peer memory (new work: the
// reference is single-GPU).  Two-shot: with n ranks, rank r owns shard r of the buffer -- it loads that
// shard from EVERY rank's copy (peer loads over NVLink / NVSwitch), reduces in the fixed order rank 0..n-1, and stores
// the result into EVERY rank's copy (peer stores).  Each rank moves (n-1)/n of the buffer in and out, the bandwidth-
// optimal amount, in one launch; the sum is bitwise identical on all ranks.  The buffer is n_sum fp32 to be summed
// (surfel gradients 40 B/surfel + the two densification sums) followed by n_max int32 to be maxed (max_radii), so the
// three NCCL launches of round 1 (all-reduce SUM, all-reduce SUM, all-reduce MAX) are one kernel.
// The caller brackets it with cross-rank barriers (symmetric-memory signal pads) on the stream.
#include "common.cuh"

namespace gsm {

struct PeerBufs { void* p[GSM_MAX_PEERS]; };

__device__ __forceinline__ float4 ld_peer(const float4* a) {  // peer data is written by other GPUs: no read-only cache
  float4 v;
  asm volatile("ld.global.relaxed.sys.v4.f32 {%0,%1,%2,%3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "l"(a) : "memory");
  return v;
}

template <int NR>
__global__ void __launch_bounds__(512) peer_allreduce_kernel(const PeerBufs bufs, int rank, size_t sum_vec, size_t max_vec) {
  // vectors of 16 bytes; [0, sum_vec) fp32 sums, [sum_vec, sum_vec + max_vec) int32 maxima
  const size_t total = sum_vec + max_vec;
  const size_t per = (total + NR - 1) / NR;
  const size_t begin = per * (size_t)rank, end = min(begin + per, total);
  for (size_t v = begin + (size_t)blockIdx.x * blockDim.x + threadIdx.x; v < end; v += (size_t)gridDim.x * blockDim.x) {
    float4 x[NR];
#pragma unroll
    for (int k = 0; k < NR; k++) x[k] = ld_peer(reinterpret_cast<const float4*>(bufs.p[k]) + v);   // NR loads in flight
    float4 r = x[0];
    if (v < sum_vec) {
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
    for (int k = 0; k < NR; k++) reinterpret_cast<float4*>(bufs.p[k])[v] = r;
  }
}

}  // namespace gsm

using namespace gsm;

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
  switch (n_ranks) {
    case 2: peer_allreduce_kernel<2><<<grid, 512, 0, stream>>>(pb, rank, sum_vec, max_vec); break;
    case 3: peer_allreduce_kernel<3><<<grid, 512, 0, stream>>>(pb, rank, sum_vec, max_vec); break;
    case 4: peer_allreduce_kernel<4><<<grid, 512, 0, stream>>>(pb, rank, sum_vec, max_vec); break;
    case 5: peer_allreduce_kernel<5><<<grid, 512, 0, stream>>>(pb, rank, sum_vec, max_vec); break;
    case 6: peer_allreduce_kernel<6><<<grid, 512, 0, stream>>>(pb, rank, sum_vec, max_vec); break;
    case 7: peer_allreduce_kernel<7><<<grid, 512, 0, stream>>>(pb, rank, sum_vec, max_vec); break;
    default: peer_allreduce_kernel<8><<<grid, 512, 0, stream>>>(pb, rank, sum_vec, max_vec); break;
  }
  GSM_CUDA(cudaGetLastError());
  return GSM_OK;
}
