// exchange.cu — the one exchange on the path: the per-Gaussian gradients of a view-sharded scene summed over
// the GPUs of an NVSwitch domain, in place, through the switch's reduction unit (NVLS).
//
// Every rank holds a replica of the gradient slab in symmetric memory; `mc` is the MULTICAST alias of those replicas.
// Rank r owns the r-th of `world` equal parts of every piece.  For each 16-byte unit of its part it issues
//   multimem.ld_reduce.add.v4.f32   the switch fetches the unit from every replica and returns the sum:  1 unit in
//   multimem.st.v4.f32              the switch writes the sum into every replica:                         1 unit out
// so a GPU's links carry (1 + 1/world) x the slab in each direction whatever the number of GPUs - where adding every
// value with multimem.red from the producing kernel (TSPLAT_FLAG_BWD_MULTIMEM) delivers `world` slabs to every GPU.
// The caller brackets the launch with barriers over the ranks (all replicas written before / all sums stored after).
#include "common.cuh"

namespace tsplat {

struct ExchangePieces {
  unsigned long long begin[4];  // first 16-byte unit of the piece within the slab
  unsigned long long count[4];  // its length in 16-byte units
};

__device__ __forceinline__ float4 mc_ld_reduce_v4(const float4* mc_addr) {
  float4 v;
  asm volatile("multimem.ld_reduce.relaxed.sys.global.add.v4.f32 {%0, %1, %2, %3}, [%4];"
               : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w)
               : "l"(mc_addr)
               : "memory");
  return v;
}
__device__ __forceinline__ void mc_st_v4(float4* mc_addr, float4 v) {
  asm volatile("multimem.st.relaxed.sys.global.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(mc_addr), "f"(v.x), "f"(v.y), "f"(v.z),
               "f"(v.w)
               : "memory");
}

constexpr int kExThreads = 512;
constexpr int kExUnroll = 8;  // reductions in flight per thread: the round trip through the switch is microseconds long

__global__ void __launch_bounds__(kExThreads) nvls_all_reduce_kernel(float4* __restrict__ mc, ExchangePieces pc, int rank,
                                                                     int world) {
  const unsigned long long stride = (unsigned long long)gridDim.x * kExThreads;
  const unsigned long long t0 = (unsigned long long)blockIdx.x * kExThreads + threadIdx.x;
#pragma unroll 1
  for (int k = 0; k < 4; k++) {
    const unsigned long long n = pc.count[k];
    const unsigned long long lo = pc.begin[k] + n * rank / world, hi = pc.begin[k] + n * (rank + 1) / world;
    unsigned long long i = lo + t0;
    for (; i + (kExUnroll - 1) * stride < hi; i += kExUnroll * stride) {
      float4 v[kExUnroll];
#pragma unroll
      for (int u = 0; u < kExUnroll; u++) v[u] = mc_ld_reduce_v4(mc + i + u * stride);
#pragma unroll
      for (int u = 0; u < kExUnroll; u++) mc_st_v4(mc + i + u * stride, v[u]);
    }
    for (; i < hi; i += stride) mc_st_v4(mc + i, mc_ld_reduce_v4(mc + i));
  }
}

}  // namespace tsplat

extern "C" int tsplat_nvls_all_reduce(void* multicast_slab, const uint64_t* piece_offset_floats,
                                      const uint64_t* piece_count_floats, int32_t num_pieces, int32_t rank, int32_t world,
                                      int32_t num_ctas, void* cuda_stream) {
  using namespace tsplat;
  if (!multicast_slab || num_pieces < 0 || num_pieces > 4 || world < 1 || rank < 0 || rank >= world ||
      (reinterpret_cast<uintptr_t>(multicast_slab) & 15u) != 0) {
    set_error("tsplat_nvls_all_reduce: bad argument (null or unaligned slab, more than 4 pieces, rank outside the world)");
    return TSPLAT_EINVAL;
  }
  ExchangePieces pc{};
  unsigned long long total = 0;
  for (int k = 0; k < num_pieces; k++) {
    if (piece_offset_floats[k] % 4 != 0 || piece_count_floats[k] % 4 != 0) {  // whole 16-byte units
      set_error("tsplat_nvls_all_reduce: piece %d is not a whole number of 16-byte units", k);
      return TSPLAT_EINVAL;
    }
    pc.begin[k] = piece_offset_floats[k] / 4;
    pc.count[k] = piece_count_floats[k] / 4;
    total += pc.count[k];
  }
  if (total == 0) return TSPLAT_OK;
  int ctas = num_ctas > 0 ? num_ctas : 64;
  const unsigned long long need = (total / world + kExThreads - 1) / kExThreads;
  if ((unsigned long long)ctas > need) ctas = (int)(need ? need : 1);
  nvls_all_reduce_kernel<<<ctas, kExThreads, 0, static_cast<cudaStream_t>(cuda_stream)>>>(
      static_cast<float4*>(multicast_slab), pc, rank, world);
  const cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) {
    set_error("nvls_all_reduce_kernel: %s", cudaGetErrorString(e));
    return TSPLAT_ECUDA;
  }
  return TSPLAT_OK;
}
