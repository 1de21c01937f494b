// dk_reduce.cuh — device which.max / top-k over a criterion vector (R/max_crit.R:107 semantics:
// first maximal index, NaN skipped).  Selection is by the strict total order
// (value descending, index ascending), so the result does not depend on block scheduling.
// Pass t finds the best entry strictly after result[t-1] in that order; k passes give the k best.
#pragma once
#include <cmath>

#include "dk_common.cuh"

#define DK_TOPK_MAX 64

namespace dk {

struct TopkEntry {
  double val;
  long long idx;  // -1 = empty
};

constexpr int RED_BLOCKS = 296;
constexpr int RED_THREADS = 256;

__host__ __device__ inline size_t topk_workspace_bytes(int) {
  return sizeof(TopkEntry) * DK_TOPK_MAX + 64 + sizeof(TopkEntry) * RED_BLOCKS;
}
inline TopkEntry* topk_result_ptr(void* ws) { return reinterpret_cast<TopkEntry*>(ws); }

__device__ __forceinline__ bool topk_better(const TopkEntry& a, const TopkEntry& b) {
  if (a.idx < 0) return false;
  if (b.idx < 0) return true;
  return a.val > b.val || (a.val == b.val && a.idx < b.idx);
}

static __global__ void __launch_bounds__(RED_THREADS)
k_topk_pass(const double* __restrict__ v, int64_t N, int t, double sign, unsigned char* __restrict__ ws) {
  TopkEntry* result = reinterpret_cast<TopkEntry*>(ws);
  unsigned* counter = reinterpret_cast<unsigned*>(ws + sizeof(TopkEntry) * DK_TOPK_MAX);
  TopkEntry* partial = reinterpret_cast<TopkEntry*>(ws + sizeof(TopkEntry) * DK_TOPK_MAX + 64);
  __shared__ TopkEntry sh[RED_THREADS / 32];
  __shared__ bool is_last;

  TopkEntry prev{0.0, -1};
  const bool have_prev = t > 0;
  if (have_prev) prev = result[t - 1];
  TopkEntry best{0.0, -1};
  if (!have_prev || prev.idx >= 0) {
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < N; i += (int64_t)gridDim.x * blockDim.x) {
      const double x = sign * v[i];  // sign = -1: the k SMALLEST values (result values carry the sign)
      if (x != x) continue;  // NaN skipped, like which.max
      if (have_prev && !(x < prev.val || (x == prev.val && i > prev.idx))) continue;
      const TopkEntry cand{x, (long long)i};
      if (topk_better(cand, best)) best = cand;
    }
  }
  auto warp_reduce = [](TopkEntry b) {
#pragma unroll
    for (int off = 16; off > 0; off >>= 1) {
      TopkEntry o;
      o.val = __shfl_xor_sync(0xffffffffu, b.val, off);
      o.idx = __shfl_xor_sync(0xffffffffu, b.idx, off);
      if (topk_better(o, b)) b = o;
    }
    return b;
  };
  best = warp_reduce(best);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (lane == 0) sh[warp] = best;
  __syncthreads();
  if (warp == 0) {
    TopkEntry b = lane < RED_THREADS / 32 ? sh[lane] : TopkEntry{0.0, -1};
    b = warp_reduce(b);
    if (lane == 0) {
      partial[blockIdx.x] = b;
      __threadfence();
      const unsigned ticket = atomicAdd(counter, 1u);
      is_last = (ticket == gridDim.x - 1);
    }
  }
  __syncthreads();
  if (is_last && warp == 0) {
    __threadfence();
    TopkEntry b{0.0, -1};
    for (int i = lane; i < (int)gridDim.x; i += 32) {
      const TopkEntry o{__ldcg(&partial[i].val), __ldcg(&partial[i].idx)};
      if (topk_better(o, b)) b = o;
    }
    b = warp_reduce(b);
    if (lane == 0) {
      result[t] = b;
      *counter = 0u;  // ready for the next pass
    }
  }
}

// ---- counter-based uniform candidates (Philox4x32-10): coordinate k of candidate j is a pure function of (seed, j, k),
// so any device can generate any range of a screen's candidates and nothing is uploaded (SURVEY 8(b) "generator spec").
// counter = (j lo, j hi, k / 2, 0), key = (seed lo, seed hi); the four 32-bit outputs give two 53-bit uniforms in [0, 1),
// the first for even k, the second for odd k.
struct GenSpec {
  unsigned long long seed;
  double lower[DK_MAX_D], width[DK_MAX_D];
};
__host__ __device__ inline void philox4x32_10(unsigned c0, unsigned c1, unsigned c2, unsigned c3, unsigned k0, unsigned k1,
                                              unsigned out[4]) {
  for (int r = 0; r < 10; r++) {
    const unsigned long long p0 = (unsigned long long)0xD2511F53u * c0, p1 = (unsigned long long)0xCD9E8D57u * c2;
    const unsigned n0 = (unsigned)(p1 >> 32) ^ c1 ^ k0, n1 = (unsigned)p1, n2 = (unsigned)(p0 >> 32) ^ c3 ^ k1, n3 = (unsigned)p0;
    c0 = n0; c1 = n1; c2 = n2; c3 = n3;
    k0 += 0x9E3779B9u; k1 += 0xBB67AE85u;
  }
  out[0] = c0; out[1] = c1; out[2] = c2; out[3] = c3;
}
__host__ __device__ inline double gen_coordinate(const GenSpec& G, long long j, int k) {
  unsigned r[4];
  philox4x32_10((unsigned)j, (unsigned)((unsigned long long)j >> 32), (unsigned)(k >> 1), 0u, (unsigned)G.seed,
                (unsigned)(G.seed >> 32), r);
  const unsigned hi = (k & 1) ? r[2] : r[0], lo = (k & 1) ? r[3] : r[1];
  const unsigned long long bits = ((unsigned long long)hi << 21) | (lo >> 11);
  const double u = (double)bits * 1.1102230246251565e-16;  // 2^-53
#ifdef __CUDA_ARCH__
  return __fma_rn(u, G.width[k], G.lower[k]);
#else
  return std::fma(u, G.width[k], G.lower[k]);
#endif
}
// rows [first, first + rows) of the candidate matrix into out (rows x d, column-major, ld = rows)
static __global__ void __launch_bounds__(256)
k_gen_uniform(const GenSpec G, int d, long long first, long long rows, double* __restrict__ out) {
  const long long total = rows * d;
  for (long long t = (long long)blockIdx.x * blockDim.x + threadIdx.x; t < total; t += (long long)gridDim.x * blockDim.x) {
    const int k = (int)(t / rows);
    const long long j = t - (long long)k * rows;
    out[t] = gen_coordinate(G, first + j, k);
  }
}

inline int launch_topk(const double* d_v, int64_t N, int k, void* ws, cudaStream_t st, double sign = 1.0) {
  DK_CUDA_TRY(cudaMemsetAsync(reinterpret_cast<unsigned char*>(ws) + sizeof(TopkEntry) * DK_TOPK_MAX, 0, 64, st));
  for (int t = 0; t < k; t++) {
    k_topk_pass<<<RED_BLOCKS, RED_THREADS, 0, st>>>(d_v, N, t, sign, reinterpret_cast<unsigned char*>(ws));
    DK_CUDA_TRY(cudaGetLastError());
  }
  return DK_OK;
}

}  // namespace dk
