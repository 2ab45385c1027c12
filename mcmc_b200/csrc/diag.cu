// diag.cu — cross-chain convergence diagnostics on the device.
//
// Replaces the ArviZ calls of the reference's compute_diagnostics (run_benchmarks.py:161-195:
// az.rhat rank-normalised split R-hat, az.ess bulk / tail, az.summary mcse_mean) for chain counts
// where a CPU pass is hopeless (65,536 chains x 256 draws x 20 dims = 3.4e8 values).  ArviZ is a
// third-party dependency; the algorithm is the published one (Vehtari et al. 2021), restated in
// oracle/diagnostics_np.py, which is the checker of these kernels.
//
//   rank_normalize_kernel      z = Phi^-1((r - 3/8)/(N + 1/4)), r = average rank over ties, from the
//                              sorted values + sorting permutation of one variable
//   split_chain_stats_kernel   per (chain, half) sequence: mean and biased autocovariance at every lag
//                              (direct O(n^2/2) products from a shared-memory tile, not an FFT:
//                              43 GFLOP for 2.6M sequences of 128 draws), summed over the block's 32
//                              sequences in a fixed shuffle order
//   split_chain_final_kernel   fixed-order sum of the block partials -> sum_seq acov[t]; Chan merge of
//                              the (count, mean, M2) of the sequence means  => deterministic results
#include "common.cuh"

namespace mcmc {

constexpr int kDiagLanes = 32;   // sequences per block (one per lane)
constexpr int kDiagWarps = 8;    // lag groups are dealt round-robin to the warps
constexpr int kLagGroup = 8;     // lags per register window

__global__ void __launch_bounds__(256) rank_normalize_kernel(const double* __restrict__ v, const int64_t* __restrict__ perm,
                                                             int64_t n, double* __restrict__ out) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const double x = v[i];
  // extent [lo, hi] of the run of values equal to x: binary searches on the sorted array, and only when a neighbour
  // ties (O(log n) per element however long the run — a constant column is one run of n)
  int64_t lo = i, hi = i;
  if (i > 0 && v[i - 1] == x) {
    int64_t a = 0, b = i - 1;             // first index in [0, i-1] with v[.] == x  (v[b] == x holds)
    while (a < b) {
      const int64_t m = a + ((b - a) >> 1);
      if (v[m] == x) b = m; else a = m + 1;
    }
    lo = a;
  }
  if (i + 1 < n && v[i + 1] == x) {
    int64_t a = i + 1, b = n - 1;         // last index in [i+1, n-1] with v[.] == x  (v[a] == x holds)
    while (a < b) {
      const int64_t m = a + ((b - a + 1) >> 1);
      if (v[m] == x) a = m; else b = m - 1;
    }
    hi = a;
  }
  const double r = 0.5 * (double)(lo + hi) + 1.0;            // average of the 1-based ranks lo+1 .. hi+1
  const double p = (r - 0.375) / ((double)n + 0.25);
  out[perm[i]] = normcdfinv(p);
}

// y[s*stride_s + c*stride_c]; rows [0, n) are the first half, rows [row2, row2 + n) the second half.
// grid = (ceil(C/32), 2).  tile: n x 32 doubles, in shared memory (SMEM) or in a global scratch slab.
template <bool SMEM>
__global__ void __launch_bounds__(kDiagLanes * kDiagWarps)
split_chain_stats_kernel(const double* __restrict__ y, int64_t C, int64_t stride_s, int64_t stride_c, int n,
                         int64_t row2, int n_lags, double* __restrict__ part_acov, double* __restrict__ part_stats,
                         double* __restrict__ scratch) {
  extern __shared__ __align__(16) double dsm[];
  __shared__ double s_red[kDiagWarps][kDiagLanes];
  __shared__ double s_mean[kDiagLanes];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int64_t blk = (int64_t)blockIdx.y * gridDim.x + blockIdx.x;
  double* tile = SMEM ? dsm : scratch + (size_t)blk * (size_t)(n + kLagGroup) * kDiagLanes;
  const int64_t c = (int64_t)blockIdx.x * kDiagLanes + lane;
  const bool valid = c < C;
  const int64_t r0 = blockIdx.y == 0 ? 0 : row2;
  double s = 0.0;
  for (int i = warp; i < n; i += kDiagWarps) {
    const double x = valid ? y[(r0 + i) * stride_s + c * stride_c] : 0.0;
    tile[(size_t)i * kDiagLanes + lane] = x;
    s += x;
  }
  for (int i = n + warp; i < n + kLagGroup; i += kDiagWarps) tile[(size_t)i * kDiagLanes + lane] = 0.0;
  s_red[warp][lane] = s;
  __syncthreads();
  if (warp == 0) {
    double m = 0.0;
#pragma unroll
    for (int w = 0; w < kDiagWarps; ++w) m += s_red[w][lane];
    s_mean[lane] = m / (double)n;
  }
  __syncthreads();
  const double mean = s_mean[lane];
  for (int i = warp; i < n; i += kDiagWarps) tile[(size_t)i * kDiagLanes + lane] -= valid ? mean : 0.0;
  __syncthreads();
  // Lags in groups of kLagGroup per warp with a sliding register window: per row one new tile element
  // and one multiplicand feed kLagGroup FMAs (the loop is fp64-pipe bound, not shared-memory bound).
  // Rows n .. n+kLagGroup-1 of the tile are zero, so products past the end of the sequence vanish.
  for (int t0 = warp * kLagGroup; t0 < n_lags; t0 += kDiagWarps * kLagGroup) {
    double acc[kLagGroup], win[kLagGroup];
#pragma unroll
    for (int k = 0; k < kLagGroup; ++k) {
      acc[k] = 0.0;
      win[k] = tile[(size_t)(t0 + k) * kDiagLanes + lane];
    }
    for (int i = 0; i < n - t0; ++i) {
      const double a = tile[(size_t)i * kDiagLanes + lane];
#pragma unroll
      for (int k = 0; k < kLagGroup; ++k) acc[k] = fma(a, win[k], acc[k]);
#pragma unroll
      for (int k = 0; k + 1 < kLagGroup; ++k) win[k] = win[k + 1];
      win[kLagGroup - 1] = tile[(size_t)(i + t0 + kLagGroup) * kDiagLanes + lane];
    }
#pragma unroll
    for (int k = 0; k < kLagGroup; ++k) {
      double a = acc[k] / (double)n;
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) a += __shfl_xor_sync(0xffffffffu, a, o);
      if (lane == 0 && t0 + k < n_lags) part_acov[(size_t)blk * n_lags + t0 + k] = a;
    }
  }
  if (threadIdx.x == 0) {
    // (count, mean, M2) of this block's sequence means, sequential Welford in lane order
    double cnt = 0.0, mu = 0.0, m2 = 0.0;
    const int64_t c0 = (int64_t)blockIdx.x * kDiagLanes;
    for (int l = 0; l < kDiagLanes && c0 + l < C; ++l) {
      cnt += 1.0;
      const double d = s_mean[l] - mu;
      mu += d / cnt;
      m2 += d * (s_mean[l] - mu);
    }
    part_stats[blk * 3 + 0] = cnt;
    part_stats[blk * 3 + 1] = mu;
    part_stats[blk * 3 + 2] = m2;
  }
}

// (count, mean, M2) merge of Chan et al.
__device__ __forceinline__ void chan_merge(double& cnt, double& mu, double& m2, double cb, double mb, double qb) {
  if (cb == 0.0) return;
  const double tot = cnt + cb, d = mb - mu;
  m2 = m2 + qb + d * d * cnt * cb / tot;
  mu = mu + d * cb / tot;
  cnt = tot;
}

// Block t < n_lags sums lag t over the nblk block partials; block n_lags merges the (count, mean, M2) triples.
// Each of the 256 threads takes a strided subset in ascending order, then a fixed shared-memory tree combines the
// 256 partial results: the order of operations depends only on nblk, so the result is deterministic — and the work
// is 256-way parallel (one thread walking 4096 partials was 1.4 ms of serial L2 latency per call, 5 calls per variable).
__global__ void __launch_bounds__(256) split_chain_final_kernel(const double* __restrict__ part_acov,
                                                                const double* __restrict__ part_stats, int64_t nblk,
                                                                int n_lags, double* __restrict__ acov_sum,
                                                                double* __restrict__ mean_stats) {
  __shared__ double s0[256], s1[256], s2[256];
  const int tid = threadIdx.x;
  const int t = blockIdx.x;
  if (t < n_lags) {
    double s = 0.0;
    for (int64_t b = tid; b < nblk; b += 256) s += part_acov[(size_t)b * n_lags + t];
    s0[tid] = s;
    __syncthreads();
    for (int o = 128; o > 0; o >>= 1) {
      if (tid < o) s0[tid] += s0[tid + o];
      __syncthreads();
    }
    if (tid == 0) acov_sum[t] = s0[0];
  } else {
    double cnt = 0.0, mu = 0.0, m2 = 0.0;
    for (int64_t b = tid; b < nblk; b += 256) chan_merge(cnt, mu, m2, part_stats[b * 3], part_stats[b * 3 + 1], part_stats[b * 3 + 2]);
    s0[tid] = cnt; s1[tid] = mu; s2[tid] = m2;
    __syncthreads();
    for (int o = 128; o > 0; o >>= 1) {
      if (tid < o) {
        double c = s0[tid], m = s1[tid], q = s2[tid];
        chan_merge(c, m, q, s0[tid + o], s1[tid + o], s2[tid + o]);
        s0[tid] = c; s1[tid] = m; s2[tid] = q;
      }
      __syncthreads();
    }
    if (tid == 0) { mean_stats[0] = s0[0]; mean_stats[1] = s1[0]; mean_stats[2] = s2[0]; }
  }
}

static const size_t kDiagSmemMax = 200 * 1024;

static size_t split_stats_ws(int64_t S_half, int64_t C, int n_lags, bool* use_smem) {
  const size_t nblk = (size_t)((C + kDiagLanes - 1) / kDiagLanes) * 2;
  const size_t tile = (size_t)(S_half + kLagGroup) * kDiagLanes * sizeof(double);
  const bool sm = tile <= kDiagSmemMax;
  if (use_smem) *use_smem = sm;
  size_t b = nblk * ((size_t)n_lags + 3) * sizeof(double) + 256;
  if (!sm) b += nblk * tile;
  return b;
}

}  // namespace mcmc

using namespace mcmc;

extern "C" {

int mcmc_rank_normalize(const double* sorted_vals, const int64_t* perm, int64_t n, double* out_z, void* stream) {
  if ((!sorted_vals || !perm || !out_z) && n > 0) { set_error("null argument"); return MCMC_EINVAL; }
  int nd = 0;
  if (cudaGetDeviceCount(&nd) != cudaSuccess || nd == 0) { cudaGetLastError(); set_error("no CUDA device (there is no CPU fallback)"); return MCMC_ECUDA; }
  if (n == 0) return MCMC_OK;
  rank_normalize_kernel<<<(unsigned)((n + 255) / 256), 256, 0, (cudaStream_t)stream>>>(sorted_vals, perm, n, out_z);
  return cuda_fail(cudaGetLastError(), "rank_normalize launch");
}

int mcmc_split_chain_stats_workspace(int64_t half_len, int64_t n_chains, int32_t n_lags, size_t* bytes) {
  if (!bytes || half_len < 1 || n_chains < 0 || n_lags < 1) { set_error("bad argument"); return MCMC_EINVAL; }
  *bytes = split_stats_ws(half_len, n_chains, n_lags, nullptr);
  return MCMC_OK;
}

int mcmc_split_chain_stats(const double* y, int64_t half_len, int64_t second_half_row, int64_t n_chains,
                           int64_t stride_s, int64_t stride_c, int32_t n_lags, double* acov_sum, double* mean_stats,
                           void* workspace, size_t workspace_bytes, void* stream) {
  if (!y || !acov_sum || !mean_stats) { set_error("null argument"); return MCMC_EINVAL; }
  if (half_len < 1 || half_len > (1 << 20) || n_lags < 1 || n_lags > half_len || n_chains < 1 || second_half_row < 0) {
    set_error("split_chain_stats: bad sizes (half_len=%lld, n_lags=%d, n_chains=%lld)", (long long)half_len, n_lags,
              (long long)n_chains);
    return MCMC_EINVAL;
  }
  int nd = 0;
  if (cudaGetDeviceCount(&nd) != cudaSuccess || nd == 0) { cudaGetLastError(); set_error("no CUDA device (there is no CPU fallback)"); return MCMC_ECUDA; }
  bool sm = true;
  const size_t need = split_stats_ws(half_len, n_chains, n_lags, &sm);
  if (!workspace || workspace_bytes < need) { set_error("split_chain_stats needs %zu workspace bytes", need); return MCMC_EWORKSPACE; }
  const int64_t gx = (n_chains + kDiagLanes - 1) / kDiagLanes;
  if (gx > 0x7fffffffLL) { set_error("too many chains"); return MCMC_ENOTSUP; }
  const size_t nblk = (size_t)gx * 2;
  double* part_acov = (double*)workspace;
  double* part_stats = part_acov + nblk * (size_t)n_lags;
  double* scratch = part_stats + nblk * 3 + 32;
  scratch = (double*)(((uintptr_t)scratch + 15) & ~(uintptr_t)15);
  const dim3 grid((unsigned)gx, 2);
  cudaStream_t s = (cudaStream_t)stream;
  if (sm) {
    const size_t tile = (size_t)(half_len + kLagGroup) * kDiagLanes * sizeof(double);
    if (int rc = cuda_fail(cudaFuncSetAttribute(split_chain_stats_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                                (int)tile), "cudaFuncSetAttribute(split_chain_stats)"))
      return rc;
    split_chain_stats_kernel<true><<<grid, kDiagLanes * kDiagWarps, tile, s>>>(y, n_chains, stride_s, stride_c, (int)half_len,
                                                                              second_half_row, n_lags, part_acov, part_stats,
                                                                              nullptr);
  } else {
    split_chain_stats_kernel<false><<<grid, kDiagLanes * kDiagWarps, 0, s>>>(y, n_chains, stride_s, stride_c, (int)half_len,
                                                                             second_half_row, n_lags, part_acov, part_stats,
                                                                             scratch);
  }
  if (int rc = cuda_fail(cudaGetLastError(), "split_chain_stats launch")) return rc;
  split_chain_final_kernel<<<n_lags + 1, 256, 0, s>>>(part_acov, part_stats, (int64_t)nblk, n_lags, acov_sum, mean_stats);
  return cuda_fail(cudaGetLastError(), "split_chain_final launch");
}

}  // extern "C"
