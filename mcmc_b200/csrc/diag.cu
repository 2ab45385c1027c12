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
//   rank_normalize_fold_kernel the bulk z AND the z of the folded draws |x - median| from ONE sort: in the sorted array the
//                              two sides of the median are already monotone in the distance, so a folded rank is the
//                              element's position on its own side plus one binary search on the other side
//   split_chain_stats_kernel   per (chain, half) sequence: mean and biased autocovariance at every lag
//                              (direct O(n^2/2) products from a shared-memory tile, not an FFT:
//                              43 GFLOP for 2.6M sequences of 128 draws), summed over the block's 32
//                              sequences in a fixed shuffle order
//   split_chain_final_kernel   fixed-order sum of the block partials -> sum_seq acov[t]; Chan merge of
//                              the (count, mean, M2) of the sequence means  => deterministic results
#include "common.cuh"

#include <cub/device/device_radix_sort.cuh>

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

// ---- bulk and folded z-scores from one sort (ArviZ: _z_scale(split) and _z_scale(abs(split - median(split))))

__global__ void iota_u32_kernel(uint32_t* __restrict__ p, int64_t n) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) p[i] = (uint32_t)i;
}

// pivot[0] = median of the sorted values (np.median: mean of the two middle elements), pivot[1] = J = first index with
// s[j] >= median (as a double; J < 2^32)
__global__ void fold_pivot_kernel(const double* __restrict__ s, int64_t n, double* __restrict__ pivot) {
  const double med = 0.5 * (s[(n - 1) / 2] + s[n / 2]);
  int64_t a = 0, b = n;
  while (a < b) {
    const int64_t m = a + ((b - a) >> 1);
    if (s[m] < med) a = m + 1; else b = m;
  }
  pivot[0] = med;
  pivot[1] = (double)a;
}

__device__ __forceinline__ double rank_to_z(double less, double eq, double n) {
  const double r = less + 0.5 * (eq + 1.0);                  // average of the 1-based ranks less+1 .. less+eq
  return normcdfinv((r - 0.375) / (n + 0.25));
}

// Thread i owns the i-th smallest value.  d_j = |s[j] - med| is non-increasing for j < J and non-decreasing for j >= J
// (a rounded subtraction is monotone, and s[j] - med is never 0 for s[j] != med), so among the folded values
//   #(d < d_i) = (own side: distance in positions to J, past the run of equal d) + (other side: one binary search)
// and the tie counts come from the same neighbours / searches.  All searches of a warp walk the same path (d_i is
// monotone in i), so they stay in L1.  Ranks are exact half-integers: the z-scores equal those of a sort of the folded
// values bit for bit.
__global__ void __launch_bounds__(256) rank_normalize_fold_kernel(const double* __restrict__ s, const uint32_t* __restrict__ perm,
                                                                  int64_t n, const double* __restrict__ pivot,
                                                                  double* __restrict__ z_bulk, double* __restrict__ z_fold) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const double x = s[i];
  const double med = pivot[0];
  const int64_t J = (int64_t)pivot[1];
  const double dn = (double)n;
  // ---- bulk: run [lo, hi] of values equal to x
  int64_t lo = i, hi = i;
  if (i > 0 && s[i - 1] == x) {
    int64_t a = 0, b = i - 1;
    while (a < b) {
      const int64_t m = a + ((b - a) >> 1);
      if (s[m] == x) b = m; else a = m + 1;
    }
    lo = a;
  }
  if (i + 1 < n && s[i + 1] == x) {
    int64_t a = i + 1, b = n - 1;
    while (a < b) {
      const int64_t m = a + ((b - a + 1) >> 1);
      if (s[m] == x) a = m; else b = m - 1;
    }
    hi = a;
  }
  const size_t dst = perm[i];
  z_bulk[dst] = rank_to_z((double)lo, (double)(hi - lo + 1), dn);
  // ---- folded
#define FD(j) fabs(s[j] - med)
  const double d = fabs(x - med);
  int64_t less, eq;
  if (i >= J) {
    // own side [J, n): non-decreasing d.  Run [a, b] of equal d around i (wider than [lo, hi] only if distinct values
    // round to the same distance)
    int64_t a = lo > J ? lo : J, b = hi;
    if (a > J && FD(a - 1) == d) {
      int64_t u = J, v = a - 1;           // first index in [J, a-1] with d == d_i
      while (u < v) {
        const int64_t m = u + ((v - u) >> 1);
        if (FD(m) == d) v = m; else u = m + 1;
      }
      a = u;
    }
    if (b + 1 < n && FD(b + 1) == d) {
      int64_t u = b + 1, v = n - 1;       // last index in [b+1, n-1] with d == d_i
      while (u < v) {
        const int64_t m = u + ((v - u + 1) >> 1);
        if (FD(m) == d) u = m; else v = m - 1;
      }
      b = u;
    }
    // other side [0, J): non-increasing d.  t = first index with d < d_i
    int64_t u = 0, v = J;
    while (u < v) {
      const int64_t m = u + ((v - u) >> 1);
      if (FD(m) < d) v = m; else u = m + 1;
    }
    const int64_t t = u;
    int64_t t2 = t;                         // first index with d <= d_i
    if (t > 0 && FD(t - 1) == d) {
      int64_t p = 0, q = t - 1;
      while (p < q) {
        const int64_t m = p + ((q - p) >> 1);
        if (FD(m) == d) q = m; else p = m + 1;
      }
      t2 = p;
    }
    less = (a - J) + (J - t);
    eq = (b - a + 1) + (t - t2);
  } else {
    // own side [0, J): non-increasing d
    int64_t a = lo, b = hi < J - 1 ? hi : J - 1;
    if (a > 0 && FD(a - 1) == d) {
      int64_t u = 0, v = a - 1;           // first index in [0, a-1] with d == d_i
      while (u < v) {
        const int64_t m = u + ((v - u) >> 1);
        if (FD(m) == d) v = m; else u = m + 1;
      }
      a = u;
    }
    if (b + 1 < J && FD(b + 1) == d) {
      int64_t u = b + 1, v = J - 1;       // last index in [b+1, J-1] with d == d_i
      while (u < v) {
        const int64_t m = u + ((v - u + 1) >> 1);
        if (FD(m) == d) u = m; else v = m - 1;
      }
      b = u;
    }
    // other side [J, n): non-decreasing d.  t = first index with d >= d_i
    int64_t u = J, v = n;
    while (u < v) {
      const int64_t m = u + ((v - u) >> 1);
      if (FD(m) < d) u = m + 1; else v = m;
    }
    const int64_t t = u;
    int64_t t2 = t;                         // one past the last index with d == d_i
    if (t < n && FD(t) == d) {
      int64_t p = t, q = n - 1;
      while (p < q) {
        const int64_t m = p + ((q - p + 1) >> 1);
        if (FD(m) == d) p = m; else q = m - 1;
      }
      t2 = p + 1;
    }
    less = ((J - 1) - b) + (t - J);
    eq = (b - a + 1) + (t2 - t);
  }
#undef FD
  z_fold[dst] = rank_to_z((double)less, (double)eq, dn);
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

// workspace of mcmc_rank_normalize_fold: [pivot: 2 doubles][iota: n u32][perm: n u32][cub temp]
static size_t rank_fold_ws(int64_t n, size_t* cub_bytes, size_t* off_iota, size_t* off_perm, size_t* off_cub) {
  size_t tb = 0;
  cub::DeviceRadixSort::SortPairs(nullptr, tb, (const double*)nullptr, (double*)nullptr, (const uint32_t*)nullptr,
                                  (uint32_t*)nullptr, n, 0, 64, (cudaStream_t)0);
  const size_t idx = (((size_t)n * sizeof(uint32_t)) + 255) & ~(size_t)255;
  if (cub_bytes) *cub_bytes = tb;
  if (off_iota) *off_iota = 256;
  if (off_perm) *off_perm = 256 + idx;
  if (off_cub) *off_cub = 256 + 2 * idx;
  return 256 + 2 * idx + tb + 256;
}

int mcmc_rank_normalize_fold_workspace(int64_t n, size_t* bytes) {
  if (!bytes || n < 0 || n >= ((int64_t)1 << 31)) { set_error("rank_normalize_fold: n outside [0, 2^31)"); return n >= ((int64_t)1 << 31) ? MCMC_ENOTSUP : MCMC_EINVAL; }
  *bytes = rank_fold_ws(n > 0 ? n : 1, nullptr, nullptr, nullptr, nullptr);
  return MCMC_OK;
}

int mcmc_rank_normalize_fold(const double* x, int64_t n, double* sorted_out, double* z_bulk, double* z_fold, void* workspace,
                             size_t workspace_bytes, void* stream) {
  if ((!x || !sorted_out || !z_bulk || !z_fold) && n > 0) { set_error("null argument"); return MCMC_EINVAL; }
  if (n < 0) { set_error("negative size"); return MCMC_EINVAL; }
  if (n >= ((int64_t)1 << 31)) { set_error("rank_normalize_fold: more than 2^31 - 1 values (use mcmc_rank_normalize)"); return MCMC_ENOTSUP; }
  int nd = 0;
  if (cudaGetDeviceCount(&nd) != cudaSuccess || nd == 0) { cudaGetLastError(); set_error("no CUDA device (there is no CPU fallback)"); return MCMC_ECUDA; }
  if (n == 0) return MCMC_OK;
  size_t tb = 0, o_iota = 0, o_perm = 0, o_cub = 0;
  const size_t need = rank_fold_ws(n, &tb, &o_iota, &o_perm, &o_cub);
  if (!workspace || workspace_bytes < need) { set_error("rank_normalize_fold needs %zu workspace bytes", need); return MCMC_EWORKSPACE; }
  if (((uintptr_t)workspace) & 255) { set_error("rank_normalize_fold: workspace must be 256-byte aligned"); return MCMC_EINVAL; }
  char* w = (char*)workspace;
  double* pivot = (double*)w;
  uint32_t* iota = (uint32_t*)(w + o_iota);
  uint32_t* perm = (uint32_t*)(w + o_perm);
  cudaStream_t s = (cudaStream_t)stream;
  const unsigned grid = (unsigned)((n + 255) / 256);
  iota_u32_kernel<<<grid, 256, 0, s>>>(iota, n);
  if (int rc = cuda_fail(cub::DeviceRadixSort::SortPairs((void*)(w + o_cub), tb, x, sorted_out, (const uint32_t*)iota, perm, n, 0, 64, s),
                         "cub::DeviceRadixSort::SortPairs"))
    return rc;
  fold_pivot_kernel<<<1, 1, 0, s>>>(sorted_out, n, pivot);
  rank_normalize_fold_kernel<<<grid, 256, 0, s>>>(sorted_out, perm, n, pivot, z_bulk, z_fold);
  return cuda_fail(cudaGetLastError(), "rank_normalize_fold launch");
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
