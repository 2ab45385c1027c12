// rowstore.cuh — write-back of one (chains_in_block x D) fp64 tile per stored iteration.
//
// samples has the reference layout (S, C, D) (HMC.py:312 stacks state.position per
// iteration), so the rows of a block's consecutive chains form ONE contiguous
// BLOCK*D*8-byte span.  Each thread parks its row in shared memory (16-byte
// stores), then one elected thread hands the whole span to the TMA engine as a
// 1-D bulk copy (cp.async.bulk.global.shared::cta -> SASS UBLKCP), double
// buffered so the next iteration's rows never wait for the previous copy.
#pragma once
#include "common.cuh"

namespace mcmc {

__device__ __forceinline__ uint32_t smem_u32(const void* p) {
  return (uint32_t)__cvta_generic_to_shared(p);
}

__device__ __forceinline__ void bulk_store_s2g(void* gdst, const void* ssrc, uint32_t bytes) {
  asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;\n" ::"l"(gdst),
               "r"(smem_u32(ssrc)), "r"(bytes)
               : "memory");
  asm volatile("cp.async.bulk.commit_group;\n" ::: "memory");
}
// at most N bulk groups may still be READING their shared-memory source
template <int N>
__device__ __forceinline__ void bulk_wait_read() {
  asm volatile("cp.async.bulk.wait_group.read %0;\n" ::"n"(N) : "memory");
}
__device__ __forceinline__ void bulk_wait_all() {
  asm volatile("cp.async.bulk.wait_group 0;\n" ::: "memory");
}
// make generic-proxy shared-memory writes visible to the async proxy (TMA)
__device__ __forceinline__ void fence_proxy_async_smem() {
  asm volatile("fence.proxy.async.shared::cta;\n" ::: "memory");
}

// Double-buffered tile writer.  All threads of the block must call put() the same
// number of times (it contains __syncthreads()).
template <int DS, int BLOCK>
struct TileWriter {
  double* stage;   // 2 * BLOCK * n doubles of shared memory, 16-byte aligned
  int n;           // row length
  int nvalid;      // rows of this block that exist (tail block < BLOCK)
  bool bulk;       // TMA path usable (alignment / size multiple of 16)
  int phase;

  // gbase: the block's tile of stored iteration 0; iter_stride_bytes: distance between the tiles of consecutive stored
  // iterations (C*n*8) — every tile address must be 16-byte aligned for the bulk path, not only the first
  __device__ __forceinline__ void init(double* stage_, int n_, int nvalid_, const double* gbase, size_t iter_stride_bytes) {
    stage = stage_;
    n = n_;
    nvalid = nvalid_;
    phase = 0;
    const size_t bytes = (size_t)nvalid * n * sizeof(double);
    bulk = ((bytes & 15) == 0) && ((((size_t)gbase) & 15) == 0) && (((size_t)n * BLOCK * 8) % 16 == 0) &&
           ((iter_stride_bytes & 15) == 0);
  }

  // row: this thread's n doubles; gtile: global address of the block's tile for this iteration
  __device__ __forceinline__ void put(const double* row, double* gtile, bool active) {
    const int nn = DS > 0 ? DS : n;
    if (bulk) {
      double* buf = stage + (size_t)phase * BLOCK * nn;
      if (threadIdx.x == 0) bulk_wait_read<1>();  // the copy issued two puts ago has drained `buf`
      __syncthreads();
      double* dst = buf + (size_t)threadIdx.x * nn;
      if ((nn & 1) == 0) {
#pragma unroll
        for (int d = 0; d < nn; d += 2) *reinterpret_cast<double2*>(dst + d) = make_double2(row[d], row[d + 1]);
      } else {
#pragma unroll
        for (int d = 0; d < nn; ++d) dst[d] = row[d];
      }
      fence_proxy_async_smem();
      __syncthreads();
      if (threadIdx.x == 0) bulk_store_s2g(gtile, buf, (uint32_t)((size_t)nvalid * nn * sizeof(double)));
      phase ^= 1;
    } else if (active) {
      double* dst = gtile + (size_t)threadIdx.x * nn;
#pragma unroll
      for (int d = 0; d < nn; ++d) dst[d] = row[d];
    }
  }

  __device__ __forceinline__ void finish() {
    if (bulk && threadIdx.x == 0) bulk_wait_all();
  }
};

// plain per-thread row store (tracking arrays, final state): 16-byte stores when aligned
template <int DS>
__device__ __forceinline__ void store_row(double* dst, const double* row, int n_rt) {
  const int nn = DS > 0 ? DS : n_rt;
  if ((nn & 1) == 0 && ((((size_t)dst) & 15) == 0)) {
#pragma unroll
    for (int d = 0; d < nn; d += 2) *reinterpret_cast<double2*>(dst + d) = make_double2(row[d], row[d + 1]);
  } else {
#pragma unroll
    for (int d = 0; d < nn; ++d) dst[d] = row[d];
  }
}

// per-thread row load (shared or global), 16-byte loads when aligned
template <int DS>
__device__ __forceinline__ void load_row(double* row, const double* src, int n_rt) {
  const int nn = DS > 0 ? DS : n_rt;
  if ((nn & 1) == 0 && ((((size_t)src) & 15) == 0)) {
#pragma unroll
    for (int d = 0; d < nn; d += 2) {
      const double2 v = *reinterpret_cast<const double2*>(src + d);
      row[d] = v.x;
      row[d + 1] = v.y;
    }
  } else {
#pragma unroll
    for (int d = 0; d < nn; ++d) row[d] = src[d];
  }
}

}  // namespace mcmc
