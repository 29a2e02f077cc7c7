// K6b: the k-medoids iterations (algo.py:184-211) on the device, one CTA, on the float32 distance matrix of K6.
// The reference runs them in numpy; token ids depend on every bit of it, so each numpy operation is restated exactly:
//   assignments = np.argmin(D[:, medoids], axis=1)          first smallest column (strict <)
//   intra = D[np.ix_(members, members)].sum(axis=1)          float32 PAIRWISE sum of numpy's add-reduce over a contiguous row
//                                                             (8 accumulators per <= 128-element leaf, halves rounded down to a
//                                                             multiple of 8 above that; pinned against numpy by the CPU test
//                                                             tests/test_pairwise_sum.py through gb_pairwise_sum_host)
//   new = members[np.argmin(intra)]                          first smallest
//   total_shift += D[medoid, new]; break if < 1e-4          float32 (a python float is a weak scalar next to np.float32)
// An empty cluster draws from the host's RNG stream (rng.integers, algo.py:199): the kernel then stops with status 1 and the
// caller repeats the call on the host from the same initial medoids -- the kernel consumes no random numbers, so the stream
// is where the reference's would be.
#include "common.cuh"
#include "geom.cuh"

#define PAM_THREADS 1024
#define PAM_MAX_N 8192
#define PAM_MAX_K 64

// numpy's pairwise sum over v[idx[0..n)]; GB_HD so that the host test can pin it against numpy itself
template <class Get>
GB_HD float gb_pw_leaf(Get get, int lo, int n) {
  if (n < 8) {
    float r = 0.f;
    for (int i = 0; i < n; ++i) r += get(lo + i);
    return r;
  }
  float r[8];
#pragma unroll
  for (int j = 0; j < 8; ++j) r[j] = get(lo + j);
  int i = 8;
  for (; i < n - (n % 8); i += 8) {
#pragma unroll
    for (int j = 0; j < 8; ++j) r[j] += get(lo + i + j);
  }
  float res = ((r[0] + r[1]) + (r[2] + r[3])) + ((r[4] + r[5]) + (r[6] + r[7]));
  for (; i < n; ++i) res += get(lo + i);
  return res;
}

template <class Get>
GB_HD float gb_pairwise_sum(Get get, int n) {
  if (n <= 128) return gb_pw_leaf(get, 0, n);
  // the recursion pw(lo, n) = pw(lo, n2) + pw(lo + n2, n - n2), n2 = (n / 2) rounded down to a multiple of 8, without a call stack
  int slo[24], sn[24];
  float sval[24];
  unsigned char sst[24];
  int sp = 0;
  slo[0] = 0; sn[0] = n; sst[0] = 0; sp = 1;
  float ret = 0.f;
  while (sp > 0) {
    const int t = sp - 1;
    if (sst[t] == 0) {
      if (sn[t] <= 128) { ret = gb_pw_leaf(get, slo[t], sn[t]); --sp; continue; }
      int n2 = sn[t] / 2; n2 -= n2 % 8;
      sst[t] = 1;
      slo[sp] = slo[t]; sn[sp] = n2; sst[sp] = 0; ++sp;
    } else if (sst[t] == 1) {
      int n2 = sn[t] / 2; n2 -= n2 % 8;
      sval[t] = ret;
      sst[t] = 2;
      slo[sp] = slo[t] + n2; sn[sp] = sn[t] - n2; sst[sp] = 0; ++sp;
    } else {
      ret = sval[t] + ret;
      --sp;
    }
  }
  return ret;
}

struct PamRowGet {
  const float* row; const u32* idx;
  GB_HD float operator()(int i) const { return row[idx[i]]; }
};
struct PamPlainGet {
  const float* v;
  GB_HD float operator()(int i) const { return v[i]; }
};

// host entry for the CPU test: numpy-order float32 sum of v[0..n)
extern "C" float gb_pairwise_sum_host(const float* v, int n) { return gb_pairwise_sum(PamPlainGet{v}, n); }

// st: [0] status (0 done, 1 empty cluster -> host), [1] iterations run, [2] float bits of the last total_shift
__global__ void __launch_bounds__(PAM_THREADS) k_pam(const float* __restrict__ D, int N, int k, const u32* __restrict__ med_init,
                                                     int max_iter, u32* __restrict__ med_out, u32* __restrict__ st) {
  extern __shared__ u32 sm[];
  u32* asg = sm;                       // [N] cluster of every sample
  u32* mem = sm + N;                   // [N] samples grouped by cluster, ascending inside each
  __shared__ u32 med[PAM_MAX_K], off[PAM_MAX_K + 1], wsum[PAM_THREADS / 32];
  __shared__ unsigned long long best[PAM_MAX_K];
  __shared__ u32 stop;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  for (int j = tid; j < k; j += PAM_THREADS) med[j] = med_init[j];
  if (tid == 0) stop = 0u;
  __syncthreads();
  int it = 0;
  float shift = 0.f;
  for (; it < max_iter; ++it) {
    for (int i = tid; i < N; i += PAM_THREADS) {
      const float* row = D + (i64)i * N;
      float bv = row[med[0]];
      u32 bj = 0;
      for (int j = 1; j < k; ++j) {
        const float v = row[med[j]];
        if (v < bv) { bv = v; bj = (u32)j; }
      }
      asg[i] = bj;
    }
    __syncthreads();
    // members of cluster j, ascending: one ordered compaction per cluster
    u32 run = 0;
    for (int j = 0; j < k; ++j) {
      if (tid == 0) off[j] = run;
      for (int base = 0; base < N; base += PAM_THREADS) {
        const int i = base + tid;
        const bool f = i < N && asg[i] == (u32)j;
        const u32 bal = __ballot_sync(0xFFFFFFFFu, f);
        if (lane == 0) wsum[warp] = __popc(bal);
        __syncthreads();
        u32 woff = 0, tot = 0;
        for (int w = 0; w < PAM_THREADS / 32; ++w) { const u32 c = wsum[w]; if (w < warp) woff += c; tot += c; }
        if (f) mem[run + woff + __popc(bal & ((1u << lane) - 1u))] = (u32)i;
        run += tot;
        __syncthreads();
      }
      if (tid == 0) { best[j] = ~0ull; if (run == off[j]) stop = 1u; }
    }
    if (tid == 0) off[k] = run;
    __syncthreads();
    if (stop) break;                     // an empty cluster: rng.integers on the host
    for (int p = tid; p < N; p += PAM_THREADS) {
      const u32 m = mem[p], j = asg[m];
      const u32 lo = off[j], c = off[j + 1] - lo;
      const float s = gb_pairwise_sum(PamRowGet{D + (i64)m * N, mem + lo}, (int)c);
      // distances are >= +0, so the bit pattern orders like the value; the position breaks ties towards the first member
      atomicMin(&best[j], ((unsigned long long)__float_as_uint(s) << 32) | (u32)p);
    }
    __syncthreads();
    if (tid == 0) {
      shift = 0.f;
      for (int j = 0; j < k; ++j) {
        const u32 nm = mem[(u32)(best[j] & 0xFFFFFFFFull)];
        shift += D[(i64)med[j] * N + nm];
        med[j] = nm;
      }
      if (shift < 1e-4f) stop = 2u;
    }
    __syncthreads();
    if (stop) { ++it; break; }
  }
  for (int j = tid; j < k; j += PAM_THREADS) med_out[j] = med[j];
  if (tid == 0) { st[0] = stop == 1u ? 1u : 0u; st[1] = (u32)it; st[2] = __float_as_uint(shift); }
}

extern "C" int gb_pam_max_n(void) { return PAM_MAX_N; }
extern "C" int gb_pam_max_k(void) { return PAM_MAX_K; }

extern "C" int gb_kmedoids_pam(const float* D, int N, int k, const uint32_t* med_init, int max_iter, uint32_t* med_out, uint32_t* st,
                               void* stream) {
  if (N <= 0 || k <= 0 || k > N) { gb_set_error("gb_kmedoids_pam: N=%d k=%d", N, k); return -2; }
  if (N > PAM_MAX_N || k > PAM_MAX_K) {
    gb_set_error("gb_kmedoids_pam: N=%d k=%d beyond the one-CTA kernel (N <= %d, k <= %d): run the iterations on the host", N, k,
                 PAM_MAX_N, PAM_MAX_K);
    return -2;
  }
  const size_t smem = sizeof(u32) * 2 * (size_t)N;
  if (smem > 48 * 1024)
    GB_CK(cudaFuncSetAttribute(k_pam, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)(sizeof(u32) * 2 * PAM_MAX_N)));
  k_pam<<<1, PAM_THREADS, smem, (cudaStream_t)stream>>>(D, N, k, med_init, max_iter, med_out, st);
  GB_LAUNCH_CK();
  return 0;
}
