// K7: device-resident merge replay of induce (BPE.tokenize, bpe.py:1116-1139 with step_helper 1316-1425): every vocabulary
// merge of a medoid-mode vocabulary applied, in learned order, to ALL held-out backbones at once without a host round trip.
// Per merge four launches on the lazily compacted token arrays (ls = len | extra << 16, 0 = dead; see k_hloop2.cu):
//   flag     one streaming pass over the class array (4 B / entry; dead entries carry the class ~0, so both classes of a pair
//            come from the stream; glue, ls only where both match): pairs with the merge's packed key -> occurrence list
//   assign   per occurrence: NeRF coordinates of the window from the pristine internal coordinates (rmsd_super_res,
//            tokenizer.py:347-363, nerf.py:84-210) and Kabsch RMSD to the key's medoids (algo.py:8-65); first minimum wins
//            (np.argmin, bpe.py:655-657); skipped for keys with a single medoid (the argmin of one distance is 0)
//   apply    overlap rule (bpe.py:1909-1914: the head of every run merges the 1st, 3rd, ... occurrence) + in-place merge: the
//            left entry takes the assigned medoid's class / token id, the right entry dies; merge events appended
//   unflag   marks removed, per-merge event range recorded
// The host passes the merges' parameters as plain arrays; nothing is read back until the whole vocabulary has been replayed.
#include "common.cuh"
#include "geom.cuh"

#define KI_AMAX 97               // atoms per window handled on the device (32 residues + 1)

enum { KI_NOCC0 = 0, KI_NOCC1, KI_EVN, KI_ERR, KI_WORDS = 8 };

struct KITok { u32* cls; u32* glue; const u32* pos; u32* ls; u32* tid; u32* cw; };

// stream word of an entry: its class (18 bits) and a 14-bit hash of its glue bins -- the flag pass filters on (left class, glue)
// and on the right class without touching any other array; false positives (1 in 16 k class-level candidates) fall to the exact check
__device__ __forceinline__ u32 ki_gh(u32 g27) { return (g27 * 0x9E3779B1u) >> 18; }
__device__ __forceinline__ u32 ki_word(u32 cls, u32 g) { return (cls & 0x3FFFFu) | (ki_gh((u32)gb_key_m(0u, g, 0u)) << 18); }

// a class-level candidate (both classes right, or the right one still unknown): the glue angles decide for most of them -- a frequent
// class pair comes with many glue bins -- and glue[j] needs no other load first, so it is read before ls (whose extra count locates the
// right neighbour).  ncu, merge 100 of the bench vocabulary: 60 M load sectors / 604 MB of DRAM reads per launch when ls came first.
__device__ __forceinline__ void ki_flag_one(const KITok& t, u32 j, u32 cL, u64 key, u32* st, int par, u32* occ_j, u32* occ_r, u8* flags,
                                            u8* rmark) {
  const u32 g = t.glue[j];
  if ((g & GB_LAST_FLAG) || (u32)(gb_key_m(0u, g, 0u)) != (u32)(key & 0x7FFFFFFull)) return;
  const u32 l = t.ls[j];
  if (l == 0u) return;
  const u32 r = j + 1u + (l >> 16);
  if (gb_key_m(cL, g, t.cls[r]) != key) return;
  const u32 q = atomicAdd(&st[KI_NOCC0 + par], 1u);
  occ_j[q] = j; occ_r[q] = r;
  flags[j] = 1; rmark[r] = 1;
}

// Flag pass.  Dead entries carry the class ~0 (ki_apply), so the class stream alone tells a thread both classes of a pair: the
// right class of entry j is the next entry of the stream that is not dead.  Each thread takes sixteen consecutive classes
// (four 16-byte loads in flight), walks them backwards and keeps only entries whose (left, right) classes are the merge's; ls /
// glue are read for those alone (and for the few whose right neighbour lies beyond the thread's sixteen).  Reading ls / glue
// wherever the LEFT class matched cost 584 MB of DRAM reads per merge against the 205 MB class array (ncu --set full, merge
// 1200 of the bench vocabulary) because a frequent left class touches most 32-byte sectors of both arrays.
#define KI_DEAD 0xFFFFFFFFu
#define KI_UNKNOWN 0xFFFFFFFEu
__global__ void __launch_bounds__(256) ki_flag(KITok t, i64 N, u64 key, u32* __restrict__ st, int par, u32* __restrict__ occ_j,
                                               u32* __restrict__ occ_r, u8* __restrict__ flags, u8* __restrict__ rmark) {
  const u32 cL = (u32)(key >> 45) & 0x3FFFFu, cR = (u32)(key >> 27) & 0x3FFFFu;
  const uint4* c4 = reinterpret_cast<const uint4*>(t.cw);
  const u32 W = cL | (ki_gh((u32)(key & 0x7FFFFFFull)) << 18);
  const i64 N16 = N >> 4, stride = (i64)gridDim.x * blockDim.x;
  for (i64 q = (i64)blockIdx.x * blockDim.x + threadIdx.x; q < N16; q += stride) {
    u32 c[16];
#pragma unroll
    for (int u = 0; u < 4; ++u) {
      const uint4 v = __ldg(c4 + 4 * q + u);
      c[4 * u] = v.x; c[4 * u + 1] = v.y; c[4 * u + 2] = v.z; c[4 * u + 3] = v.w;
    }
    u32 nxt = KI_UNKNOWN, m = 0u;
#pragma unroll
    for (int e = 15; e >= 0; --e) {
      if (c[e] == W && (nxt == cR || nxt == KI_UNKNOWN)) m |= 1u << e;
      if (c[e] != KI_DEAD) nxt = c[e] & 0x3FFFFu;
    }
    while (m) {
      const int e = __ffs(m) - 1;
      m &= m - 1u;
      ki_flag_one(t, (u32)(q << 4) + (u32)e, cL, key, st, par, occ_j, occ_r, flags, rmark);
    }
  }
  if (blockIdx.x == 0 && threadIdx.x < (u32)(N & 15)) {
    const u32 j = (u32)(N16 << 4) + threadIdx.x;
    if (t.cw[j] == W) ki_flag_one(t, j, cL, key, st, par, occ_j, occ_r, flags, rmark);
  }
}

struct KIGeo { const double* V; i64 ld; const u32* bb; const i64* offsets; const double* f3tab; };

// NeRF coordinates of the window of occurrence j into c (same order of operations as k_window_coords)
__device__ __forceinline__ void ki_window(const KITok& t, const KIGeo& g, u32 j, int A, double* c) {
  const i64 r0 = t.pos[j];
  const i64 ld = g.ld;
  const double* V = g.V;
  double nca, cac, tau;
  if (r0 == g.offsets[g.bb[r0]]) {             // the window starts its backbone: first-residue geometry of the token's class
    const double* f = g.f3tab + 3 * (i64)t.cls[j];
    nca = f[0]; cac = f[1]; tau = f[2];
  } else {
    nca = V[1 * ld + r0 - 1]; cac = V[2 * ld + r0 - 1]; tau = V[6 * ld + r0 - 1];
  }
  gb_v3 p0, p1, p2;
  gb_first_residue(nca, cac, tau, p0, p1, p2);
  int na = 0;
  c[0] = p0.x; c[1] = p0.y; c[2] = p0.z; ++na;
  if (na < A) { c[3] = p1.x; c[4] = p1.y; c[5] = p1.z; ++na; }
  if (na < A) { c[6] = p2.x; c[7] = p2.y; c[8] = p2.z; ++na; }
  for (int k = 0; na < A; ++k) {
    const i64 r = r0 + k;
    gb_v3 n = gb_place_dihedral(p0, p1, p2, V[7 * ld + r], V[0 * ld + r], V[4 * ld + r]);
    c[3 * na] = n.x; c[3 * na + 1] = n.y; c[3 * na + 2] = n.z; ++na;
    if (na >= A) break;
    gb_v3 ca = gb_place_dihedral(p1, p2, n, V[8 * ld + r], V[1 * ld + r], V[5 * ld + r]);
    c[3 * na] = ca.x; c[3 * na + 1] = ca.y; c[3 * na + 2] = ca.z; ++na;
    if (na >= A) break;
    gb_v3 cc = gb_place_dihedral(p2, n, ca, V[6 * ld + r], V[2 * ld + r], V[3 * ld + r + 1]);
    c[3 * na] = cc.x; c[3 * na + 1] = cc.y; c[3 * na + 2] = cc.z; ++na;
    p0 = n; p1 = ca; p2 = cc;
  }
}

// Two shapes of the same arithmetic.  Many occurrences (early merges, short windows): one thread per occurrence, medoids in turn.
// Few occurrences (late merges, long windows; the latency of one thread walking K Kabsch fits is what a merge then costs): one warp
// per occurrence -- lane 0 places the atoms into shared memory, the lanes take the medoids, a (distance, index) minimum picks the
// first smallest like np.argmin.  Each (occurrence, medoid) distance comes from the same gb_kabsch_rmsd call either way.
#define KI_WARP_BELOW (148 * 8 * 4 * 4)   // up to four occurrences per resident warp
__global__ void __launch_bounds__(128) ki_assign(KITok t, KIGeo g, const u32* __restrict__ st, int par, const u32* __restrict__ occ_j,
                                                 int A, const double* __restrict__ med, int K, u32* __restrict__ aux) {
  const u32 nocc = st[KI_NOCC0 + par];
  if (nocc >= (u32)KI_WARP_BELOW) {
    for (u32 i = blockIdx.x * blockDim.x + threadIdx.x; i < nocc; i += gridDim.x * blockDim.x) {
      const u32 j = occ_j[i];
      double c[3 * KI_AMAX];
      ki_window(t, g, j, A, c);
      double bv = INFINITY;
      u32 bi = 0;
      for (int m = 0; m < K; ++m) {
        const double d = gb_kabsch_rmsd(c, med + (i64)m * A * 3, A);
        if (d < bv) { bv = d; bi = (u32)m; }     // strict <: first index on ties (np.argmin)
      }
      aux[j] = bi;
    }
    return;
  }
  __shared__ double sc[4][3 * KI_AMAX];
  const int w = threadIdx.x >> 5, lane = threadIdx.x & 31;
  double* c = sc[w];
  for (u32 i = blockIdx.x * 4 + w; i < nocc; i += gridDim.x * 4) {
    const u32 j = occ_j[i];
    if (lane == 0) ki_window(t, g, j, A, c);
    __syncwarp();
    double bv = INFINITY;
    u32 bi = 0xFFFFFFFFu;
    for (int m = lane; m < K; m += 32) {
      const double d = gb_kabsch_rmsd(c, med + (i64)m * A * 3, A);
      if (d < bv) { bv = d; bi = (u32)m; }
    }
    for (int o = 16; o; o >>= 1) {
      const double ov = __shfl_xor_sync(0xFFFFFFFFu, bv, o);
      const u32 oi = __shfl_xor_sync(0xFFFFFFFFu, bi, o);
      if (ov < bv || (ov == bv && oi < bi)) { bv = ov; bi = oi; }
    }
    if (lane == 0) aux[j] = bi == 0xFFFFFFFFu ? 0u : bi;
    __syncwarp();
  }
}

__global__ void __launch_bounds__(128) ki_apply(KITok t, u32* __restrict__ st, int par, const u32* __restrict__ occ_j,
                                                const u8* __restrict__ flags, const u8* __restrict__ rmark, const u32* __restrict__ aux,
                                                const u32* __restrict__ p2cls, const u32* __restrict__ p2tid, u32* __restrict__ ev_left,
                                                u32* __restrict__ ev_right, u32* __restrict__ ev_p, u32 evcap, int single) {
  const u32 nocc = st[KI_NOCC0 + par];
  for (u32 i = blockIdx.x * blockDim.x + threadIdx.x; i < nocc; i += gridDim.x * blockDim.x) {
    u32 j = occ_j[i];
    if (rmark[j]) continue;                      // not a run head: the pair to its left is an occurrence too
    bool keep = true;
    u32 l = t.ls[j];                             // carried along the run: a merged right token's ls is zeroed below
    while (flags[j]) {
      const u32 r = j + 1u + (l >> 16), lr = t.ls[r];
      if (keep) {
        const u32 p = single ? 0u : aux[j];        // one medoid: the argmin is 0 whatever the distances (no assignment pass ran)
        const u32 q = atomicAdd(&st[KI_EVN], 1u);
        if (q < evcap) { ev_left[q] = t.pos[j]; ev_right[q] = t.pos[r]; ev_p[q] = p; } else st[KI_ERR] = 1u;
        t.cls[j] = p2cls[p];
        t.cw[j] = ki_word(p2cls[p], t.glue[r]);
        t.cw[r] = KI_DEAD;
        t.tid[j] = p2tid[p];
        t.glue[j] = t.glue[r];
        t.ls[j] = ((l & 0xFFFFu) + (lr & 0xFFFFu)) | (((l >> 16) + 1u + (lr >> 16)) << 16);
        t.ls[r] = 0u;
      }
      keep = !keep;
      j = r; l = lr;
    }
  }
}

__global__ void __launch_bounds__(256) ki_unflag(u32* __restrict__ st, int par, const u32* __restrict__ occ_j, const u32* __restrict__ occ_r,
                                                 u8* __restrict__ flags, u8* __restrict__ rmark, u32* __restrict__ rep_start,
                                                 u32* __restrict__ rep_cnt, int m, u32 ev_before_slot) {
  const u32 nocc = st[KI_NOCC0 + par];
  for (u32 i = blockIdx.x * blockDim.x + threadIdx.x; i < nocc; i += gridDim.x * blockDim.x) {
    flags[occ_j[i]] = 0; rmark[occ_r[i]] = 0;
  }
  if (blockIdx.x == 0 && threadIdx.x == 0) {
    // rep_start[m] was stored by the previous merge's unflag (= EVN after it); this merge's count closes the range
    const u32 start = rep_start[m];
    rep_cnt[m] = st[KI_EVN] - start;
    rep_start[m + 1] = st[KI_EVN];
    st[KI_NOCC0 + (par ^ 1)] = 0u;               // the other parity's counter: the next merge's flag pass starts from zero
    (void)ev_before_slot;
  }
}

// merges with host-side parameters (the caller leaves out merges whose key cannot occur); rep_start[0] = first free event slot
__global__ void __launch_bounds__(256) ki_words(KITok t, i64 N, i64 cap) {
  for (i64 j = (i64)blockIdx.x * blockDim.x + threadIdx.x; j < cap; j += (i64)gridDim.x * blockDim.x)
    t.cw[j] = j < N && t.ls[j] != 0u ? ki_word(t.cls[j], t.glue[j]) : KI_DEAD;
}

extern "C" int gb_induce_replay(uint32_t* cls, uint32_t* glue, const uint32_t* pos, uint32_t* ls, uint32_t* tid, uint8_t* flags,
                                uint8_t* rmark, uint32_t* aux, uint32_t* cw, int64_t N, const double* vals, int64_t ld, const uint32_t* bb,
                                const int64_t* offsets, const double* f3tab, int n_merges, const uint64_t* mkey_host,
                                const int* mK_host, const int* mA_host, const int* mnres_host, const int64_t* mmed_off_host,
                                const int64_t* mp2_off_host, const double* med, const uint32_t* p2cls, const uint32_t* p2tid,
                                uint32_t* occ_j, uint32_t* occ_r, uint32_t* ev_left, uint32_t* ev_right, uint32_t* ev_p,
                                int64_t evcap, uint32_t* st, uint32_t* rep_start, uint32_t* rep_cnt, void* stream) {
  cudaStream_t s = (cudaStream_t)stream;
  (void)mnres_host;
  if (N <= 0 || n_merges <= 0) return 0;
  int sms = gb_sm_count();
  if (sms <= 0) sms = 148;
  KITok t{cls, glue, pos, ls, tid, cw};
  ki_words<<<gb_blocks_for(N, 256 * 8, sms * 8), 256, 0, s>>>(t, N, N);
  KIGeo g{vals, ld, bb, offsets, f3tab};
  const int flag_blocks = gb_blocks_for(N, 256 * 16, sms * 8);
  int par = 0;
  for (int m = 0; m < n_merges; ++m) {
    if (mA_host[m] > KI_AMAX || mK_host[m] <= 0) { gb_set_error("gb_induce_replay: merge %d has %d medoids / spans %d atoms (device limit %d)", m, mK_host[m], mA_host[m], KI_AMAX); return -2; }
    ki_flag<<<flag_blocks, 256, 0, s>>>(t, N, (u64)mkey_host[m], st, par, occ_j, occ_r, flags, rmark);
    const int single = mK_host[m] == 1;
    if (!single) ki_assign<<<sms * 8, 128, 0, s>>>(t, g, st, par, occ_j, mA_host[m], med + mmed_off_host[m], mK_host[m], aux);
    ki_apply<<<sms, 128, 0, s>>>(t, st, par, occ_j, flags, rmark, aux, p2cls + mp2_off_host[m], p2tid + mp2_off_host[m], ev_left,
                                 ev_right, ev_p, (u32)evcap, single);
    ki_unflag<<<sms, 256, 0, s>>>(st, par, occ_j, occ_r, flags, rmark, rep_start, rep_cnt, m, 0u);
    par ^= 1;
  }
  GB_LAUNCH_CK();
  return 0;
}
extern "C" int gb_induce_max_atoms(void) { return KI_AMAX; }
