// Device-driven histogram-mode learn loop: BPE.step (bpe.py:1792-2166 with rmsd_key None) for N iterations without a
// host round trip.  Everything the host used to decide per iteration lives in a device state block `hs`
// (k1_sketch.cuh): the live-token count, the popped key, the reference's tie-break (python string order of
// json.dumps(key_dict, sort_keys=True), bpe.py:1149,1469 -- evaluated here by materialising the tied keys' JSON
// strings), the overlap rule, the merge-event log and a per-step record the host reads back in bulk.
// Any condition the device cannot resolve (more ties than HL_TB_CAP, more candidate slots than the cap, a content-hash
// collision, a full log) sets hs[HS_HALT]; every later kernel then returns immediately, the host handles that one
// iteration on its own path and re-enters the loop.
#include "k1_sketch.cuh"

#define HL_TB_CAP 128           // ties resolved on the device
#define HL_TIE_LIST 256         // tie list entries kept (>= HL_TB_CAP)
#define HL_BLK_LIST 512         // candidate pairs staged per CTA before one append to the global list

enum { HALT_TIES = 1, HALT_CAND = 2, HALT_COLLISION = 3, HALT_TABLE = 4, HALT_EVLOG = 5, HALT_EMPTY = 6, HALT_STEPS = 7,
       HALT_COUNT = 8, HALT_PEER = 9 };

__device__ __forceinline__ void hl_halt(u32* hs, u32 code) {
  if (atomicCAS(&hs[HS_HALT], 0u, code) == 0u) hs[HS_HALT_STEP] = hs[HS_STEP];
}

// ---- slot scan -----------------------------------------------------------------------------------------------------
// one pass over the slot counters: per-chunk maxima (bmax), global max c1, non-zero count, and the round-1 candidate
// list = slots with count >= thr1 = 0.9 * (previous iteration's max count).  The max count never grows from one
// iteration to the next, so the exact max is almost always >= thr1 and one candidate pass is conclusive; otherwise
// hl_decide asks for a second round (thr = c1 when nothing / too much was collected, else the best exact count).
__device__ __forceinline__ u32 hl_thr1(const u32* hs) {
  const u32 mp = hs[HS_MPREV];
  return mp ? max(1u, (u32)(((u64)mp * 9ull) / 10ull)) : 0xFFFFFFFFu;
}
__global__ void __launch_bounds__(256) hl_slots_scan(const uint4* __restrict__ slots4, u32* __restrict__ bmax, u32* __restrict__ hs,
                                                      u32* __restrict__ cand, u32 cap) {
  if (hs[HS_HALT]) return;
  __shared__ u32 s_mx[8], s_nz[8];
  const u32 thr = hl_thr1(hs);
  if (blockIdx.x == 0 && threadIdx.x == 0) hs[HS_THR] = thr;
  const i64 base4 = (i64)blockIdx.x * (KS_CHUNK / 4);
  uint4 v[4];
#pragma unroll
  for (int k = 0; k < 4; ++k) v[k] = slots4[base4 + k * 256 + threadIdx.x];
  u32 mx = 0, nz = 0;
#pragma unroll
  for (int k = 0; k < 4; ++k) {
    u32 a[4] = {v[k].x, v[k].y, v[k].z, v[k].w};
    mx = max(max(mx, a[0]), max(max(a[1], a[2]), a[3]));
    nz += (a[0] != 0) + (a[1] != 0) + (a[2] != 0) + (a[3] != 0);
    if (max(max(a[0], a[1]), max(a[2], a[3])) >= thr) {
#pragma unroll
      for (int e = 0; e < 4; ++e)
        if (a[e] >= thr) {
          u32 q = atomicAdd(&hs[HS_NCAND], 1u);
          if (q < cap) cand[q] = (u32)(4 * (base4 + k * 256 + threadIdx.x) + e);
        }
    }
  }
  for (int o = 16; o; o >>= 1) {
    mx = max(mx, __shfl_xor_sync(0xffffffffu, mx, o));
    nz += __shfl_xor_sync(0xffffffffu, nz, o);
  }
  if ((threadIdx.x & 31) == 0) { s_mx[threadIdx.x >> 5] = mx; s_nz[threadIdx.x >> 5] = nz; }
  __syncthreads();
  if (threadIdx.x == 0) {
    for (int w = 1; w < 8; ++w) { mx = max(mx, s_mx[w]); nz += s_nz[w]; }
    bmax[blockIdx.x] = mx;
    if (mx) atomicMax(&hs[HS_C1], mx);
    if (nz) atomicAdd(&hs[HS_NZ], nz);
  }
}
__global__ void __launch_bounds__(256) hl_slots_collect(const uint4* __restrict__ slots4, const u32* __restrict__ bmax,
                                                         u32* __restrict__ hs, int round, u32* __restrict__ cand, u32 cap) {
  if (hs[HS_HALT] || !hs[HS_ROUND2]) return;       // second round only (round 1 is collected by hl_slots_scan)
  const u32 thr = hs[HS_THR];
  (void)round;
  if (thr == 0 || bmax[blockIdx.x] < thr) return;
  const i64 base4 = (i64)blockIdx.x * (KS_CHUNK / 4);
#pragma unroll
  for (int k = 0; k < 4; ++k) {
    i64 i = base4 + k * 256 + threadIdx.x;
    uint4 v = slots4[i];
    u32 a[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
    for (int e = 0; e < 4; ++e)
      if (a[e] >= thr) {
        u32 q = atomicAdd(&hs[HS_NCAND], 1u);
        if (q < cap) cand[q] = (u32)(4 * i + e);
      }
  }
}
// ---- exact argmax over the small table + conclusiveness test (one block) -----------------------------------------
__global__ void __launch_bounds__(1024) hl_decide(u64* __restrict__ tk, u32* __restrict__ tc, u32* __restrict__ tm,
                                                   i64 M, u32* __restrict__ hs, int round, u64* __restrict__ tie_keys,
                                                   u32* __restrict__ tie_min, u32 cap, int allow_r2) {
  if (hs[HS_HALT] || (round == 2 && !hs[HS_ROUND2])) return;
  __shared__ u32 s_mx[32], s_nd[32], s_nt;
  u32 mx = 0, nd = 0;
  for (i64 i = threadIdx.x; i < M; i += blockDim.x) { u32 c = tc[i]; mx = max(mx, c); nd += c != 0; }
  for (int o = 16; o; o >>= 1) { mx = max(mx, __shfl_xor_sync(0xffffffffu, mx, o)); nd += __shfl_xor_sync(0xffffffffu, nd, o); }
  if ((threadIdx.x & 31) == 0) { s_mx[threadIdx.x >> 5] = mx; s_nd[threadIdx.x >> 5] = nd; }
  if (threadIdx.x == 0) s_nt = 0;
  __syncthreads();
  mx = 0; nd = 0;
  for (int w = 0; w < 32; ++w) { mx = max(mx, s_mx[w]); nd += s_nd[w]; }
  if (mx) {
    for (i64 i = threadIdx.x; i < M; i += blockDim.x)
      if (tc[i] == mx) {
        u32 k = atomicAdd(&s_nt, 1u);
        if (k < HL_TIE_LIST) { tie_keys[k] = tk[i]; tie_min[k] = tm[i]; }
      }
  }
  __syncthreads();
  if (threadIdx.x == 0) {
    hs[HS_M1] = mx; hs[HS_NT] = s_nt; hs[HS_ND] = nd;
    if (hs[HS_C1] == 0) { hl_halt(hs, HALT_EMPTY); return; }
    if (hs[HS_OVF] == 1 || (u64)nd * 2 > (u64)M) { hl_halt(hs, HALT_TABLE); return; }
    const bool cand_over = hs[HS_NCAND] > cap || hs[HS_OVF] == 2;
    if (round == 1) {
      if (cand_over || hs[HS_NCAND] == 0) {   // too many / no slots above 0.9 * previous max: retry with the top slots only
        hs[HS_ROUND2] = 1; hs[HS_THR] = hs[HS_C1]; hs[HS_NCAND] = 0; hs[HS_OVF] = 0; hs[HS_NCL] = 0; hs[HS_R2COUNT] += 1;
      } else if (mx < hs[HS_THR]) {           // inconclusive: a key outside the candidates could reach mx
        hs[HS_ROUND2] = 1; hs[HS_THR] = max(mx, 1u); hs[HS_NCAND] = 0; hs[HS_NCL] = 0; hs[HS_R2COUNT] += 1;
      }
    } else {
      if (cand_over || mx < hs[HS_THR]) hl_halt(hs, HALT_CAND);
    }
  }
  // sharded run: a second round would need another exchange per iteration for a ~1-in-1000 event: hand it to the host
  // (one GPU, default: the same -- the second-round kernels would otherwise be launched as no-ops in every iteration)
  if (threadIdx.x == 0 && round == 1 && (hs[HS_DIST] || !allow_r2) && hs[HS_ROUND2] && !hs[HS_HALT]) hl_halt(hs, HALT_CAND);
  // the table is consumed (ties are copied out): leave it cleared for the next candidate pass
  for (i64 i = threadIdx.x; i < M; i += blockDim.x) { tk[i] = GB_EMPTY_KEY; tc[i] = 0u; tm[i] = 0xFFFFFFFFu; }
}

#include "k_hl_json.cuh"

__global__ void __launch_bounds__(HL_TB_CAP) hl_tiebreak(u32* __restrict__ hs, u64* __restrict__ hs64, const u64* __restrict__ tie_keys,
                                                         const u32* __restrict__ tie_min, const u32* __restrict__ pos,
                                                         const u32* __restrict__ len, const u16* __restrict__ rbins, i64 R,
                                                         char* __restrict__ strbuf, int stride) {
  if (hs[HS_HALT]) return;
  __shared__ int s_best[HL_TB_CAP], s_len[HL_TB_CAP];
  const u32 nt = hs[HS_NT];
  if (nt == 0) { if (threadIdx.x == 0) hl_halt(hs, HALT_EMPTY); return; }
  if (nt > HL_TB_CAP) { if (threadIdx.x == 0) { hs[HS_NTMAX] = max(hs[HS_NTMAX], nt); hl_halt(hs, HALT_TIES); } return; }
  const u32 i = threadIdx.x;
  if (nt > 1) {
    if (i < nt) {
      u32 j = tie_min[i];
      u32 nres = hl_pair_res(len, j);
      int need = 39 * (int)nres + 200;
      if (need > stride) { hl_halt(hs, HALT_TIES); s_len[i] = 0; }
      else s_len[i] = hl_json_key(strbuf + (size_t)i * stride, rbins, R, pos[j], nres);
    }
    s_best[i] = (int)i;
    __syncthreads();
    if (hs[HS_HALT]) return;
    const unsigned warp = threadIdx.x >> 5, nw = blockDim.x >> 5;
    for (u32 s = 1; s < nt; s <<= 1) {
      for (u32 a = warp * 2 * s; a + s < nt; a += nw * 2 * s) {
        int ia = s_best[a], ib = s_best[a + s];
        bool b_less = hl_less(strbuf + (size_t)ib * stride, strbuf + (size_t)ia * stride, s_len[ib], s_len[ia]);
        if ((threadIdx.x & 31) == 0 && b_less) s_best[a] = ib;
      }
      __syncthreads();
    }
  }
  if (threadIdx.x == 0) {
    int w = nt > 1 ? s_best[0] : 0;
    u32 j = tie_min[w];
    hs64[0] = tie_keys[w];
    hs[HS_WIN_J] = j; hs[HS_WIN_R0] = pos[j]; hs[HS_WIN_NRES] = hl_pair_res(len, j); hs[HS_WIN_CNT] = hs[HS_M1];
    hs[HS_NTMAX] = max(hs[HS_NTMAX], nt);
  }
}


#include "k_hloop_dist.cuh"

// ---- pass 2 on the lazy layout: exact counts of the candidate slots' keys + the list of those pairs ---------------
template <bool CS>
__global__ void __launch_bounds__(KS_THREADS, 2) hl_cand_count(KeysHLT<CS> kf, int shift, const u32* __restrict__ cand, u32 cap,
                                                               u64* __restrict__ tk, u32* __restrict__ tc, u32* __restrict__ tm, u64 mask,
                                                               u32* __restrict__ hs, int round, u64* __restrict__ cl_key, u32* __restrict__ cl_j) {
  if (hs[HS_HALT] || (round == 2 && !hs[HS_ROUND2])) return;
  __shared__ u32 s_set[KC_SET];
  __shared__ u32 s_lin[KC_LINEAR];
  __shared__ u64 s_key[HL_BLK_LIST];
  __shared__ u32 s_j[HL_BLK_LIST];
  __shared__ u32 s_n, s_base;
  if (threadIdx.x == 0) s_n = 0;
  const i64 N = (i64)hs[HS_N];
  const u32 n = hs[HS_NCAND], clcap = hs[HS_CLCAP];
  if (n == 0) return;
  if (n > cap) { if (blockIdx.x == 0 && threadIdx.x == 0) hs[HS_OVF] = 2; return; }
  const bool linear = n <= KC_LINEAR;
  if (linear) {
    if (threadIdx.x < KC_LINEAR) s_lin[threadIdx.x] = threadIdx.x < n ? cand[threadIdx.x] : KS_EMPTY;
  } else {
    for (int i = threadIdx.x; i < KC_SET; i += KS_THREADS) s_set[i] = KS_EMPTY;
    __syncthreads();
    for (u32 i = threadIdx.x; i < n; i += KS_THREADS) {
      u32 v = cand[i], s = (v * 0x9E3779B1u) >> 19;
      while (true) {
        u32 prev = atomicCAS(&s_set[s], KS_EMPTY, v);
        if (prev == KS_EMPTY || prev == v) break;
        s = (s + 1) & (KC_SET - 1);
      }
    }
  }
  __syncthreads();
  for (i64 base = (i64)blockIdx.x * (2 * KS_TILE); base < N; base += (i64)gridDim.x * (2 * KS_TILE)) {
    const i64 ja = base + (i64)threadIdx.x * K1_VEC, jb = ja + KS_TILE;
    typename KeysHLT<CS>::Raw ra, rb;
    kf.ld(ja, N, ra);
    kf.ld(jb, N, rb);
#pragma unroll
    for (int half = 0; half < 2; ++half) {
      u64 key[K1_VEC];
      const i64 j0 = half ? jb : ja;
      if (half == 0) kf.fin(ja, N, ra, key); else kf.fin(jb, N, rb, key);
#pragma unroll
      for (int e = 0; e < K1_VEC; ++e) {
        if (key[e] == GB_EMPTY_KEY) continue;
        u32 slot = gb_slot_of(key[e], shift);
        bool hit = false;
        if (linear) {
#pragma unroll
          for (int i = 0; i < KC_LINEAR; ++i) hit |= (s_lin[i] == slot);
        } else {
          u32 s = (slot * 0x9E3779B1u) >> 19;
          while (true) {
            u32 v = s_set[s];
            if (v == slot) { hit = true; break; }
            if (v == KS_EMPTY) break;
            s = (s + 1) & (KC_SET - 1);
          }
        }
        if (!hit) continue;
        if (!gb_gtable_add(tk, tc, tm, mask, key[e], 1u, (u32)(j0 + e))) hs[HS_OVF] = 1;
        u32 q = atomicAdd(&s_n, 1u);                 // block-local list first: one global atomic per CTA at the end
        if (q < HL_BLK_LIST) { s_key[q] = key[e]; s_j[q] = (u32)(j0 + e); }
        else {
          u32 g = atomicAdd(&hs[HS_NCL], 1u);
          if (g < clcap) { cl_key[g] = key[e]; cl_j[g] = (u32)(j0 + e); }
        }
      }
    }
  }
  __syncthreads();
  const u32 m = min(s_n, (u32)HL_BLK_LIST);
  if (m) {
    if (threadIdx.x == 0) s_base = atomicAdd(&hs[HS_NCL], m);
    __syncthreads();
    for (u32 i = threadIdx.x; i < m; i += KS_THREADS) {
      u32 g = s_base + i;
      if (g < clcap) { cl_key[g] = s_key[i]; cl_j[g] = s_j[i]; }
    }
  }
}

// ---- K2 in place, driven by the candidate-pair list ---------------------------------------------------------------
// flags[j] = 1: pair (j, right(j)) is an occurrence of the winner; rmark[r] = 1: entry r is the right token of one.
// cl_r keeps right(j) so the marks can be removed after the merges changed ls.
__global__ void __launch_bounds__(256) hl_flag_list(u32* __restrict__ hs, const u64* __restrict__ hs64, const u64* __restrict__ cl_key,
                                                     const u32* __restrict__ cl_j, u32* __restrict__ cl_r, u8* __restrict__ flags,
                                                     u8* __restrict__ rmark, const u32* __restrict__ pos, const u32* __restrict__ ls,
                                                     const u16* __restrict__ rbins, i64 R) {
  if (hs[HS_HALT]) return;
  const u32 n = hs[HS_NCL];
  if (n > hs[HS_CLCAP]) { if (blockIdx.x == 0 && threadIdx.x == 0) hl_halt(hs, HALT_CAND); return; }
  const u64 win = hs64[0];
  const u32 ref_r0 = hs[HS_WIN_R0], nres = hs[HS_WIN_NRES];
  u32 nflag = 0, nbad = 0;
  for (u32 i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
    if (cl_key[i] != win) continue;
    const u32 j = cl_j[i], l = ls[j], r = j + 1u + (l >> 16);
    cl_r[i] = r;
    flags[j] = 1; rmark[r] = 1;
    ++nflag;
    const u32 r0 = pos[j];
    bool bad = ((l & 0xFFFFu) + (ls[r] & 0xFFFFu)) != nres;
    for (u32 slot = 3; slot < 9 && !bad; ++slot) {
      const u32 m = (slot >= 6) ? nres - 1 : nres;     // glue of the last residue lies outside the window
      for (u32 k = 0; k < m; ++k)
        if (rbins[(i64)slot * R + r0 + k] != rbins[(i64)slot * R + ref_r0 + k]) { bad = true; break; }
    }
    nbad += bad;
  }
  for (int o = 16; o; o >>= 1) { nflag += __shfl_xor_sync(0xffffffffu, nflag, o); nbad += __shfl_xor_sync(0xffffffffu, nbad, o); }
  if ((threadIdx.x & 31) == 0) {
    if (nflag) atomicAdd(&hs[HS_FLAGGED], nflag);
    if (nbad) atomicAdd(&hs[HS_MISMATCH], nbad);
  }
}
// overlap rule (bpe.py:1909-1914) + merge: the head of every maximal run of flagged pairs walks it left to right and
// merges the 1st, 3rd, ... occurrence in place; merge events are appended in arbitrary order (the host sorts each
// iteration's events by position when it reads the log)
__global__ void __launch_bounds__(256) hl_apply_list(u32* __restrict__ hs, const u64* __restrict__ hs64, const u64* __restrict__ cl_key,
                                                      const u32* __restrict__ cl_j, const u8* __restrict__ flags,
                                                      const u8* __restrict__ rmark, u64* __restrict__ hsh, u32* __restrict__ glue,
                                                      u32* __restrict__ ls, u32* __restrict__ tid, const u32* __restrict__ pos,
                                                      u32* __restrict__ ev_left, u32* __restrict__ ev_right, u32* __restrict__ ev_p) {
  if (hs[HS_HALT]) return;
  const bool t0 = blockIdx.x == 0 && threadIdx.x == 0;
  if (hs[HS_MISMATCH]) { if (t0) hl_halt(hs, HALT_COLLISION); return; }
  if (!hs[HS_DIST] && hs[HS_FLAGGED] != hs[HS_WIN_CNT]) { if (t0) hl_halt(hs, HALT_COUNT); return; }
  if (hs[HS_EVN] + hs[HS_FLAGGED] > hs[HS_EVCAP]) { if (t0) hl_halt(hs, HALT_EVLOG); return; }
  const u32 n = hs[HS_NCL], ev0 = hs[HS_EVN], new_tid = hs[HS_NTOK];
  const u64 win = hs64[0];
  for (u32 i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
    if (cl_key[i] != win) continue;
    u32 j = cl_j[i];
    if (rmark[j]) continue;                      // not a run head: the pair to its left is an occurrence too
    bool keep = true;
    u32 l = ls[j];                                 // carried along the run: a merged right token's ls is zeroed below
    while (flags[j]) {
      const u32 r = j + 1u + (l >> 16), lr = ls[r];
      if (keep) {
        hsh[j] = win;
        glue[j] = glue[r];
        ls[j] = ((l & 0xFFFFu) + (lr & 0xFFFFu)) | (((l >> 16) + 1u + (lr >> 16)) << 16);
        tid[j] = new_tid;
        ls[r] = 0u;                                // dead until the next compaction
        u32 q = ev0 + atomicAdd(&hs[HS_NEWT], 1u);   // HS_NEWT counts this iteration's merges
        ev_left[q] = pos[j]; ev_right[q] = pos[r]; ev_p[q] = 0u;
      }
      keep = !keep;
      j = r; l = lr;                               // runs never cross a backbone: the last token's pair is never flagged
    }
  }
}
__global__ void __launch_bounds__(256) hl_unflag_list(const u32* __restrict__ hs, const u64* __restrict__ hs64,
                                                       const u64* __restrict__ cl_key, const u32* __restrict__ cl_j,
                                                       const u32* __restrict__ cl_r, u8* __restrict__ flags, u8* __restrict__ rmark) {
  if (hs[HS_HALT]) return;
  const u32 n = hs[HS_NCL];
  const u64 win = hs64[0];
  for (u32 i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x)
    if (cl_key[i] == win) { flags[cl_j[i]] = 0; rmark[cl_r[i]] = 0; }
}
// Slot counters between full recounts: UPPER BOUNDS maintained from the merge events.  A merge creates at most two new
// pairs (left neighbour, new token) and (new token, right neighbour): their slots are incremented here, from the
// post-merge arrays; pairs that disappear are not decremented (except the winner's own occurrences, all of which are
// gone), so slots[s] >= true count of slot s at all times -- which is all the candidate logic needs: a key outside the
// slots with bound >= thr has count <= its slot's count <= bound < thr.  k_slot_count makes the bounds exact again at
// the start of every period.
__global__ void __launch_bounds__(256) hl_bump_list(const u32* __restrict__ hs, const u64* __restrict__ hs64, const u64* __restrict__ cl_key,
                                                     const u32* __restrict__ cl_j, const u64* __restrict__ hsh, const u32* __restrict__ glue,
                                                     const u32* __restrict__ ls, const u32* __restrict__ tid, const u64* __restrict__ powtab,
                                                     u64 P, u32* __restrict__ slots, int shift, const u32* __restrict__ cl_r,
                                                     u8* __restrict__ flags, u8* __restrict__ rmark) {
  if (hs[HS_HALT]) return;
  const u32 n = hs[HS_NCL], new_tid = hs[HS_NTOK];
  const i64 N = (i64)hs[HS_N];
  const u64 win = hs64[0];
  if (blockIdx.x == 0 && threadIdx.x == 0 && hs[HS_FLAGGED]) atomicSub(&slots[gb_slot_of(win, shift)], hs[HS_FLAGGED]);
  for (u32 i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
    if (cl_key[i] != win) continue;
    const u32 j = cl_j[i], l = ls[j];
    flags[j] = 0; rmark[cl_r[i]] = 0;                   // (= hl_unflag_list: nothing here reads the marks)
    if (l == 0u || tid[j] != new_tid) continue;        // the skipped occurrences of a run: their left token was absorbed
    const u64 hj = hsh[j];
    const u32 gj = glue[j];
    if (!(gj & GB_LAST_FLAG)) {                         // (new token, right neighbour)
      const i64 r = (i64)j + 1 + (l >> 16);
      if (r < N) {
        const u64 k = gb_key_h(hj, gj, hsh[r], P, __ldg(powtab + (2u * (ls[r] & 0xFFFFu) - 1u)));
        atomicAdd(&slots[gb_slot_of(k, shift)], 1u);
      }
    }
    i64 q = (i64)j - 1;                                  // (left neighbour, new token): nearest live entry to the left
    while (q >= 0 && ls[q] == 0u) --q;
    if (q >= 0) {
      const u32 gq = glue[q];
      // a freshly merged left neighbour has counted this pair as its own right pair already
      if (!(gq & GB_LAST_FLAG) && tid[q] != new_tid) {
        const u64 k = gb_key_h(hsh[q], gq, hj, P, __ldg(powtab + (2u * (l & 0xFFFFu) - 1u)));
        atomicAdd(&slots[gb_slot_of(k, shift)], 1u);
      }
    }
  }
}
// commit the iteration (step log, counters) and reset the per-iteration words for the next one
__global__ void hl_end(u32* __restrict__ hs, const u64* __restrict__ hs64, u64* __restrict__ log_key, u32* __restrict__ log_cnt,
                       u32* __restrict__ log_r0, u32* __restrict__ log_nres, u32* __restrict__ log_evstart, u32* __restrict__ log_nk) {
  if (hs[HS_HALT]) return;
  const u32 s = hs[HS_STEP], nk = hs[HS_NEWT];
  log_key[s] = hs64[0]; log_cnt[s] = hs[HS_WIN_CNT]; log_r0[s] = hs[HS_WIN_R0]; log_nres[s] = hs[HS_WIN_NRES];
  log_evstart[s] = hs[HS_EVN]; log_nk[s] = nk;
  hs[HS_EVN] += nk; hs[HS_T] -= nk; hs[HS_MPREV] = hs[HS_WIN_CNT]; hs[HS_STEP] = s + 1; hs[HS_NTOK] += 1;
  // reset of the per-iteration words (the host writes them as zeros when it (re)starts the loop)
  hs[HS_C1] = 0; hs[HS_NCAND] = 0; hs[HS_NZ] = 0; hs[HS_PAD0] = 0;
  hs[HS_M1] = 0; hs[HS_NT] = 0; hs[HS_ND] = 0; hs[HS_ROUND2] = 0;
  hs[HS_MISMATCH] = 0; hs[HS_FLAGGED] = 0; hs[HS_OVF] = 0; hs[HS_NEWT] = 0; hs[HS_NCL] = 0;
  if (s + 1 >= hs[HS_MAXSTEPS]) hl_halt(hs, HALT_STEPS);
}

// ---- occasional physical compaction of the lazily compacted arrays (runs between iterations, ignores the halt flag) --
#define HL_CH_THREADS 256
#define HL_CH_ITEMS 8
#define HL_CHUNK (HL_CH_THREADS * HL_CH_ITEMS)
__global__ void __launch_bounds__(HL_CH_THREADS) hl_cmp_count(const u32* __restrict__ hs, const u32* __restrict__ ls, u32* __restrict__ bsum) {
  const i64 N = (i64)hs[HS_N];
  const i64 base = (i64)blockIdx.x * HL_CHUNK;
  if (base >= N) return;
  __shared__ u32 wsum[HL_CH_THREADS / 32];
  u32 c = 0;
#pragma unroll
  for (int it = 0; it < HL_CH_ITEMS; ++it) {
    i64 i = base + it * HL_CH_THREADS + threadIdx.x;
    if (i < N && ls[i] != 0u) ++c;
  }
  for (int o = 16; o; o >>= 1) c += __shfl_xor_sync(0xffffffffu, c, o);
  if ((threadIdx.x & 31) == 0) wsum[threadIdx.x >> 5] = c;
  __syncthreads();
  if (threadIdx.x == 0) {
    u32 t = 0;
    for (int w = 0; w < HL_CH_THREADS / 32; ++w) t += wsum[w];
    bsum[blockIdx.x] = t;
  }
}
// in-place exclusive scan of bsum[0..nb), total -> hs[HS_CMPT] (single block)
__global__ void __launch_bounds__(1024) hl_cmp_scan(u32* __restrict__ hs, u32* __restrict__ bsum) {
  const i64 nb = ((i64)hs[HS_N] + HL_CHUNK - 1) / HL_CHUNK;
  __shared__ u32 wtot[32];
  __shared__ u32 carry_s;
  if (threadIdx.x == 0) carry_s = 0;
  __syncthreads();
  const unsigned lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  for (i64 base = 0; base < nb; base += 1024) {
    i64 i = base + threadIdx.x;
    u32 v = (i < nb) ? bsum[i] : 0u;
    u32 x = v;
    for (int o = 1; o < 32; o <<= 1) {
      u32 y = __shfl_up_sync(0xffffffffu, x, o);
      if (lane >= (unsigned)o) x += y;
    }
    if (lane == 31) wtot[warp] = x;
    __syncthreads();
    u32 woff = 0, tot = 0;
    for (unsigned w = 0; w < 32; ++w) {
      u32 t = wtot[w];
      if (w < warp) woff += t;
      tot += t;
    }
    u32 carry = carry_s;
    if (i < nb) bsum[i] = carry + woff + x - v;
    __syncthreads();
    if (threadIdx.x == 0) carry_s = carry + tot;
    __syncthreads();
  }
  if (threadIdx.x == 0) hs[HS_CMPT] = carry_s;
}
struct HLTok { const u64* hsh; const u32* glue; const u32* pos; const u32* ls; const u32* tid;
               u64* hsh_o; u32* glue_o; u32* pos_o; u32* ls_o; u32* tid_o; };
__global__ void __launch_bounds__(HL_CH_THREADS) hl_cmp_scatter(const u32* __restrict__ hs, HLTok a, const u32* __restrict__ boff) {
  const i64 N = (i64)hs[HS_N];
  const i64 base = (i64)blockIdx.x * HL_CHUNK;
  if (base >= N) return;
  __shared__ u32 wsum[HL_CH_THREADS / 32];
  u32 run = boff[blockIdx.x];
#pragma unroll
  for (int it = 0; it < HL_CH_ITEMS; ++it) {
    i64 i = base + it * HL_CH_THREADS + threadIdx.x;
    u32 l = (i < N) ? a.ls[i] : 0u;
    bool f = l != 0u;
    u32 tot;
    u32 ex = gb_block_excl_scan_flag(f, wsum, tot);
    if (f) {
      u32 o = run + ex;
      a.hsh_o[o] = a.hsh[i]; a.glue_o[o] = a.glue[i]; a.pos_o[o] = a.pos[i]; a.tid_o[o] = a.tid[i];
      a.ls_o[o] = l & 0xFFFFu;                 // every live token's right neighbour is the next entry again
    }
    run += tot;
  }
}
__global__ void hl_cmp_end(u32* __restrict__ hs) { hs[HS_N] = hs[HS_CMPT]; }

// ---- host entry: enqueue `nsteps` iterations -------------------------------------------------------------------------
enum { HA_HS = 0, HA_HS64, HA_HSH0, HA_HSH1, HA_GLUE0, HA_GLUE1, HA_POS0, HA_POS1, HA_LEN0, HA_LEN1, HA_TID0, HA_TID1,
       HA_FLAGS, HA_RMARK, HA_SCRATCH, HA_POWTAB, HA_P, HA_RBINS, HA_R, HA_SLOTS, HA_LOG2NS, HA_BMAX, HA_CAND, HA_CAP,
       HA_VARIANT, HA_CTAS, HA_TK, HA_TC, HA_TM, HA_M, HA_TIEK, HA_TIEM, HA_STRBUF, HA_STRIDE, HA_EVL, HA_EVR, HA_EVP,
       HA_LKEY, HA_LCNT, HA_LR0, HA_LNRES, HA_LEVS, HA_LNK, HA_CLKEY, HA_CLJ, HA_CLR, HA_NMAX, HA_STREAMING,
       HA_PERIOD, HA_XBUF, HA_XALL, HA_WORLD, HA_RANK, HA_TIESK, HA_TIESM, HA_SBUF, HA_LOGSTR, HA_LOGSTRLEN,
       HA_PEERS, HA_GSLOTS, HA_SRED, HA_PCTR, HA_N };

// optional CUDA-event timing of the pair-count kernel inside the loop (bench.py's roofline leg)
#include <vector>
static int g_time_k1 = 0;        // 1: events around k_slot_count, 2: around the round-1 hl_cand_count
static cudaStream_t g_cap_stream = nullptr;   // private stream the CUDA graphs are captured on
static std::vector<cudaEvent_t> g_ev;
static size_t g_ev_used = 0;
static cudaEvent_t hl_event() {
  if (g_ev_used == g_ev.size()) { cudaEvent_t e; cudaEventCreate(&e); g_ev.push_back(e); }
  return g_ev[g_ev_used++];
}

template <bool CS> static int h_loop_part_impl(const int64_t* a, int part, int parity, cudaStream_t st, bool full);
static bool g_round2 = true;           // in-loop second candidate round (else: such an iteration halts to the host path)
static bool g_incremental = true;      // slot counters: full recount once per period, upper bounds in between
static std::vector<int64_t> g_graph_args;      // argument set the captured period graphs belong to (h_loop_graphs)
// (the kernel sequence of an iteration depends on these two switches: captured graphs are dropped when they change)
extern "C" int gb_h_loop_round2(int enable) { g_round2 = enable != 0; g_graph_args.clear(); return 0; }
extern "C" int gb_h_loop_incremental(int enable) { g_incremental = enable != 0; g_graph_args.clear(); return 0; }

template <bool CS>
static int h_loop_impl(const int64_t* a, int nsteps, int parity, int step0, int compact_after, int* parity_out, cudaStream_t st) {
  u32* hs = (u32*)a[HA_HS];
  u64* hs64 = (u64*)a[HA_HS64];
  u64* hsh[2] = {(u64*)a[HA_HSH0], (u64*)a[HA_HSH1]};
  u32* glue[2] = {(u32*)a[HA_GLUE0], (u32*)a[HA_GLUE1]};
  u32* pos[2] = {(u32*)a[HA_POS0], (u32*)a[HA_POS1]};
  u32* len[2] = {(u32*)a[HA_LEN0], (u32*)a[HA_LEN1]};
  u32* tid[2] = {(u32*)a[HA_TID0], (u32*)a[HA_TID1]};
  u8* flags = (u8*)a[HA_FLAGS];
  u8* rmark = (u8*)a[HA_RMARK];
  u32* scratch = (u32*)a[HA_SCRATCH];
  const u64* powtab = (const u64*)a[HA_POWTAB];
  const u64 P = (u64)a[HA_P];
  const u16* rbins = (const u16*)a[HA_RBINS];
  const i64 R = a[HA_R];
  u32* slots = (u32*)a[HA_SLOTS];
  const int log2ns = (int)a[HA_LOG2NS];
  u32* bmax = (u32*)a[HA_BMAX];
  u32* cand = (u32*)a[HA_CAND];
  const int cap = (int)a[HA_CAP];
  const int cache = ((int)a[HA_VARIANT] >> 1) & 3;
  const int ctas = a[HA_CTAS] > 0 ? (int)a[HA_CTAS] : 2;
  u64* tk = (u64*)a[HA_TK]; u32* tc = (u32*)a[HA_TC]; u32* tm = (u32*)a[HA_TM];
  const i64 M = a[HA_M];
  u64* tie_keys = (u64*)a[HA_TIEK]; u32* tie_min = (u32*)a[HA_TIEM];
  char* strbuf = (char*)a[HA_STRBUF];
  const int stride = (int)a[HA_STRIDE];
  u64* cl_key = (u64*)a[HA_CLKEY]; u32* cl_j = (u32*)a[HA_CLJ]; u32* cl_r = (u32*)a[HA_CLR];
  const i64 Nmax = a[HA_NMAX];
  const int period = a[HA_PERIOD] > 0 ? (int)a[HA_PERIOD] : 64;
  if (log2ns < 12 || log2ns > 30 || M <= 0 || (M & (M - 1)) || cap > KC_SET / 2 || cache > 1 || (stride & 127) || Nmax < 2) {
    gb_set_error("gb_h_loop: bad arguments (log2ns %d, M %lld, cap %d, cache %d, stride %d, Nmax %lld)", log2ns, (long long)M, cap,
                 cache, stride, (long long)Nmax);
    return -2;
  }
  int sms = gb_sm_count();
  if (sms <= 0) sms = 148;
  const int shift = 64 - log2ns;
  const unsigned scan_blocks = (unsigned)(((i64)1 << log2ns) / KS_CHUNK);
  const int count_blocks = gb_blocks_for(Nmax, 2 * KS_TILE, sms * ctas);
  const int pass_blocks = gb_blocks_for(Nmax, 2 * KS_TILE, sms * 2);
  const int list_blocks = sms * 2;
  const unsigned nb = (unsigned)((Nmax + HL_CHUNK - 1) / HL_CHUNK);
  const size_t smem = cache ? 2 * sizeof(u32) * KS_CACHE : 0;
  if (cache && !(g_cap_stream && st == g_cap_stream))
    GB_CK(cudaFuncSetAttribute(k_slot_count<KeysHLT<CS>, false, 1, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  int p = parity & 1;
  bool dirty = false;              // dead entries may exist since the last compaction
  auto compact = [&]() -> int {
    const int q = p ^ 1;
    hl_cmp_count<<<nb, HL_CH_THREADS, 0, st>>>(hs, len[p], scratch);
    hl_cmp_scan<<<1, 1024, 0, st>>>(hs, scratch);
    HLTok tok{hsh[p], glue[p], pos[p], len[p], tid[p], hsh[q], glue[q], pos[q], len[q], tid[q]};
    hl_cmp_scatter<<<nb, HL_CH_THREADS, 0, st>>>(hs, tok, scratch);
    hl_cmp_end<<<1, 1, 0, st>>>(hs);
    GB_LAUNCH_CK();
    p = q;
    dirty = false;
    return 0;
  };
  for (int s = 0; s < nsteps; ++s) {
    const bool full = !g_incremental || (step0 + s) % period == 0;
    if (a[HA_PEERS]) {
      // sharded run over the peer-memory transport: the four parts back to back, no host-issued collective in between,
      // so whole periods are captured and replayed exactly like the one-GPU loop
      for (int part = 0; part < 4; ++part) { int rc = h_loop_part_impl<CS>(a, part, p, st, full); if (rc) return rc; }
      dirty = true;
      if ((step0 + s + 1) % period == 0) { int rc = compact(); if (rc) return rc; }
      continue;
    }
    KeysHLT<CS> kf{hsh[p], glue[p], len[p], powtab, P};
    if (full) {       // exact slot counts at the start of a period; upper bounds (hl_bump_list) in between
      GB_CK(cudaMemsetAsync(slots, 0, sizeof(u32) << log2ns, st));
      if (g_time_k1 == 1) GB_CK(cudaEventRecord(hl_event(), st));
      if (cache) k_slot_count<KeysHLT<CS>, false, 1, false><<<count_blocks, KS_THREADS, smem, st>>>(kf, Nmax, slots, shift, nullptr, hs);
      else k_slot_count<KeysHLT<CS>, false, 0, false><<<count_blocks, KS_THREADS, 0, st>>>(kf, Nmax, slots, shift, nullptr, hs);
      if (g_time_k1 == 1) GB_CK(cudaEventRecord(hl_event(), st));
    }
    hl_slots_scan<<<scan_blocks, 256, 0, st>>>((const uint4*)slots, bmax, hs, cand, (u32)cap);
    for (int round = 1; round <= (g_round2 ? 2 : 1); ++round) {
      if (round == 2) hl_slots_collect<<<scan_blocks, 256, 0, st>>>((const uint4*)slots, bmax, hs, round, cand, (u32)cap);
      if (g_time_k1 == 2 && round == 1) GB_CK(cudaEventRecord(hl_event(), st));
      hl_cand_count<CS><<<pass_blocks, KS_THREADS, 0, st>>>(kf, shift, cand, (u32)cap, tk, tc, tm, (u64)(M - 1), hs, round, cl_key, cl_j);
      if (g_time_k1 == 2 && round == 1) GB_CK(cudaEventRecord(hl_event(), st));
      hl_decide<<<1, 1024, 0, st>>>(tk, tc, tm, M, hs, round, tie_keys, tie_min, (u32)cap, g_round2 ? 1 : 0);
    }
    hl_tiebreak<<<1, HL_TB_CAP, 0, st>>>(hs, hs64, tie_keys, tie_min, pos[p], len[p], rbins, R, strbuf, stride);
    hl_flag_list<<<list_blocks, 256, 0, st>>>(hs, hs64, cl_key, cl_j, cl_r, flags, rmark, pos[p], len[p], rbins, R);
    hl_apply_list<<<list_blocks, 256, 0, st>>>(hs, hs64, cl_key, cl_j, flags, rmark, hsh[p], glue[p], len[p], tid[p], pos[p],
                                               (u32*)a[HA_EVL], (u32*)a[HA_EVR], (u32*)a[HA_EVP]);
    if (g_incremental)
      hl_bump_list<<<list_blocks, 256, 0, st>>>(hs, hs64, cl_key, cl_j, hsh[p], glue[p], len[p], tid[p], powtab, P, slots, shift, cl_r, flags, rmark);
    else
      hl_unflag_list<<<list_blocks, 256, 0, st>>>(hs, hs64, cl_key, cl_j, cl_r, flags, rmark);
    hl_end<<<1, 1, 0, st>>>(hs, hs64, (u64*)a[HA_LKEY], (u32*)a[HA_LCNT], (u32*)a[HA_LR0], (u32*)a[HA_LNRES], (u32*)a[HA_LEVS],
                            (u32*)a[HA_LNK]);
    GB_LAUNCH_CK();
    dirty = true;
    if ((step0 + s + 1) % period == 0) { int rc = compact(); if (rc) return rc; }
  }
  if (compact_after && (dirty || nsteps == 0)) { int rc = compact(); if (rc) return rc; }
  if (parity_out) *parity_out = p;
  return 0;
}

// CUDA graphs: one graph per buffer parity holding `period` iterations + the compaction that ends the period, captured
// once per argument set and replayed (the loop is launch-bound otherwise: ~17 launches per ~100 us iteration)
static cudaGraphExec_t g_graph_exec[2] = {nullptr, nullptr};
static bool g_use_graphs = true;
extern "C" int gb_h_loop_use_graphs(int enable) { g_use_graphs = enable != 0; return 0; }

template <bool CS>
static int h_loop_graphs(const int64_t* a, int period, cudaStream_t st) {
  std::vector<int64_t> key(a, a + HA_N);
  if (g_graph_exec[0] && key == g_graph_args) return 0;
  for (int p = 0; p < 2; ++p)
    if (g_graph_exec[p]) { cudaGraphExecDestroy(g_graph_exec[p]); g_graph_exec[p] = nullptr; }
  if (!g_cap_stream) GB_CK(cudaStreamCreateWithFlags(&g_cap_stream, cudaStreamNonBlocking));
  if (((int)a[HA_VARIANT] >> 1) & 3)    // function attributes are set outside the capture
    GB_CK(cudaFuncSetAttribute(k_slot_count<KeysHLT<CS>, false, 1, false>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                               (int)(2 * sizeof(u32) * KS_CACHE)));
  for (int p = 0; p < 2; ++p) {
    cudaGraph_t graph = nullptr;
    int pout = 0;
    GB_CK(cudaStreamBeginCapture(g_cap_stream, cudaStreamCaptureModeThreadLocal));
    int rc = h_loop_impl<CS>(a, period, p, 0, 0, &pout, g_cap_stream);
    cudaError_t e = cudaStreamEndCapture(g_cap_stream, &graph);
    if (rc) { if (graph) cudaGraphDestroy(graph); return rc; }
    GB_CK(e);
    GB_CK(cudaGraphInstantiate(&g_graph_exec[p], graph, 0));
    GB_CK(cudaGraphDestroy(graph));
  }
  g_graph_args = key;
  (void)st;
  return 0;
}

extern "C" int gb_h_loop(const int64_t* args_host, int n_args, int nsteps, int parity, int step0, int compact_after,
                         int* parity_out_host, void* stream) {
  if (n_args != HA_N) { gb_set_error("gb_h_loop: expected %d arguments, got %d", (int)HA_N, n_args); return -2; }
  if (nsteps < 0) return -2;
  cudaStream_t st = (cudaStream_t)stream;
  const bool cs = args_host[HA_STREAMING] != 0;
  const int period = args_host[HA_PERIOD] > 0 ? (int)args_host[HA_PERIOD] : 64;
  int p = parity & 1;
  // whole periods through the captured graphs
  if (g_use_graphs && !g_time_k1 && step0 % period == 0 && nsteps >= period) {
    int rc = cs ? h_loop_graphs<true>(args_host, period, st) : h_loop_graphs<false>(args_host, period, st);
    if (rc) return rc;
    while (nsteps >= period) {
      GB_CK(cudaGraphLaunch(g_graph_exec[p], st));
      p ^= 1; nsteps -= period; step0 += period;
    }
    if (nsteps == 0 && !compact_after) { if (parity_out_host) *parity_out_host = p; return 0; }
    if (nsteps == 0) compact_after = 0;      // the last period ended with a compaction
    if (nsteps == 0) { if (parity_out_host) *parity_out_host = p; return 0; }
  }
  return cs ? h_loop_impl<true>(args_host, nsteps, p, step0, compact_after, parity_out_host, st)
            : h_loop_impl<false>(args_host, nsteps, p, step0, compact_after, parity_out_host, st);
}

// ---- sharded run: one iteration in four parts; the host issues a collective between consecutive parts -------------
//   part 0: memset(slots), k_slot_count            -> allreduce(sum) slots
//   part 1: scan, pass 2, export                   -> allgather x
//   part 2: import, decide, tie strings            -> allreduce(max) sbuf
//   part 3: tie pick, flag, apply, unflag, end
// (an iteration that would need a second candidate round halts and runs on the host path: ~1 in 1000)
template <bool CS>
static int h_loop_part_impl(const int64_t* a, int part, int parity, cudaStream_t st, bool full) {
  u32* hs = (u32*)a[HA_HS];
  u64* hs64 = (u64*)a[HA_HS64];
  const int p = parity & 1;
  u64* hsh = (u64*)a[p ? HA_HSH1 : HA_HSH0];
  u32* glue = (u32*)a[p ? HA_GLUE1 : HA_GLUE0];
  u32* pos = (u32*)a[p ? HA_POS1 : HA_POS0];
  u32* len = (u32*)a[p ? HA_LEN1 : HA_LEN0];
  u32* tid = (u32*)a[p ? HA_TID1 : HA_TID0];
  u8* flags = (u8*)a[HA_FLAGS];
  u8* rmark = (u8*)a[HA_RMARK];
  const u16* rbins = (const u16*)a[HA_RBINS];
  const i64 R = a[HA_R];
  u32* slots = (u32*)a[HA_SLOTS];
  const int log2ns = (int)a[HA_LOG2NS];
  u32* bmax = (u32*)a[HA_BMAX];
  u32* cand = (u32*)a[HA_CAND];
  const int cap = (int)a[HA_CAP];
  const int cache = ((int)a[HA_VARIANT] >> 1) & 3;
  const int ctas = a[HA_CTAS] > 0 ? (int)a[HA_CTAS] : 2;
  u64* tk = (u64*)a[HA_TK]; u32* tc = (u32*)a[HA_TC]; u32* tm = (u32*)a[HA_TM];
  const i64 M = a[HA_M];
  u64* tie_keys = (u64*)a[HA_TIEK]; u32* tie_min = (u32*)a[HA_TIEM];
  u64* tie_skeys = (u64*)a[HA_TIESK]; u32* tie_smin = (u32*)a[HA_TIESM];
  char* sbuf = (char*)a[HA_SBUF];
  const int stride = (int)a[HA_STRIDE];
  u64* cl_key = (u64*)a[HA_CLKEY]; u32* cl_j = (u32*)a[HA_CLJ]; u32* cl_r = (u32*)a[HA_CLR];
  u64* x = (u64*)a[HA_XBUF];
  const u64* xall = (const u64*)a[HA_XALL];
  const int world = (int)a[HA_WORLD], rank = (int)a[HA_RANK];
  const i64 Nmax = a[HA_NMAX];
  // peer-memory transport (k_hloop_dist.cuh): a[HA_PEERS] = host table u64 [4][world] of every rank's slots / x / sbuf /
  // flags addresses (row entries of this rank = the local ones, which is also what HA_SLOTS / HA_XBUF / HA_SBUF hold)
  const u64* ptab = (const u64*)a[HA_PEERS];
  HlPeers PP;
  u32* gslots = (u32*)a[HA_GSLOTS];
  char* sred = (char*)a[HA_SRED];
  u32* pctr = (u32*)a[HA_PCTR];
  if (ptab) {
    if (world > HLP_MAXW || !gslots || !sred || !pctr) { gb_set_error("gb_h_loop_part: bad peer arguments"); return -2; }
    for (int r = 0; r < world; ++r) {
      PP.slots[r] = ptab[r]; PP.x[r] = ptab[world + r]; PP.s[r] = ptab[2 * world + r]; PP.flags[r] = ptab[3 * world + r];
    }
    if ((u64)slots != PP.slots[rank] || (u64)x != PP.x[rank] || (u64)sbuf != PP.s[rank]) {
      gb_set_error("gb_h_loop_part: local buffers are not the peer table's entries for this rank");
      return -2;
    }
  }
  if (!x || !xall || !sbuf || world < 1 || log2ns < 12 || M <= 0 || (M & (M - 1)) || cap > KC_SET / 2 || cache > 1 || (stride & 127)) {
    gb_set_error("gb_h_loop_part: bad arguments");
    return -2;
  }
  int sms = gb_sm_count();
  if (sms <= 0) sms = 148;
  const int shift = 64 - log2ns;
  const unsigned scan_blocks = (unsigned)(((i64)1 << log2ns) / KS_CHUNK);
  const int count_blocks = gb_blocks_for(Nmax, 2 * KS_TILE, sms * ctas);
  const int pass_blocks = gb_blocks_for(Nmax, 2 * KS_TILE, sms * 2);
  const int list_blocks = sms * 2;
  KeysHLT<CS> kf{hsh, glue, len, (const u64*)a[HA_POWTAB], (u64)a[HA_P]};
  switch (part) {
    case 0: {
      const size_t smem = cache ? 2 * sizeof(u32) * KS_CACHE : 0;
      if (!full) {        // the local counters are upper bounds kept by hl_bump_list since the last recount
        if (ptab) hl_peer_signal<<<1, 32, 0, st>>>(PP, world, rank, 0, pctr);
        break;
      }
      GB_CK(cudaMemsetAsync(slots, 0, sizeof(u32) << log2ns, st));
      if (g_time_k1 == 1) GB_CK(cudaEventRecord(hl_event(), st));
      if (cache) {
        if (!(g_cap_stream && st == g_cap_stream))
          GB_CK(cudaFuncSetAttribute(k_slot_count<KeysHLT<CS>, false, 1, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        k_slot_count<KeysHLT<CS>, false, 1, false><<<count_blocks, KS_THREADS, smem, st>>>(kf, Nmax, slots, shift, nullptr, hs);
      } else {
        k_slot_count<KeysHLT<CS>, false, 0, false><<<count_blocks, KS_THREADS, 0, st>>>(kf, Nmax, slots, shift, nullptr, hs);
      }
      if (g_time_k1 == 1) GB_CK(cudaEventRecord(hl_event(), st));
      if (ptab) hl_peer_signal<<<1, 32, 0, st>>>(PP, world, rank, 0, pctr);
      break;
    }
    case 1:
      if (ptab) hl_peer_sum_slots<<<sms * 2, 256, 0, st>>>(PP, world, rank, pctr, hs, (uint4*)gslots, ((i64)1 << log2ns) / 4);
      hl_slots_scan<<<scan_blocks, 256, 0, st>>>((const uint4*)(ptab ? gslots : slots), bmax, hs, cand, (u32)cap);
      if (g_time_k1 == 2) GB_CK(cudaEventRecord(hl_event(), st));
      hl_cand_count<CS><<<pass_blocks, KS_THREADS, 0, st>>>(kf, shift, cand, (u32)cap, tk, tc, tm, (u64)(M - 1), hs, 1, cl_key, cl_j);
      if (g_time_k1 == 2) GB_CK(cudaEventRecord(hl_event(), st));
      hl_table_export<<<1, 1024, 0, st>>>(tk, tc, M, hs, 1, x);
      if (ptab) hl_peer_signal<<<1, 32, 0, st>>>(PP, world, rank, 1, pctr);
      break;
    case 2:
      if (ptab) hl_peer_gather_x<<<world, 1024, 0, st>>>(PP, world, rank, pctr, hs, (u64*)xall);
      hl_table_import<<<1, 1024, 0, st>>>(tk, tc, tm, (u64)(M - 1), hs, 1, xall, world, rank);
      hl_decide<<<1, 1024, 0, st>>>(tk, tc, tm, M, hs, 1, tie_keys, tie_min, (u32)cap, 0);
      GB_CK(cudaMemsetAsync(sbuf, 0, 4 * HL_TB_CAP + (size_t)HL_TB_CAP * stride, st));
      hl_tie_strings<<<1, HL_TB_CAP, 0, st>>>(hs, tie_keys, tie_min, tie_skeys, tie_smin, pos, len, rbins, R, sbuf, stride);
      if (ptab) hl_peer_signal<<<1, 32, 0, st>>>(PP, world, rank, 2, pctr);
      break;
    case 3:
      if (ptab) hl_peer_max_s<<<32, 256, 0, st>>>(PP, world, rank, pctr, hs, (uint4*)sred, stride);
      hl_tie_pick<<<1, HL_TB_CAP, 0, st>>>(hs, hs64, tie_skeys, tie_smin, pos, len, ptab ? sred : sbuf, stride, (char*)a[HA_LOGSTR], (u32*)a[HA_LOGSTRLEN]);
      hl_flag_list<<<list_blocks, 256, 0, st>>>(hs, hs64, cl_key, cl_j, cl_r, flags, rmark, pos, len, rbins, R);
      hl_apply_list<<<list_blocks, 256, 0, st>>>(hs, hs64, cl_key, cl_j, flags, rmark, hsh, glue, len, tid, pos,
                                                 (u32*)a[HA_EVL], (u32*)a[HA_EVR], (u32*)a[HA_EVP]);
      if (g_incremental && ptab)
        hl_bump_list<<<list_blocks, 256, 0, st>>>(hs, hs64, cl_key, cl_j, hsh, glue, len, tid, (const u64*)a[HA_POWTAB], (u64)a[HA_P], slots, shift,
                                                  cl_r, flags, rmark);
      else
        hl_unflag_list<<<list_blocks, 256, 0, st>>>(hs, hs64, cl_key, cl_j, cl_r, flags, rmark);
      hl_end<<<1, 1, 0, st>>>(hs, hs64, (u64*)a[HA_LKEY], (u32*)a[HA_LCNT], (u32*)a[HA_LR0], (u32*)a[HA_LNRES], (u32*)a[HA_LEVS],
                              (u32*)a[HA_LNK]);
      break;
    default:
      gb_set_error("gb_h_loop_part: part %d outside [0, 3]", part);
      return -2;
  }
  GB_LAUNCH_CK();
  return 0;
}
extern "C" int gb_h_loop_part(const int64_t* args_host, int n_args, int part, int parity, void* stream) {
  if (n_args != HA_N) { gb_set_error("gb_h_loop_part: expected %d arguments, got %d", (int)HA_N, n_args); return -2; }
  return args_host[HA_STREAMING] ? h_loop_part_impl<true>(args_host, part, parity, (cudaStream_t)stream, true)
                                 : h_loop_part_impl<false>(args_host, part, parity, (cudaStream_t)stream, true);
}
// ---- symmetric blocks for the peer-memory transport ------------------------------------------------------------------
// gb_peer_alloc: cudaMalloc + zero + CUDA IPC handle (64 bytes, to be sent to the other ranks by any means);
// gb_peer_open: map another rank's block into this process (peer access is enabled lazily by the driver).
extern "C" int gb_peer_alloc(int64_t bytes, void** ptr_out, unsigned char* handle_out) {
  void* p = nullptr;
  GB_CK(cudaMalloc(&p, (size_t)bytes));
  GB_CK(cudaMemset(p, 0, (size_t)bytes));
  cudaIpcMemHandle_t h;
  GB_CK(cudaIpcGetMemHandle(&h, p));
  static_assert(sizeof(h) == 64, "CUDA IPC handle size");
  memcpy(handle_out, &h, sizeof(h));
  *ptr_out = p;
  return 0;
}
extern "C" int gb_peer_open(const unsigned char* handle, void** ptr_out) {
  cudaIpcMemHandle_t h;
  memcpy(&h, handle, sizeof(h));
  GB_CK(cudaIpcOpenMemHandle(ptr_out, h, cudaIpcMemLazyEnablePeerAccess));
  return 0;
}
extern "C" int gb_peer_close(void* ptr) { GB_CK(cudaIpcCloseMemHandle(ptr)); return 0; }
extern "C" int gb_peer_free(void* ptr) { GB_CK(cudaFree(ptr)); return 0; }
extern "C" int gb_peer_max_world(void) { return HLP_MAXW; }

extern "C" int64_t gb_h_loop_xbuf_elems(void) { return 1 + 2 * HLX_CAP; }
extern "C" int64_t gb_h_loop_sbuf_bytes(int stride) { return 4 * HL_TB_CAP + (int64_t)HL_TB_CAP * stride; }

// capture the CUDA graphs of both buffer parities now (warm-up), instead of at the first whole period
extern "C" int gb_h_loop_prepare(const int64_t* args_host, int n_args, void* stream) {
  if (n_args != HA_N) { gb_set_error("gb_h_loop_prepare: expected %d arguments, got %d", (int)HA_N, n_args); return -2; }
  if (!g_use_graphs) return 0;
  const int period = args_host[HA_PERIOD] > 0 ? (int)args_host[HA_PERIOD] : 64;
  return args_host[HA_STREAMING] ? h_loop_graphs<true>(args_host, period, (cudaStream_t)stream)
                                 : h_loop_graphs<false>(args_host, period, (cudaStream_t)stream);
}
extern "C" int gb_h_loop_time_k1(int mode) { g_time_k1 = mode; g_ev_used = 0; return 0; }
// elapsed ms of every timed k_slot_count launch since gb_h_loop_time_k1(1) (call after a synchronize); returns the count
extern "C" int gb_h_loop_k1_times(float* ms_host, int cap) {
  int n = 0;
  for (size_t i = 0; i + 1 < g_ev_used && n < cap; i += 2, ++n)
    if (cudaEventElapsedTime(&ms_host[n], g_ev[i], g_ev[i + 1]) != cudaSuccess) { cudaGetLastError(); ms_host[n] = -1.f; }
  g_ev_used = 0;
  return n;
}
extern "C" int gb_h_loop_nargs(void) { return (int)HA_N; }
// kernels per one-GPU iteration: scan, candidate pass, decide, tie-break, flag, apply, bump (+ unflag), end; + the recount
extern "C" int gb_h_loop_kernels_per_step(void) { return 8 + (g_round2 ? 3 : 0) + (g_incremental ? 0 : 1); }
