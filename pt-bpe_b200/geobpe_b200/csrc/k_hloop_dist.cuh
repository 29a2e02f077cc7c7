// Sharded (multi-GPU) variant of the device-driven loop: the pieces that differ from one GPU.  Every rank counts its own
// backbones; three exchanges per iteration make all ranks take the same decision without a host round trip:
//   1. allreduce(sum) of the slot counters                        -> identical scan / candidate slots everywhere
//   2. allgather of each rank's exact (key, count) candidates     -> hl_table_import adds the other ranks' counts
//   3. allreduce(max) of the tied keys' JSON strings (zero where a rank does not hold the key) -> same tie-break
// The collectives themselves are issued by the host between gb_h_loop_part calls (torch.distributed / NCCL on the
// same stream); nothing is read back.
#pragma once

#define HLX_CAP 4096            // (key, count) entries a rank contributes per round

// x u64 [1 + 2 * HLX_CAP]: x[0] = n | ovf << 32 | halt << 40, x[1 + i] = key, x[1 + HLX_CAP + i] = count
__global__ void __launch_bounds__(1024) hl_table_export(const u64* __restrict__ tk, const u32* __restrict__ tc, i64 M,
                                                         u32* __restrict__ hs, int round, u64* __restrict__ x) {
  __shared__ u32 s_n;
  if (threadIdx.x == 0) s_n = 0;
  __syncthreads();
  const bool active = !hs[HS_HALT] && !(round == 2 && !hs[HS_ROUND2]);
  if (active)
    for (i64 i = threadIdx.x; i < M; i += blockDim.x) {
      u32 c = tc[i];
      if (c) {
        u32 q = atomicAdd(&s_n, 1u);
        if (q < HLX_CAP) { x[1 + q] = tk[i]; x[1 + HLX_CAP + q] = c; }
      }
    }
  __syncthreads();
  if (threadIdx.x == 0) {
    u32 n = s_n, ovf = hs[HS_OVF];
    if (n > HLX_CAP) { n = 0; ovf = 1; }
    x[0] = (u64)n | ((u64)ovf << 32) | ((u64)hs[HS_HALT] << 40);
  }
}
// add the other ranks' counts to the local table (local first positions stay); overflow / halt of any rank is shared
__global__ void __launch_bounds__(1024) hl_table_import(u64* __restrict__ tk, u32* __restrict__ tc, u32* __restrict__ tm, u64 mask,
                                                         u32* __restrict__ hs, int round, const u64* __restrict__ xall, int world, int rank) {
  const i64 xs = 1 + 2 * HLX_CAP;
  u32 any_halt = 0, any_ovf = 0;
  for (int r = 0; r < world; ++r) {
    u64 h = xall[(i64)r * xs];
    any_ovf = max(any_ovf, (u32)((h >> 32) & 0xFF));
    if (!any_halt) any_halt = (u32)((h >> 40) & 0xFF);
  }
  if (any_halt) { if (threadIdx.x == 0) hl_halt(hs, any_halt); return; }      // every rank stops at the same iteration
  if (hs[HS_HALT] || (round == 2 && !hs[HS_ROUND2])) return;
  if (threadIdx.x == 0 && any_ovf > hs[HS_OVF]) hs[HS_OVF] = any_ovf;
  for (int r = 0; r < world; ++r) {
    if (r == rank) continue;
    const u64* x = xall + (i64)r * xs;
    const u32 n = (u32)(x[0] & 0xFFFFFFFFu);
    for (u32 i = threadIdx.x; i < n; i += blockDim.x)
      if (!gb_gtable_add(tk, tc, tm, mask, x[1 + i], (u32)x[1 + HLX_CAP + i], 0xFFFFFFFFu)) hs[HS_OVF] = 1;
  }
}

// tie strings at rank-independent indices (ties sorted by key value): sbuf = u32 len[HL_TB_CAP] then HL_TB_CAP strings;
// a rank that does not hold a tied key leaves its string zero, so allreduce(max) over the bytes completes the set
__global__ void __launch_bounds__(HL_TB_CAP) hl_tie_strings(u32* __restrict__ hs, const u64* __restrict__ tie_keys,
                                                            const u32* __restrict__ tie_min, u64* __restrict__ tie_skeys,
                                                            u32* __restrict__ tie_smin, const u32* __restrict__ pos,
                                                            const u32* __restrict__ ls, const u16* __restrict__ rbins, i64 R,
                                                            char* __restrict__ sbuf, int stride) {
  if (hs[HS_HALT]) return;
  const u32 nt = hs[HS_NT];
  if (nt == 0) { if (threadIdx.x == 0) hl_halt(hs, HALT_EMPTY); return; }
  if (nt > HL_TB_CAP) { if (threadIdx.x == 0) { hs[HS_NTMAX] = max(hs[HS_NTMAX], nt); hl_halt(hs, HALT_TIES); } return; }
  const u32 i = threadIdx.x;
  if (i >= nt) return;
  const u64 key = tie_keys[i];
  u32 sidx = 0;
  for (u32 k = 0; k < nt; ++k) sidx += tie_keys[k] < key;
  tie_skeys[sidx] = key;
  const u32 j = tie_min[i];
  tie_smin[sidx] = j;
  if (j == 0xFFFFFFFFu) return;                    // not on this rank
  const u32 nres = hl_pair_res(ls, j);
  if (39 * (int)nres + 200 > stride) { hl_halt(hs, HALT_TIES); return; }
  int len = hl_json_key(sbuf + 4 * HL_TB_CAP + (size_t)sidx * stride, rbins, R, pos[j], nres);
  reinterpret_cast<u32*>(sbuf)[sidx] = (u32)len;
}
// after the allreduce(max): same tournament on every rank; the winner's string goes to the step's string log
__global__ void __launch_bounds__(HL_TB_CAP) hl_tie_pick(u32* __restrict__ hs, u64* __restrict__ hs64, const u64* __restrict__ tie_skeys,
                                                         const u32* __restrict__ tie_smin, const u32* __restrict__ pos,
                                                         const u32* __restrict__ ls, const char* __restrict__ sbuf, int stride,
                                                         char* __restrict__ log_str, u32* __restrict__ log_strlen) {
  if (hs[HS_HALT]) return;
  __shared__ int s_best[HL_TB_CAP];
  const u32 nt = hs[HS_NT];
  const u32* lens = reinterpret_cast<const u32*>(sbuf);
  const char* strs = sbuf + 4 * HL_TB_CAP;
  s_best[threadIdx.x] = (int)threadIdx.x;
  __syncthreads();
  const unsigned warp = threadIdx.x >> 5, nw = blockDim.x >> 5;
  for (u32 s = 1; s < nt; s <<= 1) {
    for (u32 a = warp * 2 * s; a + s < nt; a += nw * 2 * s) {
      int ia = s_best[a], ib = s_best[a + s];
      bool b_less = hl_less(strs + (size_t)ib * stride, strs + (size_t)ia * stride, (int)lens[ib], (int)lens[ia]);
      if ((threadIdx.x & 31) == 0 && b_less) s_best[a] = ib;
    }
    __syncthreads();
  }
  const int w = s_best[0];
  const u32 wl = lens[w], step = hs[HS_STEP];
  if (wl == 0) { if (threadIdx.x == 0) hl_halt(hs, HALT_TIES); return; }      // no rank produced the string
  for (u32 k = threadIdx.x; k < wl; k += blockDim.x) log_str[(size_t)step * stride + k] = strs[(size_t)w * stride + k];
  if (threadIdx.x == 0) {
    log_strlen[step] = wl;
    const u32 j = tie_smin[w];
    hs64[0] = tie_skeys[w];
    hs[HS_WIN_J] = j;
    hs[HS_WIN_R0] = j != 0xFFFFFFFFu ? pos[j] : 0u;
    hs[HS_WIN_NRES] = j != 0xFFFFFFFFu ? hl_pair_res(ls, j) : 0u;
    hs[HS_WIN_CNT] = hs[HS_M1];
    hs[HS_NTMAX] = max(hs[HS_NTMAX], nt);
  }
}

// ---- peer-memory transport (NVLink / NVSwitch loads instead of NCCL calls) -------------------------------------------
// Every rank owns one cudaMalloc'd block [slots | x | sbuf | flags] that all other ranks have opened through CUDA IPC
// (gb_peer_alloc / gb_peer_open).  An exchange is: producer kernels write the local block -> hl_peer_signal (release
// store of the stage's sequence number into every peer's flag row) -> the consumer kernel spins (acquire loads on its OWN
// flag row) until every rank has signalled, then PULLS what it needs straight from the peers' blocks and reduces it
// while loading (slot sums, candidate lists, tie strings).  Three signals per iteration; because every rank waits on all
// three, a block is never rewritten before every peer has finished reading the previous iteration's contents (the next
// write of a region always sits behind a later wait).  Waits are bounded: a rank that sees no signal within
// HLP_TIMEOUT_NS halts with HALT_PEER and stops waiting, so a dead peer cannot hang the GPU.
#define HLP_MAXW 16
#define HLP_TIMEOUT_NS 20000000000ull
struct HlPeers { u64 slots[HLP_MAXW]; u64 x[HLP_MAXW]; u64 s[HLP_MAXW]; u64 flags[HLP_MAXW]; };

__device__ __forceinline__ void hlp_st_release(u32* p, u32 v) { asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory"); }
__device__ __forceinline__ u32 hlp_ld_acquire(const u32* p) {
  u32 v;
  asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ uint4 hlp_ld4(const uint4* p) {        // peer data: never from L1
  uint4 v;
  asm volatile("ld.relaxed.sys.global.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ u64 hlp_ld8(const u64* p) {
  u64 v;
  asm volatile("ld.relaxed.sys.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ u64 hlp_now() {
  u64 t;
  asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t));
  return t;
}

// flags u32 [3][HLP_MAXW] in the symmetric block: flags[stage][r] = last sequence number rank r completed for that stage;
// ctr u32 [4] local: ctr[stage] = sequence number this rank has signalled
__global__ void hl_peer_signal(HlPeers P, int world, int rank, int stage, u32* __restrict__ ctr) {
  __shared__ u32 s_seq;
  if (threadIdx.x == 0) s_seq = ++ctr[stage];
  __syncthreads();
  __threadfence_system();
  if ((int)threadIdx.x < world) hlp_st_release(reinterpret_cast<u32*>(P.flags[threadIdx.x]) + stage * HLP_MAXW + rank, s_seq);
}
// block-wide: returns once every rank has signalled `stage` for the current sequence number (false: timed out / peer halt)
__device__ __forceinline__ bool hl_peer_wait(const u32* __restrict__ myflags, const u32* __restrict__ ctr, int world, int stage,
                                             u32* __restrict__ hs) {
  __shared__ int s_ok;
  if (threadIdx.x == 0) {
    int ok = hs[HS_HALT] != HALT_PEER;
    if (ok) {
      const u32 target = ctr[stage];
      const u64 t0 = hlp_now();
      for (int q = 0; q < world && ok; ++q)
        while ((int)(hlp_ld_acquire(myflags + stage * HLP_MAXW + q) - target) < 0)
          if (hlp_now() - t0 > HLP_TIMEOUT_NS) { hl_halt(hs, HALT_PEER); hs[HS_HALT] = HALT_PEER; ok = 0; break; }
    }
    s_ok = ok;
  }
  __syncthreads();
  return s_ok != 0;
}

// exchange 1: gslots = sum over ranks of their slot counters, reduced while the peer loads arrive
__global__ void __launch_bounds__(256) hl_peer_sum_slots(HlPeers P, int world, int rank, const u32* __restrict__ ctr,
                                                          u32* __restrict__ hs, uint4* __restrict__ gslots, i64 n4) {
  if (!hl_peer_wait(reinterpret_cast<const u32*>(P.flags[rank]), ctr, world, 0, hs) || hs[HS_HALT]) return;
  for (i64 i = (i64)blockIdx.x * blockDim.x + threadIdx.x; i < n4; i += (i64)gridDim.x * blockDim.x) {
    uint4 acc = make_uint4(0u, 0u, 0u, 0u);
    for (int p = 0; p < world; ++p) {
      const uint4 v = hlp_ld4(reinterpret_cast<const uint4*>(P.slots[p]) + i);
      acc.x += v.x; acc.y += v.y; acc.z += v.z; acc.w += v.w;
    }
    gslots[i] = acc;
  }
}
// exchange 2: every rank's (key, count) candidate list -> xall (one block per peer; only the used part travels)
__global__ void __launch_bounds__(1024) hl_peer_gather_x(HlPeers P, int world, int rank, const u32* __restrict__ ctr,
                                                          u32* __restrict__ hs, u64* __restrict__ xall) {
  const i64 xs = 1 + 2 * HLX_CAP;
  const int p = blockIdx.x;
  u64* dst = xall + (i64)p * xs;
  if (!hl_peer_wait(reinterpret_cast<const u32*>(P.flags[rank]), ctr, world, 1, hs)) {
    if (threadIdx.x == 0) dst[0] = (u64)HALT_PEER << 40;
    return;
  }
  const u64* src = reinterpret_cast<const u64*>(P.x[p]);
  const u64 h = hlp_ld8(src);
  const u32 n = min((u32)(h & 0xFFFFFFFFu), (u32)HLX_CAP);
  if (threadIdx.x == 0) dst[0] = h;
  for (u32 i = threadIdx.x; i < n; i += blockDim.x) {
    dst[1 + i] = hlp_ld8(src + 1 + i);
    dst[1 + HLX_CAP + i] = hlp_ld8(src + 1 + HLX_CAP + i);
  }
}
// exchange 3: bytewise max over ranks of the tied keys' strings (rows < nt) and their lengths -> sred
__global__ void __launch_bounds__(256) hl_peer_max_s(HlPeers P, int world, int rank, const u32* __restrict__ ctr,
                                                      u32* __restrict__ hs, uint4* __restrict__ sred, int stride) {
  if (!hl_peer_wait(reinterpret_cast<const u32*>(P.flags[rank]), ctr, world, 2, hs) || hs[HS_HALT]) return;
  const u32 nt = min(hs[HS_NT], (u32)HL_TB_CAP);
  const i64 n4 = (4 * HL_TB_CAP + (i64)nt * stride) / 16;
  for (i64 i = (i64)blockIdx.x * blockDim.x + threadIdx.x; i < n4; i += (i64)gridDim.x * blockDim.x) {
    uint4 acc = make_uint4(0u, 0u, 0u, 0u);
    for (int p = 0; p < world; ++p) {
      const uint4 v = hlp_ld4(reinterpret_cast<const uint4*>(P.s[p]) + i);
      acc.x = __vmaxu4(acc.x, v.x); acc.y = __vmaxu4(acc.y, v.y); acc.z = __vmaxu4(acc.z, v.z); acc.w = __vmaxu4(acc.w, v.w);
    }
    sred[i] = acc;
  }
}
