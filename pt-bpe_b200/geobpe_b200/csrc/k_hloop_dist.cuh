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
