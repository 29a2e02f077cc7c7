// Device-driven histogram-mode learn loop, second design: BPE.step (bpe.py:1792-2166 with rmsd_key None) for N iterations
// inside ONE persistent cooperative kernel, with the pair counts maintained incrementally from the merge events exactly
// like the reference maintains _geo_dict / _priority_dict (decrement the keys of the pairs a merge destroys,
// bpe.py:1924-1950; increment the keys of the pairs it creates, bpe.py:1992-2006) instead of being recounted.
//
//   bin()      : k_slot_count (k1_sketch.cuh) counts every pair once into the key-less slot counters.
//   period     : the slot counters are scanned (max c1, candidate slots = count >= thr = 7/8 c1), ONE streaming pass over
//                the token arrays counts the keys of the candidate slots exactly (open-addressing table: key, count,
//                example window) and lists their pairs (key, left entry).  Invariant for the rest of the period:
//                every key whose slot is a candidate is in the table with its exact count; every other key has
//                count <= its slot's count < thr.
//   iteration  : P1 argmax over the table + the reference's tie-break (json string order) -- conclusive iff the best
//                count >= thr; P2 the winner's occurrences are found in the pair list (stale entries are recognised by
//                recomputing their key) and verified bin by bin against the example window; P3 the keys of the pairs the
//                merges destroy are decremented (slot counter always, table entry when the slot is a candidate) and the
//                overlap rule (bpe.py:1909-1914) picks the 1st, 3rd, ... occurrence of every run; P4 merges in place
//                (lazy layout: ls = len | extra << 16, dead entries keep a back pointer in tid); P5 the keys of the pairs
//                the merges create are incremented (slot counter; table + pair list when the slot is a candidate; a
//                non-candidate slot that reaches thr ends the period), step log, counters.
//   The period ends (and the next one starts on the device, no host round trip) when the best exact count falls below
//   thr, a non-candidate slot reaches thr, the table / list fills up or a quarter of the entries is dead (compaction).
//
// Sharded run (world > 1, one process per GPU, backbones split by contiguous ranges): slot counters, candidate bitmap and
// the exact table are REPLICATED -- every rank holds the global counts -- and kept identical by exchanging, per iteration,
// only the keys of the pairs each rank's merges destroyed / created (a few per merge) through peer memory (CUDA IPC
// blocks; release / acquire flags; loads over NVLink inside this kernel); per period the ranks' exact (key, count,
// example window) lists.  Every rank therefore takes the same decision from its own copy: no exchange for the argmax or
// the tie-break (rbins of all ranks are replicated once at start so that any rank can order any key).
//
// Phases are separated by grid-wide barriers; every branch that decides whether a barrier is executed depends only on
// state words that were written before the previous barrier and are not written again before the next one.
#include <cooperative_groups.h>
#include <vector>
#include "k1_sketch.cuh"
#include "k_hl_json.cuh"
namespace cg = cooperative_groups;

#define H2_MAXW 16               // ranks of a sharded run
#define H2_TIMEOUT_NS 20000000000ull
#define H2_THREADS 512
#define H2_TB_CAP 128            // ties resolved on the device
#define H2_BLK_LIST 512          // candidate pairs staged per CTA before one append to the global list
#define H2_CHUNK (H2_THREADS * 8)
#define H2_PROBES 1024

enum { H2_HALT_DIVERGED = 12, H2_HALT_TIES = 1, H2_HALT_CAND = 2, H2_HALT_COLLISION = 3, H2_HALT_TABLE = 4, H2_HALT_EVLOG = 5, H2_HALT_EMPTY = 6,
       H2_HALT_STEPS = 7, H2_HALT_COUNT = 8, H2_HALT_PEER = 9, H2_HALT_INTERNAL = 10, H2_HALT_GUARD = 11 };

// device state block (u32 words)
enum {
  H2_T = 0, H2_STEP, H2_HALT, H2_EVN, H2_N, H2_NTOK, H2_LOGCAP, H2_EVCAP, H2_TARGET, H2_PAR, H2_PERIOD_OK, H2_THR, H2_THR_REQ,
  H2_C1, H2_NCAND, H2_NCL, H2_CLCAP, H2_NOCC, H2_M1, H2_NT, H2_P1, H2_WIN_SLOT, H2_WIN_R0, H2_WIN_NRES, H2_WIN_CNT,
  H2_MISMATCH, H2_NEWT, H2_OVF, H2_ERR, H2_HALT_STEP, H2_NTMAX, H2_CMPT, H2_FRESH, H2_ND, H2_NZ, H2_MPREV,
  H2_ST_PERIODS, H2_ST_CROSS, H2_ST_INCONCL, H2_ST_COMPACT, H2_ST_R2, H2_ST_FULL, H2_ST_LOOPS, H2_PERIOD_LEN,
  H2_ST_MAXNCL, H2_ST_MAXND, H2_ST_MAXNCAND, H2_ST_TIES,
  H2_HALT_REQ, H2_PERIOD_REQ, H2_NDEATH, H2_NBIRTH, H2_DSEQ, H2_XSEQ, H2_ST_WAIT_US, H2_TRACE, H2_WAIT_STAGE, H2_WAIT_SEQ, H2_WAIT_SEEN0, H2_WAIT_SEEN1, H2_ST_RETRY, H2_XCROSS,
  H2_WORDS = 64
};
// hs64: [0] popped key, [1] spare, [2..] nanoseconds accumulated per phase (block 0's view)
enum { H2_TM_PS = 2, H2_TM_P1, H2_TM_P2, H2_TM_P3, H2_TM_P4, H2_TM_P5, H2_TM_CMP, H2_TM_CAND,
       H2_TW_P2 /* block 0's own work (before the barrier) */, H2_TW_P3, H2_TW_P4, H2_TW_P5, H2_TM_XCH,
       H2_CK_OUT = 16 /* checksum of the keys this rank pushed this iteration */, H2_CK_ACC, H2_CK_EXP, H2_W64 = 32 };

struct H2Args {
  u32* hs; u64* hs64;
  u64* hsh[2]; u32* glue[2]; u32* pos[2]; u32* ls[2]; u32* tid[2];
  u8* flags; u8* rmark; u32* scratch;
  const u64* powtab; u64 P; const u16* rbins; i64 R;
  u32* slots; int log2ns; u32* bmax; u32* cand; u32 cap; u32* cbit;
  u64* tk; u32* tc; u32* tr0; u32* tnres; u32* tlist; i64 M;
  u64* cl_key; u32* cl_j;
  u32* occ_j; u32* occ_r;
  char* strbuf; int stride; const u16* erank;     // erank u16 [2][1024]: order of the element strings "<v>," / "<v>]"
  u32* ev_left; u32* ev_right; u32* ev_p;
  u64* log_key; u32* log_cnt; u32* log_r0; u32* log_nres; u32* log_evs; u32* log_nk;
  u32 thr_num;                    // candidate threshold = thr_num / 16 of the top slot count
  u32 max_loops;
  // sharded run
  u32 rbase;                      // global residue index of this shard's residue 0 (rbins holds every rank's rows)
  int world, rank;
  u64 peer_d[H2_MAXW];            // every rank's delta blocks   [2 parities][H2D header | deaths | births]
  u64 peer_x[H2_MAXW];            // every rank's export blocks  [2 parities][header | key, count, r0, nres]
  u64 peer_f[H2_MAXW];            // every rank's flag rows      u32 [2 stages][H2_MAXW]
  u32 dcap, xcap;
  int verify_inbox;               // re-read the inbox until its checksum equals the sum the writers announce (bring-up aid)
  u64 timeout_ns;                 // a rank that sees no signal from a peer for this long halts (code 9) instead of spinning on
};

// ---- peer memory: release / acquire flags and loads that bypass L1 (another GPU writes these addresses) ------------------
// signal = a system-scope reduction performed at the flag's home memory (the peer's L2), not a store that could sit in a
// local cache of peer lines while this kernel keeps running
__device__ __forceinline__ void h2p_st_release(u32* p, u32 v) { asm volatile("red.release.sys.global.max.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory"); }
__device__ __forceinline__ u32 h2p_ld_acquire(const u32* p) {
  u32 v;
  asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ u32 h2p_ld32(const u32* p) {
  u32 v;
  asm volatile("ld.relaxed.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ u64 h2p_ld64(const u64* p) {
  u64 v;
  asm volatile("ld.relaxed.sys.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
  return v;
}
// delta block of one parity: u32 hdr[8] = {nocc, halt_req, period_req, n_deaths, n_births}, u64 dkey[dcap], u64 bkey[dcap],
// u32 br0[dcap], u32 bnres[dcap]
struct H2D { u32* hdr; u64* dkey; u64* bkey; u32* br0; u32* bnres; };
__device__ __forceinline__ size_t h2d_bytes(u32 dcap) { return 32 + (size_t)dcap * 24; }
// (push model: rank r WRITES its list into slot [par][r] of every rank's block; a rank only ever READS its own block, whose
//  home is its own L2 -- remote lines cached by a long-running kernel are never relied upon)
__device__ __forceinline__ H2D h2d_at(u64 base, u32 dcap, u32 par, int world, int src) {
  unsigned char* p = reinterpret_cast<unsigned char*>(base) + ((size_t)par * world + src) * h2d_bytes(dcap);
  H2D d;
  d.hdr = reinterpret_cast<u32*>(p);
  d.dkey = reinterpret_cast<u64*>(p + 32);
  d.bkey = d.dkey + dcap;
  d.br0 = reinterpret_cast<u32*>(d.bkey + dcap);
  d.bnres = d.br0 + dcap;
  return d;
}
// export block of one parity: u32 hdr[8] = {n, halt_req}, u64 key[xcap], u32 cnt[xcap], u32 r0[xcap], u32 nres[xcap]
struct H2X { u32* hdr; u64* key; u32* cnt; u32* r0; u32* nres; };
__device__ __forceinline__ size_t h2x_bytes(u32 xcap) { return 32 + (size_t)xcap * 20; }
__device__ __forceinline__ H2X h2x_at(u64 base, u32 xcap, u32 par, int world, int src) {
  unsigned char* p = reinterpret_cast<unsigned char*>(base) + ((size_t)par * world + src) * h2x_bytes(xcap);
  H2X x;
  x.hdr = reinterpret_cast<u32*>(p);
  x.key = reinterpret_cast<u64*>(p + 32);
  x.cnt = reinterpret_cast<u32*>(x.key + xcap);
  x.r0 = x.cnt + xcap;
  x.nres = x.r0 + xcap;
  return x;
}

__device__ __forceinline__ u32 h2_ld(const u32* p) { return *reinterpret_cast<const volatile u32*>(p); }
__device__ __forceinline__ u64 h2_now() {
  u64 t;
  asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t));
  return t;
}
__device__ __forceinline__ void h2_halt(u32* hs, u32 code) {
  if (atomicCAS(&hs[H2_HALT], 0u, code) == 0u) hs[H2_HALT_STEP] = hs[H2_STEP];
}
// progress marker of block 0 (readable from a second stream while the kernel runs: multi-GPU bring-up)
__device__ __forceinline__ void h2_trace(u32* hs, u32 v) {
  if (blockIdx.x == 0 && threadIdx.x == 0) *reinterpret_cast<volatile u32*>(hs + H2_TRACE) = v;      // (no fence: a marker, not a signal)
}
__device__ __forceinline__ bool h2_cbit(const u32* cbit, u32 s) { return (cbit[s >> 5] >> (s & 31u)) & 1u; }

// exact table: open addressing, keys never leave during a period (counts may fall to zero)
__device__ __forceinline__ i64 h2_table_add(const H2Args& a, u64 key, u32 r0, u32 nres) {
  const u64 mask = (u64)(a.M - 1);
  u64 s = gb_mix64(key) & mask;
  for (int probe = 0; probe < H2_PROBES; ++probe) {
    u64 prev = *reinterpret_cast<volatile u64*>(&a.tk[s]);
    if (prev == GB_EMPTY_KEY) {
      prev = atomicCAS(&a.tk[s], (u64)GB_EMPTY_KEY, key);
      if (prev == GB_EMPTY_KEY) {             // this thread inserted the key: it also provides the example window
        a.tr0[s] = r0; a.tnres[s] = nres;
        const u32 id = atomicAdd(&a.hs[H2_ND], 1u);       // dense list of the occupied slots (the argmax scans it)
        if ((i64)id < a.M) a.tlist[id] = (u32)s;
        if ((u64)id * 2u + 2u > (u64)a.M) a.hs[H2_OVF] = 1u;
        prev = key;
      }
    }
    if (prev == key) { atomicAdd(&a.tc[s], 1u); return (i64)s; }
    s = (s + 1) & mask;
  }
  a.hs[H2_OVF] = 1u;
  return -1;
}
// add c occurrences of a key counted by another rank (its example window is a global residue range: valid everywhere)
__device__ __forceinline__ void h2_table_addc(const H2Args& a, u64 key, u32 c, u32 r0, u32 nres) {
  const u64 mask = (u64)(a.M - 1);
  u64 s = gb_mix64(key) & mask;
  for (int probe = 0; probe < H2_PROBES; ++probe) {
    u64 prev = *reinterpret_cast<volatile u64*>(&a.tk[s]);
    if (prev == GB_EMPTY_KEY) {
      prev = atomicCAS(&a.tk[s], (u64)GB_EMPTY_KEY, key);
      if (prev == GB_EMPTY_KEY) {
        a.tr0[s] = r0; a.tnres[s] = nres;
        const u32 id = atomicAdd(&a.hs[H2_ND], 1u);
        if ((i64)id < a.M) a.tlist[id] = (u32)s;
        if ((u64)id * 2u + 2u > (u64)a.M) a.hs[H2_OVF] = 1u;
        prev = key;
      }
    }
    if (prev == key) { atomicAdd(&a.tc[s], c); return; }
    s = (s + 1) & mask;
  }
  a.hs[H2_OVF] = 1u;
}
__device__ __forceinline__ i64 h2_table_find(const H2Args& a, u64 key) {
  const u64 mask = (u64)(a.M - 1);
  u64 s = gb_mix64(key) & mask;
  for (int probe = 0; probe < H2_PROBES; ++probe) {
    u64 prev = *reinterpret_cast<volatile u64*>(&a.tk[s]);
    if (prev == key) return (i64)s;
    if (prev == GB_EMPTY_KEY) return -1;
    s = (s + 1) & mask;
  }
  return -1;
}

struct H2Tok { u64* hsh; u32* glue; u32* pos; u32* ls; u32* tid; };
__device__ __forceinline__ H2Tok h2_tok(const H2Args& a, int par) {
  return H2Tok{a.hsh[par], a.glue[par], a.pos[par], a.ls[par], a.tid[par]};
}
// nearest live entry to the left of j (0xFFFFFFFF: none).  A dead entry keeps the index of the token that absorbed it in tid.
__device__ __forceinline__ u32 h2_prev_live(const H2Tok& t, u32 j) {
  u32 q = j - 1u;
  while (q != 0xFFFFFFFFu && t.ls[q] == 0u) q = t.tid[q];
  return q;
}
__device__ __forceinline__ u64 h2_pair_key(const H2Args& a, const H2Tok& t, u32 left, u32 right) {
  return gb_key_h(t.hsh[left], t.glue[left], t.hsh[right], a.P, __ldg(a.powtab + (2u * (t.ls[right] & 0xFFFFu) - 1u)));
}

// a pair with key K disappears / appears
__device__ __forceinline__ void h2_death(const H2Args& a, u64 key) {
  const u32 s = gb_slot_of(key, 64 - a.log2ns);
  atomicSub(&a.slots[s], 1u);
  if (a.world > 1) {                               // the other ranks replay it on their copies
    const u32 g = atomicAdd(&a.hs[H2_NDEATH], 1u), par = a.hs[H2_DSEQ] & 1u;
    if (g < a.dcap) { for (int q = 0; q < a.world; ++q) h2d_at(a.peer_d[q], a.dcap, par, a.world, a.rank).dkey[g] = key; }
    else a.hs[H2_HALT_REQ] = H2_HALT_TABLE;
    if (a.verify_inbox) atomicAdd(&a.hs64[H2_CK_OUT], key);
  }
  if (h2_cbit(a.cbit, s)) {
    const i64 t = h2_table_find(a, key);
    if (t < 0) a.hs[H2_ERR] = 1u; else atomicSub(&a.tc[t], 1u);
  }
}
__device__ __forceinline__ void h2_birth(const H2Args& a, u64 key, u32 left, u32 r0, u32 nres, u32 thr) {
  const u32 s = gb_slot_of(key, 64 - a.log2ns);
  const u32 old = atomicAdd(&a.slots[s], 1u);
  const bool dist = a.world > 1;
  // (sharded: whatever ends a period must end it on every rank -- requests travel with the deltas, the threshold test runs on
  //  the final global counters in the next iteration's first phase)
  if (dist) {
    const u32 g = atomicAdd(&a.hs[H2_NBIRTH], 1u), par = a.hs[H2_DSEQ] & 1u;
    if (g < a.dcap) {
      for (int q = 0; q < a.world; ++q) {
        const H2D d = h2d_at(a.peer_d[q], a.dcap, par, a.world, a.rank);
        d.bkey[g] = key; d.br0[g] = r0; d.bnres[g] = nres;
      }
    } else a.hs[H2_HALT_REQ] = H2_HALT_TABLE;
    if (a.verify_inbox) atomicAdd(&a.hs64[H2_CK_OUT], 3ull * key + r0);
  }
  if (h2_cbit(a.cbit, s)) {
    if (h2_table_add(a, key, r0, nres) < 0) { if (dist) a.hs[H2_HALT_REQ] = H2_HALT_TABLE; else a.hs[H2_PERIOD_OK] = 0u; }
    const u32 g = atomicAdd(&a.hs[H2_NCL], 1u);
    if (g < a.hs[H2_CLCAP]) { a.cl_key[g] = key; a.cl_j[g] = left; }
    else { if (dist) a.hs[H2_PERIOD_REQ] = 1u; else a.hs[H2_PERIOD_OK] = 0u; a.hs[H2_ST_FULL] = 1u; }
  } else if (!dist && old + 1u >= thr) {
    a.hs[H2_PERIOD_OK] = 0u;             // a key outside the table may have reached the threshold: new period
    a.hs[H2_ST_CROSS] = 1u;
  }
}

// ---- sharded run: signal / wait, the per-iteration delta exchange, the per-period table exchange --------------------------
// stage 0 = deltas, 1 = exports; seq = this use's sequence number (its low bit selects the buffer parity)
__device__ void h2_signal_wait(const H2Args& a, int stage, u32 seq) {
  if (blockIdx.x != 0 || threadIdx.x >= 32) return;
  const int q = (int)threadIdx.x;              // lane q signals rank q and waits for rank q: the NVLink round trips overlap
  __threadfence_system();                       // (this lane's header stores to rank q precede its signal)
  if (q < a.world) h2p_st_release(reinterpret_cast<u32*>(a.peer_f[q]) + stage * H2_MAXW + a.rank, seq);
  __threadfence_system();
  const u32* mine = reinterpret_cast<const u32*>(a.peer_f[a.rank]) + stage * H2_MAXW;
  const u64 t0 = h2_now();
  if (q < a.world && a.hs[H2_HALT] != H2_HALT_PEER)
    while ((int)(h2p_ld_acquire(mine + q) - seq) < 0)
      if (h2_now() - t0 > a.timeout_ns) {
        a.hs[H2_WAIT_STAGE] = (u32)stage; a.hs[H2_WAIT_SEQ] = seq;              // what this rank was waiting for, what it saw
        a.hs[H2_WAIT_SEEN0] = h2p_ld_acquire(mine + 0); a.hs[H2_WAIT_SEEN1] = h2p_ld_acquire(mine + 1);
        h2_halt(a.hs, H2_HALT_PEER);
        break;
      }
  __syncwarp();
  if (q == 0) a.hs[H2_ST_WAIT_US] += (u32)((h2_now() - t0) / 1000ull);
}
// checksum of what has LANDED in this rank's inbox (the other ranks' slots), to be compared with the sum the writers announce
__device__ void h2_inbox_checksum(const H2Args& a, u32 par) {
  const i64 gtid = (i64)blockIdx.x * H2_THREADS + threadIdx.x, gsz = (i64)gridDim.x * H2_THREADS;
  u64 acc = 0;
  for (int r = 0; r < a.world; ++r) {
    if (r == a.rank) continue;
    const H2D d = h2d_at(a.peer_d[a.rank], a.dcap, par, a.world, r);
    const u32 nd = min(h2p_ld32(d.hdr + 3), a.dcap), nb = min(h2p_ld32(d.hdr + 4), a.dcap);
    for (i64 i = gtid; i < nd; i += gsz) acc += h2p_ld64(d.dkey + i);
    for (i64 i = gtid; i < nb; i += gsz) acc += 3ull * h2p_ld64(d.bkey + i) + h2p_ld32(d.br0 + i);
  }
  if (acc) atomicAdd(&a.hs64[H2_CK_ACC], acc);
  if (gtid == 0) {
    u64 e = 0;
    for (int r = 0; r < a.world; ++r)
      if (r != a.rank) {
        const H2D d = h2d_at(a.peer_d[a.rank], a.dcap, par, a.world, r);
        e += (u64)h2p_ld32(d.hdr + 5) | ((u64)h2p_ld32(d.hdr + 6) << 32);
      }
    a.hs64[H2_CK_EXP] = e;
  }
}
// replay the other ranks' deltas of this iteration on the local copies; block 0 thread 0 folds the headers
__device__ void h2_apply_remote(const H2Args& a, u32 par, u32 cnt) {
  u32* hs = a.hs;
  if (blockIdx.x == 0 && threadIdx.x < 32) {       // lane r folds rank r's header
    const int r = (int)threadIdx.x;
    u32 n = 0, h = 0, pr = 0, dv = 0;
    if (r < a.world) {
      const H2D d = h2d_at(a.peer_d[a.rank], a.dcap, par, a.world, r);
      n = h2p_ld32(d.hdr + 0); h = h2p_ld32(d.hdr + 1); pr = h2p_ld32(d.hdr + 2);
      dv = h2p_ld32(d.hdr + 7) != (cnt ^ (u32)*reinterpret_cast<volatile u64*>(&a.hs64[0]));
    }
    for (int o = 16; o; o >>= 1) { n += __shfl_xor_sync(0xffffffffu, n, o); pr |= __shfl_xor_sync(0xffffffffu, pr, o); dv |= __shfl_xor_sync(0xffffffffu, dv, o); }
    const unsigned hb = __ballot_sync(0xffffffffu, h != 0);
    const u32 halt = hb ? __shfl_sync(0xffffffffu, h, __ffs(hb) - 1) : 0u;      // the lowest rank's request: the same on every rank
    if (r == 0) {
      if (halt) h2_halt(hs, halt);
      else if (dv) h2_halt(hs, H2_HALT_DIVERGED);
      else if (n != cnt) h2_halt(hs, H2_HALT_COUNT);
      if (pr) hs[H2_PERIOD_OK] = 0u;
    }
  }
  const i64 gtid = (i64)blockIdx.x * H2_THREADS + threadIdx.x, gsz = (i64)gridDim.x * H2_THREADS;
  u32 nds[H2_MAXW], nbs[H2_MAXW];
#pragma unroll
  for (int r = 0; r < H2_MAXW; ++r) {              // all counts first: the loads overlap instead of gating each rank's loop
    nds[r] = nbs[r] = 0;
    if (r < a.world && r != a.rank) {
      const H2D d = h2d_at(a.peer_d[a.rank], a.dcap, par, a.world, r);
      nds[r] = h2p_ld32(d.hdr + 3); nbs[r] = h2p_ld32(d.hdr + 4);
    }
  }
#pragma unroll
  for (int r = 0; r < H2_MAXW; ++r) {
    if (r >= a.world || r == a.rank) continue;
    const H2D d = h2d_at(a.peer_d[a.rank], a.dcap, par, a.world, r);
    const u32 nd = min(nds[r], a.dcap), nb = min(nbs[r], a.dcap);
    for (i64 i = gtid; i < nd; i += gsz) {
      const u64 key = h2p_ld64(d.dkey + i);
      const u32 s = gb_slot_of(key, 64 - a.log2ns);
      atomicSub(&a.slots[s], 1u);
      if (h2_cbit(a.cbit, s)) {
        const i64 t = h2_table_find(a, key);
        if (t < 0) hs[H2_ERR] = 1u; else atomicSub(&a.tc[t], 1u);
      }
    }
    for (i64 i = gtid; i < nb; i += gsz) {
      const u64 key = h2p_ld64(d.bkey + i);
      const u32 s = gb_slot_of(key, 64 - a.log2ns);
      atomicAdd(&a.slots[s], 1u);
      if (h2_cbit(a.cbit, s)) h2_table_addc(a, key, 1u, h2p_ld32(d.br0 + i), h2p_ld32(d.bnres + i));
    }
  }
}
// threshold test on the FINAL global counters: did a pair created in the previous iteration (on any rank) lift a slot
// outside the candidate set to the threshold?  (runs beside the argmax of the next iteration; blocks 1..)
__device__ void h2_cross_check(const H2Args& a, u32 par, u32 thr) {
  if (blockIdx.x == 0 || gridDim.x < 2) return;
  const i64 gtid = (i64)(blockIdx.x - 1) * H2_THREADS + threadIdx.x, gsz = (i64)(gridDim.x - 1) * H2_THREADS;
  u32 nbs[H2_MAXW];
#pragma unroll
  for (int r = 0; r < H2_MAXW; ++r)
    nbs[r] = r < a.world ? h2p_ld32(h2d_at(a.peer_d[a.rank], a.dcap, par, a.world, r).hdr + 4) : 0u;
#pragma unroll
  for (int r = 0; r < H2_MAXW; ++r) {
    if (r >= a.world) continue;
    const H2D d = h2d_at(a.peer_d[a.rank], a.dcap, par, a.world, r);
    const u32 nb = min(nbs[r], a.dcap);
    for (i64 i = gtid; i < nb; i += gsz) {
      const u32 s = gb_slot_of(h2p_ld64(d.bkey + i), 64 - a.log2ns);
      if (!h2_cbit(a.cbit, s) && a.slots[s] >= thr) a.hs[H2_XCROSS] = 1u;
    }
  }
}
// period start: every rank publishes the exact (key, count, example window) entries of its shard; all ranks add the others'
__device__ void h2_export_table(const H2Args& a, u32 par) {
  const u32 nd = min(h2_ld(a.hs + H2_ND), (u32)a.M);
  const i64 gtid = (i64)blockIdx.x * H2_THREADS + threadIdx.x, gsz = (i64)gridDim.x * H2_THREADS;
  for (i64 i = gtid; i < nd && i < a.xcap; i += gsz) {
    const u32 sl = a.tlist[i];
    const u64 k = a.tk[sl];
    const u32 c = a.tc[sl], r0 = a.tr0[sl], nr = a.tnres[sl];
    for (int q = 0; q < a.world; ++q) {
      if (q == a.rank) continue;
      const H2X x = h2x_at(a.peer_x[q], a.xcap, par, a.world, a.rank);
      x.key[i] = k; x.cnt[i] = c; x.r0[i] = r0; x.nres[i] = nr;
    }
  }
  if (gtid == 0) {
    const u32 halt = (nd > a.xcap || a.hs[H2_OVF]) ? (u32)H2_HALT_TABLE : (a.hs[H2_NCL] > a.hs[H2_CLCAP] ? (u32)H2_HALT_CAND : 0u);
    for (int q = 0; q < a.world; ++q) {
      const H2X x = h2x_at(a.peer_x[q], a.xcap, par, a.world, a.rank);
      x.hdr[0] = min(nd, a.xcap); x.hdr[1] = halt;
    }
  }
  __threadfence_system();        // this thread's stores into the peers' blocks are performed before the barrier / the signal
}
__device__ void h2_import_tables(const H2Args& a, u32 par) {
  const i64 gtid = (i64)blockIdx.x * H2_THREADS + threadIdx.x, gsz = (i64)gridDim.x * H2_THREADS;
  if (gtid == 0) {
    u32 halt = 0;
    for (int r = 0; r < a.world; ++r) {
      const u32 h = h2p_ld32(h2x_at(a.peer_x[a.rank], a.xcap, par, a.world, r).hdr + 1);
      if (!halt) halt = h;
    }
    if (halt) h2_halt(a.hs, halt);
  }
  for (int r = 0; r < a.world; ++r) {
    if (r == a.rank) continue;
    const H2X x = h2x_at(a.peer_x[a.rank], a.xcap, par, a.world, r);
    const u32 n = min(h2p_ld32(x.hdr + 0), a.xcap);
    for (i64 i = gtid; i < n; i += gsz) {
      const u32 c = h2p_ld32(x.cnt + i);
      if (c) h2_table_addc(a, h2p_ld64(x.key + i), c, h2p_ld32(x.r0 + i), h2p_ld32(x.nres + i));
    }
  }
}

// ---- physical compaction of the lazily compacted arrays (buffer par -> par ^ 1) -------------------------------------
__device__ void h2_compact(const H2Args& a, cg::grid_group& grid, int par, u32* s_u32) {
  u32* hs = a.hs;
  const i64 N = (i64)h2_ld(hs + H2_N);
  const i64 nb = (N + H2_CHUNK - 1) / H2_CHUNK;
  const H2Tok src = h2_tok(a, par), dst = h2_tok(a, par ^ 1);
  for (i64 c = blockIdx.x; c < nb; c += gridDim.x) {
    const i64 base = c * H2_CHUNK;
    u32 cnt = 0;
#pragma unroll
    for (int it = 0; it < 8; ++it) {
      const i64 i = base + it * H2_THREADS + threadIdx.x;
      if (i < N && src.ls[i] != 0u) ++cnt;
    }
    for (int o = 16; o; o >>= 1) cnt += __shfl_xor_sync(0xffffffffu, cnt, o);
    __syncthreads();
    if ((threadIdx.x & 31) == 0) s_u32[threadIdx.x >> 5] = cnt;
    __syncthreads();
    if (threadIdx.x == 0) {
      u32 t = 0;
      for (int w = 0; w < H2_THREADS / 32; ++w) t += s_u32[w];
      a.scratch[c] = t;
    }
  }
  grid.sync();
  if (blockIdx.x == 0) {          // exclusive scan of the chunk counts
    const unsigned lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    u32 carry = 0;
    for (i64 base = 0; base < nb; base += H2_THREADS) {
      const i64 i = base + threadIdx.x;
      const u32 v = (i < nb) ? a.scratch[i] : 0u;
      u32 x = v;
      for (int o = 1; o < 32; o <<= 1) {
        u32 y = __shfl_up_sync(0xffffffffu, x, o);
        if (lane >= (unsigned)o) x += y;
      }
      __syncthreads();
      if (lane == 31) s_u32[warp] = x;
      __syncthreads();
      u32 woff = 0, tot = 0;
      for (unsigned w = 0; w < H2_THREADS / 32; ++w) {
        u32 t = s_u32[w];
        if (w < warp) woff += t;
        tot += t;
      }
      if (i < nb) a.scratch[i] = carry + woff + x - v;
      carry += tot;
    }
    if (threadIdx.x == 0) hs[H2_CMPT] = carry;
  }
  grid.sync();
  for (i64 c = blockIdx.x; c < nb; c += gridDim.x) {
    const i64 base = c * H2_CHUNK;
    u32 run = a.scratch[c];
#pragma unroll 1
    for (int it = 0; it < 8; ++it) {
      const i64 i = base + it * H2_THREADS + threadIdx.x;
      const u32 l = (i < N) ? src.ls[i] : 0u;
      const bool f = l != 0u;
      u32 tot;
      const u32 ex = gb_block_excl_scan_flag(f, s_u32, tot);
      if (f) {
        const u32 o = run + ex;
        dst.hsh[o] = src.hsh[i]; dst.glue[o] = src.glue[i]; dst.pos[o] = src.pos[i]; dst.tid[o] = src.tid[i];
        dst.ls[o] = l & 0xFFFFu;
      }
      run += tot;
    }
  }
  grid.sync();
  if (blockIdx.x == 0 && threadIdx.x == 0) {
    hs[H2_N] = hs[H2_CMPT];
    hs[H2_PAR] = (u32)(par ^ 1);
    hs[H2_ST_COMPACT] += 1u;
  }
  grid.sync();
}

// ---- period start ---------------------------------------------------------------------------------------------------
// candidate collection: slots with count >= thr, from the chunks whose maximum qualifies
__device__ void h2_collect(const H2Args& a, u32 thr) {
  const i64 nchunks = ((i64)1 << a.log2ns) / KS_CHUNK;
  const uint4* slots4 = reinterpret_cast<const uint4*>(a.slots);
  for (i64 c = blockIdx.x; c < nchunks; c += gridDim.x) {
    if (a.bmax[c] < thr) continue;
    const i64 base4 = c * (KS_CHUNK / 4);
#pragma unroll
    for (int k = 0; k < KS_CHUNK / 4 / H2_THREADS; ++k) {
      const i64 i = base4 + k * H2_THREADS + threadIdx.x;
      const uint4 v = slots4[i];
      const u32 e4[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
      for (int e = 0; e < 4; ++e)
        if (e4[e] >= thr) {
          const u32 q = atomicAdd(&a.hs[H2_NCAND], 1u);
          if (q < a.cap) a.cand[q] = (u32)(4 * i + e);
        }
    }
  }
}

// exact counts of the candidate slots' keys + the list of those pairs: one streaming pass over the token arrays
template <bool CS>
__device__ void h2_cand_pass(const H2Args& a, int par, u32 n, u32* s_set, u64* s_key, u32* s_j, u32* s_misc) {
  u32* hs = a.hs;
  const i64 N = (i64)h2_ld(hs + H2_N);
  const int shift = 64 - a.log2ns;
  const H2Tok t = h2_tok(a, par);
  KeysHLT<CS> kf{t.hsh, t.glue, t.ls, a.powtab, a.P};
  const u32 clcap = hs[H2_CLCAP];
  if (threadIdx.x == 0) s_misc[0] = 0;
  for (int i = threadIdx.x; i < KC_SET; i += H2_THREADS) s_set[i] = KS_EMPTY;
  __syncthreads();
  for (u32 i = threadIdx.x; i < n; i += H2_THREADS) {
    const u32 v = a.cand[i];
    if (i % gridDim.x == blockIdx.x) atomicOr(&a.cbit[v >> 5], 1u << (v & 31u));
    u32 s = (v * 0x9E3779B1u) >> 19;
    while (true) {
      const u32 prev = atomicCAS(&s_set[s], KS_EMPTY, v);
      if (prev == KS_EMPTY || prev == v) break;
      s = (s + 1) & (KC_SET - 1);
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
        const u32 slot = gb_slot_of(key[e], shift);
        bool hit = false;
        u32 s = (slot * 0x9E3779B1u) >> 19;
        while (true) {
          const u32 v = s_set[s];
          if (v == slot) { hit = true; break; }
          if (v == KS_EMPTY) break;
          s = (s + 1) & (KC_SET - 1);
        }
        if (!hit) continue;
        const u32 j = (u32)(j0 + e);
        const u32 l = t.ls[j];
        const u32 r = j + 1u + (l >> 16);
        h2_table_add(a, key[e], a.rbase + t.pos[j], (l & 0xFFFFu) + (t.ls[r] & 0xFFFFu));
        const u32 q = atomicAdd(&s_misc[0], 1u);           // block-local list first: one global atomic per CTA at the end
        if (q < H2_BLK_LIST) { s_key[q] = key[e]; s_j[q] = j; }
        else {
          const u32 g = atomicAdd(&hs[H2_NCL], 1u);
          if (g < clcap) { a.cl_key[g] = key[e]; a.cl_j[g] = j; }
        }
      }
    }
  }
  __syncthreads();
  const u32 m = min(s_misc[0], (u32)H2_BLK_LIST);
  if (m) {
    if (threadIdx.x == 0) s_misc[1] = atomicAdd(&hs[H2_NCL], m);
    __syncthreads();
    for (u32 i = threadIdx.x; i < m; i += H2_THREADS) {
      const u32 g = s_misc[1] + i;
      if (g < clcap) { a.cl_key[g] = s_key[i]; a.cl_j[g] = s_j[i]; }
    }
  }
}

template <bool CS>
__device__ void h2_period_start(const H2Args& a, cg::grid_group& grid, unsigned char* smem) {
  u32* hs = a.hs;
  u32* s_u32 = reinterpret_cast<u32*>(smem);
  const bool t0 = blockIdx.x == 0 && threadIdx.x == 0;
  const u64 tm0 = t0 ? h2_now() : 0ull;
  // (0) compaction when a quarter of the entries is dead
  {
    const u32 N = h2_ld(hs + H2_N), T = h2_ld(hs + H2_T);
    if ((u64)(N - T) * 4ull > (u64)N) {
      h2_compact(a, grid, (int)h2_ld(hs + H2_PAR), s_u32);
      if (t0) { const u64 t1 = h2_now(); a.hs64[H2_TM_CMP] += t1 - tm0; }
    }
  }
  const int par = (int)h2_ld(hs + H2_PAR);
  h2_trace(hs, 0x20000u | 1u);
  // (A) reset: table, candidate bitmap, counters
  if (t0) { hs[H2_C1] = 0; hs[H2_NCAND] = 0; hs[H2_NCL] = 0; hs[H2_NZ] = 0; hs[H2_OVF] = 0; hs[H2_ND] = 0; hs[H2_ST_PERIODS] += 1u; }
  const i64 gtid = (i64)blockIdx.x * H2_THREADS + threadIdx.x, gsz = (i64)gridDim.x * H2_THREADS;
  for (i64 i = gtid; i < a.M; i += gsz) { a.tk[i] = GB_EMPTY_KEY; a.tc[i] = 0u; }
  const i64 nwords = ((i64)1 << a.log2ns) / 32;
  for (i64 i = gtid; i < nwords; i += gsz) a.cbit[i] = 0u;
  grid.sync();
  // (B) per-chunk maxima of the slot counters, global max
  {
    const i64 nchunks = ((i64)1 << a.log2ns) / KS_CHUNK;
    const uint4* slots4 = reinterpret_cast<const uint4*>(a.slots);
    for (i64 c = blockIdx.x; c < nchunks; c += gridDim.x) {
      const i64 base4 = c * (KS_CHUNK / 4);
      u32 mx = 0, nz = 0;
#pragma unroll
      for (int k = 0; k < KS_CHUNK / 4 / H2_THREADS; ++k) {
        const uint4 v = slots4[base4 + k * H2_THREADS + threadIdx.x];
        mx = max(max(mx, v.x), max(max(v.y, v.z), v.w));
        nz += (v.x != 0) + (v.y != 0) + (v.z != 0) + (v.w != 0);
      }
      for (int o = 16; o; o >>= 1) { mx = max(mx, __shfl_xor_sync(0xffffffffu, mx, o)); nz += __shfl_xor_sync(0xffffffffu, nz, o); }
      __syncthreads();
      if ((threadIdx.x & 31) == 0) { s_u32[threadIdx.x >> 5] = mx; s_u32[32 + (threadIdx.x >> 5)] = nz; }
      __syncthreads();
      if (threadIdx.x == 0) {
        for (int w = 1; w < H2_THREADS / 32; ++w) { mx = max(mx, s_u32[w]); nz += s_u32[32 + w]; }
        a.bmax[c] = mx;
        if (mx) atomicMax(&hs[H2_C1], mx);
        if (nz) atomicAdd(&hs[H2_NZ], nz);
      }
    }
  }
  grid.sync();
  // (C) candidate slots
  const u32 c1 = h2_ld(hs + H2_C1), req = h2_ld(hs + H2_THR_REQ);
  if (c1 == 0 || c1 > h2_ld(hs + H2_N)) {       // no pair left / a counter wrapped (the incremental updates went wrong)
    if (t0) h2_halt(hs, c1 == 0 ? H2_HALT_EMPTY : H2_HALT_INTERNAL);
    grid.sync();
    return;
  }
  u32 thr = req ? min(req, c1) : max(1u, (u32)(((u64)c1 * a.thr_num) / 16ull));
  h2_collect(a, thr);
  grid.sync();
  u32 ncand = h2_ld(hs + H2_NCAND);
  if (ncand > a.cap && thr < c1 && !req) {     // too many slots above the fraction: the top slots only
    grid.sync();                               // (every CTA has read NCAND before it is reset)
    if (t0) hs[H2_NCAND] = 0;
    thr = c1;
    grid.sync();
    h2_collect(a, thr);
    grid.sync();
    ncand = h2_ld(hs + H2_NCAND);
  }
  if (ncand > a.cap) {
    if (t0) h2_halt(hs, H2_HALT_CAND);
    grid.sync();
    return;
  }
  // (D) exact counts + pair list of the candidate slots
  h2_trace(hs, 0x20000u | 4u);
  const u64 tm1 = t0 ? h2_now() : 0ull;
  {
    u32* s_set = reinterpret_cast<u32*>(smem);
    u64* s_key = reinterpret_cast<u64*>(smem + sizeof(u32) * KC_SET);
    u32* s_j = reinterpret_cast<u32*>(smem + sizeof(u32) * KC_SET + sizeof(u64) * H2_BLK_LIST);
    u32* s_misc = s_j + H2_BLK_LIST;
    h2_cand_pass<CS>(a, par, ncand, s_set, s_key, s_j, s_misc);
  }
  grid.sync();
  if (a.world > 1) {       // the table becomes global: every rank adds the other ranks' exact counts of the same candidate slots
    const u32 xseq = h2_ld(hs + H2_XSEQ) + 1u;
    h2_trace(hs, 0x10000u | 1u);
    h2_export_table(a, xseq & 1u);
    grid.sync();
    h2_trace(hs, 0x10000u | 2u);
    h2_signal_wait(a, 1, xseq);
    h2_trace(hs, 0x10000u | 3u);
    grid.sync();
    h2_trace(hs, 0x10000u | 4u);
    h2_import_tables(a, xseq & 1u);
    if (t0) hs[H2_XSEQ] = xseq;
    grid.sync();
  }
  if (t0) {
    if (hs[H2_OVF]) h2_halt(hs, H2_HALT_TABLE);
    else if (a.world == 1 && hs[H2_NCL] > hs[H2_CLCAP]) h2_halt(hs, H2_HALT_CAND);
    hs[H2_THR] = thr; hs[H2_THR_REQ] = 0; hs[H2_PERIOD_OK] = 1; hs[H2_FRESH] = 1; hs[H2_PERIOD_LEN] = 0;
    if (req) hs[H2_ST_R2] += 1u;
    hs[H2_ST_MAXNCL] = max(hs[H2_ST_MAXNCL], hs[H2_NCL]); hs[H2_ST_MAXND] = max(hs[H2_ST_MAXND], hs[H2_ND]);
    hs[H2_ST_MAXNCAND] = max(hs[H2_ST_MAXNCAND], hs[H2_NCAND]);
    const u64 t2 = h2_now();
    a.hs64[H2_TM_PS] += t2 - tm0; a.hs64[H2_TM_CAND] += t2 - tm1;
  }
  grid.sync();
}

// ---- P1: exact argmax over the table + the reference's tie-break (block 0) -------------------------------------------
#define H2_CODE_BYTES 32768      // shared memory for the tied keys' rank sequences (larger sets go through global memory)
__device__ void h2_decide(const H2Args& a, unsigned char* smem) {
  if (blockIdx.x != 0) return;
  u32* hs = a.hs;
  u32* s_red = reinterpret_cast<u32*>(smem);                  // [32]
  u32* s_nt = s_red + 32;                                       // [1] (+ padding)
  u32* s_tslot = s_red + 64;                                    // [H2_TB_CAP]
  int* s_best = reinterpret_cast<int*>(s_tslot + H2_TB_CAP);    // [H2_TB_CAP]
  int* s_len = s_best + H2_TB_CAP;                              // [H2_TB_CAP]
  u32* s_fail = reinterpret_cast<u32*>(s_len + H2_TB_CAP);      // [1] (+ padding)
  u16* s_codes = reinterpret_cast<u16*>(s_fail + 8);            // [H2_CODE_BYTES / 2] rank sequences of the tied keys
  if (threadIdx.x == 0) { *s_nt = 0; *s_fail = 0; }
  const u32 nd = min(h2_ld(hs + H2_ND), (u32)a.M);
  u32 mx = 0;
  for (u32 i = threadIdx.x; i < nd; i += H2_THREADS) mx = max(mx, a.tc[a.tlist[i]]);
  for (int o = 16; o; o >>= 1) mx = max(mx, __shfl_xor_sync(0xffffffffu, mx, o));
  if ((threadIdx.x & 31) == 0) s_red[threadIdx.x >> 5] = mx;
  __syncthreads();
  mx = 0;
  for (int w = 0; w < H2_THREADS / 32; ++w) mx = max(mx, s_red[w]);
  if (mx)
    for (u32 i = threadIdx.x; i < nd; i += H2_THREADS) {
      const u32 sl = a.tlist[i];
      if (a.tc[sl] == mx) {
        const u32 k = atomicAdd(s_nt, 1u);
        if (k < H2_TB_CAP) s_tslot[k] = sl;
      }
    }
  __syncthreads();
  const u32 nt = *s_nt, thr = hs[H2_THR];
  if (mx == 0 || mx < thr) {               // a key outside the table could reach mx: new period
    if (threadIdx.x == 0) {
      // right after a period start the table holds every key of the slots >= thr: lower the threshold to the best exact
      // count found (every key with at least that count then lies in a candidate slot); otherwise the usual fraction
      hs[H2_THR_REQ] = hs[H2_FRESH] ? max(mx, 1u) : 0u;
      hs[H2_P1] = 1u;
    }
    return;
  }
  if (nt > H2_TB_CAP) {
    if (threadIdx.x == 0) { hs[H2_NTMAX] = max(hs[H2_NTMAX], nt); hs[H2_P1] = 100u + H2_HALT_TIES; }
    return;
  }
  const unsigned warp = threadIdx.x >> 5, nw = H2_THREADS >> 5, lane = threadIdx.x & 31;
  if (nt > 1) {
    // The reference pops min((-count, json.dumps(key, sort_keys=True))).  For pair windows (nres >= 2) that string is
    //   {"0C:1N": [0 x m], "C:1N:1CA": [..nres-1], "CA:C": [0 x nres], "CA:C:1N": [..m], "N:CA": [0 x nres], "omega": [..nres-1],
    //    "phi": [..nres-1], "psi": [..m], "tau": [..nres]}           (m = nres - 1 when the window ends its backbone)
    // i.e. nine lists with fixed headers.  Each list element prints as "<v>, " or "<v>]": these element strings form a
    // prefix-free set, so the byte order of two key strings equals the lexicographic order of their sequences of element
    // ranks (a.erank: position of "<v>," / "<v>]" among all element strings in byte order).  One warp writes a key's rank
    // sequence in parallel; the tournament compares 32 ranks per step.
    int my_len = 0;
    for (u32 i = threadIdx.x; i < nt; i += H2_THREADS) {
      const u32 sl = s_tslot[i];
      const u32 nres = a.tnres[sl], r0 = a.tr0[sl];
      const u32 m = (__ldg(a.rbins + 4 * a.R + r0 + nres - 1) == GB_NOBIN) ? nres - 1 : nres;
      s_len[i] = (int)(3u * m + 6u * nres - 3u);
      s_best[i] = (int)m;                             // (borrowed until the tournament starts)
      my_len = max(my_len, s_len[i]);
    }
    for (int o = 16; o; o >>= 1) my_len = max(my_len, __shfl_xor_sync(0xffffffffu, my_len, o));
    __syncthreads();
    if (lane == 0) s_red[warp] = (u32)my_len;
    __syncthreads();
    u32 lmax = 0;
    for (int w = 0; w < (int)nw; ++w) lmax = max(lmax, s_red[w]);
    const u32 cstride = (lmax + 31u) & ~31u;          // ranks per key, padded to whole warp steps
    u16* cb = s_codes;
    if ((size_t)nt * cstride * sizeof(u16) > H2_CODE_BYTES) {
      cb = reinterpret_cast<u16*>(a.strbuf);
      if ((size_t)nt * cstride * sizeof(u16) > (size_t)H2_TB_CAP * a.stride) {
        if (threadIdx.x == 0) hs[H2_P1] = 100u + H2_HALT_TIES;
        return;
      }
    }
    for (u32 i = warp; i < nt; i += nw) {
      const u32 sl = s_tslot[i];
      const u32 nres = a.tnres[sl], r0 = a.tr0[sl], m = (u32)s_best[i], len = (u32)s_len[i];
      // list lengths and sources (0 = the std-bonds zero list, else the residue-major rbins slot)
      const u32 seg_n[9] = {m, nres - 1, nres, m, nres, nres - 1, nres - 1, m, nres};
      const u32 seg_s[9] = {0, 7, 0, 4, 0, 6, 8, 5, 3};
      for (u32 p = lane; p < cstride; p += 32) {
        u16 code = 0;
        if (p < len) {
          u32 k = p, g = 0;
#pragma unroll
          for (int q = 0; q < 8; ++q) if (g == (u32)q && k >= seg_n[q]) { k -= seg_n[q]; g = q + 1; }
          const u32 v = seg_s[g] ? (u32)__ldg(a.rbins + (i64)seg_s[g] * a.R + r0 + k) : 0u;
          code = __ldg(a.erank + ((k == seg_n[g] - 1u) ? 1024u : 0u) + min(v, 1023u));
        }
        cb[(size_t)i * cstride + p] = code;
      }
    }
    __syncthreads();
    if (threadIdx.x < H2_TB_CAP) s_best[threadIdx.x] = (int)threadIdx.x;
    __syncthreads();
    for (u32 s = 1; s < nt; s <<= 1) {
      for (u32 x = warp * 2 * s; x + s < nt; x += nw * 2 * s) {
        const int ia = s_best[x], ib = s_best[x + s];
        bool b_less = false;
        for (u32 off = 0; off < cstride; off += 32) {
          const u16 ca = cb[(size_t)ia * cstride + off + lane], cbv = cb[(size_t)ib * cstride + off + lane];
          const unsigned diff = __ballot_sync(0xffffffffu, ca != cbv);
          if (diff) {
            const int l = __ffs(diff) - 1;
            b_less = __shfl_sync(0xffffffffu, (int)(cbv < ca), l) != 0;
            break;
          }
        }
        if (lane == 0 && b_less) s_best[x] = ib;
      }
      __syncthreads();
    }
  }
  if (threadIdx.x == 0) {
    const u32 sl = s_tslot[nt > 1 ? s_best[0] : 0];
    a.hs64[0] = a.tk[sl];
    hs[H2_WIN_SLOT] = sl; hs[H2_WIN_R0] = a.tr0[sl]; hs[H2_WIN_NRES] = a.tnres[sl]; hs[H2_WIN_CNT] = mx;
    hs[H2_M1] = mx; hs[H2_NT] = nt; hs[H2_NTMAX] = max(hs[H2_NTMAX], nt);
    if (nt > 1) hs[H2_ST_TIES] += 1u;
    hs[H2_FRESH] = 0; hs[H2_P1] = 0;
    hs[H2_NOCC] = 0; hs[H2_MISMATCH] = 0; hs[H2_NEWT] = 0;
    hs[H2_NDEATH] = 0; hs[H2_NBIRTH] = 0; hs[H2_HALT_REQ] = 0; hs[H2_PERIOD_REQ] = 0;
  }
}

// ---- P2: the winner's occurrences from the pair list --------------------------------------------------------------
#define H2_FLAG_ENT 4           // list entries per thread and round: their loads are issued together (one chain of round trips)
__device__ void h2_flag(const H2Args& a, const H2Tok& t, u64 win, u32* s_u32) {
  u32* hs = a.hs;
  const u32 ncl = min(h2_ld(hs + H2_NCL), h2_ld(hs + H2_CLCAP));
  // (block-uniform trip count: the occurrence slots are reserved once per CTA and round -- a single hot counter bumped by
  //  every matching thread costs a serialised L2 atomic per warp)
  for (u32 base = blockIdx.x * (H2_THREADS * H2_FLAG_ENT); base < ncl; base += gridDim.x * (H2_THREADS * H2_FLAG_ENT)) {
    u64 k[H2_FLAG_ENT];
    u32 j[H2_FLAG_ENT], r[H2_FLAG_ENT], l[H2_FLAG_ENT], g[H2_FLAG_ENT];
    bool valid[H2_FLAG_ENT];
#pragma unroll
    for (int e = 0; e < H2_FLAG_ENT; ++e) {
      const u32 i = base + e * H2_THREADS + threadIdx.x;
      k[e] = i < ncl ? a.cl_key[i] : ~win;
      j[e] = i < ncl ? a.cl_j[i] : 0u;         // (loaded with the key, not after the compare: one round trip less for a match)
    }
#pragma unroll
    for (int e = 0; e < H2_FLAG_ENT; ++e) {
      valid[e] = k[e] == win;
      l[e] = valid[e] ? t.ls[j[e]] : 0u;
      g[e] = valid[e] ? t.glue[j[e]] : GB_LAST_FLAG;
    }
    u32 nv = 0;
#pragma unroll
    for (int e = 0; e < H2_FLAG_ENT; ++e) {
      valid[e] = valid[e] && l[e] != 0u && !(g[e] & GB_LAST_FLAG);      // the entry is stale when the pair no longer exists
      r[e] = j[e] + 1u + (l[e] >> 16);
      if (valid[e])
        valid[e] = gb_key_h(t.hsh[j[e]], g[e], t.hsh[r[e]], a.P, __ldg(a.powtab + (2u * (t.ls[r[e]] & 0xFFFFu) - 1u))) == win;
      nv += valid[e];
    }
    if (threadIdx.x == 0) s_u32[0] = 0;
    __syncthreads();
    u32 my = nv ? atomicAdd(&s_u32[0], nv) : 0u;
    __syncthreads();
    if (threadIdx.x == 0 && s_u32[0]) s_u32[1] = atomicAdd(&hs[H2_NOCC], s_u32[0]);
    __syncthreads();
#pragma unroll
    for (int e = 0; e < H2_FLAG_ENT; ++e)
      if (valid[e]) {
        const u32 q = s_u32[1] + my++;
        a.occ_j[q] = j[e]; a.occ_r[q] = r[e];
        a.flags[j[e]] = 1; a.rmark[r[e]] = 1;
      }
    __syncthreads();
  }
}
__device__ void h2_unflag(const H2Args& a, u32 nocc) {
  for (u32 i = blockIdx.x * H2_THREADS + threadIdx.x; i < nocc; i += gridDim.x * H2_THREADS) {
    a.flags[a.occ_j[i]] = 0; a.rmark[a.occ_r[i]] = 0;
  }
}

// ---- P3: exact window check, overlap rule, the pairs the merges destroy (one warp per occurrence) -------------------
__device__ void h2_deaths(const H2Args& a, const H2Tok& t, u32 nocc, u32 ref_r0, u32 nres, bool& pushed) {
  const unsigned lane = threadIdx.x & 31;
  const u32 wid = (blockIdx.x * H2_THREADS + threadIdx.x) >> 5, nw = gridDim.x * (H2_THREADS / 32);
  for (u32 i = wid; i < nocc; i += nw) {
    const u32 j0 = a.occ_j[i], r0tok = a.occ_r[i];
    // every occurrence's window must hold the example window's bins (a 64-bit content-hash collision is detected here)
    const u32 r0 = a.rbase + t.pos[j0];
    bool bad = ((t.ls[j0] & 0xFFFFu) + (t.ls[r0tok] & 0xFFFFu)) != nres;
    if (!bad && r0 != ref_r0) {
      for (u32 e = lane; e < 6u * nres; e += 32u) {
        const u32 slot = 3u + e / nres, k = e % nres;
        if (slot >= 6u && k == nres - 1u) continue;        // glue of the last residue lies outside the window
        bad |= __ldg(a.rbins + (i64)slot * a.R + r0 + k) != __ldg(a.rbins + (i64)slot * a.R + ref_r0 + k);
      }
    }
    bad = __any_sync(0xffffffffu, bad);
    if (lane != 0) continue;
    if (bad) atomicAdd(&a.hs[H2_MISMATCH], 1u);
    if (a.rmark[j0]) continue;                 // not a run head: the pair to its left is an occurrence too
    const u32 q = h2_prev_live(t, j0);
    if (q != 0xFFFFFFFFu && !(t.glue[q] & GB_LAST_FLAG)) { h2_death(a, h2_pair_key(a, t, q, j0)); pushed = true; }
    // the head walks its run: the 1st, 3rd, ... occurrence merges (flags 2), the others lost their left token (bpe.py:1909-1914)
    u32 j = j0, tail = j0;
    bool keep = true, last_kept = false;
    while (true) {
      const u32 r = j + 1u + (t.ls[j] >> 16);
      if (keep) a.flags[j] = 2;
      last_kept = keep; tail = r;
      keep = !keep;
      if (!a.flags[r]) break;
      j = r;
    }
    if (last_kept && !(t.glue[tail] & GB_LAST_FLAG)) {
      const u32 x = tail + 1u + (t.ls[tail] >> 16);
      // (a flagged right neighbour is the head of the next run: it accounts for this pair as its own left pair)
      if (!a.flags[x]) { h2_death(a, h2_pair_key(a, t, tail, x)); pushed = true; }
    }
  }
}

// ---- P4: merges in place ------------------------------------------------------------------------------------------------
__device__ void h2_apply(const H2Args& a, const H2Tok& t, u32 nocc, u64 win, u32 new_tid, u32 ev0, u32* s_u32) {
  for (u32 base = blockIdx.x * H2_THREADS; base < nocc; base += gridDim.x * H2_THREADS) {
    const u32 i = base + threadIdx.x;
    u32 j = 0;
    bool kept = false;
    if (i < nocc) { j = a.occ_j[i]; kept = a.flags[j] == 2; }
    if (threadIdx.x == 0) s_u32[0] = 0;
    __syncthreads();
    const u32 my = kept ? atomicAdd(&s_u32[0], 1u) : 0u;        // event slots reserved once per CTA and round
    __syncthreads();
    if (threadIdx.x == 0 && s_u32[0]) s_u32[1] = atomicAdd(&a.hs[H2_NEWT], s_u32[0]);
    __syncthreads();
    if (kept) {
      const u32 r = a.occ_r[i];
      const u32 l = t.ls[j], lr = t.ls[r];
      const u32 q = ev0 + s_u32[1] + my;
      a.ev_left[q] = t.pos[j]; a.ev_right[q] = t.pos[r]; a.ev_p[q] = 0u;
      t.hsh[j] = win;
      t.glue[j] = t.glue[r];
      t.ls[j] = ((l & 0xFFFFu) + (lr & 0xFFFFu)) | (((l >> 16) + 1u + (lr >> 16)) << 16);
      t.tid[j] = new_tid;
      t.ls[r] = 0u;                              // dead until the next compaction; tid keeps the way back to the absorbing token
      t.tid[r] = j;
    }
    __syncthreads();
  }
}

// ---- P5: the pairs the merges create, marks removed, step committed -------------------------------------------------
__device__ void h2_births_end(const H2Args& a, const H2Tok& t, u32 nocc, u64 win, u32 new_tid, u32 thr, u32 step, u32 ev0, bool& pushed) {
  u32* hs = a.hs;
  for (u32 i = blockIdx.x * H2_THREADS + threadIdx.x; i < nocc; i += gridDim.x * H2_THREADS) {
    const u32 j = a.occ_j[i];
    const u32 f = a.flags[j];
    a.flags[j] = 0; a.rmark[a.occ_r[i]] = 0;
    if (f != 2) continue;
    const u32 l = t.ls[j], gj = t.glue[j];
    // both neighbours first (their loads overlap), then the two updates
    const bool has_r = !(gj & GB_LAST_FLAG);          // (new token, right neighbour)
    const u32 x = j + 1u + (l >> 16);
    const u32 q = h2_prev_live(t, j);                   // (left neighbour, new token); a freshly merged left neighbour has counted
    const bool has_l = q != 0xFFFFFFFFu && !(t.glue[q] & GB_LAST_FLAG) && t.tid[q] != new_tid;       // this pair as its own right pair
    u64 kr = 0, kl = 0;
    u32 nr = 0, nl = 0, pj = t.pos[j], pq = 0;
    if (has_r) { kr = h2_pair_key(a, t, j, x); nr = (l & 0xFFFFu) + (t.ls[x] & 0xFFFFu); }
    if (has_l) { kl = h2_pair_key(a, t, q, j); nl = (t.ls[q] & 0xFFFFu) + (l & 0xFFFFu); pq = t.pos[q]; }
    if (has_r) { h2_birth(a, kr, j, a.rbase + pj, nr, thr); pushed = true; }
    if (has_l) { h2_birth(a, kl, q, a.rbase + pq, nl, thr); pushed = true; }
  }
  if (blockIdx.x == 0 && threadIdx.x == 0) {
    const u32 nk = hs[H2_NEWT], cnt = hs[H2_WIN_CNT];
    a.log_key[step] = win; a.log_cnt[step] = cnt; a.log_r0[step] = hs[H2_WIN_R0]; a.log_nres[step] = hs[H2_WIN_NRES];
    a.log_evs[step] = ev0; a.log_nk[step] = nk;
    a.tc[hs[H2_WIN_SLOT]] = 0u;                               // every occurrence of the popped key is gone
    atomicSub(&a.slots[gb_slot_of(win, 64 - a.log2ns)], cnt);
    const u32 T = hs[H2_T] - nk;
    hs[H2_T] = T; hs[H2_EVN] = ev0 + nk; hs[H2_NTOK] = new_tid + 1u; hs[H2_MPREV] = cnt; hs[H2_PERIOD_LEN] += 1u;
    if ((u64)(hs[H2_N] - T) * 4ull > (u64)hs[H2_N]) {                           // compaction is due
      if (a.world > 1) hs[H2_PERIOD_REQ] = 1u; else hs[H2_PERIOD_OK] = 0u;
    }
    hs[H2_STEP] = step + 1u;
    if (step + 1u >= hs[H2_LOGCAP]) h2_halt(hs, H2_HALT_STEPS);
  }
}

// ---- the loop --------------------------------------------------------------------------------------------------------
template <bool CS>
__global__ void __launch_bounds__(H2_THREADS, 1) h2_run(H2Args a, int final_compact) {
  cg::grid_group grid = cg::this_grid();
  __shared__ __align__(16) unsigned char smem[sizeof(u32) * KC_SET + sizeof(u64) * H2_BLK_LIST + sizeof(u32) * H2_BLK_LIST + 64];
  u32* hs = a.hs;
  const bool t0 = blockIdx.x == 0 && threadIdx.x == 0;
  u32 loops = 0;
  while (true) {
    // (every path below ends with a grid barrier: the control words are stable here)
    const u32 halt = h2_ld(hs + H2_HALT), step = h2_ld(hs + H2_STEP), target = h2_ld(hs + H2_TARGET), ok = h2_ld(hs + H2_PERIOD_OK);
    if (halt || step >= target) break;
    if (++loops > a.max_loops) {
      if (t0) h2_halt(hs, H2_HALT_GUARD);
      break;
    }
    if (!ok) {
      h2_period_start<CS>(a, grid, smem);
      continue;
    }
    u64 tm = t0 ? h2_now() : 0ull;
    const bool dist = a.world > 1;
    h2_trace(hs, (loops << 8) | 1u);
    const u32 dseq0 = dist ? h2_ld(hs + H2_DSEQ) : 0u;
    h2_decide(a, smem);
    if (dist) h2_cross_check(a, dseq0 & 1u, h2_ld(hs + H2_THR));      // (previous iteration's created pairs, blocks 1..)
    grid.sync();
    if (t0) { const u64 n = h2_now(); a.hs64[H2_TM_P1] += n - tm; tm = n; }
    const u32 p1 = h2_ld(hs + H2_P1);
    if (p1) {
      grid.sync();                                   // every CTA has read P1 before the control words change
      if (t0) {
        if (p1 == 1u) { hs[H2_PERIOD_OK] = 0; hs[H2_ST_INCONCL] += 1u; }
        else h2_halt(hs, p1 - 100u);
        hs[H2_P1] = 0;
      }
      grid.sync();
      continue;
    }
    if (dist && h2_ld(hs + H2_XCROSS)) {             // a created pair lifted a slot to the threshold: the decision is void
      grid.sync();                                   // (every CTA has read the word before the control words change)
      if (t0) { hs[H2_XCROSS] = 0; hs[H2_PERIOD_OK] = 0; hs[H2_ST_CROSS] += 1u; }
      grid.sync();
      continue;
    }
    if (dist && t0) hs[H2_DSEQ] = dseq0 + 1u;        // this iteration's delta exchange (low bit = buffer parity; read from P3 on)
    const int par = (int)h2_ld(hs + H2_PAR);
    const H2Tok t = h2_tok(a, par);
    const u64 win = *reinterpret_cast<volatile u64*>(&a.hs64[0]);
    const u32 ref_r0 = h2_ld(hs + H2_WIN_R0), nres = h2_ld(hs + H2_WIN_NRES), cnt = h2_ld(hs + H2_WIN_CNT);
    const u32 new_tid = h2_ld(hs + H2_NTOK), ev0 = h2_ld(hs + H2_EVN), thr = h2_ld(hs + H2_THR);
    h2_trace(hs, (loops << 8) | 2u);
    h2_flag(a, t, win, reinterpret_cast<u32*>(smem));
    if (t0) a.hs64[H2_TW_P2] += h2_now() - tm;
    grid.sync();
    if (t0) { const u64 n = h2_now(); a.hs64[H2_TM_P2] += n - tm; tm = n; }
    u32 nocc = h2_ld(hs + H2_NOCC);
    const bool evfull = (u64)ev0 + nocc > (u64)h2_ld(hs + H2_EVCAP);
    if (!dist) {
      const u32 bad = nocc != cnt ? H2_HALT_COUNT : (evfull ? H2_HALT_EVLOG : 0u);
      if (bad) {                                     // nothing has been modified yet: remove the marks, stop
        h2_unflag(a, nocc);
        grid.sync();
        if (t0) h2_halt(hs, bad);
        grid.sync();
        continue;
      }
    } else if (evfull) {                             // fatal on this rank: the iteration runs empty here, every rank stops after the exchange
      h2_unflag(a, nocc);
      grid.sync();
      if (t0) { hs[H2_HALT_REQ] = H2_HALT_EVLOG; hs[H2_NOCC] = 0; }
      nocc = 0;
    }
    h2_trace(hs, (loops << 8) | 3u);
    bool pushed = false;
    h2_deaths(a, t, nocc, ref_r0, nres, pushed);
    if (t0) a.hs64[H2_TW_P3] += h2_now() - tm;
    grid.sync();
    if (t0) { const u64 n = h2_now(); a.hs64[H2_TM_P3] += n - tm; tm = n; }
    const u32 bad3 = h2_ld(hs + H2_MISMATCH) ? H2_HALT_COLLISION : (h2_ld(hs + H2_ERR) ? H2_HALT_INTERNAL : 0u);
    if (bad3 && !dist) {                             // the token arrays are untouched (counters are rebuilt by the host)
      h2_unflag(a, nocc);
      grid.sync();
      if (t0) h2_halt(hs, bad3);
      grid.sync();
      continue;
    }
    h2_trace(hs, (loops << 8) | 4u);
    h2_apply(a, t, nocc, win, new_tid, ev0, reinterpret_cast<u32*>(smem));
    if (t0) a.hs64[H2_TW_P4] += h2_now() - tm;
    grid.sync();
    if (t0) { const u64 n = h2_now(); a.hs64[H2_TM_P4] += n - tm; tm = n; }
    h2_trace(hs, (loops << 8) | 5u);
    h2_births_end(a, t, nocc, win, new_tid, thr, step, ev0, pushed);
    if (dist && pushed) __threadfence_system();              // this thread's pushes (P3 and P5) have landed in the peers' blocks before the signal
    if (t0) a.hs64[H2_TW_P5] += h2_now() - tm;
    grid.sync();
    if (t0) { const u64 n = h2_now(); a.hs64[H2_TM_P5] += n - tm; tm = n; hs[H2_ST_LOOPS] += 1u; }
    if (dist) {
      // publish this rank's deltas (header last), wait for every rank's, replay the others' on the local copies
      const u32 seq = dseq0 + 1u;
      if (blockIdx.x == 0 && threadIdx.x < 32) {     // lane q writes this rank's header into rank q's block
        const int q = (int)threadIdx.x;
        const u64 ck = a.hs64[H2_CK_OUT];
        if (q < a.world) {
          const H2D d = h2d_at(a.peer_d[q], a.dcap, seq & 1u, a.world, a.rank);
          d.hdr[0] = hs[H2_NOCC]; d.hdr[1] = hs[H2_HALT_REQ] ? hs[H2_HALT_REQ] : (bad3 ? bad3 : 0u); d.hdr[2] = hs[H2_PERIOD_REQ];
          d.hdr[3] = hs[H2_NDEATH]; d.hdr[4] = hs[H2_NBIRTH];
          d.hdr[5] = (u32)ck; d.hdr[6] = (u32)(ck >> 32); d.hdr[7] = cnt ^ (u32)win;
        }
        __syncwarp();
        if (q == 0) a.hs64[H2_CK_OUT] = 0ull;
      }
      h2_trace(hs, (loops << 8) | 6u);
      h2_signal_wait(a, 0, seq);
      h2_trace(hs, (loops << 8) | 7u);
      grid.sync();
      // the other ranks' lists must have LANDED completely before they are replayed: compare the checksum of what is visible here
      // with the sum the writers announce; re-read until they agree (bounded)
      bool landed = !a.verify_inbox;
      for (int attempt = 0; attempt < 2000 && !landed; ++attempt) {
        if (t0) a.hs64[H2_CK_ACC] = 0ull;
        grid.sync();
        h2_inbox_checksum(a, seq & 1u);
        grid.sync();
        landed = *reinterpret_cast<volatile u64*>(&a.hs64[H2_CK_ACC]) == *reinterpret_cast<volatile u64*>(&a.hs64[H2_CK_EXP]);
        if (!landed) {                               // (every CTA has compared before the accumulator is reset)
          grid.sync();
          if (t0) hs[H2_ST_RETRY] += 1u;
        }
      }
      if (!landed) {
        grid.sync();
        if (t0) h2_halt(hs, H2_HALT_DIVERGED);
        grid.sync();
        continue;
      }
      h2_trace(hs, (loops << 8) | 8u);
      h2_apply_remote(a, seq & 1u, cnt);
      grid.sync();
      if (t0) { const u64 n = h2_now(); a.hs64[H2_TM_XCH] += n - tm; }
    }
  }
  if (final_compact) {          // the host reads dense arrays (ls = plain residue counts)
    const u32 N = h2_ld(hs + H2_N), T = h2_ld(hs + H2_T);
    if (N != T) {
      const u64 tm0 = t0 ? h2_now() : 0ull;
      h2_compact(a, grid, (int)h2_ld(hs + H2_PAR), reinterpret_cast<u32*>(smem));
      if (t0) { hs[H2_PERIOD_OK] = 0; a.hs64[H2_TM_CMP] += h2_now() - tm0; }
    }
  }
}

// ---- host entry ------------------------------------------------------------------------------------------------------
enum { HB_HS = 0, HB_HS64, HB_HSH0, HB_HSH1, HB_GLUE0, HB_GLUE1, HB_POS0, HB_POS1, HB_LEN0, HB_LEN1, HB_TID0, HB_TID1,
       HB_FLAGS, HB_RMARK, HB_SCRATCH, HB_POWTAB, HB_P, HB_RBINS, HB_R, HB_SLOTS, HB_LOG2NS, HB_BMAX, HB_CAND, HB_CAP, HB_CBIT,
       HB_TK, HB_TC, HB_TR0, HB_TNRES, HB_TLIST, HB_M, HB_CLKEY, HB_CLJ, HB_OCCJ, HB_OCCR, HB_STRBUF, HB_STRIDE, HB_ERANK, HB_EVL, HB_EVR, HB_EVP,
       HB_LKEY, HB_LCNT, HB_LR0, HB_LNRES, HB_LEVS, HB_LNK, HB_THRNUM, HB_STREAMING, HB_CTAS,
       HB_RBASE, HB_WORLD, HB_RANK, HB_PEERS /* host table u64 [3][world]: delta / export / flag block of every rank */, HB_DCAP, HB_XCAP, HB_TIMEOUT_MS, HB_VERIFY, HB_N };

__global__ void h2_set_target(u32* hs, u32 nsteps) { hs[H2_TARGET] = hs[H2_STEP] + nsteps; }

extern "C" int gb_h_loop2_nargs(void) { return (int)HB_N; }
extern "C" int gb_h_loop2_state_words(void) { return (int)H2_WORDS; }

static int g_h2_per_sm[2] = {0, 0};    // co-resident CTAs per SM of h2_run<false>, h2_run<true> (occupancy query, once)

extern "C" int gb_h_loop2(const int64_t* b, int n_args, int nsteps, int final_compact, void* stream) {
  if (n_args != HB_N) { gb_set_error("gb_h_loop2: expected %d arguments, got %d", (int)HB_N, n_args); return -2; }
  if (nsteps < 0) return -2;
  cudaStream_t st = (cudaStream_t)stream;
  H2Args a;
  a.hs = (u32*)b[HB_HS]; a.hs64 = (u64*)b[HB_HS64];
  a.hsh[0] = (u64*)b[HB_HSH0]; a.hsh[1] = (u64*)b[HB_HSH1];
  a.glue[0] = (u32*)b[HB_GLUE0]; a.glue[1] = (u32*)b[HB_GLUE1];
  a.pos[0] = (u32*)b[HB_POS0]; a.pos[1] = (u32*)b[HB_POS1];
  a.ls[0] = (u32*)b[HB_LEN0]; a.ls[1] = (u32*)b[HB_LEN1];
  a.tid[0] = (u32*)b[HB_TID0]; a.tid[1] = (u32*)b[HB_TID1];
  a.flags = (u8*)b[HB_FLAGS]; a.rmark = (u8*)b[HB_RMARK]; a.scratch = (u32*)b[HB_SCRATCH];
  a.powtab = (const u64*)b[HB_POWTAB]; a.P = (u64)b[HB_P]; a.rbins = (const u16*)b[HB_RBINS]; a.R = b[HB_R];
  a.slots = (u32*)b[HB_SLOTS]; a.log2ns = (int)b[HB_LOG2NS]; a.bmax = (u32*)b[HB_BMAX]; a.cand = (u32*)b[HB_CAND];
  a.cap = (u32)b[HB_CAP]; a.cbit = (u32*)b[HB_CBIT];
  a.tk = (u64*)b[HB_TK]; a.tc = (u32*)b[HB_TC]; a.tr0 = (u32*)b[HB_TR0]; a.tnres = (u32*)b[HB_TNRES]; a.tlist = (u32*)b[HB_TLIST]; a.M = b[HB_M];
  a.cl_key = (u64*)b[HB_CLKEY]; a.cl_j = (u32*)b[HB_CLJ]; a.occ_j = (u32*)b[HB_OCCJ]; a.occ_r = (u32*)b[HB_OCCR];
  a.strbuf = (char*)b[HB_STRBUF]; a.stride = (int)b[HB_STRIDE]; a.erank = (const u16*)b[HB_ERANK];
  a.ev_left = (u32*)b[HB_EVL]; a.ev_right = (u32*)b[HB_EVR]; a.ev_p = (u32*)b[HB_EVP];
  a.log_key = (u64*)b[HB_LKEY]; a.log_cnt = (u32*)b[HB_LCNT]; a.log_r0 = (u32*)b[HB_LR0]; a.log_nres = (u32*)b[HB_LNRES];
  a.log_evs = (u32*)b[HB_LEVS]; a.log_nk = (u32*)b[HB_LNK];
  a.thr_num = (u32)b[HB_THRNUM];
  a.max_loops = (u32)(4 * nsteps + 64);
  a.rbase = (u32)b[HB_RBASE]; a.world = b[HB_WORLD] > 1 ? (int)b[HB_WORLD] : 1; a.rank = (int)b[HB_RANK];
  a.dcap = (u32)b[HB_DCAP]; a.xcap = (u32)b[HB_XCAP];
  a.verify_inbox = b[HB_VERIFY] != 0;
  a.timeout_ns = b[HB_TIMEOUT_MS] > 0 ? (u64)b[HB_TIMEOUT_MS] * 1000000ull : H2_TIMEOUT_NS;
  for (int r = 0; r < H2_MAXW; ++r) a.peer_d[r] = a.peer_x[r] = a.peer_f[r] = 0;
  if (a.world > 1) {
    const u64* pt = (const u64*)b[HB_PEERS];
    if (!pt || a.world > H2_MAXW || a.rank < 0 || a.rank >= a.world || a.dcap < 1024 || a.xcap < 1024) {
      gb_set_error("gb_h_loop2: bad sharding arguments (world %d, rank %d, dcap %u, xcap %u)", a.world, a.rank, a.dcap, a.xcap);
      return -2;
    }
    for (int r = 0; r < a.world; ++r) { a.peer_d[r] = pt[r]; a.peer_x[r] = pt[a.world + r]; a.peer_f[r] = pt[2 * a.world + r]; }
  }
  if (a.log2ns < 12 || a.log2ns > 30 || a.M < 1024 || (a.M & (a.M - 1)) || a.cap < 1 || a.cap > KC_SET / 2 || (a.stride & 127) ||
      a.thr_num < 1 || a.thr_num > 16 || !a.hs || !a.hs64 || !a.cbit || !a.occ_j || !a.tlist || !a.erank) {
    gb_set_error("gb_h_loop2: bad arguments (log2ns %d, M %lld, cap %u, stride %d, thr_num %u)", a.log2ns, (long long)a.M, a.cap,
                 a.stride, a.thr_num);
    return -2;
  }
  const int cs = b[HB_STREAMING] ? 1 : 0;
  const int sms = gb_sm_count();
  if (!g_h2_per_sm[cs]) {
    int per_sm = 0;
    if (cs) GB_CK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, h2_run<true>, H2_THREADS, 0));
    else GB_CK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, h2_run<false>, H2_THREADS, 0));
    if (sms <= 0 || per_sm <= 0) { gb_set_error("gb_h_loop2: no co-resident CTAs (sms %d, per SM %d)", sms, per_sm); return -3; }
    g_h2_per_sm[cs] = per_sm;
  }
  const int per_sm = g_h2_per_sm[cs];
  int want = (int)b[HB_CTAS] > 0 ? (int)b[HB_CTAS] : 2;
  if (want > per_sm) want = per_sm;
  const int grid = sms * want;
  h2_set_target<<<1, 1, 0, st>>>(a.hs, (u32)nsteps);
  void* params[] = {&a, &final_compact};
  if (cs) GB_CK(cudaLaunchCooperativeKernel((const void*)h2_run<true>, dim3(grid), dim3(H2_THREADS), params, 0, st));
  else GB_CK(cudaLaunchCooperativeKernel((const void*)h2_run<false>, dim3(grid), dim3(H2_THREADS), params, 0, st));
  GB_LAUNCH_CK();
  return 0;
}

// ---- content hashes of all live tokens under another base (recovery from a detected 64-bit collision) ---------------
// H(s_1 .. s_n) = sum s_i P^(n-i) over the token's symbol sequence r_1 g_1 r_2 ... r_k (residue symbols, glue symbols + 1):
// the same value the merges build with gb_key_h ((H(L) P + g) P^|R| + H(R)).  Dense arrays (len = plain residue count).
__global__ void __launch_bounds__(256) k_rehash_tokens(const u64* __restrict__ sym, const u32* __restrict__ glue_res,
                                                        const u32* __restrict__ pos, const u32* __restrict__ len, i64 T, u64 P,
                                                        u64* __restrict__ hsh) {
  for (i64 j = (i64)blockIdx.x * blockDim.x + threadIdx.x; j < T; j += (i64)gridDim.x * blockDim.x) {
    const u32 r0 = pos[j], n = len[j] & 0xFFFFu;
    u64 h = sym[r0];
    for (u32 k = 1; k < n; ++k) h = (h * P + ((u64)(glue_res[r0 + k - 1] & 0x3FFFFFFFu) + 1ull)) * P + sym[r0 + k];
    hsh[j] = h;
  }
}
extern "C" int gb_rehash_tokens(const uint64_t* sym, const uint32_t* glue_res, const uint32_t* pos, const uint32_t* len, int64_t T,
                                uint64_t P, uint64_t* hsh, void* stream) {
  if (T <= 0) return 0;
  int sms = gb_sm_count();
  k_rehash_tokens<<<gb_blocks_for(T, 256, (sms > 0 ? sms : 148) * 8), 256, 0, (cudaStream_t)stream>>>(
      (const u64*)sym, glue_res, pos, len, T, (u64)P, (u64*)hsh);
  GB_LAUNCH_CK();
  return 0;
}

// sizes of the symmetric block a sharded run allocates per rank (gb_peer_alloc): [delta x 2 | export x 2 | flags]
extern "C" int64_t gb_h_loop2_delta_bytes(int64_t dcap, int world) { return 2 * (int64_t)world * (32 + dcap * 24); }
extern "C" int64_t gb_h_loop2_export_bytes(int64_t xcap, int world) { return 2 * (int64_t)world * (32 + xcap * 20); }
extern "C" int64_t gb_h_loop2_flag_bytes(void) { return 2 * H2_MAXW * 4; }
extern "C" int gb_h_loop2_max_world(void) { return H2_MAXW; }
